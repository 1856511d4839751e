"""Generate the committed known-answer fixtures under tests/golden/ from the oracle.

The reference holds no golden vectors for this path and Paddle cannot be imported
here, so these fixtures pin the ORACLE (regression) and give the GPU tests committed vectors to
hit. Vectors from the reference's own Python are made by tools/make_reference_golden.py
(tests/golden/ref_*.npz); tests/test_reference_golden.py also replays the fixtures written here
through the reference's layer.

  python tools/make_golden.py
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import decoder_layer as O  # noqa: E402
from oracle import philox  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def layer_case(name, b, s, h, nh, f, n, dropout, dtype, round_bf16, seed=1024):
    cfg = O.LayerConfig(d_model=h, nhead=nh, dim_feedforward=f, dropout=dropout, num_partitions=n)
    params = O.init_params(cfg, seed, dtype=dtype)
    g = torch.Generator().manual_seed(seed + 7)
    for k in params:  # non-trivial biases and LayerNorm affine
        if k.endswith("bias"):
            params[k] = (0.02 * torch.randn(params[k].shape, generator=g)).to(dtype)
        elif k.startswith("norm"):
            params[k] = (1 + 0.1 * torch.randn(params[k].shape, generator=g)).to(dtype)
    x = torch.randn(b, s, h, generator=g).to(dtype)
    dy = torch.randn(b, s, h, generator=g).to(dtype)
    if round_bf16:
        rb = lambda t: t.to(torch.bfloat16).to(dtype)  # noqa: E731
        x, dy = rb(x), rb(dy)
        params = {k: (v if k.startswith("norm") else rb(v)) for k, v in params.items()}
    masks = O.make_masks(cfg, b, s, seed) if dropout > 0 else None
    y, dx, grads = O.decoder_layer_fwd_bwd(x, dy, params, cfg, masks)
    out = {"x": x.numpy(), "dy": dy.numpy(), "y": y.numpy(), "dx": dx.numpy(),
           "cfg": np.array([b, s, h, nh, f, n, int(round(dropout * 1000)), seed], dtype=np.int64)}
    for k, v in params.items():
        out["param." + k] = v.numpy()
    for k, v in grads.items():
        out["grad." + k] = v.numpy()
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    print(name, {k: v.shape for k, v in out.items() if k in ("x", "y")})


def main():
    os.makedirs(OUT, exist_ok=True)
    # oracle regression pins (fp64, tiny, hand-checkable sizes), with and without simulated TP
    layer_case("oracle_tiny_fp64_n1", 2, 8, 16, 4, 64, 1, 0.0, torch.float64, False)
    layer_case("oracle_tiny_fp64_n2_dropout", 2, 8, 16, 4, 64, 2, 0.1, torch.float64, False)
    # vectors the CUDA path is compared with on the GPU box (head_dim 64, bf16-representable inputs)
    layer_case("layer_gpu_small_fp32", 2, 8, 128, 2, 512, 1, 0.0, torch.float32, True)
    layer_case("layer_gpu_small_fp32_dropout", 2, 40, 128, 2, 512, 1, 0.1, torch.float32, True)
    # Philox keep masks
    np.savez_compressed(os.path.join(OUT, "philox_masks.npz"),
                        words=philox.philox4x32_10(1024, 3, np.arange(16, dtype=np.uint64)),
                        hidden=philox.dropout_mask(1024, 0, 4 * 64, 0.1),
                        attn=philox.attn_dropout_mask(1147, 2, 1, 2, 24, 0.25))


if __name__ == "__main__":
    main()
