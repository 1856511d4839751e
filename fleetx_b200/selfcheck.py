"""smoke(): one small decoder-layer forward+backward on cuda:0 through the C ABI, checked
against the oracle (CPU restatement of the reference layer). Used by __graft_entry__.smoke()."""
from __future__ import annotations

import torch


def layer_from_oracle_params(cfg, params, group=None, device="cuda"):
    """Build a TransformerDecoderLayer holding this rank's shard of oracle-format parameters."""
    from .layers import TransformerDecoderLayer

    layer = TransformerDecoderLayer(cfg.d_model, cfg.nhead, cfg.dim_feedforward, dropout=cfg.dropout,
                                    attn_dropout=cfg.attn_dropout, act_dropout=cfg.act_dropout,
                                    normalize_before=cfg.normalize_before, num_partitions=cfg.num_partitions,
                                    mp_group=group, device=device)
    layer.load_full_state(params)
    return layer


def rel_err(a: torch.Tensor, b: torch.Tensor) -> float:
    a = a.detach().to("cpu", torch.float64)
    b = b.detach().to("cpu", torch.float64)
    denom = b.abs().max().item()
    return (a - b).abs().max().item() / (denom if denom > 0 else 1.0)


def smoke(verbose: bool = True) -> dict:
    from oracle import decoder_layer as O  # the checker; never on the product path
    from .rng import set_seed

    if not torch.cuda.is_available():
        raise RuntimeError("smoke() needs a CUDA device (B200); there is no CPU path")
    torch.cuda.set_device(0)
    b, s = 2, 256
    cfg = O.LayerConfig(d_model=256, nhead=4, dim_feedforward=1024, dropout=0.1)
    seed = 1024
    params = {k: v.to(torch.bfloat16).float() for k, v in O.init_params(cfg, seed).items()}
    g = torch.Generator().manual_seed(7)
    x = torch.randn(b, s, cfg.d_model, generator=g).to(torch.bfloat16)
    dy = torch.randn(b, s, cfg.d_model, generator=g).to(torch.bfloat16)

    set_seed(seed)
    layer = layer_from_oracle_params(cfg, params)
    layer.train()
    xd = x.cuda().requires_grad_(True)
    y = layer(xd)
    y.backward(dy.cuda())
    torch.cuda.synchronize()

    masks = O.make_masks(cfg, b, s, seed)  # the same (seed, offset) draws, restated on the CPU
    y_ref, dx_ref, g_ref = O.decoder_layer_fwd_bwd(x.float(), dy.float(), params, cfg, masks)
    errs = {"y": rel_err(y, y_ref), "dx": rel_err(xd.grad, dx_ref)}
    for name, p in layer.named_parameters():
        errs["d" + name] = rel_err(p.grad, g_ref[name])
    worst = max(errs.values())
    if verbose:
        print("smoke: max relative error vs oracle: %.3e (%s)" % (worst, max(errs, key=errs.get)))
    if not worst < 2e-2:  # north_star: <= 2e-2 relative for bf16
        raise AssertionError(f"smoke: parity with the oracle failed: {errs}")
    return errs
