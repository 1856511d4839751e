"""FusedAdamW (fx_multi_sq_norm + fx_multi_adamw through fleetx_b200.optim) against the oracle's restatement
of paddle AdamW + ClipGradByGlobalNorm. fp32 state must agree to 1e-5 relative (the second moment to 5e-5),
bf16 parameters are the rounded master weights."""
import pytest
import torch

from conftest import rel_err
from oracle import optim as OO

pytestmark = pytest.mark.gpu


def test_fused_adamw_matches_oracle_over_three_steps():
    from fleetx_b200.optim import FusedAdamW

    g = torch.Generator().manual_seed(0)
    shapes = {"linear1.weight": (300, 520), "linear1.bias": (520,), "norm1.weight": (264,), "big.weight": (70000 * 2,),
              "qkv.weight": (128, 384)}
    dtypes = {"norm1.weight": torch.float32}
    params = {k: torch.randn(*s, generator=g) * 0.1 for k, s in shapes.items()}
    dev_params = []
    for k, v in params.items():
        p = torch.nn.Parameter(v.to(dtypes.get(k, torch.bfloat16)).cuda())
        dev_params.append((k, p))
    ref = {k: p.detach().float().cpu().double() for k, p in dev_params}  # masters start from the rounded values
    opt = FusedAdamW(None, named_parameters=dev_params, lr=1e-2, weight_decay=0.1, grad_clip=1.0)
    state = {}
    for t in range(1, 4):
        grads = {k: (torch.randn(*s, generator=g) * (3.0 if t == 1 else 0.01)).to(dtypes.get(k, torch.bfloat16)) for k, s in shapes.items()}
        for k, p in dev_params:
            p.grad = grads[k].cuda()
        opt.step()
        gn = OO.adamw_step(ref, {k: v.double() for k, v in grads.items()}, state, t, 1e-2, weight_decay=0.1, clip_norm=1.0)
        torch.cuda.synchronize()
        assert abs(opt.global_grad_norm() - gn) < 1e-4 * gn  # step 1 clips (norm >> 1), steps 2-3 do not
        for i, (k, p) in enumerate(dev_params):
            st = opt.state[i]
            assert rel_err(st["m"].view(shapes[k]), state[k]["m"]) < 1e-5, (t, k)
            # 1 - beta2 in fp32 (0.999f = 0.99900001287...) is 1.3e-5 off the fp64 value the oracle uses
            assert rel_err(st["v"].view(shapes[k]), state[k]["v"]) < 5e-5, (t, k)
            if st["master"] is not None:
                assert rel_err(st["master"], ref[k]) < 1e-5, (t, k)
                assert torch.equal(p.detach(), st["master"].to(torch.bfloat16))
            else:
                assert rel_err(p, ref[k]) < 1e-5, (t, k)
    # "bias" / "norm" parameters never decay: with zero gradients and zero state they must not move
    z = [("x.bias", torch.nn.Parameter(torch.ones(64, device="cuda"))), ("x.weight", torch.nn.Parameter(torch.ones(64, device="cuda")))]
    o2 = FusedAdamW(None, named_parameters=z, lr=0.5, weight_decay=0.1, grad_clip=None)
    for _, p in z:
        p.grad = torch.zeros_like(p)
    o2.step()
    torch.cuda.synchronize()
    assert torch.equal(z[0][1].detach(), torch.ones(64, device="cuda"))
    assert torch.allclose(z[1][1].detach(), torch.full((64,), 0.95, device="cuda"))
