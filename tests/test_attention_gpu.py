"""Fused causal attention (fx_attn_causal_fwd / _bwd through the C ABI) against the oracle's
_fuse_prepare_qkv + core_attn restatement (hybrid_model.py:165-186, 268-298).

Tolerance: bf16 output of fp32 math -> max|a-b| / max|b| <= 1e-2 for the forward (north_star's
bound is 2e-2); gradients go through one more bf16 rounding of P / dS, bound 2e-2.
"""
import pytest
import torch

from conftest import rel_err
from oracle import decoder_layer as O
from oracle import philox as OP

pytestmark = pytest.mark.gpu

FWD_TOL = 1e-2
BWD_TOL = 2e-2

SHAPES = [
    (1, 128, 1, 64), (2, 256, 2, 64), (1, 384, 3, 128), (2, 200, 2, 64), (1, 1024, 2, 64),
    (1, 8, 1, 64), (1, 136, 5, 128), (1, 640, 2, 128),
]


def _qkv(b, s, nh, hd, seed=0):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(b, s, nh * 3 * hd, generator=g).to(torch.bfloat16)


def _oracle(qkv, nh, hd, p, keep, dout=None):
    mix = qkv.float().requires_grad_(True)
    q, k, v = O.fuse_prepare_qkv(mix, nh, hd)
    out = O.core_attn(q, k, v, hd, p, keep)
    if dout is None:
        return out.detach(), None
    out.backward(dout.float())
    return out.detach(), mix.grad


@pytest.mark.parametrize("b,s,nh,hd", SHAPES)
@pytest.mark.parametrize("p", [0.0, 0.1])
def test_attn_fwd(b, s, nh, hd, p):
    from fleetx_b200 import ops

    qkv = _qkv(b, s, nh, hd)
    seed, offset = 1147, 5
    keep = torch.from_numpy(OP.attn_dropout_mask(seed, offset, b, nh, s, p)) if p > 0 else None
    ref, _ = _oracle(qkv, nh, hd, p, keep)
    out, lse, amask = ops.attn_causal_fwd(qkv.cuda(), nh, hd, hd ** -0.5, p, seed, offset)
    torch.cuda.synchronize()
    assert rel_err(out, ref) < FWD_TOL
    # lse against the oracle's scores
    q, k, _ = O.fuse_prepare_qkv(qkv.float(), nh, hd)
    sc = torch.matmul(q, k.transpose(-1, -2)) * hd ** -0.5
    sc = sc.masked_fill(torch.triu(torch.ones(s, s, dtype=torch.bool), 1), float("-inf"))
    assert rel_err(lse, torch.logsumexp(sc, -1)) < 1e-3


def test_attn_causality():
    """Output at position t must not depend on inputs at positions > t."""
    from fleetx_b200 import ops

    b, s, nh, hd = 1, 256, 2, 64
    qkv = _qkv(b, s, nh, hd, seed=3)
    out1, _, _ = ops.attn_causal_fwd(qkv.cuda(), nh, hd, hd ** -0.5)
    qkv2 = qkv.clone()
    qkv2[:, 150:] = _qkv(b, s - 150, nh, hd, seed=4)
    out2, _, _ = ops.attn_causal_fwd(qkv2.cuda(), nh, hd, hd ** -0.5)
    assert torch.equal(out1[:, :150], out2[:, :150])
    assert not torch.equal(out1[:, 150:], out2[:, 150:])


@pytest.mark.parametrize("b,s,nh,hd", SHAPES)
@pytest.mark.parametrize("p", [0.0, 0.1])
def test_attn_bwd(b, s, nh, hd, p):
    from fleetx_b200 import ops

    qkv = _qkv(b, s, nh, hd, seed=1)
    dout = torch.randn(b, s, nh * hd, generator=torch.Generator().manual_seed(2)).to(torch.bfloat16)
    seed, offset = 1147, 9
    keep = torch.from_numpy(OP.attn_dropout_mask(seed, offset, b, nh, s, p)) if p > 0 else None
    _, dref = _oracle(qkv, nh, hd, p, keep, dout)
    qd = qkv.cuda()
    out, lse, amask = ops.attn_causal_fwd(qd, nh, hd, hd ** -0.5, p, seed, offset)
    dqkv = ops.attn_causal_bwd(dout.cuda(), qd, out, lse, amask, nh, hd, hd ** -0.5, p)
    torch.cuda.synchronize()
    d = dqkv.float().cpu().reshape(b, s, nh, 3, hd)
    r = dref.reshape(b, s, nh, 3, hd)
    for i, name in enumerate("qkv"):
        assert rel_err(d[..., i, :], r[..., i, :]) < BWD_TOL, name


def test_attn_rejects_bad_arguments():
    from fleetx_b200 import FleetXError, ops

    with pytest.raises(FleetXError):
        ops.attn_causal_fwd(_qkv(1, 64, 1, 32).cuda(), 1, 32, 1.0)  # head_dim 32 unsupported
    with pytest.raises(FleetXError):
        ops.attn_causal_fwd(_qkv(1, 64, 1, 64), 1, 64, 1.0)  # CPU tensor
