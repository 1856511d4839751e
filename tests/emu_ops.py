"""TEST DOUBLE for fleetx_b200.ops — never part of the product.

The decoder layer's host side (fleetx_b200/functional.py) is a hand-sequenced forward and a hand-derived
backward: which kernel runs when, what is saved, where the collectives sit, how recompute and sequence
parallelism reorder things. On the GPU that logic is only visible through bf16 kernels (2e-2 tolerance).
Here every `ops.*` call is replaced by a plain torch-CPU statement of what that kernel computes, in the
dtype of its inputs, so the SAME host code can run in fp64 on CPU and be held to the reference's own
outputs and gradients (tests/golden/ref_*.npz) at 1e-10 — a missing term or a misplaced collective in the
adjoint cannot hide behind bf16 rounding.

`install(monkeypatch)` swaps the functions into fleetx_b200.ops for one test; the product never imports this
file, and the real ops still refuse CPU tensors (tests/test_abi.py::test_no_cpu_fallback).
"""
from __future__ import annotations

import math

import numpy as np
import torch

from oracle import philox

FX_EPI_NONE, FX_EPI_BIAS, FX_EPI_BIAS_GELU, FX_EPI_DGELU = 0, 1, 2, 3
_K = math.sqrt(2.0 / math.pi)

# parity hook: keep mask(s) [b, nh, s, s] to use instead of the Philox draw (one per attn_causal_fwd call)
ATTN_KEEP_QUEUE: list = []
CALLS: list = []  # names of the emulated kernels in launch order


def _gelu(z):
    return 0.5 * z * (1.0 + torch.tanh(_K * (z + 0.044715 * z ** 3)))


def _dgelu(z):
    t = torch.tanh(_K * (z + 0.044715 * z ** 3))
    return 0.5 * (1.0 + t) + 0.5 * z * (1.0 - t * t) * _K * (1.0 + 3 * 0.044715 * z * z)


def _hidden_keep(shape, p, seed, offset, mask_in):
    if mask_in is not None:
        return mask_in.reshape(shape).bool()
    if p == 0:
        return None
    n = int(np.prod(shape))
    return torch.from_numpy(np.asarray(philox.dropout_mask(seed, offset, n, p))).reshape(shape).bool()


# ------------------------------------------------------------------------------- fx_layernorm_*
def layernorm_fwd(x, gamma, beta, eps=1e-5, out=None):
    CALLS.append("layernorm_fwd")
    h = x.shape[-1]
    x2 = x.reshape(-1, h)
    mean = x2.mean(-1)
    rstd = torch.rsqrt(((x2 - mean[:, None]) ** 2).mean(-1) + eps)
    y = (x2 - mean[:, None]) * rstd[:, None] * gamma + beta
    if out is not None:
        out[0].copy_(y.view(out[0].shape)), out[1].copy_(mean), out[2].copy_(rstd)
        return out[0], out[1], out[2]
    return y.view(x.shape), mean, rstd


def layernorm_bwd(dy, x, gamma, mean, rstd, dresidual=None):
    CALLS.append("layernorm_bwd")
    h = x.shape[-1]
    x2, dy2 = x.reshape(-1, h), dy.reshape(-1, h)
    xhat = (x2 - mean[:, None]) * rstd[:, None]
    dxhat = dy2 * gamma
    dx = rstd[:, None] * (dxhat - dxhat.mean(-1, keepdim=True) - xhat * (dxhat * xhat).mean(-1, keepdim=True))
    if dresidual is not None:
        dx = dx + dresidual.reshape(-1, h)
    return dx.view(x.shape), (dy2 * xhat).sum(0), dy2.sum(0)


# ------------------------------------------------------------------------------- fx_bias_dropout_residual_*
def bias_dropout_residual_fwd(y, bias, residual, p, seed, offset, mask_in=None, return_mask=False):
    CALLS.append("bias_dropout_residual_fwd")
    v = y if bias is None else y + bias
    keep = _hidden_keep(tuple(y.shape), p, seed, offset, mask_in)
    if keep is not None:
        v = v * keep.to(v.dtype) / (1.0 - p)
    out = residual + v
    if return_mask:
        return out, (torch.ones_like(y, dtype=torch.uint8) if keep is None else keep.to(torch.uint8))
    return out


def bias_dropout_residual_bwd(dout, p, seed, offset, mask_in=None):
    CALLS.append("bias_dropout_residual_bwd")
    keep = _hidden_keep(tuple(dout.shape), p, seed, offset, mask_in)
    return dout.clone() if keep is None else dout * keep.to(dout.dtype) / (1.0 - p)


def colsum(x, out_dtype=None, out=None):
    CALLS.append("colsum")
    r = x.reshape(-1, x.shape[-1]).sum(0)
    if out is not None:
        out.copy_(r)
        return out
    return r


# ------------------------------------------------------------------------------- fx_gemm_bf16
import contextlib


@contextlib.contextmanager
def sm_budget(ctas):
    yield


def gemm(a, b, *, trans_a=False, trans_b=False, out=None, out_dtype=None, accumulate=False, epilogue=FX_EPI_NONE,
         bias=None, aux=None, split_k=False):
    CALLS.append("gemm")
    acc = (a.t() if trans_a else a) @ (b.t() if trans_b else b)
    if epilogue == FX_EPI_BIAS:
        c = acc + bias
    elif epilogue == FX_EPI_BIAS_GELU:
        z = acc + bias
        if aux is not None:
            aux.copy_(z)
        c = _gelu(z)
    elif epilogue == FX_EPI_DGELU:
        c = acc * _dgelu(aux)
    else:
        c = acc
    if out is None:
        assert not accumulate
        return c
    assert tuple(out.shape) == tuple(c.shape)
    if accumulate:
        out.add_(c)
    else:
        out.copy_(c)
    return out


# ------------------------------------------------------------------------------- fx_attn_causal_*
def _unpack(qkv, nh, hd):
    b, s = qkv.shape[0], qkv.shape[1]
    t = qkv.reshape(b, s, nh, 3, hd).permute(3, 0, 2, 1, 4)  # [3, b, nh, s, hd]
    return t[0], t[1], t[2]


def attn_causal_fwd(qkv, nh, hd, scale, p_drop=0.0, seed=0, offset=0, batch0=0, out=None, lse=None, mask=None):
    dst = (out, lse, mask)
    CALLS.append("attn_causal_fwd")
    b, s = qkv.shape[0], qkv.shape[1]
    q, k, v = _unpack(qkv, nh, hd)
    causal = torch.ones(s, s, dtype=torch.bool).triu(1)
    sc = (q @ k.transpose(-1, -2) * scale).masked_fill(causal, float("-inf"))
    lse = torch.logsumexp(sc, dim=-1)
    pm = torch.exp(sc - lse[..., None])
    keep = None
    if p_drop > 0:
        if ATTN_KEEP_QUEUE:  # a whole-batch mask: calls on parts of the batch take their sequences, the last one pops it
            full = ATTN_KEEP_QUEUE[0]
            keep = full[batch0:batch0 + b].bool()
            if batch0 + b >= full.shape[0]:
                ATTN_KEEP_QUEUE.pop(0)
        else:
            keep = torch.from_numpy(np.asarray(philox.attn_dropout_mask(seed, offset, batch0 + b, nh, s, p_drop)))[batch0:].bool()
        pm = pm * keep.to(pm.dtype) / (1.0 - p_drop)
    out = (pm @ v).permute(0, 2, 1, 3).reshape(b, s, nh * hd)
    if dst[0] is not None:
        dst[0].copy_(out.view(dst[0].shape))
        out = dst[0]
    if dst[1] is not None:
        dst[1].copy_(lse)
        lse = dst[1]
    if dst[2] is not None and keep is not None:
        dst[2].copy_(keep.view(dst[2].shape))
        keep = dst[2]
    return out, lse, keep


def attn_causal_bwd(dout, qkv, out, lse, mask, nh, hd, scale, p_drop=0.0, dqkv=None):
    dst = dqkv
    CALLS.append("attn_causal_bwd")
    b, s = qkv.shape[0], qkv.shape[1]
    q, k, v = _unpack(qkv, nh, hd)
    do = dout.reshape(b, s, nh, hd).permute(0, 2, 1, 3)
    o = out.reshape(b, s, nh, hd).permute(0, 2, 1, 3)
    causal = torch.ones(s, s, dtype=torch.bool).triu(1)
    sc = (q @ k.transpose(-1, -2) * scale).masked_fill(causal, float("-inf"))
    pm = torch.exp(sc - lse[..., None])
    drop = 1.0 if mask is None else mask.to(pm.dtype) / (1.0 - p_drop)
    dv = (pm * drop).transpose(-1, -2) @ do
    dp = (do @ v.transpose(-1, -2)) * drop
    delta = (do * o).sum(-1, keepdim=True)          # = rowsum(P_dropped * dP_dropped): what the kernels use
    ds = pm * (dp - delta)
    dq = ds @ k * scale
    dk = ds.transpose(-1, -2) @ q * scale
    dqkv = torch.stack([dq, dk, dv], dim=0).permute(1, 3, 2, 0, 4)  # [b, s, nh, 3, hd]
    if dst is not None:
        dst.copy_(dqkv.reshape(dst.shape))
        return dst
    return dqkv.reshape(qkv.shape)


NAMES = ("layernorm_fwd", "layernorm_bwd", "bias_dropout_residual_fwd", "bias_dropout_residual_bwd", "colsum", "gemm",
         "attn_causal_fwd", "attn_causal_bwd", "sm_budget")


def install(monkeypatch):
    import fleetx_b200.ops as real

    for n in NAMES:
        monkeypatch.setattr(real, n, globals()[n])
    ATTN_KEEP_QUEUE.clear()
    CALLS.clear()


def install_permanently():
    """For spawned gloo workers (their process ends with the test)."""
    import fleetx_b200.ops as real

    for n in NAMES:
        setattr(real, n, globals()[n])
    ATTN_KEEP_QUEUE.clear()
    CALLS.clear()
