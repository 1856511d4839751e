"""Raw operator calls: torch tensors in, C-ABI calls out.

torch is plumbing here (device memory, the current stream); every function hands raw device
pointers to libfleetx_b200.so. Nothing in this module computes on the host.
"""
from __future__ import annotations

import torch

from . import _lib
from ._lib import FX_BF16, FX_EPI_BIAS, FX_EPI_BIAS_GELU, FX_EPI_DGELU, FX_EPI_NONE, FX_F32, check

__all__ = [
    "layernorm_fwd", "layernorm_bwd", "bias_dropout_residual_fwd", "bias_dropout_residual_bwd",
    "colsum", "gemm", "attn_causal_fwd", "attn_causal_bwd",
    "FX_EPI_NONE", "FX_EPI_BIAS", "FX_EPI_BIAS_GELU", "FX_EPI_DGELU",
]


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _p(t) -> int | None:
    return None if t is None else t.data_ptr()


def _need(t: torch.Tensor, dtype, name: str) -> torch.Tensor:
    if not t.is_cuda:
        raise _lib.FleetXError(f"{name}: expected a CUDA tensor (there is no CPU path)")
    if t.dtype != dtype:
        raise _lib.FleetXError(f"{name}: expected {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise _lib.FleetXError(f"{name}: expected a contiguous tensor")
    return t


_ws_cache: dict = {}


def _workspace(nbytes: int, device) -> torch.Tensor:
    """Per-device, per-stream scratch buffer, grown on demand (caller-owned per the ABI)."""
    key = (device, torch.cuda.current_stream(device).cuda_stream)
    buf = _ws_cache.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
        _ws_cache[key] = buf
    return buf


# ------------------------------------------------------------------------------- LayerNorm
def layernorm_fwd(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, eps: float = 1e-5, out=None):
    """out: optional (y, mean, rstd) to write into (row slices of larger buffers)."""
    h = x.shape[-1]
    x2 = _need(x, torch.bfloat16, "x").view(-1, h)
    rows = x2.shape[0]
    _need(gamma, torch.float32, "gamma"), _need(beta, torch.float32, "beta")
    if out is not None:
        y, mean, rstd = out
        _need(y, torch.bfloat16, "y"), _need(mean, torch.float32, "mean"), _need(rstd, torch.float32, "rstd")
    else:
        y = torch.empty_like(x2)
        mean = torch.empty(rows, dtype=torch.float32, device=x.device)
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
    check(_lib.lib().fx_layernorm_fwd(_stream(), _p(x2), _p(gamma), _p(beta), _p(y), _p(mean), _p(rstd),
                                      rows, h, float(eps)), "fx_layernorm_fwd")
    return y.view(x.shape), mean, rstd


def layernorm_bwd(dy, x, gamma, mean, rstd, dresidual=None):
    h = x.shape[-1]
    x2 = _need(x, torch.bfloat16, "x").view(-1, h)
    dy2 = _need(dy, torch.bfloat16, "dy").view(-1, h)
    rows = x2.shape[0]
    dr = None if dresidual is None else _need(dresidual, torch.bfloat16, "dresidual")
    dx = torch.empty_like(x2)
    dgamma = torch.empty(h, dtype=torch.float32, device=x.device)
    dbeta = torch.empty(h, dtype=torch.float32, device=x.device)
    ws = _workspace(_lib.lib().fx_layernorm_bwd_workspace_bytes(rows, h), x.device)
    check(_lib.lib().fx_layernorm_bwd(_stream(), _p(dy2), _p(x2), _p(gamma), _p(mean), _p(rstd), _p(dr),
                                      _p(dx), _p(dgamma), _p(dbeta), _p(ws), rows, h), "fx_layernorm_bwd")
    return dx.view(x.shape), dgamma, dbeta


# ------------------------------------------------------------------------------- dropout
def bias_dropout_residual_fwd(y, bias, residual, p: float, seed: int, offset: int,
                              mask_in=None, return_mask: bool = False):
    h = y.shape[-1]
    y2 = _need(y, torch.bfloat16, "y").view(-1, h)
    r2 = _need(residual, torch.bfloat16, "residual").view(-1, h)
    if bias is not None:
        _need(bias, torch.bfloat16, "bias")
    if mask_in is not None:
        _need(mask_in, torch.uint8, "mask_in")
    out = torch.empty_like(y2)
    mask_out = torch.empty(y2.shape, dtype=torch.uint8, device=y.device) if return_mask else None
    check(_lib.lib().fx_bias_dropout_residual_fwd(_stream(), _p(y2), _p(bias), _p(r2), _p(out), y2.shape[0], h,
                                                  float(p), int(seed), int(offset), _p(mask_in), _p(mask_out)),
          "fx_bias_dropout_residual_fwd")
    out = out.view(y.shape)
    return (out, mask_out.view(y.shape)) if return_mask else out


def bias_dropout_residual_bwd(dout, p: float, seed: int, offset: int, mask_in=None):
    h = dout.shape[-1]
    d2 = _need(dout, torch.bfloat16, "dout").view(-1, h)
    if mask_in is not None:
        _need(mask_in, torch.uint8, "mask_in")
    dy = torch.empty_like(d2)
    check(_lib.lib().fx_bias_dropout_residual_bwd(_stream(), _p(d2), _p(dy), d2.shape[0], h, float(p),
                                                  int(seed), int(offset), _p(mask_in)),
          "fx_bias_dropout_residual_bwd")
    return dy.view(dout.shape)


def colsum(x: torch.Tensor, out_dtype=torch.bfloat16, out: torch.Tensor | None = None) -> torch.Tensor:
    n = x.shape[-1]
    x2 = _need(x, torch.bfloat16, "x").view(-1, n)
    rows = x2.shape[0]
    if out is None:
        out = torch.empty(n, dtype=out_dtype, device=x.device)
    else:
        out_dtype = out.dtype
        if out.numel() != n or not out.is_contiguous() or out_dtype not in (torch.float32, torch.bfloat16):
            raise _lib.FleetXError("colsum: bad output tensor")
    ws = _workspace(_lib.lib().fx_colsum_workspace_bytes(rows, n), x.device)
    check(_lib.lib().fx_colsum_bf16(_stream(), _p(x2), n, _p(out), FX_F32 if out_dtype == torch.float32 else FX_BF16,
                                    _p(ws), rows, n), "fx_colsum_bf16")
    return out


# ------------------------------------------------------------------------------- GEMM
_gemm_profile = None


def profile_gemm(enable: bool):
    """bench.py hook: while enabled, every gemm() is bracketed by CUDA events on its stream.
    Turning it off returns the records [(flops, start_event, end_event), ...]."""
    global _gemm_profile
    if enable:
        _gemm_profile = []
        return None
    rec, _gemm_profile = _gemm_profile, None
    return rec or []


import contextlib


@contextlib.contextmanager
def sm_budget(ctas: int):
    """While active, the persistent GEMM kernels launched by this thread use at most `ctas` CTAs (0: no limit) —
    for the stretch of a step where a cross-rank collective holds some SMs on another stream."""
    old = _lib.lib().fx_set_sm_budget(int(ctas))
    try:
        yield
    finally:
        _lib.lib().fx_set_sm_budget(old)


def gemm(a: torch.Tensor, b: torch.Tensor, *, trans_a: bool = False, trans_b: bool = False,
         out: torch.Tensor | None = None, out_dtype=torch.bfloat16, accumulate: bool = False,
         epilogue: int = FX_EPI_NONE, bias=None, aux=None, split_k: bool = False) -> torch.Tensor:
    """C[M,N] = op(A) · op(B); a, b are 2-D contiguous bf16 (see fx_gemm_bf16 for the layouts). split_k: let the
    library cut K into ranges when the output has too few tiles to fill the machine (weight gradients at a high
    mp degree; fx_gemm_bf16_splitk)."""
    for name, t in (("a", a), ("b", b)):
        if not t.is_cuda or t.dtype != torch.bfloat16 or t.dim() != 2 or t.stride(1) != 1:
            raise _lib.FleetXError(f"gemm: {name} must be a 2-D CUDA bf16 tensor with unit column stride")
    M, K = (a.shape[1], a.shape[0]) if trans_a else (a.shape[0], a.shape[1])
    Kb, N = (b.shape[1], b.shape[0]) if trans_b else (b.shape[0], b.shape[1])
    if K != Kb:
        raise _lib.FleetXError(f"gemm: contraction mismatch {K} vs {Kb}")
    if out is None:
        if accumulate:
            raise _lib.FleetXError("gemm: accumulate needs an existing output")
        out = torch.empty(M, N, dtype=out_dtype, device=a.device)
    c_dtype = FX_F32 if out.dtype == torch.float32 else FX_BF16
    if out.dtype not in (torch.float32, torch.bfloat16) or not out.is_contiguous() or tuple(out.shape) != (M, N):
        raise _lib.FleetXError("gemm: bad output tensor")
    if bias is not None:
        _need(bias, torch.bfloat16, "bias")
    if aux is not None:
        _need(aux, torch.bfloat16, "aux")
    if _gemm_profile is not None:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    splits = 1
    if split_k and not accumulate and epilogue == FX_EPI_NONE:
        splits = _lib.lib().fx_gemm_splitk_splits(M, N, K)
    if splits > 1:
        ws = _workspace(_lib.lib().fx_gemm_splitk_workspace_bytes(M, N, splits), a.device)
        check(_lib.lib().fx_gemm_bf16_splitk(_stream(), int(trans_a), int(trans_b), M, N, K, _p(a), a.stride(0), _p(b),
                                             b.stride(0), _p(out), c_dtype, splits, _p(ws)), "fx_gemm_bf16_splitk")
    else:
        check(_lib.lib().fx_gemm_bf16(_stream(), int(trans_a), int(trans_b), M, N, K, _p(a), a.stride(0), _p(b),
                                      b.stride(0), _p(out), N, c_dtype, int(accumulate), int(epilogue), _p(bias),
                                      _p(aux), N if aux is not None else 0), "fx_gemm_bf16")
    if _gemm_profile is not None:
        e1.record()
        _gemm_profile.append((2.0 * M * N * K, e0, e1))
    return out


# ------------------------------------------------------------------------------- attention
def attn_causal_fwd(qkv: torch.Tensor, nh: int, hd: int, scale: float, p_drop: float = 0.0,
                    seed: int = 0, offset: int = 0, batch0: int = 0, out=None, lse=None, mask=None,
                    mask_ready: bool = False):
    """qkv: bf16 [b, s, nh*3*hd] (packed per head q|k|v). Returns (out [b, s, nh*hd], lse [b, nh, s],
    mask): mask holds the dropout keep bits (None when p_drop == 0) and goes to attn_causal_bwd. batch0: these are
    sequences [batch0, batch0 + b) of a larger batch (same keep bits as the whole-batch call)."""
    _need(qkv, torch.bfloat16, "qkv")
    b, s = qkv.shape[0], qkv.shape[1]
    if qkv.numel() != b * s * nh * 3 * hd:
        raise _lib.FleetXError("attn_causal_fwd: qkv shape does not match (nh, hd)")
    if out is None:
        out = torch.empty(b, s, nh * hd, dtype=torch.bfloat16, device=qkv.device)
    else:  # caller-provided destinations (the batch part of larger tensors)
        _need(out, torch.bfloat16, "out")
        if out.numel() != b * s * nh * hd:
            raise _lib.FleetXError("attn_causal_fwd: bad out tensor")
    if lse is None:
        lse = torch.empty(b, nh, s, dtype=torch.float32, device=qkv.device)
    else:
        _need(lse, torch.float32, "lse")
        if lse.numel() != b * nh * s:
            raise _lib.FleetXError("attn_causal_fwd: bad lse tensor")
    if p_drop > 0:
        nbytes = _lib.lib().fx_attn_dropout_mask_bytes(b, s, nh)
        if mask is None:
            mask = torch.empty(nbytes, dtype=torch.uint8, device=qkv.device)
        elif _need(mask, torch.uint8, "mask").numel() != nbytes:
            raise _lib.FleetXError("attn_causal_fwd: bad mask tensor")
    else:
        mask = None
    check(_lib.lib().fx_attn_causal_fwd_at(_stream(), _p(qkv), _p(out), _p(lse), _p(mask), b, s, nh, hd, float(scale),
                                           float(p_drop), int(seed), int(offset), int(batch0), int(mask_ready)),
          "fx_attn_causal_fwd")
    return out, lse, mask


def attn_dropout_mask(b: int, s: int, nh: int, p_drop: float, seed: int, offset: int, device, batch0: int = 0):
    """The keep bits of attention dropout on their own (pass the result to attn_causal_fwd with mask_ready=True)."""
    mask = torch.empty(_lib.lib().fx_attn_dropout_mask_bytes(b, s, nh), dtype=torch.uint8, device=device)
    check(_lib.lib().fx_attn_dropout_mask(_stream(), _p(mask), b, s, nh, float(p_drop), int(seed), int(offset),
                                          int(batch0)), "fx_attn_dropout_mask")
    return mask


def attn_causal_bwd(dout, qkv, out, lse, mask, nh: int, hd: int, scale: float, p_drop: float = 0.0, dqkv=None):
    _need(dout, torch.bfloat16, "dout"), _need(qkv, torch.bfloat16, "qkv"), _need(out, torch.bfloat16, "out")
    _need(lse, torch.float32, "lse")
    b, s = qkv.shape[0], qkv.shape[1]
    if dqkv is None:
        dqkv = torch.empty_like(qkv)
    elif _need(dqkv, torch.bfloat16, "dqkv").numel() != qkv.numel():
        raise _lib.FleetXError("attn_causal_bwd: bad dqkv tensor")
    ws = _workspace(_lib.lib().fx_attn_causal_bwd_workspace_bytes(b, s, nh, hd), qkv.device)
    check(_lib.lib().fx_attn_causal_bwd(_stream(), _p(dout), _p(qkv), _p(out), _p(lse), _p(mask), _p(dqkv), _p(ws), b,
                                        s, nh, hd, float(scale), float(p_drop)), "fx_attn_causal_bwd")
    return dqkv


# ------------------------------------------------------------------------------- KV cache (inference)
def kv_cache_append(qkv: torch.Tensor, cache_k: torch.Tensor, cache_v: torch.Tensor, nh: int, hd: int, L: int) -> None:
    """qkv bf16 [b, t, nh*3*hd]; cache_k / cache_v bf16 [b, nh, cap, hd]: write the new k, v into rows [L, L + t)."""
    _need(qkv, torch.bfloat16, "qkv"), _need(cache_k, torch.bfloat16, "cache_k"), _need(cache_v, torch.bfloat16, "cache_v")
    b, t = qkv.shape[0], qkv.shape[1]
    check(_lib.lib().fx_kv_cache_append(_stream(), _p(qkv), _p(cache_k), _p(cache_v), b, t, nh, hd, cache_k.shape[2], int(L)),
          "fx_kv_cache_append")


def attn_decode(qkv: torch.Tensor, cache_k: torch.Tensor, cache_v: torch.Tensor, nh: int, hd: int, L: int, scale: float):
    """Attention of the t new tokens over cache rows [0, L + i] (after kv_cache_append). -> bf16 [b, t, nh*hd]."""
    _need(qkv, torch.bfloat16, "qkv"), _need(cache_k, torch.bfloat16, "cache_k"), _need(cache_v, torch.bfloat16, "cache_v")
    b, t = qkv.shape[0], qkv.shape[1]
    out = torch.empty(b, t, nh * hd, dtype=torch.bfloat16, device=qkv.device)
    check(_lib.lib().fx_attn_decode(_stream(), _p(qkv), _p(cache_k), _p(cache_v), _p(out), b, t, nh, hd, cache_k.shape[2], int(L),
                                    float(scale)), "fx_attn_decode")
    return out
