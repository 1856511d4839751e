"""Forward / backward of the decoder layer as explicit kernel sequences, plus autograd glue.

`decoder_layer_fwd` / `decoder_layer_bwd` are the hot path: every tensor op in them is a call
into libfleetx_b200.so (fleetx_b200.ops); torch only allocates buffers and carries the stream.
The order of operations follows TransformerDecoderLayer.forward (hybrid_model.py:525-566) and
MultiHeadAttention.forward (:300-347); the backward is the hand-derived adjoint of that order
(Paddle autograd in the reference), with the tensor-parallel collectives of
fleet.meta_parallel.{Column,Row}ParallelLinear placed where Paddle places them:

  forward   allreduce after out_proj's local matmul and after linear2's   (_mp_allreduce)
  backward  allreduce of dX entering qkv_proj and linear1                  (_c_identity backward)
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple

import torch

from . import ops
from .ops import FX_EPI_BIAS, FX_EPI_BIAS_GELU, FX_EPI_DGELU, FX_EPI_NONE


@dataclass
class LayerParams:
    """Per-rank parameter shards, names as in the reference layer."""
    norm1_w: torch.Tensor  # fp32 [h]
    norm1_b: torch.Tensor
    qkv_w: torch.Tensor    # bf16 [h, 3h/n]
    qkv_b: torch.Tensor    # bf16 [3h/n]
    out_w: torch.Tensor    # bf16 [h/n, h]
    out_b: torch.Tensor    # bf16 [h]
    norm2_w: torch.Tensor
    norm2_b: torch.Tensor
    lin1_w: torch.Tensor   # bf16 [h, f/n]
    lin1_b: torch.Tensor   # bf16 [f/n]
    lin2_w: torch.Tensor   # bf16 [f/n, h]
    lin2_b: torch.Tensor   # bf16 [h]

    def tensors(self):
        return (self.norm1_w, self.norm1_b, self.qkv_w, self.qkv_b, self.out_w, self.out_b,
                self.norm2_w, self.norm2_b, self.lin1_w, self.lin1_b, self.lin2_w, self.lin2_b)


@dataclass
class LayerMeta:
    nh_local: int
    hd: int
    p_attn: float
    p_hidden1: float
    p_hidden2: float
    normalize_before: bool = True
    eps: float = 1e-5
    # (seed, offset) of the three dropout calls: attention ('local_seed'), dropout1, dropout2
    # ('global_seed'); see fleetx_b200.rng
    rng_attn: Tuple[int, int] = (0, 0)
    rng_h1: Tuple[int, int] = (0, 0)
    rng_h2: Tuple[int, int] = (0, 0)
    # parity-test hooks: injected keep masks for dropout1 / dropout2 (uint8 [T, h])
    mask_h1: Optional[torch.Tensor] = None
    mask_h2: Optional[torch.Tensor] = None
    # activation recompute (hybrid_model.py:332-335 'core_attn', :537-539 'full_attn', :406-408 'full'):
    # None keeps every activation; otherwise the named part of the forward is re-run inside backward
    recompute: Optional[str] = None
    # sequence_parallel (hybrid_model.py:462, 475-520): activations between the blocks are [s/n, b, h] shards
    sequence_parallel: bool = False


import os as _os

_side_streams: dict = {}
_SIDE = _os.environ.get("FX_SIDE", "1") != "0"  # tuning knob: 0 runs the side-stream kernels in line


def _side_stream(device) -> "torch.cuda.Stream":
    """One extra stream per device for small kernels that hide beside a resident GEMM (integer / memory-bound work
    next to tensor-core work): launched AFTER the GEMM, so the GEMM's CTAs take their SMs first."""
    key = torch.device(device).index if torch.device(device).index is not None else torch.cuda.current_device()
    st = _side_streams.get(key)
    if st is None:
        st = torch.cuda.Stream(device=device)
        _side_streams[key] = st
    return st


class _Beside:
    """Small memory-bound kernels (the bias-gradient column sums) on the side stream, submitted right AFTER the GEMM
    that reads the same matrix, so they run in that GEMM's shadow (its CTAs are resident first; the small kernel's
    blocks need no shared memory and fit beside them)."""

    def __init__(self, device):
        self.on = _SIDE and torch.device(device).type == "cuda"
        self.st = _side_stream(device) if self.on else None
        self.used = False

    def mark(self):
        """Call BEFORE submitting the GEMM: the point at which the small kernel's inputs are ready."""
        if not self.on:
            return None
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream())
        return ev

    def colsum(self, ev, x, out):
        if not self.on:
            return ops.colsum(x, out=out)
        with torch.cuda.stream(self.st):
            self.st.wait_event(ev)
            ops.colsum(x, out=out)
        x.record_stream(self.st)
        self.used = True
        return out

    def join(self):
        if self.on and self.used:
            torch.cuda.current_stream().wait_stream(self.st)
            self.used = False


def _attn_mask_beside(b, s, M: LayerMeta, device, batch0=0):
    """Draw the attention keep bits (a pure function of seed / offset / shape) on the side stream; call right after
    launching the QKV projection. Returns (mask or None, event to wait for before the attention kernel)."""
    if not _SIDE or not M.p_attn > 0 or not hasattr(ops, "attn_dropout_mask") or not torch.device(device).type == "cuda":
        return None, None
    st = _side_stream(device)
    nbytes_owner = torch.cuda.current_stream()
    with torch.cuda.stream(st):
        mask = ops.attn_dropout_mask(b, s, M.nh_local, M.p_attn, *M.rng_attn, device, batch0=batch0)
        ev = torch.cuda.Event()
        ev.record(st)
    mask.record_stream(nbytes_owner)  # allocated under the side stream, consumed on the caller's
    return mask, ev

# Tuning knob, OFF by default: pipelining two batch halves through the forward allreduces measured
# slower than the plain order at 2, 4 and 8 GPUs (the allreduce kernel and the other half's GEMMs
# compete for the same SMs, and half-batch GEMMs quantise worse): 2.93 -> 3.04 ms at 4, 2.49 -> 2.59 at 8.
_PIPELINE_FWD = _os.environ.get("FX_PIPELINE_FWD", "0") == "1"


def _fwd_pipelined(x: torch.Tensor, P: LayerParams, M: LayerMeta, group, nchunks: int = 2):
    """Forward for a symmetric-memory group (pre-LN): the batch is cut into `nchunks` row blocks after
    attention, and each block's fused allreduce(+bias+dropout+residual) runs on the group's stream
    while the main stream already computes the other block's GEMMs — the transfer of one half hides
    under the math of the other. Same kernels, same results as the plain path."""
    b, s, h = x.shape
    T = b * s
    dev = x.device
    x2 = x.reshape(T, h)
    scale = M.hd ** -0.5
    rows = (b // nchunks) * s
    f_local = P.lin1_w.shape[1]
    bf, f32 = torch.bfloat16, torch.float32

    ln1, mean1, rstd1 = ops.layernorm_fwd(x2, P.norm1_w, P.norm1_b, M.eps)
    qkv = ops.gemm(ln1, P.qkv_w, epilogue=FX_EPI_BIAS, bias=P.qkv_b)
    ctx, lse, attn_mask = ops.attn_causal_fwd(qkv.view(b, s, -1), M.nh_local, M.hd, scale, M.p_attn, *M.rng_attn)
    ctx2 = ctx.view(T, -1)

    part = group.partial_buffer(T, h, dev)  # symmetric: the row-parallel GEMMs write their partial sums here
    x1 = torch.empty(T, h, dtype=bf, device=dev)
    ln2 = torch.empty(T, h, dtype=bf, device=dev)
    mean2 = torch.empty(T, dtype=f32, device=dev)
    rstd2 = torch.empty(T, dtype=f32, device=dev)
    z = torch.empty(T, f_local, dtype=bf, device=dev)
    act = torch.empty(T, f_local, dtype=bf, device=dev)
    out = torch.empty(T, h, dtype=bf, device=dev)
    sl = [slice(c * rows, (c + 1) * rows) for c in range(nchunks)]
    m1 = [None if M.mask_h1 is None else M.mask_h1[k] for k in sl]
    m2 = [None if M.mask_h2 is None else M.mask_h2[k] for k in sl]

    pend = []
    for c, k in enumerate(sl):
        ops.gemm(ctx2[k], P.out_w, out=part[k])                                     # :340 local matmul
        pend.append(group.all_reduce_bdr_start(part[k], x1[k], P.out_b, x2[k], M.p_hidden1, *M.rng_h1, m1[c],
                                               c * rows * h))                       # :546-547
    pend2 = []
    for c, k in enumerate(sl):
        group.all_reduce_bdr_finish(pend[c])
        ops.layernorm_fwd(x1[k], P.norm2_w, P.norm2_b, M.eps, out=(ln2[k], mean2[k], rstd2[k]))  # :554
        ops.gemm(ln2[k], P.lin1_w, out=act[k], epilogue=FX_EPI_BIAS_GELU, bias=P.lin1_b, aux=z[k])  # :558-559
        ops.gemm(act[k], P.lin2_w, out=part[k])                                     # :558 local matmul
        pend2.append(group.all_reduce_bdr_start(part[k], out[k], P.lin2_b, x1[k], M.p_hidden2, *M.rng_h2, m2[c],
                                                c * rows * h))                      # :556-561
    for ev in pend2:
        group.all_reduce_bdr_finish(ev)
    saved = dict(x=x2, ln1=ln1, mean1=mean1, rstd1=rstd1, qkv=qkv, ctx=ctx2, lse=lse, attn_mask=attn_mask, r1=x1, x1=x1,
                 ln2=ln2, mean2=mean2, rstd2=rstd2, z=z, act=act, r2=out, mean3=None, rstd3=None, shape=(b, s, h))
    return out.view(b, s, h), saved


def _row_owner_path(T: int, h: int, M: LayerMeta, group) -> bool:
    return bool(group.nranks > 1 and getattr(group, "row_owner", False) and M.normalize_before and T % group.nranks == 0)


def _chunks(b: int, group) -> int:
    nc = int(getattr(group, "pipeline_chunks", 1))
    return nc if nc > 1 and b % nc == 0 else 1


def _batch_slice(t, b: int, c: int, bc: int):
    """Sequences [c*bc, (c+1)*bc) of a per-batch tensor (lse [b, nh, s]; the keep bits: a flat byte buffer of b equal
    parts from the kernel, or a list of per-chunk buffers)."""
    if t is None:
        return None
    if isinstance(t, (list, tuple)):
        return t[c]
    m = t.reshape(b, -1)[c * bc:(c + 1) * bc]
    return m.reshape((bc,) + tuple(t.shape[1:])) if t.dim() > 1 else m.reshape(-1)


def _rows_cat(parts):
    """Row chunks back to back in memory (views of one communication buffer) as one tensor; else concatenated."""
    if len(parts) == 1:
        return parts[0]
    p0 = parts[0]
    step = p0.numel() * p0.element_size()
    if all(t.is_contiguous() and t.data_ptr() == p0.data_ptr() + i * step and t.shape == p0.shape for i, t in enumerate(parts)):
        return p0.as_strided((p0.shape[0] * len(parts), p0.shape[1]), (p0.shape[1], 1))
    return torch.cat(parts, dim=0)


def decoder_layer_fwd_rows(x: torch.Tensor, P: LayerParams, M: LayerMeta, group):
    """decoder_layer_fwd for mp degree > 1 with the row-wise work behind each allreduce done by the rank that owns the
    rows (group.fwd_ln_* / fwd_out_*; csrc/comm_rows.cu): dropout1 + residual + LayerNorm2 and dropout2 + residual run
    on T/n rows per rank inside the collective, x1 and its statistics exist only as this rank's rows. The batch goes
    through in group.pipeline_chunks chunks, so that one chunk's collective (on the group's stream) runs underneath
    the other chunk's GEMMs and attention. Same tensors, same (seed, offset) draws — the masks are indexed by position
    in the whole tensor / batch — as the replicated order."""
    b, s, h = x.shape
    T = b * s
    nc = _chunks(b, group)
    bc, Tc = b // nc, (b // nc) * s
    dev, dt = x.device, x.dtype
    x2 = x.reshape(T, h)
    scale = M.hd ** -0.5
    ln1, mean1, rstd1 = ops.layernorm_fwd(x2, P.norm1_w, P.norm1_b, M.eps)          # :534 (replicated: HBM is cheaper than NVLink)
    A = group.partial_buffer(T, h, dev)
    qkv = torch.empty(T, P.qkv_w.shape[1], dtype=dt, device=dev)
    ctx = torch.empty(T, P.out_w.shape[0], dtype=dt, device=dev)
    lse = torch.empty(b, M.nh_local, s, dtype=mean1.dtype, device=dev)
    z = torch.empty(T, P.lin1_w.shape[1], dtype=dt, device=dev)
    act = torch.empty_like(z)
    ln2_keep = torch.empty(T, h, dtype=dt, device=dev)
    out = torch.empty(T, h, dtype=dt, device=dev)
    rows = [slice(c * Tc, (c + 1) * Tc) for c in range(nc)]
    m1 = [None if M.mask_h1 is None else M.mask_h1.reshape(T, h)[r] for r in rows]
    m2 = [None if M.mask_h2 is None else M.mask_h2.reshape(T, h)[r] for r in rows]
    masks, pend = [], []
    with ops.sm_budget(group.free_sms() if nc > 1 else 0):  # from the first collective on, its SMs are taken
        for c, r in enumerate(rows):
            ops.gemm(ln1[r], P.qkv_w, out=qkv[r], epilogue=FX_EPI_BIAS, bias=P.qkv_b)   # :166
            mask_pre, mask_ev = _attn_mask_beside(bc, s, M, dev, batch0=c * bc)          # keep bits: beside the projection
            if mask_ev is not None:
                torch.cuda.current_stream().wait_event(mask_ev)
            _, _, mk = ops.attn_causal_fwd(qkv[r].view(bc, s, -1), M.nh_local, M.hd, scale, M.p_attn, *M.rng_attn,
                                           batch0=c * bc, out=ctx[r].view(bc, s, -1), lse=lse[c * bc:(c + 1) * bc],
                                           **({} if mask_pre is None else dict(mask=mask_pre, mask_ready=True)))
            masks.append(mk)
            ops.gemm(ctx[r], P.out_w, out=A[r])                                      # :340 local matmul
            # _mp_allreduce + bias, dropout1 + residual (:546-547), norm2 (:554)
            pend.append(group.fwd_ln_start(A[r], P.out_b, x2[r], M.p_hidden1, *M.rng_h1, m1[c], P.norm2_w, P.norm2_b,
                                           M.eps, drop_off=c * Tc * h))
        x1_own, mean2, rstd2, pend2 = [], [], [], []
        for c, r in enumerate(rows):
            x1_c, mean_c, rstd_c, ln2_c = group.fwd_ln_finish(pend[c])
            x1_own.append(x1_c), mean2.append(mean_c), rstd2.append(rstd_c)
            group.keep_copy(ln2_c, ln2_keep[r])  # linear1's weight gradient needs it in backward
            ops.gemm(ln2_c, P.lin1_w, out=act[r], epilogue=FX_EPI_BIAS_GELU, bias=P.lin1_b, aux=z[r])  # :558-559
            ops.gemm(act[r], P.lin2_w, out=A[r])                                     # :558 local matmul
            pend2.append(group.fwd_out_start(A[r], P.lin2_b, x1_c, M.p_hidden2, *M.rng_h2, m2[c],
                                             drop_off=c * Tc * h))                   # :556-561
        for c, r in enumerate(rows):
            group.fwd_out_finish(pend2[c], out[r])
    saved = dict(x=x2, ln1=ln1, mean1=mean1, rstd1=rstd1, qkv=qkv, ctx=ctx, lse=lse,
                 attn_mask=(masks if nc > 1 else masks[0]), x1_own=x1_own, ln2=ln2_keep, mean2=mean2, rstd2=rstd2, z=z,
                 act=act, shape=(b, s, h), rows=True, chunks=nc)
    return out.view(b, s, h), saved


def decoder_layer_bwd_rows(dout: torch.Tensor, saved: dict, P: LayerParams, M: LayerMeta, group):
    """Adjoint of decoder_layer_fwd_rows, chunk by chunk. The two allreduces of dX (backward of _c_identity) carry
    LayerNorm backward, the skip-path adds and dropout1's backward on the owner's rows, and run on the group's stream
    underneath the other chunk's kernels and the (whole-batch) weight-gradient GEMMs; gradients of parameters whose
    kernels saw only this rank's rows (LayerNorm weight / bias, out_proj / linear2 bias) are summed over the ranks in
    one small fp32 allreduce that rides behind the last collective."""
    b, s, h = saved["shape"]
    T = b * s
    nc = saved["chunks"]
    bc, Tc = b // nc, (b // nc) * s
    d = dout.reshape(T, h)
    dev, dt = d.device, d.dtype
    scale = M.hd ** -0.5
    g = {}
    r0, r1 = group.own_rows(T)
    rows = [slice(c * Tc, (c + 1) * Tc) for c in range(nc)]
    m1 = [None if M.mask_h1 is None else M.mask_h1.reshape(T, h)[r] for r in rows]
    A = group.partial_buffer(T, h, dev)
    acc = group.acc_buffer(h, dev, P.norm1_w.dtype)  # fp32 [6, h]: dgamma2 | dbeta2 | d out_b | dgamma1 | dbeta1 | d lin2_b

    d_y2 = ops.bias_dropout_residual_bwd(d, M.p_hidden2, *M.rng_h2, mask_in=M.mask_h2)   # replicated: the GEMMs need all rows
    ops.colsum(d_y2[r0:r1], out=acc[5])                                              # own rows; summed over ranks at the end
    d_z = torch.empty_like(saved["z"])
    d_qkv = torch.empty_like(saved["qkv"])
    dx = torch.empty(T, h, dtype=dt, device=dev)
    with ops.sm_budget(group.free_sms()):  # a collective holds its SMs until every rank is through
        pend = []
        for c, r in enumerate(rows):
            ops.gemm(d_y2[r], P.lin2_w, trans_b=True, out=d_z[r], epilogue=FX_EPI_DGELU, aux=saved["z"][r])
            ops.gemm(d_z[r], P.lin1_w, trans_b=True, out=A[r])
            pend.append(group.bwd_ln_bdr_start(A[r], saved["x1_own"][c], saved["mean2"][c], saved["rstd2"][c], P.norm2_w,
                                               d[r], M.p_hidden1, *M.rng_h1, m1[c], acc, drop_off=c * Tc * h,
                                               accumulate=c > 0))
        side = _Beside(dev)
        ev = side.mark()
        g["lin2_w"] = ops.gemm(saved["act"], d_y2, trans_a=True, split_k=True)
        del d_y2
        g["lin1_b"] = side.colsum(ev, d_z, torch.empty(d_z.shape[1], dtype=dt, device=dev))  # in the GEMM's shadow
        g["lin1_w"] = ops.gemm(group.wait_kept(saved["ln2"]), d_z, trans_a=True, split_k=True)
        del d_z
        d_ao, pend2 = [], []
        for c, r in enumerate(rows):
            d_x1_c, d_ao_c = group.bwd_ln_bdr_finish(pend[c])
            d_ao.append(d_ao_c)
            d_ctx = ops.gemm(d_ao_c, P.out_w, trans_b=True)
            ops.attn_causal_bwd(d_ctx.view(bc, s, -1), saved["qkv"][r].view(bc, s, -1), saved["ctx"][r].view(bc, s, -1),
                                _batch_slice(saved["lse"], b, c, bc), _batch_slice(saved["attn_mask"], b, c, bc),
                                M.nh_local, M.hd, scale, M.p_attn, dqkv=d_qkv[r].view(bc, s, -1))
            del d_ctx
            ops.gemm(d_qkv[r], P.qkv_w, trans_b=True, out=A[r])
            pend2.append(group.bwd_ln_start(A[r], saved["x"][r], saved["mean1"][r], saved["rstd1"][r], P.norm1_w, d_x1_c,
                                            acc, accumulate=c > 0, last=c == nc - 1))
        ev = side.mark()
        g["out_w"] = ops.gemm(saved["ctx"], _rows_cat(d_ao), trans_a=True, split_k=True)
        del d_ao
        g["qkv_b"] = side.colsum(ev, d_qkv, torch.empty(d_qkv.shape[1], dtype=dt, device=dev))
        g["qkv_w"] = ops.gemm(saved["ln1"], d_qkv, trans_a=True, split_k=True)
        del d_qkv
        tot = None
        for c, r in enumerate(rows):
            t_ = group.bwd_ln_finish(pend2[c], dx[r])
            tot = t_ if t_ is not None else tot
        side.join()
    g["norm2_w"], g["norm2_b"], g["norm1_w"], g["norm1_b"] = tot[0], tot[1], tot[3], tot[4]
    g["out_b"], g["lin2_b"] = tot[2].to(P.out_b.dtype), tot[5].to(P.lin2_b.dtype)
    return dx.view(b, s, h), g


def decoder_layer_fwd(x: torch.Tensor, P: LayerParams, M: LayerMeta, group):
    """x: bf16 [b, s, h] -> (out bf16 [b, s, h], saved activations for backward)."""
    b, s, h = x.shape
    T = b * s
    if _row_owner_path(T, h, M, group):
        return decoder_layer_fwd_rows(x, P, M, group)
    if (_PIPELINE_FWD and getattr(group, "supports_chunks", False) and group.nranks > 1 and M.normalize_before and b % 2 == 0
            and group.partial_buffer(T, h, x.device).data_ptr() == group.partial_buffer(1, h, x.device).data_ptr()
            and T * h <= group.max_elems):
        return _fwd_pipelined(x, P, M, group)
    x2 = x.reshape(T, h)
    scale = M.hd ** -0.5

    if M.normalize_before:
        ln1, mean1, rstd1 = ops.layernorm_fwd(x2, P.norm1_w, P.norm1_b, M.eps)      # :534
        attn_in = ln1
    else:
        ln1 = mean1 = rstd1 = None
        attn_in = x2
    qkv = ops.gemm(attn_in, P.qkv_w, epilogue=FX_EPI_BIAS, bias=P.qkv_b)            # :166
    mask_pre, mask_ev = _attn_mask_beside(b, s, M, x.device)                         # keep bits: beside the projection
    if mask_ev is not None:
        torch.cuda.current_stream().wait_event(mask_ev)
    ctx, lse, attn_mask = ops.attn_causal_fwd(qkv.view(b, s, -1), M.nh_local, M.hd, scale, M.p_attn, *M.rng_attn,
                                              **({} if mask_pre is None else dict(mask=mask_pre, mask_ready=True)))  # :167-173, :268-298
    ctx2 = ctx.view(T, -1)
    attn_out = ops.gemm(ctx2, P.out_w, out=group.partial_buffer(T, h, x.device))    # :340 local matmul
    # _mp_allreduce, + bias, dropout1 + residual (:546-547) — one fused kernel on a symmetric-memory group
    r1 = group.all_reduce_bias_dropout_residual(attn_out, P.out_b, x2, M.p_hidden1, *M.rng_h1, mask_in=M.mask_h1)
    del attn_out
    if M.normalize_before:
        x1 = r1
        ln2, mean2, rstd2 = ops.layernorm_fwd(x1, P.norm2_w, P.norm2_b, M.eps)      # :554
        ffn_in = ln2
    else:
        x1, mean1, rstd1 = ops.layernorm_fwd(r1, P.norm1_w, P.norm1_b, M.eps)       # :550
        ln2 = mean2 = rstd2 = None
        ffn_in = x1
    z = torch.empty(T, P.lin1_w.shape[1], dtype=x.dtype, device=x.device)  # bf16: the kernels accept nothing else
    act = ops.gemm(ffn_in, P.lin1_w, epilogue=FX_EPI_BIAS_GELU, bias=P.lin1_b, aux=z)  # :558-559
    y2 = ops.gemm(act, P.lin2_w, out=group.partial_buffer(T, h, x.device))          # :558 local matmul
    # _mp_allreduce, + bias, dropout2 + residual (:556-561)
    r2 = group.all_reduce_bias_dropout_residual(y2, P.lin2_b, x1, M.p_hidden2, *M.rng_h2, mask_in=M.mask_h2)
    del y2
    if M.normalize_before:
        out = r2
        mean3 = rstd3 = None
    else:
        out, mean3, rstd3 = ops.layernorm_fwd(r2, P.norm2_w, P.norm2_b, M.eps)      # :564
    saved = dict(x=x2, ln1=ln1, mean1=mean1, rstd1=rstd1, qkv=qkv, ctx=ctx2, lse=lse, attn_mask=attn_mask, r1=r1, x1=x1,
                 ln2=ln2, mean2=mean2, rstd2=rstd2, z=z, act=act, r2=r2, mean3=mean3, rstd3=rstd3,
                 shape=(b, s, h))
    return out.view(b, s, h), saved


def decoder_layer_bwd(dout: torch.Tensor, saved: dict, P: LayerParams, M: LayerMeta, group):
    """dout bf16 [b, s, h] -> (dx, dict of per-rank parameter gradients, keyed like LayerParams)."""
    if saved.get("rows"):
        return decoder_layer_bwd_rows(dout, saved, P, M, group)
    b, s, h = saved["shape"]
    T = b * s
    d = dout.reshape(T, h)
    scale = M.hd ** -0.5
    g = {}

    if not M.normalize_before:
        # out = LN2(r2)
        d, g["norm2_w"], g["norm2_b"] = ops.layernorm_bwd(d, saved["r2"], P.norm2_w, saved["mean3"], saved["rstd3"])
    # r2 = x1 + dropout2(y2 + b2): d is d(r2); the skip-path share d(x1) += d is folded in later
    d_y2 = ops.bias_dropout_residual_bwd(d, M.p_hidden2, *M.rng_h2, mask_in=M.mask_h2)
    d_z = ops.gemm(d_y2, P.lin2_w, trans_b=True, epilogue=FX_EPI_DGELU, aux=saved["z"])  # (dY · W2^T) ∘ gelu'(z)
    d_ffn_in = ops.gemm(d_z, P.lin1_w, trans_b=True, out=group.partial_buffer(T, h, d.device))
    # backward of _c_identity: allreduce of dX. Nothing on the weight-gradient side depends on it, so the
    # transfer runs on the group's own stream underneath those GEMMs.
    pending = group.all_reduce_start(d_ffn_in)
    side = _Beside(d.device)
    ev = side.mark()
    g["lin2_w"] = ops.gemm(saved["act"], d_y2, trans_a=True, split_k=True)                       # dW2 = act^T · dY
    g["lin2_b"] = side.colsum(ev, d_y2, torch.empty(h, dtype=d.dtype, device=d.device))
    del d_y2
    ffn_in = saved["ln2"] if M.normalize_before else saved["x1"]
    g["lin1_w"] = ops.gemm(ffn_in, d_z, trans_a=True, split_k=True)
    g["lin1_b"] = side.colsum(ev, d_z, torch.empty(d_z.shape[1], dtype=d.dtype, device=d.device))
    del d_z
    d_ffn_in = group.all_reduce_finish(pending)
    if M.normalize_before:
        # x1 feeds LN2 and the skip: d(x1) = LN2'(d_ffn_in) + d
        d_x1, g["norm2_w"], g["norm2_b"] = ops.layernorm_bwd(d_ffn_in, saved["x1"], P.norm2_w, saved["mean2"],
                                                             saved["rstd2"], dresidual=d)
        d_r1 = d_x1
    else:
        # x1 = LN1(r1) feeds linear1 and the skip: d(x1) = d_ffn_in + d (an add = dropout-free residual kernel)
        d_x1 = ops.bias_dropout_residual_fwd(d_ffn_in, None, d, 0.0, 0, 0)
        d_r1, g["norm1_w"], g["norm1_b"] = ops.layernorm_bwd(d_x1, saved["r1"], P.norm1_w, saved["mean1"], saved["rstd1"])
    del d_ffn_in
    # r1 = x + dropout1(attn_out + bo)
    d_ao = ops.bias_dropout_residual_bwd(d_r1, M.p_hidden1, *M.rng_h1, mask_in=M.mask_h1)
    d_ctx = ops.gemm(d_ao, P.out_w, trans_b=True)
    d_qkv = ops.attn_causal_bwd(d_ctx.view(b, s, -1), saved["qkv"].view(b, s, -1), saved["ctx"].view(b, s, -1),
                                saved["lse"], saved["attn_mask"], M.nh_local, M.hd, scale, M.p_attn).view(T, -1)
    del d_ctx
    d_attn_in = ops.gemm(d_qkv, P.qkv_w, trans_b=True, out=group.partial_buffer(T, h, d.device))
    pending = group.all_reduce_start(d_attn_in)                                     # backward of _c_identity
    ev = side.mark()
    g["out_w"] = ops.gemm(saved["ctx"], d_ao, trans_a=True, split_k=True)
    g["out_b"] = side.colsum(ev, d_ao, torch.empty(h, dtype=d.dtype, device=d.device))
    del d_ao
    attn_in = saved["ln1"] if M.normalize_before else saved["x"]
    g["qkv_w"] = ops.gemm(attn_in, d_qkv, trans_a=True, split_k=True)
    g["qkv_b"] = side.colsum(ev, d_qkv, torch.empty(d_qkv.shape[1], dtype=d.dtype, device=d.device))
    del d_qkv
    d_attn_in = group.all_reduce_finish(pending)
    side.join()
    if M.normalize_before:
        dx, g["norm1_w"], g["norm1_b"] = ops.layernorm_bwd(d_attn_in, saved["x"], P.norm1_w, saved["mean1"],
                                                           saved["rstd1"], dresidual=d_r1)
    else:
        dx = ops.bias_dropout_residual_fwd(d_attn_in, None, d_r1, 0.0, 0, 0)
    return dx.view(b, s, h), g


RECOMPUTE_GRANULARITIES = ("full", "full_attn", "core_attn")


def prune_saved(saved: dict, M: LayerMeta) -> dict:
    """Drop what `M.recompute` says backward will rebuild. Dropout needs no RNG state to be stashed: every
    mask is a pure function of the (seed, offset) pairs in M (the reference restores the tracker state
    instead, hybrid_model.py:406-408 via paddle recompute's preserve_rng_state)."""
    if M.recompute is None:
        return saved
    if M.recompute == "full":        # the whole layer is re-run: keep the input only
        return dict(x=saved["x"], shape=saved["shape"])
    keep = dict(saved)
    for k in ("ctx", "lse", "attn_mask"):   # 'core_attn': scores / softmax / PV are re-run from q, k, v
        keep[k] = None
    if M.recompute == "full_attn":   # the attention block is re-run from its input: qkv goes too
        keep["qkv"] = None
    return keep


def rebuild_saved(saved: dict, P: LayerParams, M: LayerMeta, group) -> dict:
    """Inverse of prune_saved, run at the start of backward: same kernels, same (seed, offset) pairs, so
    the rebuilt activations are the forward's bit for bit."""
    if M.recompute is None:
        return saved
    b, s, h = saved["shape"]
    if M.recompute == "full":
        return decoder_layer_fwd(saved["x"].view(b, s, h), P, M, group)[1]
    full = dict(saved)
    if full["qkv"] is None:
        attn_in = full["ln1"] if M.normalize_before else full["x"]
        full["qkv"] = ops.gemm(attn_in, P.qkv_w, epilogue=FX_EPI_BIAS, bias=P.qkv_b)
    ctx, full["lse"], full["attn_mask"] = ops.attn_causal_fwd(full["qkv"].view(b, s, -1), M.nh_local, M.hd,
                                                               M.hd ** -0.5, M.p_attn, *M.rng_attn)
    full["ctx"] = ctx.view(b * s, -1)
    return full


def saved_bytes(saved: dict) -> int:
    """Bytes of activation memory a `saved` dict pins between forward and backward (distinct buffers)."""
    seen, n = set(), 0
    for v in saved.values():
        if isinstance(v, torch.Tensor) and v.data_ptr() not in seen:
            seen.add(v.data_ptr())
            n += v.numel() * v.element_size()
    return n


# ---------------------------------------------------------------------------- sequence parallel
def decoder_layer_fwd_sp(x_shard: torch.Tensor, P: LayerParams, M: LayerMeta, group):
    """sequence_parallel=True (hybrid_model.py:475-520 with Column/RowSequenceParallelLinear,
    sequence_parallel_utils.py:150-326): x_shard is THIS rank's [s/n, b, h] slice of the sequence-first
    activations; LayerNorm, dropout and the residual adds run on the shard only, an all-gather stands in
    front of each column-parallel matmul and a reduce-scatter (fused with bias + dropout + residual)
    behind each row-parallel one. Pre-LN (:533-534, :553-554) and post-LN (:549-550, :563-564) orders.
    Returns ([s/n, b, h], saved)."""
    sl, b, h = x_shard.shape
    n = group.nranks
    s = sl * n
    T, Tl = b * s, sl * b
    dev = x_shard.device
    xs = x_shard.reshape(Tl, h)
    scale = M.hd ** -0.5
    pre = M.normalize_before

    if pre:
        attn_in, mean1, rstd1 = ops.layernorm_fwd(xs, P.norm1_w, P.norm1_b, M.eps)   # :534 on the shard
    else:
        attn_in, mean1, rstd1 = xs, None, None
    a_full = group.sp_all_gather(attn_in, b, s)                                      # AllGatherOp, utils:225
    qkv = ops.gemm(a_full, P.qkv_w, epilogue=FX_EPI_BIAS, bias=P.qkv_b)
    # the gathered tensor is what the weight gradient needs in backward (the reference keeps it too, as the
    # input of F.linear); it lives in the symmetric buffer, which the next GEMM overwrites: keep a copy
    a_keep = group.sp_keep(a_full)
    ctx, lse, attn_mask = ops.attn_causal_fwd(qkv.view(b, s, -1), M.nh_local, M.hd, scale, M.p_attn, *M.rng_attn)
    ctx2 = ctx.view(T, -1)
    part = ops.gemm(ctx2, P.out_w, out=group.partial_buffer(T, h, dev))
    # ReduceScatterOp (utils:319) + bias (:322) + dropout1 + residual (:546-547), one kernel
    r1 = group.sp_reduce_scatter_bdr(part, b, s, P.out_b, xs, M.p_hidden1, *M.rng_h1, mask_in=M.mask_h1)
    if pre:
        x1 = r1
        ffn_in, mean2, rstd2 = ops.layernorm_fwd(x1, P.norm2_w, P.norm2_b, M.eps)    # :554
    else:
        x1, mean1, rstd1 = ops.layernorm_fwd(r1, P.norm1_w, P.norm1_b, M.eps)        # :550
        ffn_in, mean2, rstd2 = x1, None, None
    f_full = group.sp_all_gather(ffn_in, b, s)
    z = torch.empty(T, P.lin1_w.shape[1], dtype=x_shard.dtype, device=dev)
    act = ops.gemm(f_full, P.lin1_w, epilogue=FX_EPI_BIAS_GELU, bias=P.lin1_b, aux=z)
    f_keep = group.sp_keep(f_full)
    part = ops.gemm(act, P.lin2_w, out=group.partial_buffer(T, h, dev))
    r2 = group.sp_reduce_scatter_bdr(part, b, s, P.lin2_b, x1, M.p_hidden2, *M.rng_h2, mask_in=M.mask_h2)
    if pre:
        out, mean3, rstd3 = r2, None, None
    else:
        out, mean3, rstd3 = ops.layernorm_fwd(r2, P.norm2_w, P.norm2_b, M.eps)       # :564
    saved = dict(x=xs, a_full=a_keep, mean1=mean1, rstd1=rstd1, qkv=qkv, ctx=ctx2, lse=lse, attn_mask=attn_mask, r1=r1,
                 x1=x1, f_full=f_keep, mean2=mean2, rstd2=rstd2, z=z, act=act, r2=r2, mean3=mean3, rstd3=rstd3,
                 shape=(b, s, h), sp=True)
    return out.view(sl, b, h), saved


def prune_saved_sp(saved: dict, M: LayerMeta) -> dict:
    """prune_saved for the sequence-parallel layer (hybrid_model.py:332-335, 406-408, 537-539 together with
    :475-520): 'full' keeps this rank's input shard only."""
    if M.recompute is None:
        return saved
    if M.recompute == "full":
        return dict(x=saved["x"], shape=saved["shape"], sp=True)
    keep = dict(saved)
    for k in ("ctx", "lse", "attn_mask"):
        keep[k] = None
    if M.recompute == "full_attn":
        keep["qkv"] = None
    return keep


def rebuild_saved_sp(saved: dict, P: LayerParams, M: LayerMeta, group) -> dict:
    if M.recompute is None:
        return saved
    b, s, h = saved["shape"]
    if M.recompute == "full":
        return decoder_layer_fwd_sp(saved["x"].view(s // group.nranks, b, h), P, M, group)[1]
    full = dict(saved)
    if full["qkv"] is None:  # from the gathered input of qkv_proj, which backward keeps anyway (weight gradient)
        full["qkv"] = ops.gemm(full["a_full"], P.qkv_w, epilogue=FX_EPI_BIAS, bias=P.qkv_b)
    ctx, full["lse"], full["attn_mask"] = ops.attn_causal_fwd(full["qkv"].view(b, s, -1), M.nh_local, M.hd,
                                                               M.hd ** -0.5, M.p_attn, *M.rng_attn)
    full["ctx"] = ctx.view(b * s, -1)
    return full


def decoder_layer_bwd_sp(dout_shard: torch.Tensor, saved: dict, P: LayerParams, M: LayerMeta, group):
    """Adjoint of decoder_layer_fwd_sp: every all-gather becomes a reduce-scatter and back
    (sequence_parallel_utils.py:99-126); the two reduce-scatters run on the group's stream underneath the
    weight-gradient GEMMs of the column-parallel linears; gradients of parameters that saw only this rank's rows (LayerNorm weight / bias,
    out_proj / linear2 bias) are summed over the group (all_reduce_gradient_hook, utils:136-140, :297-298,
    hybrid_model.py:515-520)."""
    b, s, h = saved["shape"]
    n = group.nranks
    T, Tl = b * s, (s // n) * b
    dev = dout_shard.device
    d = dout_shard.reshape(Tl, h)
    scale = M.hd ** -0.5
    pre = M.normalize_before
    g = {}

    if not pre:  # out = LN2(r2)
        d, norm2_w, norm2_b = ops.layernorm_bwd(d, saved["r2"], P.norm2_w, saved["mean3"], saved["rstd3"])
    d_y2s = ops.bias_dropout_residual_bwd(d, M.p_hidden2, *M.rng_h2, mask_in=M.mask_h2)
    lin2_b = ops.colsum(d_y2s, out_dtype=torch.float32)
    d_y2 = group.sp_all_gather(d_y2s, b, s)                                          # backward of ReduceScatterOp
    d_z = ops.gemm(d_y2, P.lin2_w, trans_b=True, epilogue=FX_EPI_DGELU, aux=saved["z"])
    g["lin2_w"] = ops.gemm(saved["act"], d_y2, trans_a=True, split_k=True)
    d_ffn_part = ops.gemm(d_z, P.lin1_w, trans_b=True, out=group.partial_buffer(T, h, dev))
    # backward of AllGatherOp; nothing in linear1's weight gradient depends on it, so the transfer runs on the
    # group's stream underneath that GEMM
    pending = group.sp_reduce_scatter_start(d_ffn_part, b, s)
    g["lin1_b"] = ops.colsum(d_z)
    g["lin1_w"] = ops.gemm(saved["f_full"], d_z, trans_a=True, split_k=True)
    del d_z
    d_ffn_in = group.sp_reduce_scatter_finish(pending)
    if pre:  # x1 feeds LN2 and the skip: d(x1) = LN2'(d_ffn_in) + d
        d_r1, norm2_w, norm2_b = ops.layernorm_bwd(d_ffn_in, saved["x1"], P.norm2_w, saved["mean2"], saved["rstd2"],
                                                   dresidual=d)
    else:    # x1 = LN1(r1) feeds linear1 and the skip: d(x1) = d_ffn_in + d, then through LN1
        d_x1 = ops.bias_dropout_residual_fwd(d_ffn_in, None, d, 0.0, 0, 0)
        d_r1, norm1_w, norm1_b = ops.layernorm_bwd(d_x1, saved["r1"], P.norm1_w, saved["mean1"], saved["rstd1"])
    d_aos = ops.bias_dropout_residual_bwd(d_r1, M.p_hidden1, *M.rng_h1, mask_in=M.mask_h1)
    out_b = ops.colsum(d_aos, out_dtype=torch.float32)
    d_ao = group.sp_all_gather(d_aos, b, s)
    d_ctx = ops.gemm(d_ao, P.out_w, trans_b=True)
    g["out_w"] = ops.gemm(saved["ctx"], d_ao, trans_a=True, split_k=True)
    d_qkv = ops.attn_causal_bwd(d_ctx.view(b, s, -1), saved["qkv"].view(b, s, -1), saved["ctx"].view(b, s, -1),
                                saved["lse"], saved["attn_mask"], M.nh_local, M.hd, scale, M.p_attn).view(T, -1)
    del d_ctx
    d_in_part = ops.gemm(d_qkv, P.qkv_w, trans_b=True, out=group.partial_buffer(T, h, dev))
    pending = group.sp_reduce_scatter_start(d_in_part, b, s)
    g["qkv_b"] = ops.colsum(d_qkv)
    g["qkv_w"] = ops.gemm(saved["a_full"], d_qkv, trans_a=True, split_k=True)
    del d_qkv
    d_in = group.sp_reduce_scatter_finish(pending)
    if pre:
        dxs, norm1_w, norm1_b = ops.layernorm_bwd(d_in, saved["x"], P.norm1_w, saved["mean1"], saved["rstd1"],
                                                  dresidual=d_r1)
    else:
        dxs = ops.bias_dropout_residual_fwd(d_in, None, d_r1, 0.0, 0, 0)
    # one small fp32 allreduce for the six shard-local parameter gradients
    parts = [norm1_w, norm1_b, norm2_w, norm2_b, out_b, lin2_b]
    flat = group.sp_sum_small(parts)
    o = 0
    for name, t in zip(("norm1_w", "norm1_b", "norm2_w", "norm2_b", "out_b", "lin2_b"), parts):
        v = flat[o:o + t.numel()].view(t.shape)
        g[name] = v if name.startswith("norm") else v.to(P.out_b.dtype)
        o += t.numel()
    return dxs.view(s // n, b, h), g


_GRAD_ORDER = ("norm1_w", "norm1_b", "qkv_w", "qkv_b", "out_w", "out_b",
               "norm2_w", "norm2_b", "lin1_w", "lin1_b", "lin2_w", "lin2_b")


class DecoderLayerFunction(torch.autograd.Function):
    """One autograd node for the whole layer: no framework ops between the kernels."""

    @staticmethod
    def forward(ctx, x, meta, group, *params):
        P = LayerParams(*params)
        if meta.recompute is not None and meta.recompute not in RECOMPUTE_GRANULARITIES:
            raise ValueError(f"recompute_granularity must be one of {RECOMPUTE_GRANULARITIES}, got {meta.recompute!r}")
        if meta.sequence_parallel:
            out, saved = decoder_layer_fwd_sp(x, P, meta, group)
            ctx.saved, ctx.P, ctx.meta, ctx.group = prune_saved_sp(saved, meta), P, meta, group
            return out
        out, saved = decoder_layer_fwd(x, P, meta, group)
        ctx.saved, ctx.P, ctx.meta, ctx.group = prune_saved(saved, meta), P, meta, group
        return out

    @staticmethod
    def backward(ctx, dout):
        if ctx.meta.sequence_parallel:
            saved = rebuild_saved_sp(ctx.saved, ctx.P, ctx.meta, ctx.group)
            dx, g = decoder_layer_bwd_sp(dout.contiguous(), saved, ctx.P, ctx.meta, ctx.group)
            ctx.saved = None
            return (dx, None, None) + tuple(g[k] for k in _GRAD_ORDER)
        saved = rebuild_saved(ctx.saved, ctx.P, ctx.meta, ctx.group)
        dx, g = decoder_layer_bwd(dout.contiguous(), saved, ctx.P, ctx.meta, ctx.group)
        ctx.saved = None
        return (dx, None, None) + tuple(g[k] for k in _GRAD_ORDER)


# ---------------------------------------------------------------------------- standalone linears
class ColumnParallelLinearFunction(torch.autograd.Function):
    """y = x·W_shard + b_shard; backward allreduces dX over the mp group (_c_identity)."""

    @staticmethod
    def forward(ctx, x, weight, bias, group):
        x2 = x.reshape(-1, x.shape[-1])
        y = ops.gemm(x2, weight, epilogue=FX_EPI_BIAS if bias is not None else FX_EPI_NONE, bias=bias)
        ctx.save_for_backward(x2, weight)
        ctx.has_bias, ctx.group, ctx.in_shape = bias is not None, group, x.shape
        return y.view(*x.shape[:-1], weight.shape[1])

    @staticmethod
    def backward(ctx, dy):
        x2, weight = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1]).contiguous()
        dw = ops.gemm(x2, dy2, trans_a=True, split_k=True)
        db = ops.colsum(dy2) if ctx.has_bias else None
        dx = ops.gemm(dy2, weight, trans_b=True)
        ctx.group.all_reduce_(dx)
        return dx.view(ctx.in_shape), dw, db, None


class RowParallelLinearFunction(torch.autograd.Function):
    """y = allreduce(x_shard·W_shard) + b; the bias is added once, after the allreduce."""

    @staticmethod
    def forward(ctx, x, weight, bias, group):
        x2 = x.reshape(-1, x.shape[-1])
        if group.nranks == 1:
            y = ops.gemm(x2, weight, epilogue=FX_EPI_BIAS if bias is not None else FX_EPI_NONE, bias=bias)
        else:
            y = ops.gemm(x2, weight)
            group.all_reduce_(y)
            if bias is not None:
                y = _add_bias(y, bias)
        ctx.save_for_backward(x2, weight)
        ctx.has_bias, ctx.in_shape = bias is not None, x.shape
        return y.view(*x.shape[:-1], weight.shape[1])

    @staticmethod
    def backward(ctx, dy):
        x2, weight = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1]).contiguous()
        dw = ops.gemm(x2, dy2, trans_a=True, split_k=True)
        db = ops.colsum(dy2) if ctx.has_bias else None
        dx = ops.gemm(dy2, weight, trans_b=True)
        return dx.view(ctx.in_shape), dw, db, None


class AddBiasFunction(torch.autograd.Function):
    """y + bias over the last dim (RowSequenceParallelLinear adds its bias after the reduce-scatter,
    sequence_parallel_utils.py:322), with the library's own kernels on both sides."""

    @staticmethod
    def forward(ctx, y, bias):
        return _add_bias(y.reshape(-1, y.shape[-1]).contiguous(), bias).view(y.shape)

    @staticmethod
    def backward(ctx, dy):
        return dy, ops.colsum(dy.reshape(-1, dy.shape[-1]).contiguous())


def _add_bias(y: torch.Tensor, bias: torch.Tensor) -> torch.Tensor:
    """y + bias after the allreduce, through the residual kernel with p = 0 and a zero residual
    (cached per shape, so no memset on the hot path)."""
    return ops.bias_dropout_residual_fwd(y, bias, _zeros_like_cached(y), 0.0, 0, 0)


class DropoutFunction(torch.autograd.Function):
    """nn.Dropout(p, mode="upscale_in_train") on its own (the embedding dropout of the sequence-parallel model,
    hybrid_model.py:618-619), through the residual kernel with a zero residual; backward regenerates the mask."""

    @staticmethod
    def forward(ctx, x, p, seed, offset):
        ctx.cfg = (float(p), int(seed), int(offset))
        x2 = x.reshape(-1, x.shape[-1]).contiguous()
        return ops.bias_dropout_residual_fwd(x2, None, _zeros_like_cached(x2), p, seed, offset).view(x.shape)

    @staticmethod
    def backward(ctx, dout):
        p, seed, offset = ctx.cfg
        d2 = dout.reshape(-1, dout.shape[-1]).contiguous()
        return ops.bias_dropout_residual_bwd(d2, p, seed, offset).view(dout.shape), None, None, None


_zero_cache: dict = {}


def _zeros_like_cached(t: torch.Tensor) -> torch.Tensor:
    key = (t.device, tuple(t.shape))
    z = _zero_cache.get(key)
    if z is None:
        z = torch.zeros_like(t)
        _zero_cache[key] = z
    return z


class CausalSelfAttentionFunction(torch.autograd.Function):
    """core_attn on the packed qkv_proj output (hybrid_model.py:165-186, 268-298)."""

    @staticmethod
    def forward(ctx, qkv, nh_local, hd, p_drop, seed, offset, recompute=False):
        out, lse, mask = ops.attn_causal_fwd(qkv, nh_local, hd, hd ** -0.5, p_drop, seed, offset)
        if recompute:  # 'core_attn' recompute (hybrid_model.py:332-335): keep q, k, v only
            ctx.save_for_backward(qkv)
            ctx.mask = None
        else:
            ctx.save_for_backward(qkv, out, lse)
            ctx.mask = mask
        ctx.cfg = (nh_local, hd, p_drop, seed, offset, recompute)
        return out

    @staticmethod
    def backward(ctx, dout):
        nh_local, hd, p_drop, seed, offset, recompute = ctx.cfg
        if recompute:
            (qkv,) = ctx.saved_tensors
            out, lse, mask = ops.attn_causal_fwd(qkv, nh_local, hd, hd ** -0.5, p_drop, seed, offset)
        else:
            qkv, out, lse = ctx.saved_tensors
            mask = ctx.mask
        dqkv = ops.attn_causal_bwd(dout.contiguous(), qkv, out, lse, mask, nh_local, hd, hd ** -0.5, p_drop)
        return dqkv, None, None, None, None, None, None
