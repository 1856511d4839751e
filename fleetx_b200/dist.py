"""Model-parallel (mp) group plumbing for the tensor-parallel linears.

The reference gets its mp group from fleet.get_hybrid_communicate_group() and issues the
collectives through Paddle's ProcessGroupNCCL: `_mp_allreduce` after RowParallelLinear's local
matmul, and the allreduce that is the backward of `_c_identity` in front of ColumnParallelLinear
(fleet.meta_parallel; mirrored in-tree at sequence_parallel_utils.py:150-326). Here the group is a
small object with `rank`, `nranks` and `all_reduce_(tensor)`:

  SingleGroup      mp_degree 1, all_reduce_ is a no-op
  TorchDistGroup   one process per GPU, torch.distributed (NCCL over NVLink on GPUs; gloo in the
                   CPU tests) — the plumbing backend

Sharding helpers mirror what each mp rank holds.
"""
from __future__ import annotations

from typing import Optional

import torch


class _GroupOps:
    """What the layer needs from a model-parallel group beyond all_reduce_(): where a row-parallel
    GEMM should write its partial result, and the allreduce fused with what follows it. The base
    versions compose all_reduce_() with the library's elementwise kernels."""

    def partial_buffer(self, rows: int, cols: int, device) -> torch.Tensor:
        return torch.empty(rows, cols, dtype=torch.bfloat16, device=device)

    def all_reduce(self, t: torch.Tensor) -> torch.Tensor:
        return self.all_reduce_(t)

    # Split form, so the caller can put independent work (the weight-gradient GEMMs) between the two
    # halves and have it overlap the transfer. The base version is simply synchronous.
    def all_reduce_start(self, t: torch.Tensor):
        return self.all_reduce(t)

    def all_reduce_finish(self, handle) -> torch.Tensor:
        return handle

    def all_reduce_bias_dropout_residual(self, y, bias, residual, p, seed, offset, mask_in=None):
        from . import ops

        self.all_reduce_(y)
        return ops.bias_dropout_residual_fwd(y, bias, residual, p, seed, offset, mask_in=mask_in)

    # ---- sequence parallelism (sequence_parallel_utils.py:36-126). Shards are sequence-first [s/n, b, h]
    # flattened to [s/n * b, h]; full tensors are batch-first [b * s, h] (what the GEMM / attention kernels
    # take). The base versions go through torch.distributed and torch permutes (the plumbing backend).
    def sp_all_gather(self, shard: torch.Tensor, b: int, s: int) -> torch.Tensor:
        """[s/n * b, h] shard of this rank -> full [b * s, h], batch-first (all_gather, :47-55)."""
        n, h = self.nranks, shard.shape[-1]
        if n == 1:
            return shard.view(s, b, h).transpose(0, 1).reshape(b * s, h)
        full = torch.empty(s, b, h, dtype=shard.dtype, device=shard.device)  # rank-major rows = sequence-first
        self._dist.all_gather_into_tensor(full, shard.contiguous().view(s // n, b, h), group=self.pg)
        return full.transpose(0, 1).reshape(b * s, h)

    def sp_reduce_scatter(self, part: torch.Tensor, b: int, s: int) -> torch.Tensor:
        """partial sums [b * s, h], batch-first -> this rank's rows of the sum, [s/n * b, h] (:58-70)."""
        n, h = self.nranks, part.shape[-1]
        seq_first = part.view(b, s, h).transpose(0, 1).contiguous()  # [s, b, h]
        if n == 1:
            return seq_first.view(s * b, h)
        if self._dist.get_backend(self.pg) == "nccl":
            out = torch.empty(s // n, b, h, dtype=part.dtype, device=part.device)
            self._dist.reduce_scatter_tensor(out, seq_first, op=self._dist.ReduceOp.SUM, group=self.pg)
            return out.view(-1, h)
        self._dist.all_reduce(seq_first, op=self._dist.ReduceOp.SUM, group=self.pg)  # gloo has no reduce_scatter
        sl = s // n
        return seq_first[self.rank * sl:(self.rank + 1) * sl].reshape(-1, h)

    def sp_reduce_scatter_bdr(self, part, b, s, bias, residual, p, seed, offset, mask_in=None):
        """reduce-scatter, then residual + dropout(sum + bias) on the local shard (dropout indexed locally)."""
        from . import ops

        y = self.sp_reduce_scatter(part, b, s)
        if residual is None:
            return y
        return ops.bias_dropout_residual_fwd(y, bias, residual, p, seed, offset, mask_in=mask_in)

    def sp_keep(self, t: torch.Tensor) -> torch.Tensor:
        """A gathered tensor that must outlive the next collective (it may sit in a communication buffer)."""
        return t

    def sp_sum_small(self, parts) -> torch.Tensor:
        """Sum over the ranks of a few small fp32 vectors (the gradients of parameters whose kernels saw only this
        rank's rows: all_reduce_gradient_hook, sequence_parallel_utils.py:136-140), returned as one flat tensor."""
        flat = torch.cat([t.reshape(-1) for t in parts])
        return self.all_reduce_(flat)

    # Split form of the plain reduce-scatter (backward of AllGatherOp): the caller puts the weight-gradient
    # GEMM that does not depend on it between the two halves. The base version is synchronous.
    def sp_reduce_scatter_start(self, part, b, s):
        return self.sp_reduce_scatter(part, b, s)

    def sp_reduce_scatter_finish(self, handle):
        return handle



    # ---- row-owner form of the allreduce (csrc/comm_rows.cu): rank r owns rows [r0, r1) of the [rows, h] tensor of a
    # call and does the row-wise work that follows the allreduce on them only; the tensor every rank needs next comes
    # back complete. Every operation is split into start / finish so that the caller can put independent work (the
    # other batch chunk, the weight-gradient GEMMs) in between. The base versions state the same contract with
    # all_reduce_() and the library's row-wise kernels (the plumbing backend; what the CPU tests of the host
    # sequencing run over gloo). `part` may be a row slice [rows, h] of a larger [T, h] tensor (batch chunks);
    # mask_in / residual / dres are then the matching slices and drop_off is the element offset of the slice inside
    # [T, h] (the dropout mask is indexed by position in the whole tensor).
    row_owner = False     # whether decoder_layer_fwd / _bwd take the row-owner path on this group
    pipeline_chunks = 1   # batch chunks that path pipelines through the collectives

    def own_rows(self, rows: int):
        per = (rows + self.nranks - 1) // self.nranks
        r0 = min(self.rank * per, rows)
        return r0, min(r0 + per, rows)

    def acc_buffer(self, h: int, device, dtype=torch.float32) -> torch.Tensor:
        """[6, h] (fp32 on the GPU): partial sums over the own rows of dgamma2 | dbeta2 | d(out_proj.bias) | dgamma1 |
        dbeta1 | d(linear2.bias), summed over the ranks by all_reduce_acc() at the end of backward."""
        return torch.zeros(6, h, dtype=dtype, device=device)

    def all_reduce_acc(self, acc: torch.Tensor) -> torch.Tensor:
        return self.all_reduce_(acc.clone())

    def free_sms(self) -> int:
        """SMs left to the compute kernels while one of this group's collectives runs on another stream (0: all)."""
        return 0

    def keep_copy(self, t: torch.Tensor, dst: torch.Tensor) -> None:
        """dst <- t, where t came back from a fused call and may be a view of a communication buffer that the next
        call rewrites (dst must be waited for with wait_kept before it is read)."""
        dst.copy_(t)

    def wait_kept(self, t: torch.Tensor) -> torch.Tensor:
        return t

    def _no_generated_mask_in_chunks(self, p, mask_in, drop_off):
        if p > 0 and mask_in is None and drop_off:
            raise NotImplementedError("the plumbing statement of the row-owner collectives draws dropout masks from "
                                      "element 0: batch chunks need injected masks (or pipeline_chunks = 1)")

    def _gather_rows(self, own: torch.Tensor, rows: int) -> torch.Tensor:
        n, h = self.nranks, own.shape[-1]
        per = (rows + n - 1) // n
        if per * n != rows:
            raise ValueError(f"row-owner path needs rows ({rows}) divisible by the mp degree ({n})")
        full = torch.empty(rows, h, dtype=own.dtype, device=own.device)
        self._dist.all_gather_into_tensor(full, own.contiguous(), group=self.pg)
        return full

    def fwd_ln_start(self, part, bias, residual, p, seed, offset, mask_in, gamma, beta, eps, drop_off=0):
        """allreduce(part) + bias -> dropout -> + residual = x1; LayerNorm(x1).
        finish -> (x1_own, mean_own, rstd_own, ln_full)."""
        from . import ops

        self._no_generated_mask_in_chunks(p, mask_in, drop_off)
        r0, r1 = self.own_rows(part.shape[0])
        x1 = self.all_reduce_bias_dropout_residual(part, bias, residual, p, seed, offset, mask_in=mask_in)
        ln, mean, rstd = ops.layernorm_fwd(x1, gamma, beta, eps)
        return x1[r0:r1].contiguous(), mean[r0:r1].contiguous(), rstd[r0:r1].contiguous(), ln

    def fwd_ln_finish(self, handle):
        return handle

    def fwd_out_start(self, part, bias, residual_own, p, seed, offset, mask_in, drop_off=0):
        """allreduce(part) + bias -> dropout -> + residual (own rows given). finish(handle, out) writes the full
        result into `out` (a [rows, h] destination owned by the caller)."""
        self._no_generated_mask_in_chunks(p, mask_in, drop_off)
        # the dropout mask is indexed by position in the whole tensor: gather the residual rows and apply the
        # row-wise kernel to the full tensor (this is the plumbing statement, not the fast path)
        residual_full = self._gather_rows(residual_own, part.shape[0])
        return self.all_reduce_bias_dropout_residual(part, bias, residual_full, p, seed, offset, mask_in=mask_in)

    def fwd_out_finish(self, handle, out):
        out.copy_(handle)

    def bwd_ln_bdr_start(self, part, x_own, mean_own, rstd_own, gamma, dres, p, seed, offset, mask_in, acc,
                         drop_off=0, accumulate=False):
        """d_x1 = LayerNorm'(allreduce(part)) + dres on the own rows; d_ao = dropout'(d_x1) gathered for every rank;
        acc[0:3] (+)= dgamma | dbeta | column sums of d_ao over the own rows. finish -> (d_x1_own, d_ao_full)."""
        from . import ops

        self._no_generated_mask_in_chunks(p, mask_in, drop_off)
        rows, h = part.shape
        r0, r1 = self.own_rows(rows)
        self.all_reduce_(part)
        d_x1, dg, db = ops.layernorm_bwd(part[r0:r1].contiguous(), x_own, gamma, mean_own, rstd_own,
                                         dresidual=dres[r0:r1].contiguous())
        d_ao = ops.bias_dropout_residual_bwd(self._gather_rows(d_x1, rows), p, seed, offset, mask_in=mask_in)
        for i, v in enumerate((dg, db, ops.colsum(d_ao[r0:r1].contiguous(), out_dtype=acc.dtype))):
            acc[i].copy_(acc[i] + v.to(acc.dtype) if accumulate else v)
        return (d_x1, d_ao)

    def bwd_ln_bdr_finish(self, handle):
        return handle

    def bwd_ln_start(self, part, x, mean, rstd, gamma, dres_own, acc, accumulate=False, last=True):
        """dx = LayerNorm'(allreduce(part)) + dres on the own rows, gathered for every rank; acc[3:5] (+)= dgamma |
        dbeta over the own rows. last: this is the final accumulation into acc of the backward pass, so the sum of
        acc over the ranks rides along. finish(handle, out) writes dx into `out`; -> summed acc or None."""
        from . import ops

        rows, h = part.shape
        r0, r1 = self.own_rows(rows)
        self.all_reduce_(part)
        dx, dg, db = ops.layernorm_bwd(part[r0:r1].contiguous(), x[r0:r1].contiguous(), gamma,
                                       mean[r0:r1].contiguous(), rstd[r0:r1].contiguous(), dresidual=dres_own)
        for i, v in ((3, dg), (4, db)):
            acc[i].copy_(acc[i] + v.to(acc.dtype) if accumulate else v)
        return (self._gather_rows(dx, rows), self.all_reduce_acc(acc) if last else None)

    def bwd_ln_finish(self, handle, out):
        out.copy_(handle[0])
        return handle[1]


class SingleGroup(_GroupOps):
    rank = 0
    nranks = 1

    def all_reduce_(self, t: torch.Tensor) -> torch.Tensor:
        return t

    def barrier(self):
        return None


class TorchDistGroup(_GroupOps):
    """Wraps a torch.distributed process group (default: WORLD)."""

    def __init__(self, process_group=None):
        import torch.distributed as dist

        if not dist.is_initialized():
            raise RuntimeError("torch.distributed is not initialised")
        self._dist = dist
        self.pg = process_group
        self.rank = dist.get_rank(process_group)
        self.nranks = dist.get_world_size(process_group)

    def all_reduce_(self, t: torch.Tensor) -> torch.Tensor:
        if self.nranks > 1:
            self._dist.all_reduce(t, op=self._dist.ReduceOp.SUM, group=self.pg)
        return t

    def barrier(self):
        self._dist.barrier(group=self.pg)

    def all_reduce_start(self, t: torch.Tensor):
        if self.nranks == 1:
            return (t, None)
        return (t, self._dist.all_reduce(t, op=self._dist.ReduceOp.SUM, group=self.pg, async_op=True))

    def all_reduce_finish(self, handle) -> torch.Tensor:
        t, work = handle
        if work is not None:
            work.wait()  # the current stream waits for the collective
        return t


class SymmetricTPGroup(TorchDistGroup):
    """Model-parallel group whose allreduces are this library's own kernel over NVLink/NVSwitch peer
    memory (csrc/comm.cu: switch-side reduce-scatter with multimem.ld_reduce, peer-load all-gather,
    bias + dropout + residual fused in). torch.distributed only does the plumbing: it allocates the
    symmetric buffer, exchanges the mappings and gives us the raw peer / multicast / flag addresses.

    One symmetric buffer of `max_elems` bf16 is reused by every allreduce of the layer: the GEMM
    writes its partial result into it and the kernel leaves the reduced tensor in ordinary memory."""

    def __init__(self, process_group=None, max_elems: int = 1 << 26, num_blocks: int = 148):
        super().__init__(process_group)
        import ctypes

        import torch.distributed._symmetric_memory as symm_mem

        from . import _lib

        dev = torch.device("cuda", torch.cuda.current_device())
        pg = self.pg if self.pg is not None else self._dist.group.WORLD
        self._buf = symm_mem.empty(max_elems, dtype=torch.bfloat16, device=dev)
        self._hdl = symm_mem.rendezvous(self._buf, pg.group_name)
        if not self._hdl.multicast_ptr:
            raise RuntimeError("NVSwitch multicast is not available for this group")
        n = self.nranks
        self._peer_bufs = (ctypes.c_uint64 * n)(*[int(p) for p in self._hdl.buffer_ptrs])
        self._peer_sigs = (ctypes.c_uint64 * n)(*[int(p) for p in self._hdl.signal_pad_ptrs])
        self._mc = int(self._hdl.multicast_ptr)
        self._lib = _lib
        self.fallbacks = {}  # name of a call that went to torch.distributed -> how often
        self._epoch = 0
        self._comm_stream = torch.cuda.Stream(device=dev, priority=-1)  # high priority: its blocks go first
        self.num_blocks = num_blocks
        self.max_elems = max_elems
        # row-owner allreduce (csrc/comm_rows.cu): two more symmetric buffers receive the gathered results (B0: LayerNorm
        # output / d(attention out), B1: layer output / dx) and a small fp32 one holds the parameter-gradient partial sums.
        # The flag pages of the first buffer serve every kernel (one epoch sequence).
        import os

        self.row_owner = os.environ.get("FX_TP_ROWS", "1") != "0"
        self.rows_blocks = int(os.environ.get("FX_TP_ROWS_BLOCKS", "32"))
        self._acc_h = min(8192, int(_lib.lib().fx_tp_rows_max_h()))
        if True:  # the buffers exist whichever order the layer takes (the small fp32 one also serves sequence parallel)
            self._bufB, self._mcB = [], []
            for _ in range(2):
                t = symm_mem.empty(max_elems, dtype=torch.bfloat16, device=dev)
                hd_ = symm_mem.rendezvous(t, pg.group_name)
                if not hd_.multicast_ptr:
                    raise RuntimeError("NVSwitch multicast is not available for this group")
                self._bufB.append(t)
                self._mcB.append(int(hd_.multicast_ptr))
                self._keep_handles = getattr(self, "_keep_handles", []) + [hd_]
            self._bufC = symm_mem.empty(6 * self._acc_h, dtype=torch.float32, device=dev)
            hc = symm_mem.rendezvous(self._bufC, pg.group_name)
            self._mcC = int(hc.multicast_ptr)
            self._keep_handles.append(hc)
            self._b0_copies = []   # events of copies still being taken out of B0
            self._kept = {}        # data_ptr of a keep_copy destination -> event its readers wait for
            self._copy_stream = torch.cuda.Stream(device=dev)
            # batch chunks pipelined through the collectives: measured slower than one pass at 2 and 8 GPUs (half-size
            # collectives take 0.7x the time of whole ones, the GEMMs beside them quantise worse), so 1 by default
            self.pipeline_chunks = int(os.environ.get("FX_TP_CHUNKS", "1"))
        # the flag page must start at zero on every rank before the first call
        pad = self._hdl.get_signal_pad(self.rank, (self._hdl.signal_pad_size // 4,), torch.int32)
        pad.zero_()
        torch.cuda.synchronize(dev)
        self._dist.barrier(group=self.pg)

    def _fallback(self, what: str) -> None:
        """A call this group cannot serve with the library's own kernels (tensor outside the symmetric buffer, too
        large, wrong dtype, h beyond the row-owner limit) goes to torch.distributed — never quietly: it is counted
        in self.fallbacks (bench.py prints it) and warned about once per kind; FX_STRICT_COMM=1 makes it an error."""
        import os
        import warnings

        if os.environ.get("FX_STRICT_COMM", "0") == "1":
            raise RuntimeError(f"SymmetricTPGroup.{what}: not servable by the library's collectives (FX_STRICT_COMM=1)")
        if what not in self.fallbacks:
            warnings.warn(f"SymmetricTPGroup.{what}: falling back to torch.distributed + separate kernels for this call "
                          "shape (tensor not in the symmetric buffer / beyond its size); it is much slower", stacklevel=3)
        self.fallbacks[what] = self.fallbacks.get(what, 0) + 1

    def partial_buffer(self, rows: int, cols: int, device) -> torch.Tensor:
        if rows * cols > self.max_elems:
            self._fallback("partial_buffer")
            return super().partial_buffer(rows, cols, device)
        return self._buf[: rows * cols].view(rows, cols)

    supports_chunks = True  # the layer may pipeline batch halves through all_reduce_bdr_start / _finish

    def sp_all_gather(self, shard: torch.Tensor, b: int, s: int) -> torch.Tensor:
        h = shard.shape[-1]
        if b * s * h > self.max_elems or shard.dtype != torch.bfloat16 or not shard.is_contiguous():
            self._fallback("sp_all_gather")
            return super().sp_all_gather(shard, b, s)
        self._epoch += 1
        self._lib.check(self._lib.lib().fx_tp_allgather_rows(
            torch.cuda.current_stream().cuda_stream, self._peer_sigs, self._mc, self.rank, self.nranks, self._epoch,
            self.num_blocks, shard.data_ptr(), b, s, h, 0), "fx_tp_allgather_rows")
        return self._buf[: b * s * h].view(b * s, h)

    def sp_reduce_scatter_bdr(self, part, b, s, bias, residual, p, seed, offset, mask_in=None):
        if not self._is_symmetric(part):
            self._fallback("sp_reduce_scatter_bdr")
            return super().sp_reduce_scatter_bdr(part, b, s, bias, residual, p, seed, offset, mask_in)
        h = part.shape[-1]
        out = torch.empty(s // self.nranks * b, h, dtype=torch.bfloat16, device=part.device)
        self._epoch += 1
        self._lib.check(self._lib.lib().fx_tp_reduce_scatter_bdr(
            torch.cuda.current_stream().cuda_stream, self._peer_sigs, self._mc, self.rank, self.nranks, self._epoch,
            self.num_blocks, out.data_ptr(), b, s, h, None if bias is None else bias.data_ptr(),
            None if residual is None else residual.data_ptr(), float(p), int(seed), int(offset),
            None if mask_in is None else mask_in.data_ptr(), self._symm_offset(part)), "fx_tp_reduce_scatter_bdr")
        return out

    def sp_reduce_scatter(self, part, b, s):
        return self.sp_reduce_scatter_bdr(part, b, s, None, None, 0.0, 0, 0)

    def sp_keep(self, t: torch.Tensor) -> torch.Tensor:
        if self.nranks == 1 or not self._is_symmetric(t):
            return t
        keep = torch.empty_like(t)
        keep.copy_(t)  # a device-to-device copy out of the symmetric buffer (the next row-parallel GEMM rewrites it)
        return keep

    def sp_sum_small(self, parts) -> torch.Tensor:
        n = sum(t.numel() for t in parts)
        if n > self._bufC.numel() or n % 4 or any(t.dtype != torch.float32 for t in parts):
            self._fallback("sp_sum_small")
            return super().sp_sum_small(parts)
        # through the library's own small allreduce: an NCCL kernel between the spinning collectives could deadlock
        o = 0
        for t in parts:
            self._bufC[o:o + t.numel()].copy_(t.reshape(-1))
            o += t.numel()
        return self.all_reduce_acc(self._bufC[:n])

    def sp_reduce_scatter_start(self, part, b, s):
        if not self._is_symmetric(part):
            self._fallback("sp_reduce_scatter_start")
            return super().sp_reduce_scatter_start(part, b, s)
        h = part.shape[-1]
        out = torch.empty(s // self.nranks * b, h, dtype=torch.bfloat16, device=part.device)  # on the caller's stream
        main = torch.cuda.current_stream()
        ready = torch.cuda.Event()
        ready.record(main)
        with torch.cuda.stream(self._comm_stream):
            self._comm_stream.wait_event(ready)
            self._epoch += 1
            self._lib.check(self._lib.lib().fx_tp_reduce_scatter_bdr(
                self._comm_stream.cuda_stream, self._peer_sigs, self._mc, self.rank, self.nranks, self._epoch,
                self.num_blocks, out.data_ptr(), b, s, h, None, None, 0.0, 0, 0, None, self._symm_offset(part)),
                "fx_tp_reduce_scatter_bdr")
            done = torch.cuda.Event()
            done.record(self._comm_stream)
        return ("symm_rs", out, done)

    def sp_reduce_scatter_finish(self, handle):
        if isinstance(handle, tuple) and len(handle) == 3 and handle[0] == "symm_rs":
            torch.cuda.current_stream().wait_event(handle[2])
            return handle[1]
        return handle

    def native_comm(self, rows: int, cols: int, device):
        """fx_tp_comm for the native whole-layer entry points (fleetx_b200.native): this group's symmetric
        buffer, flag pages and call counter as plain addresses. The returned object owns the ctypes storage."""
        import ctypes

        if rows * cols > self.max_elems:
            raise RuntimeError(f"symmetric buffer of {self.max_elems} elements is too small for [{rows}, {cols}]")

        class _Comm:
            pass

        c = _Comm()
        c.epoch = ctypes.c_uint32(self._epoch)
        c.comm = self._lib.TpComm(
            peer_bufs=ctypes.cast(self._peer_bufs, ctypes.POINTER(ctypes.c_uint64)),
            peer_signal_pads=ctypes.cast(self._peer_sigs, ctypes.POINTER(ctypes.c_uint64)),
            multicast_ptr=self._mc, local_buf=self._buf.data_ptr(), rank=self.rank, world=self.nranks,
            num_blocks=self.num_blocks, epoch=ctypes.pointer(c.epoch))
        return c

    # ---- row-owner allreduce: one kernel per site (csrc/comm_rows.cu), always launched on the group's stream
    def _rows_ok(self, part: torch.Tensor) -> bool:
        return (self.row_owner and self._is_symmetric(part) and part.shape[0] % self.nranks == 0
                and part.shape[1] <= self._acc_h)

    def acc_buffer(self, h: int, device, dtype=torch.float32) -> torch.Tensor:
        if not self.row_owner or h > self._acc_h or dtype != torch.float32:
            return super().acc_buffer(h, device, dtype)
        return self._bufC[: 6 * h].view(6, h)  # zeroed slice by slice by the kernels that accumulate into it

    def all_reduce_acc(self, acc: torch.Tensor) -> torch.Tensor:
        if acc.data_ptr() != self._bufC.data_ptr():
            self._fallback("all_reduce_acc")
            return super().all_reduce_acc(acc)
        out = torch.empty_like(acc)
        self._epoch += 1
        self._lib.check(self._lib.lib().fx_tp_allreduce_f32(
            torch.cuda.current_stream().cuda_stream, self._peer_sigs, self._mcC, self.rank, self.nranks, self._epoch,
            out.data_ptr(), acc.numel()), "fx_tp_allreduce_f32")
        return out

    def free_sms(self) -> int:
        n = torch.cuda.get_device_properties(torch.cuda.current_device()).multi_processor_count
        return max(n - self.rows_blocks, 1) if self.row_owner else 0

    def _rows_call(self, stream, mode, part, out_region, **kw):
        import ctypes

        rows, h = part.shape
        off = self._symm_offset(part)  # a row slice of partial_buffer(T, h): the result lands on the same rows
        d = self._lib.TpRowsDesc()
        d.peer_signal_pads = ctypes.cast(self._peer_sigs, ctypes.POINTER(ctypes.c_uint64))
        d.multicast_in = self._mc + off * 2
        d.multicast_out = self._mcB[out_region] + off * 2
        d.rank, d.world, d.num_blocks, d.mode = self.rank, self.nranks, self.rows_blocks, mode
        self._epoch += 1
        d.epoch = self._epoch
        d.rows, d.h = rows, h
        for k, v in kw.items():
            setattr(d, k, v.data_ptr() if isinstance(v, torch.Tensor) else v)
        self._lib.check(self._lib.lib().fx_tp_rows_fused(stream.cuda_stream, ctypes.byref(d)), "fx_tp_rows_fused")
        return self._bufB[out_region][off: off + rows * h].view(rows, h)

    def _on_comm_stream(self, fn, writes_b0=False):
        """Run fn(stream) on the group's stream behind everything the caller's stream holds so far."""
        main = torch.cuda.current_stream()
        ready = torch.cuda.Event()
        ready.record(main)
        with torch.cuda.stream(self._comm_stream):
            self._comm_stream.wait_event(ready)
            if writes_b0:  # copies still being taken out of B0 (keep_copy) go first
                for ev in self._b0_copies:
                    self._comm_stream.wait_event(ev)
                self._b0_copies = []
            res = fn(self._comm_stream)
            done = torch.cuda.Event()
            done.record(self._comm_stream)
        return ("rows", res, done)

    @staticmethod
    def _is_handle(handle) -> bool:
        return isinstance(handle, tuple) and len(handle) == 3 and isinstance(handle[0], str) and handle[0] == "rows"

    def fwd_ln_start(self, part, bias, residual, p, seed, offset, mask_in, gamma, beta, eps, drop_off=0):
        if not self._rows_ok(part):
            self._fallback("fwd_ln_start")
            return super().fwd_ln_start(part, bias, residual, p, seed, offset, mask_in, gamma, beta, eps, drop_off)
        rows, h = part.shape
        r0, r1 = self.own_rows(rows)
        x1 = torch.empty(r1 - r0, h, dtype=torch.bfloat16, device=part.device)  # allocated on the caller's stream
        mean = torch.empty(r1 - r0, dtype=torch.float32, device=part.device)
        rstd = torch.empty_like(mean)
        return self._on_comm_stream(lambda st: (x1, mean, rstd, self._rows_call(
            st, self._lib.FX_ROWS_FWD_LN, part, 0, bias=bias, residual=residual, p=float(p), seed=int(seed),
            offset=int(offset), dropout_elem_offset=int(drop_off), mask_in=mask_in, gamma=gamma, beta=beta,
            eps=float(eps), x_own=x1, mean=mean, rstd=rstd)), writes_b0=True)

    def fwd_ln_finish(self, handle):
        if not self._is_handle(handle):
            return super().fwd_ln_finish(handle)
        torch.cuda.current_stream().wait_event(handle[2])
        return handle[1]

    def keep_copy(self, t: torch.Tensor, dst: torch.Tensor) -> None:
        """The copy runs on a third stream, underneath whatever the caller launches next; the next kernel that
        rewrites B0 waits for it, readers of dst go through wait_kept()."""
        if self._symm_offset_in(t, self._bufB[0]) is None:
            return super().keep_copy(t, dst)
        ready = torch.cuda.Event()
        ready.record(torch.cuda.current_stream())
        with torch.cuda.stream(self._copy_stream):
            self._copy_stream.wait_event(ready)
            dst.copy_(t, non_blocking=True)
            done = torch.cuda.Event()
            done.record(self._copy_stream)
        # dst was allocated on the caller's stream: if it is dropped before anyone waits for the copy (a forward
        # without backward), the allocator must not hand its memory out while the copy stream still writes it
        dst.record_stream(self._copy_stream)
        self._b0_copies.append(done)
        self._kept[dst.data_ptr()] = done

    def wait_kept(self, t: torch.Tensor) -> torch.Tensor:
        lo = t.data_ptr()
        hi = lo + t.numel() * t.element_size()
        for ptr in [k for k in self._kept if lo <= k < hi]:  # every copy that landed inside t (one per batch chunk)
            torch.cuda.current_stream().wait_event(self._kept.pop(ptr))
        return t

    def fwd_out_start(self, part, bias, residual_own, p, seed, offset, mask_in, drop_off=0):
        if not self._rows_ok(part):
            self._fallback("fwd_out_start")
            return super().fwd_out_start(part, bias, residual_own, p, seed, offset, mask_in, drop_off)
        return self._on_comm_stream(lambda st: self._rows_call(
            st, self._lib.FX_ROWS_FWD_OUT, part, 1, bias=bias, residual=residual_own, p=float(p), seed=int(seed),
            offset=int(offset), dropout_elem_offset=int(drop_off), mask_in=mask_in))

    def fwd_out_finish(self, handle, out):
        if not self._is_handle(handle):
            return super().fwd_out_finish(handle, out)
        torch.cuda.current_stream().wait_event(handle[2])
        out.copy_(handle[1])  # the caller owns its result; B1 is rewritten by a later call

    def bwd_ln_bdr_start(self, part, x_own, mean_own, rstd_own, gamma, dres, p, seed, offset, mask_in, acc,
                         drop_off=0, accumulate=False):
        if not self._rows_ok(part) or acc.data_ptr() != self._bufC.data_ptr():
            self._fallback("bwd_ln_bdr_start")
            return super().bwd_ln_bdr_start(part, x_own, mean_own, rstd_own, gamma, dres, p, seed, offset, mask_in, acc,
                                            drop_off, accumulate)
        d_x1 = torch.empty_like(x_own)  # allocated on the caller's stream
        return self._on_comm_stream(lambda st: (d_x1, self._rows_call(
            st, self._lib.FX_ROWS_BWD_LN_BDR, part, 0, p=float(p), seed=int(seed), offset=int(offset),
            dropout_elem_offset=int(drop_off), mask_in=mask_in, gamma=gamma, x_own=x_own, mean=mean_own, rstd=rstd_own,
            dres=dres, dx_own=d_x1, acc=acc[0:3], acc_accumulate=int(accumulate))), writes_b0=True)

    def bwd_ln_bdr_finish(self, handle):
        if not self._is_handle(handle):
            return super().bwd_ln_bdr_finish(handle)
        torch.cuda.current_stream().wait_event(handle[2])
        return handle[1]

    def bwd_ln_start(self, part, x, mean, rstd, gamma, dres_own, acc, accumulate=False, last=True):
        if not self._rows_ok(part) or acc.data_ptr() != self._bufC.data_ptr():
            self._fallback("bwd_ln_start")
            return super().bwd_ln_start(part, x, mean, rstd, gamma, dres_own, acc, accumulate, last)

        # last: every partial sum is in once this kernel's rows are done, so the sum of acc over the ranks is its tail
        tot = torch.empty_like(acc) if last else None  # allocated on the caller's stream
        extra = dict(multicast_sum=self._mcC, sum_out=tot, sum_n=acc.numel()) if last else {}
        return self._on_comm_stream(lambda st: (self._rows_call(
            st, self._lib.FX_ROWS_BWD_LN, part, 1, gamma=gamma, x_full=x, mean=mean, rstd=rstd, dres=dres_own,
            acc=acc[3:5], acc_accumulate=int(accumulate), **extra), tot))

    def bwd_ln_finish(self, handle, out):
        if not self._is_handle(handle):
            return super().bwd_ln_finish(handle, out)
        main = torch.cuda.current_stream()
        main.wait_event(handle[2])
        dx, tot = handle[1]
        out.copy_(dx)  # the caller owns dx; B1 is rewritten by a later call
        return tot

    def _symm_offset_in(self, t: torch.Tensor, buf: torch.Tensor):
        if t.dtype != buf.dtype or not t.is_contiguous():
            return None
        off = t.data_ptr() - buf.data_ptr()
        if off < 0 or off // 2 + t.numel() > buf.numel():
            return None
        return off // 2

    def _symm_offset(self, t: torch.Tensor):
        """Element offset of t inside the symmetric buffer, or None if t is not a contiguous view of it."""
        if t.dtype != torch.bfloat16 or not t.is_contiguous():
            return None
        off = t.data_ptr() - self._buf.data_ptr()
        if off < 0 or off % 16 or off // 2 + t.numel() > self.max_elems:
            return None
        return off // 2

    def _is_symmetric(self, t: torch.Tensor) -> bool:
        return self._symm_offset(t) is not None

    def _launch(self, y, out, bias, residual, p, seed, offset, mask_in, dropout_elem_offset=0):
        rows, h = y.shape[0], y.shape[1]
        self._epoch += 1
        lib = self._lib.lib()
        self._lib.check(lib.fx_tp_allreduce_bdr(
            torch.cuda.current_stream().cuda_stream, self._peer_bufs, self._peer_sigs, self._mc, self.rank, self.nranks,
            self._epoch, self.num_blocks, None if out is None else out.data_ptr(), rows, h,
            None if bias is None else bias.data_ptr(), None if residual is None else residual.data_ptr(),
            float(p), int(seed), int(offset), None if mask_in is None else mask_in.data_ptr(),
            self._symm_offset(y), int(dropout_elem_offset)), "fx_tp_allreduce_bdr")
        return out

    def all_reduce(self, t: torch.Tensor) -> torch.Tensor:
        if not self._is_symmetric(t):
            self._fallback("all_reduce")
            return self.all_reduce_(t)
        t2 = t.view(-1, t.shape[-1])
        if self.nranks == 2:  # one-shot peer-load variant: the peer still reads my copy, so the sum goes elsewhere
            return self._launch(t2, torch.empty_like(t2), None, None, 0.0, 0, 0, None).view(t.shape)
        # in place: the switch writes the sum into every rank's copy of the symmetric buffer, which the
        # caller consumes before the next row-parallel GEMM of this rank overwrites it (stream order)
        self._launch(t2, None, None, None, 0.0, 0, 0, None)
        return t

    def all_reduce_start(self, t: torch.Tensor):
        if not self._is_symmetric(t):
            self._fallback("all_reduce_start")
            return super().all_reduce_start(t)
        t2 = t.view(-1, t.shape[-1])
        out = torch.empty_like(t2) if self.nranks == 2 else None  # allocated on the caller's stream
        main = torch.cuda.current_stream()
        ready = torch.cuda.Event()
        ready.record(main)
        with torch.cuda.stream(self._comm_stream):
            self._comm_stream.wait_event(ready)
            self._launch(t2, out, None, None, 0.0, 0, 0, None)
            done = torch.cuda.Event()
            done.record(self._comm_stream)
        return ("symm", (out.view(t.shape) if out is not None else t), done)

    def all_reduce_finish(self, handle) -> torch.Tensor:
        if isinstance(handle, tuple) and len(handle) == 3 and handle[0] == "symm":
            torch.cuda.current_stream().wait_event(handle[2])
            return handle[1]
        return super().all_reduce_finish(handle)

    def all_reduce_bias_dropout_residual(self, y, bias, residual, p, seed, offset, mask_in=None):
        if not self._is_symmetric(y):
            self._fallback("all_reduce_bias_dropout_residual")
            return super().all_reduce_bias_dropout_residual(y, bias, residual, p, seed, offset, mask_in)
        y2 = y.view(-1, y.shape[-1])
        out = torch.empty_like(y2)
        return self._launch(y2, out, bias, residual.view_as(y2), p, seed, offset, mask_in).view(y.shape)

    def all_reduce_bdr_start(self, y, out, bias, residual, p, seed, offset, mask_in, dropout_elem_offset):
        """Fused allreduce + bias + dropout + residual of a row slice, on the group's stream. y is a view
        of the symmetric buffer, out / residual / mask_in are the matching row slices."""
        main = torch.cuda.current_stream()
        ready = torch.cuda.Event()
        ready.record(main)
        with torch.cuda.stream(self._comm_stream):
            self._comm_stream.wait_event(ready)
            self._launch(y, out, bias, residual, p, seed, offset, mask_in, dropout_elem_offset)
            done = torch.cuda.Event()
            done.record(self._comm_stream)
        return done

    def all_reduce_bdr_finish(self, done) -> None:
        torch.cuda.current_stream().wait_event(done)


_DEFAULT_GROUP = None


def set_mp_group(group) -> None:
    """Install the group the parallel linears use when `mp_group=None` (the reference's layers
    look theirs up from the hybrid communicate group the same way)."""
    global _DEFAULT_GROUP
    _DEFAULT_GROUP = group


def get_mp_group():
    return _DEFAULT_GROUP if _DEFAULT_GROUP is not None else SingleGroup()


def resolve_group(mp_group=None, num_partitions: Optional[int] = None):
    g = mp_group if mp_group is not None else get_mp_group()
    if num_partitions is not None and num_partitions != g.nranks:
        raise ValueError(
            f"num_partitions={num_partitions} does not match the model-parallel group size {g.nranks}")
    return g


# ------------------------------------------------------------------------------ sharding
def shard_columns(t: torch.Tensor, rank: int, nranks: int) -> torch.Tensor:
    """Column-parallel split of a [in, out] weight (dim 1) or a [out] bias (dim 0)."""
    dim = t.dim() - 1
    if t.shape[dim] % nranks:
        raise ValueError(f"dimension {t.shape[dim]} is not divisible by mp degree {nranks}")
    return t.chunk(nranks, dim=dim)[rank].contiguous()


def shard_rows(t: torch.Tensor, rank: int, nranks: int) -> torch.Tensor:
    """Row-parallel split of a [in, out] weight along dim 0."""
    if t.shape[0] % nranks:
        raise ValueError(f"dimension {t.shape[0]} is not divisible by mp degree {nranks}")
    return t.chunk(nranks, dim=0)[rank].contiguous()


# ------------------------------------------------------------------------------ autograd collectives
class _CIdentity(torch.autograd.Function):
    """Paddle's `_c_identity`: identity forward, allreduce(sum) of the gradient backward — what
    ColumnParallelLinear puts in front of its matmul."""

    @staticmethod
    def forward(ctx, x, group):
        ctx.group = group
        return x

    @staticmethod
    def backward(ctx, dy):
        dy = dy.contiguous().clone()
        ctx.group.all_reduce_(dy)
        return dy, None


class _MpAllReduce(torch.autograd.Function):
    """Paddle's `_mp_allreduce`: allreduce(sum) forward, identity backward — what
    RowParallelLinear puts behind its matmul."""

    @staticmethod
    def forward(ctx, x, group):
        y = x.contiguous().clone()
        group.all_reduce_(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        return dy, None


class _CConcat(torch.autograd.Function):
    """Paddle's `_c_concat`: all-gather along the LAST dim forward (rank r's slice lands at columns
    [r * w, (r + 1) * w)), take this rank's slice of the gradient backward — what parallel_matmul puts behind the
    vocabulary-split logits when parallel_output=False (hybrid_model.py:59-63). Not on the training hot path (the
    pretraining criterion consumes the split logits), so torch.distributed does the exchange."""

    @staticmethod
    def forward(ctx, x, group):
        ctx.group, ctx.width = group, x.shape[-1]
        n = group.nranks
        if n == 1:
            return x
        x2 = x.contiguous()
        flat = torch.empty((n * x2.shape[0],) + tuple(x2.shape[1:]), dtype=x2.dtype, device=x2.device)
        group._dist.all_gather_into_tensor(flat, x2, group=group.pg)  # rank-major along dim 0
        # [n, ..., w] -> [..., n * w]
        return flat.view((n,) + tuple(x2.shape)).movedim(0, -2).reshape(*x2.shape[:-1], n * x2.shape[-1])

    @staticmethod
    def backward(ctx, dy):
        g, w = ctx.group, ctx.width
        if g.nranks == 1:
            return dy, None
        return dy[..., g.rank * w:(g.rank + 1) * w].contiguous(), None


def c_concat(x: torch.Tensor, group) -> torch.Tensor:
    return _CConcat.apply(x, group)


def c_identity(x: torch.Tensor, group) -> torch.Tensor:
    return _CIdentity.apply(x, group)


def mp_allreduce(x: torch.Tensor, group) -> torch.Tensor:
    return _MpAllReduce.apply(x, group)


# ---------------------------------------------------------------------------- sequence-parallel operators
# sequence_parallel_utils.py:36-126: scatter / all_gather / reduce_scatter along dim 0 of a sequence-first
# [s, b, h] tensor and the four autograd operators built on them. These keep the reference's layout on both
# sides (no permutation): the building blocks of ColumnSequenceParallelLinear / RowSequenceParallelLinear
# outside the fused layer path, and what the CPU (gloo) tests drive.
def sp_scatter(x: torch.Tensor, group) -> torch.Tensor:
    n = group.nranks
    if x.shape[0] % n != 0:
        raise ValueError(f"Input sequence length {x.shape[0]} can't be divided exactly by sequence parallelism {n}")
    sl = x.shape[0] // n
    return x[group.rank * sl:(group.rank + 1) * sl].contiguous()


def sp_all_gather_dim0(x: torch.Tensor, group) -> torch.Tensor:
    n = group.nranks
    if n == 1:
        return x
    out = torch.empty((x.shape[0] * n,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    group._dist.all_gather_into_tensor(out, x.contiguous(), group=group.pg)
    return out


def sp_reduce_scatter_dim0(x: torch.Tensor, group) -> torch.Tensor:
    n = group.nranks
    if n == 1:
        return x
    if x.shape[0] % n != 0:
        raise ValueError(f"Input sequence length {x.shape[0]} can't be divided exactly by sequence parallelism {n}")
    if group._dist.get_backend(group.pg) == "nccl":
        out = torch.empty((x.shape[0] // n,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
        group._dist.reduce_scatter_tensor(out, x.contiguous(), op=group._dist.ReduceOp.SUM, group=group.pg)
        return out
    t = x.clone()
    group._dist.all_reduce(t, op=group._dist.ReduceOp.SUM, group=group.pg)  # gloo has no reduce_scatter
    return sp_scatter(t, group)


class ScatterOp(torch.autograd.Function):
    """[s, b, h] -> [s/n, b, h]; backward all-gathers (sequence_parallel_utils.py:73-83)."""

    @staticmethod
    def forward(ctx, x, group):
        ctx.group = group
        return sp_scatter(x, group)

    @staticmethod
    def backward(ctx, g):
        return sp_all_gather_dim0(g, ctx.group), None


class GatherOp(torch.autograd.Function):
    """[s/n, b, h] -> [s, b, h]; backward scatters (:86-96)."""

    @staticmethod
    def forward(ctx, x, group):
        ctx.group = group
        return sp_all_gather_dim0(x, group)

    @staticmethod
    def backward(ctx, g):
        return sp_scatter(g, ctx.group), None


class AllGatherOp(torch.autograd.Function):
    """all-gather forward, reduce-scatter backward (:99-112): in front of ColumnSequenceParallelLinear."""

    @staticmethod
    def forward(ctx, x, group):
        ctx.group = group
        return sp_all_gather_dim0(x, group)

    @staticmethod
    def backward(ctx, g):
        return sp_reduce_scatter_dim0(g, ctx.group), None


class ReduceScatterOp(torch.autograd.Function):
    """reduce-scatter forward, all-gather backward (:115-126): behind RowSequenceParallelLinear's matmul."""

    @staticmethod
    def forward(ctx, x, group):
        ctx.group = group
        return sp_reduce_scatter_dim0(x, group)

    @staticmethod
    def backward(ctx, g):
        return sp_all_gather_dim0(g, ctx.group), None


def all_reduce_gradient_hook(group):
    """sequence_parallel_utils.py:136-140: parameters that see only this rank's sequence shard (LayerNorm
    weight / bias, the bias of RowSequenceParallelLinear) sum their gradients over the mp group."""
    def hook(grad):
        return group.all_reduce_(grad.contiguous())
    return hook
