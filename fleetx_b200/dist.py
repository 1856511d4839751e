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


class SingleGroup:
    rank = 0
    nranks = 1

    def all_reduce_(self, t: torch.Tensor) -> torch.Tensor:
        return t

    def barrier(self):
        return None


class TorchDistGroup:
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


def c_identity(x: torch.Tensor, group) -> torch.Tensor:
    return _CIdentity.apply(x, group)


def mp_allreduce(x: torch.Tensor, group) -> torch.Tensor:
    return _MpAllReduce.apply(x, group)
