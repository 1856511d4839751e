"""TEST INFRASTRUCTURE — a stand-in `paddle` package over torch-CPU, just large enough to EXECUTE the
reference's own Python for the hot path (ppfleetx/models/language_model/gpt/dygraph/hybrid_model.py,
sequence_parallel_utils.py, single_model.py and ppfleetx/utils/env.py) in this container, where
PaddlePaddle itself cannot be installed.

What this pins and what it does not:
  * the reference files run UNMODIFIED from /root/reference (imported by path, never copied): the order
    of operations in TransformerDecoderLayer / MultiHeadAttention / GPTModelHybrid, the reshapes and
    transposes of the fused QKV layout, where each dropout sits and under which RNG-tracker stream, the
    post-/pre-LayerNorm variants, the KV-cache concatenation, and — for sequence parallelism — the
    in-tree ScatterOp / GatherOp / AllGatherOp / ReduceScatterOp, Column/RowSequenceParallelLinear and
    the gradient hooks, executed over a simulated multi-rank process group (one thread per rank);
  * the Paddle PRIMITIVES those files call are torch's own built-ins here (torch.layer_norm,
    torch.softmax, gelu(approximate="tanh"), matmul, embedding, cross_entropy): independent of the
    hand-written formulas in oracle/decoder_layer.py, but not Paddle's kernels;
  * classes that live in upstream Paddle and not in the reference tree — fleet.meta_parallel.
    ColumnParallelLinear / RowParallelLinear / VocabParallelEmbedding / ParallelCrossEntropy and the
    RNG state tracker — are restated here from their documented behaviour (marked [U] below, as in
    SURVEY.md); the in-tree Column/RowSequenceParallelLinear mirror their argument lists.

Used by tools/make_reference_golden.py (writes tests/golden/ref_*.npz) and by tests/ when
/root/reference is present. Never imported by the product package.

Dropout masks are not drawn here: `set_dropout_provider(fn)` installs fn(stream, rank, shape, p) -> keep
mask, called once per dropout in program order with the RNG-tracker stream that is active ("global_seed",
"local_seed", or None outside the tracker).
"""
from __future__ import annotations

import contextlib
import importlib
import math
import os
import sys
import threading
import types
from typing import Callable, List, Optional

import torch
import torch.nn.functional as TF

REFERENCE_ROOT = os.environ.get("FLEETX_REFERENCE", "/root/reference")

_tls = threading.local()
_state = {"default_dtype": torch.float32, "dropout_provider": None, "world": None, "dropout_log": []}


# ------------------------------------------------------------------------------------ tensors
class Tensor(torch.Tensor):
    """torch tensor whose `.shape` is a mutable list, as paddle.Tensor.shape is (the reference edits it in
    place: sequence_parallel_utils.py:51-52, :63-67)."""

    @property
    def shape(self):  # type: ignore[override]
        return list(torch.Tensor.size(self))

    @property
    def stop_gradient(self):
        return not self.requires_grad

    @stop_gradient.setter
    def stop_gradient(self, v):
        if self.is_leaf:
            self.requires_grad_(not v)

    def numpy(self):  # detached view, as paddle's
        return torch.Tensor.numpy(self.detach().as_subclass(torch.Tensor))


def _wrap(t: torch.Tensor) -> Tensor:
    return t if isinstance(t, Tensor) else t.as_subclass(Tensor)


def _dtype(d):
    if d is None:
        return _state["default_dtype"]
    if isinstance(d, torch.dtype):
        return d
    return {"float32": torch.float32, "float64": torch.float64, "float16": torch.float16,
            "bfloat16": torch.bfloat16, "int64": torch.int64, "int32": torch.int32, "bool": torch.bool}[str(d)]


def _resolve_shape(x, shape):
    """paddle.reshape: 0 copies the input dimension at that position, -1 is inferred."""
    return [x.size(i) if d == 0 else int(d) for i, d in enumerate(shape)]


# ------------------------------------------------------------------------------------ simulated ranks
class SimWorld:
    """n ranks = n threads of this process; collectives rendezvous on a barrier."""

    def __init__(self, n: int):
        self.n = n
        self.slots: List[Optional[torch.Tensor]] = [None] * n
        self.barrier = threading.Barrier(n, timeout=120)
        self.calls = []  # (rank, kind, shape) in issue order, for tests

    def exchange(self, rank: int, t: torch.Tensor, kind: str) -> List[torch.Tensor]:
        self.calls.append((rank, kind, tuple(t.shape)))
        self.slots[rank] = t.detach().as_subclass(torch.Tensor).clone()
        self.barrier.wait()
        got = list(self.slots)
        self.barrier.wait()
        return got


def current_rank() -> int:
    return getattr(_tls, "rank", 0)


def world_size() -> int:
    w = _state["world"]
    return w.n if w is not None else 1


def run_ranks(n: int, fn: Callable[[int], object]) -> list:
    """Run fn(rank) on n simulated mp ranks (threads); returns the per-rank results."""
    if n == 1:
        _tls.rank = 0
        _state["world"] = None
        return [fn(0)]
    world = SimWorld(n)
    _state["world"] = world
    out, err = [None] * n, [None] * n

    def body(r):
        _tls.rank = r
        _tls.stream = None
        try:
            out[r] = fn(r)
        except BaseException as e:  # noqa: BLE001
            err[r] = e
            world.barrier.abort()

    ts = [threading.Thread(target=body, args=(r,)) for r in range(n)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    _state["world"] = None
    for e in err:
        if e is not None and not isinstance(e, threading.BrokenBarrierError):
            raise e
    for e in err:
        if e is not None:
            raise e
    return out


class _Task:
    def wait(self):
        return None


class _ProcessGroup:
    """The three calls the reference makes on group.process_group (sequence_parallel_utils.py:55, 69, 139)."""

    def all_gather(self, tensor_in, tensor_out):
        parts = _state["world"].exchange(current_rank(), tensor_in, "all_gather")
        with torch.no_grad():
            tensor_out.copy_(torch.cat(parts, dim=0))
        return _Task()

    def _reduce_scatter_base(self, tensor_out, tensor_in):
        parts = _state["world"].exchange(current_rank(), tensor_in, "reduce_scatter")
        total = parts[0].clone()
        for p in parts[1:]:
            total = total + p
        n, r = world_size(), current_rank()
        rows = total.shape[0] // n
        with torch.no_grad():
            tensor_out.copy_(total[r * rows:(r + 1) * rows])
        return _Task()

    def allreduce(self, tensor):
        parts = _state["world"].exchange(current_rank(), tensor, "allreduce")
        total = parts[0].clone()
        for p in parts[1:]:
            total = total + p
        with torch.no_grad():
            tensor.copy_(total)
        return _Task()


class _Group:
    process_group = _ProcessGroup()

    @property
    def nranks(self):
        return world_size()

    @property
    def rank(self):
        return current_rank()


class _HCG:
    """fleet.get_hybrid_communicate_group(): only model parallelism is simulated (dp = pp = 1)."""
    _group = _Group()

    def get_model_parallel_group(self):
        return self._group

    def get_model_parallel_world_size(self):
        return world_size()

    def get_model_parallel_rank(self):
        return current_rank()

    def get_stage_id(self):
        return 0

    def get_data_parallel_rank(self):
        return 0

    def get_data_parallel_world_size(self):
        return 1

    def get_sharding_parallel_rank(self):
        return 0

    def get_sharding_parallel_world_size(self):
        return 1


def _allreduce_sum(t: torch.Tensor) -> torch.Tensor:
    if world_size() == 1:
        return t
    parts = _state["world"].exchange(current_rank(), t, "allreduce")
    total = parts[0].clone()
    for p in parts[1:]:
        total = total + p
    return _wrap(total)


class _CIdentity(torch.autograd.Function):
    """[U] paddle.distributed.collective._c_identity: identity forward, allreduce(sum) over mp backward."""

    @staticmethod
    def forward(ctx, x):
        return x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        return _allreduce_sum(g)


class _MpAllreduce(torch.autograd.Function):
    """[U] paddle.distributed.collective._mp_allreduce: allreduce(sum) forward, identity backward."""

    @staticmethod
    def forward(ctx, x):
        return _allreduce_sum(x)

    @staticmethod
    def backward(ctx, g):
        return g


class _CConcat(torch.autograd.Function):
    """[U] _c_concat: gather the last dim over mp forward, take this rank's slice backward."""

    @staticmethod
    def forward(ctx, x):
        if world_size() == 1:
            return x.view_as(x)
        parts = _state["world"].exchange(current_rank(), x, "c_concat")
        return _wrap(torch.cat(parts, dim=-1))

    @staticmethod
    def backward(ctx, g):
        n, r = world_size(), current_rank()
        return g.chunk(n, dim=-1)[r]


class PyLayer(torch.autograd.Function):
    """paddle.autograd.PyLayer: same static forward(ctx, ...) / backward(ctx, ...) protocol as
    torch.autograd.Function; gradients are handed to backward as paddle-like tensors (list `.shape`)."""

    def __init_subclass__(cls, **kw):
        super().__init_subclass__(**kw)
        user_bwd = cls.__dict__.get("backward")
        if user_bwd is not None:
            fn = user_bwd.__func__ if isinstance(user_bwd, staticmethod) else user_bwd

            def backward(ctx, *grads, _fn=fn):
                return _fn(ctx, *[_wrap(g) if isinstance(g, torch.Tensor) else g for g in grads])

            cls.backward = staticmethod(backward)


# ------------------------------------------------------------------------------------ RNG tracker, dropout
class _RNGTracker:
    """[U] fleet.meta_parallel.get_rng_state_tracker(): named generator states; `rng_state(name)` makes one
    current for the duration of the block. Only the NAME matters here — masks come from the provider."""

    def __init__(self):
        self.seeds = {}

    def reset(self):
        self.seeds = {}

    def add(self, name, seed):
        key = (name, current_rank())  # every rank is its own process in reality: one tracker each
        if key in self.seeds:
            raise ValueError("state {} already exists".format(name))
        self.seeds[key] = seed

    def seed_of(self, name, rank=None):
        return self.seeds[(name, current_rank() if rank is None else rank)]

    @contextlib.contextmanager
    def rng_state(self, name="model_parallel_rng"):
        prev = getattr(_tls, "stream", None)
        _tls.stream = name
        try:
            yield
        finally:
            _tls.stream = prev


_tracker = _RNGTracker()


def set_dropout_provider(fn):
    """fn(stream, rank, shape, p) -> keep mask (bool / uint8 tensor of `shape`) or None for 'keep all'."""
    _state["dropout_provider"] = fn
    _state["dropout_log"] = []


def dropout_log():
    """[(stream, rank, shape, p)] of every training-mode dropout executed since set_dropout_provider."""
    return list(_state["dropout_log"])


def _dropout(x, p=0.5, axis=None, training=True, mode="upscale_in_train", name=None):
    assert axis is None and mode == "upscale_in_train", "only what the reference uses"
    if not training or p == 0:
        return x
    stream = getattr(_tls, "stream", None)
    _state["dropout_log"].append((stream, current_rank(), tuple(x.shape), float(p)))
    fn = _state["dropout_provider"]
    if fn is None:
        raise RuntimeError("paddle_shim: dropout with p > 0 needs set_dropout_provider()")
    keep = fn(stream, current_rank(), tuple(x.shape), float(p))
    if keep is None:
        return x
    return x * keep.to(x.dtype) / (1.0 - p)


# ------------------------------------------------------------------------------------ nn
class ParamAttr:
    def __init__(self, name=None, initializer=None, learning_rate=1.0, regularizer=None, trainable=True,
                 do_model_average=True, need_clip=True):
        self.name, self.initializer, self.trainable = name, initializer, trainable


class _Normal:
    def __init__(self, mean=0.0, std=1.0, name=None):
        self.mean, self.std = mean, std

    def __call__(self, t):
        g = getattr(_tls, "init_gen", None)
        with torch.no_grad():
            t.copy_(torch.randn(t.shape, generator=g, dtype=torch.float64).mul_(self.std).add_(self.mean).to(t.dtype))


class _Constant:
    def __init__(self, value=0.0):
        self.value = value

    def __call__(self, t):
        with torch.no_grad():
            t.fill_(self.value)


class _Helper:
    def get_default_dtype(self):
        return _state["default_dtype"]


class Layer(torch.nn.Module):
    """paddle.nn.Layer over torch.nn.Module: create_parameter, sublayers, state-dict names are the same."""

    def __init__(self, name_scope=None, dtype="float32"):
        super().__init__()
        self._helper = _Helper()

    def create_parameter(self, shape, attr=None, dtype=None, is_bias=False, default_initializer=None):
        if attr is False:
            return None
        p = torch.nn.Parameter(torch.zeros(*shape, dtype=_dtype(dtype)))
        init = attr.initializer if isinstance(attr, ParamAttr) else attr if callable(attr) else None
        init = init or default_initializer or (_Constant(0.0) if is_bias else _Normal(0.0, 0.02))
        init(p)
        return p

    def add_sublayer(self, name, sublayer):
        self.add_module(name, sublayer)
        return sublayer

    def add_parameter(self, name, parameter):
        self.register_parameter(name, parameter)
        return parameter

    def sublayers(self, include_self=False):
        return [m for m in self.modules() if include_self or m is not self]

    def set_state_dict(self, sd):
        own = dict(self.named_parameters())
        assert set(own) == set(sd), (sorted(own), sorted(sd))
        with torch.no_grad():
            for k, v in sd.items():
                assert tuple(own[k].shape) == tuple(v.shape), (k, tuple(own[k].shape), tuple(v.shape))
                own[k].copy_(torch.as_tensor(v))


class LayerList(torch.nn.ModuleList):
    pass


class Linear(Layer):
    """paddle.nn.Linear: weight [in, out] (the transpose of torch's), y = x·W + b."""

    def __init__(self, in_features, out_features, weight_attr=None, bias_attr=None, name=None):
        super().__init__()
        self.weight = self.create_parameter([in_features, out_features], attr=weight_attr)
        self.bias = self.create_parameter([out_features], attr=bias_attr, is_bias=True)

    def forward(self, x):
        return _linear(x, self.weight, self.bias)


class LayerNorm(Layer):
    def __init__(self, normalized_shape, epsilon=1e-5, weight_attr=None, bias_attr=None, name=None):
        super().__init__()
        if isinstance(normalized_shape, int):
            normalized_shape = [normalized_shape]
        self._shape, self._eps = list(normalized_shape), epsilon
        self.weight = self.create_parameter(self._shape, attr=weight_attr, default_initializer=_Constant(1.0))
        self.bias = self.create_parameter(self._shape, attr=bias_attr, is_bias=True)

    def forward(self, x):
        return torch.layer_norm(x, self._shape, self.weight, self.bias, self._eps)


class Dropout(Layer):
    def __init__(self, p=0.5, axis=None, mode="upscale_in_train", name=None):
        super().__init__()
        self.p, self.axis, self.mode = p, axis, mode

    def forward(self, x):
        return _dropout(x, self.p, self.axis, self.training, self.mode)


class Embedding(Layer):
    def __init__(self, num_embeddings, embedding_dim, padding_idx=None, sparse=False, weight_attr=None, name=None):
        super().__init__()
        self.weight = self.create_parameter([num_embeddings, embedding_dim], attr=weight_attr)

    def forward(self, ids):
        return TF.embedding(ids, self.weight)


class CrossEntropyLoss(Layer):
    """paddle.nn.CrossEntropyLoss(reduction='none') on logits [..., V] and labels [..., 1] -> [..., 1]."""

    def __init__(self, weight=None, ignore_index=-100, reduction="mean", soft_label=False, axis=-1, name=None):
        super().__init__()
        assert reduction == "none" and not soft_label and axis == -1
        self.ignore_index = ignore_index

    def forward(self, logits, label):
        v = logits.shape[-1]
        lab = label.reshape(-1)
        loss = TF.cross_entropy(logits.reshape(-1, v), lab, reduction="none", ignore_index=self.ignore_index)
        return loss.reshape(list(label.shape))


def _linear(x, weight, bias=None, name=None):
    y = torch.matmul(x, weight)
    return y if bias is None else y + bias


# ---- [U] fleet.meta_parallel layers (upstream Paddle, not in the reference tree)
class ColumnParallelLinear(Layer):
    """[U] weight [in, out/n] (columns split over mp), bias [out/n]; forward = _c_identity(x)·W + b,
    gathered over mp only when gather_output. Argument list as mirrored by sequence_parallel_utils.py:151-159."""

    def __init__(self, in_features, out_features, weight_attr=None, has_bias=None, gather_output=True,
                 fuse_matmul_bias=False, mp_group=None, name=None):
        super().__init__()
        n = world_size()
        self.world_size, self.is_mp, self.gather_output = n, n > 1, gather_output
        assert out_features % n == 0
        self.weight = self.create_parameter([in_features, out_features // n], attr=weight_attr)
        self.weight.is_distributed = self.is_mp
        self.bias = self.create_parameter([out_features // n], is_bias=True) if has_bias else None
        if self.bias is not None:
            self.bias.is_distributed = self.is_mp

    def forward(self, x):
        xp = _CIdentity.apply(x) if self.is_mp else x
        y = _linear(xp, self.weight, self.bias)
        return _CConcat.apply(y) if (self.gather_output and self.is_mp) else y


class RowParallelLinear(Layer):
    """[U] weight [in/n, out] (rows split over mp), bias [out] replicated; forward = allreduce(x_r·W_r) + b,
    the bias added once after the allreduce. Argument list as mirrored by sequence_parallel_utils.py:236-244."""

    def __init__(self, in_features, out_features, weight_attr=None, has_bias=True, input_is_parallel=False,
                 fuse_matmul_bias=False, mp_group=None, name=None):
        super().__init__()
        n = world_size()
        self.world_size, self.is_mp, self.input_is_parallel = n, n > 1, input_is_parallel
        assert in_features % n == 0
        self.weight = self.create_parameter([in_features // n, out_features], attr=weight_attr)
        self.weight.is_distributed = self.is_mp
        self.bias = self.create_parameter([out_features], is_bias=True) if has_bias else None

    def forward(self, x):
        if self.is_mp and not self.input_is_parallel:
            x = x.chunk(self.world_size, dim=-1)[current_rank()]
        if not self.is_mp:
            return _linear(x, self.weight, self.bias)
        y = _MpAllreduce.apply(_linear(x, self.weight))
        return y + self.bias if self.bias is not None else y


class VocabParallelEmbedding(Layer):
    """[U] weight [vocab/n, h]: rank r owns ids [r·vocab/n, (r+1)·vocab/n); rows of other ranks' ids are
    zero, then allreduce(sum) over mp."""

    def __init__(self, num_embeddings, embedding_dim, weight_attr=None, mp_group=None, name=None):
        super().__init__()
        n = world_size()
        self.world_size, self.is_mp = n, n > 1
        assert num_embeddings % n == 0
        self.per_part = num_embeddings // n
        self.vocab_start_index = current_rank() * self.per_part
        self.weight = self.create_parameter([self.per_part, embedding_dim], attr=weight_attr)
        self.weight.is_distributed = self.is_mp

    def forward(self, ids):
        if not self.is_mp:
            return TF.embedding(ids, self.weight)
        local = ids - self.vocab_start_index
        mine = (local >= 0) & (local < self.per_part)
        out = TF.embedding(local.clamp(0, self.per_part - 1), self.weight) * mine.unsqueeze(-1).to(self.weight.dtype)
        return _MpAllreduce.apply(out)


class _ParallelSoftmaxCE(torch.autograd.Function):
    """[U] _c_softmax_with_cross_entropy on logits whose last dim is split over mp."""

    @staticmethod
    def forward(ctx, logits, label):
        n, r = world_size(), current_rank()
        vl = logits.shape[-1]
        gmax = logits.max(dim=-1, keepdim=True).values
        parts = _state["world"].exchange(r, gmax, "allreduce_max")
        gmax = torch.stack(parts).max(dim=0).values
        e = torch.exp(logits.as_subclass(torch.Tensor) - gmax)
        gsum = torch.stack(_state["world"].exchange(r, e.sum(dim=-1, keepdim=True), "allreduce")).sum(dim=0)
        local = label.as_subclass(torch.Tensor) - r * vl
        mine = (local >= 0) & (local < vl)
        tgt = torch.gather(logits.as_subclass(torch.Tensor), -1, local.clamp(0, vl - 1)) * mine.to(logits.dtype)
        tgt = torch.stack(_state["world"].exchange(r, tgt, "allreduce")).sum(dim=0)
        soft = e / gsum
        ctx.save_for_backward(soft, local, mine)
        return _wrap(torch.log(gsum) + gmax - tgt)

    @staticmethod
    def backward(ctx, g):
        soft, local, mine = ctx.saved_tensors
        onehot = torch.zeros_like(soft).scatter_(-1, local.clamp(0, soft.shape[-1] - 1), 1.0) * mine.to(soft.dtype)
        return (soft - onehot) * g, None


class ParallelCrossEntropy(Layer):
    def __init__(self, mp_group=None, name=None):
        super().__init__()

    def forward(self, logits, label):
        return _ParallelSoftmaxCE.apply(logits, label)


class LayerDesc:
    def __init__(self, layer_func, *inputs, **kwargs):
        self.layer_func, self.inputs, self.kwargs = layer_func, inputs, kwargs


class SharedLayerDesc(LayerDesc):
    def __init__(self, key, layer_func, forward_func=None, shared_weight_attr="weight", *inputs, **kwargs):
        super().__init__(layer_func, *inputs, **kwargs)


class PipelineLayer(Layer):
    """Placeholder base class: pipeline parallelism is outside the path."""


class LRScheduler:
    """[U] paddle.optimizer.lr.LRScheduler: the constructor stores last_epoch and takes one step(); step()
    advances last_epoch by one (or sets it) and caches get_lr()."""

    def __init__(self, learning_rate=0.1, last_epoch=-1, verbose=False):
        self.base_lr = self.last_lr = float(learning_rate)
        self.last_epoch, self.verbose = last_epoch, verbose
        self.step()

    def __call__(self):
        return self.last_lr

    def step(self, epoch=None):
        self.last_epoch = self.last_epoch + 1 if epoch is None else epoch
        self.last_lr = self.get_lr()

    def get_lr(self):
        raise NotImplementedError


def _recompute(function, *args, **kwargs):
    """[U] fleet.utils.recompute: same values and gradients as calling the function (it only trades memory)."""
    return function(*args, **kwargs)


# ------------------------------------------------------------------------------------ functional / tensor ops
def _transpose(x, perm, name=None):
    return x.permute(*perm)


def _reshape(x, shape, name=None):
    return x.reshape(_resolve_shape(x, shape))


def _split(x, num_or_sections, axis=0, name=None):
    if isinstance(num_or_sections, int):
        return list(torch.chunk(x, num_or_sections, dim=axis))
    return list(torch.split(x, num_or_sections, dim=axis))


def _matmul(x, y, transpose_x=False, transpose_y=False, alpha=1.0, name=None):
    if transpose_x:
        x = x.transpose(-1, -2)
    if transpose_y:
        y = y.transpose(-1, -2)
    out = torch.matmul(x, y)
    return out if alpha == 1.0 else out * alpha


def _softmax(x, axis=-1, dtype=None, name=None):
    return torch.softmax(x, dim=axis)


def _softmax_mask_fuse_upper_triangle(x):
    """[U] paddle.incubate.softmax_mask_fuse_upper_triangle: softmax over the last dim with everything above
    the diagonal masked out (x: [b, heads, s, s])."""
    s = x.shape[-1]
    upper = torch.ones(s, s, dtype=torch.bool).triu(1)
    return torch.softmax(x.masked_fill(upper, float("-inf")), dim=-1)


def _gelu(x, approximate=False, name=None):
    return TF.gelu(x, approximate="tanh" if approximate else "none")


def _to_tensor(data, dtype=None, place=None, stop_gradient=True):
    t = torch.as_tensor(data)
    if dtype is not None:
        t = t.to(_dtype(dtype))
    t = _wrap(t.clone())
    if not stop_gradient:
        t.requires_grad_(True)
    return t


def _fill_constant_batch_size_like(input, shape, dtype, value, input_dim_idx=0, output_dim_idx=0):
    shape = list(shape)
    shape[output_dim_idx] = input.shape[input_dim_idx]
    return _wrap(torch.full(shape, value, dtype=_dtype(dtype)))


def _ones_like(x, dtype=None):
    return _wrap(torch.ones_like(x, dtype=_dtype(dtype) if dtype is not None else None))


def _shape_of(x):
    return _wrap(torch.tensor(list(x.size()), dtype=torch.int64))


def _arange(start, end=None, step=1, dtype=None):
    f = lambda v: int(v) if isinstance(v, torch.Tensor) else v  # noqa: E731
    if end is None:
        start, end = 0, start
    return _wrap(torch.arange(f(start), f(end), f(step), dtype=_dtype(dtype) if dtype is not None else torch.int64))


def _ones(shape, dtype=None):
    return _wrap(torch.ones([int(d) for d in shape], dtype=_dtype(dtype)))


def _empty(shape, dtype=None):
    return _wrap(torch.empty([int(d) for d in shape], dtype=_dtype(dtype)))


def _module(name: str, **attrs) -> types.ModuleType:
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    m.__path__ = []  # behaves as a package for `import a.b`
    sys.modules[name] = m
    parent, _, leaf = name.rpartition(".")
    if parent:
        setattr(sys.modules[parent], leaf, m)
    return m


_installed = False


def install(default_dtype="float64") -> None:
    """Register the stand-in `paddle` modules and empty `ppfleetx` parent packages (so that importing one
    reference file does not execute ppfleetx/__init__.py and with it the whole framework)."""
    global _installed
    _state["default_dtype"] = _dtype(default_dtype)
    if _installed:
        return
    if "paddle" in sys.modules and not getattr(sys.modules["paddle"], "__fleetx_shim__", False):
        raise RuntimeError("a real paddle is already imported")
    hcg = _HCG()
    fwk = types.SimpleNamespace(in_dygraph_mode=lambda: True)
    _module("paddle", __fleetx_shim__=True, Tensor=Tensor, ParamAttr=ParamAttr, to_tensor=_to_tensor,
            reshape=_reshape, reshape_=_reshape, transpose=_transpose, split=_split, matmul=_matmul,
            concat=lambda xs, axis=0, name=None: torch.cat(list(xs), dim=axis), ones_like=_ones_like,
            cumsum=lambda x, axis=None, dtype=None: torch.cumsum(x, dim=axis), arange=_arange, shape=_shape_of,
            ones=_ones, empty=_empty, sum=lambda x, axis=None, keepdim=False: x.sum() if axis is None else x.sum(axis, keepdim),
            seed=lambda s: None, set_default_dtype=lambda d: _state.__setitem__("default_dtype", _dtype(d)),
            get_default_dtype=lambda: str(_state["default_dtype"]).replace("torch.", ""),
            in_dynamic_mode=lambda: True, is_compiled_with_cuda=lambda: False, is_compiled_with_rocm=lambda: False,
            set_device=lambda d: None, no_grad=torch.no_grad, framework=fwk)
    _module("paddle.nn", Layer=Layer, LayerList=LayerList, Linear=Linear, LayerNorm=LayerNorm, Dropout=Dropout,
            Embedding=Embedding, CrossEntropyLoss=CrossEntropyLoss)
    _module("paddle.nn.initializer", Normal=_Normal, Constant=_Constant)
    _module("paddle.nn.functional", softmax=_softmax, dropout=_dropout, gelu=_gelu, linear=_linear,
            embedding=lambda x, weight, **kw: TF.embedding(x, weight))
    _module("paddle.nn.layer")
    _module("paddle.nn.layer.transformer",
            _convert_param_attr_to_list=lambda attr, n: list(attr) if isinstance(attr, (list, tuple)) else [attr] * n)
    _module("paddle.tensor", concat=sys.modules["paddle"].concat, reshape=lambda x, shape, name=None: _reshape(x, shape),
            transpose=lambda x, perm, name=None: _transpose(x, perm), matmul=_matmul,
            triu=lambda x, diagonal=0, name=None: torch.triu(x, diagonal))
    _module("paddle.fluid")
    _module("paddle.fluid.layers", matmul=_matmul, fill_constant_batch_size_like=_fill_constant_batch_size_like,
            expand_as=lambda x, target: x.expand_as(target))
    _module("paddle.fluid.core", ops=types.SimpleNamespace())
    _module("paddle.fluid.dygraph")
    _module("paddle.fluid.dygraph.layers", Layer=Layer)
    _module("paddle.framework", in_dygraph_mode=lambda: True)
    _module("paddle.autograd", PyLayer=PyLayer)
    _module("paddle.incubate", softmax_mask_fuse_upper_triangle=_softmax_mask_fuse_upper_triangle)
    _module("paddle.incubate.nn", FusedLinear=Linear)
    _module("paddle.incubate.nn.functional", fused_linear=_linear)
    _module("paddle.common_ops_import", convert_dtype=lambda d: str(d).replace("torch.", ""))
    _module("paddle.optimizer")
    _module("paddle.optimizer.lr", LRScheduler=LRScheduler)
    _module("paddle.distributed", get_world_size=world_size, get_rank=current_rank)
    _module("paddle.distributed.auto_parallel", shard_tensor=lambda x, *a, **k: x, shard_op=lambda f, *a, **k: f)
    _module("paddle.distributed.collective", _c_identity=lambda x, group=None: _CIdentity.apply(x) if world_size() > 1 else x,
            _c_concat=lambda x, group=None: _CConcat.apply(x),
            _mp_allreduce=lambda x, group=None, **kw: _MpAllreduce.apply(x))
    _module("paddle.distributed.fleet", get_hybrid_communicate_group=lambda: hcg,
            DistributedStrategy=lambda: types.SimpleNamespace())
    _module("paddle.distributed.fleet.base")
    _module("paddle.distributed.fleet.base.topology")
    _module("paddle.distributed.fleet.meta_parallel", get_rng_state_tracker=lambda: _tracker,
            ColumnParallelLinear=ColumnParallelLinear, RowParallelLinear=RowParallelLinear,
            VocabParallelEmbedding=VocabParallelEmbedding, ParallelCrossEntropy=ParallelCrossEntropy,
            LayerDesc=LayerDesc, SharedLayerDesc=SharedLayerDesc, PipelineLayer=PipelineLayer)
    _module("paddle.distributed.fleet.utils", recompute=_recompute)
    sys.modules["paddle"].nn.initializer = sys.modules["paddle.nn.initializer"]
    # parent packages of the reference files, WITHOUT running their __init__.py
    pkg = "ppfleetx"
    path = os.path.join(REFERENCE_ROOT, "ppfleetx")
    for part in ("", "models", "language_model", "gpt", "dygraph"):
        if part:
            pkg, path = pkg + "." + part, os.path.join(path, part)
        m = _module(pkg)
        m.__path__ = [path]
    a = _module("ppfleetx.models.language_model.gpt.auto")
    a.__path__ = [os.path.join(REFERENCE_ROOT, "ppfleetx", "models", "language_model", "gpt", "auto")]
    for sub in ("utils", "optims"):
        u = _module("ppfleetx." + sub)
        u.__path__ = [os.path.join(REFERENCE_ROOT, "ppfleetx", sub)]
    _installed = True


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "ppfleetx", "models", "language_model", "gpt", "dygraph",
                                       "hybrid_model.py"))


def load_reference(name: str = "hybrid_model"):
    """Import one of the reference's files for the path (hybrid_model, single_model, sequence_parallel_utils)
    straight from REFERENCE_ROOT; `env` gives ppfleetx/utils/env.py, `lr_scheduler` ppfleetx/optims/lr_scheduler.py."""
    install(str(_state["default_dtype"]).replace("torch.", ""))
    if name == "env":
        return importlib.import_module("ppfleetx.utils.env")
    if name == "auto_model":
        return importlib.import_module("ppfleetx.models.language_model.gpt.auto.auto_model")
    if name == "lr_scheduler":
        return importlib.import_module("ppfleetx.optims.lr_scheduler")
    return importlib.import_module("ppfleetx.models.language_model.gpt.dygraph." + name)


def rng_tracker():
    return _tracker
