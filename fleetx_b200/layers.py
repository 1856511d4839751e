"""Host-side mirror of the reference's layer API for the hot path — same class names, constructor
arguments, parameter names and [in, out] weight layout, so it stays a drop-in:

  ColumnParallelLinear / RowParallelLinear   fleet.meta_parallel.* (signatures mirrored in-tree at
                                             sequence_parallel_utils.py:151-159, 236-244)
  MultiHeadAttention                         hybrid_model.py:69-347
  TransformerDecoderLayer                    hybrid_model.py:439-571

torch.nn.Module stands in for paddle.nn.Layer (Paddle is not installable here; INTEGRATION.md
shows the Paddle custom-op binding over the same C ABI). All arithmetic runs in
libfleetx_b200.so; a CPU tensor or a missing library is an error, never a fallback.
"""
from __future__ import annotations

import collections
from typing import Optional

import torch
from torch import nn

from . import functional as Fn
from .dist import resolve_group
from .rng import get_rng_state_tracker

_INIT_STD = 0.02  # initializer_range (hybrid_model.py:671-673)


def _init_weight(shape, device, dtype) -> torch.Tensor:
    return (torch.randn(*shape, dtype=torch.float32) * _INIT_STD).to(device=device, dtype=dtype)


_causal_checked: dict = {}


def accept_causal_mask(mask: Optional[torch.Tensor], s: int, what: str) -> None:
    """The explicit-mask branch of core_attn (hybrid_model.py:273-275: softmax(product + attn_mask)) for the mask the
    reference's own model builds in eval, -1e4 * triu(ones, 1) (:716-726): in fp32 that is the fused causal softmax
    up to exp(-1e4) == 0, so the causal kernel serves it. The tensor is checked once (cached by storage / version);
    any other additive mask (a padding mask folded in) would need a kernel that reads it and is refused."""
    if mask is None:
        return
    key = (mask.data_ptr(), tuple(mask.shape), mask._version, str(mask.device))
    ok = _causal_checked.get(key)
    if ok is None:
        m = mask.detach()
        ok = m.dim() >= 2 and m.shape[-1] == s and m.shape[-2] == s and all(d == 1 for d in m.shape[:-2])
        if ok:
            m2 = m.reshape(s, s).float()
            ok = bool(((m2.tril() == 0).all() & (m2.triu(1) + torch.tril(torch.full_like(m2, -1e4)) <= -9999.0).all()).item())
        if len(_causal_checked) > 64:
            _causal_checked.clear()
        _causal_checked[key] = ok
    if not ok:
        raise NotImplementedError(f"{what}: only the causal additive mask (-1e4 * triu(1), what the reference's model "
                                  "passes in eval) is served, by the fused causal kernel; other masks are not built")


class ColumnParallelLinear(nn.Module):
    """Y = X·W + b with W [in, out] split by columns over the mp group; gather_output is not
    supported on this path (the decoder layer always passes gather_output=False,
    hybrid_model.py:125-131, 497-503)."""

    def __init__(self, in_features, out_features, weight_attr=None, has_bias=None, gather_output=True,
                 fuse_matmul_bias=False, mp_group=None, name=None, device="cuda", dtype=torch.bfloat16):
        super().__init__()
        self.model_parallel_group = resolve_group(mp_group)
        self.world_size = self.model_parallel_group.nranks
        self._name = name
        self.is_mp = self.world_size > 1
        self.gather_output = gather_output
        if gather_output and self.world_size > 1:
            raise NotImplementedError("ColumnParallelLinear(gather_output=True) is outside the decoder-layer path")
        assert out_features % self.world_size == 0, (
            f"Number of column of the weight for linear ({out_features}) must be divisible by "
            f"model parallel size ({self.world_size})")
        self.output_size_per_partition = out_features // self.world_size
        self.fuse_matmul_bias = fuse_matmul_bias  # the bias add is always fused into the GEMM epilogue here
        self.weight = nn.Parameter(_init_weight((in_features, self.output_size_per_partition), device, dtype))
        self.weight.is_distributed = self.is_mp
        if has_bias:
            self.bias = nn.Parameter(torch.zeros(self.output_size_per_partition, device=device, dtype=dtype))
            self.bias.is_distributed = self.is_mp
        else:
            self.bias = None

    def forward(self, x):
        return Fn.ColumnParallelLinearFunction.apply(x, self.weight, self.bias, self.model_parallel_group)


class RowParallelLinear(nn.Module):
    """Y = allreduce(X_shard·W_shard) + b with W [in, out] split by rows; input_is_parallel must be
    True on this path (hybrid_model.py:157-163, 505-511)."""

    def __init__(self, in_features, out_features, weight_attr=None, has_bias=True, input_is_parallel=False,
                 fuse_matmul_bias=False, mp_group=None, name=None, device="cuda", dtype=torch.bfloat16):
        super().__init__()
        self.model_parallel_group = resolve_group(mp_group)
        self.world_size = self.model_parallel_group.nranks
        self._name = name
        self.is_mp = self.world_size > 1
        self.input_is_parallel = input_is_parallel
        if not input_is_parallel and self.world_size > 1:
            raise NotImplementedError("RowParallelLinear(input_is_parallel=False) is outside the decoder-layer path")
        assert in_features % self.world_size == 0, (
            f"Number of row of the weight for linear ({in_features}) must be divisible by "
            f"model parallel size ({self.world_size})")
        self.input_size_per_partition = in_features // self.world_size
        self.fuse_matmul_bias = fuse_matmul_bias
        self.weight = nn.Parameter(_init_weight((self.input_size_per_partition, out_features), device, dtype))
        self.weight.is_distributed = self.is_mp
        if has_bias:
            self.bias = nn.Parameter(torch.zeros(out_features, device=device, dtype=dtype))  # replicated
        else:
            self.bias = None

    def forward(self, x):
        return Fn.RowParallelLinearFunction.apply(x, self.weight, self.bias, self.model_parallel_group)


class ColumnSequenceParallelLinear(ColumnParallelLinear):
    """sequence_parallel_utils.py:150-232: the input is this rank's [s/n, b, h] sequence shard; an
    all-gather along the sequence (AllGatherOp: reduce-scatter in backward) stands in front of the matmul.
    Stand-alone form; inside TransformerDecoderLayer(sequence_parallel=True) the gather is the library's
    own kernel and the layer runs as one autograd node."""

    def __init__(self, in_features, out_features, weight_attr=None, has_bias=None, gather_output=True,
                 fuse_matmul_bias=False, mp_group=None, name=None, device="cuda", dtype=torch.bfloat16):
        assert gather_output is False, "If sequence_parallel is True, gather_output is False"  # utils:171-172
        super().__init__(in_features, out_features, weight_attr, has_bias, False, fuse_matmul_bias, mp_group, name,
                         device, dtype)

    def forward(self, x):
        from .dist import AllGatherOp, SingleGroup

        x = AllGatherOp.apply(x, self.model_parallel_group) if self.is_mp else x  # utils:224-227
        return Fn.ColumnParallelLinearFunction.apply(x, self.weight, self.bias, SingleGroup())


class RowSequenceParallelLinear(RowParallelLinear):
    """sequence_parallel_utils.py:235-326: local matmul, reduce-scatter along the sequence
    (ReduceScatterOp: all-gather in backward), then the bias, whose gradient — it only sees this rank's rows
    — is summed over the mp group by a hook (:297-298)."""

    def __init__(self, in_features, out_features, weight_attr=None, has_bias=True, input_is_parallel=False,
                 fuse_matmul_bias=False, mp_group=None, name=None, device="cuda", dtype=torch.bfloat16):
        assert input_is_parallel is True, "If sequence_parallel is True, input_is_parallel should be true."  # utils:262-263
        super().__init__(in_features, out_features, weight_attr, has_bias, True, fuse_matmul_bias, mp_group, name,
                         device, dtype)
        if self.bias is not None and self.is_mp:
            from .dist import all_reduce_gradient_hook

            self.bias.register_hook(all_reduce_gradient_hook(self.model_parallel_group))

    def forward(self, x):
        from .dist import ReduceScatterOp, SingleGroup

        y = Fn.RowParallelLinearFunction.apply(x, self.weight, None, SingleGroup())
        if self.is_mp:
            y = ReduceScatterOp.apply(y, self.model_parallel_group)                # utils:319
        return Fn.AddBiasFunction.apply(y, self.bias) if self.bias is not None else y  # utils:322


class LayerNorm(nn.Module):
    """nn.LayerNorm(d_model, epsilon) with fp32 affine parameters (they stay fp32 under AMP-O2)."""

    def __init__(self, normalized_shape, epsilon=1e-5, device="cuda"):
        super().__init__()
        self._epsilon = epsilon
        self.weight = nn.Parameter(torch.ones(normalized_shape, device=device, dtype=torch.float32))
        self.bias = nn.Parameter(torch.zeros(normalized_shape, device=device, dtype=torch.float32))


class MultiHeadAttention(nn.Module):
    """hybrid_model.py:69-163. Only the fused-qkv, self-attention, training/eval-without-cache
    path the decoder layer takes (fuse=True, use_cache=False) is built."""

    Cache = collections.namedtuple("Cache", ["k", "v"])
    StaticCache = collections.namedtuple("StaticCache", ["k", "v"])

    def __init__(self, embed_dim, num_heads, dropout=0., kdim=None, vdim=None, need_weights=False,
                 weight_attr=None, bias_attr=None, fuse=True, num_partitions=1, fused_linear=False,
                 use_recompute=False, recompute_granularity="full", sequence_parallel=False,
                 mp_group=None, device="cuda", dtype=torch.bfloat16):
        super().__init__()
        self.embed_dim = embed_dim
        self.kdim = kdim if kdim is not None else embed_dim
        self.vdim = vdim if vdim is not None else embed_dim
        self.num_heads = num_heads
        self.dropout = dropout
        self.need_weights = need_weights
        self.fuse = fuse
        self.use_recompute = use_recompute
        self.recompute_granularity = recompute_granularity
        self.sequence_parallel = sequence_parallel  # honoured by TransformerDecoderLayer; see forward()
        if not fuse:
            raise NotImplementedError("only the fused qkv projection (fuse=True, the reference default) is built")
        if need_weights:
            raise NotImplementedError("need_weights=True would materialise the s x s matrix; not on this path")
        self.head_dim = embed_dim // num_heads
        assert self.head_dim * num_heads == self.embed_dim, "embed_dim must be divisible by num_heads"
        assert self.num_heads % num_partitions == 0, "num_heads {} must be divisible by num_partitions {}".format(
            self.num_heads, num_partitions)
        self.model_parallel_group = resolve_group(mp_group, num_partitions)
        self.num_heads = self.num_heads // num_partitions
        assert self.kdim == embed_dim and self.vdim == embed_dim
        self.qkv_proj = ColumnParallelLinear(embed_dim, 3 * embed_dim, weight_attr=weight_attr, has_bias=True,
                                             gather_output=False, fuse_matmul_bias=fused_linear,
                                             mp_group=self.model_parallel_group, device=device, dtype=dtype)
        self.out_proj = RowParallelLinear(embed_dim, embed_dim, weight_attr=weight_attr, has_bias=True,
                                          input_is_parallel=True, fuse_matmul_bias=fused_linear,
                                          mp_group=self.model_parallel_group, device=device, dtype=dtype)

    def gen_cache(self, key, value=None, type=Cache):
        """hybrid_model.py:243-266: an empty incremental cache ([b, nh_local, 0, hd]) for `key`'s batch, or one
        holding the given k, v. StaticCache (encoder-decoder attention) is outside this model."""
        if type == MultiHeadAttention.StaticCache:
            raise NotImplementedError("StaticCache is for encoder-decoder attention; GPT has none")
        if value is None:
            b = key.shape[0]
            mk = lambda: torch.empty(b, self.num_heads, 0, self.head_dim, dtype=torch.bfloat16, device=key.device)  # noqa: E731
            return self.Cache(mk(), mk())
        return self.Cache(key, value)

    _CAP_STEP = 256  # cache buffers grow in steps, so appending is in place most of the time
    _cache_fill: dict = {}  # (k buffer, v buffer) addresses -> rows filled: who may append in place

    def _cache_buffers(self, cache, need: int):
        """The [b, nh, cap, hd] buffers behind cache.k / cache.v (views [:, :, :L] of them) with cap >= need:
        the ones the cache already lives in when they are large enough, else new ones holding the L old rows."""
        k, v = cache.k, cache.v
        L = k.shape[2]
        bk, bv = k._base, v._base
        ok = (bk is not None and bv is not None and bk.dim() == 4 and bk.is_contiguous() and bv.is_contiguous()
              and bk.shape[2] >= need and bk.shape == bv.shape and bk.data_ptr() == k.data_ptr()
              and bv.data_ptr() == v.data_ptr() and bk.shape[:2] == k.shape[:2] and bk.shape[3] == k.shape[3]
              # only the NEWEST view of a buffer may grow in place: decoding again from an older Cache (beam search,
              # speculative decoding) must not overwrite rows a later Cache still shows — the reference concatenates
              # into fresh tensors (hybrid_model.py:178-184)
              and self._cache_fill.get((bk.data_ptr(), bv.data_ptr())) == L)
        if ok:
            return bk, bv
        cap = max((need + self._CAP_STEP - 1) // self._CAP_STEP * self._CAP_STEP, 2 * L)
        nk = torch.empty(k.shape[0], k.shape[1], cap, k.shape[3], dtype=torch.bfloat16, device=k.device)
        nv = torch.empty_like(nk)
        if L:
            nk[:, :, :L].copy_(k)
            nv[:, :, :L].copy_(v)
        if len(self._cache_fill) > 256:
            self._cache_fill.clear()  # entries of long-gone buffers; a miss only costs one copy
        self._cache_fill[(nk.data_ptr(), nv.data_ptr())] = L
        return nk, nv

    def attend_cached(self, mix: torch.Tensor, cache):
        """core attention of the t new tokens with an incremental cache (:178-184, :268-298 in eval): the new
        k, v go behind the cached ones; returns (context [b, t, h/n], Cache with L + t rows)."""
        from . import ops

        b, t = mix.shape[0], mix.shape[1]
        L = cache.k.shape[2]
        bk, bv = self._cache_buffers(cache, L + t)
        ops.kv_cache_append(mix, bk, bv, self.num_heads, self.head_dim, L)
        self._cache_fill[(bk.data_ptr(), bv.data_ptr())] = L + t  # rows of the buffers in use now
        scale = self.head_dim ** -0.5
        if L == 0 and t >= 64:  # a prompt: the tensor-core kernel
            ctx = ops.attn_causal_fwd(mix, self.num_heads, self.head_dim, scale)[0]
        else:
            ctx = ops.attn_decode(mix, bk, bv, self.num_heads, self.head_dim, L, scale)
        return ctx, self.Cache(bk[:, :, :L + t], bv[:, :, :L + t])

    def forward(self, query, key=None, value=None, attn_mask=None, use_cache=False, cache=None):
        if use_cache or cache is not None:
            if self.training and self.dropout > 0:
                raise NotImplementedError("use_cache is the inference path: call eval() first")
            from . import ops

            with torch.no_grad():
                q2 = query.reshape(-1, query.shape[-1]).contiguous()
                mix = ops.gemm(q2, self.qkv_proj.weight, epilogue=ops.FX_EPI_BIAS, bias=self.qkv_proj.bias)
                ctx, new_cache = self.attend_cached(mix.view(query.shape[0], query.shape[1], -1),
                                                    cache if cache is not None else self.gen_cache(query))
                out = self.out_proj(ctx)
            return (out, new_cache) if use_cache else out
        accept_causal_mask(attn_mask, query.shape[1], "MultiHeadAttention(attn_mask)")  # :273-275 with the model's mask
        if (key is not None and key is not query) or (value is not None and value is not query):
            raise NotImplementedError("self-attention only")
        if self.sequence_parallel:
            raise NotImplementedError("sequence_parallel runs through TransformerDecoderLayer (one fused node); the "
                                      "stand-alone attention module takes batch-first [b, s, h] input only")
        mix = self.qkv_proj(query)  # [b, s, 3h/n], packed per head q|k|v
        p = self.dropout if self.training else 0.0
        seed, offset = (0, 0)
        if p > 0:
            with get_rng_state_tracker().rng_state("local_seed"):  # :280
                seed, offset = get_rng_state_tracker().draw()
        recompute = self.use_recompute and self.recompute_granularity == "core_attn" and self.training  # :332-335
        out = Fn.CausalSelfAttentionFunction.apply(mix.contiguous(), self.num_heads, self.head_dim, p, seed, offset,
                                                   recompute)
        return self.out_proj(out)


class TransformerDecoderLayer(nn.Module):
    """hybrid_model.py:439-571. forward() runs the whole layer as one autograd node made of
    libfleetx_b200 kernels."""

    def __init__(self, d_model, nhead, dim_feedforward, dropout=0.1, activation="gelu", attn_dropout=None,
                 act_dropout=None, normalize_before=True, weight_attr=None, bias_attr=None, num_partitions=1,
                 fused_linear=False, recompute_attn=False, use_recompute=False, recompute_granularity="full",
                 sequence_parallel=False, mp_group=None, device="cuda", dtype=torch.bfloat16):
        config = {k: v for k, v in locals().items() if k not in ("self", "__class__")}
        super().__init__()
        self._config = config
        attn_dropout = dropout if attn_dropout is None else attn_dropout
        act_dropout = dropout if act_dropout is None else act_dropout
        self.normalize_before = normalize_before
        self.use_recompute = use_recompute
        self.recompute_granularity = recompute_granularity
        self.sequence_parallel = sequence_parallel
        if activation != "gelu":
            # the reference stores `activation` but hard-codes F.gelu(approximate=True) (:523, :558)
            raise NotImplementedError("the reference layer always applies tanh-GELU")
        self.model_parallel_group = resolve_group(mp_group, num_partitions)
        assert dim_feedforward % num_partitions == 0
        self.self_attn = MultiHeadAttention(d_model, nhead, dropout=attn_dropout, weight_attr=weight_attr,
                                            bias_attr=bias_attr, num_partitions=num_partitions,
                                            fused_linear=fused_linear, use_recompute=use_recompute,
                                            recompute_granularity=recompute_granularity,
                                            mp_group=self.model_parallel_group, device=device, dtype=dtype)
        self.linear1 = ColumnParallelLinear(d_model, dim_feedforward, weight_attr=weight_attr, gather_output=False,
                                            has_bias=True, fuse_matmul_bias=fused_linear,
                                            mp_group=self.model_parallel_group, device=device, dtype=dtype)
        self.linear2 = RowParallelLinear(dim_feedforward, d_model, weight_attr=weight_attr, input_is_parallel=True,
                                         has_bias=True, fuse_matmul_bias=fused_linear,
                                         mp_group=self.model_parallel_group, device=device, dtype=dtype)
        self.norm1 = LayerNorm(d_model, epsilon=1e-5, device=device)
        self.norm2 = LayerNorm(d_model, epsilon=1e-5, device=device)
        self.dropout1_p = dropout
        self.dropout2_p = act_dropout
        self.activation = activation
        # parity-test hooks (uint8 keep masks [b*s, h]); None = use the Philox generator
        self._inject_mask1: Optional[torch.Tensor] = None
        self._inject_mask2: Optional[torch.Tensor] = None

    def _params(self):
        a = self.self_attn
        return (self.norm1.weight, self.norm1.bias, a.qkv_proj.weight, a.qkv_proj.bias, a.out_proj.weight,
                a.out_proj.bias, self.norm2.weight, self.norm2.bias, self.linear1.weight, self.linear1.bias,
                self.linear2.weight, self.linear2.bias)

    def _meta(self) -> Fn.LayerMeta:
        tracker = get_rng_state_tracker()
        train = self.training
        p_attn = self.self_attn.dropout if train else 0.0
        p1 = self.dropout1_p if train else 0.0
        p2 = self.dropout2_p if train else 0.0
        rng_attn = rng_h1 = rng_h2 = (0, 0)
        # same consumption order as the reference forward: attention ('local_seed', :280), then
        # dropout1 and dropout2 ('global_seed', :546 and :556)
        if p_attn > 0:
            rng_attn = tracker.draw("local_seed")
        if p1 > 0 and self._inject_mask1 is None:
            rng_h1 = tracker.draw("global_seed")
        if p2 > 0 and self._inject_mask2 is None:
            rng_h2 = tracker.draw("global_seed")
        # use_recompute / recompute_granularity as in the reference (hybrid_model.py:332-335, 406-408,
        # 537-539): only in training, and only the granularity named
        recompute = self.recompute_granularity if (self.use_recompute and train) else None
        return Fn.LayerMeta(nh_local=self.self_attn.num_heads, hd=self.self_attn.head_dim, p_attn=p_attn,
                            p_hidden1=p1, p_hidden2=p2, normalize_before=self.normalize_before,
                            rng_attn=rng_attn, rng_h1=rng_h1, rng_h2=rng_h2, recompute=recompute,
                            sequence_parallel=self.sequence_parallel,
                            mask_h1=self._inject_mask1 if p1 > 0 else None,
                            mask_h2=self._inject_mask2 if p2 > 0 else None)

    def _forward_cached(self, tgt, use_cache, cache):
        """hybrid_model.py:525-566 with use_cache=True (inference, no dropout, no autograd): the same kernels as
        the training forward, the attention going through the KV cache. Returns (out, incremental_cache)."""
        from . import ops

        if self.training and (self.dropout1_p > 0 or self.dropout2_p > 0 or self.self_attn.dropout > 0):
            raise NotImplementedError("use_cache is the inference path: call eval() first")
        a, g = self.self_attn, self.model_parallel_group
        b, t, h = tgt.shape
        with torch.no_grad():
            x = tgt.reshape(b * t, h).contiguous()
            pre = self.normalize_before
            attn_in = ops.layernorm_fwd(x, self.norm1.weight, self.norm1.bias, self.norm1._epsilon)[0] if pre else x
            mix = ops.gemm(attn_in, a.qkv_proj.weight, epilogue=ops.FX_EPI_BIAS, bias=a.qkv_proj.bias)
            ctx, new_cache = a.attend_cached(mix.view(b, t, -1), cache if cache is not None else a.gen_cache(tgt))
            part = ops.gemm(ctx.view(b * t, -1), a.out_proj.weight, out=g.partial_buffer(b * t, h, x.device))
            r1 = g.all_reduce_bias_dropout_residual(part, a.out_proj.bias, x, 0.0, 0, 0)
            x1 = r1 if pre else ops.layernorm_fwd(r1, self.norm1.weight, self.norm1.bias, self.norm1._epsilon)[0]
            ffn_in = ops.layernorm_fwd(x1, self.norm2.weight, self.norm2.bias, self.norm2._epsilon)[0] if pre else x1
            act = ops.gemm(ffn_in, self.linear1.weight, epilogue=ops.FX_EPI_BIAS_GELU, bias=self.linear1.bias)
            part = ops.gemm(act, self.linear2.weight, out=g.partial_buffer(b * t, h, x.device))
            r2 = g.all_reduce_bias_dropout_residual(part, self.linear2.bias, x1, 0.0, 0, 0)
            out = r2 if pre else ops.layernorm_fwd(r2, self.norm2.weight, self.norm2.bias, self.norm2._epsilon)[0]
        out = out.view(b, t, h)
        return (out, new_cache) if use_cache else out

    def forward(self, tgt, memory=None, tgt_mask=None, use_cache=False, cache=None):
        if use_cache or cache is not None:
            return self._forward_cached(tgt, use_cache, cache)
        accept_causal_mask(tgt_mask, tgt.shape[0] * getattr(self.model_parallel_group, "nranks", 1) if self.sequence_parallel
                           else tgt.shape[1], "TransformerDecoderLayer(tgt_mask)")  # :273-275 with the model's mask
        if not tgt.is_cuda:
            raise RuntimeError("TransformerDecoderLayer needs CUDA tensors: there is no CPU path")
        return Fn.DecoderLayerFunction.apply(tgt.contiguous(), self._meta(), self.model_parallel_group,
                                             *self._params())

    def gen_cache(self, memory):
        return self.self_attn.gen_cache(memory, type=self.self_attn.Cache)

    # -------------------------------------------------------------------------- checkpoints
    def load_full_state(self, full: dict) -> None:
        """Load FULL (unsharded) reference-named parameters, keeping this rank's shard — the
        per-mp-rank layout of the reference's checkpoints (eager_engine.py:581-615)."""
        from .dist import shard_columns, shard_rows

        g = self.model_parallel_group
        own = dict(self.named_parameters())
        for name, t in full.items():
            if name in ("self_attn.qkv_proj.weight", "self_attn.qkv_proj.bias", "linear1.weight", "linear1.bias"):
                t = shard_columns(t, g.rank, g.nranks)
            elif name in ("self_attn.out_proj.weight", "linear2.weight"):
                t = shard_rows(t, g.rank, g.nranks)
            with torch.no_grad():
                own[name].copy_(t.to(own[name].dtype))
