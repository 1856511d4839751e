"""GPTModelHybrid and what surrounds the decoder stack, mirrored from
ppfleetx/models/language_model/gpt/dygraph/hybrid_model.py:

  parallel_matmul                   :45-66    logits against the (tied) word embedding
  TransformerDecoder                :350-437  N decoder layers + final LayerNorm
  GPTEmbeddings                     :574-622  vocab-parallel word embedding + position embedding + dropout
  GPTModelHybrid                    :625-742
  GPTForPretrainingHybrid           :745-788
  GPTPretrainingCriterionHybird     :791-829  (the reference's spelling) per-token cross entropy, masked mean

Same class names, constructor arguments and parameter names; torch.nn.Module stands in for paddle.nn.Layer.
The arithmetic is libfleetx_b200.so (csrc/model.cu for the embedding and the cross entropy, the GEMM for the
logits, the decoder layer as before); torch carries buffers, autograd edges and the small [T]-vector
collectives of the vocab-parallel cross entropy.
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import nn

from . import _lib, ops
from ._lib import check
from .dist import c_concat, c_identity, resolve_group
from .layers import LayerNorm, TransformerDecoderLayer, _init_weight
from .rng import get_rng_state_tracker


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _p(t):
    return None if t is None else t.data_ptr()


def _need_cuda(t: torch.Tensor, dtype, name: str) -> torch.Tensor:
    if not t.is_cuda:
        raise _lib.FleetXError(f"{name}: expected a CUDA tensor (there is no CPU path)")
    if t.dtype != dtype:
        raise _lib.FleetXError(f"{name}: expected {dtype}, got {t.dtype}")
    return t.contiguous()


# ---------------------------------------------------------------------------------------- embedding
class EmbeddingFunction(torch.autograd.Function):
    """GPTEmbeddings.forward (:603-622) as one node: word lookup in this rank's vocabulary slice (zeros for
    ids of other ranks, summed over the mp group), + position embedding, dropout."""

    @staticmethod
    def forward(ctx, wte, wpe, ids, pos_ids, vocab_start, p, seed, offset, group):
        b, s = ids.shape
        T, h = b * s, wte.shape[1]
        ids1 = _need_cuda(ids, torch.int64, "input_ids").view(-1)
        pos1 = None if pos_ids is None else _need_cuda(pos_ids, torch.int64, "position_ids").view(-1)
        out = torch.empty(T, h, dtype=torch.bfloat16, device=wte.device)
        lib = _lib.lib()
        if group.nranks == 1:
            check(lib.fx_embedding_fwd(_stream(), _p(ids1), _p(pos1), _p(wte), vocab_start, wte.shape[0], None, _p(wpe),
                                       _p(out), T, s, h, float(p), int(seed), int(offset)), "fx_embedding_fwd")
        else:
            part = group.partial_buffer(T, h, wte.device)
            check(lib.fx_embedding_fwd(_stream(), _p(ids1), None, _p(wte), vocab_start, wte.shape[0], None, None,
                                       _p(part), T, s, h, 0.0, 0, 0), "fx_embedding_fwd")
            part = group.all_reduce(part)  # VocabParallelEmbedding's allreduce
            check(lib.fx_embedding_fwd(_stream(), None, _p(pos1), None, 0, 0, _p(part), _p(wpe), _p(out), T, s, h,
                                       float(p), int(seed), int(offset)), "fx_embedding_fwd")
        ctx.save_for_backward(ids1, pos1 if pos1 is not None else torch.empty(0, device=wte.device))
        ctx.cfg = (b, s, h, vocab_start, wte.shape[0], wpe.shape[0], float(p), int(seed), int(offset), pos1 is not None)
        return out.view(b, s, h)

    @staticmethod
    def backward(ctx, dout):
        ids1, pos1 = ctx.saved_tensors
        b, s, h, vocab_start, rows, max_pos, p, seed, offset, has_pos = ctx.cfg
        dev = dout.device
        dwte = torch.zeros(rows, h, dtype=torch.float32, device=dev)
        dwpe = torch.zeros(max_pos, h, dtype=torch.float32, device=dev)
        lib = _lib.lib()
        check(lib.fx_embedding_bwd(_stream(), _p(dout.contiguous()), _p(ids1), _p(pos1) if has_pos else None, _p(dwte),
                                   vocab_start, rows, _p(dwpe), b * s, s, h, p, seed, offset), "fx_embedding_bwd")
        gw = torch.empty(rows, h, dtype=torch.bfloat16, device=dev)
        gp = torch.empty(max_pos, h, dtype=torch.bfloat16, device=dev)
        check(lib.fx_f32_to_bf16(_stream(), _p(dwte), _p(gw), dwte.numel()), "fx_f32_to_bf16")
        check(lib.fx_f32_to_bf16(_stream(), _p(dwpe), _p(gp), dwpe.numel()), "fx_f32_to_bf16")
        return gw, gp, None, None, None, None, None, None, None


# ---------------------------------------------------------------------------------------- logits
class LogitsFunction(torch.autograd.Function):
    """paddle.matmul(x, word_embedding_weight, transpose_y=True) (:55, :64): [T, h] x [Vl, h]^T -> [T, Vl]."""

    @staticmethod
    def forward(ctx, x, wte):
        x2 = x.reshape(-1, x.shape[-1])
        ctx.save_for_backward(x2, wte)
        ctx.in_shape = x.shape
        return ops.gemm(x2, wte, trans_b=True).view(*x.shape[:-1], wte.shape[0])

    @staticmethod
    def backward(ctx, dlogits):
        x2, wte = ctx.saved_tensors
        d2 = dlogits.reshape(-1, dlogits.shape[-1])
        dx = ops.gemm(d2, wte)                   # [T, Vl] x [Vl, h]
        dw = ops.gemm(d2, x2, trans_a=True)      # [Vl, T] x [T, h]
        return dx.view(ctx.in_shape), dw


def parallel_matmul(lm_output, logit_weights, parallel_output, mp_group=None):
    """hybrid_model.py:45-66: _c_identity, the matmul against this rank's vocabulary slice, and — with
    parallel_output=False (:59-63, the eval / generation callers) — _c_concat of the slices into full logits."""
    group = resolve_group(mp_group)
    if group.nranks > 1:
        lm_output = c_identity(lm_output, group)
        logits = LogitsFunction.apply(lm_output, logit_weights)
        return logits if parallel_output else c_concat(logits, group)
    return LogitsFunction.apply(lm_output, logit_weights)


# ---------------------------------------------------------------------------------------- cross entropy
class MaskedCrossEntropyFunction(torch.autograd.Function):
    """GPTPretrainingCriterionHybird.forward (:801-829): softmax cross entropy per token over the (possibly
    mp-split) vocabulary, then sum(loss * mask) / sum(mask). dlogits is written over the logits."""

    @staticmethod
    def forward(ctx, logits, labels, loss_mask, vocab_start, group):
        vl = logits.shape[-1]
        lg = _need_cuda(logits, torch.bfloat16, "prediction_scores").view(-1, vl)
        T = lg.shape[0]
        lab = _need_cuda(labels, torch.int64, "masked_lm_labels").view(-1)
        mask = None if loss_mask is None else _need_cuda(loss_mask, torch.float32, "loss_mask").view(-1)
        dev = lg.device
        stats = torch.empty(3, T, dtype=torch.float32, device=dev)  # max | sum | target
        lib = _lib.lib()
        check(lib.fx_cross_entropy_stats(_stream(), _p(lg), _p(lab), vocab_start, T, vl, _p(stats[0]), _p(stats[1]),
                                         _p(stats[2])), "fx_cross_entropy_stats")
        gmax = stats[0]
        if group.nranks > 1:  # ParallelCrossEntropy: max, then sum of the rescaled partial sums and of the target logit
            gmax = stats[0].clone()
            group._dist.all_reduce(gmax, op=group._dist.ReduceOp.MAX, group=group.pg)
            check(lib.fx_cross_entropy_rescale(_stream(), _p(stats[0]), _p(gmax), _p(stats[1]), T), "fx_cross_entropy_rescale")
            group.all_reduce_(stats[1:])
        out2 = torch.empty(2, dtype=torch.float32, device=dev)
        check(lib.fx_cross_entropy_loss(_stream(), _p(gmax), _p(stats[1]), _p(stats[2]), _p(mask), None, _p(out2), T),
              "fx_cross_entropy_loss")
        ctx.save_for_backward(lg, lab, gmax, stats, out2, mask if mask is not None else torch.empty(0, device=dev))
        ctx.cfg = (vocab_start, T, vl, mask is not None, logits.shape)
        return out2[0]

    @staticmethod
    def backward(ctx, dloss):
        lg, lab, gmax, stats, out2, mask = ctx.saved_tensors
        vocab_start, T, vl, has_mask, shape = ctx.cfg
        dl = dloss.to(torch.float32).reshape(1).contiguous()
        if getattr(ctx, "consumed", False):
            raise RuntimeError("MaskedCrossEntropyFunction: dlogits are written over the logits, so this node can be "
                               "differentiated once (no retain_graph / double backward)")
        ctx.consumed = True
        check(_lib.lib().fx_cross_entropy_bwd(_stream(), _p(lg), _p(lg), _p(lab), vocab_start, T, vl, _p(gmax), _p(stats[1]),
                                              _p(mask) if has_mask else None, _p(out2), _p(dl)), "fx_cross_entropy_bwd")
        return lg.view(shape), None, None, None, None


# ---------------------------------------------------------------------------------------- modules
class VocabParallelEmbedding(nn.Module):
    """fleet.meta_parallel.VocabParallelEmbedding as GPTEmbeddings uses it (:589-593): the vocabulary is split
    in contiguous slices over the mp ranks. Looked up inside EmbeddingFunction; this module holds the slice."""

    def __init__(self, num_embeddings, embedding_dim, weight_attr=None, mp_group=None, device="cuda", dtype=torch.bfloat16):
        super().__init__()
        self.model_parallel_group = resolve_group(mp_group)
        n, r = self.model_parallel_group.nranks, self.model_parallel_group.rank
        assert num_embeddings % n == 0, "The length of the vocabulary must be divisible by the parallelism degree of MP"
        self.origin_num_embeddings = num_embeddings
        self.per_part_size = num_embeddings // n
        self.vocab_start_index = r * self.per_part_size
        self.weight = nn.Parameter(_init_weight((self.per_part_size, embedding_dim), device, dtype))
        self.weight.is_distributed = n > 1


class GPTEmbeddings(nn.Module):
    """hybrid_model.py:574-622."""

    def __init__(self, vocab_size, hidden_size=768, hidden_dropout_prob=0.1, max_position_embeddings=512,
                 type_vocab_size=16, initializer_range=0.02, sequence_parallel=False, mp_group=None, device="cuda",
                 dtype=torch.bfloat16):
        super().__init__()
        self.sequence_parallel = sequence_parallel
        self.word_embeddings = VocabParallelEmbedding(vocab_size, hidden_size, mp_group=mp_group, device=device, dtype=dtype)
        self.position_embeddings = nn.Embedding(max_position_embeddings, hidden_size, device=device, dtype=dtype)
        with torch.no_grad():
            self.position_embeddings.weight.copy_(_init_weight((max_position_embeddings, hidden_size), device, dtype))
        self.dropout_p = hidden_dropout_prob

    def forward(self, input_ids, position_ids=None):
        p = self.dropout_p if self.training else 0.0
        we = self.word_embeddings
        if self.sequence_parallel:
            # :614-619: [b, s, h] -> [s, b, h], keep this rank's sequence shard, then dropout under 'local_seed'
            # (distinct masks on the mp ranks, each for its own rows)
            from .dist import ScatterOp
            from .functional import DropoutFunction

            emb = EmbeddingFunction.apply(we.weight, self.position_embeddings.weight, input_ids, position_ids,
                                          we.vocab_start_index, 0.0, 0, 0, we.model_parallel_group)
            emb = ScatterOp.apply(emb.transpose(0, 1).contiguous(), we.model_parallel_group)
            if p > 0:
                seed, offset = get_rng_state_tracker().draw("local_seed")
                emb = DropoutFunction.apply(emb, p, seed, offset)
            return emb
        # self.dropout(embeddings) runs under Paddle's default generator, which env.py:39 seeds with the same
        # value as the 'global_seed' stream: identical on all mp ranks. Drawn from that stream here.
        seed, offset = get_rng_state_tracker().draw("global_seed") if p > 0 else (0, 0)
        return EmbeddingFunction.apply(we.weight, self.position_embeddings.weight, input_ids, position_ids,
                                       we.vocab_start_index, p, seed, offset, we.model_parallel_group)


class _LayerNormFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, eps):
        y, mean, rstd = ops.layernorm_fwd(x.contiguous(), weight, bias, eps)
        ctx.save_for_backward(x, weight, mean, rstd)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, weight, mean, rstd = ctx.saved_tensors
        dx, dg, db = ops.layernorm_bwd(dy.contiguous(), x.contiguous(), weight, mean, rstd)
        return dx, dg, db, None


class TransformerDecoder(nn.Module):
    """hybrid_model.py:350-437: the stack of decoder layers and the final LayerNorm."""

    def __init__(self, decoder_layers, num_layers, norm=None, hidden_size=None, use_recompute=False,
                 recompute_granularity="full", sequence_parallel=False, device="cuda", mp_group=None):
        super().__init__()
        self.num_layers = num_layers
        self.layers = decoder_layers
        self.use_recompute = use_recompute
        self.recompute_granularity = recompute_granularity
        self.sequence_parallel = sequence_parallel
        if norm == "LayerNorm":
            self.norm = LayerNorm(hidden_size, epsilon=1e-5, device=device)
            if sequence_parallel:  # :371-373: the final norm sees only this rank's rows
                from .dist import all_reduce_gradient_hook, resolve_group

                g = resolve_group(mp_group)
                self.norm.weight.register_hook(all_reduce_gradient_hook(g))
                self.norm.bias.register_hook(all_reduce_gradient_hook(g))
        elif norm is not None:
            raise ValueError("Only support LayerNorm")
        else:
            self.norm = None

    def forward(self, tgt, memory=None, tgt_mask=None, memory_mask=None, use_cache=False, cache=None):
        output = tgt
        new_caches = []
        for i, mod in enumerate(self.layers):  # 'full' recompute (:406-408) is honoured inside each layer
            if use_cache or cache is not None:  # :396-417
                output, new_cache = mod(output, memory, tgt_mask=tgt_mask, use_cache=True,
                                        cache=None if cache is None else cache[i])
                new_caches.append(new_cache)
            else:
                output = mod(output, memory, tgt_mask, use_cache, cache)
        if self.norm is not None:
            output = _LayerNormFunction.apply(output, self.norm.weight, self.norm.bias, self.norm._epsilon)
        return output if not (use_cache or cache is not None) else (output, new_caches)

    def gen_cache(self, memory, do_zip=False):
        """hybrid_model.py:425-437."""
        cache = [layer.gen_cache(memory) for layer in self.layers]
        return list(zip(*cache)) if do_zip else cache


class GPTModelHybrid(nn.Module):
    """hybrid_model.py:625-742. Training / evaluation without KV cache; the causal mask is the fused one in
    both modes (the reference's eval mask -1e4·triu(1) is the same causal mask, :716-726)."""

    def __init__(self, vocab_size=51200, hidden_size=768, num_layers=12, num_attention_heads=12, ffn_hidden_size=3072,
                 hidden_dropout_prob=0.1, attention_probs_dropout_prob=0.1, max_position_embeddings=512,
                 type_vocab_size=16, initializer_range=0.02, num_partitions=1, use_recompute=False, fused_linear=False,
                 recompute_granularity="full", sequence_parallel=False, mp_group=None, device="cuda",
                 dtype=torch.bfloat16):
        super().__init__()
        self.initializer_range = initializer_range
        self.hidden_size = hidden_size
        self.vocab_size = vocab_size
        self.model_parallel_group = resolve_group(mp_group, num_partitions)
        if self.model_parallel_group.nranks <= 1:
            sequence_parallel = False  # :650-652
        self.sequence_parallel = sequence_parallel
        self.embeddings = GPTEmbeddings(vocab_size, hidden_size, hidden_dropout_prob, max_position_embeddings,
                                        type_vocab_size, initializer_range, sequence_parallel,
                                        mp_group=self.model_parallel_group, device=device, dtype=dtype)
        layers = nn.ModuleList()
        for _ in range(num_layers):
            layers.append(TransformerDecoderLayer(
                d_model=hidden_size, nhead=num_attention_heads, dim_feedforward=ffn_hidden_size,
                dropout=hidden_dropout_prob, activation="gelu", attn_dropout=attention_probs_dropout_prob,
                act_dropout=hidden_dropout_prob, num_partitions=num_partitions, fused_linear=fused_linear,
                use_recompute=use_recompute, recompute_granularity=recompute_granularity,
                sequence_parallel=sequence_parallel, mp_group=self.model_parallel_group, device=device, dtype=dtype))
        self.decoder = TransformerDecoder(layers, num_layers, norm="LayerNorm", hidden_size=hidden_size,
                                          use_recompute=use_recompute, recompute_granularity=recompute_granularity,
                                          sequence_parallel=sequence_parallel, device=device,
                                          mp_group=self.model_parallel_group)

    def forward(self, input_ids, position_ids=None, attention_mask=None, use_cache=False, cache=None):
        if attention_mask is not None:
            # :716-729: in eval the reference adds its causal mask (-1e4 * triu(1)) to a given attention_mask and hands
            # the sum to the layers; a mask that already IS that causal mask is served by the fused causal kernel
            from .layers import accept_causal_mask

            accept_causal_mask(attention_mask, input_ids.shape[-1], "GPTModelHybrid(attention_mask)")
        if position_ids is None and cache is not None:  # :697-705: positions continue behind the cached tokens
            past = cache[0].k.shape[-2]
            position_ids = torch.arange(past, past + input_ids.shape[-1], dtype=torch.int64, device=input_ids.device)
            position_ids = position_ids.unsqueeze(0).expand_as(input_ids).contiguous()
        embedding_output = self.embeddings(input_ids=input_ids, position_ids=position_ids)  # :710-711
        out = self.decoder(embedding_output, memory=None, tgt_mask=None, use_cache=use_cache, cache=cache)
        if self.sequence_parallel:  # :738-740: [s/n, b, h] shards -> [s, b, h] -> [b, s, h]
            if use_cache or cache is not None:
                raise NotImplementedError("use_cache with sequence_parallel is not built")
            from .dist import GatherOp

            out = GatherOp.apply(out, self.model_parallel_group).transpose(0, 1).contiguous()
        return out


class GPTForPretrainingHybrid(nn.Module):
    """hybrid_model.py:745-788: the model with the logits head tied to the word embedding."""

    def __init__(self, gpt: GPTModelHybrid):
        super().__init__()
        self.gpt = gpt

    def forward(self, input_ids, position_ids=None, attention_mask=None, masked_positions=None, use_cache=False,
                cache=None):
        outputs = self.gpt(input_ids, position_ids=position_ids, attention_mask=attention_mask,
                           use_cache=use_cache, cache=cache)
        encoder_outputs, cached_kvs = outputs if (use_cache or cache is not None) else (outputs, None)  # :771-775
        logits = parallel_matmul(encoder_outputs, self.gpt.embeddings.word_embeddings.weight, True,
                                 self.gpt.model_parallel_group)
        return (logits, cached_kvs) if cached_kvs is not None else logits


def load_full_state(model: "GPTForPretrainingHybrid", full: dict) -> None:
    """Load FULL (unsharded) reference-named parameters into `model`, keeping this rank's shards: vocabulary
    rows of the word embedding, and per decoder layer what TransformerDecoderLayer.load_full_state keeps."""
    gpt = model.gpt
    we = gpt.embeddings.word_embeddings
    with torch.no_grad():
        w = full["gpt.embeddings.word_embeddings.weight"]
        we.weight.copy_(w[we.vocab_start_index:we.vocab_start_index + we.per_part_size].to(we.weight.dtype))
        gpt.embeddings.position_embeddings.weight.copy_(full["gpt.embeddings.position_embeddings.weight"].to(we.weight.dtype))
        gpt.decoder.norm.weight.copy_(full["gpt.decoder.norm.weight"])
        gpt.decoder.norm.bias.copy_(full["gpt.decoder.norm.bias"])
    for i, layer in enumerate(gpt.decoder.layers):
        pre = f"gpt.decoder.layers.{i}."
        layer.load_full_state({k[len(pre):]: v for k, v in full.items() if k.startswith(pre)})


class GPTPretrainingCriterionHybird(nn.Module):
    """hybrid_model.py:791-829 (spelling as in the reference)."""

    def __init__(self, topo=None, mp_group=None):
        super().__init__()
        self.model_parallel_group = resolve_group(mp_group)

    def forward(self, prediction_scores, masked_lm_labels, loss_mask):
        g = self.model_parallel_group
        vocab_start = g.rank * prediction_scores.shape[-1] if g.nranks > 1 else 0
        return MaskedCrossEntropyFunction.apply(prediction_scores, masked_lm_labels, loss_mask, vocab_start, g)
