"""CPU restatement of the model around the decoder stack — TEST INFRASTRUCTURE ONLY.

PARITY as for oracle/decoder_layer.py: held to the reference's own GPTForPretrainingHybrid + criterion executed
over oracle/paddle_shim.py (tests/golden/ref_model_*.npz, tests/test_reference_golden.py: mp 1 / 2, sequence
parallel, eval + KV cache); Paddle's kernels for the primitives are not installable here and stay unpinned. Restated from
ppfleetx/models/language_model/gpt/dygraph/hybrid_model.py:
  GPTEmbeddings.forward            :603-622   word + position embedding, dropout
  TransformerDecoder.forward       :381-423   layers, then the final LayerNorm
  GPTForPretrainingHybrid.forward  :764-788   logits = h · word_embedding^T (parallel_matmul :45-66)
  GPTPretrainingCriterionHybird    :801-829   CrossEntropyLoss(reduction="none") * loss_mask, / loss_mask.sum()
in plain torch (fp32 / fp64); the vocab-parallel forms (VocabParallelEmbedding, ParallelCrossEntropy) compute
the same numbers as the single-rank forms, which is what is stated here; gradients come from torch autograd.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Optional

import torch
import torch.nn.functional as F

from . import decoder_layer as L


@dataclass
class ModelConfig:
    vocab_size: int
    hidden_size: int
    num_layers: int
    num_attention_heads: int
    ffn_hidden_size: int
    max_position_embeddings: int
    hidden_dropout_prob: float = 0.1
    attention_probs_dropout_prob: float = 0.1
    num_partitions: int = 1

    def layer_cfg(self) -> L.LayerConfig:
        # hybrid_model.py:662-679: dropout = act_dropout = hidden_dropout_prob, attn_dropout separate
        return L.LayerConfig(d_model=self.hidden_size, nhead=self.num_attention_heads,
                             dim_feedforward=self.ffn_hidden_size, dropout=self.hidden_dropout_prob,
                             attn_dropout=self.attention_probs_dropout_prob, act_dropout=self.hidden_dropout_prob,
                             num_partitions=self.num_partitions)


def init_params(cfg: ModelConfig, seed: int = 1024, dtype=torch.float32) -> Dict[str, torch.Tensor]:
    """Reference parameter names (state_dict of GPTForPretrainingHybrid); Normal(0, 0.02) weights (:589-600)."""
    g = torch.Generator().manual_seed(seed)
    p = {
        "gpt.embeddings.word_embeddings.weight": torch.randn(cfg.vocab_size, cfg.hidden_size, generator=g) * 0.02,
        "gpt.embeddings.position_embeddings.weight": torch.randn(cfg.max_position_embeddings, cfg.hidden_size, generator=g) * 0.02,
    }
    for i in range(cfg.num_layers):
        for k, v in L.init_params(cfg.layer_cfg(), seed + 1 + i).items():
            p[f"gpt.decoder.layers.{i}.{k}"] = v
    p["gpt.decoder.norm.weight"] = torch.ones(cfg.hidden_size)
    p["gpt.decoder.norm.bias"] = torch.zeros(cfg.hidden_size)
    return {k: v.to(dtype) for k, v in p.items()}


@dataclass
class ModelMasks:
    embedding: Optional[torch.Tensor] = None          # [b, s, h] keep mask of the embedding dropout
    layers: Optional[List[L.DropoutMasks]] = None


def make_masks(cfg: ModelConfig, b: int, s: int, seed: int = 1024) -> ModelMasks:
    """The (seed, offset) draws of one training forward of fleetx_b200.model: the embedding dropout takes offset
    0 of 'global_seed'; layer i takes offset i of 'local_seed' and offsets 1 + 2i, 2 + 2i of 'global_seed'."""
    from . import philox

    lc = cfg.layer_cfg()
    emb = None
    if cfg.hidden_dropout_prob > 0:
        emb = torch.from_numpy(philox.dropout_mask(L.rng_seeds(seed)["global_seed"], 0, b * s * cfg.hidden_size,
                                                   cfg.hidden_dropout_prob)).reshape(b, s, cfg.hidden_size)
    layers = [L.make_masks(lc, b, s, seed, offsets=(i, 1 + 2 * i, 2 + 2 * i)) for i in range(cfg.num_layers)]
    return ModelMasks(embedding=emb, layers=layers)


def gpt_logits(input_ids: torch.Tensor, params: Dict[str, torch.Tensor], cfg: ModelConfig,
               masks: Optional[ModelMasks] = None, training: bool = True,
               position_ids: Optional[torch.Tensor] = None) -> torch.Tensor:
    b, s = input_ids.shape
    if position_ids is None:  # :604-607
        position_ids = torch.arange(s).unsqueeze(0).expand(b, s)
    wte = params["gpt.embeddings.word_embeddings.weight"]
    x = F.embedding(input_ids, wte) + F.embedding(position_ids, params["gpt.embeddings.position_embeddings.weight"])
    p = cfg.hidden_dropout_prob if training else 0.0
    x = L.dropout(x, p, masks.embedding if masks is not None else None)                        # :619-621
    lc = cfg.layer_cfg()
    for i in range(cfg.num_layers):
        lp = {k: params[f"gpt.decoder.layers.{i}.{k}"] for k in L.PARAM_NAMES}
        x = L.decoder_layer_forward(x, lp, lc, masks.layers[i] if masks is not None else None, training=training)
    x = L.layer_norm(x, params["gpt.decoder.norm.weight"], params["gpt.decoder.norm.bias"])      # :421-422
    return x @ wte.t()                                                                           # :778-781


def pretraining_loss(logits: torch.Tensor, labels: torch.Tensor, loss_mask: torch.Tensor) -> torch.Tensor:
    """hybrid_model.py:817-829."""
    lg = logits.reshape(-1, logits.shape[-1])
    lg = lg if lg.dtype == torch.float64 else lg.float()  # "Its data type should be float32" (:806)
    tok = F.cross_entropy(lg, labels.reshape(-1), reduction="none")
    m = loss_mask.reshape(-1).to(tok.dtype)
    return (tok * m).sum() / m.sum()


def loss_and_grads(input_ids, labels, loss_mask, params, cfg: ModelConfig, masks: Optional[ModelMasks] = None):
    ps = {k: v.detach().clone().requires_grad_(True) for k, v in params.items()}
    loss = pretraining_loss(gpt_logits(input_ids, ps, cfg, masks), labels, loss_mask)
    loss.backward()
    return loss.detach(), {k: v.grad.detach() for k, v in ps.items()}


def make_masks_sp(cfg: ModelConfig, b: int, s: int, seed: int = 1024) -> ModelMasks:
    """Masks of one training forward of the SEQUENCE-PARALLEL model (hybrid_model.py:614-619, 475-520), stated as
    full-size masks for the plain oracle model: the embedding dropout runs on each rank's [s/n, b, h] shard under
    ITS 'local_seed' stream (offset 0; distinct per rank), the attention dropout of layer i takes offset 1 + i of
    that stream, and dropout1 / dropout2 of layer i take offsets 2i, 2i + 1 of 'global_seed' on the shard — the
    same shard-sized mask on every rank (decoder_layer.make_masks_sp)."""
    from . import philox

    n, h = cfg.num_partitions, cfg.hidden_size
    sl = s // n
    emb = None
    if cfg.hidden_dropout_prob > 0:
        shards = [torch.from_numpy(philox.dropout_mask(L.rng_seeds(seed, r)["local_seed"], 0, sl * b * h,
                                                       cfg.hidden_dropout_prob)).reshape(sl, b, h) for r in range(n)]
        emb = L.sp_gather(shards)
    lc = cfg.layer_cfg()
    layers = [L.make_masks_sp(lc, b, s, seed, offsets=(1 + i, 2 * i, 2 * i + 1)) for i in range(cfg.num_layers)]
    return ModelMasks(embedding=emb, layers=layers)
