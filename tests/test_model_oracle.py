"""Pins of the model oracle (oracle/gpt_model.py) that need no GPU: the criterion against its definition, the
embedding against hand values, and mp equivalence (the n = 2 simulated decoder stack gives the n = 1 loss)."""
import math

import torch

from oracle import gpt_model as G


def test_pretraining_loss_is_the_masked_mean_of_token_cross_entropies():
    g = torch.Generator().manual_seed(0)
    logits = torch.randn(2, 5, 11, generator=g, dtype=torch.float64)
    labels = torch.randint(0, 11, (2, 5), generator=g)
    mask = torch.tensor([[1., 1, 0, 1, 0], [0, 1, 1, 1, 1]], dtype=torch.float64)
    want = 0.0
    for i in range(2):
        for j in range(5):
            row = logits[i, j]
            want += mask[i, j].item() * (math.log(sum(math.exp(v) for v in row.tolist())) - row[labels[i, j]].item())
    want /= mask.sum().item()
    assert abs(G.pretraining_loss(logits, labels, mask).item() - want) < 1e-12


def test_model_oracle_embedding_and_mp_equivalence():
    cfg1 = G.ModelConfig(vocab_size=32, hidden_size=16, num_layers=2, num_attention_heads=4, ffn_hidden_size=32,
                         max_position_embeddings=8, hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
    params = G.init_params(cfg1, 5, dtype=torch.float64)
    ids = torch.tensor([[1, 2, 3, 1], [0, 31, 7, 7]])
    labels = torch.tensor([[2, 3, 1, 0], [31, 7, 7, 1]])
    mask = torch.ones(2, 4, dtype=torch.float64)
    # zero layers: the model is LayerNorm(wte[id] + wpe[pos]) · wte^T
    cfg0 = G.ModelConfig(**{**cfg1.__dict__, "num_layers": 0})
    x = params["gpt.embeddings.word_embeddings.weight"][ids] + params["gpt.embeddings.position_embeddings.weight"][:4]
    want = torch.nn.functional.layer_norm(x, (16,), eps=1e-5) @ params["gpt.embeddings.word_embeddings.weight"].t()
    assert torch.allclose(G.gpt_logits(ids, params, cfg0, training=False), want, atol=1e-12)
    l1, g1 = G.loss_and_grads(ids, labels, mask, params, cfg1)
    cfg2 = G.ModelConfig(**{**cfg1.__dict__, "num_partitions": 2})
    l2, g2 = G.loss_and_grads(ids, labels, mask, params, cfg2)
    assert abs(l1.item() - l2.item()) < 1e-12
    assert all(torch.allclose(g1[k], g2[k], atol=1e-10) for k in g1)
