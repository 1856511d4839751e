"""GPTForPretrainingHybrid + GPTPretrainingCriterionHybird (fleetx_b200.model over the C ABI: embedding,
decoder stack, final LayerNorm, tied logits GEMM, masked softmax cross entropy) against the oracle model
(oracle/gpt_model.py restating hybrid_model.py:574-829). Tolerance: 2e-2 relative per tensor for bf16 compute
against the fp32 oracle on the same bf16-rounded parameters; the loss within 2e-3 relative."""
import pytest
import torch

from conftest import rel_err
from oracle import gpt_model as G

pytestmark = pytest.mark.gpu


def _build(cfg, params, train=True):
    from fleetx_b200.model import GPTForPretrainingHybrid, GPTModelHybrid, GPTPretrainingCriterionHybird

    gpt = GPTModelHybrid(vocab_size=cfg.vocab_size, hidden_size=cfg.hidden_size, num_layers=cfg.num_layers,
                         num_attention_heads=cfg.num_attention_heads, ffn_hidden_size=cfg.ffn_hidden_size,
                         hidden_dropout_prob=cfg.hidden_dropout_prob,
                         attention_probs_dropout_prob=cfg.attention_probs_dropout_prob,
                         max_position_embeddings=cfg.max_position_embeddings)
    model = GPTForPretrainingHybrid(gpt)
    own = dict(model.named_parameters())
    assert set(own) == set(params), set(own) ^ set(params)  # the reference's parameter names
    with torch.no_grad():
        for k, v in params.items():
            own[k].copy_(v.to(own[k].dtype))
    model.train(train)
    return model, GPTPretrainingCriterionHybird()


@pytest.mark.parametrize("p_drop", [0.0, 0.1])
@pytest.mark.parametrize("hd", [64, 128])
def test_pretraining_loss_and_gradients_match_oracle(p_drop, hd):
    from fleetx_b200.rng import set_seed

    cfg = G.ModelConfig(vocab_size=512, hidden_size=2 * hd, num_layers=2, num_attention_heads=2, ffn_hidden_size=8 * hd,
                        max_position_embeddings=64, hidden_dropout_prob=p_drop, attention_probs_dropout_prob=p_drop)
    params = {k: v.to(torch.bfloat16).float() for k, v in G.init_params(cfg, 1024).items()}
    g = torch.Generator().manual_seed(3)
    for k in params:  # non-trivial biases / LayerNorm affine
        if k.endswith("bias"):
            params[k] = (0.02 * torch.randn(params[k].shape, generator=g)).to(torch.bfloat16).float()
        elif "norm" in k and k.endswith("weight"):
            params[k] = (1 + 0.1 * torch.randn(params[k].shape, generator=g)).float()
    b, s = 2, 32
    ids = torch.randint(0, cfg.vocab_size, (b, s), generator=g)
    ids[0, :4] = ids[1, :4] = 7  # repeated tokens: the embedding gradient must accumulate
    labels = torch.randint(0, cfg.vocab_size, (b, s), generator=g)
    loss_mask = (torch.rand(b, s, generator=g) > 0.2).float()

    set_seed(1024)
    model, crit = _build(cfg, params)
    logits = model(ids.cuda())
    loss = crit(logits, labels.cuda(), loss_mask.cuda())
    loss.backward()
    torch.cuda.synchronize()

    masks = G.make_masks(cfg, b, s, 1024) if p_drop > 0 else None
    loss_ref, g_ref = G.loss_and_grads(ids, labels, loss_mask, params, cfg, masks)
    assert abs(loss.item() - loss_ref.item()) / abs(loss_ref.item()) < 2e-3, (loss.item(), loss_ref.item())
    for name, prm in model.named_parameters():
        assert prm.grad is not None, name
        assert rel_err(prm.grad, g_ref[name]) < 2e-2, name


def test_cross_entropy_kernels_against_torch():
    """fx_cross_entropy_* on a GPT-345M-sized vocabulary (50304) against torch's fp32 cross entropy."""
    from fleetx_b200.dist import SingleGroup
    from fleetx_b200.model import MaskedCrossEntropyFunction

    T, V = 256, 50304
    g = torch.Generator().manual_seed(0)
    logits = (torch.randn(T, V, generator=g) * 3).to(torch.bfloat16)
    labels = torch.randint(0, V, (T,), generator=g)
    mask = (torch.rand(T, generator=g) > 0.3).float()
    lg = logits.cuda().requires_grad_(True)
    loss = MaskedCrossEntropyFunction.apply(lg, labels.cuda(), mask.cuda(), 0, SingleGroup())
    ref_in = logits.float().requires_grad_(True)
    tok = torch.nn.functional.cross_entropy(ref_in, labels, reduction="none")
    ref = (tok * mask).sum() / mask.sum()
    ref.backward()
    assert abs(loss.item() - ref.item()) < 1e-4 * abs(ref.item())
    (loss * 2.0).backward()  # upstream gradient 2
    assert rel_err(lg.grad, 2.0 * ref_in.grad) < 1e-2


def test_model_rejects_what_is_not_built():
    from fleetx_b200.model import GPTModelHybrid

    m = GPTModelHybrid(vocab_size=512, hidden_size=128, num_layers=1, num_attention_heads=2, ffn_hidden_size=512,
                       max_position_embeddings=64)
    with pytest.raises(NotImplementedError):
        m(torch.zeros(1, 8, dtype=torch.int64).cuda(), use_cache=True)
