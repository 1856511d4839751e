"""KV-cache inference path (hybrid_model.py:165-186, 243-266, 396-417, 697-705): prefill a
prompt with use_cache=True, then decode token by token; every step's output must equal the cache-free causal
forward over the whole sequence at the same positions (bf16 tolerance), which itself is checked against the
oracle elsewhere. Also the decode kernel alone against a torch fp32 statement of cached attention."""
import pytest
import torch

from conftest import rel_err

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("b,L,t,nh,hd", [(2, 0, 5, 2, 64), (2, 37, 1, 3, 64), (1, 300, 1, 2, 128), (2, 129, 4, 2, 128), (1, 1023, 1, 4, 64)])
def test_decode_attention_against_torch(b, L, t, nh, hd):
    from fleetx_b200 import ops

    g = torch.Generator().manual_seed(L + t)
    past = torch.randn(b, L, nh * 3 * hd, generator=g).to(torch.bfloat16) if L else None
    new = torch.randn(b, t, nh * 3 * hd, generator=g).to(torch.bfloat16)
    cap = 1280
    ck = torch.zeros(b, nh, cap, hd, dtype=torch.bfloat16, device="cuda")
    cv = torch.zeros_like(ck)
    if L:
        ops.kv_cache_append(past.cuda(), ck, cv, nh, hd, 0)
    ops.kv_cache_append(new.cuda(), ck, cv, nh, hd, L)
    out = ops.attn_decode(new.cuda(), ck, cv, nh, hd, L, hd ** -0.5)
    torch.cuda.synchronize()
    allt = new if past is None else torch.cat([past, new], 1)
    x = allt.float().view(b, L + t, nh, 3, hd)
    q, k, v = x[:, L:, :, 0].transpose(1, 2), x[..., 1, :].transpose(1, 2), x[..., 2, :].transpose(1, 2)  # [b, nh, *, hd]
    sc = q @ k.transpose(-1, -2) * hd ** -0.5
    causal = torch.arange(L + t)[None, :] > (L + torch.arange(t))[:, None]
    ref = (sc.masked_fill(causal, float("-inf")).softmax(-1) @ v).transpose(1, 2).reshape(b, t, nh * hd)
    assert rel_err(out, ref) < 1e-2
    # the cache rows are the k | v thirds of the packed projections
    assert torch.equal(ck[:, :, :L + t].cpu(), allt.view(b, L + t, nh, 3, hd)[..., 1, :].transpose(1, 2))


@pytest.mark.parametrize("hd", [64, 128])
def test_model_decoding_with_cache_matches_cache_free_forward(hd):
    from fleetx_b200.model import GPTForPretrainingHybrid, GPTModelHybrid

    torch.manual_seed(1)
    model = GPTForPretrainingHybrid(GPTModelHybrid(vocab_size=512, hidden_size=2 * hd, num_layers=2, num_attention_heads=2,
                                                   ffn_hidden_size=8 * hd, max_position_embeddings=256))
    model.eval()
    g = torch.Generator().manual_seed(2)
    ids = torch.randint(0, 512, (2, 80), generator=g).cuda()
    prompt = 72
    with torch.no_grad():
        full = model(ids)                                         # cache-free causal forward, [2, 80, 512]
        logits, cache = model(ids[:, :prompt], use_cache=True)    # prefill
        assert cache[0].k.shape == (2, 2, prompt, hd)
        assert rel_err(logits, full[:, :prompt].float().cpu()) < 2e-2
        for pos in range(prompt, 80):                             # decode one token at a time
            logits, cache = model(ids[:, pos:pos + 1], use_cache=True, cache=cache)
            assert cache[1].v.shape == (2, 2, pos + 1, hd)
            assert rel_err(logits[:, 0], full[:, pos].float().cpu()) < 2e-2, pos
        assert torch.equal(logits.argmax(-1)[:, 0], full[:, -1].argmax(-1))


def test_cached_layer_api():
    from fleetx_b200.layers import MultiHeadAttention, TransformerDecoderLayer

    layer = TransformerDecoderLayer(128, 2, 512)
    layer.train()
    x = torch.randn(1, 8, 128).to(torch.bfloat16).cuda()
    with pytest.raises(NotImplementedError):   # inference path: eval() first
        layer(x, use_cache=True)
    layer.eval()
    c0 = layer.gen_cache(x)
    assert isinstance(c0, MultiHeadAttention.Cache) and c0.k.shape == (1, 2, 0, 64)
    y, c1 = layer(x, use_cache=True, cache=c0)
    assert y.shape == x.shape and c1.k.shape == (1, 2, 8, 64)
    with pytest.raises(NotImplementedError):
        layer.self_attn.gen_cache(x, type=MultiHeadAttention.StaticCache)


def test_decoding_twice_from_one_cache_does_not_clobber_the_other_branch():
    """Beam / speculative decoding: two different tokens appended to the SAME earlier Cache. The reference concatenates
    into fresh tensors (hybrid_model.py:178-184), so both branches keep their own rows; here only the newest view of
    a buffer may grow in place, the second branch gets a copy."""
    from fleetx_b200.layers import TransformerDecoderLayer

    torch.manual_seed(0)
    layer = TransformerDecoderLayer(128, 2, 512)
    layer.eval()
    prompt = torch.randn(1, 8, 128).to(torch.bfloat16).cuda()
    ta = torch.randn(1, 1, 128).to(torch.bfloat16).cuda()
    tb = torch.randn(1, 1, 128).to(torch.bfloat16).cuda()
    _, c1 = layer(prompt, use_cache=True, cache=layer.gen_cache(prompt))
    ya, ca = layer(ta, use_cache=True, cache=c1)     # grows c1's buffer in place
    ka = ca.k.clone()
    yb, cb = layer(tb, use_cache=True, cache=c1)     # same parent: must not write into ca's row 8
    assert torch.equal(ca.k, ka) and not torch.equal(cb.k[:, :, 8], ca.k[:, :, 8])
    assert torch.equal(cb.k[:, :, :8], ca.k[:, :, :8]) and ca.k.shape == cb.k.shape == (1, 2, 9, 64)
    ya2, _ = layer(ta, use_cache=True, cache=c1)     # and branch a again gives branch a's answer
    assert torch.equal(ya2, ya)
