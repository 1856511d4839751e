"""The layer at BASELINE.json's full sizes, checked through properties that do not need a full-size oracle
run: causality, batch independence, exact homogeneity of the backward in dy, dropout statistics, and one
sequence of the batch against the oracle (batch independence makes that a full-size check).

Shapes (per rank): the bench workload / C3 (GPT-1.3B layer, h 2048, 32 heads of 64, ffn 8192, s 2048, b 8),
C4's rank shard (6.7B, n = 8: h 4096, 4 local heads of 128, 2048 local ffn columns, s 2048) and C5's
(13B, n = 8: h 5120, 5 local heads of 128 — not a power of two —, 2560 local ffn columns, s 4096)."""
import pytest
import torch

from conftest import rel_err
from oracle import decoder_layer as O

pytestmark = pytest.mark.gpu

# name, b, s, h, nh_local, hd, f_local
SHAPES = [
    ("c3_bench", 8, 2048, 2048, 32, 64, 8192),
    ("c4_rank", 1, 2048, 4096, 4, 128, 2048),
    ("c5_rank", 1, 4096, 5120, 5, 128, 2560),
]


def _params(h, nhl, hd, fl, dev="cuda"):
    from fleetx_b200 import functional as Fn

    g = torch.Generator().manual_seed(1024)
    q3 = 3 * nhl * hd

    def w(*shape):
        return (torch.randn(*shape, generator=g) * 0.02).to(torch.bfloat16).to(dev)

    ones = lambda: (1 + 0.1 * torch.randn(h, generator=g)).float().to(dev)  # noqa: E731
    small = lambda n: (0.02 * torch.randn(n, generator=g)).to(torch.bfloat16).to(dev)  # noqa: E731
    return Fn.LayerParams(norm1_w=ones(), norm1_b=(0.02 * torch.randn(h, generator=g)).float().to(dev), qkv_w=w(h, q3),
                          qkv_b=small(q3), out_w=w(nhl * hd, h), out_b=small(h), norm2_w=ones(),
                          norm2_b=(0.02 * torch.randn(h, generator=g)).float().to(dev), lin1_w=w(h, fl), lin1_b=small(fl),
                          lin2_w=w(fl, h), lin2_b=small(h))


def _meta(nhl, hd, p):
    from fleetx_b200 import functional as Fn

    return Fn.LayerMeta(nh_local=nhl, hd=hd, p_attn=p, p_hidden1=p, p_hidden2=p, rng_attn=(1147, 0), rng_h1=(1024, 0),
                        rng_h2=(1024, 1))


@pytest.mark.parametrize("name,b,s,h,nhl,hd,fl", SHAPES)
def test_fullsize_causality_batch_independence_and_homogeneity(name, b, s, h, nhl, hd, fl):
    from fleetx_b200 import functional as Fn
    from fleetx_b200.dist import SingleGroup

    P, M, grp = _params(h, nhl, hd, fl), _meta(nhl, hd, 0.0), SingleGroup()
    g = torch.Generator().manual_seed(5)
    x = torch.randn(b, s, h, generator=g).to(torch.bfloat16).cuda()
    dy = torch.randn(b, s, h, generator=g).to(torch.bfloat16).cuda()
    y, saved = Fn.decoder_layer_fwd(x, P, M, grp)
    dx, grads = Fn.decoder_layer_bwd(dy, saved, P, M, grp)

    # causality: positions < t never see positions >= t (every kernel is row-wise or causally masked: exact)
    t = s // 2 + 37
    x2 = x.clone()
    x2[:, t:] = torch.randn(b, s - t, h, generator=g).to(torch.bfloat16).cuda()
    y2, _ = Fn.decoder_layer_fwd(x2, P, M, grp)
    assert torch.equal(y2[:, :t], y[:, :t]) and not torch.equal(y2[:, t:], y[:, t:])
    del x2, y2

    # batch independence: a sequence's output does not depend on its batch neighbours (exact)
    if b > 1:
        x3 = x.clone()
        x3[1:] = torch.randn(b - 1, s, h, generator=g).to(torch.bfloat16).cuda()
        y3, _ = Fn.decoder_layer_fwd(x3, P, M, grp)
        assert torch.equal(y3[0], y[0])
        del x3, y3

    # backward is homogeneous in dy, and scaling by 2 is exact in bf16 and fp32: every gradient doubles bit for
    # bit, except what is downstream of dQ at head_dim 64 (fp32 partials added in no fixed order: last bf16 bit)
    dx2, grads2 = Fn.decoder_layer_bwd((dy.float() * 2).to(torch.bfloat16), saved, P, M, grp)
    for k in ("lin2_w", "lin2_b", "lin1_w", "lin1_b", "norm2_w", "norm2_b", "out_w", "out_b"):
        assert torch.equal(grads2[k].float(), grads[k].float() * 2), k
    for k, a, ref in [("dx", dx2, dx)] + [(k, grads2[k], grads[k]) for k in ("qkv_w", "qkv_b", "norm1_w", "norm1_b")]:
        if hd == 128:
            assert torch.equal(a.float(), ref.float() * 2), k
        else:
            assert rel_err(a, (ref.float() * 2).cpu()) < 4e-3, k


def test_fullsize_one_sequence_against_the_oracle():
    """Sequence 0 of the b = 8 bench batch (forward, p = 0) against the CPU oracle on that sequence alone."""
    from fleetx_b200 import functional as Fn
    from fleetx_b200.dist import SingleGroup

    name, b, s, h, nhl, hd, fl = SHAPES[0]
    P, M = _params(h, nhl, hd, fl), _meta(nhl, hd, 0.0)
    g = torch.Generator().manual_seed(6)
    x = torch.randn(b, s, h, generator=g).to(torch.bfloat16)
    y, _ = Fn.decoder_layer_fwd(x.cuda(), P, M, SingleGroup())
    cfg = O.LayerConfig(d_model=h, nhead=nhl, dim_feedforward=fl, dropout=0.0)
    names = dict(zip(("norm1_w", "norm1_b", "qkv_w", "qkv_b", "out_w", "out_b", "norm2_w", "norm2_b", "lin1_w",
                      "lin1_b", "lin2_w", "lin2_b"), O.PARAM_NAMES))
    params = {names[k]: getattr(P, k).float().cpu() for k in names}
    y_ref = O.decoder_layer_forward(x[:1].float(), params, cfg, training=False)
    assert rel_err(y[:1], y_ref) < 2e-2


def test_fullsize_dropout_statistics_and_replay():
    """p = 0.1 at the bench shape: the keep rate of the hidden masks is within 4 sigma, two runs with the same
    (seed, offset) pairs agree bit for bit in the forward, a different offset gives a different output."""
    from fleetx_b200 import functional as Fn, ops
    from fleetx_b200.dist import SingleGroup

    name, b, s, h, nhl, hd, fl = SHAPES[0]
    P, M = _params(h, nhl, hd, fl), _meta(nhl, hd, 0.1)
    g = torch.Generator().manual_seed(8)
    x = torch.randn(b, s, h, generator=g).to(torch.bfloat16).cuda()
    y1, _ = Fn.decoder_layer_fwd(x, P, M, SingleGroup())
    y2, _ = Fn.decoder_layer_fwd(x, P, M, SingleGroup())
    assert torch.equal(y1, y2)
    M3 = _meta(nhl, hd, 0.1)
    M3.rng_h2 = (1024, 7)
    y3, _ = Fn.decoder_layer_fwd(x, P, M3, SingleGroup())
    assert not torch.equal(y1, y3)
    _, mask = ops.bias_dropout_residual_fwd(x.view(-1, h), None, x.view(-1, h), 0.1, 1024, 0, return_mask=True)
    n = mask.numel()
    keep = mask.float().mean().item()
    p_eff = round(0.1 * 32768) / 32768  # the threshold is quantised to 15 bits
    sigma = (p_eff * (1 - p_eff) / n) ** 0.5
    assert abs(keep - (1 - p_eff)) < 4 * sigma


# One sequence per shape, forward AND backward, dropout 0.1 on all three streams, every tensor against the oracle
# with the generator's masks restated in numpy (oracle/philox.py). A rank shard of C4 / C5 is stated to the oracle
# as a layer of its own (LayerConfig.head_dim): same arithmetic as rank 0 of the mp-8 layer before the allreduce.
@pytest.mark.parametrize("name,b,s,h,nhl,hd,fl", SHAPES)
def test_fullsize_fwd_bwd_with_dropout_against_the_oracle(name, b, s, h, nhl, hd, fl):
    from fleetx_b200 import functional as Fn
    from fleetx_b200.dist import SingleGroup

    p = 0.1
    P, M = _params(h, nhl, hd, fl), _meta(nhl, hd, p)
    g = torch.Generator().manual_seed(11)
    x = torch.randn(1, s, h, generator=g).to(torch.bfloat16)
    dy = torch.randn(1, s, h, generator=g).to(torch.bfloat16)
    y, saved = Fn.decoder_layer_fwd(x.cuda(), P, M, SingleGroup())
    dx, grads = Fn.decoder_layer_bwd(dy.cuda(), saved, P, M, SingleGroup())
    torch.cuda.synchronize()

    from oracle import philox

    cfg = O.LayerConfig(d_model=h, nhead=nhl, dim_feedforward=fl, dropout=p, head_dim=hd)
    names = dict(zip(("norm1_w", "norm1_b", "qkv_w", "qkv_b", "out_w", "out_b", "norm2_w", "norm2_b", "lin1_w",
                      "lin1_b", "lin2_w", "lin2_b"), O.PARAM_NAMES))
    params = {names[k]: getattr(P, k).float().cpu() for k in names}
    masks = O.DropoutMasks(
        attn=[torch.from_numpy(philox.attn_dropout_mask(M.rng_attn[0], M.rng_attn[1], 1, nhl, s, p))],
        hidden1=torch.from_numpy(philox.dropout_mask(M.rng_h1[0], M.rng_h1[1], s * h, p)).reshape(1, s, h),
        hidden2=torch.from_numpy(philox.dropout_mask(M.rng_h2[0], M.rng_h2[1], s * h, p)).reshape(1, s, h))
    y_ref, dx_ref, g_ref = O.decoder_layer_fwd_bwd(x.float(), dy.float(), params, cfg, masks)
    errs = {"y": rel_err(y, y_ref), "dx": rel_err(dx, dx_ref)}
    for k, full in names.items():
        errs[k] = rel_err(grads[k], g_ref[full])
    assert max(errs.values()) < 2e-2, errs  # north_star: <= 2e-2 relative per tensor for bf16
