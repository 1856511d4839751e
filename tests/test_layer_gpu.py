"""Whole decoder layer (TransformerDecoderLayer over the C ABI) against the oracle layer
(hybrid_model.py:525-566 restated). Tolerance: north_star's 2e-2 relative (max|a-b| / max|b|),
per tensor, for bf16 compute against the fp32 oracle fed the same bf16-rounded inputs."""
import pytest
import torch

from conftest import rel_err
from oracle import decoder_layer as O

pytestmark = pytest.mark.gpu
TOL = 2e-2


def _setup(cfg, b, s, seed=1024):
    params = {k: v.to(torch.bfloat16).float() for k, v in O.init_params(cfg, seed).items()}
    # non-trivial biases / LN affine so their gradients and uses are exercised
    g = torch.Generator().manual_seed(seed + 1)
    for k in params:
        if k.endswith("bias"):
            params[k] = (0.02 * torch.randn(params[k].shape, generator=g)).to(torch.bfloat16).float()
        elif k.startswith("norm") and k.endswith("weight"):
            params[k] = (1 + 0.1 * torch.randn(params[k].shape, generator=g)).float()
    x = torch.randn(b, s, cfg.d_model, generator=g).to(torch.bfloat16)
    dy = torch.randn(b, s, cfg.d_model, generator=g).to(torch.bfloat16)
    return params, x, dy


def _run(cfg, params, x, dy, seed, train=True, inject=None):
    from fleetx_b200.rng import set_seed
    from selfcheck import layer_from_oracle_params

    set_seed(seed)
    layer = layer_from_oracle_params(cfg, params)
    layer.train(train)
    if inject is not None:
        layer._inject_mask1, layer._inject_mask2 = inject
    xd = x.cuda().requires_grad_(True)
    y = layer(xd)
    y.backward(dy.cuda())
    torch.cuda.synchronize()
    return y, xd.grad, {n: p.grad for n, p in layer.named_parameters()}


def _check(y, dx, grads, y_ref, dx_ref, g_ref):
    assert rel_err(y, y_ref) < TOL, "y"
    assert rel_err(dx, dx_ref) < TOL, "dx"
    for n, g in grads.items():
        assert rel_err(g, g_ref[n]) < TOL, n


SHAPES = [  # (b, s, h, nh, ffn)
    (2, 8, 128, 2, 512),      # tiny
    (2, 256, 256, 4, 1024),
    (1, 200, 256, 2, 1024),   # ragged s, head_dim 128
    (1, 1024, 1024, 16, 4096),  # C1: GPT-345M layer, seq 1024, micro-batch 1
]


@pytest.mark.parametrize("b,s,h,nh,f", SHAPES)
@pytest.mark.parametrize("normalize_before", [True, False])
def test_layer_no_dropout_exact_path(b, s, h, nh, f, normalize_before):
    cfg = O.LayerConfig(d_model=h, nhead=nh, dim_feedforward=f, dropout=0.0, normalize_before=normalize_before)
    params, x, dy = _setup(cfg, b, s)
    y, dx, grads = _run(cfg, params, x, dy, 1024)
    _check(y, dx, grads, *O.decoder_layer_fwd_bwd(x.float(), dy.float(), params, cfg))


@pytest.mark.parametrize("b,s,h,nh,f", SHAPES[:3])
def test_layer_with_generated_dropout(b, s, h, nh, f):
    """Dropout 0.1 everywhere: the oracle replays the library's (seed, offset) draws through its
    own restatement of the Philox generator."""
    cfg = O.LayerConfig(d_model=h, nhead=nh, dim_feedforward=f, dropout=0.1)
    params, x, dy = _setup(cfg, b, s)
    y, dx, grads = _run(cfg, params, x, dy, 1024)
    masks = O.make_masks(cfg, b, s, 1024)
    _check(y, dx, grads, *O.decoder_layer_fwd_bwd(x.float(), dy.float(), params, cfg, masks))


def test_layer_with_injected_masks():
    b, s, h, nh, f = 2, 128, 256, 4, 1024
    cfg = O.LayerConfig(d_model=h, nhead=nh, dim_feedforward=f, dropout=0.2, attn_dropout=0.0)
    params, x, dy = _setup(cfg, b, s)
    g = torch.Generator().manual_seed(5)
    m1 = (torch.rand(b, s, h, generator=g) > 0.2).to(torch.uint8)
    m2 = (torch.rand(b, s, h, generator=g) > 0.2).to(torch.uint8)
    y, dx, grads = _run(cfg, params, x, dy, 1024, inject=(m1.reshape(-1, h).cuda(), m2.reshape(-1, h).cuda()))
    masks = O.DropoutMasks(attn=None, hidden1=m1, hidden2=m2)
    _check(y, dx, grads, *O.decoder_layer_fwd_bwd(x.float(), dy.float(), params, cfg, masks))


def test_layer_eval_mode_and_recompute_determinism():
    from fleetx_b200.rng import get_rng_state_tracker, set_seed
    from selfcheck import layer_from_oracle_params

    b, s, h, nh, f = 2, 128, 256, 4, 1024
    cfg = O.LayerConfig(d_model=h, nhead=nh, dim_feedforward=f, dropout=0.1)
    params, x, dy = _setup(cfg, b, s)
    set_seed(1024)
    layer = layer_from_oracle_params(cfg, params)
    xd = x.cuda()
    layer.eval()
    with torch.no_grad():
        y_eval = layer(xd)
    y_ref = O.decoder_layer_forward(x.float(), params, cfg, training=False)
    assert rel_err(y_eval, y_ref) < TOL
    # recompute (hybrid_model.py:406-408): restoring the tracker state reproduces the masks bit for bit
    layer.train()
    state = get_rng_state_tracker().get_states_tracker()
    with torch.no_grad():
        y1 = layer(xd)
        y_other = layer(xd)
    get_rng_state_tracker().set_states_tracker(state)
    with torch.no_grad():
        y2 = layer(xd)
    assert torch.equal(y1, y2)
    assert not torch.equal(y1, y_other)


def test_layer_with_the_models_explicit_causal_mask():
    """The explicit-mask branch (hybrid_model.py:273-275) with the mask the reference's model builds in eval
    (-1e4 * triu(1), :716-726): served by the fused causal kernel, compared with the oracle's softmax(product + mask);
    BASELINE configs[0]'s shape. Any other additive mask is refused."""
    from selfcheck import layer_from_oracle_params

    b, s, h, nh, f = 1, 1024, 1024, 16, 4096
    cfg = O.LayerConfig(d_model=h, nhead=nh, dim_feedforward=f, dropout=0.0)
    params, x, dy = _setup(cfg, b, s)
    mask = torch.triu(torch.ones(s, s) * -1e4, diagonal=1)
    y_ref, dx_ref, g_ref = O.decoder_layer_fwd_bwd(x.float(), dy.float(), params, cfg, None, attn_mask=mask)
    layer = layer_from_oracle_params(cfg, params)
    layer.train()
    xd = x.cuda().requires_grad_(True)
    y = layer(xd, tgt_mask=mask.cuda())
    y.backward(dy.cuda())
    _check(y, xd.grad, {n: p.grad for n, p in layer.named_parameters()}, y_ref, dx_ref, g_ref)
    y4 = layer(x.cuda(), tgt_mask=mask.cuda()[None, None])  # [1, 1, s, s] as the model broadcasts it
    assert torch.equal(y4, y)
    bad = mask.clone()
    bad[5, 2] = -1e4  # a padding-style entry below the diagonal
    with pytest.raises(NotImplementedError):
        layer(x.cuda(), tgt_mask=bad.cuda())


def test_standalone_parallel_linears_and_attention_module():
    from fleetx_b200.layers import ColumnParallelLinear, MultiHeadAttention, RowParallelLinear

    torch.manual_seed(0)
    col = ColumnParallelLinear(256, 512, has_bias=True, gather_output=False)
    row = RowParallelLinear(512, 256, has_bias=True, input_is_parallel=True)
    with torch.no_grad():
        col.bias.normal_(0, 0.02)
        row.bias.normal_(0, 0.02)
    x = torch.randn(4, 64, 256).to(torch.bfloat16)
    xd = x.cuda().requires_grad_(True)
    y = row(col(xd))
    dy = torch.randn(4, 64, 256).to(torch.bfloat16)
    y.backward(dy.cuda())
    xr = x.float().requires_grad_(True)
    ps = [p.detach().float().cpu().requires_grad_(True) for p in (col.weight, col.bias, row.weight, row.bias)]
    mid = O.column_parallel_linear(xr, ps[0], ps[1]).to(torch.bfloat16).float()  # the bf16 hand-off between the two GEMMs
    # autograd through the rounding: use straight-through by recomputing without it for gradients
    yr = O.row_parallel_linear([O.column_parallel_linear(xr, ps[0], ps[1])], [ps[2]], ps[3])
    yr.backward(dy.float())
    assert rel_err(y, yr) < TOL and rel_err(xd.grad, xr.grad) < TOL
    for p, r in zip((col.weight, col.bias, row.weight, row.bias), ps):
        assert rel_err(p.grad, r.grad) < TOL
    del mid

    mha = MultiHeadAttention(256, 4, dropout=0.0)
    q = torch.randn(2, 128, 256).to(torch.bfloat16)
    out = mha(q.cuda())
    w = {n: p.detach().float().cpu() for n, p in mha.named_parameters()}
    mix = O.column_parallel_linear(q.float(), w["qkv_proj.weight"], w["qkv_proj.bias"])
    qq, kk, vv = O.fuse_prepare_qkv(mix, 4, 64)
    ref = O.row_parallel_linear([O.core_attn(qq, kk, vv, 64)], [w["out_proj.weight"]], w["out_proj.bias"])
    assert rel_err(out, ref) < TOL
    # use_cache (hybrid_model.py:178-184): dropout is 0 here, so the module may serve the inference path as it is;
    # a prompt through the cache gives the same output and leaves its k, v behind
    out_c, cache = mha(q.cuda(), use_cache=True)
    assert rel_err(out_c, ref) < TOL and cache.k.shape == (2, 4, 128, 64)


@pytest.mark.parametrize("granularity", ["full", "full_attn", "core_attn"])
@pytest.mark.parametrize("hd128", [False, True])
def test_layer_recompute_matches_plain(granularity, hd128):
    """use_recompute (hybrid_model.py:332-335 core_attn, :537-539 full_attn, :406-408 full): the re-run
    forward must rebuild the dropped activations exactly (same kernels, same Philox (seed, offset)), so
    outputs are bit-identical and gradients differ at most by the last bf16 bit of dQ (head_dim 64: the
    single-kernel attention backward adds dQ partials in no fixed order), and fewer bytes stay pinned."""
    from fleetx_b200 import functional as Fn
    from fleetx_b200.rng import set_seed
    from selfcheck import layer_from_oracle_params

    b, s, h, f = 2, 256, 256, 1024
    nh = 2 if hd128 else 4
    cfg = O.LayerConfig(d_model=h, nhead=nh, dim_feedforward=f, dropout=0.1)
    params, x, dy = _setup(cfg, b, s)

    def run(use, gran):
        set_seed(77)
        layer = layer_from_oracle_params(cfg, params)
        layer.use_recompute, layer.recompute_granularity = use, gran
        layer.train()
        xd = x.cuda().requires_grad_(True)
        y = layer(xd)
        pinned = Fn.saved_bytes(y.grad_fn.saved)
        y.backward(dy.cuda())
        torch.cuda.synchronize()
        return y.detach(), xd.grad, {n: p.grad for n, p in layer.named_parameters()}, pinned

    y0, dx0, g0, pinned0 = run(False, "full")
    y1, dx1, g1, pinned1 = run(True, granularity)
    assert torch.equal(y0, y1)
    tol = 0.0 if hd128 else 4e-3  # head_dim 128 runs the deterministic split backward: exact
    assert rel_err(dx1, dx0.float().cpu()) <= tol
    for n in g0:
        assert rel_err(g1[n], g0[n].float().cpu()) <= tol, n
    assert pinned1 < pinned0
    if granularity == "full":
        assert pinned1 == x.numel() * 2  # only the layer input stays


def test_layer_recompute_rejects_unknown_granularity():
    from selfcheck import layer_from_oracle_params

    cfg = O.LayerConfig(d_model=128, nhead=2, dim_feedforward=512, dropout=0.0)
    params, x, _ = _setup(cfg, 1, 8)
    layer = layer_from_oracle_params(cfg, params)
    layer.use_recompute, layer.recompute_granularity = True, "everything"
    layer.train()
    with pytest.raises(ValueError):
        layer(x.cuda())


def test_mha_core_attn_recompute():
    from fleetx_b200.layers import MultiHeadAttention
    from fleetx_b200.rng import set_seed

    x = torch.randn(2, 128, 256).to(torch.bfloat16)
    dy = torch.randn(2, 128, 256).to(torch.bfloat16)

    def run(use):
        torch.manual_seed(3)
        set_seed(5)
        m = MultiHeadAttention(256, 2, dropout=0.1, use_recompute=use, recompute_granularity="core_attn")  # hd 128
        m.train()
        xd = x.cuda().requires_grad_(True)
        y = m(xd)
        y.backward(dy.cuda())
        return y.detach(), xd.grad

    y0, d0 = run(False)
    y1, d1 = run(True)
    assert torch.equal(y0, y1) and torch.equal(d0, d1)
