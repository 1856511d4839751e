"""Parity of each CUDA kernel (called through the C ABI) against the oracle's op of the same name.

Tolerances: bf16 outputs carry 2^-8 relative rounding; north_star allows <= 2e-2 relative vs the
fp32 reference. Each test states the bound it uses. Masks and indices are compared bit for bit.
"""
import numpy as np
import pytest
import torch

from conftest import rel_err
from oracle import decoder_layer as O
from oracle import philox as OP

pytestmark = pytest.mark.gpu

BF16_TOL = 1e-2  # max|a-b| / max|b| for a bf16 result of fp32 math


def _dev():
    return torch.device("cuda:0")


def _randn(*shape, seed=0, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(torch.bfloat16)


# ------------------------------------------------------------------------------------ GEMM
GEMM_SHAPES = [
    (128, 128, 64), (128, 256, 64), (256, 512, 256), (384, 768, 1024), (1024, 3072, 1024),
    (136, 48, 72),      # ragged: every dim only a multiple of 8
    (8, 8, 8),          # smallest legal problem
    (2048, 2048, 2048),
]


@pytest.mark.parametrize("trans_a,trans_b", [(False, False), (False, True), (True, False), (True, True)])
@pytest.mark.parametrize("M,N,K", GEMM_SHAPES)
def test_gemm_layouts(M, N, K, trans_a, trans_b):
    from fleetx_b200 import ops

    a = _randn(M, K, seed=1)
    b = _randn(K, N, seed=2)
    ref = a.float() @ b.float()
    a_store = a.t().contiguous() if trans_a else a
    b_store = b.t().contiguous() if trans_b else b
    c = ops.gemm(a_store.to(_dev()), b_store.to(_dev()), trans_a=trans_a, trans_b=trans_b)
    torch.cuda.synchronize()
    assert c.shape == (M, N)
    assert rel_err(c, ref) < BF16_TOL
    # fp32 output path: only the bf16 inputs are rounded, so it must be much tighter
    c32 = ops.gemm(a_store.to(_dev()), b_store.to(_dev()), trans_a=trans_a, trans_b=trans_b,
                   out_dtype=torch.float32)
    assert rel_err(c32, ref) < 2e-4


def test_gemm_bias_gelu_dgelu_accumulate():
    from fleetx_b200 import ops

    M, N, K = 256, 384, 320
    a, w, bias = _randn(M, K, seed=3), _randn(K, N, seed=4, scale=0.05), _randn(N, seed=5)
    ad, wd, bd = a.to(_dev()), w.to(_dev()), bias.to(_dev())
    z_ref = O.column_parallel_linear(a.float(), w.float(), bias.float())

    y = ops.gemm(ad, wd, epilogue=ops.FX_EPI_BIAS, bias=bd)
    assert rel_err(y, z_ref) < BF16_TOL

    z = torch.empty(M, N, dtype=torch.bfloat16, device=_dev())
    act = ops.gemm(ad, wd, epilogue=ops.FX_EPI_BIAS_GELU, bias=bd, aux=z)
    assert rel_err(z, z_ref) < BF16_TOL
    assert torch.equal(z, y)  # same accumulate, same rounding
    assert rel_err(act, O.gelu_tanh(z.float().cpu())) < BF16_TOL

    # dgelu: C = (dA · W2^T) * gelu'(Z) — check the epilogue factor against autograd of the oracle
    da = _randn(M, N, seed=6)
    zz = z.float().cpu().requires_grad_(True)
    O.gelu_tanh(zz).backward(da.float())
    ident = torch.eye(N, dtype=torch.bfloat16)
    dz = ops.gemm(da.to(_dev()), ident.to(_dev()), epilogue=ops.FX_EPI_DGELU, aux=z)
    assert rel_err(dz, zz.grad) < BF16_TOL

    # accumulate into fp32 and bf16 outputs
    base32 = torch.randn(M, N)
    c32 = base32.clone().to(_dev())
    ops.gemm(ad, wd, out=c32, accumulate=True)
    assert rel_err(c32, base32 + a.float() @ w.float()) < 1e-4
    base16 = _randn(M, N, seed=7)
    c16 = base16.clone().to(_dev())
    ops.gemm(ad, wd, out=c16, accumulate=True)
    assert rel_err(c16, base16.float() + a.float() @ w.float()) < BF16_TOL


@pytest.mark.parametrize("M", [1, 3, 12, 130])
def test_gemm_any_row_count(M):
    """Decoding runs the linears on a handful of token rows: M need not be a multiple of 8 (nor of the tile)."""
    from fleetx_b200 import ops

    g = torch.Generator().manual_seed(M)
    a = torch.randn(M, 256, generator=g).to(torch.bfloat16)
    w = (torch.randn(256, 512, generator=g) * 0.05).to(torch.bfloat16)
    bias = (torch.randn(512, generator=g) * 0.1).to(torch.bfloat16)
    out = ops.gemm(a.to(_dev()), w.to(_dev()), epilogue=ops.FX_EPI_BIAS, bias=bias.to(_dev()))
    assert rel_err(out, a.float() @ w.float() + bias.float()) < BF16_TOL
    out_t = ops.gemm(a.to(_dev()), w.t().contiguous().to(_dev()), trans_b=True)
    assert rel_err(out_t, a.float() @ w.float()) < BF16_TOL


@pytest.mark.parametrize("M,N,K,splits", [(256, 2048, 16384, 0), (2048, 768, 16384, 0), (264, 520, 4096 + 64, 3),
                                          (128, 256, 1024, 16), (2048, 1024, 16384, 0)])
def test_gemm_split_k(M, N, K, splits):
    """Weight-gradient shapes of a high mp degree (few output tiles, long K): K cut into ranges, fp32 partial tiles
    summed in a fixed order. splits = 0: the library's own proposal for the shape (must be > 1 for these). Also a
    ragged case (M, N not multiples of the tile, K not a multiple of splits * 64) and more splits than k-blocks / 4."""
    from fleetx_b200 import _lib, ops

    lib = _lib.lib()
    a = _randn(K, M, seed=21)       # wgrad layout: both operands stored [K, *]
    b = _randn(K, N, seed=22)
    ref = a.float().t() @ b.float()
    ad, bd = a.to(_dev()), b.to(_dev())
    if splits == 0:
        assert lib.fx_gemm_splitk_splits(M, N, K) > 1
        c = ops.gemm(ad, bd, trans_a=True, split_k=True)
        c32 = ops.gemm(ad, bd, trans_a=True, split_k=True, out_dtype=torch.float32)
    else:
        ws = torch.empty(lib.fx_gemm_splitk_workspace_bytes(M, N, splits), dtype=torch.uint8, device=_dev())
        st = torch.cuda.current_stream().cuda_stream
        c = torch.empty(M, N, dtype=torch.bfloat16, device=_dev())
        c32 = torch.empty(M, N, dtype=torch.float32, device=_dev())
        for out, dt in ((c, _lib.FX_BF16), (c32, _lib.FX_F32)):
            _lib.check(lib.fx_gemm_bf16_splitk(st, 1, 0, M, N, K, ad.data_ptr(), M, bd.data_ptr(), N, out.data_ptr(), dt,
                                               splits, ws.data_ptr()), "fx_gemm_bf16_splitk")
    torch.cuda.synchronize()
    assert rel_err(c, ref) < BF16_TOL
    assert rel_err(c32, ref) < 2e-4
    # deterministic: the same call again gives the same bits
    c2 = ops.gemm(ad, bd, trans_a=True, split_k=True) if splits == 0 else None
    if c2 is not None:
        assert torch.equal(c, c2)
    # and the plain kernel agrees to bf16 rounding
    assert rel_err(ops.gemm(ad, bd, trans_a=True), ref) < BF16_TOL


def test_gemm_under_an_sm_budget():
    """fx_set_sm_budget: the persistent grid shrinks, the result does not change."""
    from fleetx_b200 import ops

    a, b = _randn(4096, 1024, seed=31).to(_dev()), _randn(1024, 2048, seed=32).to(_dev())
    full = ops.gemm(a, b)
    with ops.sm_budget(100):
        part = ops.gemm(a, b)
    assert torch.equal(full, part)
    assert torch.equal(ops.gemm(a, b), full)  # and the budget is gone again


def test_gemm_rejects_bad_arguments():
    from fleetx_b200 import FleetXError, ops

    a = torch.zeros(16, 16, dtype=torch.bfloat16, device=_dev())
    b = torch.zeros(16, 12, dtype=torch.bfloat16, device=_dev())  # N = 12 is not a multiple of 8
    with pytest.raises(FleetXError):
        ops.gemm(a, b)
    with pytest.raises(FleetXError):  # A stored [K, M]: M is then a contiguous dimension and must be one too
        ops.gemm(torch.zeros(16, 12, dtype=torch.bfloat16, device=_dev()), torch.zeros(16, 16, dtype=torch.bfloat16, device=_dev()),
                 trans_a=True)
    with pytest.raises(FleetXError):
        ops.gemm(torch.zeros(16, 16, dtype=torch.bfloat16), torch.zeros(16, 16, dtype=torch.bfloat16))  # CPU tensors


# ------------------------------------------------------------------------------------ LayerNorm
@pytest.mark.parametrize("rows,h", [(16, 128), (256, 1024), (1000, 2048), (64, 4096), (48, 5120), (7, 8), (5, 16384)])
def test_layernorm_fwd_bwd(rows, h):
    from fleetx_b200 import ops

    x = _randn(rows, h, seed=10) * 2 + 0.5
    x = x.to(torch.bfloat16)
    g = torch.Generator().manual_seed(11)
    gamma = 1 + 0.1 * torch.randn(h, generator=g)
    beta = 0.1 * torch.randn(h, generator=g)
    dy = _randn(rows, h, seed=12)
    dres = _randn(rows, h, seed=13)

    xr = x.float().requires_grad_(True)
    gr, br = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    y_ref = O.layer_norm(xr, gr, br)
    y_ref.backward(dy.float())

    y, mean, rstd = ops.layernorm_fwd(x.to(_dev()), gamma.to(_dev()), beta.to(_dev()))
    assert rel_err(y, y_ref) < BF16_TOL
    assert rel_err(mean, x.float().mean(-1)) < 1e-5
    dx, dgamma, dbeta = ops.layernorm_bwd(dy.to(_dev()), x.to(_dev()), gamma.to(_dev()), mean, rstd)
    assert rel_err(dx, xr.grad) < BF16_TOL
    assert rel_err(dgamma, gr.grad) < 1e-4
    assert rel_err(dbeta, br.grad) < 1e-4
    dx2, _, _ = ops.layernorm_bwd(dy.to(_dev()), x.to(_dev()), gamma.to(_dev()), mean, rstd, dresidual=dres.to(_dev()))
    assert rel_err(dx2, xr.grad + dres.float()) < BF16_TOL


# ------------------------------------------------------------------------------------ dropout
@pytest.mark.parametrize("rows,h,p", [(64, 256, 0.1), (33, 1024, 0.5), (128, 2048, 0.0), (9, 8, 0.25)])
def test_bias_dropout_residual(rows, h, p):
    from fleetx_b200 import ops

    y, bias, res = _randn(rows, h, seed=20), _randn(h, seed=21), _randn(rows, h, seed=22)
    seed, offset = 1024, 3
    out, mask = ops.bias_dropout_residual_fwd(y.to(_dev()), bias.to(_dev()), res.to(_dev()), p, seed, offset,
                                              return_mask=True)
    keep = OP.dropout_mask(seed, offset, rows * h, p).reshape(rows, h)
    assert np.array_equal(mask.cpu().numpy(), keep)  # generator parity, bit for bit
    ref = res.float() + O.dropout(y.float() + bias.float(), p, torch.from_numpy(keep))
    assert rel_err(out, ref) < BF16_TOL

    dout = _randn(rows, h, seed=23)
    dy = ops.bias_dropout_residual_bwd(dout.to(_dev()), p, seed, offset)
    assert rel_err(dy, O.dropout(dout.float(), p, torch.from_numpy(keep))) < BF16_TOL

    # injected mask overrides the generator (parity mode)
    inj = (torch.rand(rows, h, generator=torch.Generator().manual_seed(24)) > 0.3).to(torch.uint8)
    out2 = ops.bias_dropout_residual_fwd(y.to(_dev()), None, res.to(_dev()), 0.3, seed, offset, mask_in=inj.to(_dev()))
    assert rel_err(out2, res.float() + O.dropout(y.float(), 0.3, inj)) < BF16_TOL


def test_dropout_statistics_and_streams():
    from fleetx_b200 import ops

    rows, h, p = 512, 2048, 0.1
    z = torch.zeros(rows, h, dtype=torch.bfloat16, device=_dev())
    _, m0 = ops.bias_dropout_residual_fwd(z, None, z, p, 1024, 0, return_mask=True)
    _, m0b = ops.bias_dropout_residual_fwd(z, None, z, p, 1024, 0, return_mask=True)
    _, m1 = ops.bias_dropout_residual_fwd(z, None, z, p, 1024, 1, return_mask=True)
    _, m2 = ops.bias_dropout_residual_fwd(z, None, z, p, 1024 + 123 + 10, 0, return_mask=True)
    n = rows * h
    keep_rate = m0.float().mean().item()
    sigma = (p * (1 - p) / n) ** 0.5
    assert abs(keep_rate - (1 - p)) < 4 * sigma + 1.0 / 32768  # p is quantised to 1/32768
    assert torch.equal(m0, m0b)  # same (seed, offset) -> same mask: recompute-safe
    assert not torch.equal(m0, m1) and not torch.equal(m0, m2)
    agree = (m0 == m1).float().mean().item()
    assert abs(agree - (p * p + (1 - p) ** 2)) < 0.01  # independent streams


@pytest.mark.parametrize("rows,n", [(1024, 2048), (100, 48), (8, 8), (4096, 6144)])
def test_colsum(rows, n):
    from fleetx_b200 import ops

    x = _randn(rows, n, seed=30)
    ref = x.float().sum(0)
    assert rel_err(ops.colsum(x.to(_dev()), out_dtype=torch.float32), ref) < 1e-5
    assert rel_err(ops.colsum(x.to(_dev())), ref) < BF16_TOL
