"""fx_layernorm_bwd_noalloc — the entry point for callers that are not allowed to allocate GPU
memory (a runtime that fixed all its buffers before the step, a static-graph custom op that only
gets its declared inputs and outputs, a loop where cudaMalloc's device-wide synchronisation is not
acceptable). It keeps its partial-sum scratch in a fixed array inside the library instead of taking
a workspace.

The test calls it through the C ABI twice in a row, with different inputs and shapes, and compares
every output with the regular entry point (fx_layernorm_bwd + caller workspace) on the same inputs,
and with the oracle's LayerNorm under autograd.
The second call sees whatever the first left in the shared array, so a stale-scratch bug would
show up there. Tolerances are the ones the LayerNorm tests use: bf16 dx 1e-2, fp32 dgamma/dbeta
1e-4 (the two entry points sum the per-CTA partials in a different order). Collected last on
purpose (file name), so a failure here cannot hide another test behind `-x`.
"""
import pytest
import torch

from conftest import rel_err

pytestmark = pytest.mark.gpu


def _run_both(rows, h, seed, with_residual):
    from fleetx_b200 import _lib, ops

    lib = _lib.lib()
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn(rows, h, generator=g) * 2 + 0.5).to(torch.bfloat16).to(dev)
    dy = torch.randn(rows, h, generator=g).to(torch.bfloat16).to(dev)
    dres = torch.randn(rows, h, generator=g).to(torch.bfloat16).to(dev) if with_residual else None
    gamma = (1 + 0.1 * torch.randn(h, generator=g)).to(dev)
    beta = (0.1 * torch.randn(h, generator=g)).to(dev)
    _, mean, rstd = ops.layernorm_fwd(x, gamma, beta)

    want = ops.layernorm_bwd(dy, x, gamma, mean, rstd, dresidual=dres)  # existing entry point + workspace

    dx = torch.empty_like(x)
    dgamma = torch.empty(h, dtype=torch.float32, device=dev)
    dbeta = torch.empty(h, dtype=torch.float32, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    rc = lib.fx_layernorm_bwd_noalloc(stream, dy.data_ptr(), x.data_ptr(), gamma.data_ptr(), mean.data_ptr(),
                                      rstd.data_ptr(), None if dres is None else dres.data_ptr(), dx.data_ptr(),
                                      dgamma.data_ptr(), dbeta.data_ptr(), rows, h)
    _lib.check(rc, "fx_layernorm_bwd_noalloc")
    torch.cuda.synchronize()
    # and the oracle's LayerNorm under autograd on the same bf16-rounded inputs, so the test stands on its own
    from oracle import decoder_layer as O

    xr = x.float().cpu().requires_grad_(True)
    gr = gamma.cpu().clone().requires_grad_(True)
    br = beta.cpu().clone().requires_grad_(True)
    O.layer_norm(xr, gr, br).backward(dy.float().cpu())
    dx_o = xr.grad + (dres.float().cpu() if dres is not None else 0)
    return (dx, dgamma, dbeta), want, (dx_o, gr.grad, br.grad)


def test_no_alloc_layernorm_bwd_matches_regular_entry_point_twice_in_a_row():
    # two calls back to back; the second is smaller in both dimensions, so most of the shared
    # scratch still holds the first call's partial sums when it runs
    for rows, h, seed, with_residual in ((3000, 2048, 1, True), (77, 1024, 2, False)):
        (dx, dgamma, dbeta), (dx_ref, dgamma_ref, dbeta_ref), (dx_o, dgamma_o, dbeta_o) = _run_both(rows, h, seed, with_residual)
        assert rel_err(dx, dx_ref) < 1e-2
        assert rel_err(dgamma, dgamma_ref) < 1e-4
        assert rel_err(dbeta, dbeta_ref) < 1e-4
        # against the oracle (CPU fp32 autograd of the reference's LayerNorm): the LayerNorm tests' tolerances
        assert rel_err(dx, dx_o) < 1e-2
        assert rel_err(dgamma, dgamma_o) < 1e-3
        assert rel_err(dbeta, dbeta_o) < 1e-3


def test_no_alloc_rejects_sizes_beyond_its_scratch():
    from fleetx_b200 import _lib

    lib = _lib.lib()
    z = torch.zeros(16, dtype=torch.float32, device="cuda:0")
    p = z.data_ptr()
    rc = lib.fx_layernorm_bwd_noalloc(None, p, p, p, p, p, None, p, p, p, 4, 8192)  # h > 4096
    assert rc == -1
    assert b"exceeds the built-in scratch" in lib.fx_last_error()
