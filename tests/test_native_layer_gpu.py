"""fx_decoder_layer_fwd / fx_decoder_layer_bwd (one native call per direction) against the Python-sequenced
path of fleetx_b200.functional — same kernels in the same order, so results must be identical bit for bit
(except dQ-derived tensors at head_dim 64, where the single-kernel attention backward adds partials in no
fixed order) — and against the oracle layer."""
import pytest
import torch

from conftest import rel_err
from oracle import decoder_layer as O

pytestmark = pytest.mark.gpu
TOL = 2e-2


def _make(cfg, b, s, p_drop, normalize_before):
    from fleetx_b200 import functional as Fn
    from fleetx_b200.dist import SingleGroup
    from selfcheck import layer_from_oracle_params
    from test_layer_gpu import _setup

    params, x, dy = _setup(cfg, b, s)
    layer = layer_from_oracle_params(cfg, params)
    P = Fn.LayerParams(*layer._params())
    M = Fn.LayerMeta(nh_local=cfg.nhead, hd=cfg.d_model // cfg.nhead, p_attn=p_drop, p_hidden1=p_drop, p_hidden2=p_drop,
                     normalize_before=normalize_before, rng_attn=(1147, 3), rng_h1=(1024, 0), rng_h2=(1024, 8))
    return params, x, dy, P, M, SingleGroup()


@pytest.mark.parametrize("b,s,h,nh,f", [(2, 8, 128, 2, 512), (2, 256, 256, 4, 1024), (1, 200, 256, 2, 1024)])
@pytest.mark.parametrize("normalize_before", [True, False])
@pytest.mark.parametrize("p_drop", [0.0, 0.1])
def test_native_layer_equals_python_path(b, s, h, nh, f, normalize_before, p_drop):
    from fleetx_b200 import functional as Fn
    from fleetx_b200.native import NativeLayer

    cfg = O.LayerConfig(d_model=h, nhead=nh, dim_feedforward=f, dropout=p_drop, normalize_before=normalize_before)
    params, x, dy, P, M, group = _make(cfg, b, s, p_drop, normalize_before)
    xd, dyd = x.cuda(), dy.cuda()
    y_py, saved = Fn.decoder_layer_fwd(xd, P, M, group)
    dx_py, g_py = Fn.decoder_layer_bwd(dyd, saved, P, M, group)
    nat = NativeLayer(b, s, P, M)
    y_n = nat.forward(xd)
    dx_n, g_n = nat.backward(dyd)
    torch.cuda.synchronize()
    assert torch.equal(y_n, y_py)
    exact = (h // nh) == 128  # deterministic split attention backward
    for name, a, ref in [("dx", dx_n, dx_py)] + [(k, g_n[k], g_py[k]) for k in g_py]:
        if exact or name in ("lin2_w", "lin2_b", "lin1_w", "lin1_b", "norm2_w", "norm2_b", "out_w", "out_b"):
            assert torch.equal(a, ref), name  # nothing downstream of dQ
        else:
            assert rel_err(a, ref.float().cpu()) < 4e-3, name
    if p_drop == 0.0:  # and against the oracle
        y_ref, dx_ref, g_ref = O.decoder_layer_fwd_bwd(x.float(), dy.float(), params, cfg, None)
        assert rel_err(y_n, y_ref) < TOL and rel_err(dx_n, dx_ref) < TOL


def test_native_layer_rejects_bad_descriptors():
    import ctypes as C

    from fleetx_b200 import _lib

    lib = _lib.lib()
    d = _lib.LayerDesc()
    assert lib.fx_decoder_layer_saved_bytes(C.byref(d)) == 0  # empty shape
    assert b"empty shape" in lib.fx_last_error()
    d.b, d.s, d.h, d.nh_local, d.hd, d.f_local = 1, 8, 128, 2, 32, 512
    assert lib.fx_decoder_layer_workspace_bytes(C.byref(d)) == 0
    assert b"head_dim" in lib.fx_last_error()
