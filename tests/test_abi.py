"""The C-ABI shared library loads on a CPU-only box and exports exactly what include/*.h declares.
No compute is launched here: only argument validation and host-side helpers are called."""
import ctypes as C
import os
import re

import pytest

from fleetx_b200 import _lib


def test_library_exports_every_declared_symbol():
    declared = _lib.declared_symbols()
    assert len(declared) >= 14
    assert sorted(_lib.SIGNATURES) == declared, "ctypes table and header disagree"
    lib = _lib.lib()
    for name in declared:
        assert hasattr(lib, name), name


def test_header_cites_reference_lines_for_every_entry_point():
    text = open(_lib.HEADER_PATH).read()
    assert "hybrid_model.py" in text
    assert not re.search(r"\btorch\b|at::Tensor", text), "no torch types in the ABI"


def test_host_side_helpers_and_argument_validation():
    lib = _lib.lib()
    assert lib.fx_version() >= 100
    assert lib.fx_layernorm_bwd_workspace_bytes(16384, 2048) == 592 * 2 * 2048 * 4
    assert lib.fx_layernorm_bwd_workspace_bytes(10, 128) == 10 * 2 * 128 * 4
    assert lib.fx_layernorm_bwd_workspace_bytes(0, 128) == 0
    # head_dim 64: -lse2, -delta' and the fp32 dQ accumulator [b, nh, s_pad, 64]; head_dim 128: the two rows only
    assert lib.fx_attn_causal_bwd_workspace_bytes(2, 100, 4, 64) == (2 + 64) * (2 * 4 * 128 * 4)
    assert lib.fx_attn_causal_bwd_workspace_bytes(2, 100, 4, 128) == 2 * (2 * 4 * 128 * 4)
    assert lib.fx_colsum_workspace_bytes(0, 8) == 0
    assert lib.fx_attn_dropout_mask_bytes(2, 100, 3) == 2 * 3 * 100 * 4 * 4  # 1 bit per (row, col), cols padded to 128
    # bad arguments are rejected before any CUDA call, with a message
    rc = lib.fx_gemm_bf16(None, 0, 0, 16, 12, 16, None, 16, None, 16, None, 16, 0, 0, 0, None, None, 0)  # N = 12
    assert rc == -1 and b"multiples of 8" in lib.fx_last_error()
    rc = lib.fx_attn_causal_fwd(None, None, None, None, None, 1, 64, 1, 32, 1.0, 0.0, 0, 0)
    assert rc == -1 and b"head_dim" in lib.fx_last_error()
    rc = lib.fx_layernorm_fwd(None, None, None, None, None, None, None, 4, 12, 1e-5)
    assert rc == -1
    rc = lib.fx_bias_dropout_residual_fwd(None, None, None, None, None, 4, 16, 1.5, 0, 0, None, None)
    assert rc == -1 and b"dropout p" in lib.fx_last_error()


def test_no_cpu_fallback():
    import torch

    from fleetx_b200 import FleetXError, ops
    from fleetx_b200.layers import TransformerDecoderLayer

    with pytest.raises(FleetXError):
        ops.layernorm_fwd(torch.zeros(4, 16, dtype=torch.bfloat16), torch.ones(16), torch.zeros(16))
    layer = TransformerDecoderLayer(128, 2, 512, device="cpu")
    with pytest.raises(RuntimeError):
        layer(torch.zeros(1, 8, 128, dtype=torch.bfloat16))


def test_product_code_never_imports_the_oracle():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    pkg = os.path.join(root, "fleetx_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):  # no exceptions: the checker (selfcheck.py) lives outside the package
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, flags=re.M), f


@pytest.mark.parametrize("fill", [0, 8])
def test_every_compute_entry_point_rejects_null_and_empty_arguments(fill):
    """All pointers NULL, every integer `fill`: each entry point must return a negative code with a message (and
    never touch the pointers); the size helpers return a number. No CUDA call is reached."""
    lib = _lib.lib()
    # host-side helpers that compute nothing on the device: version / counters, a split proposal, a size limit,
    # the CTA budget of the calling thread
    helpers = {"fx_last_error", "fx_version", "fx_launch_count", "fx_gemm_splitk_splits", "fx_tp_rows_max_h",
               "fx_set_sm_budget"}
    for name, (res, args) in sorted(_lib.SIGNATURES.items()):
        if name in helpers:
            continue
        vals = []
        for a in args:
            if a in (C.c_void_p, C.c_char_p) or (hasattr(a, "_type_") and not isinstance(a._type_, str)):
                vals.append(None)
            elif a in (C.c_float, C.c_double):
                vals.append(0.0)
            else:
                vals.append(fill)
        rc = getattr(lib, name)(*vals)
        if name.endswith("_bytes"):
            assert rc >= 0, name
        else:
            assert rc in (-1, -3), (name, rc)
            # fx_attn_causal_fwd_at shares its checks (and messages) with fx_attn_causal_fwd
            who = name[:-3] if name.endswith("_at") else name
            assert who.encode() in lib.fx_last_error(), (name, lib.fx_last_error())


def test_whole_layer_buffer_plans_on_the_host():
    """fx_decoder_layer_saved_bytes / _workspace_bytes are pure host arithmetic: the activation memory the native
    entry points ask the caller for, at the bench shape (GPT-1.3B layer, b 8, s 2048) and its 8-way shard."""
    lib = _lib.lib()
    up = lambda n: (n + 255) // 256 * 256  # noqa: E731

    def desc(b, s, h, nh_local, hd, f_local, p_attn, pre=1):
        d = _lib.LayerDesc()
        d.b, d.s, d.h, d.nh_local, d.hd, d.f_local, d.normalize_before, d.p_attn = b, s, h, nh_local, hd, f_local, pre, p_attn
        for name, typ in _lib.LayerDesc._fields_:  # the plans need a complete descriptor; nothing is dereferenced
            if name.endswith(("_w", "_b")):
                setattr(d, name, C.cast(0x1000, typ))
        return d

    for n in (1, 8):
        b, s, h, nhl, hd, fl = 8, 2048, 2048, 32 // n, 64, 8192 // n
        T, hl = b * s, nhl * hd
        d = desc(b, s, h, nhl, hd, fl, 0.1)
        mask = lib.fx_attn_dropout_mask_bytes(b, s, nhl)
        want = (3 * up(T * h * 2) + 4 * up(T * 4) + up(T * 3 * hl * 2) + up(T * hl * 2) + up(b * nhl * s * 4) + up(mask)
                + 2 * up(T * fl * 2))  # ln1, x1, ln2 | 4 statistics rows | qkv | ctx | lse | keep bits | z, act
        assert lib.fx_decoder_layer_saved_bytes(C.byref(d)) == want
        d0 = desc(b, s, h, nhl, hd, fl, 0.0)
        assert lib.fx_decoder_layer_saved_bytes(C.byref(d0)) == want - up(mask)  # no keep bits without dropout
        ws = lib.fx_decoder_layer_workspace_bytes(C.byref(d))
        assert ws >= 5 * up(T * h * 2) + up(T * fl * 2) + up(T * hl * 2) + up(T * 3 * hl * 2) and ws % 256 == 0
    post = desc(8, 2048, 2048, 32, 64, 8192, 0.0, pre=0)  # post-LN keeps r1, r2 and a third pair of statistics
    pre = desc(8, 2048, 2048, 32, 64, 8192, 0.0, pre=1)
    assert lib.fx_decoder_layer_saved_bytes(C.byref(post)) == lib.fx_decoder_layer_saved_bytes(C.byref(pre))
    bad = desc(8, 2048, 2048, 32, 96, 8192, 0.0)
    assert lib.fx_decoder_layer_saved_bytes(C.byref(bad)) == 0 and b"head_dim" in lib.fx_last_error()
