"""ctypes binding of libfleetx_b200.so (the C ABI declared in include/fleetx_b200.h).

There is deliberately no fallback: if the shared library is missing or a declared symbol cannot
be resolved, `lib()` raises. Compute calls without a CUDA device fail inside the library.
"""
from __future__ import annotations

import ctypes as C
import os
import re
import threading

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.path.join(_ROOT, "fleetx_b200", "lib", "libfleetx_b200.so")
HEADER_PATH = os.path.join(_ROOT, "include", "fleetx_b200.h")

FX_OK = 0
FX_BF16, FX_F32 = 0, 1
FX_EPI_NONE, FX_EPI_BIAS, FX_EPI_BIAS_GELU, FX_EPI_DGELU = 0, 1, 2, 3

_vp, _i, _f, _u64, _sz = C.c_void_p, C.c_int, C.c_float, C.c_uint64, C.c_size_t

class TpComm(C.Structure):
    """fx_tp_comm (include/fleetx_b200.h)."""
    _fields_ = [("peer_bufs", C.POINTER(C.c_uint64)), ("peer_signal_pads", C.POINTER(C.c_uint64)),
                ("multicast_ptr", C.c_uint64), ("local_buf", C.c_void_p), ("rank", C.c_int), ("world", C.c_int),
                ("num_blocks", C.c_int), ("epoch", C.POINTER(C.c_uint32))]


class TpRowsDesc(C.Structure):
    """fx_tp_rows_desc (include/fleetx_b200.h)."""
    _fields_ = [("peer_signal_pads", C.POINTER(C.c_uint64)), ("multicast_in", C.c_uint64), ("multicast_out", C.c_uint64),
                ("rank", C.c_int), ("world", C.c_int), ("num_blocks", C.c_int), ("mode", C.c_int), ("epoch", C.c_uint32),
                ("rows", C.c_int), ("h", C.c_int), ("bias", C.c_void_p), ("residual", C.c_void_p), ("p", C.c_float),
                ("seed", C.c_uint64), ("offset", C.c_uint64), ("dropout_elem_offset", C.c_uint64), ("mask_in", C.c_void_p),
                ("gamma", C.c_void_p),
                ("beta", C.c_void_p), ("eps", C.c_float), ("x_own", C.c_void_p), ("x_full", C.c_void_p),
                ("mean", C.c_void_p), ("rstd", C.c_void_p), ("dres", C.c_void_p), ("dx_own", C.c_void_p),
                ("acc", C.c_void_p), ("acc_accumulate", C.c_int), ("multicast_sum", C.c_uint64), ("sum_out", C.c_void_p),
                ("sum_n", C.c_int)]


FX_ROWS_FWD_LN, FX_ROWS_FWD_OUT, FX_ROWS_BWD_LN_BDR, FX_ROWS_BWD_LN = 0, 1, 2, 3


class LayerDesc(C.Structure):
    """fx_layer_desc (include/fleetx_b200.h)."""
    _fields_ = [("b", C.c_int), ("s", C.c_int), ("h", C.c_int), ("nh_local", C.c_int), ("hd", C.c_int),
                ("f_local", C.c_int), ("normalize_before", C.c_int), ("eps", C.c_float), ("p_attn", C.c_float),
                ("p_hidden1", C.c_float), ("p_hidden2", C.c_float), ("seed_attn", C.c_uint64),
                ("offset_attn", C.c_uint64), ("seed_h1", C.c_uint64), ("offset_h1", C.c_uint64),
                ("seed_h2", C.c_uint64), ("offset_h2", C.c_uint64)] + [
        (n, C.c_void_p) for n in ("norm1_w", "norm1_b", "norm2_w", "norm2_b", "qkv_w", "qkv_b", "out_w", "out_b",
                                  "lin1_w", "lin1_b", "lin2_w", "lin2_b")] + [("tp", C.POINTER(TpComm))]


class AdamTensor(C.Structure):
    """fx_adam_tensor (include/fleetx_b200.h)."""
    _fields_ = [("param", C.c_void_p), ("master", C.c_void_p), ("grad", C.c_void_p), ("m", C.c_void_p), ("v", C.c_void_p),
                ("n", C.c_int64), ("decay", C.c_float), ("param_f32", C.c_int32), ("grad_f32", C.c_int32),
                ("is_distributed", C.c_int32)]


class LayerGrads(C.Structure):
    """fx_layer_grads (include/fleetx_b200.h)."""
    _fields_ = [(n, C.c_void_p) for n in ("norm1_w", "norm1_b", "norm2_w", "norm2_b", "qkv_w", "qkv_b", "out_w",
                                          "out_b", "lin1_w", "lin1_b", "lin2_w", "lin2_b")]


# name -> (restype, argtypes); must list every function the header declares
SIGNATURES = {
    "fx_last_error": (C.c_char_p, []),
    "fx_version": (_i, []),
    "fx_launch_count": (_u64, []),
    "fx_layernorm_fwd": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _f]),
    "fx_layernorm_bwd_workspace_bytes": (_sz, [_i, _i]),
    "fx_layernorm_bwd": (_i, [_vp] * 11 + [_i, _i]),
    "fx_layernorm_bwd_noalloc": (_i, [_vp] * 10 + [_i, _i]),
    "fx_bias_dropout_residual_fwd": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _f, _u64, _u64, _vp, _vp]),
    "fx_bias_dropout_residual_bwd": (_i, [_vp, _vp, _vp, _i, _i, _f, _u64, _u64, _vp]),
    "fx_colsum_workspace_bytes": (_sz, [_i, _i]),
    "fx_colsum_bf16": (_i, [_vp, _vp, _i, _vp, _i, _vp, _i, _i]),
    "fx_gemm_bf16": (_i, [_vp, _i, _i, _i, _i, _i, _vp, _i, _vp, _i, _vp, _i, _i, _i, _i, _vp, _vp, _i]),
    "fx_gemm_splitk_splits": (_i, [_i, _i, _i]),
    "fx_gemm_splitk_workspace_bytes": (_sz, [_i, _i, _i]),
    "fx_gemm_bf16_splitk": (_i, [_vp, _i, _i, _i, _i, _i, _vp, _i, _vp, _i, _vp, _i, _i, _vp]),
    "fx_set_sm_budget": (_i, [_i]),
    "fx_tp_allreduce_bdr": (_i, [_vp, _vp, _vp, _u64, _i, _i, C.c_uint32, _i, _vp, _i, _i, _vp, _vp, _f, _u64, _u64, _vp, _u64, _u64]),
    "fx_tp_allgather_rows": (_i, [_vp, _vp, _u64, _i, _i, C.c_uint32, _i, _vp, _i, _i, _i, _u64]),
    "fx_tp_reduce_scatter_bdr": (_i, [_vp, _vp, _u64, _i, _i, C.c_uint32, _i, _vp, _i, _i, _i, _vp, _vp, _f, _u64, _u64, _vp, _u64]),
    "fx_tp_rows_max_h": (_i, []),
    "fx_tp_rows_fused": (_i, [_vp, C.POINTER(TpRowsDesc)]),
    "fx_tp_allreduce_f32": (_i, [_vp, _vp, _u64, _i, _i, C.c_uint32, _vp, _i]),
    "fx_attn_dropout_mask_bytes": (_sz, [_i, _i, _i]),
    "fx_attn_causal_fwd": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _f, _f, _u64, _u64]),
    "fx_attn_causal_fwd_at": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _f, _f, _u64, _u64, _i, _i]),
    "fx_attn_dropout_mask": (_i, [_vp, _vp, _i, _i, _i, _f, _u64, _u64, _i]),
    "fx_attn_causal_bwd_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "fx_attn_causal_bwd": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _f, _f]),
    "fx_embedding_fwd": (_i, [_vp, _vp, _vp, _vp, C.c_int64, C.c_int64, _vp, _vp, _vp, C.c_int64, _i, _i, _f, _u64, _u64]),
    "fx_embedding_bwd": (_i, [_vp, _vp, _vp, _vp, _vp, C.c_int64, C.c_int64, _vp, C.c_int64, _i, _i, _f, _u64, _u64]),
    "fx_f32_to_bf16": (_i, [_vp, _vp, _vp, C.c_int64]),
    "fx_cross_entropy_stats": (_i, [_vp, _vp, _vp, C.c_int64, C.c_int64, _i, _vp, _vp, _vp]),
    "fx_cross_entropy_rescale": (_i, [_vp, _vp, _vp, _vp, C.c_int64]),
    "fx_cross_entropy_loss": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, C.c_int64]),
    "fx_cross_entropy_bwd": (_i, [_vp, _vp, _vp, _vp, C.c_int64, C.c_int64, _i, _vp, _vp, _vp, _vp, _vp]),
    "fx_kv_cache_append": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i]),
    "fx_attn_decode": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _f]),
    "fx_multi_sq_norm": (_i, [_vp, _vp, _vp, _i, _i, _vp]),
    "fx_multi_adamw": (_i, [_vp, _vp, _vp, _i, _i, _f, _f, _f, _f, C.c_int64, _vp, _f]),
    "fx_decoder_layer_saved_bytes": (_sz, [C.POINTER(LayerDesc)]),
    "fx_decoder_layer_workspace_bytes": (_sz, [C.POINTER(LayerDesc)]),
    "fx_decoder_layer_fwd": (_i, [_vp, C.POINTER(LayerDesc), _vp, _vp, _vp, _vp]),
    "fx_decoder_layer_bwd": (_i, [_vp, C.POINTER(LayerDesc), _vp, _vp, _vp, _vp, C.POINTER(LayerGrads), _vp]),
}


class FleetXError(RuntimeError):
    pass


def declared_symbols(header_path: str = HEADER_PATH):
    """Function names declared in the public header (used by the export test)."""
    text = open(header_path).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(fx_[a-z0-9_]+)\s*\(", text)))


_lock = threading.Lock()
_lib = None


def lib() -> C.CDLL:
    global _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise FleetXError(
                    f"{LIB_PATH} is missing: run `python __graft_entry__.py` (nvcc, sm_100a) first; "
                    "there is no CPU fallback"
                )
            handle = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
            for name, (res, args) in SIGNATURES.items():
                try:
                    fn = getattr(handle, name)
                except AttributeError as e:
                    raise FleetXError(f"libfleetx_b200.so does not export {name}") from e
                fn.restype = res
                fn.argtypes = args
            _lib = handle
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != FX_OK:
        msg = lib().fx_last_error().decode(errors="replace")
        raise FleetXError(f"{what or 'fleetx_b200'} failed (code {rc}): {msg}")


def launch_count() -> int:
    return int(lib().fx_launch_count())
