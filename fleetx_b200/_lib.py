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
    "fx_tp_allreduce_bdr": (_i, [_vp, _vp, _vp, _u64, _i, _i, C.c_uint32, _i, _vp, _i, _i, _vp, _vp, _f, _u64, _u64, _vp, _u64, _u64]),
    "fx_attn_dropout_mask_bytes": (_sz, [_i, _i, _i]),
    "fx_attn_causal_fwd": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _f, _f, _u64, _u64]),
    "fx_attn_causal_bwd_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "fx_attn_causal_bwd": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _f, _f]),
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
