"""Debug: phase timing inside the attention backward (private -DFX_TRACE build of the library)."""
import ctypes, os, subprocess, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
so = os.path.join(ROOT, "tools", "micro", "libfx_trace.so")
from fleetx_b200 import _lib
_lib.LIB_PATH = so
from fleetx_b200 import ops
lib = _lib.lib()
lib.fx_debug_trace.argtypes = [ctypes.c_void_p, ctypes.c_int]
p = float(sys.argv[1]) if len(sys.argv) > 1 else 0.0
b, s, nh, hd = 8, 2048, 32, 64
qkv = torch.randn(b, s, nh * 3 * hd, device="cuda").to(torch.bfloat16)
out, lse, amask = ops.attn_causal_fwd(qkv, nh, hd, hd ** -0.5, p, 1, 0)
for _ in range(2):
    ops.attn_causal_bwd(out, qkv, out, lse, amask, nh, hd, hd ** -0.5, p)
torch.cuda.synchronize()
lib.fx_debug_trace(None, 1)
ops.attn_causal_bwd(out, qkv, out, lse, amask, nh, hd, hd ** -0.5, p)
torch.cuda.synchronize()
buf = (ctypes.c_ulonglong * 16)()
lib.fx_debug_trace(buf, 0)
tiles = b * nh * 272
names = ["maskgen+bar", "wait st_full", "ldtm+arrive", "compute", "wait pt_empty", "sts+fence+arrive"]
tot = 0
for i, n in enumerate(names):
    print(f"{n:18s} {buf[i] / tiles:9.1f} cycles/tile")
    tot += buf[i] / tiles
print("sum", tot)
for i, n in ((14, "tma: wait q_empty"), (15, "tma: issue"), (6, "ew: first wait"), (7, "ew: epilogue"), (13, "mma: first q_full"), (12, "mma: loop/other"), (8, "mma: wait q_full"), (9, "mma: issue ST/DPT"), (10, "mma: wait pt_full"), (11, "mma: issue dV/dK")):
    print(f"{n:18s} {buf[i] / tiles:9.1f} cycles/tile")
