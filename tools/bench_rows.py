"""Achieved HBM bandwidth of the memory-bound kernels of the rows after the layer at
GPT-345M / GPT-1.3B sizes: embedding, cross entropy, fused AdamW, KV-cache decode attention.
Algorithmic bytes per call are stated per line; peak = MEASURED_PEAKS.json HBM figure (fallback 6545.6 GB/s)."""
import json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fleetx_b200 import ops
from fleetx_b200.dist import SingleGroup
from fleetx_b200.model import EmbeddingFunction, MaskedCrossEntropyFunction
from fleetx_b200.optim import FusedAdamW

try:
    PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbps"]
except Exception:
    PEAK = 6545.6
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timeit(fn, n=10):
    for _ in range(3):
        fn()
    ms = []
    for _ in range(n):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ms.append(a.elapsed_time(b))
    return sorted(ms)[len(ms) // 2]


def report(op, ms, byt, note):
    print(json.dumps({"op": op, "ms": round(ms, 4), "GBs": round(byt / ms / 1e6, 1), "frac_of_hbm_peak": round(byt / ms / 1e6 / PEAK, 3),
                      "algorithmic_bytes": int(byt), "note": note}), flush=True)


V, H, S, B = 50304, 1024, 1024, 8
T = B * S
wte = (torch.randn(V, H, device="cuda") * 0.02).to(torch.bfloat16).requires_grad_(True)
wpe = (torch.randn(S, H, device="cuda") * 0.02).to(torch.bfloat16).requires_grad_(True)
ids = torch.randint(0, V, (B, S), device="cuda")
g = SingleGroup()
report("embedding_fwd", timeit(lambda: EmbeddingFunction.apply(wte, wpe, ids, None, 0, 0.1, 1, 0, g)), T * H * 2 * 3,
       "345M: read word row + position row, write out (bf16), per token 6h B")
logits = (torch.randn(T, V, device="cuda") * 2).to(torch.bfloat16)
labels = torch.randint(0, V, (T,), device="cuda")
mask = torch.ones(T, device="cuda")
from fleetx_b200 import _lib
lib = _lib.lib()
st = lambda: torch.cuda.current_stream().cuda_stream  # noqa: E731
stats = torch.empty(3, T, dtype=torch.float32, device="cuda")
out2 = torch.empty(2, dtype=torch.float32, device="cuda")
def ce_fwd():
    _lib.check(lib.fx_cross_entropy_stats(st(), logits.data_ptr(), labels.data_ptr(), 0, T, V, stats[0].data_ptr(),
                                          stats[1].data_ptr(), stats[2].data_ptr()))
    _lib.check(lib.fx_cross_entropy_loss(st(), stats[0].data_ptr(), stats[1].data_ptr(), stats[2].data_ptr(), mask.data_ptr(),
                                         None, out2.data_ptr(), T))
report("cross_entropy_fwd", timeit(ce_fwd), T * V * 2, "345M, batch 8: one sweep over the bf16 logits [8192, 50304]")
dl = torch.empty_like(logits)
def ce_bwd():
    _lib.check(lib.fx_cross_entropy_bwd(st(), logits.data_ptr(), dl.data_ptr(), labels.data_ptr(), 0, T, V, stats[0].data_ptr(),
                                        stats[1].data_ptr(), mask.data_ptr(), out2.data_ptr(), None))
report("cross_entropy_bwd", timeit(ce_bwd), T * V * 2 * 2, "read logits, write dlogits")

params = [("w%d.weight" % i, torch.nn.Parameter((torch.randn(4096, 1024, device="cuda") * 0.02).to(torch.bfloat16))) for i in range(80)]
opt = FusedAdamW(None, named_parameters=params, lr=1e-4)
for _, p in params:
    p.grad = torch.randn_like(p)
n = sum(p.numel() for _, p in params)
report("fused_adamw_step", timeit(opt.step), n * (2 + 2 + 12 + 14),
       "335M bf16 parameters: norm pass reads grad (2 B), update reads grad + master, m, v (14 B) and writes master, m, v, param (14 B)")

for b_, L, nh, hd in ((8, 2047, 32, 64), (32, 2047, 32, 64), (8, 4095, 40, 128)):
    cap = L + 1
    ck = torch.randn(b_, nh, cap, hd, device="cuda").to(torch.bfloat16)
    cv = torch.randn_like(ck)
    q = torch.randn(b_, 1, nh * 3 * hd, device="cuda").to(torch.bfloat16)
    report("attn_decode", timeit(lambda: ops.attn_decode(q, ck, cv, nh, hd, L, hd ** -0.5)), 2 * b_ * nh * cap * hd * 2,
           f"one new token over {cap} cached rows, batch {b_}, {nh} heads of {hd}: K and V read once")
