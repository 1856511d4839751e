"""Does a small kernel on a second stream hide under a resident persistent GEMM? (1 GPU)
Times GEMM alone, the small kernel alone, and both launched together on two streams."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fleetx_b200 import ops

dev = torch.device("cuda:0")
T, h = 16384, 2048
a = torch.randn(T, h, device=dev).to(torch.bfloat16)
w = (torch.randn(h, 3 * h, device=dev) * 0.02).to(torch.bfloat16)
bias = torch.zeros(3 * h, device=dev, dtype=torch.bfloat16)
d = torch.randn(T, h, device=dev).to(torch.bfloat16)
wide = torch.randn(T, 3 * h, device=dev).to(torch.bfloat16)
side = torch.cuda.Stream(device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

def gemm():
    ops.gemm(a, w, epilogue=ops.FX_EPI_BIAS, bias=bias)

smalls = {
    "attn_dropout_mask": lambda: ops.attn_dropout_mask(8, 2048, 32, 0.1, 1147, 0, dev),
    "bdr_bwd": lambda: ops.bias_dropout_residual_bwd(d, 0.1, 1024, 0),
    "colsum_6144": lambda: ops.colsum(wide),
    "ln_fwd": lambda: ops.layernorm_fwd(d, torch.ones(h, device=dev), torch.zeros(h, device=dev)),
}

def timeit(fn, iters=10):
    ts = []
    for i in range(iters + 2):
        flush.zero_()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        if i >= 2:
            ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]

def both(small):
    def run():
        main = torch.cuda.current_stream()
        side.wait_stream(main)
        with torch.cuda.stream(side):
            small()
        gemm()
        main.wait_stream(side)
    return run

def both_small_first_delay(small):
    def run():  # GEMM first, then the small kernel: must squeeze in beside resident GEMM CTAs
        main = torch.cuda.current_stream()
        gemm()
        ev = torch.cuda.Event(); ev.record(main)
        # side stream does NOT wait for the gemm: it only needs launch order behind it
        with torch.cuda.stream(side):
            small()
        main.wait_stream(side)
    return run

t_g = timeit(gemm)
print(json.dumps({"gemm_qkv_us": round(t_g, 1)}))
for name, fn in smalls.items():
    t_s = timeit(fn)
    t_b = timeit(both(fn))
    t_b2 = timeit(both_small_first_delay(fn))
    print(json.dumps({"small": name, "alone_us": round(t_s, 1), "small_then_gemm_us": round(t_b, 1), "gemm_then_small_us": round(t_b2, 1),
                      "serial_us": round(t_g + t_s, 1)}), flush=True)
