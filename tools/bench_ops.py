"""Micro-benchmarks of single kernels (CUDA events, L2 flushed between iterations).

Usage (GPU box): python tools/bench_ops.py [gemm] [ln] [bdr] [attn]
Prints one JSON line per case; roofline peaks come from MEASURED_PEAKS.json when present.
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fleetx_b200 import ops  # noqa: E402


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d["hbm_gbs"], d["bf16_tflops"], "measured"
    return 6650.0, 1590.0, "fallback"


def timeit(fn, iters=10, warmup=3, flush=None, reps=8):
    """Median ms per call. `reps` back-to-back launches per timed region hide the Python/ctypes
    launch latency (the GPU would otherwise idle between the start event and the kernel)."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) / reps)
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def main():
    which = set(sys.argv[1:]) or {"gemm", "ln", "bdr", "attn"}
    dev = torch.device("cuda:0")
    hbm, tf, src = peaks()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    if "gemm" in which:
        cases = []
        for T in (2048, 16384):
            h = 2048
            cases += [("qkv_fwd", T, 3 * h, h, False, False), ("out_fwd", T, h, h, False, False),
                      ("ffn1_fwd", T, 4 * h, h, False, False), ("ffn2_fwd", T, h, 4 * h, False, False),
                      ("ffn1_dgrad", T, h, 4 * h, False, True), ("ffn1_wgrad", h, 4 * h, T, True, False)]
        cases += [("square8k", 8192, 8192, 8192, False, True)]
        for ta in (False, True):
            for tb in (False, True):
                cases += [("lay4k", 4096, 4096, 4096, ta, tb), ("laywg", 2048, 8192, 16384, ta, tb)]
        for name, M, N, K, ta, tb in cases:
            a = torch.randn((K, M) if ta else (M, K), device=dev).to(torch.bfloat16)
            b = torch.randn((N, K) if tb else (K, N), device=dev).to(torch.bfloat16)
            out = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
            med, best = timeit(lambda: ops.gemm(a, b, trans_a=ta, trans_b=tb, out=out), flush=flush)
            med_t, best_t = timeit(lambda: torch.matmul(a.t() if ta else a, b.t() if tb else b, out=out), flush=flush)
            fl = 2.0 * M * N * K
            print(json.dumps({"op": "gemm", "case": name, "M": M, "N": N, "K": K, "ta": ta, "tb": tb,
                              "ms": round(med, 4), "tflops": round(fl / med / 1e9, 1),
                              "frac_of_" + src: round(fl / med / 1e9 / tf, 3),
                              "cublas_ms": round(med_t, 4), "cublas_tflops": round(fl / med_t / 1e9, 1)}), flush=True)
    if "ln" in which:
        for rows, h in ((16384, 2048), (2048, 2048), (16384, 4096)):
            x = torch.randn(rows, h, device=dev).to(torch.bfloat16)
            g = torch.ones(h, device=dev)
            bt = torch.zeros(h, device=dev)
            med, _ = timeit(lambda: ops.layernorm_fwd(x, g, bt), flush=flush)
            y, mean, rstd = ops.layernorm_fwd(x, g, bt)
            byt = rows * h * 4 + rows * 8
            print(json.dumps({"op": "ln_fwd", "rows": rows, "h": h, "ms": round(med, 4), "GBs": round(byt / med / 1e6, 1),
                              "frac_of_" + src: round(byt / med / 1e6 / hbm, 3)}), flush=True)
            med, _ = timeit(lambda: ops.layernorm_bwd(y, x, g, mean, rstd, dresidual=x), flush=flush)
            byt = rows * h * 8
            print(json.dumps({"op": "ln_bwd", "rows": rows, "h": h, "ms": round(med, 4), "GBs": round(byt / med / 1e6, 1),
                              "frac_of_" + src: round(byt / med / 1e6 / hbm, 3)}), flush=True)
    if "bdr" in which:
        for rows, h in ((16384, 2048), (2048, 2048)):
            x = torch.randn(rows, h, device=dev).to(torch.bfloat16)
            bias = torch.zeros(h, device=dev, dtype=torch.bfloat16)
            med, _ = timeit(lambda: ops.bias_dropout_residual_fwd(x, bias, x, 0.1, 1, 0), flush=flush)
            byt = rows * h * 6
            print(json.dumps({"op": "bdr_fwd", "rows": rows, "h": h, "ms": round(med, 4), "GBs": round(byt / med / 1e6, 1),
                              "frac_of_" + src: round(byt / med / 1e6 / hbm, 3)}), flush=True)
            med, _ = timeit(lambda: ops.colsum(x), flush=flush)
            byt = rows * h * 2
            print(json.dumps({"op": "colsum", "rows": rows, "h": h, "ms": round(med, 4), "GBs": round(byt / med / 1e6, 1),
                              "frac_of_" + src: round(byt / med / 1e6 / hbm, 3)}), flush=True)
    if "attn" in which:
        for b, s, nh, hd, pd in ((8, 2048, 32, 64, 0.1), (8, 2048, 32, 64, 0.0), (1, 2048, 32, 64, 0.1), (2, 4096, 5, 128, 0.1), (8, 2048, 4, 128, 0.1), (8, 2048, 4, 128, 0.0)):
            qkv = torch.randn(b, s, nh * 3 * hd, device=dev).to(torch.bfloat16)
            try:
                med, _ = timeit(lambda: ops.attn_causal_fwd(qkv, nh, hd, hd ** -0.5, pd, 1, 0), flush=flush, reps=2)
                fl = 2.0 * b * nh * s * s * hd  # causal half of 4·b·nh·s²·hd
                print(json.dumps({"op": "attn_fwd", "b": b, "s": s, "nh": nh, "hd": hd, "p": pd, "ms": round(med, 4),
                                  "tflops": round(fl / med / 1e9, 1)}), flush=True)
                out, lse, amask = ops.attn_causal_fwd(qkv, nh, hd, hd ** -0.5, pd, 1, 0)
                med, _ = timeit(lambda: ops.attn_causal_bwd(out, qkv, out, lse, amask, nh, hd, hd ** -0.5, pd), flush=flush, reps=2)
                print(json.dumps({"op": "attn_bwd", "b": b, "s": s, "nh": nh, "hd": hd, "p": pd, "ms": round(med, 4),
                                  "tflops": round(2.5 * fl / med / 1e9, 1)}), flush=True)
            except Exception as e:  # noqa: BLE001
                print(json.dumps({"op": "attn", "error": str(e)[:200]}), flush=True)


if __name__ == "__main__":
    main()
