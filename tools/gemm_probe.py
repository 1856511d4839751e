"""Experiment: what makes some GEMM shapes slow? (padding of leading dims, BN, GROUP)"""
import json, os, subprocess, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fleetx_b200 import ops
from tools.bench_ops import timeit

dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def run(M, N, K, pad, ta=False, tb=True):
    a_full = torch.randn((K, M + pad) if ta else (M, K + pad), device=dev).to(torch.bfloat16)
    b_full = torch.randn((N, K + pad) if tb else (K, N + pad), device=dev).to(torch.bfloat16)
    a = a_full[:, :M] if ta else a_full[:, :K]
    b = b_full[:, :K] if tb else b_full[:, :N]
    out = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
    med, _ = timeit(lambda: ops.gemm(a, b, trans_a=ta, trans_b=tb, out=out), flush=flush)
    print(json.dumps({"M": M, "N": N, "K": K, "pad": pad, "ta": ta, "tb": tb, "BN": os.environ.get("FX_GEMM_BN"), "GROUP": os.environ.get("FX_GEMM_GROUP"),
                      "ms": round(med, 4), "tflops": round(2.0 * M * N * K / med / 1e9, 1)}), flush=True)
for (M, N, K) in ((4096, 4096, 4096), (2048, 8192, 2048), (2048, 8192, 16384), (8192, 8192, 8192), (4736, 4096, 4096), (4096, 4608, 4096)):
    for pad in (0, 64):
        run(M, N, K, pad)
