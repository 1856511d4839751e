"""Layer fwd+bwd at the bench shape three ways: Python-sequenced kernels, the native whole-layer entry
points, and a CUDA-graph replay of the native calls (what launch gaps cost)."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fleetx_b200 import functional as Fn
from fleetx_b200.dist import SingleGroup
from fleetx_b200.layers import TransformerDecoderLayer
from fleetx_b200.native import NativeLayer
from fleetx_b200.rng import set_seed

H, NH, FFN, S, B = 2048, 32, 8192, 2048, 8
set_seed(1024)
layer = TransformerDecoderLayer(H, NH, FFN, dropout=0.1)
layer.train()
P = Fn.LayerParams(*[p.detach() for p in layer._params()])
M = layer._meta()
x = torch.randn(B, S, H, device="cuda").to(torch.bfloat16)
dy = torch.randn(B, S, H, device="cuda").to(torch.bfloat16)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
group = SingleGroup()

def timeit(fn, n=20):
    for _ in range(3):
        fn()
    ms = []
    for _ in range(n):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ms.append(a.elapsed_time(b))
    return sum(ms) / len(ms)

def py_step():
    out, saved = Fn.decoder_layer_fwd(x, P, M, group)
    Fn.decoder_layer_bwd(dy, saved, P, M, group)

nat = NativeLayer(B, S, P, M)
def nat_step():
    nat.forward(x)
    nat.backward(dy)

print("python-sequenced  %.3f ms" % timeit(py_step))
print("native entry      %.3f ms" % timeit(nat_step))
nat_step(); torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    nat_step()
print("native, graph     %.3f ms" % timeit(g.replay))
