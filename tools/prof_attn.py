"""Small attention run for ncu: python tools/prof_attn.py [p_drop] [hd]"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fleetx_b200 import ops
p = float(sys.argv[1]) if len(sys.argv) > 1 else 0.1
hd = int(sys.argv[2]) if len(sys.argv) > 2 else 64
b, s, nh = 4, 2048, 32
qkv = torch.randn(b, s, nh * 3 * hd, device="cuda").to(torch.bfloat16)
for _ in range(3):
    out, lse, amask = ops.attn_causal_fwd(qkv, nh, hd, hd ** -0.5, p, 1, 0)
    dq = ops.attn_causal_bwd(out, qkv, out, lse, amask, nh, hd, hd ** -0.5, p)
torch.cuda.synchronize()
print("ok")
