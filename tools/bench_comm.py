"""torchrun --nproc-per-node N tools/bench_comm.py — fused allreduce(+bias+dropout+residual) kernel
vs NCCL allreduce + separate elementwise kernel, [16384, 2048] bf16."""
import json, os, sys
import torch
import torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fleetx_b200 import ops
from fleetx_b200.dist import SymmetricTPGroup, TorchDistGroup

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dev = torch.device("cuda", rank)
dist.init_process_group("nccl", device_id=dev)
rows, h = 16384, 2048
nccl = TorchDistGroup()
res = torch.randn(rows, h, device=dev).to(torch.bfloat16)
bias = torch.zeros(h, device=dev, dtype=torch.bfloat16)
y = torch.randn(rows, h, device=dev).to(torch.bfloat16)

def timeit(fn, iters=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()

def nccl_bdr():
    nccl.all_reduce_(y)
    ops.bias_dropout_residual_fwd(y, bias, res, 0.1, 1, 0)
t_nccl_ar = timeit(lambda: nccl.all_reduce_(y))
t_nccl = timeit(nccl_bdr)
if rank == 0: print(json.dumps({"world": world, "nccl_allreduce_ms": round(t_nccl_ar, 4), "nccl_plus_bdr_ms": round(t_nccl, 4)}), flush=True)
for nb in (64, 96, 128, 148):
    sym = SymmetricTPGroup(max_elems=rows * h, num_blocks=nb)  # the barrier counters assume a fixed grid per group
    buf = sym.partial_buffer(rows, h, dev)
    buf.copy_(y)
    t_f = timeit(lambda: sym.all_reduce_bias_dropout_residual(buf, bias, res, 0.1, 1, 0))
    t_p = timeit(lambda: sym.all_reduce(buf))
    if rank == 0: print(json.dumps({"world": world, "blocks": nb, "fused_bdr_ms": round(t_f, 4), "fused_plain_ms": round(t_p, 4)}), flush=True)
dist.destroy_process_group()
