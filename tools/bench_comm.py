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
sym = SymmetricTPGroup(max_elems=rows * h)
buf = sym.partial_buffer(rows, h, dev)
buf.copy_(y)
for nb in (16, 32, 64, 96, 148):
    sym.num_blocks = nb  # the rank barriers do not depend on the grid, so one group serves every block count
    t_f = timeit(lambda: sym.all_reduce_bias_dropout_residual(buf, bias, res, 0.1, 1, 0))
    t_p = timeit(lambda: sym.all_reduce(buf))
    t_ag = timeit(lambda: sym.sp_all_gather(res[: rows // world], 1, rows))
    t_rs = timeit(lambda: sym.sp_reduce_scatter(buf, 1, rows))
    if rank == 0: print(json.dumps({"world": world, "blocks": nb, "fused_bdr_ms": round(t_f, 4), "fused_plain_ms": round(t_p, 4),
                                    "allgather_ms": round(t_ag, 4), "reduce_scatter_ms": round(t_rs, 4)}), flush=True)
# row-owner kernels (csrc/comm_rows.cu): allreduce with bias + dropout + residual (+ LayerNorm) on the owner's rows
gamma = torch.ones(h, device=dev); beta = torch.zeros(h, device=dev)
r0, r1 = sym.own_rows(rows)
x_own = res[r0:r1].contiguous()
mean = torch.zeros(r1 - r0, device=dev); rstd = torch.ones(r1 - r0, device=dev)
acc = sym.acc_buffer(h, dev)
out_full = torch.empty(rows, h, dtype=torch.bfloat16, device=dev)
mean_f = torch.zeros(rows, device=dev); rstd_f = torch.ones(rows, device=dev)
for nb in (16, 24, 32, 48, 64):
    sym.rows_blocks = nb
    t_ln = timeit(lambda: sym.fwd_ln_finish(sym.fwd_ln_start(buf, bias, res, 0.1, 1, 0, None, gamma, beta, 1e-5)))
    t_out = timeit(lambda: sym.fwd_out_finish(sym.fwd_out_start(buf, bias, x_own, 0.1, 1, 0, None), out_full))
    t_b1 = timeit(lambda: sym.bwd_ln_bdr_finish(sym.bwd_ln_bdr_start(buf, x_own, mean, rstd, gamma, res, 0.1, 1, 0, None, acc)))
    t_b2 = timeit(lambda: sym.bwd_ln_finish(sym.bwd_ln_start(buf, res, mean_f, rstd_f, gamma, x_own, acc), out_full))
    if rank == 0: print(json.dumps({"world": world, "rows_blocks": nb, "fwd_ln_ms": round(t_ln, 4), "fwd_out_ms(+copy)": round(t_out, 4),
                                    "bwd_ln_bdr_ms": round(t_b1, 4), "bwd_ln_ms(+f32 allreduce +copy)": round(t_b2, 4)}), flush=True)
dist.destroy_process_group()
