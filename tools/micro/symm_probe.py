"""Probe: does torch symmetric memory (P2P buffers + NVSwitch multicast) come up on this box?
torchrun --nproc-per-node N tools/micro/symm_probe.py"""
import os
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
try:
    t = symm_mem.empty(1 << 20, dtype=torch.bfloat16, device=f"cuda:{rank}")
    hdl = symm_mem.rendezvous(t, dist.group.WORLD.group_name)
    print(rank, "world", hdl.world_size, "buffer_ptrs", [hex(p) for p in hdl.buffer_ptrs], "mc", hex(hdl.multicast_ptr),
          "signal", [hex(p) for p in hdl.signal_pad_ptrs], "pad_size", hdl.signal_pad_size, flush=True)
    t.fill_(rank + 1)
    hdl.barrier()
    peer = hdl.get_buffer((rank + 1) % world, (16,), torch.bfloat16)
    print(rank, "peer value", peer[0].item(), flush=True)
    hdl.barrier()
except Exception as e:  # noqa: BLE001
    print(rank, "FAILED", repr(e)[:500], flush=True)
dist.destroy_process_group()
