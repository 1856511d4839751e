#!/usr/bin/env python
"""bench.py — decoder-layer fwd+bwd throughput (BASELINE.json's metric) on N B200s of one node.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--impl reference]
  N > 1:  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
              --master-port P bench.py --gpus N --steps K --warmup W

Workload: the GPT-1.3B-shape TransformerDecoderLayer (h=2048, 32 heads, ffn 8192, seq 2048),
micro-batch 8 (the reference's own 1.3B micro_batch_size, pretrain_gpt_1.3B_single_card.yaml),
dropout 0.1/0.1, bf16, synthetic N(0,1) activations, N(0,0.02) weights, tensor parallel over
the N GPUs (mp_degree = N, so total work is fixed: "strong" scaling).

One "step" = one forward + one backward of the layer over the whole batch. `value` is timed with
inputs resident in HBM (CUDA events per step, L2 flushed between steps, max over ranks);
`e2e` is the same step through the public API (TransformerDecoderLayer + autograd) with HOST
buffers: every step's x and dy are copied from pinned memory and its dx is copied back inside the timed
region, on copy streams, double-buffered against the neighbouring steps' kernels (an input pipeline).

--impl reference times the CPU restatement of the reference layer (oracle/, torch-CPU fp32; the
reference's own Paddle path cannot run here — see BASELINE.md) on a bounded sample.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# BASELINE.json configs[2..4] (C3, C4, C5). c3 is the default, the layer the metric is quoted on.
CONFIGS = {
    "c3": dict(name="GPT-1.3B", h=2048, nh=32, ffn=8192, seq=2048),
    "c4": dict(name="GPT-6.7B", h=4096, nh=32, ffn=16384, seq=2048),
    "c5": dict(name="GPT-13B", h=5120, nh=40, ffn=20480, seq=4096),
}
H, NH, FFN, SEQ = 2048, 32, 8192, 2048  # set from --config in main()
DROPOUT = 0.1
SEED = 1024
METRIC = "decoder_layer_fwd_bwd_tokens_per_sec"
PARITY_TOL = 2e-2  # north_star: <= 2e-2 relative for bf16 against the fp32 reference


def select_config(name: str) -> None:
    global H, NH, FFN, SEQ
    c = CONFIGS[name]
    H, NH, FFN, SEQ = c["h"], c["nh"], c["ffn"], c["seq"]


def layer_flops(tokens: int, s: int, h: int) -> float:
    """72·T·h² + 6·T·s·h (causal attention at half of dense, bwd = 2x fwd)."""
    return 72.0 * tokens * h * h + 6.0 * tokens * s * h


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"bf16_tflops": d.get("bf16_tflops", 1590.0),
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", 1400.0),
                "hbm_gbs": d.get("hbm_gbs", 6650.0), "source": "measured"}
    return {"bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "hbm_gbs": 6650.0, "source": "fallback"}


# ------------------------------------------------------------------------------------ CPU arm
def cpu_layer_sample(batch_tokens_note: str, steps: int, warmup: int, budget_s: float):
    """Time the oracle port (torch-CPU fp32, every host core) on ONE sequence of the workload (s = SEQ, never
    shortened: sequences are independent, so tokens/s of one sequence is the per-token cost of the batch).
    If K + W such steps would not fit `budget_s`, fewer steps are timed (and the count is returned) — the
    workload itself is fixed. Returns (tokens_per_s, cores, sample_description, ms_per_step, steps_timed)."""
    import torch

    from oracle import decoder_layer as O

    # torch.distributed.run exports OMP_NUM_THREADS=1: take the cores back, this is a whole-host baseline
    cores = os.cpu_count() or 1
    try:
        cores = len(os.sched_getaffinity(0))
    except Exception:  # noqa: BLE001
        pass
    torch.set_num_threads(cores)
    cfg = O.LayerConfig(d_model=H, nhead=NH, dim_feedforward=FFN, dropout=DROPOUT)
    params = O.init_params(cfg, SEED)
    s = SEQ
    g = torch.Generator().manual_seed(SEED)
    x = torch.randn(1, s, H, generator=g)
    dy = torch.randn(1, s, H, generator=g)
    masks = O.DropoutMasks(
        attn=[(torch.rand(1, NH, s, s, generator=g) > DROPOUT).to(torch.uint8)],
        hidden1=(torch.rand(1, s, H, generator=g) > DROPOUT).to(torch.uint8),
        hidden2=(torch.rand(1, s, H, generator=g) > DROPOUT).to(torch.uint8))
    t0 = time.perf_counter()
    O.decoder_layer_fwd_bwd(x, dy, params, cfg, masks)  # first warm-up step
    first = time.perf_counter() - t0
    warm = max(warmup - 1, 0)
    if first * (steps + warm) > budget_s:
        warm = 0
        steps = max(1, min(steps, int(budget_s / first) - 1))
    for _ in range(warm):
        O.decoder_layer_fwd_bwd(x, dy, params, cfg, masks)
    ts = []
    for _ in range(steps):
        t0 = time.perf_counter()
        O.decoder_layer_fwd_bwd(x, dy, params, cfg, masks)
        ts.append(time.perf_counter() - t0)
    mean = sum(ts) / len(ts)
    note = (f"1 sequence x {s} tokens of the same layer (b=1 of the b={batch_tokens_note} batch; sequences are "
            f"independent, so this is the batch's per-token cost), torch-CPU fp32 oracle port on {torch.get_num_threads()} "
            f"threads, dropout masks pre-drawn, {steps} timed step(s)")
    return s / mean, torch.get_num_threads(), note, mean * 1e3, steps


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = max(1, args.steps)
    # bound the whole run to a few minutes whatever K/W the driver passes
    v, cores, sample, ms, steps = cpu_layer_sample(str(args.batch), steps, args.warmup, budget_s=240.0)
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": "tokens/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, args.gpus),
        "cpu_baseline": {"value": v, "unit": "tokens/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference Paddle CPU path unavailable on this box (no paddle); this is the CPU restatement "
                "of the same layer (oracle/decoder_layer.py)",
    }
    print(json.dumps(line), flush=True)


def workload_config(args, n):
    return {"workload": f"{CONFIGS[args.config]['name']} decoder layer fwd+bwd ({args.config}): h={H}, heads={NH}, "
                        f"ffn={FFN}, seq={SEQ}, micro_batch={args.batch}, dropout={DROPOUT}",
            "global_batch": args.batch, "seq_len": SEQ,
            "parallelism": f"tp{n}" + ("+sp (sequence_parallel=True: activations between the blocks are [s/n, b, h] "
                                       "shards)" if getattr(args, "sequence_parallel", 0) and n > 1 else ""),
            "l2": "flushed between timed steps (256 MiB write); working set > L2"}


# ------------------------------------------------------------------------------------ GPU arm
class ClockSampler:
    QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                                       "-lms", "20", "-i", str(index)], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:  # noqa: BLE001
            self.p = None
        self.marks = []

    def mark(self):
        """Remember how many samples exist now (start / end of the timed region)."""
        try:
            self.f.flush()
            self.marks.append(sum(1 for _ in open(self.f.name)))
        except Exception:  # noqa: BLE001
            self.marks.append(0)

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:  # noqa: BLE001
            self.p.kill()
        self.f.flush()
        rows = [r.strip().split(",") for r in open(self.f.name).read().strip().splitlines() if r.strip()]
        os.unlink(self.f.name)
        window = "warm-up + timed region"
        if len(self.marks) == 2 and self.marks[1] - self.marks[0] >= 3:
            rows = rows[self.marks[0]:self.marks[1]]
            window = "timed region"
        out["window"] = window
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            try:
                sm.append(float(r[0]))
                smax.append(float(r[1]))
            except Exception:  # noqa: BLE001
                continue
            for name, val in zip(names, r[3:7]):
                if val.strip().lower().startswith("active"):
                    reasons.add(name)
        if sm:
            sm.sort()
            out["sm_mhz"] = sm[len(sm) // 2]  # median over the window named in "window"
            out["sm_max_mhz"] = max(smax)
            out["samples"] = len(sm)
        out["reasons"] = sorted(reasons)
        return out


def parity_block(args, group, world, rank, dev, sp_selected):
    """Correctness evidence next to the throughput number (the driver's GPU test box has one GPU, so this is
    where the multi-GPU paths are checked): through the SAME group object the timed steps used, one small
    layer call (b2 s256 h512, dropout 0.1, twice) on every collective path and one sequence of the bench shape
    on the selected path, y / dx / every gradient shard against the oracle's simulation of mp degree N with
    the same (seed, offset) dropout draws. Max over ranks; a figure above PARITY_TOL fails the run."""
    import torch
    import torch.distributed as dist

    import selfcheck

    cores = os.cpu_count() or 1
    try:
        cores = len(os.sched_getaffinity(0))
    except Exception:  # noqa: BLE001
        pass
    torch.set_num_threads(max(1, cores // world))  # every rank runs the oracle on its share of the host
    t0 = time.time()
    paths = ["single"] if world == 1 else ["allreduce", "sp"]
    cases = []
    worst, worst_name = 0.0, ""
    h_small, nh_small = (512, 8) if 8 % world == 0 else (64 * world, world)
    plan = [("small b2 s256 h%d" % h_small, pth, 2, 256, h_small, nh_small, 4 * h_small, 2) for pth in paths]
    sel = "sp" if sp_selected else paths[0]
    s_big = SEQ if args.config == "c3" else min(SEQ, 1024)  # the oracle's host time stays in seconds
    plan.append((f"one bench-shape sequence b1 s{s_big} h{H} heads{NH} ffn{FFN}", sel, 1, s_big, H, NH, FFN, 1))
    for label, pth, b_, s_, h_, nh_, f_, reps in plan:
        errs = selfcheck.layer_parity(group, b_, s_, h_, nh_, f_, dropout=DROPOUT, seed=SEED,
                                      sequence_parallel=(pth == "sp"), device=dev, repeats=reps)
        name = max(errs, key=lambda k: (errs[k] if errs[k] == errs[k] else float("inf")))
        v = errs[name]
        cases.append({"case": label, "path": pth, "max_rel_err": v, "worst": name})
        if not v <= worst:  # also true for NaN
            worst, worst_name = v, f"{name} ({label}, {pth})"
    # max over ranks
    if world > 1:
        gathered = [None] * world
        dist.all_gather_object(gathered, (worst, worst_name, cases))
        for r, (w, nm, cs) in enumerate(gathered):
            if not w <= worst:
                worst, worst_name = w, f"rank {r}: {nm}"
            for mine, theirs in zip(cases, cs):
                if not theirs["max_rel_err"] <= mine["max_rel_err"]:
                    mine["max_rel_err"], mine["worst"] = theirs["max_rel_err"], theirs["worst"]
    ok = bool(worst <= PARITY_TOL)
    return {"max_rel_err": worst, "worst": worst_name, "tol": PARITY_TOL, "ok": ok, "paths": paths,
            "ranks": world, "cases": cases, "seconds": round(time.time() - t0, 1),
            "against": "oracle/decoder_layer.py (CPU fp32 restatement of hybrid_model.py:525-566, mp simulated) with "
                       "oracle/philox.py masks; per tensor max|a-b|/max|b|, max over ranks"}


class _TraceLib:
    """--trace: a stand-in for the ctypes handle that brackets every compute entry point (first argument = the
    stream) with CUDA events on that stream, so one step's kernels can be laid out on a time line in situ —
    un-serialised, with the gaps between them — on any number of ranks (ncu cannot follow a multi-rank run)."""

    def __init__(self, real):
        self._real = real
        self.rec = []

    def __getattr__(self, name):
        import torch

        fn = getattr(self._real, name)
        if (not name.startswith("fx_") or name.endswith("_bytes")
                or name in ("fx_last_error", "fx_version", "fx_launch_count", "fx_set_sm_budget", "fx_gemm_splitk_splits",
                            "fx_tp_rows_max_h")):  # host-side helpers: no stream argument
            return fn

        def wrapped(*a):
            if not a:
                return fn()
            st = torch.cuda.ExternalStream(a[0]) if a[0] else torch.cuda.default_stream()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            rc = fn(*a)
            e1.record(st)
            note = "x".join(str(int(v)) for v in a[3:6]) + (" ta" if a[1] else "") + (" tb" if a[2] else "") if name == "fx_gemm_bf16" else ""
            self.rec.append((name, note, int(a[0] or 0), e0, e1))
            return rc

        return wrapped


def trace_steps(step_fn, barrier, path, rank, n_steps=3):
    """Run n_steps traced steps and write the last one's time line (rank 0) to `path`."""
    import torch

    from fleetx_b200 import _lib

    real = _lib.lib()
    tl = _TraceLib(real)
    _lib._lib = tl
    try:
        out = None
        for i in range(n_steps):
            tl.rec.clear()
            barrier()
            e_start = torch.cuda.Event(enable_timing=True)
            e_end = torch.cuda.Event(enable_timing=True)
            e_start.record()
            step_fn()
            e_end.record()
            torch.cuda.synchronize()
            rows = [{"op": nm, "note": note, "stream": st, "start_us": round(e_start.elapsed_time(a) * 1e3, 1),
                     "dur_us": round(a.elapsed_time(b_) * 1e3, 1)} for nm, note, st, a, b_ in tl.rec]
            out = {"step_us": round(e_start.elapsed_time(e_end) * 1e3, 1), "busy_us_main": None, "ops": rows}
        main = min((r["stream"] for r in out["ops"]), default=0) if out else 0
        if rank == 0 and out is not None:
            os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
            with open(path, "w") as f:
                json.dump(out, f, indent=0)
    finally:
        _lib._lib = real


def run_gpu(args):
    import torch
    import torch.distributed as dist

    from fleetx_b200 import _lib, functional as Fn, ops
    from fleetx_b200.dist import SingleGroup, SymmetricTPGroup, TorchDistGroup
    from fleetx_b200.layers import TransformerDecoderLayer
    from fleetx_b200.rng import get_rng_state_tracker, set_seed

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("bench.py --gpus N>1 must be launched with torch.distributed.run (one rank per GPU)")
        args.gpus = world
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
        collective = "fused allreduce kernel over NVSwitch peer memory (csrc/comm.cu)"
        if args.collective == "nccl":
            group = TorchDistGroup()
            collective = "NCCL allreduce (torch.distributed)"
        else:
            # no quiet backend switch: without NVSwitch multicast / symmetric memory this raises, and
            # `--collective nccl` is the explicit way to time the torch.distributed path
            group = SymmetricTPGroup(max_elems=args.batch * SEQ * H)
    else:
        group = SingleGroup()
        collective = "none (1 GPU)"
    n = world
    b, s = args.batch, SEQ
    T = b * s
    set_seed(SEED, mp_rank=rank)

    # full parameters from one seed on every rank, then keep this rank's shard
    layer = TransformerDecoderLayer(H, NH, FFN, dropout=DROPOUT, num_partitions=n, mp_group=group, device=dev)
    gen = torch.Generator().manual_seed(SEED)
    full = {
        "self_attn.qkv_proj.weight": torch.randn(H, 3 * H, generator=gen) * 0.02,
        "self_attn.out_proj.weight": torch.randn(H, H, generator=gen) * 0.02,
        "linear1.weight": torch.randn(H, FFN, generator=gen) * 0.02,
        "linear2.weight": torch.randn(FFN, H, generator=gen) * 0.02,
    }
    layer.load_full_state(full)
    layer.train()
    sp = bool(args.sequence_parallel) and world > 1
    layer.sequence_parallel = sp
    x_host = torch.randn(b, s, H, generator=gen).to(torch.bfloat16).pin_memory()
    dy_host = torch.randn(b, s, H, generator=gen).to(torch.bfloat16).pin_memory()
    dx_host = torch.empty_like(x_host).pin_memory()
    sl = s // world
    if sp:  # this rank's [s/n, b, h] sequence shard (sequence-first, the reference's layout)
        xs_host = x_host[:, rank * sl:(rank + 1) * sl].transpose(0, 1).contiguous().pin_memory()
        dys_host = dy_host[:, rank * sl:(rank + 1) * sl].transpose(0, 1).contiguous().pin_memory()
        dxs_host = torch.empty_like(xs_host).pin_memory()
        x, dy = xs_host.to(dev), dys_host.to(dev)
    else:
        x = x_host.to(dev)
        dy = dy_host.to(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    P = Fn.LayerParams(*[p.detach() for p in layer._params()])

    def step_device():
        """The hot path with inputs resident in HBM: explicit forward + backward kernel sequence."""
        meta = layer._meta()
        if sp:
            out, saved = Fn.decoder_layer_fwd_sp(x, P, meta, group)
            dx, grads = Fn.decoder_layer_bwd_sp(dy, saved, P, meta, group)
        else:
            out, saved = Fn.decoder_layer_fwd(x, P, meta, group)
            dx, grads = Fn.decoder_layer_bwd(dy, saved, P, meta, group)
        return out, dx, grads

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # nvidia-smi takes a moment to start: launch it before the warm-up so it is sampling (every 20 ms)
    # by the time the timed region begins
    sampler = ClockSampler(local) if rank == 0 else None
    for _ in range(max(args.warmup, 3)):
        step_device()
    barrier()
    if sampler is not None:
        # wait (host only: the other ranks sit in the barrier below) until the sampler prints its first line
        t_wait = time.time()
        while time.time() - t_wait < 3.0 and os.path.getsize(sampler.f.name) == 0:
            time.sleep(0.02)
    barrier()
    for _ in range(3):  # bring clocks back up after the wait; still warm-up
        step_device()
    barrier()

    # ---- timed region: K steps, CUDA events per step on the launching stream, L2 flushed between steps
    if sampler is not None:
        sampler.mark()
    ops.profile_gemm(True)
    launches0 = _lib.launch_count()
    evs = []
    barrier()
    for _ in range(args.steps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        step_device()
        e1.record()
        evs.append((e0, e1))
    barrier()
    launches = _lib.launch_count() - launches0
    gemm_records = ops.profile_gemm(False)
    if sampler is not None:
        sampler.mark()
    clocks = sampler.stop() if sampler else None
    total_ms = sum(a.elapsed_time(c) for a, c in evs)
    if args.trace:
        trace_steps(step_device, barrier, args.trace, rank)
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    ms_per_step = total_ms / args.steps
    tokens_per_s = T * args.steps / (total_ms * 1e-3)

    # dominant kernel = the tcgen05 GEMM (all four linears, fwd + dgrad + wgrad): live event timing
    gemm_flops = sum(r[0] for r in gemm_records)
    gemm_ms = sum(r[1].elapsed_time(r[2]) for r in gemm_records)
    peaks = load_peaks()
    # A timed region of well under a second runs at boost clocks and never reaches the power-capped regime the
    # sustained cuBLAS figure was measured in: the burst figure is the denominator then, the sustained one is
    # printed next to it (frac_sustained).
    burst = total_ms < 1000.0
    peak_tf = peaks["bf16_tflops"] if burst else peaks["bf16_tflops_sustained"]
    peak_other = peaks["bf16_tflops_sustained"] if burst else peaks["bf16_tflops"]
    peak_name = "burst" if burst else "sustained"
    gemm_tf = gemm_flops / (gemm_ms * 1e-3) / 1e12 if gemm_ms > 0 else 0.0
    layer_tf = layer_flops(T, s, H) / n / (ms_per_step * 1e-3) / 1e12

    # ---- e2e: public API (module + autograd), host buffers, copies inside the timed region.
    # Every step copies ITS x and dy from pinned host memory and copies ITS dx back; as an input pipeline
    # does, the copies run on their own streams, double-buffered: H2D of step i+1 and D2H of step i-1
    # overlap the kernels of step i. The first H2D and the last D2H are inside the timed region too.
    # With N ranks the layer input is replicated: each rank copies 1/N of the rows from its host and the
    # ranks all-gather over NVLink (what a TP input pipeline does); dx is identical everywhere, so each rank
    # returns its 1/N of the rows. Whole-job host traffic per step is then the same at every N.
    h2d_s, d2h_s = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
    r0, r1 = rank * (T // world), (rank + 1) * (T // world)
    xs_buf = [torch.empty(r1 - r0, H, dtype=torch.bfloat16, device=dev) for _ in range(2)]
    dys_buf = [torch.empty(r1 - r0, H, dtype=torch.bfloat16, device=dev) for _ in range(2)]
    xh2, dyh2, dxh2 = x_host.view(T, H), dy_host.view(T, H), dx_host.view(T, H)
    in_shape = (sl, b, H) if sp else (b, s, H)
    xd_buf = [torch.empty(in_shape, dtype=torch.bfloat16, device=dev) for _ in range(2)]
    dyd_buf = [torch.empty(in_shape, dtype=torch.bfloat16, device=dev) for _ in range(2)]

    def run_e2e(k_steps):
        main = torch.cuda.current_stream()
        ready = [torch.cuda.Event() for _ in range(2)]
        done = [None, None]
        keep = []  # gradients being copied out on d2h_s stay referenced until the final sync

        def h2d(i):
            slot = i % 2
            with torch.cuda.stream(h2d_s):
                if done[slot] is not None:
                    h2d_s.wait_event(done[slot])  # the step that last used these buffers has finished
                if sp:  # the layer consumes and returns this rank's shard: nothing to gather
                    xd_buf[slot].copy_(xs_host, non_blocking=True)
                    dyd_buf[slot].copy_(dys_host, non_blocking=True)
                elif world == 1:
                    xd_buf[slot].copy_(x_host, non_blocking=True)
                    dyd_buf[slot].copy_(dy_host, non_blocking=True)
                else:
                    xs_buf[slot].copy_(xh2[r0:r1], non_blocking=True)
                    dys_buf[slot].copy_(dyh2[r0:r1], non_blocking=True)
                ready[slot].record(h2d_s)

        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(main)
        h2d_s.wait_stream(main)
        d2h_s.wait_stream(main)
        h2d(0)
        for i in range(k_steps):
            slot = i % 2
            if i + 1 < k_steps:
                h2d(i + 1)
            main.wait_event(ready[slot])
            if world > 1 and not sp:
                # The all-gather runs on the MAIN stream, never next to the layer's own collectives: their
                # blocks spin on each other across ranks and must all be resident, so an NCCL kernel holding SMs
                # on one rank while waiting for a peer whose SMs are full of spinning blocks would deadlock.
                dist.all_gather_into_tensor(xd_buf[slot].view(-1), xs_buf[slot].view(-1))
                dist.all_gather_into_tensor(dyd_buf[slot].view(-1), dys_buf[slot].view(-1))
            flush.zero_()
            xin = xd_buf[slot].detach().requires_grad_(True)
            y = layer(xin)
            y.backward(dyd_buf[slot])
            ev = torch.cuda.Event()
            ev.record(main)
            done[slot] = ev
            keep.append(xin.grad)
            with torch.cuda.stream(d2h_s):
                d2h_s.wait_event(ev)
                if sp:
                    dxs_host.copy_(xin.grad, non_blocking=True)
                else:
                    dxh2[r0:r1].copy_(xin.grad.view(T, H)[r0:r1], non_blocking=True)
            for p_ in layer.parameters():
                p_.grad = None
        main.wait_stream(d2h_s)
        main.wait_stream(h2d_s)
        e1.record(main)
        torch.cuda.synchronize(dev)
        return e0.elapsed_time(e1)

    e2e_steps = 0 if args.profile_mode else max(3, min(args.steps, 10))
    e2e_ms = 0.0
    if e2e_steps:
        run_e2e(3)
        barrier()
        e2e_ms = run_e2e(e2e_steps)
        barrier()
    t = torch.tensor([e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_tokens_per_s = T * e2e_steps / (float(t.item()) * 1e-3) if e2e_steps else None

    # ---- parity block (after both timed regions, never inside it): the same group object, small + bench shapes
    parity = None if (args.no_parity or args.profile_mode) else parity_block(args, group, world, rank, dev, sp)

    if rank == 0:
        cpu = None
        if n == 1 and not args.no_cpu_baseline and not args.profile_mode:
            v, cores, sample, _, _ = cpu_layer_sample(str(b), 2, 1, budget_s=60.0)
            cpu = {"value": v, "unit": "tokens/s", "cores": cores, "kind": "port", "sample": sample}
        # measured DRAM bytes per GEMM launch from the ncu --set full capture of THIS config at THIS N
        # (profiles/gemm_traffic.json, keyed "<config>/n<N>/b<batch>"); null when no such capture exists
        traffic = None
        tp = os.path.join(ROOT, "profiles", "gemm_traffic.json")
        if os.path.exists(tp):
            try:
                ent = json.load(open(tp)).get(f"{args.config}/n{n}/b{b}")
                traffic = ent.get("dram_bytes_per_launch_mean") if ent else None
            except Exception:  # noqa: BLE001
                traffic = None
        n_gemm = len(gemm_records)
        line = {
            "metric": METRIC, "value": tokens_per_s, "unit": "tokens/s", "n_gpus": n, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": dict(workload_config(args, n), collective=collective,
                           collective_fallbacks=getattr(group, "fallbacks", {}) or None),
            "e2e": {"value": e2e_tokens_per_s, "unit": "tokens/s", "h2d_bytes_per_step": 2 * T * H * 2,
                    "d2h_bytes_per_step": T * H * 2, "steps": e2e_steps,
                    "pipeline": "double-buffered copies on their own streams: H2D of step i+1 and D2H of step i-1 "
                                "overlap the kernels of step i; first H2D and last D2H inside the timed region"
                                + ("" if world == 1 else f"; each of the {world} ranks moves 1/{world} of the rows over PCIe and "
                                   "the replicated input is all-gathered over NVLink (byte counts are whole-job totals)")},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "tensor", "kernel": "fx::gemm::gemm_kernel (tcgen05, all linears fwd+bwd)",
                         "achieved": gemm_tf, "peak": peak_tf, "unit": "TFLOP/s", "frac": gemm_tf / peak_tf,
                         "frac_sustained" if burst else "frac_burst": gemm_tf / peak_other,
                         "traffic": traffic,
                         "peak_source": f"{peaks['source']} cuBLAS bf16, {peak_name} ({peak_tf:.1f}); the other "
                                        f"denominator is {peak_other:.1f}",
                         "launches_timed": n_gemm, "share_of_step": gemm_ms / total_ms if total_ms else None},
            "layer_roofline": {"bound": "tensor", "achieved": layer_tf, "peak": peak_tf, "unit": "TFLOP/s",
                               "frac": layer_tf / peak_tf,
                               "frac_sustained" if burst else "frac_burst": layer_tf / peak_other,
                               "flops_per_step_per_gpu": layer_flops(T, s, H) / n,
                               "frac_of_nominal_2250": layer_tf / 2250.0},
            "cpu_baseline": cpu,
            "parity": parity,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    if parity is not None and not parity["ok"]:
        raise SystemExit(f"bench.py: parity with the oracle failed: {parity}")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--config", default="c3", choices=sorted(CONFIGS),
                    help="BASELINE.json layer shape: c3 = GPT-1.3B (default, the metric's), c4 = 6.7B, c5 = 13B / seq 4096")
    ap.add_argument("--trace", default="", help="after the timed region, write a per-call time line of one step (rank 0) here")
    ap.add_argument("--no-parity", action="store_true", help="skip the oracle parity block after the timed region")
    ap.add_argument("--impl", default="fleetx_b200", choices=["fleetx_b200", "reference"])
    ap.add_argument("--collective", default="fused", choices=["fused", "nccl"])
    ap.add_argument("--sequence-parallel", type=int, default=int(os.environ.get("FX_BENCH_SP", "0")),
                    help="1: run the layer with sequence_parallel=True at N > 1 (hybrid_model.py:462)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--profile-mode", action="store_true",
                    help="device-timed steps only (no e2e / CPU legs): the command ncu wraps")
    args = ap.parse_args()
    select_config(args.config)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
