"""The layer's host code (fleetx_b200/functional.py: hand-sequenced forward, hand-derived backward, recompute,
tensor- and sequence-parallel collectives) run on CPU in fp64 over the test double of tests/emu_ops.py and held
to the REFERENCE'S OWN outputs and gradients (tests/golden/ref_layer_*.npz, made by executing hybrid_model.py)
at 1e-10. The CUDA kernels themselves are checked on the GPU (-m gpu); this file is about the sequencing."""
import json
import os
import socket

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

import emu_ops

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TOL = 1e-10
NAMES = {"norm1_w": "norm1.weight", "norm1_b": "norm1.bias", "qkv_w": "self_attn.qkv_proj.weight",
         "qkv_b": "self_attn.qkv_proj.bias", "out_w": "self_attn.out_proj.weight", "out_b": "self_attn.out_proj.bias",
         "norm2_w": "norm2.weight", "norm2_b": "norm2.bias", "lin1_w": "linear1.weight", "lin1_b": "linear1.bias",
         "lin2_w": "linear2.weight", "lin2_b": "linear2.bias"}
COL, COLBIAS, ROW = ("qkv_w", "lin1_w"), ("qkv_b", "lin1_b"), ("out_w", "lin2_w")


def load(name):
    d = np.load(os.path.join(GOLD, name + ".npz"))
    return {k: torch.from_numpy(d[k]) for k in d.files if k != "meta"}, json.loads(str(d["meta"]))


def shard(short, t, n, r):
    if short in COL:
        return t.chunk(n, dim=1)[r].contiguous()
    if short in COLBIAS or short in ROW:
        return t.chunk(n, dim=0)[r].contiguous()
    return t


def params_and_meta(d, m, rank=0):
    from fleetx_b200 import functional as Fn

    n = m["n"]
    P = Fn.LayerParams(**{k: shard(k, d["param." + v], n, rank) for k, v in NAMES.items()})
    p = m["dropout"]
    M = Fn.LayerMeta(nh_local=m["nh"] // n, hd=m["h"] // m["nh"], p_attn=p, p_hidden1=p, p_hidden2=p,
                     normalize_before=m["normalize_before"], sequence_parallel=m["sequence_parallel"])
    if p > 0:
        h = m["h"]
        M.mask_h1 = d["mask.hidden1"].reshape(-1, h).contiguous()
        M.mask_h2 = d["mask.hidden2"].reshape(-1, h).contiguous()
    return P, M


def check(got, want, what):
    err = (got.double() - want.double()).abs().max().item()
    assert err <= TOL, f"{what}: max abs error {err:.3e}"


def check_grads(g, d, m, rank=0):
    for short, long in NAMES.items():
        check(g[short], shard(short, d["grad." + long], m["n"], rank), "grad " + long)


def _cpu_group(base):
    class G(base):
        def partial_buffer(self, rows, cols, device):  # the product hands out bf16 (GPU) buffers
            return torch.empty(rows, cols, dtype=torch.float64)
    return G


# ------------------------------------------------------------------------------------ one rank
@pytest.mark.parametrize("name", ["ref_layer_n1", "ref_layer_n1_nodrop", "ref_layer_postln_n1"])
@pytest.mark.parametrize("recompute", [None, "full", "full_attn", "core_attn"])
def test_forward_and_hand_derived_backward_match_the_reference(monkeypatch, name, recompute):
    from fleetx_b200 import functional as Fn
    from fleetx_b200.dist import SingleGroup

    emu_ops.install(monkeypatch)
    d, m = load(name)
    P, M = params_and_meta(d, m)
    M.recompute = recompute
    group = _cpu_group(SingleGroup)()
    if m["dropout"] > 0:  # the forward draws once; a recomputing backward draws the same mask again
        emu_ops.ATTN_KEEP_QUEUE.extend([d["mask.attn.0"]] * 2)
    y, saved = Fn.decoder_layer_fwd(d["in.x"], P, M, group)
    check(y, d["out.y"], "y")
    kept = Fn.prune_saved(saved, M)
    if recompute == "full":
        assert set(kept) == {"x", "shape"}
    elif recompute is not None:
        assert kept["ctx"] is None and kept["lse"] is None and (kept["qkv"] is None) == (recompute == "full_attn")
    dx, g = Fn.decoder_layer_bwd(d["in.dy"], Fn.rebuild_saved(kept, P, M, group), P, M, group)
    check(dx, d["out.dx"], "dx")
    check_grads(g, d, m)


def test_kernel_order_of_the_pre_ln_layer(monkeypatch):
    """The launch sequence of one forward + backward: what DESIGN.md §4 lists and the ncu launch list shows."""
    from fleetx_b200 import functional as Fn
    from fleetx_b200.dist import SingleGroup

    emu_ops.install(monkeypatch)
    d, m = load("ref_layer_n1_nodrop")
    P, M = params_and_meta(d, m)
    group = _cpu_group(SingleGroup)()
    _, saved = Fn.decoder_layer_fwd(d["in.x"], P, M, group)
    fwd = list(emu_ops.CALLS)
    assert fwd == ["layernorm_fwd", "gemm", "attn_causal_fwd", "gemm", "bias_dropout_residual_fwd",
                   "layernorm_fwd", "gemm", "gemm", "bias_dropout_residual_fwd"]
    emu_ops.CALLS.clear()
    Fn.decoder_layer_bwd(d["in.dy"], saved, P, M, group)
    bwd = list(emu_ops.CALLS)
    assert bwd.count("gemm") == 8 and bwd.count("colsum") == 4 and bwd.count("layernorm_bwd") == 2
    assert bwd.count("attn_causal_bwd") == 1 and bwd.count("bias_dropout_residual_bwd") == 2
    assert bwd[0] == "bias_dropout_residual_bwd" and bwd[-1] == "layernorm_bwd"


def test_autograd_node_gives_the_same_numbers(monkeypatch):
    from fleetx_b200 import functional as Fn
    from fleetx_b200.dist import SingleGroup

    emu_ops.install(monkeypatch)
    d, m = load("ref_layer_n1")
    P, M = params_and_meta(d, m)
    emu_ops.ATTN_KEEP_QUEUE.append(d["mask.attn.0"])
    x = d["in.x"].clone().requires_grad_(True)
    ps = [t.clone().requires_grad_(True) for t in P.tensors()]
    y = Fn.DecoderLayerFunction.apply(x, M, _cpu_group(SingleGroup)(), *ps)
    y.backward(d["in.dy"])
    check(y.detach(), d["out.y"], "y")
    check(x.grad, d["out.dx"], "dx")
    order = ("norm1_w", "norm1_b", "qkv_w", "qkv_b", "out_w", "out_b", "norm2_w", "norm2_b", "lin1_w", "lin1_b",
             "lin2_w", "lin2_b")
    for short, t in zip(order, ps):
        check(t.grad, d["grad." + NAMES[short]], "grad " + NAMES[short])


# ------------------------------------------------------------------------------------ n ranks over gloo
def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, name, ret, rows=0):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(1)
        from fleetx_b200 import functional as Fn
        from fleetx_b200.dist import TorchDistGroup

        emu_ops.install_permanently()
        d, m = load(name)
        assert m["n"] == world
        P, M = params_and_meta(d, m, rank)
        group = _cpu_group(TorchDistGroup)()
        group.row_owner = bool(rows)  # the row-owner order of the layer over the plumbing statement of its collectives
        group.pipeline_chunks = int(rows)  # ... in that many batch chunks
        if m["dropout"] > 0:
            emu_ops.ATTN_KEEP_QUEUE.append(d[f"mask.attn.{rank}"])
        x, dy = d["in.x"], d["in.dy"]
        if m["sequence_parallel"]:
            sl = m["s"] // world
            cut = lambda t: t[:, rank * sl:(rank + 1) * sl].transpose(0, 1).contiguous()  # noqa: E731  [s/n, b, h]
            # plain, and every recompute granularity (hybrid_model.py:332-335, 406-408, 537-539 under :475-520)
            for rc in (None, "full", "full_attn", "core_attn"):
                M.recompute = rc
                if m["dropout"] > 0 and rc is not None:  # forward + its re-run in backward each take the attention mask
                    emu_ops.ATTN_KEEP_QUEUE.extend([d[f"mask.attn.{rank}"]] * 2)
                y, saved = Fn.decoder_layer_fwd_sp(cut(x), P, M, group)
                kept = Fn.prune_saved_sp(saved, M)
                if rc == "full":
                    assert set(kept) == {"x", "shape", "sp"}
                del saved
                dx, g = Fn.decoder_layer_bwd_sp(cut(dy), Fn.rebuild_saved_sp(kept, P, M, group), P, M, group)
                check(y, cut(d["out.y"]), f"y shard ({rc})")
                check(dx, cut(d["out.dx"]), f"dx shard ({rc})")
                check_grads(g, d, m, rank)
            M.recompute = None
        else:
            y, saved = Fn.decoder_layer_fwd(x, P, M, group)
            assert bool(saved.get("rows")) == bool(rows) and saved.get("chunks", 0) == (min(rows, x.shape[0]) if rows else 0)
            dx, g = Fn.decoder_layer_bwd(dy, saved, P, M, group)
            check(y, d["out.y"], "y")
            check(dx, d["out.dx"], "dx")
        check_grads(g, d, m, rank)
        ret[rank] = "ok"
    except BaseException as e:  # noqa: BLE001
        ret[rank] = f"{type(e).__name__}: {e}"
        raise
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("name,world,rows", [("ref_layer_tp2", 2, 0), ("ref_layer_tp4", 4, 0),
                                             ("ref_layer_sp2", 2, 0), ("ref_layer_sp4", 4, 0),
                                             ("ref_layer_postln_tp2", 2, 0), ("ref_layer_postln_sp2", 2, 0),
                                             ("ref_layer_tp2", 2, 1), ("ref_layer_tp4", 4, 1),
                                             ("ref_layer_tp2", 2, 2), ("ref_layer_tp4", 4, 2)])
def test_tensor_and_sequence_parallel_sequencing_over_gloo(name, world, rows):
    """Each process runs ONE mp rank of the product's own forward / backward (allreduce after the row-parallel
    matmuls, allreduce of dX in front of the column-parallel ones; sequence parallel: all-gather / reduce-scatter,
    shard-local dropout, summed LayerNorm / bias gradients) and must land on the reference's numbers, which came
    from n simulated ranks of hybrid_model.py / sequence_parallel_utils.py. rows = 1 / 2: the row-owner order
    (functional.decoder_layer_fwd_rows / _bwd_rows: each rank does the row-wise work behind every allreduce on its
    own rows, gathered results, parameter-gradient partial sums summed at the end) with the batch in that many
    chunks."""
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, name, ret, rows)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
    for p in procs:
        if p.is_alive():
            p.terminate()
    assert dict(ret) == {r: "ok" for r in range(world)}, dict(ret)


def test_module_draws_its_dropout_streams_in_the_reference_order():
    """TransformerDecoderLayer._meta(): one draw per dropout, in the order the executed reference logged them
    (fixture meta: 'local_seed' for attention, then 'global_seed' twice), nothing drawn in eval mode."""
    from fleetx_b200.layers import TransformerDecoderLayer
    from fleetx_b200.rng import get_rng_state_tracker, set_seed

    _, m = load("ref_layer_n1")
    assert [st for st, _ in m["dropout_streams_rank0"]] == ["local_seed", "global_seed", "global_seed"]
    layer = TransformerDecoderLayer(128, 2, 512, dropout=0.1, device="cpu")
    set_seed(1024, mp_rank=3)
    layer.train()
    m1, m2 = layer._meta(), layer._meta()
    assert (m1.rng_attn, m1.rng_h1, m1.rng_h2) == ((1024 + 123 + 30, 0), (1024, 0), (1024, 1))
    assert (m2.rng_attn, m2.rng_h1, m2.rng_h2) == ((1024 + 123 + 30, 1), (1024, 2), (1024, 3))
    layer.eval()
    before = get_rng_state_tracker().get_states_tracker()
    e = layer._meta()
    assert (e.p_attn, e.p_hidden1, e.p_hidden2) == (0.0, 0.0, 0.0)
    assert get_rng_state_tracker().get_states_tracker() == before
    with pytest.raises(RuntimeError, match="no CPU path"):
        layer(torch.zeros(1, 8, 128, dtype=torch.bfloat16))
