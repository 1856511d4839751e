"""world_size-2 (and 4) gloo runs on CPU of the multi-rank host logic: each process holds one mp
rank's parameter shard, computes its share of the layer with the oracle's ops, and exchanges
through fleetx_b200.dist (TorchDistGroup + the _c_identity / _mp_allreduce autograd pair). The
result and every per-rank gradient must equal the single-process oracle."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import decoder_layer as O


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank_forward(x, shard, full, cfg, group, masks, rank):
    from fleetx_b200.dist import c_identity, mp_allreduce

    nhl, hd = cfg.nhead // cfg.num_partitions, cfg.hd
    t = O.layer_norm(x, full["norm1.weight"], full["norm1.bias"])
    mix = O.column_parallel_linear(c_identity(t, group), shard["self_attn.qkv_proj.weight"], shard["self_attn.qkv_proj.bias"])
    q, k, v = O.fuse_prepare_qkv(mix, nhl, hd)
    ctx = O.core_attn(q, k, v, hd, cfg.attn_dropout, masks.attn[rank] if masks.attn else None)
    t = mp_allreduce(ctx @ shard["self_attn.out_proj.weight"], group) + full["self_attn.out_proj.bias"]
    t = x + O.dropout(t, cfg.dropout, masks.hidden1)
    res = t
    t = O.layer_norm(t, full["norm2.weight"], full["norm2.bias"])
    a = O.gelu_tanh(O.column_parallel_linear(c_identity(t, group), shard["linear1.weight"], shard["linear1.bias"]))
    t = mp_allreduce(a @ shard["linear2.weight"], group) + full["linear2.bias"]
    return res + O.dropout(t, cfg.act_dropout, masks.hidden2)


def _worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from fleetx_b200.dist import TorchDistGroup, shard_columns, shard_rows
        from fleetx_b200.rng import get_rng_state_tracker, set_seed

        torch.set_num_threads(1)
        group = TorchDistGroup()
        assert (group.rank, group.nranks) == (rank, world)
        cfg = O.LayerConfig(d_model=64, nhead=8, dim_feedforward=256, dropout=0.1, num_partitions=world)
        full = O.init_params(cfg, 1024, dtype=torch.float64)
        g = torch.Generator().manual_seed(11)
        x = torch.randn(2, 16, 64, generator=g, dtype=torch.float64)
        dy = torch.randn(2, 16, 64, generator=g, dtype=torch.float64)
        masks = O.make_masks(cfg, 2, 16, 1024)
        # the tracker on this rank must hand out exactly the seeds the oracle used for its masks
        set_seed(1024, mp_rank=rank)
        st = get_rng_state_tracker().get_states_tracker()
        assert st["local_seed"][0] == O.rng_seeds(1024, rank)["local_seed"] and st["global_seed"][0] == 1024

        shard = {}
        for k, v in full.items():
            if k in ("self_attn.qkv_proj.weight", "self_attn.qkv_proj.bias", "linear1.weight", "linear1.bias"):
                shard[k] = shard_columns(v, rank, world)
            elif k in ("self_attn.out_proj.weight", "linear2.weight"):
                shard[k] = shard_rows(v, rank, world)
        shard = {k: v.clone().requires_grad_(True) for k, v in shard.items()}
        rep = {k: v.clone().requires_grad_(True) for k, v in full.items() if k not in shard}
        xr = x.clone().requires_grad_(True)
        y = _rank_forward(xr, shard, rep, cfg, group, masks, rank)
        y.backward(dy)

        y_ref, dx_ref, g_ref = O.decoder_layer_fwd_bwd(x, dy, full, cfg, masks)
        ok = torch.allclose(y, y_ref, atol=1e-11) and torch.allclose(xr.grad, dx_ref, atol=1e-11)
        ref_shard = O.shard_params(g_ref, cfg, rank)
        for k, v in shard.items():
            ok = ok and torch.allclose(v.grad, ref_shard[k], atol=1e-10)
        for k, v in rep.items():  # replicated parameters see the full gradient on every rank
            ok = ok and torch.allclose(v.grad, g_ref[k], atol=1e-10)
        # parallel_matmul(parallel_output=False) (hybrid_model.py:45-66): logits against this rank's vocabulary slice,
        # _c_concat along the last dim; backward = this rank's columns of the gradient, then _c_identity's allreduce
        from fleetx_b200.dist import c_concat, c_identity

        gv = torch.Generator().manual_seed(5)
        wte = torch.randn(8 * world, 64, generator=gv, dtype=torch.float64)     # [vocab, h], split by rows
        hid = torch.randn(2, 16, 64, generator=gv, dtype=torch.float64)
        dlog = torch.randn(2, 16, 8 * world, generator=gv, dtype=torch.float64)
        w_r = wte[rank * 8:(rank + 1) * 8].clone().requires_grad_(True)
        hr = hid.clone().requires_grad_(True)
        logits = c_concat(c_identity(hr, group) @ w_r.t(), group)
        logits.backward(dlog)
        hf, wf = hid.clone().requires_grad_(True), wte.clone().requires_grad_(True)
        ref = hf @ wf.t()
        ref.backward(dlog)
        ok = ok and torch.allclose(logits, ref, atol=1e-12) and torch.allclose(hr.grad, hf.grad, atol=1e-11)
        ok = ok and torch.allclose(w_r.grad, wf.grad[rank * 8:(rank + 1) * 8], atol=1e-11)
        # plain collective plumbing
        t = torch.full((4,), float(rank + 1), dtype=torch.float64)
        group.all_reduce_(t)
        ok = ok and bool((t == sum(range(1, world + 1))).all())
        ret[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_tp_ranks_over_gloo_match_single_process_oracle(world):
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=240)
    for p in procs:
        if p.is_alive():
            p.kill()
            pytest.fail("gloo worker hung")
    assert all(ret.get(r) for r in range(world)), dict(ret)


# ---------------------------------------------------------------------------- sequence parallel
def _rank_forward_sp(x_shard, shard, rep, cfg, group, masks_sp, rank):
    """One mp rank of the sequence-parallel layer (hybrid_model.py:525-566 with Column/RowSequenceParallel-
    Linear, sequence_parallel_utils.py:150-326) on sequence-first [s/n, b, h] activations, composed from
    the oracle's ops and fleetx_b200.dist's four operators. masks_sp: shard-sized hidden masks [s/n, b, h]."""
    from fleetx_b200.dist import AllGatherOp, ReduceScatterOp

    nhl, hd = cfg.nhead // cfg.num_partitions, cfg.hd
    t = O.layer_norm(x_shard, rep["norm1.weight"], rep["norm1.bias"])
    full = AllGatherOp.apply(t, group)                                   # [s, b, h]
    mix = O.column_parallel_linear(full, shard["self_attn.qkv_proj.weight"], shard["self_attn.qkv_proj.bias"])
    q, k, v = O.fuse_prepare_qkv(mix.transpose(0, 1), nhl, hd)           # the oracle's attention is batch-first
    ctx = O.core_attn(q, k, v, hd, cfg.attn_dropout, masks_sp["attn"][rank] if masks_sp["attn"] else None)
    part = (ctx @ shard["self_attn.out_proj.weight"]).transpose(0, 1)    # back to [s, b, h]
    t = ReduceScatterOp.apply(part.contiguous(), group) + rep["self_attn.out_proj.bias"]
    t = x_shard + O.dropout(t, cfg.dropout, masks_sp["h1"])
    res = t
    t = O.layer_norm(t, rep["norm2.weight"], rep["norm2.bias"])
    full = AllGatherOp.apply(t, group)
    a = O.gelu_tanh(O.column_parallel_linear(full, shard["linear1.weight"], shard["linear1.bias"]))
    t = ReduceScatterOp.apply((a @ shard["linear2.weight"]).contiguous(), group) + rep["linear2.bias"]
    return res + O.dropout(t, cfg.act_dropout, masks_sp["h2"])


def _worker_sp(rank, world, port, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from fleetx_b200.dist import (GatherOp, ScatterOp, TorchDistGroup, all_reduce_gradient_hook, shard_columns,
                                      shard_rows)

        torch.set_num_threads(1)
        group = TorchDistGroup()
        b, s = 2, 16
        cfg = O.LayerConfig(d_model=64, nhead=8, dim_feedforward=256, dropout=0.1, num_partitions=world)
        full = O.init_params(cfg, 1024, dtype=torch.float64)
        g = torch.Generator().manual_seed(11)
        x = torch.randn(b, s, 64, generator=g, dtype=torch.float64)
        dy = torch.randn(b, s, 64, generator=g, dtype=torch.float64)
        masks = O.make_masks_sp(cfg, b, s, 1024)           # full-size masks the oracle layer consumes
        masks_sp = {"attn": masks.attn, "h1": O.sp_scatter(masks.hidden1, world, 0), "h2": O.sp_scatter(masks.hidden2, world, 0)}
        # every rank's shard of the tiled mask is the same shard-sized mask
        ok = torch.equal(O.sp_scatter(masks.hidden1, world, rank), masks_sp["h1"])

        shard = {}
        for k, v in full.items():
            if k in ("self_attn.qkv_proj.weight", "self_attn.qkv_proj.bias", "linear1.weight", "linear1.bias"):
                shard[k] = shard_columns(v, rank, world)
            elif k in ("self_attn.out_proj.weight", "linear2.weight"):
                shard[k] = shard_rows(v, rank, world)
        shard = {k: v.clone().requires_grad_(True) for k, v in shard.items()}
        rep = {k: v.clone().requires_grad_(True) for k, v in full.items() if k not in shard}
        for v in rep.values():  # LayerNorm weight / bias and the row-parallel biases see only this rank's rows
            v.register_hook(all_reduce_gradient_hook(group))

        # ScatterOp / GatherOp as the model uses them around the layer stack (hybrid_model.py:709, :738)
        x_sf = x.transpose(0, 1).contiguous().requires_grad_(True)       # [s, b, h] on every rank
        xs = ScatterOp.apply(x_sf, group)
        ys = _rank_forward_sp(xs, shard, rep, cfg, group, masks_sp, rank)
        y = GatherOp.apply(ys, group).transpose(0, 1)                    # [b, s, h]
        y.backward(dy)

        y_ref, dx_ref, g_ref = O.decoder_layer_fwd_bwd(x, dy, full, cfg, masks)
        ok = ok and torch.allclose(y, y_ref, atol=1e-11)
        # GatherOp's backward scatters dy, ScatterOp's backward all-gathers dx: every rank ends with the full dx
        ok = ok and torch.allclose(x_sf.grad.transpose(0, 1), dx_ref, atol=1e-11)
        ref_shard = O.shard_params(g_ref, cfg, rank)
        for k, v in shard.items():
            ok = ok and torch.allclose(v.grad, ref_shard[k], atol=1e-10)
        for k, v in rep.items():
            ok = ok and torch.allclose(v.grad, g_ref[k], atol=1e-10)
        # and the oracle's own scatter / gather are inverse to each other
        ok = ok and torch.equal(O.sp_gather([O.sp_scatter(x, world, r) for r in range(world)]), x)
        ret[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_sequence_parallel_ranks_over_gloo_match_oracle(world):
    """sequence_parallel=True equals the plain layer with the shard mask repeated over the ranks."""
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_worker_sp, args=(r, world, port, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=240)
    for p in procs:
        if p.is_alive():
            p.kill()
            pytest.fail("gloo worker hung")
    assert all(ret.get(r) for r in range(world)), dict(ret)
