"""The checker: the CUDA layer against the oracle (CPU restatement of the reference layer).

Lives next to __graft_entry__.py, outside the product package, because it imports oracle/: only
__graft_entry__.smoke(), bench.py's parity block (after the timed region) and tests/ use it.

smoke()          one small decoder-layer forward+backward on cuda:0 through the C ABI
layer_parity()   the same comparison for any shape / mp group / sequence-parallel mode, per tensor"""
from __future__ import annotations

import torch


def layer_from_oracle_params(cfg, params, group=None, device="cuda"):
    """Build a TransformerDecoderLayer holding this rank's shard of oracle-format parameters."""
    from fleetx_b200.layers import TransformerDecoderLayer

    layer = TransformerDecoderLayer(cfg.d_model, cfg.nhead, cfg.dim_feedforward, dropout=cfg.dropout,
                                    attn_dropout=cfg.attn_dropout, act_dropout=cfg.act_dropout,
                                    normalize_before=cfg.normalize_before, num_partitions=cfg.num_partitions,
                                    mp_group=group, device=device)
    layer.load_full_state(params)
    return layer


def rel_err(a: torch.Tensor, b: torch.Tensor) -> float:
    a = a.detach().to("cpu", torch.float64)
    b = b.detach().to("cpu", torch.float64)
    denom = b.abs().max().item()
    return (a - b).abs().max().item() / (denom if denom > 0 else 1.0)


def smoke(verbose: bool = True) -> dict:
    from oracle import decoder_layer as O  # the checker; never on the product path
    from fleetx_b200.rng import set_seed

    if not torch.cuda.is_available():
        raise RuntimeError("smoke() needs a CUDA device (B200); there is no CPU path")
    torch.cuda.set_device(0)
    b, s = 2, 256
    cfg = O.LayerConfig(d_model=256, nhead=4, dim_feedforward=1024, dropout=0.1)
    seed = 1024
    params = {k: v.to(torch.bfloat16).float() for k, v in O.init_params(cfg, seed).items()}
    g = torch.Generator().manual_seed(7)
    x = torch.randn(b, s, cfg.d_model, generator=g).to(torch.bfloat16)
    dy = torch.randn(b, s, cfg.d_model, generator=g).to(torch.bfloat16)

    set_seed(seed)
    layer = layer_from_oracle_params(cfg, params)
    layer.train()
    xd = x.cuda().requires_grad_(True)
    y = layer(xd)
    y.backward(dy.cuda())
    torch.cuda.synchronize()

    masks = O.make_masks(cfg, b, s, seed)  # the same (seed, offset) draws, restated on the CPU
    y_ref, dx_ref, g_ref = O.decoder_layer_fwd_bwd(x.float(), dy.float(), params, cfg, masks)
    errs = {"y": rel_err(y, y_ref), "dx": rel_err(xd.grad, dx_ref)}
    for name, p in layer.named_parameters():
        errs["d" + name] = rel_err(p.grad, g_ref[name])
    worst = max(errs.values())
    if verbose:
        print("smoke: max relative error vs oracle: %.3e (%s)" % (worst, max(errs, key=errs.get)))
    if not worst < 2e-2:  # north_star: <= 2e-2 relative for bf16
        raise AssertionError(f"smoke: parity with the oracle failed: {errs}")
    return errs


def perturbed_params(cfg, seed: int):
    """Oracle-format FULL parameters, bf16-rounded, with non-trivial biases and LayerNorm affine terms (the
    reference initialises those to 0 / 1, which would hide a wrong bias or gamma/beta gradient path)."""
    from oracle import decoder_layer as O

    params = O.init_params(cfg, seed)
    g = torch.Generator().manual_seed(seed + 17)
    for name in O.PARAM_NAMES:
        t = params[name]
        if name.endswith(".bias"):
            t = 0.02 * torch.randn(t.shape, generator=g)
        elif name.startswith("norm"):
            t = 1.0 + 0.1 * torch.randn(t.shape, generator=g)
        params[name] = t if name.startswith("norm") else t.to(torch.bfloat16).float()
    return params


def layer_parity(group, b: int, s: int, h: int, nh: int, ffn: int, *, dropout: float = 0.1, seed: int = 1024,
                 sequence_parallel: bool = False, normalize_before: bool = True, recompute=None,
                 device="cuda", repeats: int = 1) -> dict:
    """fwd+bwd of TransformerDecoderLayer over `group` (this process = mp rank group.rank of group.nranks)
    against the oracle's simulation of the same mp degree with the same (seed, offset) dropout draws.
    Returns {tensor name: max|a-b| / max|b|} for y, dx and this rank's shard of every parameter gradient
    (worst over `repeats` calls: later calls reuse the symmetric buffer and its flag epochs)."""
    from oracle import decoder_layer as O
    from fleetx_b200.rng import set_seed

    world, rank = group.nranks, group.rank
    cfg = O.LayerConfig(d_model=h, nhead=nh, dim_feedforward=ffn, dropout=dropout, num_partitions=world,
                        normalize_before=normalize_before)
    params = perturbed_params(cfg, seed)
    g = torch.Generator().manual_seed(seed + 3)
    x = torch.randn(b, s, h, generator=g).to(torch.bfloat16)
    dy = torch.randn(b, s, h, generator=g).to(torch.bfloat16)
    masks = None
    if dropout > 0:
        masks = O.make_masks_sp(cfg, b, s, seed) if sequence_parallel and world > 1 else O.make_masks(cfg, b, s, seed)
    y_ref, dx_ref, g_ref = O.decoder_layer_fwd_bwd(x.float(), dy.float(), params, cfg, masks)
    g_shard = O.shard_params(g_ref, cfg, rank)
    del g_ref, masks
    layer = layer_from_oracle_params(cfg, params, group=group, device=device)
    layer.train()
    layer.sequence_parallel = bool(sequence_parallel)
    if recompute is not None:
        layer.use_recompute, layer.recompute_granularity = True, recompute
    errs: dict = {}
    for _ in range(repeats):
        set_seed(seed, mp_rank=rank)
        for p_ in layer.parameters():
            p_.grad = None
        if sequence_parallel:
            xin = O.sp_scatter(x, world, rank).to(device).requires_grad_(True)
            y = layer(xin)
            y.backward(O.sp_scatter(dy, world, rank).to(device))
            cur = {"y": rel_err(y, O.sp_scatter(y_ref, world, rank)), "dx": rel_err(xin.grad, O.sp_scatter(dx_ref, world, rank))}
        else:
            xin = x.to(device).requires_grad_(True)
            y = layer(xin)
            y.backward(dy.to(device))
            cur = {"y": rel_err(y, y_ref), "dx": rel_err(xin.grad, dx_ref)}
        torch.cuda.synchronize()
        for name, p_ in layer.named_parameters():
            cur["d" + name] = rel_err(p_.grad, g_shard[name])
        for k, v in cur.items():
            errs[k] = max(errs.get(k, 0.0), v)
    return errs
