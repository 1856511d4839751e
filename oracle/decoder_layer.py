"""ORACLE (test infrastructure, never on the product path).

CPU restatement, op by op, of the reference's GPT hybrid-parallel decoder layer:
ppfleetx/models/language_model/gpt/dygraph/hybrid_model.py — MultiHeadAttention (:69-347) and
TransformerDecoderLayer (:439-571) — plus the tensor-parallel linears it is built from
(fleet.meta_parallel.{Column,Row}ParallelLinear; their in-tree mirrors are
sequence_parallel_utils.py:150-326 with allreduce in place of reduce-scatter / all-gather).

PARITY: pinned to the reference's own Python, unpinned against Paddle's kernels. The arithmetic of the
primitives lives in PaddlePaddle (unpinned; Dockerfile:7 installs nightly paddlepaddle-gpu==0.0.0.post112,
ppfleetx/utils/version.py:18-21 requires >= 2.3.0), which is absent from /root/reference and not
installable here, and the reference holds no test, golden vector or fixture for this path (its only test
is codestyle/test_docstring_checker.py). The Python that composes the primitives does run here:
tools/make_reference_golden.py executes hybrid_model.py / sequence_parallel_utils.py unmodified over the
torch-backed `paddle` stand-in of oracle/paddle_shim.py (mp 1 / 2 / 4, sequence parallel 2 / 4) and
tests/test_reference_golden.py holds this restatement to those outputs and gradients (1e-10, fp64). The
same math is stated three times in-tree (hybrid_model.py, single_model.py:98-116,189-213,397-427,
auto_model.py:181-206,382-420); all three are executed there and agree (the auto-parallel one after cutting
the fused qkv_proj into its separate q / k / v projections).

Written in torch-CPU (fp32 / fp64): non-fused, one function per reference op, tensor
parallelism simulated explicitly (n shards in one process, allreduce = Python sum). Gradients
come from torch autograd over this forward, the restatement of Paddle's autograd. Dropout masks
are INPUTS (uint8 keep masks), so any generator can be checked against it.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional

import torch
import torch.nn.functional as F

EPS = 1e-5  # nn.LayerNorm(d_model, epsilon=1e-5), hybrid_model.py:513-514


# ------------------------------------------------------------------------------------ ops
def layer_norm(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, eps: float = EPS):
    """nn.LayerNorm over the last dim, biased variance (hybrid_model.py:513-514, :534, :554)."""
    mu = x.mean(dim=-1, keepdim=True)
    var = ((x - mu) ** 2).mean(dim=-1, keepdim=True)
    return (x - mu) / torch.sqrt(var + eps) * gamma + beta


def gelu_tanh(z: torch.Tensor) -> torch.Tensor:
    """F.gelu(x, approximate=True) (hybrid_model.py:558)."""
    return 0.5 * z * (1.0 + torch.tanh(math.sqrt(2.0 / math.pi) * (z + 0.044715 * z ** 3)))


def dropout(x: torch.Tensor, p: float, keep: Optional[torch.Tensor]) -> torch.Tensor:
    """F.dropout(..., mode="upscale_in_train") (hybrid_model.py:281-285, :521-522).

    keep: uint8/bool mask of x's shape, or None for p == 0 / eval."""
    if keep is None or p == 0.0:
        return x
    return x * keep.to(x.dtype) / (1.0 - p)


def column_parallel_linear(x, w_shard, b_shard):
    """fleet.meta_parallel.ColumnParallelLinear.forward with gather_output=False:
    _c_identity(x) then x·W_shard + b_shard (weights are [in, out/n]); mirrored in-tree at
    sequence_parallel_utils.py:150-232. Call sites hybrid_model.py:125-131, :497-503."""
    y = x @ w_shard
    return y if b_shard is None else y + b_shard


def row_parallel_linear(x_shards: List[torch.Tensor], w_shards: List[torch.Tensor], bias):
    """fleet.meta_parallel.RowParallelLinear.forward with input_is_parallel=True:
    local x_i·W_i without bias, allreduce(sum) over the mp group, THEN + bias once
    (mirror: sequence_parallel_utils.py:235-326). Call sites hybrid_model.py:157-163, :505-511."""
    y = None
    for xs, ws in zip(x_shards, w_shards):
        part = xs @ ws
        y = part if y is None else y + part  # == mp allreduce
    return y if bias is None else y + bias


def fuse_prepare_qkv(mix: torch.Tensor, nh_local: int, hd: int):
    """_fuse_prepare_qkv (hybrid_model.py:165-186): reshape [b,s,nh/n,3*hd], transpose to
    [b,nh/n,s,3*hd], split the last dim into q, k, v."""
    b, s, _ = mix.shape
    mix = mix.reshape(b, s, nh_local, 3 * hd).permute(0, 2, 1, 3)
    return mix[..., :hd], mix[..., hd:2 * hd], mix[..., 2 * hd:]


def core_attn(q, k, v, hd: int, p_drop: float = 0.0, keep: Optional[torch.Tensor] = None,
              attn_mask: Optional[torch.Tensor] = None):
    """core_attn (hybrid_model.py:268-298). Training passes attn_mask=None and the reference
    calls incubate.softmax_mask_fuse_upper_triangle (:277) — causal mask + softmax. With a mask
    (:273-275) it is softmax(product + mask); the model's mask is -1e4 * triu(1) (:716-726),
    identical in fp32 up to exp(-1e4) == 0. Returns out merged to [b, s, nh/n*hd] (:293-296)."""
    s = q.shape[2]
    product = torch.matmul(q, k.transpose(-1, -2)) * (hd ** -0.5)
    if attn_mask is not None:
        weights = F.softmax(product + attn_mask, dim=-1)
    else:
        causal = torch.triu(torch.ones(s, s, dtype=torch.bool), diagonal=1)
        weights = F.softmax(product.masked_fill(causal, float("-inf")), dim=-1)
    weights = dropout(weights, p_drop, keep)
    out = torch.matmul(weights, v)  # [b, nh/n, s, hd]
    b, nhl = out.shape[0], out.shape[1]
    return out.permute(0, 2, 1, 3).reshape(b, s, nhl * hd)


# ------------------------------------------------------------------------------------ layer
@dataclass
class LayerConfig:
    d_model: int
    nhead: int
    dim_feedforward: int
    dropout: float = 0.1
    attn_dropout: Optional[float] = None
    act_dropout: Optional[float] = None
    normalize_before: bool = True
    num_partitions: int = 1
    # None: d_model // nhead, as in the reference (:113). An explicit value states ONE mp rank's shard of a larger
    # layer as a layer of its own (nhead = that rank's heads, dim_feedforward = its FFN columns, weights
    # [h, 3*nhead*hd] / [nhead*hd, h]): what the full-size tests compare a rank's kernels against.
    head_dim: Optional[int] = None

    def __post_init__(self):
        # hybrid_model.py:469-470
        if self.attn_dropout is None:
            self.attn_dropout = self.dropout
        if self.act_dropout is None:
            self.act_dropout = self.dropout
        assert self.head_dim is not None or self.d_model % self.nhead == 0  # :114-115
        assert self.nhead % self.num_partitions == 0  # :117-118
        assert self.dim_feedforward % self.num_partitions == 0

    @property
    def hd(self) -> int:
        return self.head_dim if self.head_dim is not None else self.d_model // self.nhead


PARAM_NAMES = (
    "norm1.weight", "norm1.bias", "self_attn.qkv_proj.weight", "self_attn.qkv_proj.bias",
    "self_attn.out_proj.weight", "self_attn.out_proj.bias", "norm2.weight", "norm2.bias",
    "linear1.weight", "linear1.bias", "linear2.weight", "linear2.bias",
)


def init_params(cfg: LayerConfig, seed: int = 1024, dtype=torch.float32) -> Dict[str, torch.Tensor]:
    """FULL (unsharded) parameters. Weights Normal(0, 0.02) (hybrid_model.py:671-673), biases 0,
    LayerNorm gamma 1 / beta 0. Shapes per SURVEY §8(a1), Paddle [in, out] layout."""
    g = torch.Generator().manual_seed(seed)
    h, f = cfg.d_model, cfg.dim_feedforward

    def w(i, o):
        return (torch.randn(i, o, generator=g, dtype=torch.float32) * 0.02).to(dtype)

    return {
        "norm1.weight": torch.ones(h, dtype=dtype), "norm1.bias": torch.zeros(h, dtype=dtype),
        "self_attn.qkv_proj.weight": w(h, 3 * h), "self_attn.qkv_proj.bias": torch.zeros(3 * h, dtype=dtype),
        "self_attn.out_proj.weight": w(h, h), "self_attn.out_proj.bias": torch.zeros(h, dtype=dtype),
        "norm2.weight": torch.ones(h, dtype=dtype), "norm2.bias": torch.zeros(h, dtype=dtype),
        "linear1.weight": w(h, f), "linear1.bias": torch.zeros(f, dtype=dtype),
        "linear2.weight": w(f, h), "linear2.bias": torch.zeros(h, dtype=dtype),
    }


def shard_params(params: Dict[str, torch.Tensor], cfg: LayerConfig, rank: int) -> Dict[str, torch.Tensor]:
    """What mp rank `rank` holds. Column-parallel weights/biases are split by output columns —
    Wqkv's columns are ordered per head [q|k|v] (hybrid_model.py:167-173), so a plain column
    split hands each rank whole heads; row-parallel weights are split by input rows, their bias
    and the LayerNorm parameters are replicated."""
    n = cfg.num_partitions
    out = {}
    for name, t in params.items():
        if name in ("self_attn.qkv_proj.weight", "linear1.weight"):
            out[name] = t.chunk(n, dim=1)[rank].contiguous()
        elif name in ("self_attn.qkv_proj.bias", "linear1.bias"):
            out[name] = t.chunk(n, dim=0)[rank].contiguous()
        elif name in ("self_attn.out_proj.weight", "linear2.weight"):
            out[name] = t.chunk(n, dim=0)[rank].contiguous()
        else:
            out[name] = t
    return out


@dataclass
class DropoutMasks:
    """Keep masks (uint8). attn[r]: [b, nh/n, s, s] for mp rank r ('local_seed' stream, distinct
    per rank, env.py:42); hidden1 / hidden2: [b, s, h] ('global_seed' stream, identical on all
    ranks, env.py:43). None = no dropout."""
    attn: Optional[List[torch.Tensor]] = None
    hidden1: Optional[torch.Tensor] = None
    hidden2: Optional[torch.Tensor] = None


def decoder_layer_forward(x: torch.Tensor, params: Dict[str, torch.Tensor], cfg: LayerConfig,
                          masks: Optional[DropoutMasks] = None, training: bool = True,
                          attn_mask: Optional[torch.Tensor] = None,
                          intermediates: Optional[dict] = None) -> torch.Tensor:
    """TransformerDecoderLayer.forward (hybrid_model.py:525-566) with use_cache=False, executed as
    cfg.num_partitions simulated mp ranks. `params` are the FULL parameters (autograd then yields
    full-size gradients whose slices are the per-rank gradients)."""
    n = cfg.num_partitions
    masks = masks or DropoutMasks()
    nhl, hd = cfg.nhead // n, cfg.hd
    shards = [shard_params(params, cfg, r) for r in range(n)]
    p_attn = cfg.attn_dropout if training else 0.0
    p1 = cfg.dropout if training else 0.0
    p2 = cfg.act_dropout if training else 0.0

    residual = x  # :531
    t = layer_norm(x, params["norm1.weight"], params["norm1.bias"]) if cfg.normalize_before else x  # :533-534

    # ---- self_attn (:541 -> MultiHeadAttention.forward :300-347)
    ctx = []
    for r in range(n):
        mix = column_parallel_linear(t, shards[r]["self_attn.qkv_proj.weight"], shards[r]["self_attn.qkv_proj.bias"])  # :166
        q, k, v = fuse_prepare_qkv(mix, nhl, hd)  # :167-173
        keep = masks.attn[r] if (masks.attn is not None and p_attn > 0) else None
        ctx.append(core_attn(q, k, v, hd, p_attn, keep, attn_mask))  # :335
    t = row_parallel_linear(ctx, [s_["self_attn.out_proj.weight"] for s_ in shards], params["self_attn.out_proj.bias"])  # :340
    if intermediates is not None:
        intermediates["attn_out"] = t

    t = residual + dropout(t, p1, masks.hidden1 if p1 > 0 else None)  # :546-547
    if not cfg.normalize_before:
        t = layer_norm(t, params["norm1.weight"], params["norm1.bias"])  # :549-550

    residual = t  # :552
    if cfg.normalize_before:
        t = layer_norm(t, params["norm2.weight"], params["norm2.bias"])  # :553-554

    # ---- FFN (:556-559): dropout2(linear2(gelu(linear1(t))))
    acts = [gelu_tanh(column_parallel_linear(t, shards[r]["linear1.weight"], shards[r]["linear1.bias"])) for r in range(n)]
    t = row_parallel_linear(acts, [s_["linear2.weight"] for s_ in shards], params["linear2.bias"])
    t = dropout(t, p2, masks.hidden2 if p2 > 0 else None)

    t = residual + t  # :561
    if not cfg.normalize_before:
        t = layer_norm(t, params["norm2.weight"], params["norm2.bias"])  # :563-564
    return t


def decoder_layer_fwd_bwd(x, dy, params, cfg, masks=None, attn_mask=None):
    """Forward + backward (torch autograd = restated Paddle autograd). Returns
    (y, dx, {param name: full-size gradient})."""
    x = x.detach().clone().requires_grad_(True)
    ps = {k: v.detach().clone().requires_grad_(True) for k, v in params.items()}
    y = decoder_layer_forward(x, ps, cfg, masks, training=True, attn_mask=attn_mask)
    y.backward(dy)
    return y.detach(), x.grad.detach(), {k: v.grad.detach() for k, v in ps.items()}


# ------------------------------------------------------------------------------------ seeds
def rng_seeds(seed: int, mp_rank: int = 0, pp_rank: int = 0, data_world_rank: int = 0):
    """ppfleetx/utils/env.py:27-46: the two dropout streams of the RNG state tracker."""
    return {"global_seed": seed + data_world_rank,
            "local_seed": seed + 123 + mp_rank * 10 + pp_rank * 1000}


def make_masks(cfg: LayerConfig, b: int, s: int, seed: int = 1024, offsets=(0, 0, 1)) -> DropoutMasks:
    """Masks the library generates for one layer call: attention dropout at offset offsets[0] of
    each rank's 'local_seed' stream, dropout1 / dropout2 at offsets[1], offsets[2] of the shared
    'global_seed' stream (generator restated in oracle/philox.py)."""
    from . import philox

    n = cfg.num_partitions
    nhl = cfg.nhead // n
    attn = None
    if cfg.attn_dropout > 0:
        attn = [torch.from_numpy(philox.attn_dropout_mask(rng_seeds(seed, r)["local_seed"], offsets[0], b, nhl, s,
                                                          cfg.attn_dropout)) for r in range(n)]
    gseed = rng_seeds(seed)["global_seed"]
    h1 = h2 = None
    if cfg.dropout > 0:
        h1 = torch.from_numpy(philox.dropout_mask(gseed, offsets[1], b * s * cfg.d_model, cfg.dropout)).reshape(b, s, cfg.d_model)
    if cfg.act_dropout > 0:
        h2 = torch.from_numpy(philox.dropout_mask(gseed, offsets[2], b * s * cfg.d_model, cfg.act_dropout)).reshape(b, s, cfg.d_model)
    return DropoutMasks(attn=attn, hidden1=h1, hidden2=h2)


# ------------------------------------------------------------------------------------ sequence parallel
# sequence_parallel=True (hybrid_model.py:475-520, sequence_parallel_utils.py:36-126, 150-326). The layer
# then works on sequence-first activations and rank r of n keeps only rows [r·s/n, (r+1)·s/n) between the
# blocks. Summed over the ranks this is the SAME arithmetic as the plain tensor-parallel layer, with one
# difference the reference makes on purpose or not: dropout1 / dropout2 are applied to the [s/n, b, h]
# shard under the 'global_seed' stream (hybrid_model.py:546, :556), whose state is identical on all mp
# ranks, so every rank applies the same mask to its own rows. The oracle therefore restates the
# sequence-parallel layer as: the plain layer on the batch-first full tensor, with hidden masks that repeat
# one shard-sized mask n times along the sequence.
def sp_scatter(x_bsh: torch.Tensor, n: int, r: int) -> torch.Tensor:
    """Rank r's shard [s/n, b, h] (sequence-first, utils:36-45) of a batch-first [b, s, h] tensor."""
    b, s, h = x_bsh.shape
    assert s % n == 0, f"Input sequence length {s} can't be divided exactly by sequence parallelism {n}"
    sl = s // n
    return x_bsh[:, r * sl:(r + 1) * sl].transpose(0, 1).contiguous()


def sp_gather(shards: List[torch.Tensor]) -> torch.Tensor:
    """Inverse of sp_scatter: n shards [s/n, b, h] -> batch-first [b, s, h] (utils:47-55)."""
    return torch.cat(shards, dim=0).transpose(0, 1).contiguous()


def sp_tile_mask(mask_shard: torch.Tensor, n: int) -> torch.Tensor:
    """One [s/n, b, h] keep mask, used by every rank on its own rows -> the equivalent [b, s, h] mask."""
    return sp_gather([mask_shard] * n)


def make_masks_sp(cfg: LayerConfig, b: int, s: int, seed: int = 1024, offsets=(0, 0, 1)) -> DropoutMasks:
    """Masks of one sequence-parallel layer call: attention as in make_masks; dropout1 / dropout2 drawn for
    a [s/n, b, h] shard (element index = position in the shard) and repeated over the n ranks."""
    from . import philox

    n = cfg.num_partitions
    m = make_masks(cfg, b, s, seed, offsets)
    gseed = rng_seeds(seed)["global_seed"]
    shard_elems = (s // n) * b * cfg.d_model
    if cfg.dropout > 0:
        sh = torch.from_numpy(philox.dropout_mask(gseed, offsets[1], shard_elems, cfg.dropout)).reshape(s // n, b, cfg.d_model)
        m.hidden1 = sp_tile_mask(sh, n)
    if cfg.act_dropout > 0:
        sh = torch.from_numpy(philox.dropout_mask(gseed, offsets[2], shard_elems, cfg.act_dropout)).reshape(s // n, b, cfg.d_model)
        m.hidden2 = sp_tile_mask(sh, n)
    return m
