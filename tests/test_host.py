"""Host-side logic that needs no GPU: RNG tracker, sharding helpers, layer construction / API."""
import pytest
import torch

from fleetx_b200 import dist as D
from fleetx_b200.rng import RNGStatesTracker, get_rng_state_tracker, set_seed
from oracle import decoder_layer as O


def test_tracker_streams_and_errors():
    t = RNGStatesTracker()
    t.add("global_seed", 1024)
    t.add("local_seed", 1147)
    with pytest.raises(ValueError):
        t.add("global_seed", 5)       # duplicate name
    with pytest.raises(ValueError):
        t.add("other", 1024)          # duplicate seed
    with pytest.raises(ValueError):
        with t.rng_state("missing"):
            pass
    with t.rng_state("global_seed"):
        assert t.draw() == (1024, 0)
        assert t.draw() == (1024, 1)
    assert t.draw("local_seed") == (1147, 0)
    state = t.get_states_tracker()
    a = t.draw("global_seed")
    t.set_states_tracker(state)       # recompute: restoring the state replays the same draws
    assert t.draw("global_seed") == a


def test_set_seed_matches_env_py_and_oracle():
    for mp_rank in (0, 3):
        set_seed(1024, mp_rank=mp_rank, pp_rank=1, data_world_rank=2)
        st = get_rng_state_tracker().get_states_tracker()
        want = O.rng_seeds(1024, mp_rank, 1, 2)
        assert st["global_seed"][0] == want["global_seed"] and st["local_seed"][0] == want["local_seed"]


def test_sharding_helpers_match_oracle_shards():
    cfg = O.LayerConfig(d_model=64, nhead=8, dim_feedforward=256, num_partitions=4)
    full = O.init_params(cfg)
    for r in range(4):
        ref = O.shard_params(full, cfg, r)
        assert torch.equal(D.shard_columns(full["self_attn.qkv_proj.weight"], r, 4), ref["self_attn.qkv_proj.weight"])
        assert torch.equal(D.shard_columns(full["linear1.bias"], r, 4), ref["linear1.bias"])
        assert torch.equal(D.shard_rows(full["linear2.weight"], r, 4), ref["linear2.weight"])
    with pytest.raises(ValueError):
        D.shard_columns(torch.zeros(4, 6), 0, 4)


def test_layer_api_surface_on_cpu():
    from fleetx_b200.layers import ColumnParallelLinear, MultiHeadAttention, RowParallelLinear, TransformerDecoderLayer

    layer = TransformerDecoderLayer(128, 2, 512, dropout=0.1, device="cpu")
    names = sorted(n for n, _ in layer.named_parameters())
    assert names == sorted(O.PARAM_NAMES)                      # checkpoint-compatible names
    shapes = {n: tuple(p.shape) for n, p in layer.named_parameters()}
    assert shapes["self_attn.qkv_proj.weight"] == (128, 384)   # Paddle [in, out] layout
    assert shapes["linear2.weight"] == (512, 128)
    assert layer.norm1.weight.dtype == torch.float32 and layer.linear1.weight.dtype == torch.bfloat16
    assert layer._config["d_model"] == 128 and layer.self_attn.num_heads == 2 and layer.self_attn.head_dim == 64
    with pytest.raises(ValueError):
        TransformerDecoderLayer(128, 2, 512, num_partitions=2, device="cpu")   # group size is 1
    assert TransformerDecoderLayer(128, 2, 512, sequence_parallel=True, device="cpu").sequence_parallel
    assert not TransformerDecoderLayer(128, 2, 512, sequence_parallel=True, normalize_before=False,
                                       device="cpu").normalize_before  # post-LN sequence parallel is built too
    from fleetx_b200.layers import ColumnSequenceParallelLinear, RowSequenceParallelLinear
    with pytest.raises(AssertionError):   # sequence_parallel_utils.py:171-172, :262-263
        ColumnSequenceParallelLinear(16, 32, has_bias=True, gather_output=True, device="cpu")
    with pytest.raises(AssertionError):
        RowSequenceParallelLinear(32, 16, input_is_parallel=False, device="cpu")
    assert RowSequenceParallelLinear(32, 16, input_is_parallel=True, device="cpu").bias.shape == (16,)
    with pytest.raises(AssertionError):
        MultiHeadAttention(100, 3, device="cpu")
    assert ColumnParallelLinear(16, 32, has_bias=True, gather_output=False, device="cpu").bias.shape == (32,)
    assert RowParallelLinear(32, 16, input_is_parallel=True, device="cpu").weight.shape == (32, 16)
    # load_full_state keeps this rank's shard
    cfg = O.LayerConfig(d_model=128, nhead=2, dim_feedforward=512)
    full = O.init_params(cfg)
    layer.load_full_state(full)
    assert torch.equal(layer.linear1.weight.float(), full["linear1.weight"].to(torch.bfloat16).float())


def test_bench_flop_model_matches_survey():
    import bench

    # 327.2 MFLOP/token for 1.3B/s2048, 81.8 for 345M/s1024
    assert abs(bench.layer_flops(1, 2048, 2048) / 1e6 - 327.2) < 0.1
    assert abs(bench.layer_flops(1, 1024, 1024) / 1e6 - 81.8) < 0.1


def test_lr_scheduler_matches_the_reference_schedule():
    """lr_scheduler.py:20-48 at pretrain_gpt_base.yaml:75-80 (max 5e-5, min 1e-5, warm-up 1 % of 360000)."""
    import math

    from fleetx_b200.optim import CosineAnnealingWithWarmupDecay

    sch = CosineAnnealingWithWarmupDecay(5e-5, 1e-5, 0.01, 360000)
    assert sch.get_lr() == 0.0
    sch.step(1800)
    assert abs(sch.get_lr() - 2.5e-5) < 1e-12          # half way through the warm-up
    sch.step(3600)
    assert abs(sch.get_lr() - 5e-5) < 1e-12
    sch.step(3600 + (360000 - 3600) // 2)
    assert abs(sch.get_lr() - 3e-5) < 1e-9             # middle of the cosine
    sch.step(360000)
    assert abs(sch.get_lr() - (1e-5 + 0.5 * (math.cos(math.pi) + 1) * 4e-5)) < 1e-12
    sch.step(360001)
    assert sch.get_lr() == 1e-5


def test_split_k_proposal_and_cta_budget_are_host_arithmetic():
    """fx_gemm_splitk_splits / fx_set_sm_budget touch no device memory: the weight-gradient shapes of mp 8 get a
    split, the 1-GPU shapes do not; the budget is per thread and returns the previous value."""
    from fleetx_b200 import _lib

    lib = _lib.lib()
    try:
        one = [lib.fx_gemm_splitk_splits(*shp) for shp in ((2048, 8192, 16384), (8192, 2048, 16384), (2048, 6144, 16384))]
    except Exception:  # pragma: no cover
        pytest.skip("needs the CUDA runtime's device count")
    assert one == [1, 1, 1]                                   # >= one wave of tiles: no split
    assert lib.fx_gemm_splitk_splits(256, 2048, 16384) > 1    # dW_out at mp 8: 16 tiles
    assert lib.fx_gemm_splitk_splits(2048, 768, 16384) > 1    # dW_qkv at mp 8: 48 tiles
    assert lib.fx_gemm_splitk_splits(256, 2048, 2048) == 1    # short K: not worth the partial tiles
    assert lib.fx_gemm_splitk_workspace_bytes(256, 2048, 4) == 4 * 256 * 2048 * 4
    assert lib.fx_gemm_splitk_workspace_bytes(256, 2048, 1) == 0
    old = lib.fx_set_sm_budget(100)
    assert lib.fx_set_sm_budget(old) == 100


def test_chunk_helpers_and_causal_mask_acceptance():
    from fleetx_b200 import functional as Fn
    from fleetx_b200.layers import accept_causal_mask

    buf = torch.arange(24, dtype=torch.float32).view(6, 4)
    parts = [buf[0:2], buf[2:4], buf[4:6]]
    cat = Fn._rows_cat(parts)
    assert cat.data_ptr() == buf.data_ptr() and torch.equal(cat, buf)          # back-to-back views: no copy
    assert torch.equal(Fn._rows_cat([buf[0:2], buf[4:6]]), torch.cat([buf[0:2], buf[4:6]]))
    lse = torch.arange(2 * 3 * 5, dtype=torch.float32).view(2, 3, 5)
    assert torch.equal(Fn._batch_slice(lse, 2, 1, 1), lse[1:2])
    flat = torch.arange(8, dtype=torch.uint8)
    assert torch.equal(Fn._batch_slice(flat, 2, 1, 1), flat[4:])
    assert Fn._batch_slice(["a", "b"], 2, 1, 1) == "b" and Fn._batch_slice(None, 2, 0, 1) is None

    class G:
        pipeline_chunks = 2
    assert Fn._chunks(8, G()) == 2 and Fn._chunks(3, G()) == 1 and Fn._chunks(8, object()) == 1

    s = 6
    causal = torch.triu(torch.ones(s, s) * -1e4, diagonal=1)
    accept_causal_mask(None, s, "t")
    accept_causal_mask(causal, s, "t")
    accept_causal_mask(causal[None, None], s, "t")
    bad = causal.clone()
    bad[3, 1] = -1e4
    for m in (bad, causal[:, :4], torch.zeros(s, s)):
        with pytest.raises(NotImplementedError):
            accept_causal_mask(m, s, "t")
