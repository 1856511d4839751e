"""The oracle against vectors produced by the REFERENCE'S OWN Python (tests/golden/ref_*.npz).

tools/make_reference_golden.py executed hybrid_model.py / sequence_parallel_utils.py / env.py from
/root/reference, unmodified, over the torch-backed `paddle` stand-in of oracle/paddle_shim.py (n > 1: n
simulated mp ranks) and committed inputs, masks, outputs and gradients in fp64. Here the oracle — the
restatement every GPU parity test is checked against — has to reproduce them to 1e-10, which pins the
order of operations, layouts, dropout placement / streams, tensor- and sequence-parallel semantics, the
model around the layer, eval mode and the KV cache to the reference's code. The Paddle primitives
themselves (its kernels) stay unexercised: Paddle cannot be installed here.

When /root/reference is present (the build container, not the GPU box) the fixtures are also recomputed
and compared, and the fixtures the GPU tests use are replayed through the reference's layer.
"""
import json
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import decoder_layer as O  # noqa: E402
from oracle import gpt_model as GM  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
TOL = 1e-10
HAVE_REFERENCE = os.path.isfile("/root/reference/ppfleetx/models/language_model/gpt/dygraph/hybrid_model.py")

LAYER_FIXTURES = ["ref_layer_n1", "ref_layer_n1_nodrop", "ref_layer_postln_n1", "ref_layer_tp2", "ref_layer_tp4",
                  "ref_layer_sp2", "ref_layer_sp4", "ref_layer_postln_tp2", "ref_layer_postln_sp2"]
MODEL_FIXTURES = ["ref_model_n1", "ref_model_tp2", "ref_model_sp2"]


def load(name):
    d = np.load(os.path.join(GOLD, name + ".npz"))
    meta = json.loads(str(d["meta"]))
    return {k: torch.from_numpy(d[k]) for k in d.files if k != "meta"}, meta


def close(a, b, what, tol=TOL):
    err = (a.double() - b.double()).abs().max().item()
    assert err <= tol, f"{what}: max abs error {err:.3e}"


def layer_masks(d, meta, prefix="mask."):
    if meta["dropout"] == 0:
        return None
    n = meta["n"]
    h1, h2 = d[prefix + "hidden1"], d[prefix + "hidden2"]
    if meta["sequence_parallel"]:  # one shard-shaped mask, the same on every rank (drawn under 'global_seed')
        h1, h2 = O.sp_tile_mask(h1, n), O.sp_tile_mask(h2, n)
    return O.DropoutMasks(attn=[d[f"{prefix}attn.{r}"] for r in range(n)], hidden1=h1, hidden2=h2)


@pytest.mark.parametrize("name", LAYER_FIXTURES)
def test_oracle_layer_matches_the_reference_layer(name):
    d, m = load(name)
    cfg = O.LayerConfig(d_model=m["h"], nhead=m["nh"], dim_feedforward=m["f"], dropout=m["dropout"],
                        normalize_before=m["normalize_before"], num_partitions=m["n"])
    params = {k: d["param." + k] for k in O.PARAM_NAMES}
    y, dx, grads = O.decoder_layer_fwd_bwd(d["in.x"], d["in.dy"], params, cfg, layer_masks(d, m))
    close(y, d["out.y"], "y")
    close(dx, d["out.dx"], "dx")
    for k in O.PARAM_NAMES:
        close(grads[k], d["grad." + k], "grad " + k)


@pytest.mark.parametrize("name", LAYER_FIXTURES)
def test_reference_draws_dropout_under_the_streams_the_library_uses(name):
    """hybrid_model.py:280 (attention: 'local_seed'), :546 / :556 (dropout1 / dropout2: 'global_seed'); with
    sequence parallelism the hidden masks have the SHARD's shape — every rank applies the same mask to its
    own rows, the quirk oracle.make_masks_sp restates."""
    d, m = load(name)
    streams = m["dropout_streams_rank0"]
    if m["dropout"] == 0:
        assert streams == []
        return
    b, s, h, nh, n = m["b"], m["s"], m["h"], m["nh"], m["n"]
    hidden = [s // n, b, h] if m["sequence_parallel"] else [b, s, h]
    assert streams == [["local_seed", [b, nh // n, s, s]], ["global_seed", hidden], ["global_seed", hidden]]


@pytest.mark.parametrize("name", MODEL_FIXTURES)
def test_oracle_model_matches_the_reference_model(name):
    d, m = load(name)
    n, L_ = m["n"], m["layers"]
    cfg = GM.ModelConfig(vocab_size=m["vocab"], hidden_size=m["h"], num_layers=L_, num_attention_heads=m["nh"],
                         ffn_hidden_size=m["f"], max_position_embeddings=m["max_pos"],
                         hidden_dropout_prob=m["dropout"], attention_probs_dropout_prob=m["dropout"], num_partitions=n)
    params = {k[len("param."):]: v for k, v in d.items() if k.startswith("param.")}
    lm = dict(m, sequence_parallel=m["sequence_parallel"])
    if m["sequence_parallel"]:  # per-rank [s/n, b, h] shards under each rank's 'local_seed' (hybrid_model.py:614-619)
        emb = O.sp_gather([d[f"mask.embedding.{r}"] for r in range(n)])
    else:
        emb = d["mask.embedding.0"]
    masks = GM.ModelMasks(embedding=emb, layers=[layer_masks(d, lm, f"mask.layers.{i}.") for i in range(L_)])
    ids, labels, loss_mask = d["in.ids"], d["in.labels"], d["in.loss_mask"]
    logits = GM.gpt_logits(ids, params, cfg, masks)
    close(logits, d["out.logits"], "logits")
    loss, grads = GM.loss_and_grads(ids, labels, loss_mask, params, cfg, masks)
    close(loss, d["out.loss"][0], "loss")
    for k, g in grads.items():
        close(g, d["grad." + k], "grad " + k)
    # where the reference draws each mask: embedding under no tracker stream (sequence parallel: 'local_seed',
    # on the shard), then per layer attention / dropout1 / dropout2 as in the layer test
    streams = [st for st, _ in m["dropout_streams_rank0"]]
    assert streams == (["local_seed" if m["sequence_parallel"] else None] + ["local_seed", "global_seed", "global_seed"] * L_)


def test_oracle_eval_mode_and_kv_cache_match_the_reference():
    d, m = load("ref_model_eval_cache")
    cfg = GM.ModelConfig(vocab_size=m["vocab"], hidden_size=m["h"], num_layers=m["layers"], num_attention_heads=m["nh"],
                         ffn_hidden_size=m["f"], max_position_embeddings=m["max_pos"])
    params = {k[len("param."):]: v for k, v in d.items() if k.startswith("param.")}
    logits = GM.gpt_logits(d["in.ids"], params, cfg, None, training=False)
    close(logits, d["out.logits_full"], "eval logits")
    # the reference's own cached decoding (prompt, then one token per call) equals its full causal forward:
    # what tests/test_kvcache_gpu.py demands of the CUDA path
    close(d["out.logits_stepped"], d["out.logits_full"], "reference: stepped vs full", 1e-12)
    b, s, p = m["b"], m["s"], m["prompt"]
    assert m["cache_len_after_each_call"] == list(range(p, s + 1))
    assert m["gen_cache_shape"] == [b, m["nh"], 0, m["h"] // m["nh"]]  # Cache(k, v): [b, nh, L, hd], L = 0
    assert list(d["out.cache0_k"].shape) == [b, m["nh"], s, m["h"] // m["nh"]]


def test_rng_seeds_match_env_py_executed():
    d, m = load("ref_env_seeds")
    for r in range(4):
        sd = O.rng_seeds(m["seed"], mp_rank=r)
        assert [sd["global_seed"], sd["local_seed"]] == d["seeds"][r].tolist()
    from fleetx_b200 import rng

    for r in range(4):
        rng.set_seed(m["seed"], mp_rank=r)
        st = rng.get_rng_state_tracker().get_states_tracker()
        assert [st["global_seed"][0], st["local_seed"][0]] == d["seeds"][r].tolist()


def test_lr_schedule_matches_lr_scheduler_py_executed():
    d, m = load("ref_lr_schedule")
    from fleetx_b200.optim import CosineAnnealingWithWarmupDecay

    sch = CosineAnnealingWithWarmupDecay(m["max_lr"], m["min_lr"], m["warmup_rate"], m["decay_steps"])
    for k, want in zip(d["steps"].tolist(), d["lr"].tolist()):
        sch.step(k)
        assert sch.get_lr() == want, (k, sch.get_lr(), want)


# ------------------------------------------------------------------------------------ live (build container only)
@pytest.mark.skipif(not HAVE_REFERENCE, reason="the reference tree is not on this machine")
def test_committed_fixtures_are_what_the_reference_produces_now():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "make_reference_golden.py"), "--check"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]


@pytest.mark.skipif(not HAVE_REFERENCE, reason="the reference tree is not on this machine")
@pytest.mark.parametrize("name", ["layer_gpu_small_fp32", "layer_gpu_small_fp32_dropout", "oracle_tiny_fp64_n1"])
def test_fixtures_of_the_gpu_tests_replayed_through_the_reference_layer(name):
    """tests/test_golden_gpu.py compares the CUDA path with layer_gpu_small_*.npz (written by the oracle). Run
    the reference's TransformerDecoderLayer on the same inputs, parameters and masks: same numbers."""
    from oracle import paddle_shim as S

    z = np.load(os.path.join(GOLD, name + ".npz"))
    b, s, h, nh, f, n, drop_milli, seed = [int(v) for v in z["cfg"]]
    assert n == 1
    p = drop_milli / 1000.0
    S.install("float64")
    hm = S.load_reference("hybrid_model")
    cfg = O.LayerConfig(d_model=h, nhead=nh, dim_feedforward=f, dropout=p)
    masks = O.make_masks(cfg, b, s, seed) if p > 0 else None
    order = iter([masks.attn[0], masks.hidden1, masks.hidden2]) if masks else iter([])
    S.set_dropout_provider(lambda stream, rank, shape, p_: next(order))

    def rank_fn(r):
        layer = hm.TransformerDecoderLayer(h, nh, f, dropout=p)
        layer.train()
        layer.set_state_dict({k: torch.from_numpy(z["param." + k]).double() for k in O.PARAM_NAMES})
        x = S._to_tensor(torch.from_numpy(z["x"]).double(), stop_gradient=False)
        y = layer(x)
        y.backward(torch.from_numpy(z["dy"]).double())
        return y.detach(), x.grad, {k: q.grad for k, q in layer.named_parameters()}

    y, dx, grads = S.run_ranks(1, rank_fn)[0]
    tol = 1e-10 if z["y"].dtype == np.float64 else 2e-5  # the fp32 fixtures carry fp32 rounding
    close(y, torch.from_numpy(z["y"]), "y", tol)
    close(dx, torch.from_numpy(z["dx"]), "dx", tol)
    for k in O.PARAM_NAMES:
        g = torch.from_numpy(z["grad." + k])
        close(grads[k], g, "grad " + k, tol * max(1.0, g.abs().max().item()))


@pytest.mark.skipif(not HAVE_REFERENCE, reason="the reference tree is not on this machine")
def test_single_card_statement_of_the_layer_agrees_with_the_hybrid_one():
    """single_model.py:340-427 is the reference's own 1-GPU statement of the same layer (nn.Linear instead of the
    mp linears, no RNG tracker): bit-identical to the hybrid layer's fixture on the same inputs and masks."""
    from oracle import paddle_shim as S

    d, m = load("ref_layer_n1")
    S.install("float64")
    sm = S.load_reference("single_model")
    order = iter([d["mask.attn.0"], d["mask.hidden1"], d["mask.hidden2"]])
    S.set_dropout_provider(lambda stream, rank, shape, p: next(order))

    def rank_fn(r):
        layer = sm.TransformerDecoderLayer(m["h"], m["nh"], m["f"], dropout=m["dropout"])
        layer.train()
        layer.set_state_dict({k: d["param." + k] for k in O.PARAM_NAMES})
        x = S._to_tensor(d["in.x"], stop_gradient=False)
        y = layer(x, None)
        y.backward(d["in.dy"])
        return y.detach(), x.grad, {k: q.grad for k, q in layer.named_parameters()}

    y, dx, grads = S.run_ranks(1, rank_fn)[0]
    assert [st for st, _, _, _ in S.dropout_log()] == [None, None, None]  # no tracker on a single card
    close(y, d["out.y"], "y", 0.0)
    close(dx, d["out.dx"], "dx", 0.0)
    for k in O.PARAM_NAMES:
        close(grads[k], d["grad." + k], "grad " + k, 0.0)


@pytest.mark.skipif(not HAVE_REFERENCE, reason="the reference tree is not on this machine")
def test_host_mirror_has_the_reference_signatures_and_parameter_names():
    """The drop-in boundary: every class of the host mirror takes the reference class's arguments, in
    order, with the same defaults (a mirror may add trailing keyword arguments — device, dtype, mp_group — and may
    give a default where the reference has none), and holds parameters of the same names and shapes."""
    import inspect

    from oracle import paddle_shim as S

    S.install("float64")
    hm, spu = S.load_reference("hybrid_model"), S.load_reference("sequence_parallel_utils")
    import fleetx_b200.layers as Lm
    import fleetx_b200.model as Mm

    pairs = [("MultiHeadAttention", hm, Lm), ("TransformerDecoderLayer", hm, Lm), ("TransformerDecoder", hm, Mm),
             ("GPTEmbeddings", hm, Mm), ("GPTModelHybrid", hm, Mm), ("GPTForPretrainingHybrid", hm, Mm),
             ("GPTPretrainingCriterionHybird", hm, Mm), ("ColumnSequenceParallelLinear", spu, Lm),
             ("RowSequenceParallelLinear", spu, Lm)]
    empty = inspect.Parameter.empty
    for name, ref_mod, my_mod in pairs:
        for meth in ("__init__", "forward"):
            ref = list(inspect.signature(getattr(getattr(ref_mod, name), meth)).parameters.values())
            mine = list(inspect.signature(getattr(getattr(my_mod, name), meth)).parameters.values())
            assert [p.name for p in mine[:len(ref)]] == [p.name for p in ref], (name, meth)
            for r, m in zip(ref, mine):
                assert r.default is empty or r.default == m.default, (name, meth, r.name, r.default, m.default)
            assert all(p.default is not empty for p in mine[len(ref):]), (name, meth)
    # the ops ScatterOp / GatherOp / AllGatherOp / ReduceScatterOp and the hook exist under the reference's names
    for op in ("ScatterOp", "GatherOp", "AllGatherOp", "ReduceScatterOp", "all_reduce_gradient_hook"):
        assert hasattr(spu, op) and (hasattr(Lm, op) or hasattr(__import__("fleetx_b200.dist", fromlist=[op]), op)), op

    kw = dict(vocab_size=64, hidden_size=128, num_layers=2, num_attention_heads=2, ffn_hidden_size=512,
              max_position_embeddings=32)
    ref_model = S.run_ranks(1, lambda r: hm.GPTForPretrainingHybrid(hm.GPTModelHybrid(**kw)))[0]
    my_model = Mm.GPTForPretrainingHybrid(Mm.GPTModelHybrid(**kw, device="cpu"))
    ref_shapes = {k: tuple(v.shape) for k, v in ref_model.named_parameters()}
    my_shapes = {k: tuple(v.shape) for k, v in my_model.named_parameters()}
    assert ref_shapes == my_shapes


@pytest.mark.skipif(not HAVE_REFERENCE, reason="the reference tree is not on this machine")
def test_baseline_config_0_through_the_reference_layer():
    """BASELINE.json configs[0]: GPT-345M decoder layer (h 1024, 16 heads, ffn 4096), seq 1024, micro-batch 1, fp32,
    forward + backward, dropout 0.1 with the library's Philox masks: the reference's layer (executed over the
    stand-in, torch-CPU fp32) and the oracle agree to fp32 rounding."""
    from oracle import paddle_shim as S

    b, s, h, nh, f, seed = 1, 1024, 1024, 16, 4096, 1024
    cfg = O.LayerConfig(d_model=h, nhead=nh, dim_feedforward=f, dropout=0.1)
    params = O.init_params(cfg, seed, dtype=torch.float32)
    g = torch.Generator().manual_seed(seed + 1)
    x = torch.randn(b, s, h, generator=g)
    dy = torch.randn(b, s, h, generator=g)
    masks = O.make_masks(cfg, b, s, seed)
    y_o, dx_o, g_o = O.decoder_layer_fwd_bwd(x, dy, params, cfg, masks)

    S.install("float32")
    try:
        hm = S.load_reference("hybrid_model")
        order = iter([masks.attn[0], masks.hidden1, masks.hidden2])
        S.set_dropout_provider(lambda stream, rank, shape, p: next(order))

        def rank_fn(r):
            layer = hm.TransformerDecoderLayer(h, nh, f, dropout=0.1)
            layer.train()
            layer.set_state_dict(params)
            xt = S._to_tensor(x, stop_gradient=False)
            y = layer(xt)
            y.backward(dy)
            return y.detach(), xt.grad, {k: q.grad for k, q in layer.named_parameters()}

        y, dx, grads = S.run_ranks(1, rank_fn)[0]
    finally:
        S.install("float64")
    rel = lambda a, b_: ((a.double() - b_.double()).abs().max() / b_.double().abs().max()).item()  # noqa: E731
    assert rel(y, y_o) < 1e-5 and rel(dx, dx_o) < 1e-4
    for k in O.PARAM_NAMES:
        assert rel(grads[k], g_o[k]) < 1e-4, k


@pytest.mark.skipif(not HAVE_REFERENCE, reason="the reference tree is not on this machine")
def test_auto_parallel_statement_of_the_layer_agrees_after_unfusing_qkv():
    """auto_model.py:327-420 is the third in-tree statement of the layer, with SEPARATE q / k / v projections. Cut
    the fused qkv_proj of the hybrid fixture into them — columns of head i are [q_i | k_i | v_i], hybrid_model.py:
    167-173 — and it reproduces the hybrid layer's outputs and gradients: the packed layout the CUDA attention
    kernels read (fx_attn_causal_fwd: [b, s, nh, 3, hd]) is the reference's."""
    from oracle import paddle_shim as S

    d, m = load("ref_layer_n1")
    h, nh, f = m["h"], m["nh"], m["f"]
    hd = h // nh
    S.install("float64")
    am = S.load_reference("auto_model")

    class Mesh:  # auto-parallel annotations are no-ops here
        mp = dp = None

        def __getitem__(self, i):
            return None

    w = d["param.self_attn.qkv_proj.weight"].reshape(h, nh, 3, hd)
    bq = d["param.self_attn.qkv_proj.bias"].reshape(nh, 3, hd)
    sd = {k: d["param." + k] for k in O.PARAM_NAMES if "qkv_proj" not in k}
    for j, name in enumerate(("q_proj", "k_proj", "v_proj")):
        sd[f"self_attn.{name}.weight"] = w[:, :, j, :].reshape(h, h).contiguous()
        sd[f"self_attn.{name}.bias"] = bq[:, j, :].reshape(h).contiguous()
    order = iter([d["mask.attn.0"], d["mask.hidden1"], d["mask.hidden2"]])
    S.set_dropout_provider(lambda stream, rank, shape, p: next(order))

    def rank_fn(r):
        layer = am.TransformerDecoderLayer(h, nh, f, dropout=m["dropout"], mesh=Mesh(), mesh_idx=0)
        layer.train()
        layer.set_state_dict(sd)
        x = S._to_tensor(d["in.x"], stop_gradient=False)
        y = layer(x, None)
        y.backward(d["in.dy"])
        return y.detach(), x.grad, {k: q.grad for k, q in layer.named_parameters()}

    y, dx, grads = S.run_ranks(1, rank_fn)[0]
    close(y, d["out.y"], "y")
    close(dx, d["out.dx"], "dx")
    for k in O.PARAM_NAMES:
        if "qkv_proj" not in k:
            close(grads[k], d["grad." + k], "grad " + k)
    gw = torch.stack([grads[f"self_attn.{n}.weight"].reshape(h, nh, hd) for n in ("q_proj", "k_proj", "v_proj")], dim=2)
    gb = torch.stack([grads[f"self_attn.{n}.bias"].reshape(nh, hd) for n in ("q_proj", "k_proj", "v_proj")], dim=1)
    close(gw.reshape(h, 3 * h), d["grad.self_attn.qkv_proj.weight"], "grad qkv weight (re-fused)")
    close(gb.reshape(3 * h), d["grad.self_attn.qkv_proj.bias"], "grad qkv bias (re-fused)")
