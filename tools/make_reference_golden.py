"""Golden vectors from the REFERENCE'S OWN Python, executed in this container.

  python tools/make_reference_golden.py            # rewrites tests/golden/ref_*.npz
  python tools/make_reference_golden.py --check    # recomputes and compares with the committed files

The classes come unmodified from /root/reference (hybrid_model.py: TransformerDecoderLayer,
MultiHeadAttention, GPTModelHybrid, GPTForPretrainingHybrid, GPTPretrainingCriterionHybird;
sequence_parallel_utils.py: the sequence-parallel linears and operators); `paddle` is the torch-backed
stand-in of oracle/paddle_shim.py (read its header for what that does and does not pin). n > 1 runs n
simulated mp ranks. Everything is fp64 at hand-checkable sizes; dropout keep masks are drawn here, stored
in the fixture and logged with the RNG-tracker stream the reference drew them under.

This script needs /root/reference, so it runs only in the build container; the fixtures it writes are
what travels. It does not import the oracle: tests/test_reference_golden.py compares the oracle (and,
through it, the CUDA path) with these files.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import paddle_shim as S  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
F64 = torch.float64

LAYER_PARAMS = ("norm1.weight", "norm1.bias", "self_attn.qkv_proj.weight", "self_attn.qkv_proj.bias",
                "self_attn.out_proj.weight", "self_attn.out_proj.bias", "norm2.weight", "norm2.bias",
                "linear1.weight", "linear1.bias", "linear2.weight", "linear2.bias")


# ------------------------------------------------------------------------------------ helpers
def split_kind(name: str) -> str:
    """How a parameter of the hybrid model is laid out over mp ranks."""
    if name.endswith(("qkv_proj.weight", "linear1.weight")):
        return "col"       # [in, out/n]
    if name.endswith(("qkv_proj.bias", "linear1.bias")):
        return "colbias"   # [out/n]
    if name.endswith(("out_proj.weight", "linear2.weight")):
        return "row"       # [in/n, out]
    if name.endswith("word_embeddings.weight"):
        return "row"       # [vocab/n, h]
    return "rep"


def shard(name: str, t: torch.Tensor, n: int, r: int) -> torch.Tensor:
    k = split_kind(name)
    if k == "col":
        return t.chunk(n, dim=1)[r].contiguous()
    if k in ("colbias", "row"):
        return t.chunk(n, dim=0)[r].contiguous()
    return t


def unshard(name: str, parts):
    k = split_kind(name)
    if k == "col":
        return torch.cat(parts, dim=1)
    if k in ("colbias", "row"):
        return torch.cat(parts, dim=0)
    for p in parts[1:]:  # replicated parameters must end with identical gradients on every rank
        assert torch.allclose(p, parts[0], rtol=0, atol=1e-12), name
    return parts[0]


class MaskBook:
    """Keep masks per (stream, rank, call index). 'global_seed' is ONE state shared by all mp ranks
    (ppfleetx/utils/env.py:43: global_seed = seed + data_world_rank), and so is the default generator
    (env.py:38: paddle.seed(seed + data_world_rank)): their draws do not depend on the mp rank. 'local_seed'
    (env.py:42: + mp_rank * 10) differs per rank."""

    def __init__(self, seed: int):
        self.seed, self.lock = seed, threading.Lock()
        self.count, self.masks = {}, {}

    def __call__(self, stream, rank, shape, p):
        with self.lock:
            idx = self.count.get((stream, rank), 0)
            self.count[(stream, rank)] = idx + 1
            key = (stream, rank if stream == "local_seed" else -1, idx)
            if key not in self.masks:
                sid = {"global_seed": 1, "local_seed": 2, None: 3}[stream]
                g = torch.Generator().manual_seed(self.seed * 1000003 + sid * 10007 + (key[1] + 1) * 101 + idx)
                self.masks[key] = torch.rand(shape, generator=g) >= p
            m = self.masks[key]
            assert tuple(m.shape) == tuple(shape), (key, m.shape, shape)
            return m

    def get(self, stream, rank, idx):
        return self.masks[(stream, rank if stream == "local_seed" else -1, idx)]


def rand_layer_params(h, f, g, prefix=""):
    w = lambda i, o: torch.randn(i, o, generator=g, dtype=F64) * 0.02  # noqa: E731
    v = lambda n_: torch.randn(n_, generator=g, dtype=F64) * 0.02  # noqa: E731
    p = {"norm1.weight": 1 + 5 * v(h), "norm1.bias": v(h),
         "self_attn.qkv_proj.weight": w(h, 3 * h), "self_attn.qkv_proj.bias": v(3 * h),
         "self_attn.out_proj.weight": w(h, h), "self_attn.out_proj.bias": v(h),
         "norm2.weight": 1 + 5 * v(h), "norm2.bias": v(h),
         "linear1.weight": w(h, f), "linear1.bias": v(f), "linear2.weight": w(f, h), "linear2.bias": v(h)}
    return {prefix + k: t for k, t in p.items()}


def save(name, arrays, meta, check):
    arrays = {k: (v.detach().numpy() if isinstance(v, torch.Tensor) else np.asarray(v)) for k, v in arrays.items()}
    arrays["meta"] = np.array(json.dumps(meta, sort_keys=True))
    path = os.path.join(OUT, name + ".npz")
    if check:
        old = np.load(path)
        assert sorted(old.files) == sorted(arrays), (name, sorted(set(old.files) ^ set(arrays)))
        for k, v in arrays.items():
            if k == "meta":
                assert str(old[k]) == str(v), (name, str(old[k]), str(v))
            else:
                np.testing.assert_allclose(old[k], v, rtol=0, atol=1e-12, err_msg=f"{name}:{k}")
        print("ok  ", name)
    else:
        np.savez_compressed(path, **arrays)
        print("wrote", name, "%.1f KiB" % (os.path.getsize(path) / 1024))


# ------------------------------------------------------------------------------------ one decoder layer
def layer_case(name, check, *, b=2, s=8, h=16, nh=4, f=64, n=1, sp=False, dropout=0.1, pre_ln=True, seed=11):
    hm = S.load_reference("hybrid_model")
    g = torch.Generator().manual_seed(seed)
    params = rand_layer_params(h, f, g)
    x = torch.randn(b, s, h, generator=g, dtype=F64)
    dy = torch.randn(b, s, h, generator=g, dtype=F64)
    book = MaskBook(seed)
    S.set_dropout_provider(book)
    sl = s // n

    def rank_fn(r):
        layer = hm.TransformerDecoderLayer(h, nh, f, dropout=dropout, normalize_before=pre_ln, num_partitions=n,
                                           sequence_parallel=sp)
        layer.train()
        layer.set_state_dict({k: shard(k, v, n, r) for k, v in params.items()})
        if sp:  # this rank's sequence-first shard [s/n, b, h] (hybrid_model.py:612-617 produces this layout)
            xin = x[:, r * sl:(r + 1) * sl].transpose(0, 1).contiguous()
            dyin = dy[:, r * sl:(r + 1) * sl].transpose(0, 1).contiguous()
        else:
            xin, dyin = x, dy
        xt = S._to_tensor(xin, stop_gradient=False)
        y = layer(xt)
        y.backward(dyin)
        return (y.detach().as_subclass(torch.Tensor), xt.grad.as_subclass(torch.Tensor),
                {k: p.grad.clone() for k, p in layer.named_parameters()})

    res = S.run_ranks(n, rank_fn)
    if sp:
        y = torch.cat([r_[0] for r_ in res], dim=0).transpose(0, 1).contiguous()
        dx = torch.cat([r_[1] for r_ in res], dim=0).transpose(0, 1).contiguous()
    else:
        y, dx = res[0][0], res[0][1]
        for r_ in res[1:]:  # replicated activations: every rank must hold the same result
            assert torch.allclose(r_[0], y, rtol=0, atol=1e-12) and torch.allclose(r_[1], dx, rtol=0, atol=1e-12)
    arrays = {"in.x": x, "in.dy": dy, "out.y": y, "out.dx": dx}
    for k in LAYER_PARAMS:
        arrays["param." + k] = params[k]
        arrays["grad." + k] = unshard(k, [r_[2][k] for r_ in res])
    log = S.dropout_log()
    if dropout > 0:
        for r in range(n):
            arrays[f"mask.attn.{r}"] = book.get("local_seed", r, 0).to(torch.uint8)
        arrays["mask.hidden1"] = book.get("global_seed", 0, 0).to(torch.uint8)  # shard-shaped when sp
        arrays["mask.hidden2"] = book.get("global_seed", 0, 1).to(torch.uint8)
    meta = {"kind": "layer", "b": b, "s": s, "h": h, "nh": nh, "f": f, "n": n, "sequence_parallel": sp,
            "dropout": dropout, "normalize_before": pre_ln,
            "dropout_streams_rank0": [[st, list(shape)] for st, r, shape, _ in log if r == 0],
            "source": "hybrid_model.py TransformerDecoderLayer via oracle/paddle_shim.py"}
    save(name, arrays, meta, check)


# ------------------------------------------------------------------------------------ the model + criterion
def model_case(name, check, *, b=2, s=8, h=16, nh=4, f=64, vocab=64, layers=2, max_pos=16, n=1, sp=False,
               dropout=0.1, seed=23):
    hm = S.load_reference("hybrid_model")
    g = torch.Generator().manual_seed(seed)
    params = {"gpt.embeddings.word_embeddings.weight": torch.randn(vocab, h, generator=g, dtype=F64) * 0.5,
              "gpt.embeddings.position_embeddings.weight": torch.randn(max_pos, h, generator=g, dtype=F64) * 0.5}
    for i in range(layers):
        params.update(rand_layer_params(h, f, g, f"gpt.decoder.layers.{i}."))
    params["gpt.decoder.norm.weight"] = 1 + 0.1 * torch.randn(h, generator=g, dtype=F64)
    params["gpt.decoder.norm.bias"] = 0.02 * torch.randn(h, generator=g, dtype=F64)
    ids = torch.randint(0, vocab, (b, s), generator=g)
    labels = torch.randint(0, vocab, (b, s), generator=g)
    loss_mask = (torch.rand(b, s, generator=g) > 0.2).to(F64)
    book = MaskBook(seed)
    S.set_dropout_provider(book)

    def rank_fn(r):
        gpt = hm.GPTModelHybrid(vocab_size=vocab, hidden_size=h, num_layers=layers, num_attention_heads=nh,
                                ffn_hidden_size=f, hidden_dropout_prob=dropout, attention_probs_dropout_prob=dropout,
                                max_position_embeddings=max_pos, num_partitions=n, sequence_parallel=sp)
        model = hm.GPTForPretrainingHybrid(gpt)
        crit = hm.GPTPretrainingCriterionHybird()
        model.train()
        model.set_state_dict({k: shard(k, v, n, r) for k, v in params.items()})
        logits = model(S._to_tensor(ids))
        loss = crit(logits, S._to_tensor(labels), S._to_tensor(loss_mask))
        loss.backward()
        return (loss.detach().as_subclass(torch.Tensor), logits.detach().as_subclass(torch.Tensor),
                {k: p.grad.clone() for k, p in model.named_parameters()})

    res = S.run_ranks(n, rank_fn)
    loss = res[0][0]
    for r_ in res[1:]:
        assert torch.allclose(r_[0], loss, rtol=0, atol=1e-12)
    arrays = {"in.ids": ids, "in.labels": labels, "in.loss_mask": loss_mask, "out.loss": loss.reshape(1),
              "out.logits": torch.cat([r_[1] for r_ in res], dim=-1)}  # parallel_output=True: vocab split
    for k in params:
        arrays["param." + k] = params[k]
        arrays["grad." + k] = unshard(k, [r_[2][k] for r_ in res])
    log = S.dropout_log()
    if dropout > 0:
        emb_stream = "local_seed" if (sp and n > 1) else None
        for r in range(n if (sp and n > 1) else 1):
            arrays[f"mask.embedding.{r}"] = book.get(emb_stream, r, 0).to(torch.uint8)
        for i in range(layers):
            for r in range(n):
                idx = i + (1 if emb_stream == "local_seed" else 0)  # the embedding took this stream's first draw
                arrays[f"mask.layers.{i}.attn.{r}"] = book.get("local_seed", r, idx).to(torch.uint8)
            arrays[f"mask.layers.{i}.hidden1"] = book.get("global_seed", 0, 2 * i).to(torch.uint8)
            arrays[f"mask.layers.{i}.hidden2"] = book.get("global_seed", 0, 2 * i + 1).to(torch.uint8)
    meta = {"kind": "model", "b": b, "s": s, "h": h, "nh": nh, "f": f, "vocab": vocab, "layers": layers,
            "max_pos": max_pos, "n": n, "sequence_parallel": bool(sp and n > 1), "dropout": dropout,
            "dropout_streams_rank0": [[st, list(shape)] for st, r, shape, _ in log if r == 0],
            "source": "hybrid_model.py GPTForPretrainingHybrid + GPTPretrainingCriterionHybird via oracle/paddle_shim.py"}
    save(name, arrays, meta, check)


# ------------------------------------------------------------------------------------ eval + KV cache
def cache_case(name, check, *, b=2, s=8, prompt=5, h=16, nh=4, f=64, vocab=64, layers=2, max_pos=16, seed=37):
    """Eval mode (explicit -1e4 causal mask, hybrid_model.py:716-726): logits of the whole sequence at once, and
    of prompt + token-by-token decoding through use_cache / Cache (:179-186, :243-266, :697-705)."""
    hm = S.load_reference("hybrid_model")
    g = torch.Generator().manual_seed(seed)
    params = {"gpt.embeddings.word_embeddings.weight": torch.randn(vocab, h, generator=g, dtype=F64) * 0.5,
              "gpt.embeddings.position_embeddings.weight": torch.randn(max_pos, h, generator=g, dtype=F64) * 0.5}
    for i in range(layers):
        params.update(rand_layer_params(h, f, g, f"gpt.decoder.layers.{i}."))
    params["gpt.decoder.norm.weight"] = 1 + 0.1 * torch.randn(h, generator=g, dtype=F64)
    params["gpt.decoder.norm.bias"] = 0.02 * torch.randn(h, generator=g, dtype=F64)
    ids = torch.randint(0, vocab, (b, s), generator=g)
    S.set_dropout_provider(None)

    def rank_fn(r):
        gpt = hm.GPTModelHybrid(vocab_size=vocab, hidden_size=h, num_layers=layers, num_attention_heads=nh,
                                ffn_hidden_size=f, max_position_embeddings=max_pos)
        model = hm.GPTForPretrainingHybrid(gpt)
        model.eval()
        model.set_state_dict(params)
        with torch.no_grad():
            full = model(S._to_tensor(ids))
            step_logits, cache = model(S._to_tensor(ids[:, :prompt]), use_cache=True)
            outs = [step_logits]
            cache_len = [int(cache[0].k.shape[2])]
            for t in range(prompt, s):
                step_logits, cache = model(S._to_tensor(ids[:, t:t + 1]), use_cache=True, cache=cache)
                outs.append(step_logits)
                cache_len.append(int(cache[0].k.shape[2]))
            empty = gpt.decoder.layers[0].gen_cache(S._to_tensor(torch.zeros(b, 1, h, dtype=F64)))
        return (full.as_subclass(torch.Tensor), torch.cat(outs, dim=1).as_subclass(torch.Tensor), cache_len,
                cache[0].k.as_subclass(torch.Tensor), list(empty.k.shape))

    full, stepped, cache_len, k0, empty_shape = S.run_ranks(1, rank_fn)[0]
    arrays = {"in.ids": ids, "out.logits_full": full, "out.logits_stepped": stepped, "out.cache0_k": k0}
    for k in params:
        arrays["param." + k] = params[k]
    meta = {"kind": "cache", "b": b, "s": s, "prompt": prompt, "h": h, "nh": nh, "f": f, "vocab": vocab,
            "layers": layers, "max_pos": max_pos, "cache_len_after_each_call": cache_len,
            "gen_cache_shape": empty_shape,
            "source": "hybrid_model.py GPTForPretrainingHybrid in eval mode, use_cache, via oracle/paddle_shim.py"}
    save(name, arrays, meta, check)


def seeds_case(name, check):
    """ppfleetx/utils/env.py set_seed executed for mp ranks 0..3: the two tracker seeds per rank."""
    S.install("float64")
    env = S.load_reference("env")
    env.get_data_world_rank = lambda: 0  # dp = sharding = 1 in the simulation (env.py reads it from the hcg)
    S.rng_tracker().reset()

    def rank_fn(r):
        env.set_seed(1024)
        t = S.rng_tracker()
        return [t.seed_of("global_seed"), t.seed_of("local_seed")]

    import paddle.distributed as pdist
    out = S.run_ranks(4, rank_fn)
    assert pdist.get_world_size() == 1  # outside run_ranks
    save(name, {"seeds": np.array(out, dtype=np.int64)},
         {"kind": "seeds", "seed": 1024, "columns": ["global_seed", "local_seed"], "rows": "mp_rank 0..3",
          "source": "ppfleetx/utils/env.py set_seed via oracle/paddle_shim.py"}, check)


def lr_case(name, check):
    """ppfleetx/optims/lr_scheduler.py CosineAnnealingWithWarmupDecay.get_lr at pretrain_gpt_base.yaml:75-80
    (max 5e-5, min 1e-5, warm-up 1 % of 360000 decay steps), evaluated at chosen values of last_epoch."""
    lr = S.load_reference("lr_scheduler")
    sch = lr.CosineAnnealingWithWarmupDecay(5e-5, 1e-5, 0.01, 360000)
    steps = [0, 1, 1800, 3599, 3600, 3601, 100000, 181800, 359999, 360000, 360001, 500000]
    vals = []
    for k in steps:
        sch.last_epoch = k
        vals.append(sch.get_lr())
    save(name, {"steps": np.array(steps, dtype=np.int64), "lr": np.array(vals, dtype=np.float64)},
         {"kind": "lr", "max_lr": 5e-5, "min_lr": 1e-5, "warmup_rate": 0.01, "decay_steps": 360000,
          "source": "ppfleetx/optims/lr_scheduler.py via oracle/paddle_shim.py"}, check)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--check", action="store_true")
    a = ap.parse_args()
    if not S.reference_available():
        sys.exit("needs the reference tree at " + S.REFERENCE_ROOT)
    os.makedirs(OUT, exist_ok=True)
    S.install("float64")
    c = a.check
    layer_case("ref_layer_n1", c)
    layer_case("ref_layer_n1_nodrop", c, dropout=0.0)
    layer_case("ref_layer_postln_n1", c, pre_ln=False)
    layer_case("ref_layer_tp2", c, n=2)
    layer_case("ref_layer_tp4", c, n=4)
    layer_case("ref_layer_sp2", c, n=2, sp=True)
    layer_case("ref_layer_sp4", c, n=4, sp=True)
    layer_case("ref_layer_postln_tp2", c, n=2, pre_ln=False)
    layer_case("ref_layer_postln_sp2", c, n=2, sp=True, pre_ln=False)
    model_case("ref_model_n1", c)
    model_case("ref_model_tp2", c, n=2)
    model_case("ref_model_sp2", c, n=2, sp=True)
    cache_case("ref_model_eval_cache", c)
    seeds_case("ref_env_seeds", c)
    lr_case("ref_lr_schedule", c)


if __name__ == "__main__":
    main()
