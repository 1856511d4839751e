"""GPT-345M full model (BASELINE.json configs[1]: 24 layers, h 1024, 16 heads, seq 1024) forward + backward on
one GPU through fleetx_b200.model: tokens/s and model FLOP utilisation. No optimizer step
unless --adamw is given."""
import argparse, json, os, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fleetx_b200.model import GPTForPretrainingHybrid, GPTModelHybrid, GPTPretrainingCriterionHybird
from fleetx_b200.rng import set_seed

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=8)
ap.add_argument("--steps", type=int, default=10)
ap.add_argument("--adamw", action="store_true")
args = ap.parse_args()
V, H, NL, NH, FFN, S = 50304, 1024, 24, 16, 4096, 1024
b = args.batch
set_seed(1024)
torch.manual_seed(0)
model = GPTForPretrainingHybrid(GPTModelHybrid(vocab_size=V, hidden_size=H, num_layers=NL, num_attention_heads=NH,
                                               ffn_hidden_size=FFN, max_position_embeddings=S))
model.train()
crit = GPTPretrainingCriterionHybird()
opt = None
if args.adamw:
    from fleetx_b200.optim import FusedAdamW
    opt = FusedAdamW.for_model(model, lr=5e-5, weight_decay=0.01, grad_clip=1.0)
ids = torch.randint(0, V, (b, S), device="cuda")
labels = torch.randint(0, V, (b, S), device="cuda")
mask = torch.ones(b, S, device="cuda")

def step():
    loss = crit(model(ids), labels, mask)
    loss.backward()
    if opt is not None:
        opt.step()
        opt.clear_grad()
    else:
        for p in model.parameters():
            p.grad = None
    return loss

for _ in range(3):
    l = step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(args.steps):
    l = step()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / args.steps
T = b * S
nparams = sum(p.numel() for p in model.parameters())
# 6·N·T for the dense parameters (the tied embedding counted once as the logits GEMM) + causal attention
flops = 6.0 * (nparams - S * H) * T + NL * 6.0 * T * S * H
print(json.dumps({"workload": f"GPT-345M full model fwd+bwd{' + AdamW' if opt else ''}, seq {S}, batch {b}, bf16, 1 GPU",
                  "ms_per_step": round(ms, 3), "tokens_per_s": round(T / ms * 1e3, 1), "loss": round(l.item(), 4),
                  "params": nparams, "model_tflops": round(flops / ms / 1e9, 1)}))
