"""CPU restatement of the optimizer step — TEST INFRASTRUCTURE ONLY. PARITY UNPINNED (the
arithmetic is paddle.optimizer.AdamW and paddle.nn.ClipGradByGlobalNorm, not installable here).

ppfleetx/optims/optimizer.py:29-50: FusedAdamW = paddle AdamW with apply_decay_param_fun (no decay for names
containing "bias" or "norm"); pretrain_gpt_base.yaml:69-84: beta1 0.9, beta2 0.999, epsilon 1e-8,
weight_decay 0.01, ClipGradByGlobalNorm(1.0). Paddle's AdamW, per parameter with gradient g at step t:
    g   <- g * clip_norm / max(global_norm, clip_norm),  global_norm = sqrt(sum over ALL parameters of |g|^2)
    p   <- p * (1 - lr * coeff)                          (decoupled weight decay, decaying parameters only)
    m   <- b1 m + (1 - b1) g ;  v <- b2 v + (1 - b2) g^2
    p   <- p - lr / (1 - b1^t) * m / (sqrt(v) / sqrt(1 - b2^t) + eps)
"""
from __future__ import annotations

from typing import Dict

import torch


def adamw_step(params: Dict[str, torch.Tensor], grads: Dict[str, torch.Tensor], state: Dict[str, dict], t: int, lr: float,
               weight_decay: float = 0.01, beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-8,
               clip_norm: float = 1.0) -> float:
    """In-place fp64/fp32 update of `params`; `state[name] = {"m": ..., "v": ...}`. Returns the global norm."""
    gn = torch.sqrt(sum((g.double() ** 2).sum() for g in grads.values())).item()
    scale = clip_norm / max(gn, clip_norm) if clip_norm > 0 else 1.0
    for name, p in params.items():
        g = grads[name].to(p.dtype) * scale
        st = state.setdefault(name, {"m": torch.zeros_like(p), "v": torch.zeros_like(p)})
        if not any(nd in name for nd in ("bias", "norm")):
            p.mul_(1 - lr * weight_decay)
        st["m"].mul_(beta1).add_(g, alpha=1 - beta1)
        st["v"].mul_(beta2).addcmul_(g, g, value=1 - beta2)
        denom = st["v"].sqrt() / (1 - beta2 ** t) ** 0.5 + eps
        p.sub_(lr / (1 - beta1 ** t) * st["m"] / denom)
    return gn
