"""FusedAdamW (ppfleetx/optims/optimizer.py:29-50) with ClipGradByGlobalNorm (pretrain_gpt_base.yaml:81-83):
the whole model is stepped by two kernel launches of libfleetx_b200.so (csrc/optim.cu) over device-resident
tensor / chunk tables; the clip factor never visits the host.

Same knobs as the reference: learning_rate, parameters, grad_clip (a clip norm or None), weight_decay, beta1,
beta2, epsilon, multi_precision (fp32 master weights for bf16 parameters, AMP-O2). Parameters whose name
contains "bias" or "norm" do not decay (optimizer.py:40-44). The reference tests `p.name`, which under Paddle is
the parameter's framework-assigned unique name, not its state-dict key (upstream behaviour, not checkable here);
this class tests whatever names it is given — `for_model` passes the state-dict keys ("linear1.bias",
"norm1.weight", ...), a caller that wants other names passes `named_parameters` itself. `tensor_fusion` is
accepted and irrelevant here: the update is already one launch for all tensors.
"""
from __future__ import annotations

import ctypes as C
from typing import Iterable, Optional

import torch

from . import _lib
from ._lib import AdamTensor, check

CHUNK = 1 << 16  # elements per block


class FusedAdamW:
    def __init__(self, parameters: Iterable, learning_rate: float = 5e-5, grad_clip: Optional[float] = 1.0,
                 weight_decay: float = 0.01, beta1: float = 0.9, beta2: float = 0.999, epsilon: float = 1e-8,
                 multi_precision: bool = True, tensor_fusion: bool = False, mp_group=None, lr: Optional[float] = None,
                 named_parameters=None):
        named = list(named_parameters) if named_parameters is not None else [(f"param_{i}", p) for i, p in enumerate(parameters)]
        self.params = [(n, p) for n, p in named if p.requires_grad]
        if not self.params:
            raise ValueError("FusedAdamW: no parameters")
        self.lr = float(lr if lr is not None else learning_rate)
        self.clip_norm = float(grad_clip) if grad_clip else 0.0
        self.weight_decay, self.beta1, self.beta2, self.eps = float(weight_decay), float(beta1), float(beta2), float(epsilon)
        self.mp_group = mp_group
        self.step_count = 0
        dev = self.params[0][1].device
        if dev.type != "cuda":
            raise _lib.FleetXError("FusedAdamW needs CUDA parameters (there is no CPU path)")
        self.device = dev
        self.state = []
        chunks = []
        for i, (name, p) in enumerate(self.params):
            if p.dtype not in (torch.bfloat16, torch.float32) or not p.is_contiguous():
                raise _lib.FleetXError(f"FusedAdamW: parameter {name} must be contiguous bf16 or fp32")
            master = p.detach().float().clone() if (multi_precision and p.dtype == torch.bfloat16) else None
            self.state.append(dict(m=torch.zeros(p.numel(), dtype=torch.float32, device=dev),
                                   v=torch.zeros(p.numel(), dtype=torch.float32, device=dev), master=master))
            chunks += [(i, c) for c in range((p.numel() + CHUNK - 1) // CHUNK)]
        self.n_chunks = len(chunks)
        self._chunks = torch.tensor(chunks, dtype=torch.int32, device=dev)
        self._table_host = (AdamTensor * len(self.params))()
        self._table_pinned = torch.empty(C.sizeof(self._table_host), dtype=torch.uint8).pin_memory()
        self._table_dev = torch.empty(C.sizeof(self._table_host), dtype=torch.uint8, device=dev)
        self._sq = torch.zeros(2, dtype=torch.float32, device=dev)
        self._copied = None
        for i, (name, p) in enumerate(self.params):
            st, t = self.state[i], self._table_host[i]
            t.param, t.m, t.v, t.n = p.data_ptr(), st["m"].data_ptr(), st["v"].data_ptr(), p.numel()
            t.master = st["master"].data_ptr() if st["master"] is not None else None
            t.decay = 0.0 if any(nd in name for nd in ("bias", "norm")) else self.weight_decay
            t.param_f32 = int(p.dtype == torch.float32)
            t.is_distributed = int(bool(getattr(p, "is_distributed", False)))

    @classmethod
    def for_model(cls, model: torch.nn.Module, **kw):
        """Decay is decided by parameter NAME, as in the reference: pass the model so the names are known."""
        return cls(None, named_parameters=model.named_parameters(), **kw)

    def set_lr(self, lr: float) -> None:
        self.lr = float(lr)

    def _upload_table(self) -> None:
        self._grad_keep = []  # contiguous copies of strided gradients stay alive until the next step's table is built
        for i, (name, p) in enumerate(self.params):
            if p.grad is None:
                raise _lib.FleetXError(f"FusedAdamW.step: parameter {name} has no gradient")
            g = p.grad
            if not g.is_contiguous():
                g = g.contiguous()
                self._grad_keep.append(g)  # the kernels read this block after this function returns
            self._table_host[i].grad = g.data_ptr()
            self._table_host[i].grad_f32 = int(g.dtype == torch.float32)
        if self._copied is not None:
            self._copied.synchronize()  # the previous step's upload has left the pinned staging buffer
        C.memmove(self._table_pinned.data_ptr(), C.addressof(self._table_host), C.sizeof(self._table_host))
        self._table_dev.copy_(self._table_pinned, non_blocking=True)
        self._copied = torch.cuda.Event()
        self._copied.record()

    def step(self) -> None:
        self._upload_table()
        self.step_count += 1
        lib, st = _lib.lib(), torch.cuda.current_stream().cuda_stream
        sq = None
        if self.clip_norm > 0:
            check(lib.fx_multi_sq_norm(st, self._table_dev.data_ptr(), self._chunks.data_ptr(), self.n_chunks, CHUNK,
                                       self._sq.data_ptr()), "fx_multi_sq_norm")
            if self.mp_group is not None and self.mp_group.nranks > 1:
                self.mp_group.all_reduce_(self._sq[:1])  # distributed parameters: sum over the mp ranks
            sq = self._sq.data_ptr()
        check(lib.fx_multi_adamw(st, self._table_dev.data_ptr(), self._chunks.data_ptr(), self.n_chunks, CHUNK, self.lr,
                                 self.beta1, self.beta2, self.eps, self.step_count, sq, self.clip_norm), "fx_multi_adamw")

    def clear_grad(self) -> None:
        for _, p in self.params:
            p.grad = None

    def global_grad_norm(self) -> float:
        """The norm the last step clipped against (host read: for logging and tests)."""
        return float(self._sq.sum().sqrt().item())


class CosineAnnealingWithWarmupDecay:
    """ppfleetx/optims/lr_scheduler.py:20-48 (host arithmetic only): linear warm-up over warmup_rate * decay_steps
    steps to max_lr, cosine decay to min_lr at decay_steps, min_lr afterwards. step() then get_lr(), as
    paddle's LRScheduler; last_epoch starts at 0."""

    def __init__(self, max_lr, min_lr, warmup_rate, decay_steps, last_epoch=0, verbose=False, **kwargs):
        self.decay_steps = decay_steps
        self.warmup_step = warmup_rate * decay_steps
        self.max_lr, self.min_lr = max_lr, min_lr
        self.last_epoch = last_epoch

    def get_lr(self) -> float:
        import math

        if self.warmup_step > 0 and self.last_epoch <= self.warmup_step:
            return float(self.max_lr) * self.last_epoch / self.warmup_step
        if self.last_epoch > self.decay_steps:
            return self.min_lr
        ratio = float(self.last_epoch - self.warmup_step) / float(self.decay_steps - self.warmup_step)
        return self.min_lr + 0.5 * (math.cos(math.pi * ratio) + 1.0) * (self.max_lr - self.min_lr)

    def step(self, epoch=None) -> None:
        self.last_epoch = self.last_epoch + 1 if epoch is None else epoch
