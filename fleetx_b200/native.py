"""The whole decoder layer through the two native entry points fx_decoder_layer_fwd / _bwd.

Same kernels, same order and the same results as fleetx_b200.functional (which sequences them from
Python and overlaps the backward allreduces with the weight-gradient GEMMs); this is the call a host
without Python makes — one C call per direction, every buffer planned up front, nothing allocated.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib
from ._lib import LayerDesc, LayerGrads, TpComm, check
from .functional import LayerMeta, LayerParams

_FIELDS = ("norm1_w", "norm1_b", "qkv_w", "qkv_b", "out_w", "out_b", "norm2_w", "norm2_b", "lin1_w", "lin1_b",
           "lin2_w", "lin2_b")


class NativeLayer:
    """Plans the saved-activation and workspace buffers for one (shape, parameters) pair and runs the layer.

    group: None / SingleGroup for one rank, or a fleetx_b200.dist.SymmetricTPGroup (its symmetric buffer,
    flag pages and call counter are handed to the library)."""

    def __init__(self, b: int, s: int, P: LayerParams, M: LayerMeta, group=None):
        if M.mask_h1 is not None or M.mask_h2 is not None:
            raise _lib.FleetXError("the native layer entry draws its dropout masks from (seed, offset); "
                                   "injected masks are a Python-path test hook")
        self.P, self.M, self.group = P, M, group
        self.b, self.s, self.h = b, s, P.qkv_w.shape[0]
        dev = P.qkv_w.device
        d = LayerDesc()
        d.b, d.s, d.h = b, s, self.h
        d.nh_local, d.hd, d.f_local = M.nh_local, M.hd, P.lin1_w.shape[1]
        d.normalize_before = int(M.normalize_before)
        d.eps, d.p_attn, d.p_hidden1, d.p_hidden2 = M.eps, M.p_attn, M.p_hidden1, M.p_hidden2
        (d.seed_attn, d.offset_attn), (d.seed_h1, d.offset_h1), (d.seed_h2, d.offset_h2) = M.rng_attn, M.rng_h1, M.rng_h2
        for n in _FIELDS:
            setattr(d, n, getattr(P, n).data_ptr())
        self._tp = None
        if group is not None and getattr(group, "nranks", 1) > 1:
            self._tp = group.native_comm(b * s, self.h, dev)  # keeps the ctypes arrays alive
            d.tp = C.pointer(self._tp.comm)
        self.desc = d
        lib = _lib.lib()
        self.saved = torch.empty(lib.fx_decoder_layer_saved_bytes(C.byref(d)), dtype=torch.uint8, device=dev)
        self.workspace = torch.empty(lib.fx_decoder_layer_workspace_bytes(C.byref(d)), dtype=torch.uint8, device=dev)
        if self.saved.numel() == 0:
            check(-1, "fx_decoder_layer_saved_bytes")

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if not x.is_cuda or x.dtype != torch.bfloat16 or not x.is_contiguous() or tuple(x.shape) != (self.b, self.s, self.h):
            raise _lib.FleetXError("NativeLayer.forward: x must be a contiguous CUDA bf16 [b, s, h] tensor")
        out = torch.empty_like(x)
        self._sync_epoch_in()
        check(_lib.lib().fx_decoder_layer_fwd(torch.cuda.current_stream().cuda_stream, C.byref(self.desc), x.data_ptr(),
                                              out.data_ptr(), self.saved.data_ptr(), self.workspace.data_ptr()),
              "fx_decoder_layer_fwd")
        self._sync_epoch_out()
        self._x = x
        return out

    def backward(self, dout: torch.Tensor):
        """-> (dx, {name: grad}) with the gradient dtypes of the parameters."""
        dout = dout.contiguous()
        dx = torch.empty_like(dout)
        grads = {n: torch.empty_like(getattr(self.P, n)) for n in _FIELDS}
        g = LayerGrads()
        for n in _FIELDS:
            setattr(g, n, grads[n].data_ptr())
        self._sync_epoch_in()
        check(_lib.lib().fx_decoder_layer_bwd(torch.cuda.current_stream().cuda_stream, C.byref(self.desc), dout.data_ptr(),
                                              self._x.data_ptr(), self.saved.data_ptr(), dx.data_ptr(), C.byref(g),
                                              self.workspace.data_ptr()), "fx_decoder_layer_bwd")
        self._sync_epoch_out()
        return dx, grads

    # the symmetric group counts its allreduces (the kernels' flag epochs); keep its counter and the C one in step
    def _sync_epoch_in(self):
        if self._tp is not None:
            self._tp.epoch.value = self.group._epoch

    def _sync_epoch_out(self):
        if self._tp is not None:
            self.group._epoch = self._tp.epoch.value
