"""Photometric L1 loss of the training step as one fused kernel (forward value + per-pixel gradient): host side of
`l1_loss_forward` / `l1_loss_backward` (include/gsplat_b200.h).  Mirrors utils/loss_utils.py:19-20 (`l1_loss`, called
at train.py:121).  No CPU fallback."""
import ctypes

import torch

from diff_gaussian_rasterization import _C as _gs

vp = ctypes.c_void_p
_configured = False
_ws = {}


def _lib():
    global _configured
    L = _gs.lib()
    if not _configured:
        L.l1_loss_workspace_bytes.restype = ctypes.c_size_t
        L.l1_loss_workspace_bytes.argtypes = []
        L.l1_loss_forward.restype = ctypes.c_int
        L.l1_loss_forward.argtypes = [vp, vp, ctypes.c_int64, vp, vp, vp, ctypes.c_size_t, vp]
        L.l1_loss_backward.restype = ctypes.c_int
        L.l1_loss_backward.argtypes = [vp, vp, ctypes.c_int64, vp, vp]
        _configured = True
    return L


class _L1(torch.autograd.Function):
    """Through the C++ extension (csrc/ext/ext.cpp: l1_forward / l1_backward call l1_loss_forward / l1_loss_backward of the C ABI)."""

    @staticmethod
    def forward(ctx, image, target):
        if not image.is_cuda or not target.is_cuda:
            raise _gs.GsError("l1_loss needs CUDA tensors (there is no CPU path)")
        if image.shape != target.shape:
            raise _gs.GsError("l1_loss: image and target shapes differ")
        loss, unit = _gs._call(_gs._ext.l1_forward, image.detach().float(), target.detach().float())
        ctx.save_for_backward(unit)
        return loss

    @staticmethod
    def backward(ctx, g):
        (unit,) = ctx.saved_tensors
        return _gs._call(_gs._ext.l1_backward, unit, g.detach().float()), None


def l1_loss(network_output, gt):
    """mean |network_output - gt| (utils/loss_utils.py:19-20); gradient to `network_output` only."""
    return _L1.apply(network_output, gt)
