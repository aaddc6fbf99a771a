"""Skinning field (forward): per-Gaussian LBS weights from the skinning MLP on the tensor cores -- host side of
`skinning_forward` (include/gsplat_b200.h).  Mirrors the weight part of SkinningField (models/deformer/rigid.py:131-236,
registry key `skinning_field`, non-distilled path): `lbs_network` = VanillaCondMLP(3 -> 128 x4 -> 25), hierarchical
softmax.  The result feeds `avatar_pose.fused_pose(..., weights, ...)`.  `query_weights` is the tcgen05 forward;
`query_weights_autograd` adds a backward that re-evaluates the MLP with torch ops on the GPU (`_torch_ref.py`; a native
backward is the next step).  No CPU fallback."""
import ctypes

import torch
import torch.nn as nn

from diff_gaussian_rasterization import _C as _gs
from . import _torch_ref

vp = ctypes.c_void_p


class SkinArgs(ctypes.Structure):
    _fields_ = [("P", ctypes.c_int32), ("hidden", ctypes.c_int32), ("n_out", ctypes.c_int32), ("pad_", ctypes.c_int32),
                ("aabb_min", vp), ("aabb_max", vp), ("xyz", vp), ("W", vp * 5), ("b", vp * 5), ("weights", vp), ("logits", vp)]


_configured = False


def _lib():
    global _configured
    L = _gs.lib()
    if not _configured:
        L.skinning_workspace_bytes.restype = ctypes.c_size_t
        L.skinning_workspace_bytes.argtypes = [ctypes.POINTER(SkinArgs)]
        L.skinning_forward.restype = ctypes.c_int
        L.skinning_forward.argtypes = [ctypes.POINTER(SkinArgs), vp, ctypes.c_size_t, vp]
        _configured = True
    return L


class SkinningFieldFused(nn.Module):
    def __init__(self, d_out=25, hidden=128, aabb_min=(-1, -1, -1), aabb_max=(1, 1, 1)):
        super().__init__()
        self.d_out, self.hidden = d_out, hidden
        dims = [3, hidden, hidden, hidden, hidden, d_out]  # n_hidden_layers: 4 (configs/rigid/skinning_field.yaml)
        for l in range(5):
            setattr(self, f"lin{l}", nn.Linear(dims[l], dims[l + 1]))
        self.register_buffer("aabb_min", torch.tensor(aabb_min, dtype=torch.float32))
        self.register_buffer("aabb_max", torch.tensor(aabb_max, dtype=torch.float32))
        self._ws = None

    @torch.no_grad()
    def query_weights(self, xyz, return_logits=False):
        """xyz [P,3] canonical means -> skinning weights [P,24] (rows sum to 1)."""
        if not xyz.is_cuda:
            raise _gs.GsError("SkinningFieldFused needs CUDA tensors (there is no CPU path)")
        L = _lib()
        dev, P = xyz.device, xyz.shape[0]
        c = lambda t: t.detach().contiguous().float()
        xyz = c(xyz)
        a = SkinArgs()
        a.P, a.hidden, a.n_out = P, self.hidden, self.d_out
        p = _gs._ptr
        with torch.cuda.device(dev):
            keep = []
            for l in range(5):
                lin = getattr(self, f"lin{l}")
                wt, bt = c(lin.weight), c(lin.bias)
                keep += [wt, bt]
                a.W[l], a.b[l] = p(wt, "W"), p(bt, "b")
            a.aabb_min, a.aabb_max, a.xyz = p(self.aabb_min, "aabb_min"), p(self.aabb_max, "aabb_max"), p(xyz, "xyz")
            weights = torch.empty(P, 24, dtype=torch.float32, device=dev)
            logits = torch.empty(P, self.d_out, dtype=torch.float32, device=dev) if return_logits else None
            a.weights = weights.data_ptr()
            a.logits = logits.data_ptr() if logits is not None else None
            need = L.skinning_workspace_bytes(ctypes.byref(a))
            if self._ws is None or self._ws.numel() < need or self._ws.device != dev:
                self._ws = torch.empty(need, dtype=torch.uint8, device=dev)
            if P:
                _gs._check(L.skinning_forward(ctypes.byref(a), self._ws.data_ptr(), self._ws.numel(),
                                              torch.cuda.current_stream(dev).cuda_stream), "skinning_forward")
        return (weights, logits) if return_logits else weights

    def query_weights_autograd(self, xyz):
        """Differentiable `query_weights`: gradients to xyz and to the five linear layers (rigid.py:205-212)."""
        params = [t for l in range(5) for t in (getattr(self, f"lin{l}").weight, getattr(self, f"lin{l}").bias)]
        return _SkinFn.apply(self, xyz, *params)


class _SkinFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, mod, xyz, *params):
        ctx.mod = mod
        ctx.save_for_backward(xyz, *params)
        return mod.query_weights(xyz)

    @staticmethod
    def backward(ctx, g_w):
        mod = ctx.mod
        leaves = [t.detach().float().requires_grad_(True) for t in ctx.saved_tensors]
        with torch.enable_grad():
            w = _torch_ref.skinning_weights(leaves[0], leaves[1::2], leaves[2::2], mod.aabb_min, mod.aabb_max)
            need = [i for i, n in enumerate(ctx.needs_input_grad[1:]) if n]
            grads = torch.autograd.grad([w], [leaves[i] for i in need], [g_w], allow_unused=True)
        out = [None] * len(leaves)
        for i, g in zip(need, grads):
            out[i] = g
        return (None, *out)
