"""Skinning field (forward): per-Gaussian LBS weights from the skinning MLP on the tensor cores -- host side of
`skinning_forward` (include/gsplat_b200.h).  Mirrors the weight part of SkinningField (models/deformer/rigid.py:131-236,
registry key `skinning_field`, non-distilled path): `lbs_network` = VanillaCondMLP(3 -> 128 x4 -> 25), hierarchical
softmax.  The result feeds `avatar_pose.fused_pose(..., weights, ...)`.  `query_weights` is the tcgen05 forward;
`query_weights_autograd` adds the hand-written backward `skinning_backward` (softmax adjoint + tcgen05 dX / dW GEMMs over the
activations the training forward kept): no torch ops, no cuBLAS.  No CPU fallback."""
import ctypes

import torch
import torch.nn as nn

from diff_gaussian_rasterization import _C as _gs

vp = ctypes.c_void_p


class SkinArgs(ctypes.Structure):
    _fields_ = [("P", ctypes.c_int32), ("hidden", ctypes.c_int32), ("n_out", ctypes.c_int32), ("pad_", ctypes.c_int32),
                ("aabb_min", vp), ("aabb_max", vp), ("xyz", vp), ("W", vp * 5), ("b", vp * 5), ("weights", vp), ("logits", vp)]


class SkinBackwardArgs(ctypes.Structure):
    _fields_ = [("fwd", SkinArgs), ("dL_dweights", vp), ("dL_dxyz", vp), ("dL_dW", vp * 5), ("dL_db", vp * 5)]


_configured = False


def _lib():
    global _configured
    L = _gs.lib()
    if not _configured:
        L.skinning_workspace_bytes.restype = ctypes.c_size_t
        L.skinning_workspace_bytes.argtypes = [ctypes.POINTER(SkinArgs)]
        L.skinning_forward.restype = ctypes.c_int
        L.skinning_forward.argtypes = [ctypes.POINTER(SkinArgs), vp, ctypes.c_size_t, vp]
        L.skinning_saved_bytes.restype = ctypes.c_size_t
        L.skinning_saved_bytes.argtypes = [ctypes.c_int32]
        L.skinning_forward_train.restype = ctypes.c_int
        L.skinning_forward_train.argtypes = [ctypes.POINTER(SkinArgs), vp, ctypes.c_size_t, vp, ctypes.c_size_t, vp]
        L.skinning_backward.restype = ctypes.c_int
        L.skinning_backward.argtypes = [ctypes.POINTER(SkinBackwardArgs), vp, ctypes.c_size_t, vp, ctypes.c_size_t, vp]
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

    def _fill(self, a, xyz, keep):
        c = lambda t: t.detach().contiguous().float()
        p = _gs._ptr
        xyz = c(xyz)
        a.P, a.hidden, a.n_out = xyz.shape[0], self.hidden, self.d_out
        for l in range(5):
            lin = getattr(self, f"lin{l}")
            wt, bt = c(lin.weight), c(lin.bias)
            keep += [wt, bt]
            a.W[l], a.b[l] = p(wt, "W"), p(bt, "b")
        keep.append(xyz)
        a.aabb_min, a.aabb_max, a.xyz = p(self.aabb_min, "aabb_min"), p(self.aabb_max, "aabb_max"), p(xyz, "xyz")

    def _workspace(self, L, a, dev):
        need = L.skinning_workspace_bytes(ctypes.byref(a))
        if self._ws is None or self._ws.numel() < need or self._ws.device != dev:
            self._ws = torch.empty(need, dtype=torch.uint8, device=dev)
        return self._ws

    @torch.no_grad()
    def query_weights(self, xyz, return_logits=False, _saved=None):
        """xyz [P,3] canonical means -> skinning weights [P,24] (rows sum to 1)."""
        if not xyz.is_cuda:
            raise _gs.GsError("SkinningFieldFused needs CUDA tensors (there is no CPU path)")
        L = _lib()
        dev, P = xyz.device, xyz.shape[0]
        a, keep = SkinArgs(), []
        with torch.cuda.device(dev):
            self._fill(a, xyz, keep)
            weights = torch.empty(P, 24, dtype=torch.float32, device=dev)
            logits = torch.empty(P, self.d_out, dtype=torch.float32, device=dev) if return_logits else None
            a.weights = weights.data_ptr()
            a.logits = logits.data_ptr() if logits is not None else None
            ws = self._workspace(L, a, dev)
            st = torch.cuda.current_stream(dev).cuda_stream
            if P and _saved is None:
                _gs._check(L.skinning_forward(ctypes.byref(a), ws.data_ptr(), ws.numel(), st), "skinning_forward")
            elif P:
                _gs._check(L.skinning_forward_train(ctypes.byref(a), ws.data_ptr(), ws.numel(), _saved.data_ptr(), _saved.numel(), st),
                           "skinning_forward_train")
        return (weights, logits) if return_logits else weights

    @torch.no_grad()
    def query_weights_backward(self, xyz, logits, saved, g_w):
        """Hand-written backward (skinning_backward): -> (dxyz, [dW0, db0, ..., dW4, db4])."""
        L = _lib()
        dev, P = xyz.device, xyz.shape[0]
        b, keep = SkinBackwardArgs(), []
        with torch.cuda.device(dev):
            self._fill(b.fwd, xyz, keep)
            logits, g_w = logits.contiguous(), g_w.detach().contiguous().float()
            b.fwd.logits = logits.data_ptr()
            b.dL_dweights = g_w.data_ptr()
            dxyz = torch.empty(P, 3, dtype=torch.float32, device=dev)
            b.dL_dxyz = dxyz.data_ptr()
            grads = []
            for l in range(5):
                lin = getattr(self, f"lin{l}")
                gw_, gb_ = torch.empty_like(lin.weight, dtype=torch.float32), torch.empty_like(lin.bias, dtype=torch.float32)
                b.dL_dW[l], b.dL_db[l] = gw_.data_ptr(), gb_.data_ptr()
                grads += [gw_, gb_]
            ws = self._workspace(L, b.fwd, dev)
            _gs._check(L.skinning_backward(ctypes.byref(b), ws.data_ptr(), ws.numel(), saved.data_ptr(), saved.numel(),
                                           torch.cuda.current_stream(dev).cuda_stream), "skinning_backward")
        return dxyz, grads

    def query_weights_autograd(self, xyz):
        """Differentiable `query_weights`: gradients to xyz and to the five linear layers (rigid.py:205-212)."""
        params = [t for l in range(5) for t in (getattr(self, f"lin{l}").weight, getattr(self, f"lin{l}").bias)]
        return _SkinFn.apply(self, xyz, *params)


class _SkinFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, mod, xyz, *params):
        ctx.mod = mod
        saved = torch.empty(max(_lib().skinning_saved_bytes(xyz.shape[0]), 16), dtype=torch.uint8, device=xyz.device)
        w, logits = mod.query_weights(xyz, return_logits=True, _saved=saved)
        ctx.save_for_backward(xyz, logits, saved)
        return w

    @staticmethod
    def backward(ctx, g_w):
        xyz, logits, saved = ctx.saved_tensors
        dxyz, grads = ctx.mod.query_weights_backward(xyz, logits, saved, g_w)
        need = ctx.needs_input_grad  # (mod, xyz, W0, b0, ..., W4, b4)
        return (None, *[g if n else None for g, n in zip([dxyz] + grads, need[1:])])
