"""Non-rigid deformation (forward) of the canonical Gaussians on the tensor cores: host side of `nonrigid_forward`
(include/gsplat_b200.h).  Mirrors HashGridwithMLP (models/deformer/non_rigid.py:203-300, registry key `hashgrid`):
multiresolution hash grid (16 levels x 2 features) + VanillaCondMLP (32+C -> 128 -> 128 -> 128 -> 10+feature_dim,
LeakyReLU) + delta application; `iteration < delay` is the identity exactly as in the reference.  The forward is the
tcgen05 kernel (`deform`); `deform_autograd` wraps it in an autograd.Function whose backward re-evaluates the stage with
torch ops on the GPU (`_torch_ref.py`) to obtain parameter / input gradients -- a native backward is the next step.
There is no CPU fallback."""
import ctypes
import math

import numpy as np
import torch
import torch.nn as nn

from diff_gaussian_rasterization import _C as _gs
from . import _torch_ref

vp = ctypes.c_void_p
NR_MAX_LEVELS = 32


class NonRigidArgs(ctypes.Structure):
    _fields_ = [("P", ctypes.c_int32), ("n_levels", ctypes.c_int32), ("n_feat", ctypes.c_int32), ("hidden", ctypes.c_int32),
                ("n_cond", ctypes.c_int32), ("n_out", ctypes.c_int32),
                ("level_offset", ctypes.c_uint32 * (NR_MAX_LEVELS + 1)), ("level_res", ctypes.c_uint32 * NR_MAX_LEVELS),
                ("level_scale", ctypes.c_float * NR_MAX_LEVELS),
                ("aabb_min", vp), ("aabb_max", vp), ("xyz", vp), ("scaling", vp), ("rotation", vp), ("table", vp), ("cond", vp),
                ("W0", vp), ("b0", vp), ("W1", vp), ("b1", vp), ("W2", vp), ("b2", vp), ("W3", vp), ("b3", vp),
                ("xyz_out", vp), ("scaling_out", vp), ("rotation_out", vp), ("deltas", vp), ("features", vp)]


def grid_levels(n_levels=16, base_resolution=16, per_level_scale=None, max_resolution=2048, log2_hashmap_size=16):
    """Level sizes as tiny-cuda-nn computes them (see oracle/nonrigid_oracle.py for the derivation)."""
    if per_level_scale is None:  # models/network_utils.py:332-336
        per_level_scale = float(np.exp(np.log(max_resolution / base_resolution) / (n_levels - 1)))
    off, res_l, scale_l = [0], [], []
    for l in range(n_levels):
        scale = np.float32(np.exp2(np.float32(l) * np.log2(np.float32(per_level_scale))) * np.float32(base_resolution) - np.float32(1.0))
        res = int(math.ceil(float(scale))) + 1
        n = min((min(res ** 3, 1 << 32) + 7) // 8 * 8, 1 << log2_hashmap_size)
        off.append(off[-1] + n)
        res_l.append(res)
        scale_l.append(float(scale))
    return np.asarray(off, np.uint32), np.asarray(res_l, np.uint32), np.asarray(scale_l, np.float32)


_configured = False


def _lib():
    global _configured
    L = _gs.lib()
    if not _configured:
        L.nonrigid_workspace_bytes.restype = ctypes.c_size_t
        L.nonrigid_workspace_bytes.argtypes = [ctypes.POINTER(NonRigidArgs)]
        L.nonrigid_forward.restype = ctypes.c_int
        L.nonrigid_forward.argtypes = [ctypes.POINTER(NonRigidArgs), vp, ctypes.c_size_t, vp]
        _configured = True
    return L


class HashGridMLPFused(nn.Module):
    """Drop-in for the compute of HashGridwithMLP: `deform(xyz, scaling, rotation, cond)` -> (xyz', scaling', rotation',
    deltas).  Parameters: `table` [n_entries, 2] (FP32; tiny-cuda-nn keeps FP16 parameters by default), `lin0..lin3`."""

    def __init__(self, n_cond=144, feature_dim=0, aabb_min=(-1, -1, -1), aabb_max=(1, 1, 1), n_levels=16, base_resolution=16,
                 max_resolution=2048, log2_hashmap_size=16, hidden=128, delay=0):
        super().__init__()
        self.levels = grid_levels(n_levels, base_resolution, None, max_resolution, log2_hashmap_size)
        self.n_levels, self.n_cond, self.n_out, self.hidden, self.delay = n_levels, n_cond, 10 + feature_dim, hidden, delay
        n_entries = int(self.levels[0][-1])
        self.table = nn.Parameter((torch.rand(n_entries, 2) * 2 - 1) * 1e-4)  # tcnn initialises U(-1e-4, 1e-4)
        self.lin0 = nn.Linear(2 * n_levels + n_cond, hidden)
        self.lin1 = nn.Linear(hidden, hidden)
        self.lin2 = nn.Linear(hidden, hidden)
        self.lin3 = nn.Linear(hidden, self.n_out)
        self.register_buffer("aabb_min", torch.tensor(aabb_min, dtype=torch.float32))
        self.register_buffer("aabb_max", torch.tensor(aabb_max, dtype=torch.float32))
        self._ws = None

    @torch.no_grad()
    def deform(self, xyz, scaling, rotation, cond, iteration=None, return_features=False):
        if iteration is not None and iteration < self.delay:  # non_rigid.py:227-232: identity before `delay`
            return xyz, scaling, rotation, None
        if not xyz.is_cuda:
            raise _gs.GsError("HashGridMLPFused needs CUDA tensors (there is no CPU path)")
        L = _lib()
        dev, P = xyz.device, xyz.shape[0]
        c = lambda t: t.detach().contiguous().float()
        xyz, scaling, rotation, cond = c(xyz), c(scaling), c(rotation), c(cond).reshape(-1)
        a = NonRigidArgs()
        a.P, a.n_levels, a.n_feat, a.hidden, a.n_cond, a.n_out = P, self.n_levels, 2, self.hidden, self.n_cond, self.n_out
        off, res, scale = self.levels
        for l in range(self.n_levels):
            a.level_offset[l], a.level_res[l], a.level_scale[l] = int(off[l]), int(res[l]), float(scale[l])
        a.level_offset[self.n_levels] = int(off[self.n_levels])
        p = _gs._ptr
        with torch.cuda.device(dev):
            f32 = dict(dtype=torch.float32, device=dev)
            out_xyz, out_sc, out_rot = torch.empty(P, 3, **f32), torch.empty(P, 3, **f32), torch.empty(P, 4, **f32)
            deltas = torch.empty(P, self.n_out, **f32)
            feats = torch.empty(P, 2 * self.n_levels, **f32) if return_features else None
            a.aabb_min, a.aabb_max = p(self.aabb_min, "aabb_min"), p(self.aabb_max, "aabb_max")
            a.xyz, a.scaling, a.rotation = p(xyz, "xyz"), p(scaling, "scaling"), p(rotation, "rotation")
            a.table, a.cond = p(c(self.table), "table"), p(cond, "cond")
            ws_t = [c(l.weight) for l in (self.lin0, self.lin1, self.lin2, self.lin3)]
            bs_t = [c(l.bias) for l in (self.lin0, self.lin1, self.lin2, self.lin3)]
            a.W0, a.W1, a.W2, a.W3 = (p(t, "W") for t in ws_t)
            a.b0, a.b1, a.b2, a.b3 = (p(t, "b") for t in bs_t)
            a.xyz_out, a.scaling_out, a.rotation_out = out_xyz.data_ptr(), out_sc.data_ptr(), out_rot.data_ptr()
            a.deltas = deltas.data_ptr()
            a.features = feats.data_ptr() if feats is not None else None
            need = L.nonrigid_workspace_bytes(ctypes.byref(a))
            if self._ws is None or self._ws.numel() < need or self._ws.device != dev:
                self._ws = torch.empty(need, dtype=torch.uint8, device=dev)
            if P:
                _gs._check(L.nonrigid_forward(ctypes.byref(a), self._ws.data_ptr(), self._ws.numel(),
                                              torch.cuda.current_stream(dev).cuda_stream), "nonrigid_forward")
        if return_features:
            return out_xyz, out_sc, out_rot, deltas, feats
        return out_xyz, out_sc, out_rot, deltas

    def deform_autograd(self, xyz, scaling, rotation, cond, iteration=None):
        """Differentiable `deform`: forward on the tcgen05 kernel, gradients to xyz / scaling / rotation / cond, the hash
        table and the four linear layers (non_rigid.py:248-282 under autograd in the reference)."""
        if iteration is not None and iteration < self.delay:
            return xyz, scaling, rotation, None
        params = [self.table] + [t for l in (self.lin0, self.lin1, self.lin2, self.lin3) for t in (l.weight, l.bias)]
        return _NonRigidFn.apply(self, xyz, scaling, rotation, cond, *params)


class _NonRigidFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, mod, xyz, scaling, rotation, cond, *params):
        ctx.mod = mod
        ctx.save_for_backward(xyz, scaling, rotation, cond, *params)
        return mod.deform(xyz, scaling, rotation, cond)

    @staticmethod
    def backward(ctx, g_xyz, g_sc, g_rot, g_d):
        mod = ctx.mod
        leaves = [t.detach().float().requires_grad_(True) for t in ctx.saved_tensors]
        xyz, scaling, rotation, cond, table = leaves[:5]
        Ws, bs = leaves[5::2], leaves[6::2]
        with torch.enable_grad():
            outs = _torch_ref.nonrigid(xyz, scaling, rotation, cond, table, Ws, bs, mod.aabb_min, mod.aabb_max, mod.levels)
            pairs = [(o, g) for o, g in zip(outs, (g_xyz, g_sc, g_rot, g_d)) if g is not None]
            need = [i for i, n in enumerate(ctx.needs_input_grad[1:]) if n]
            grads = torch.autograd.grad([o for o, _ in pairs], [leaves[i] for i in need], [g for _, g in pairs],
                                        allow_unused=True)
        out = [None] * len(leaves)
        for i, g in zip(need, grads):
            out[i] = g
        return (None, *out)
