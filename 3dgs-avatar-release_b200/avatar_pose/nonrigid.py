"""Non-rigid deformation (forward) of the canonical Gaussians on the tensor cores: host side of `nonrigid_forward`
(include/gsplat_b200.h).  Mirrors HashGridwithMLP (models/deformer/non_rigid.py:203-300, registry key `hashgrid`):
multiresolution hash grid (16 levels x 2 features) + VanillaCondMLP (32+C -> 128 -> 128 -> 128 -> 10+feature_dim,
LeakyReLU) + delta application; `iteration < delay` is the identity exactly as in the reference.  The forward is the
tcgen05 kernel (`deform`); `deform_autograd` wraps it in an autograd.Function whose backward is the hand-written
`nonrigid_backward` (tcgen05 dX / dW GEMMs over the activations the training forward kept, hash-grid scatter with
red.global.add) -- no torch ops, no cuBLAS.  There is no CPU fallback."""
import ctypes
import math

import numpy as np
import torch
import torch.nn as nn

from diff_gaussian_rasterization import _C as _gs

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


class NonRigidBackwardArgs(ctypes.Structure):
    _fields_ = [("fwd", NonRigidArgs), ("dL_dxyz_out", vp), ("dL_dscaling_out", vp), ("dL_drotation_out", vp), ("dL_ddeltas", vp),
                ("dL_dxyz", vp), ("dL_dscaling", vp), ("dL_drotation", vp), ("dL_dcond", vp), ("dL_dtable", vp),
                ("dL_dW0", vp), ("dL_db0", vp), ("dL_dW1", vp), ("dL_db1", vp), ("dL_dW2", vp), ("dL_db2", vp),
                ("dL_dW3", vp), ("dL_db3", vp)]


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
        L.nonrigid_saved_bytes.restype = ctypes.c_size_t
        L.nonrigid_saved_bytes.argtypes = [ctypes.c_int32]
        L.nonrigid_forward_train.restype = ctypes.c_int
        L.nonrigid_forward_train.argtypes = [ctypes.POINTER(NonRigidArgs), vp, ctypes.c_size_t, vp, ctypes.c_size_t, vp]
        L.nonrigid_backward.restype = ctypes.c_int
        L.nonrigid_backward.argtypes = [ctypes.POINTER(NonRigidBackwardArgs), vp, ctypes.c_size_t, vp, ctypes.c_size_t, vp]
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

    def _fill(self, a, xyz, scaling, rotation, cond, keep):
        """Fills the input part of NonRigidArgs; `keep` collects the tensors whose storage the pointers refer to."""
        c = lambda t: t.detach().contiguous().float()
        P = xyz.shape[0]
        a.P, a.n_levels, a.n_feat, a.hidden, a.n_cond, a.n_out = P, self.n_levels, 2, self.hidden, self.n_cond, self.n_out
        off, res, scale = self.levels
        for l in range(self.n_levels):
            a.level_offset[l], a.level_res[l], a.level_scale[l] = int(off[l]), int(res[l]), float(scale[l])
        a.level_offset[self.n_levels] = int(off[self.n_levels])
        p = _gs._ptr
        xyz, scaling, rotation, cond, table = c(xyz), c(scaling), c(rotation), c(cond).reshape(-1), c(self.table)
        ws_t = [c(l.weight) for l in (self.lin0, self.lin1, self.lin2, self.lin3)]
        bs_t = [c(l.bias) for l in (self.lin0, self.lin1, self.lin2, self.lin3)]
        keep += [xyz, scaling, rotation, cond, table] + ws_t + bs_t
        a.aabb_min, a.aabb_max = p(self.aabb_min, "aabb_min"), p(self.aabb_max, "aabb_max")
        a.xyz, a.scaling, a.rotation = p(xyz, "xyz"), p(scaling, "scaling"), p(rotation, "rotation")
        a.table, a.cond = p(table, "table"), p(cond, "cond")
        a.W0, a.W1, a.W2, a.W3 = (p(t, "W") for t in ws_t)
        a.b0, a.b1, a.b2, a.b3 = (p(t, "b") for t in bs_t)

    def _workspace(self, L, a, dev):
        need = L.nonrigid_workspace_bytes(ctypes.byref(a))
        if self._ws is None or self._ws.numel() < need or self._ws.device != dev:
            self._ws = torch.empty(need, dtype=torch.uint8, device=dev)
        return self._ws

    @torch.no_grad()
    def deform(self, xyz, scaling, rotation, cond, iteration=None, return_features=False, _saved=None):
        if iteration is not None and iteration < self.delay:  # non_rigid.py:227-232: identity before `delay`
            return xyz, scaling, rotation, None
        if not xyz.is_cuda:
            raise _gs.GsError("HashGridMLPFused needs CUDA tensors (there is no CPU path)")
        L = _lib()
        dev, P = xyz.device, xyz.shape[0]
        a, keep = NonRigidArgs(), []
        with torch.cuda.device(dev):
            self._fill(a, xyz, scaling, rotation, cond, keep)
            f32 = dict(dtype=torch.float32, device=dev)
            out_xyz, out_sc, out_rot = torch.empty(P, 3, **f32), torch.empty(P, 3, **f32), torch.empty(P, 4, **f32)
            deltas = torch.empty(P, self.n_out, **f32)
            feats = torch.empty(P, 2 * self.n_levels, **f32) if return_features else None
            a.xyz_out, a.scaling_out, a.rotation_out = out_xyz.data_ptr(), out_sc.data_ptr(), out_rot.data_ptr()
            a.deltas = deltas.data_ptr()
            a.features = feats.data_ptr() if feats is not None else None
            ws = self._workspace(L, a, dev)
            st = torch.cuda.current_stream(dev).cuda_stream
            if P and _saved is None:
                _gs._check(L.nonrigid_forward(ctypes.byref(a), ws.data_ptr(), ws.numel(), st), "nonrigid_forward")
            elif P:
                _gs._check(L.nonrigid_forward_train(ctypes.byref(a), ws.data_ptr(), ws.numel(), _saved.data_ptr(), _saved.numel(), st),
                           "nonrigid_forward_train")
        if return_features:
            return out_xyz, out_sc, out_rot, deltas, feats
        return out_xyz, out_sc, out_rot, deltas

    def saved_buffer(self, P, dev):
        """Storage for the activations the training forward keeps for `nonrigid_backward` (1.6 KB per Gaussian)."""
        return torch.empty(max(_lib().nonrigid_saved_bytes(P), 16), dtype=torch.uint8, device=dev)

    @torch.no_grad()
    def deform_backward(self, xyz, scaling, rotation, cond, deltas, saved, g_xyz, g_sc, g_rot, g_d):
        """Hand-written backward (nonrigid_backward): -> (dxyz, dscaling, drotation, dcond, dtable, [dW0, db0, ..., dW3, db3])."""
        L = _lib()
        dev, P = xyz.device, xyz.shape[0]
        b, keep = NonRigidBackwardArgs(), []
        c = lambda t: None if t is None else t.detach().contiguous().float()
        with torch.cuda.device(dev):
            self._fill(b.fwd, xyz, scaling, rotation, cond, keep)
            f32 = dict(dtype=torch.float32, device=dev)
            deltas, g_xyz, g_sc, g_rot, g_d = c(deltas), c(g_xyz), c(g_sc), c(g_rot), c(g_d)
            b.fwd.deltas = deltas.data_ptr()
            pn = lambda t: None if t is None else t.data_ptr()
            b.dL_dxyz_out, b.dL_dscaling_out, b.dL_drotation_out, b.dL_ddeltas = pn(g_xyz), pn(g_sc), pn(g_rot), pn(g_d)
            dxyz, dsc, drot = torch.empty(P, 3, **f32), torch.empty(P, 3, **f32), torch.empty(P, 4, **f32)
            dcond = torch.empty(self.n_cond, **f32)
            dtable = torch.empty_like(self.table, dtype=torch.float32)
            lins = (self.lin0, self.lin1, self.lin2, self.lin3)
            dWb = [torch.empty_like(t, dtype=torch.float32) for l in lins for t in (l.weight, l.bias)]
            b.dL_dxyz, b.dL_dscaling, b.dL_drotation = dxyz.data_ptr(), dsc.data_ptr(), drot.data_ptr()
            b.dL_dcond, b.dL_dtable = dcond.data_ptr(), dtable.data_ptr()
            (b.dL_dW0, b.dL_db0, b.dL_dW1, b.dL_db1, b.dL_dW2, b.dL_db2, b.dL_dW3, b.dL_db3) = (t.data_ptr() for t in dWb)
            ws = self._workspace(L, b.fwd, dev)
            _gs._check(L.nonrigid_backward(ctypes.byref(b), ws.data_ptr(), ws.numel(), saved.data_ptr(), saved.numel(),
                                           torch.cuda.current_stream(dev).cuda_stream), "nonrigid_backward")
        return dxyz, dsc, drot, dcond, dtable, dWb

    def deform_autograd(self, xyz, scaling, rotation, cond, iteration=None):
        """Differentiable `deform`: forward on the tcgen05 kernel, gradients to xyz / scaling / rotation / cond, the hash
        table and the four linear layers (non_rigid.py:248-282 under autograd in the reference) from `nonrigid_backward`."""
        if iteration is not None and iteration < self.delay:
            return xyz, scaling, rotation, None
        params = [self.table] + [t for l in (self.lin0, self.lin1, self.lin2, self.lin3) for t in (l.weight, l.bias)]
        return _NonRigidFn.apply(self, xyz, scaling, rotation, cond, *params)


class _NonRigidFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, mod, xyz, scaling, rotation, cond, *params):
        ctx.mod = mod
        saved = mod.saved_buffer(xyz.shape[0], xyz.device)
        out = mod.deform(xyz, scaling, rotation, cond, _saved=saved)
        ctx.save_for_backward(xyz, scaling, rotation, cond, out[3], saved)
        ctx.set_materialize_grads(False)
        return out

    @staticmethod
    def backward(ctx, g_xyz, g_sc, g_rot, g_d):
        mod = ctx.mod
        xyz, scaling, rotation, cond, deltas, saved = ctx.saved_tensors
        dxyz, dsc, drot, dcond, dtable, dWb = mod.deform_backward(xyz, scaling, rotation, cond, deltas, saved, g_xyz, g_sc, g_rot, g_d)
        need = ctx.needs_input_grad  # (mod, xyz, scaling, rotation, cond, table, W0, b0, ..., W3, b3)
        grads = [dxyz, dsc, drot, dcond.view(cond.shape), dtable] + dWb
        return (None, *[g if n else None for g, n in zip(grads, need[1:])])
