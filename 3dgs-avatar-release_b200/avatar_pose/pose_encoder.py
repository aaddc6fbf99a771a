"""Hierarchical pose encoder in one launch: host side of `pose_encoder_forward` (include/gsplat_b200.h).  Mirrors
HierarchicalPoseEncoder (models/network_utils.py:119-178; built from cfg.pose_encoder at models/deformer/non_rigid.py:206
and called at :236 as `pose_encoder(rots, Jtrs)`): `forward(rots [1,24,9], Jtrs [1,24,3]) -> [1,144]`, the condition the
non-rigid MLP folds into its first bias.  Parameters are stored packed ([24,19,19] ...) so that a forward is ONE kernel;
`load_reference_state_dict` maps the reference's names (layer_0.*, layers.{j}.0.*, layers.{j}.2.*).  The backward
re-evaluates the ~11k multiply-adds with torch ops on the GPU.  No CPU fallback."""
import ctypes

import torch
import torch.nn as nn
import torch.nn.functional as F

from diff_gaussian_rasterization import _C as _gs

vp = ctypes.c_void_p
KTREE_PARENTS = [-1, 0, 0, 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 9, 9, 12, 13, 14, 16, 17, 18, 19, 20, 21]


class PoseEncoderArgs(ctypes.Structure):
    _fields_ = [("num_joints", ctypes.c_int32), ("dim_per_joint", ctypes.c_int32), ("rel_joints", ctypes.c_int32),
                ("pad_", ctypes.c_int32), ("parents", ctypes.c_int32 * 24), ("rots", vp), ("Jtrs", vp), ("W0", vp), ("b0", vp),
                ("Wa", vp), ("ba", vp), ("Wb", vp), ("bb", vp), ("out", vp)]


_configured = False


def _lib():
    global _configured
    L = _gs.lib()
    if not _configured:
        L.pose_encoder_forward.restype = ctypes.c_int
        L.pose_encoder_forward.argtypes = [ctypes.POINTER(PoseEncoderArgs), vp]
        _configured = True
    return L


def _torch_encode(rots, Jtrs, W0, b0, Wa, ba, Wb, bb, rel_joints):
    """Differentiable re-evaluation (network_utils.py:152-178), used by the backward only."""
    J = Jtrs
    if rel_joints:
        J = torch.cat([Jtrs[:1], Jtrs[1:] - Jtrs[KTREE_PARENTS[1:]]]).detach()  # the reference computes this under no_grad
    glob = F.linear(torch.cat([rots.reshape(-1), J.reshape(-1)]), W0, b0)
    out = [None] * 24
    for j in range(24):
        p = KTREE_PARENTS[j]
        d = J[j] if (p == -1 or rel_joints) else J[j] - J[p]
        x = torch.cat([rots[j], J[j], d.norm().reshape(1), glob if p == -1 else out[p]])
        out[j] = F.linear(F.relu(F.linear(x, Wa[j], ba[j])), Wb[j], bb[j])
    return torch.cat(out)


class HierarchicalPoseEncoderFused(nn.Module):
    def __init__(self, num_joints=24, rel_joints=False, dim_per_joint=6, out_dim=-1, **kwargs):
        super().__init__()
        if num_joints != 24 or dim_per_joint != 6 or out_dim > 0:
            raise ValueError("HierarchicalPoseEncoderFused covers the configured encoder: 24 joints, 6 per joint, no output layer")
        self.rel_joints, self.n_output_dims = bool(rel_joints), 144
        ref = [nn.Linear(288, 6)] + [nn.Linear(19, 19) for _ in range(24)] + [nn.Linear(19, 6) for _ in range(24)]  # same init
        self.W0, self.b0 = nn.Parameter(ref[0].weight.detach().clone()), nn.Parameter(ref[0].bias.detach().clone())
        self.Wa = nn.Parameter(torch.stack([l.weight.detach() for l in ref[1:25]]))
        self.ba = nn.Parameter(torch.stack([l.bias.detach() for l in ref[1:25]]))
        self.Wb = nn.Parameter(torch.stack([l.weight.detach() for l in ref[25:]]))
        self.bb = nn.Parameter(torch.stack([l.bias.detach() for l in ref[25:]]))

    @torch.no_grad()
    def load_reference_state_dict(self, sd, prefix=""):
        """Copies a state_dict of the reference's HierarchicalPoseEncoder (layer_0.*, layers.{j}.0.*, layers.{j}.2.*)."""
        g = lambda k: torch.as_tensor(sd[prefix + k])
        self.W0.copy_(g("layer_0.weight")); self.b0.copy_(g("layer_0.bias"))
        for j in range(24):
            self.Wa[j].copy_(g(f"layers.{j}.0.weight")); self.ba[j].copy_(g(f"layers.{j}.0.bias"))
            self.Wb[j].copy_(g(f"layers.{j}.2.weight")); self.bb[j].copy_(g(f"layers.{j}.2.bias"))

    def params(self):
        return [self.W0, self.b0, self.Wa, self.ba, self.Wb, self.bb]

    @torch.no_grad()
    def encode(self, rots, Jtrs):
        if not rots.is_cuda:
            raise _gs.GsError("HierarchicalPoseEncoderFused needs CUDA tensors (there is no CPU path)")
        L = _lib()
        dev = rots.device
        c = lambda t: t.detach().contiguous().float()
        rots, Jtrs = c(rots.reshape(24, 9)), c(Jtrs.reshape(24, 3))
        a = PoseEncoderArgs()
        a.num_joints, a.dim_per_joint, a.rel_joints = 24, 6, int(self.rel_joints)
        for j, p in enumerate(KTREE_PARENTS):
            a.parents[j] = p
        keep = [c(t) for t in self.params()]
        p = _gs._ptr
        a.rots, a.Jtrs = p(rots, "rots"), p(Jtrs, "Jtrs")
        a.W0, a.b0, a.Wa, a.ba, a.Wb, a.bb = (p(t, "param") for t in keep)
        with torch.cuda.device(dev):
            out = torch.empty(1, 144, dtype=torch.float32, device=dev)
            a.out = out.data_ptr()
            _gs._check(L.pose_encoder_forward(ctypes.byref(a), torch.cuda.current_stream(dev).cuda_stream), "pose_encoder_forward")
        return out

    def forward(self, rots, Jtrs, skinning_weight=None):
        if torch.is_grad_enabled() and (any(t.requires_grad for t in self.params()) or rots.requires_grad or Jtrs.requires_grad):
            return _EncFn.apply(self, rots, Jtrs, *self.params())
        return self.encode(rots, Jtrs)


class _EncFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, mod, rots, Jtrs, *params):
        ctx.mod = mod
        ctx.save_for_backward(rots, Jtrs, *params)
        return mod.encode(rots, Jtrs)

    @staticmethod
    def backward(ctx, g):
        leaves = [t.detach().float().requires_grad_(True) for t in ctx.saved_tensors]
        with torch.enable_grad():
            y = _torch_encode(leaves[0].reshape(24, 9), leaves[1].reshape(24, 3), *leaves[2:], ctx.mod.rel_joints)
            need = [i for i, n in enumerate(ctx.needs_input_grad[1:]) if n]
            grads = torch.autograd.grad([y], [leaves[i] for i in need], [g.reshape(-1)], allow_unused=True)
        out = [None] * len(leaves)
        for i, gr in zip(need, grads):
            out[i] = gr
        return (None, *out)
