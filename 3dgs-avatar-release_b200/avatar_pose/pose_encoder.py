"""Hierarchical pose encoder in one launch: host side of `pose_encoder_forward` (include/gsplat_b200.h).  Mirrors
HierarchicalPoseEncoder (models/network_utils.py:119-178; built from cfg.pose_encoder at models/deformer/non_rigid.py:206
and called at :236 as `pose_encoder(rots, Jtrs)`): `forward(rots [1,24,9], Jtrs [1,24,3]) -> [1,144]`, the condition the
non-rigid MLP folds into its first bias.  Parameters are stored packed ([24,19,19] ...) so that a forward is ONE kernel;
`load_reference_state_dict` maps the reference's names (layer_0.*, layers.{j}.0.*, layers.{j}.2.*).  The backward is ONE
kernel too (`pose_encoder_backward`: the reference's autograd runs several hundred launches for it).  No CPU fallback."""
import ctypes

import torch
import torch.nn as nn

from diff_gaussian_rasterization import _C as _gs

vp = ctypes.c_void_p
KTREE_PARENTS = [-1, 0, 0, 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 9, 9, 12, 13, 14, 16, 17, 18, 19, 20, 21]


class PoseEncoderArgs(ctypes.Structure):
    _fields_ = [("num_joints", ctypes.c_int32), ("dim_per_joint", ctypes.c_int32), ("rel_joints", ctypes.c_int32),
                ("pad_", ctypes.c_int32), ("parents", ctypes.c_int32 * 24), ("rots", vp), ("Jtrs", vp), ("W0", vp), ("b0", vp),
                ("Wa", vp), ("ba", vp), ("Wb", vp), ("bb", vp), ("out", vp)]


class PoseEncoderBackwardArgs(ctypes.Structure):
    _fields_ = [("fwd", PoseEncoderArgs), ("dL_dout", vp), ("dL_drots", vp), ("dL_dJtrs", vp), ("dL_dW0", vp), ("dL_db0", vp),
                ("dL_dWa", vp), ("dL_dba", vp), ("dL_dWb", vp), ("dL_dbb", vp)]


_configured = False


def _lib():
    global _configured
    L = _gs.lib()
    if not _configured:
        L.pose_encoder_forward.restype = ctypes.c_int
        L.pose_encoder_forward.argtypes = [ctypes.POINTER(PoseEncoderArgs), vp]
        L.pose_encoder_backward.restype = ctypes.c_int
        L.pose_encoder_backward.argtypes = [ctypes.POINTER(PoseEncoderBackwardArgs), vp]
        _configured = True
    return L


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

    def _fill(self, a, rots, Jtrs, keep):
        c = lambda t: t.detach().contiguous().float()
        rots, Jtrs = c(rots.reshape(24, 9)), c(Jtrs.reshape(24, 3))
        a.num_joints, a.dim_per_joint, a.rel_joints = 24, 6, int(self.rel_joints)
        for j, p in enumerate(KTREE_PARENTS):
            a.parents[j] = p
        params = [c(t) for t in self.params()]
        keep += [rots, Jtrs] + params
        p = _gs._ptr
        a.rots, a.Jtrs = p(rots, "rots"), p(Jtrs, "Jtrs")
        a.W0, a.b0, a.Wa, a.ba, a.Wb, a.bb = (p(t, "param") for t in params)

    @torch.no_grad()
    def encode(self, rots, Jtrs):
        if not rots.is_cuda:
            raise _gs.GsError("HierarchicalPoseEncoderFused needs CUDA tensors (there is no CPU path)")
        L = _lib()
        dev = rots.device
        a, keep = PoseEncoderArgs(), []
        with torch.cuda.device(dev):
            self._fill(a, rots, Jtrs, keep)
            out = torch.empty(1, 144, dtype=torch.float32, device=dev)
            a.out = out.data_ptr()
            _gs._check(L.pose_encoder_forward(ctypes.byref(a), torch.cuda.current_stream(dev).cuda_stream), "pose_encoder_forward")
        return out

    @torch.no_grad()
    def encode_backward(self, rots, Jtrs, g):
        """Hand-written backward (pose_encoder_backward, one launch): -> (drots [24,9], dJtrs [24,3] or None with rel_joints (the
        reference cuts that gradient, network_utils.py:155-159), [dW0, db0, dWa, dba, dWb, dbb])."""
        L = _lib()
        dev = rots.device
        b, keep = PoseEncoderBackwardArgs(), []
        with torch.cuda.device(dev):
            self._fill(b.fwd, rots, Jtrs, keep)
            g = g.detach().contiguous().float().reshape(-1)
            f32 = dict(dtype=torch.float32, device=dev)
            drots = torch.empty(24, 9, **f32)
            dJ = None if self.rel_joints else torch.empty(24, 3, **f32)
            dP = [torch.empty_like(t, dtype=torch.float32) for t in self.params()]
            b.dL_dout, b.dL_drots, b.dL_dJtrs = g.data_ptr(), drots.data_ptr(), (dJ.data_ptr() if dJ is not None else None)
            b.dL_dW0, b.dL_db0, b.dL_dWa, b.dL_dba, b.dL_dWb, b.dL_dbb = (t.data_ptr() for t in dP)
            _gs._check(L.pose_encoder_backward(ctypes.byref(b), torch.cuda.current_stream(dev).cuda_stream), "pose_encoder_backward")
        return drots, dJ, dP

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
        rots, Jtrs = ctx.saved_tensors[:2]
        drots, dJ, dP = ctx.mod.encode_backward(rots, Jtrs, g)
        grads = [drots.view(rots.shape), None if dJ is None else dJ.view(Jtrs.shape)] + dP
        return (None, *[gr if n else None for gr, n in zip(grads, ctx.needs_input_grad[1:])])
