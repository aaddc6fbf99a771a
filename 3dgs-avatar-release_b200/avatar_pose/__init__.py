"""Fused posing of canonical Gaussians (host side of pose_forward / pose_backward, include/gsplat_b200.h).

`fused_pose(...)` is ONE autograd op for what the reference runs as ~60 small PyTorch kernels per frame:
rigid LBS of means and rotations (models/deformer/rigid.py:49-71, 215-236), covariance from the posed rotation
(scene/gaussian_model.py:28-32,154-157), activations (:132-152) and SH2RGB colour (models/texture/texture.py:21-38).
`SMPLNNFused` / `SH2RGBFused` in .modules mirror the reference's registry entries (`smpl_nn`, `sh2rgb`).
There is no CPU fallback: tensors must live on a CUDA device and the native library must be built.
"""
import ctypes

import torch

from diff_gaussian_rasterization import _C as _gs

vp = ctypes.c_void_p


class PoseArgs(ctypes.Structure):
    _fields_ = [("P", ctypes.c_int32), ("sh_degree", ctypes.c_int32), ("M", ctypes.c_int32),
                ("cano_view_dir", ctypes.c_int32), ("scaling_modifier", ctypes.c_float),
                ("xyz", vp), ("rotation", vp), ("scaling", vp), ("opacity", vp), ("features_dc", vp),
                ("features_rest", vp), ("weights", vp), ("bone_transforms", vp), ("camera_center", vp),
                ("means3D", vp), ("cov3D", vp), ("colors", vp), ("opacity_out", vp), ("T_fwd", vp),
                ("rotation_bar", vp), ("view_noise", vp)]


class PoseBackwardArgs(ctypes.Structure):
    _fields_ = [("fwd", PoseArgs), ("dL_dmeans3D", vp), ("dL_dcov3D", vp), ("dL_dcolors", vp), ("dL_dopacity", vp),
                ("dL_dxyz", vp), ("dL_drotation", vp), ("dL_dscaling", vp), ("dL_dopacity_logit", vp),
                ("dL_dfeatures_dc", vp), ("dL_dfeatures_rest", vp), ("dL_dweights", vp), ("dL_dbone_transforms", vp)]


_configured = False
_last_slab = [None]
_slab_buffer = [None]


def set_grad_slab_buffer(buf):
    """Makes every following pose backward write its flat gradient slab into `buf[:n]` (a fixed FP32 CUDA buffer, e.g. the
    symmetric-memory buffer of avatar_pose.dist.SymmetricGradSlab, so that the all-reduce kernel finds the gradients in place)
    instead of a fresh allocation.  None restores the default."""
    _slab_buffer[0] = buf


def last_grad_slab():
    """The flat FP32 buffer the last pose backward wrote all per-Gaussian parameter gradients into
    (xyz 3 | rotation 4 | scaling 3 | opacity 1 | features_dc 3 | features_rest 3*(M-1), attribute-major).  Data-parallel
    training all-reduces THIS buffer with one NCCL call; the .grad tensors of the parameters are views into it."""
    return _last_slab[0]


def _lib():
    global _configured
    L = _gs.lib()
    if not _configured:
        L.pose_forward.restype = ctypes.c_int
        L.pose_forward.argtypes = [ctypes.POINTER(PoseArgs), vp]
        L.pose_backward.restype = ctypes.c_int
        L.pose_backward.argtypes = [ctypes.POINTER(PoseBackwardArgs), vp]
        _configured = True
    return L


def _p(t, name):
    return _gs._ptr(t, name)


def _fill_fwd(xyz, rotation, scaling, opacity, f_dc, f_rest, weights, bones, cam, sh_degree, mod, cano, view_noise=None):
    P = xyz.shape[0]
    M = 1 + (f_rest.shape[1] if f_rest is not None and f_rest.numel() else 0)
    a = PoseArgs()
    a.P, a.sh_degree, a.M, a.cano_view_dir, a.scaling_modifier = P, int(sh_degree), M, int(bool(cano)), float(mod)
    a.xyz, a.rotation, a.scaling, a.opacity = _p(xyz, "xyz"), _p(rotation, "rotation"), _p(scaling, "scaling"), _p(opacity, "opacity")
    a.features_dc, a.features_rest = _p(f_dc, "features_dc"), _p(f_rest, "features_rest")
    a.weights, a.bone_transforms, a.camera_center = _p(weights, "weights"), _p(bones, "bone_transforms"), _p(cam, "camera_center")
    a.view_noise = _p(view_noise, "view_noise") if view_noise is not None else None
    if weights.shape[-1] != 24 or bones.numel() != 24 * 16:
        raise _gs.GsError("weights must be [P,24] and bone_transforms [24,4,4]")
    return a


class _FusedPose(torch.autograd.Function):
    """One forward kernel + one backward kernel, through the C++ extension (csrc/ext/ext.cpp: pose_forward / pose_backward fill
    PoseArgs / PoseBackwardArgs and call the C ABI; the ctypes route cost ~25 us of Python per call)."""

    @staticmethod
    def forward(ctx, xyz, rotation, scaling, opacity, f_dc, f_rest, weights, bones, cam, sh_degree, mod, cano,
                want_aux, view_noise=None):
        if not xyz.is_cuda:
            raise _gs.GsError("fused_pose needs CUDA tensors (there is no CPU path)")
        c = lambda t: t.detach().contiguous()
        saved = tuple(map(c, (xyz, rotation, scaling, opacity, f_dc, f_rest, weights, bones, cam)))
        P = saved[0].shape[0]
        vn_ = None if view_noise is None else view_noise.detach().contiguous().float().reshape(-1)
        out, T_fwd, Rbar = _gs._call(_gs._ext.pose_forward, *saved, int(sh_degree), float(mod), bool(cano), bool(want_aux), vn_)
        flat = out.view(-1)  # one allocation: means 3 | cov 6 | colour 3 | opacity 1 (SoA slabs)
        means, cov = flat[:3 * P].view(P, 3), flat[3 * P:9 * P].view(P, 6)
        colors, opac = flat[9 * P:12 * P].view(P, 3), flat[12 * P:].view(P, 1)
        ctx.save_for_backward(*saved)
        ctx.cfg = (int(sh_degree), float(mod), bool(cano))
        ctx.view_noise = vn_
        ctx.need_w = weights.requires_grad
        ctx.need_b = bones.requires_grad
        if want_aux:
            ctx.mark_non_differentiable(T_fwd, Rbar)
            return means, cov, colors, opac, T_fwd, Rbar
        return means, cov, colors, opac

    @staticmethod
    def backward(ctx, g_means, g_cov, g_col, g_op, *unused):
        xyz, rot, sc, op, dc, rest, w, b, cam = ctx.saved_tensors
        sh_degree, mod, cano = ctx.cfg
        P = xyz.shape[0]
        nrest = rest.shape[1] if rest.numel() else 0
        # one flat gradient slab (the buffer a data-parallel all-reduce runs over): xyz 3 | rot 4 | scale 3 | opacity 1 | dc 3 |
        # rest 3*nrest, attribute-major; written into the registered slab buffer (set_grad_slab_buffer) when there is one
        flat, gw, gb = _gs._call(_gs._ext.pose_backward, xyz, rot, sc, op, dc, rest, w, b, cam, sh_degree, mod, cano, ctx.view_noise,
                                 g_means, g_cov, g_col, g_op, bool(ctx.need_w), bool(ctx.need_b), _slab_buffer[0])
        _last_slab[0] = flat
        gxyz, grot, gsc = flat[:3 * P].view(P, 3), flat[3 * P:7 * P].view(P, 4), flat[7 * P:10 * P].view(P, 3)
        gop, gdc = flat[10 * P:11 * P].view(op.shape), flat[11 * P:14 * P].view(P, 1, 3)
        grest = flat[14 * P:].view(P, nrest, 3)
        return gxyz, grot, gsc, gop, gdc, grest, gw, (gb.view(b.shape) if gb is not None else None), None, None, None, None, None, None


def fused_pose(xyz, rotation, scaling, opacity, features_dc, features_rest, weights, bone_transforms, camera_center,
               sh_degree=3, scaling_modifier=1.0, cano_view_dir=True, return_aux=False, view_noise=None):
    """-> (means3D [P,3], cov3D [P,6], colors [P,3], opacity [P,1]) (+ T_fwd [P,4,4], rotation_bar [P,3,3]).

    Inputs are the raw GaussianModel parameters (`_xyz`, `_rotation`, `_scaling`, `_opacity`, `_features_dc`,
    `_features_rest`), per-Gaussian skinning weights [P,24], `camera.bone_transforms` [24,4,4] and
    `camera.camera_center` [3]; `view_noise` [3,3]: the training-time view augmentation of SH2RGB (texture.py:28-33: the matrix the
    reference multiplies the canonical view direction with, i.e. augm_rots(...).T), None in eval mode.  Gradients flow to the six parameter tensors, to `weights` and to `bone_transforms` (the route by
    which the image loss reaches a per-frame pose correction, rigid.py:215-236)."""
    return _FusedPose.apply(xyz, rotation, scaling, opacity, features_dc, features_rest, weights, bone_transforms,
                            camera_center, sh_degree, scaling_modifier, cano_view_dir, return_aux, view_noise)
