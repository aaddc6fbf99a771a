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
    @staticmethod
    def forward(ctx, xyz, rotation, scaling, opacity, f_dc, f_rest, weights, bones, cam, sh_degree, mod, cano,
                want_aux, view_noise=None):
        L = _lib()
        dev = xyz.device
        if not xyz.is_cuda:
            raise _gs.GsError("fused_pose needs CUDA tensors (there is no CPU path)")
        c = lambda t: t.detach().contiguous()
        xyz_, rot_, sc_, op_, dc_, rest_, w_, b_, cam_ = map(c, (xyz, rotation, scaling, opacity, f_dc, f_rest, weights, bones, cam))
        P = xyz_.shape[0]
        with torch.cuda.device(dev):
            f32 = dict(dtype=torch.float32, device=dev)
            out = torch.empty(P, 13, **f32)  # one allocation: means 3 | cov 6 | colour 3 | opacity 1 (SoA slabs)
            flat = out.view(-1)
            means, cov = flat[:3 * P].view(P, 3), flat[3 * P:9 * P].view(P, 6)
            colors, opac = flat[9 * P:12 * P].view(P, 3), flat[12 * P:].view(P, 1)
            T_fwd = torch.empty(P, 4, 4, **f32) if want_aux else None
            Rbar = torch.empty(P, 3, 3, **f32) if want_aux else None
            vn_ = None if view_noise is None else view_noise.detach().contiguous().float().reshape(-1)
            a = _fill_fwd(xyz_, rot_, sc_, op_, dc_, rest_, w_, b_, cam_, sh_degree, mod, cano, vn_)
            a.means3D, a.cov3D, a.colors, a.opacity_out = means.data_ptr(), cov.data_ptr(), colors.data_ptr(), opac.data_ptr()
            a.T_fwd = T_fwd.data_ptr() if want_aux else None
            a.rotation_bar = Rbar.data_ptr() if want_aux else None
            if P:
                _gs._check(L.pose_forward(ctypes.byref(a), torch.cuda.current_stream(dev).cuda_stream), "pose_forward")
        ctx.save_for_backward(xyz_, rot_, sc_, op_, dc_, rest_, w_, b_, cam_)
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
        L = _lib()
        xyz, rot, sc, op, dc, rest, w, b, cam = ctx.saved_tensors
        sh_degree, mod, cano = ctx.cfg
        dev = xyz.device
        P = xyz.shape[0]
        with torch.cuda.device(dev):
            f32 = dict(dtype=torch.float32, device=dev)
            z = lambda g, shape: (torch.zeros(shape, **f32) if g is None else g.contiguous())
            g_means, g_cov, g_col, g_op = z(g_means, (P, 3)), z(g_cov, (P, 6)), z(g_col, (P, 3)), z(g_op, (P, 1))
            nrest = rest.shape[1] if rest.numel() else 0
            # one flat gradient slab (the buffer a data-parallel all-reduce runs over): xyz 3 | rot 4 | scale 3 |
            # opacity 1 | dc 3 | rest 3*nrest
            widths = [3, 4, 3, 1, 3, 3 * nrest]
            total = P * sum(widths)
            buf = _slab_buffer[0]
            flat = buf[:total] if (buf is not None and buf.numel() >= total and buf.device == dev) else torch.empty(total, **f32)
            parts, off = [], 0
            for wd in widths:
                parts.append(flat[off:off + P * wd])
                off += P * wd
            gxyz, grot, gsc, gop = parts[0].view(P, 3), parts[1].view(P, 4), parts[2].view(P, 3), parts[3].view(P, 1)
            gdc, grest = parts[4].view(P, 1, 3), parts[5].view(P, nrest, 3)
            gw = torch.empty(P, 24, **f32) if ctx.need_w else None
            ba = PoseBackwardArgs()
            ba.fwd = _fill_fwd(xyz, rot, sc, op, dc, rest, w, b, cam, sh_degree, mod, cano, ctx.view_noise)
            ba.dL_dmeans3D, ba.dL_dcov3D, ba.dL_dcolors, ba.dL_dopacity = (g_means.data_ptr(), g_cov.data_ptr(),
                                                                           g_col.data_ptr(), g_op.data_ptr())
            ba.dL_dxyz, ba.dL_drotation, ba.dL_dscaling, ba.dL_dopacity_logit = (gxyz.data_ptr(), grot.data_ptr(),
                                                                                 gsc.data_ptr(), gop.data_ptr())
            ba.dL_dfeatures_dc = gdc.data_ptr()
            ba.dL_dfeatures_rest = grest.data_ptr() if nrest else None
            ba.dL_dweights = gw.data_ptr() if gw is not None else None
            gb = torch.empty(24, 4, 4, **f32) if ctx.need_b else None
            ba.dL_dbone_transforms = gb.data_ptr() if gb is not None else None
            if P:
                _gs._check(L.pose_backward(ctypes.byref(ba), torch.cuda.current_stream(dev).cuda_stream), "pose_backward")
        _last_slab[0] = flat
        gop = gop.view(op.shape)
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
