"""nn.Module drop-ins for the posing side of 3DGS-Avatar, mirroring the reference's module boundary
(models/gaussian_converter.py:41-59, models/deformer/deformer.py:13-19, registry keys of rigid.py:244-251,
non_rigid.py:302-310, texture.py:128-134):

    converter = FusedGaussianConverter(rigid=SMPLNNFused(smpl_verts, skinning_weights), non_rigid=None)
    deformed, loss_reg, colors_precomp = converter(gaussians, camera, iteration, compute_loss)

`gaussians` is the reference's GaussianModel (or anything exposing `_xyz, _rotation, _scaling, _opacity, _features_dc,
_features_rest, active_sh_degree, max_sh_degree`); `camera` exposes `bone_transforms [24,4,4]`, `camera_center [3]` and,
for the non-rigid stage, `rots`, `Jtrs`.  The returned object duck-types what gaussian_renderer/__init__.py:76-114 and
utils/loss_utils.py:69-74 read from a deformed GaussianModel.  One fused kernel replaces non-rigid clone + rigid clone +
LBS + build_rotation + covariance + activations + SH2RGB."""
import torch
import torch.nn as nn

from . import fused_pose
from .nonrigid import HashGridMLPFused
from .skinning import SkinningFieldFused


class PosedGaussians:
    """Lazy views over the fused outputs (what render() reads: gaussian_renderer/__init__.py:76-114)."""

    def __init__(self, means, cov, opacity, scaling_raw, rotation_raw, features, T_fwd, rotation_bar, active_sh_degree, max_sh_degree):
        self._xyz, self._cov, self._opacity_act = means, cov, opacity
        self._scaling, self._rotation, self._features = scaling_raw, rotation_raw, features
        self.fwd_transform, self.rotation_precomp = T_fwd, rotation_bar
        self.active_sh_degree, self.max_sh_degree = active_sh_degree, max_sh_degree

    get_xyz = property(lambda s: s._xyz)
    get_opacity = property(lambda s: s._opacity_act)
    get_scaling = property(lambda s: torch.exp(s._scaling))                           # scene/gaussian_model.py:132-134
    get_rotation = property(lambda s: torch.nn.functional.normalize(s._rotation))     # :136-138
    get_features = property(lambda s: s._features)

    def get_covariance(self, scaling_modifier=1):
        """[P,6]; R_bar (mod*S)(mod*S)^T R_bar^T = mod^2 * the covariance the fused kernel wrote with mod = 1."""
        return self._cov if scaling_modifier == 1 else self._cov * (scaling_modifier * scaling_modifier)


class SMPLNNFused(nn.Module):
    """`smpl_nn` (rigid.py:35-71): skinning weights of the nearest SMPL vertex; cached while `_xyz` is unchanged."""

    def __init__(self, smpl_verts, skinning_weights):
        super().__init__()
        self.register_buffer("smpl_verts", torch.as_tensor(smpl_verts, dtype=torch.float32))
        self.register_buffer("skinning_weights", torch.as_tensor(skinning_weights, dtype=torch.float32))
        self._key, self._cached = None, None

    @torch.no_grad()
    def query_weights(self, xyz, return_index=False):
        """[P,24] weights of the nearest SMPL vertex: one launch of smpl_nn_kernel (exact brute-force 1-NN + gather)."""
        import ctypes
        from diff_gaussian_rasterization import _C as _gs
        key = (xyz.data_ptr(), xyz._version, xyz.shape[0])
        if key != self._key or return_index:
            if not xyz.is_cuda:
                raise _gs.GsError("SMPLNNFused needs CUDA tensors (there is no CPU path)")
            L = _gs.lib()
            L.smpl_nn_forward.restype = ctypes.c_int
            L.smpl_nn_forward.argtypes = [ctypes.c_int32, ctypes.c_void_p, ctypes.c_int32] + [ctypes.c_void_p] * 5
            x = xyz.detach().contiguous().float()
            P, dev = x.shape[0], x.device
            verts, vw = self.smpl_verts.contiguous(), self.skinning_weights.contiguous()
            with torch.cuda.device(dev):
                out = torch.empty(P, 24, dtype=torch.float32, device=dev)
                idx = torch.empty(P, dtype=torch.int32, device=dev)
                if P:
                    _gs._check(L.smpl_nn_forward(P, x.data_ptr(), verts.shape[0], verts.data_ptr(), vw.data_ptr(), idx.data_ptr(),
                                                 out.data_ptr(), torch.cuda.current_stream(dev).cuda_stream), "smpl_nn_forward")
            self._cached, self._key, self._idx = out, key, idx
        return (self._cached, self._idx) if return_index else self._cached


def sample_view_noise(scale, rng=None):
    """The matrix the reference multiplies the canonical view direction with in training mode (texture.py:28-33, :94-99):
    augm_rots(scale, scale, scale).T (utils/sh_utils.py:169-214) -- roll and yaw ~ clip(N(0,1) * scale, +-2 scale) degrees, pitch
    ~ U(0,1) * scale (the reference draws the pitch with rand, not randn), R = Rx . Ry . Rz.  [3,3] float32 CPU tensor."""
    import numpy as np
    rng = np.random if rng is None else rng
    clip = lambda v: min(2 * scale, max(-2 * scale, v))
    ax, ay, az = clip(rng.randn() * scale), clip(rng.rand() * scale), clip(rng.randn() * scale)
    sx, cx = np.sin(np.pi / 180 * ax), np.cos(np.pi / 180 * ax)
    sy, cy = np.sin(np.pi / 180 * ay), np.cos(np.pi / 180 * ay)
    sz, cz = np.sin(np.pi / 180 * az), np.cos(np.pi / 180 * az)
    rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return torch.tensor(rx @ (ry @ rz), dtype=torch.float32).transpose(0, 1).contiguous()


class FusedGaussianConverter(nn.Module):
    def __init__(self, rigid, non_rigid=None, pose_encoder=None, cano_view_dir=True, texture=None, view_noise=0.0, pose_noise=0.0):
        """rigid: SMPLNNFused | SkinningFieldFused (anything with query_weights(xyz) -> [P,24]);
        non_rigid: HashGridMLPFused or None; pose_encoder(rots, Jtrs) -> [1,C] (HierarchicalPoseEncoder) for its condition;
        texture: None = `sh2rgb` (SH colour inside the fused posing kernel) or a ColorMLPFused = `mlp` (texture.py:128-134),
        in which case the Gaussians carry feature vectors (`use_sh: false`, gaussian_model.py:45-46,182) instead of SH.
        Training-mode randomness of the reference (module in .train() mode only): view_noise (degrees, cfg.model.texture.view_noise:
        rotation of the canonical view direction, texture.py:28-33 / :94-99) and pose_noise (cfg.pipeline.pose_noise: with
        probability 1/2 Gaussian noise on camera.rots, the input of the pose encoder, gaussian_converter.py:47-50)."""
        super().__init__()
        self.rigid, self.non_rigid, self.pose_encoder, self.cano_view_dir = rigid, non_rigid, pose_encoder, cano_view_dir
        self.texture, self.view_noise, self.pose_noise = texture, float(view_noise), float(pose_noise)

    def forward(self, gaussians, camera, iteration, compute_loss=True):
        g = gaussians
        xyz, scaling, rotation = g._xyz, g._scaling, g._rotation
        loss_reg = {}
        deltas = None
        vn = None
        if self.training:
            import copy
            import numpy as np
            if self.pose_noise > 0 and np.random.uniform() <= 0.5 and getattr(camera, "rots", None) is not None:
                camera = copy.copy(camera)  # gaussian_converter.py:47-50
                camera.rots = camera.rots + torch.randn(camera.rots.shape, device=camera.rots.device) * self.pose_noise
            if self.view_noise > 0 and self.cano_view_dir:
                vn = sample_view_noise(self.view_noise).to(xyz.device)
                if self.texture is not None:
                    camera = copy.copy(camera)
                    camera.view_noise = vn  # read by ColorMLPFused.forward in training mode
        if self.non_rigid is not None and iteration >= self.non_rigid.delay:  # non_rigid.py:227-232
            cond = self.pose_encoder(camera.rots, camera.Jtrs) if self.pose_encoder is not None else \
                torch.zeros(self.non_rigid.n_cond, device=xyz.device)
            deform = self.non_rigid.deform_autograd if torch.is_grad_enabled() else self.non_rigid.deform
            xyz, scaling, rotation, deltas = deform(xyz, scaling, rotation, cond)
            if compute_loss:  # non_rigid.py:288-297
                loss_reg = {"nr_xyz": deltas[:, :3].norm(p=2, dim=1).mean(), "nr_scale": deltas[:, 3:6].norm(p=1, dim=1).mean(),
                            "nr_rot": deltas[:, 7:10].norm(p=1, dim=1).mean()}
        query = getattr(self.rigid, "query_weights_autograd", None) if torch.is_grad_enabled() else None
        weights = (query or self.rigid.query_weights)(xyz)  # at the non-rigidly displaced position (rigid.py:217-220)
        feats = torch.cat([g._features_dc, g._features_rest], 1)
        if self.texture is None:
            f_dc, f_rest, deg = g._features_dc, g._features_rest, g.active_sh_degree
        else:  # colours come from the texture MLP: the posing kernel gets a constant degree-0 stand-in
            f_dc, f_rest, deg = xyz.new_zeros(xyz.shape[0], 1, 3), xyz.new_zeros(xyz.shape[0], 0, 3), 0
        means, cov, colors, opacity, T_fwd, Rbar = fused_pose(
            xyz, rotation, scaling, g._opacity, f_dc, f_rest, weights, camera.bone_transforms,
            camera.camera_center, deg, 1.0, self.cano_view_dir, return_aux=True, view_noise=vn if self.texture is None else None)
        deformed = PosedGaussians(means, cov, opacity, scaling, rotation, feats, T_fwd, Rbar, g.active_sh_degree,
                                  getattr(g, "max_sh_degree", 3))
        if self.texture is not None:
            tex = self.texture
            if tex.non_rigid_dim > 0:  # non_rigid.py:59-60 (zeros before the delay), :87,119-120
                deformed.non_rigid_feature = deltas[:, 10:10 + tex.non_rigid_dim] if deltas is not None else \
                    xyz.new_zeros(xyz.shape[0], tex.non_rigid_dim)
            colors = tex(deformed, camera)  # ColorMLP.forward(deformed_gaussians, camera), gaussian_converter.py:57
        return deformed, loss_reg, colors
