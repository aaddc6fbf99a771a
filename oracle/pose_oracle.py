"""torch-CPU ORACLE of the posing stage (rigid LBS -> covariance -> SH colour, + optional non-rigid).

TEST INFRASTRUCTURE ONLY (tests/, smoke(), bench.py CPU legs).  It restates, op for op, the reference's
PyTorch chain; the reference modules themselves cannot be imported (pytorch3d / tinycudann / trimesh / igl
missing, device="cuda" hard-coded in utils/general_utils.py:74,92,195), see SURVEY.md §8(c).
`eval_sh` is restated here and pinned against the reference's own utils/sh_utils.py in
tests/test_oracle_pose.py (that file is importable on CPU in the build container).

All functions are differentiable torch code, so autograd through this module is the gradient oracle of
the fused posing kernels as well.
"""
import torch

# utils/sh_utils.py:27-44
C0 = 0.28209479177387814
C1 = 0.4886025119029199
C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
      1.445305721320277, -0.5900435899266435]


def eval_sh(deg, sh, dirs):
    """utils/sh_utils.py:58-113 (degrees 0-3).  sh [..., C, (deg+1)^2], dirs [..., 3] -> [..., C]."""
    assert 0 <= deg <= 3
    result = C0 * sh[..., 0]
    if deg > 0:
        x, y, z = dirs[..., 0:1], dirs[..., 1:2], dirs[..., 2:3]
        result = (result - C1 * y * sh[..., 1] + C1 * z * sh[..., 2] - C1 * x * sh[..., 3])
        if deg > 1:
            xx, yy, zz = x * x, y * y, z * z
            xy, yz, xz = x * y, y * z, x * z
            result = (result + C2[0] * xy * sh[..., 4] + C2[1] * yz * sh[..., 5] +
                      C2[2] * (2.0 * zz - xx - yy) * sh[..., 6] + C2[3] * xz * sh[..., 7] +
                      C2[4] * (xx - yy) * sh[..., 8])
            if deg > 2:
                result = (result + C3[0] * y * (3 * xx - yy) * sh[..., 9] + C3[1] * xy * z * sh[..., 10] +
                          C3[2] * y * (4 * zz - xx - yy) * sh[..., 11] +
                          C3[3] * z * (2 * zz - 3 * xx - 3 * yy) * sh[..., 12] +
                          C3[4] * x * (4 * zz - xx - yy) * sh[..., 13] + C3[5] * z * (xx - yy) * sh[..., 14] +
                          C3[6] * x * (xx - 3 * yy) * sh[..., 15])
    return result


def build_rotation(r):
    """utils/general_utils.py:87-108: normalise (w,x,y,z), rotation matrix [P,3,3]."""
    norm = torch.sqrt(r[:, 0] * r[:, 0] + r[:, 1] * r[:, 1] + r[:, 2] * r[:, 2] + r[:, 3] * r[:, 3])
    q = r / norm[:, None]
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = torch.stack([
        1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
        2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
        2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], dim=-1).reshape(-1, 3, 3)
    return R


def quaternion_multiply(a, b):
    """Hamilton product a (x) b, real part first: what pytorch3d.transforms.quaternion_multiply computes
    (models/deformer/non_rigid.py:114,281; pytorch3d 0.7.1 is not vendored -- its sign standardisation to a
    non-negative real part is omitted because build_rotation is invariant to it).  R(a (x) b) = R(a) R(b)."""
    aw, ax, ay, az = a.unbind(-1)
    bw, bx, by, bz = b.unbind(-1)
    return torch.stack([aw * bw - ax * bx - ay * by - az * bz, aw * bx + ax * bw + ay * bz - az * by,
                        aw * by - ax * bz + ay * bw + az * bx, aw * bz + ax * by - ay * bx + az * bw], dim=-1)


def quaternion_multiply_general_utils(r, s):
    """utils/general_utils.py:184-192 restated as written: the cross-product part has the opposite sign of the Hamilton
    product, i.e. it returns s (x) r.  Not on the hot path (non_rigid.py:281 uses pytorch3d's; the MLP variant :185
    uses this one); kept for reference."""
    r0, r1, r2, r3 = r.unbind(-1)
    s0, s1, s2, s3 = s.unbind(-1)
    return torch.stack([r0 * s0 - r1 * s1 - r2 * s2 - r3 * s3, r0 * s1 + r1 * s0 - r2 * s3 + r3 * s2,
                        r0 * s2 + r1 * s3 + r2 * s0 - r3 * s1, r0 * s3 - r1 * s2 + r2 * s1 + r3 * s0], dim=-1)


def aabb_normalize(x, coord_min, coord_max, sym=True):
    """utils/dataset_utils.py:66-76."""
    x = (x - coord_min) / (coord_max - coord_min)
    return 2 * x - 1. if sym else x


def lbs(xyz, rotation, pts_W, bone_transforms):
    """models/deformer/rigid.py:49-71 (SMPLNN.forward with the weights already queried) ==
    :215-236 (SkinningField.forward after get_forward_transform).
    Returns T_fwd [P,4,4], x_bar [P,3], rotation_bar [P,3,3]."""
    n_pts = xyz.shape[0]
    T_fwd = torch.matmul(pts_W, bone_transforms.view(-1, 16)).view(n_pts, 4, 4).to(xyz.dtype)  # .float() in the reference
    homo = torch.ones(n_pts, 1, dtype=xyz.dtype, device=xyz.device)
    x_hat_homo = torch.cat([xyz, homo], dim=-1).view(n_pts, 4, 1)
    x_bar = torch.matmul(T_fwd, x_hat_homo)[:, :3, 0]
    rotation_hat = build_rotation(rotation)
    rotation_bar = torch.matmul(T_fwd[:, :3, :3], rotation_hat)
    return T_fwd, x_bar, rotation_bar


def covariance(rotation_bar, scaling_log, scaling_modifier=1.0):
    """scene/gaussian_model.py:154-157,28-32 + utils/general_utils.py:194-207,73-85:
    L = R_bar diag(mod*exp(s)); Sigma = L L^T; packed [00,01,02,11,12,22]."""
    s = scaling_modifier * torch.exp(scaling_log)
    L = rotation_bar * s[:, None, :]  # R @ diag(s)
    cov = L @ L.transpose(1, 2)
    return torch.stack([cov[:, 0, 0], cov[:, 0, 1], cov[:, 0, 2], cov[:, 1, 1], cov[:, 1, 2], cov[:, 2, 2]], dim=-1)


def sh2rgb(features, x_bar, T_fwd, camera_center, active_sh_degree, cano_view_dir=True, view_noise=None):
    """models/texture/texture.py:21-38.  features [P,16,3].  `view_noise` [3,3] (training mode, :28-33): the matrix the canonical
    view direction is multiplied with (augm_rots(...).T), None in eval mode.
    T_fwd is used detached, as set_fwd_transform(T_fwd.detach()) rigid.py:223."""
    shs_view = features.transpose(1, 2).reshape(features.shape[0], 3, -1)
    dir_pp = x_bar - camera_center.reshape(1, 3)
    if cano_view_dir:
        R_bwd = T_fwd.detach()[:, :3, :3].transpose(1, 2)
        dir_pp = torch.matmul(R_bwd, dir_pp.unsqueeze(-1)).squeeze(-1)
        if view_noise is not None:
            dir_pp = torch.matmul(dir_pp, view_noise.to(dir_pp.dtype))
    dir_n = dir_pp / (dir_pp.norm(dim=1, keepdim=True) + 1e-12)
    return torch.clamp_min(eval_sh(active_sh_degree, shs_view, dir_n) + 0.5, 0.0)


def pose_forward(xyz, rotation, scaling, opacity, features, pts_W, bone_transforms, camera_center,
                 active_sh_degree=3, scaling_modifier=1.0, cano_view_dir=True, view_noise=None):
    """The whole rigid posing stage (config 1 of BASELINE.json): what gaussian_renderer/__init__.py:73-129
    hands to the rasterizer.  Returns dict(means3D, cov3D, colors, opacity, T_fwd, rotation_bar)."""
    T_fwd, x_bar, R_bar = lbs(xyz, rotation, pts_W, bone_transforms)
    cov = covariance(R_bar, scaling, scaling_modifier)
    col = sh2rgb(features, x_bar, T_fwd, camera_center, active_sh_degree, cano_view_dir, view_noise)
    return dict(means3D=x_bar, cov3D=cov, colors=col, opacity=torch.sigmoid(opacity), T_fwd=T_fwd, rotation_bar=R_bar)
