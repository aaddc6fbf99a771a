"""Deterministic synthetic inputs for the hot path.

Everything is generated on the CPU with seeded numpy generators and returned as float32 numpy arrays, so
the CPU oracle, the rebuilt reference and the CUDA path see identical bits.

Conventions follow the reference:
  * camera matrices: scene/cameras.py:35-40 + utils/graphics_utils.py:38-71 (row-vector storage, i.e. the
    flat float[16] is the column-major standard matrix);
  * bone transforms: forward kinematics with the rest-pose joint removed, as
    models/pose_correction/lbs.py:208-263 (batch_rigid_transform);
  * Gaussian initialisation mirrors scene/gaussian_model.py:172-199 (kNN-derived isotropic log-scales,
    identity quaternion and opacity 0.1 for the "random-init" training config).
"""
import math

import numpy as np

# SMPL kinematic tree (models/network_utils.py:128-129)
PARENTS = np.array([-1, 0, 0, 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 9, 9, 12, 13, 14, 16, 17, 18, 19, 20, 21])

# A hand-made 24-joint humanoid rest pose in metres (legs slightly apart, arms ~45 degrees down: a
# "star pose"-like canonical body).  Not SMPL data: the licensed model files are absent.
_J = np.array([
    [0.00, 0.92, 0.00],    # 0 pelvis
    [0.10, 0.84, 0.00],    # 1 l_hip
    [-0.10, 0.84, 0.00],   # 2 r_hip
    [0.00, 1.04, -0.02],   # 3 spine1
    [0.19, 0.47, 0.00],    # 4 l_knee
    [-0.19, 0.47, 0.00],   # 5 r_knee
    [0.00, 1.17, -0.01],   # 6 spine2
    [0.27, 0.08, -0.03],   # 7 l_ankle
    [-0.27, 0.08, -0.03],  # 8 r_ankle
    [0.00, 1.23, 0.00],    # 9 spine3
    [0.30, 0.02, 0.09],    # 10 l_foot
    [-0.30, 0.02, 0.09],   # 11 r_foot
    [0.00, 1.44, -0.02],   # 12 neck
    [0.08, 1.36, -0.01],   # 13 l_collar
    [-0.08, 1.36, -0.01],  # 14 r_collar
    [0.00, 1.53, 0.02],    # 15 head
    [0.19, 1.38, -0.02],   # 16 l_shoulder
    [-0.19, 1.38, -0.02],  # 17 r_shoulder
    [0.38, 1.20, -0.03],   # 18 l_elbow
    [-0.38, 1.20, -0.03],  # 19 r_elbow
    [0.56, 1.03, -0.02],   # 20 l_wrist
    [-0.56, 1.03, -0.02],  # 21 r_wrist
    [0.62, 0.97, -0.02],   # 22 l_hand
    [-0.62, 0.97, -0.02],  # 23 r_hand
], dtype=np.float64)
_BONE_RADIUS = {0: 0.12}


def skeleton():
    """(joints [24,3] float64 metres, parents [24])."""
    return _J.copy(), PARENTS.copy()


def _segments():
    """The 23 bones as (child index, a, b, radius); plus a head blob."""
    segs = []
    for c in range(1, 24):
        p = PARENTS[c]
        r = 0.05
        if c in (1, 2, 3, 6, 9):
            r = 0.12
        elif c in (4, 5):
            r = 0.08
        elif c in (7, 8, 18, 19):
            r = 0.055
        elif c in (12, 13, 14, 16, 17):
            r = 0.07
        elif c == 15:
            r = 0.10
        elif c in (10, 11, 20, 21, 22, 23):
            r = 0.04
        segs.append((c, _J[p], _J[c], r))
    return segs


def canonical_gaussians(P=50_000, seed=0, random_init=False, sh_degree=3):
    """Canonical-space Gaussians: points in capsules around the bones (a surface-ish avatar proxy).

    Returns dict of float32 arrays: xyz [P,3], scaling (log) [P,3], rotation (w,x,y,z) [P,4], opacity (logit)
    [P,1], features_dc [P,1,3], features_rest [P,15,3], plus bone (nearest segment's child joint) [P].
    """
    from scipy.spatial import cKDTree

    rng = np.random.default_rng(seed)
    segs = _segments()
    lens = np.array([np.linalg.norm(b - a) + 2 * r for (_, a, b, r) in segs])
    vol = lens * np.array([r * r for (_, _, _, r) in segs])
    cnt = np.floor(P * vol / vol.sum()).astype(int)
    cnt[0] += P - cnt.sum()
    pts, bone = [], []
    for (c, a, b, r), n in zip(segs, cnt):
        t = rng.uniform(-0.15, 1.15, size=(n, 1))
        d = rng.normal(size=(n, 3))
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        rad = r * rng.uniform(0.6, 1.0, size=(n, 1)) ** 0.5  # shell-weighted
        pts.append(a + t * (b - a) + d * rad)
        bone.append(np.full(n, c))
    xyz = np.concatenate(pts).astype(np.float32)
    bone = np.concatenate(bone)
    perm = rng.permutation(P)
    xyz, bone = xyz[perm], bone[perm]
    # scene/gaussian_model.py:186-187: scales = log(sqrt(mean sq. dist to 3 NN)) replicated x3
    d, _ = cKDTree(xyz).query(xyz, k=4)
    dist2 = np.clip((d[:, 1:] ** 2).mean(axis=1), 1e-7, None)
    scaling = np.repeat(np.log(np.sqrt(dist2))[:, None], 3, axis=1)
    if random_init:
        rot = np.zeros((P, 4))
        rot[:, 0] = 1.0
        opacity = np.full((P, 1), math.log(0.1 / 0.9))
    else:
        scaling = scaling + rng.normal(0, 0.3, size=(P, 3))
        rot = rng.normal(size=(P, 4))
        rot /= np.linalg.norm(rot, axis=1, keepdims=True)
        o = rng.uniform(0.05, 0.95, size=(P, 1))
        opacity = np.log(o / (1 - o))
    n_rest = (sh_degree + 1) ** 2 - 1
    f_dc = rng.normal(0, 1.0, size=(P, 1, 3))
    f_rest = rng.normal(0, 0.1, size=(P, n_rest, 3))
    f32 = lambda a: np.ascontiguousarray(a, dtype=np.float32)
    return dict(xyz=f32(xyz), scaling=f32(scaling), rotation=f32(rot), opacity=f32(opacity), features_dc=f32(f_dc),
                features_rest=f32(f_rest), bone=bone)


def skinning_weights(xyz, sigma=0.08, k=4):
    """[P,24] dense weights: softmax(-d^2/sigma^2) over the k nearest bones (distance to the bone segment;
    a bone's weight goes to its PARENT joint, which is the joint that moves it), zero elsewhere."""
    P = xyz.shape[0]
    x = xyz.astype(np.float64)
    d2 = np.full((P, 24), np.inf)
    for (c, a, b, _r) in _segments():
        ab = b - a
        t = np.clip(((x - a) @ ab) / (ab @ ab), 0, 1)
        q = a + t[:, None] * ab
        j = PARENTS[c]
        d2[:, j] = np.minimum(d2[:, j], ((x - q) ** 2).sum(1))
    finite = np.isfinite(d2)
    d2f = np.where(finite, d2, 1e9)
    idx = np.argsort(d2f, axis=1)[:, :k]
    sel = np.take_along_axis(d2f, idx, 1)
    w = np.exp(-(sel - sel[:, :1]) / sigma ** 2)
    w /= w.sum(1, keepdims=True)
    W = np.zeros((P, 24))
    np.put_along_axis(W, idx, w, 1)
    return W.astype(np.float32)


def _rodrigues(aa):
    th = np.linalg.norm(aa)
    if th < 1e-12:
        return np.eye(3)
    k = aa / th
    K = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.eye(3) + math.sin(th) * K + (1 - math.cos(th)) * (K @ K)


def bone_transforms_from_rotmats(R, trans=None):
    """FK exactly as batch_rigid_transform (models/pose_correction/lbs.py:208-263): per-joint world
    transforms with the rest-pose joint location removed -> [24,4,4] float32."""
    J = _J
    G = np.zeros((24, 4, 4))
    for j in range(24):
        T = np.eye(4)
        T[:3, :3] = R[j]
        T[:3, 3] = J[j] - (J[PARENTS[j]] if j > 0 else 0)
        G[j] = T if j == 0 else G[PARENTS[j]] @ T
    out = G.copy()
    for j in range(24):
        out[j, :3, 3] = G[j, :3, 3] - G[j, :3, :3] @ J[j]
    if trans is not None:
        out[:, :3, 3] += np.asarray(trans)
    return out.astype(np.float32)


def pose(seed=1, sigma=0.35):
    """One synthetic SMPL-like pose: per-joint axis-angle ~ N(0, sigma^2), root = identity.
    Returns dict(rotmats [24,3,3] f64, bone_transforms [24,4,4] f32, rots [1,24,9] f32, Jtrs [1,24,3] f32)."""
    rng = np.random.default_rng(seed)
    aa = rng.normal(0, sigma, size=(24, 3))
    aa[0] = 0
    R = np.stack([_rodrigues(a) for a in aa])
    return _pose_dict(R)


def _pose_dict(R):
    lo, hi = _J.min(0) - 0.2, _J.max(0) + 0.2
    Jn = 2 * (_J - lo) / (hi - lo) - 1
    return dict(rotmats=R, bone_transforms=bone_transforms_from_rotmats(R), rots=R.reshape(1, 24, 9).astype(np.float32),
                Jtrs=Jn.reshape(1, 24, 3).astype(np.float32))


def _quat_from_R(R):
    from scipy.spatial.transform import Rotation
    return Rotation.from_matrix(R)


def pose_sequence(n_frames=64, n_keys=5, seed=1, sigma=0.35):
    """Smooth "AMASS-like" sequence: slerp between n_keys seeded key poses (seed+k)."""
    from scipy.spatial.transform import Rotation, Slerp
    keys = [pose(seed + k, sigma)["rotmats"] for k in range(n_keys)]
    tk = np.linspace(0, 1, n_keys)
    tf = np.linspace(0, 1, n_frames)
    Rs = np.zeros((n_frames, 24, 3, 3))
    for j in range(24):
        sl = Slerp(tk, Rotation.from_matrix(np.stack([k[j] for k in keys])))
        Rs[:, j] = sl(tf).as_matrix()
    return [_pose_dict(Rs[f]) for f in range(n_frames)]


def aabb():
    """Canonical bounding box used by AABB.normalize (utils/dataset_utils.py:66-76)."""
    lo, hi = _J.min(0) - 0.25, _J.max(0) + 0.25
    return lo.astype(np.float32), hi.astype(np.float32)


def camera(W=512, H=512, fx=537.0, fy=537.0, dist=3.5, azimuth_deg=0.0, target=None, znear=0.01, zfar=100.0):
    """ZJU-377-style camera looking at the pelvis from `dist` metres.  Returns a dict with the fields
    gaussian_renderer/__init__.py:82-98 reads: FoVx, FoVy, image_height, image_width,
    world_view_transform [4,4], full_proj_transform [4,4], camera_center [3] (float32, row-vector storage)."""
    target = _J[0] if target is None else np.asarray(target, dtype=np.float64)
    az = math.radians(azimuth_deg)
    eye = target + dist * np.array([math.sin(az), 0.0, math.cos(az)])
    fwd = target - eye
    fwd /= np.linalg.norm(fwd)
    up = np.array([0.0, -1.0, 0.0])  # image y points down: world +y (up) maps to -y in the image
    right = np.cross(up, fwd)
    right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    Rw2c = np.stack([right, down, fwd])  # rows: camera axes in world coordinates
    R = Rw2c.T  # the reference stores R transposed (getWorld2View2 uses R.transpose())
    T = -Rw2c @ eye
    Rt = np.zeros((4, 4))
    Rt[:3, :3] = R.T
    Rt[:3, 3] = T
    Rt[3, 3] = 1.0
    w2v = np.float32(Rt)  # getWorld2View2 with translate=0, scale=1 (utils/graphics_utils.py:38-49)
    fovx, fovy = 2 * math.atan(W / (2 * fx)), 2 * math.atan(H / (2 * fy))
    tx, ty = math.tan(fovx / 2), math.tan(fovy / 2)
    top, right_ = ty * znear, tx * znear
    Pm = np.zeros((4, 4), dtype=np.float32)  # utils/graphics_utils.py:51-71
    Pm[0, 0] = 2.0 * znear / (2 * right_)
    Pm[1, 1] = 2.0 * znear / (2 * top)
    Pm[3, 2] = 1.0
    Pm[2, 2] = zfar / (zfar - znear)
    Pm[2, 3] = -(zfar * znear) / (zfar - znear)
    wvt = np.ascontiguousarray(w2v.T)  # scene/cameras.py:35
    proj_t = np.ascontiguousarray(Pm.T)
    full = (wvt @ proj_t).astype(np.float32)  # :37-39
    center = np.linalg.inv(wvt.astype(np.float64))[3, :3].astype(np.float32)  # :40
    return dict(FoVx=fovx, FoVy=fovy, image_width=W, image_height=H, world_view_transform=wvt,
                full_proj_transform=full, camera_center=center, tanfovx=tx, tanfovy=ty, znear=znear, zfar=zfar)


def ring_cameras(n, **kw):
    return [camera(azimuth_deg=360.0 * i / n, **kw) for i in range(n)]


def l1_target(H, W, seed=2):
    """Seeded random target image for the L1 loss of train.py:121."""
    return np.random.default_rng(seed).uniform(0, 1, size=(3, H, W)).astype(np.float32)
