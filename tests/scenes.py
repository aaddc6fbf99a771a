"""Seeded test scenes shared by the CPU and GPU tests (numpy in, numpy out)."""
import numpy as np
import torch

import synth
from oracle import pose_oracle


def avatar_scene(P=3000, W=128, H=128, seed=0, pose_seed=1, random_init=False, fx=None, dist=3.5, azimuth=0.0):
    """Posed avatar (colors_precomp + cov3D_precomp path = what gaussian_renderer hands to the rasterizer)."""
    g = synth.canonical_gaussians(P, seed=seed, random_init=random_init)
    wts = synth.skinning_weights(g["xyz"])
    ps = synth.pose(pose_seed)
    fx = fx if fx is not None else 537.0 * W / 512.0
    cam = synth.camera(W=W, H=H, fx=fx, fy=fx, dist=dist, azimuth_deg=azimuth)
    T = torch.from_numpy
    feat = torch.cat([T(g["features_dc"]), T(g["features_rest"])], 1)
    with torch.no_grad():
        posed = pose_oracle.pose_forward(T(g["xyz"]), T(g["rotation"]), T(g["scaling"]), T(g["opacity"]), feat, T(wts),
                                         T(ps["bone_transforms"]), T(cam["camera_center"]))
    return dict(cam=cam, W=W, H=H, bg=np.zeros(3, np.float32), means3D=posed["means3D"].numpy(),
                opacities=posed["opacity"].numpy().reshape(-1, 1), colors_precomp=posed["colors"].numpy(),
                cov3D_precomp=posed["cov3D"].numpy(), gaussians=g, weights=wts, pose=ps)


def generic_scene(P=2000, W=160, H=96, seed=3, bg=(0.2, 0.5, 0.9), spread=1.0, scale_lo=-4.5, scale_hi=-2.0):
    """Random cloud in front of (and partly behind / beside) the camera, for the SH + scale/rotation path,
    non-zero background, non-multiple-of-16 sizes and culling edge cases."""
    rng = np.random.default_rng(seed)
    cam = synth.camera(W=W, H=H, fx=0.9 * W, fy=0.9 * W, dist=3.0, azimuth_deg=25.0)
    means = rng.normal(0, spread, size=(P, 3)).astype(np.float32) + np.array([0, 0.92, 0], np.float32)
    means[: P // 20] += np.array([0, 0, 6.0], np.float32)  # some behind the camera (culled)
    scales = np.exp(rng.uniform(scale_lo, scale_hi, size=(P, 3))).astype(np.float32)
    rots = rng.normal(size=(P, 4)).astype(np.float32)
    rots /= np.linalg.norm(rots, axis=1, keepdims=True)
    op = rng.uniform(0.02, 0.99, size=(P, 1)).astype(np.float32)
    shs = np.concatenate([rng.normal(0, 1.0, size=(P, 1, 3)), rng.normal(0, 0.3, size=(P, 15, 3))], 1).astype(np.float32)
    return dict(cam=cam, W=W, H=H, bg=np.asarray(bg, np.float32), means3D=means, opacities=op, shs=shs, scales=scales,
                rotations=rots)


def oracle_forward(sc, sh_degree=3, scale_modifier=1.0):
    from oracle import gs_oracle
    cam = sc["cam"]
    return gs_oracle.forward(sc["means3D"], sc["opacities"], sc["W"], sc["H"], cam["tanfovx"], cam["tanfovy"],
                             cam["world_view_transform"], cam["full_proj_transform"], cam["camera_center"], sc["bg"],
                             colors_precomp=sc.get("colors_precomp"), shs=sc.get("shs"), sh_degree=sh_degree,
                             scales=sc.get("scales"), rotations=sc.get("rotations"), scale_modifier=scale_modifier,
                             cov3D_precomp=sc.get("cov3D_precomp"))


def l1_grad(img, seed=2):
    """dL/dpix of the L1 loss against a seeded random target (train.py:121)."""
    tgt = synth.l1_target(img.shape[1], img.shape[2], seed)
    return (np.sign(img - tgt) / img.size).astype(np.float32)
