"""Runs fwd+bwd of a chosen subset of the default-configuration chain (non-rigid hash grid MLP / skinning field / colour MLP /
pose encoder -> fused posing -> rasterizer with opacity pass -> L1 + mask loss), eagerly or as a CUDA graph.

    python profiles/default_chain_run.py P W iters graph|eager nr+sk+tex+enc

Stages that are left out are replaced by the reference's simpler registry choice (identity non-rigid, fixed skinning weights,
SH colour in the posing kernel).  With CUDA_LAUNCH_BLOCKING=1 and `eager` a device fault surfaces at the call that caused it."""
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "3dgs-avatar-release_b200"), ROOT):
    sys.path.insert(0, p)
import numpy as np
import torch

import synth
from avatar_pose.graphs import GraphedStep
from avatar_pose.loss import l1_loss
from avatar_pose.modules import FusedGaussianConverter
from avatar_pose.nonrigid import HashGridMLPFused
from avatar_pose.pose_encoder import HierarchicalPoseEncoderFused
from avatar_pose.skinning import SkinningFieldFused
from avatar_pose.texture import ColorMLPFused
from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer


class FixedWeights(torch.nn.Module):
    def __init__(self, w):
        super().__init__()
        self.register_buffer("w", w)

    def query_weights(self, xyz):
        return self.w


def main():
    P, W, iters, mode, variant = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4], set(sys.argv[5].split("+"))
    H = W
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    g = synth.canonical_gaussians(P, seed=0, random_init=True)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    tex = "tex" in variant
    gm = types.SimpleNamespace(active_sh_degree=0 if tex else 3, max_sh_degree=0 if tex else 3)
    for k in ("xyz", "rotation", "scaling", "opacity"):
        setattr(gm, "_" + k, t(g[k]).requires_grad_(True))
    if tex:  # feature vectors (use_sh: false)
        gm._features_dc = (0.5 * torch.randn(P, 1, 1, device=dev)).requires_grad_(True)
        gm._features_rest = (0.5 * torch.randn(P, 31, 1, device=dev)).requires_grad_(True)
    else:    # spherical harmonics, degree 3
        gm._features_dc = t(g["features_dc"]).requires_grad_(True)
        gm._features_rest = t(g["features_rest"]).requires_grad_(True)
    lo, hi = (g["xyz"].min(0) - 0.05).tolist(), (g["xyz"].max(0) + 0.05).tolist()
    nr = HashGridMLPFused(n_cond=144, feature_dim=16 if tex else 0, aabb_min=lo, aabb_max=hi, delay=0) if "nr" in variant else None
    if nr is not None:
        with torch.no_grad():
            nr.lin3.weight.mul_(0.01)
    rigid = SkinningFieldFused(aabb_min=lo, aabb_max=hi) if "sk" in variant else FixedWeights(t(synth.skinning_weights(g["xyz"])))
    enc = HierarchicalPoseEncoderFused() if "enc" in variant else None
    conv = FusedGaussianConverter(rigid, nr, enc, texture=ColorMLPFused(n_frames=8) if tex else None).to(dev).eval()
    fr = dict(cam=synth.camera(W=W, H=H, fx=537.0 * W / 512.0, fy=537.0 * H / 512.0), pose=synth.pose(1), target=synth.l1_target(H, W, seed=2))
    cam = fr["cam"]
    bones, view, proj, campos = (t(x) for x in (fr["pose"]["bone_transforms"], cam["world_view_transform"], cam["full_proj_transform"], cam["camera_center"]))
    rots, Jtrs, target = t(fr["pose"]["rots"]), t(fr["pose"]["Jtrs"]), t(fr["target"])
    bg = torch.zeros(3, device=dev)
    params = list(conv.parameters()) + [getattr(gm, "_" + k) for k in ("xyz", "rotation", "scaling", "opacity", "features_dc", "features_rest")]
    sync = os.environ.get("CHAIN_SYNC") == "1"

    def mark(what):
        if sync:
            torch.cuda.synchronize()
            print("  ok:", what, flush=True)

    def step():
        for p in params:
            p.grad = None
        camera = types.SimpleNamespace(bone_transforms=bones, camera_center=campos, rots=rots, Jtrs=Jtrs, latent_idx=3)
        deformed, _, colors = conv(gm, camera, 100)
        mark("converter forward")
        rs = GaussianRasterizationSettings(H, W, cam["tanfovx"], cam["tanfovy"], bg, 1.0, view, proj, gm.active_sh_degree, campos, False, False)
        means2D = torch.zeros_like(deformed.get_xyz, requires_grad=True)
        img, radii, opim = GaussianRasterizer(rs)(means3D=deformed.get_xyz, means2D=means2D, opacities=deformed.get_opacity,
                                                  colors_precomp=colors, cov3D_precomp=deformed.get_covariance(1.0), return_opacity=True)
        mark("rasterizer forward")
        loss = l1_loss(img, target) + 0.1 * l1_loss(opim, target[:1])
        mark("loss")
        loss.backward()
        mark("backward")
        return loss

    fn = step if mode == "eager" else GraphedStep(step).replay
    for _ in range(iters):
        loss = fn()
    torch.cuda.synchronize()
    fin = all(p.grad is None or bool(torch.isfinite(p.grad).all()) for p in params)
    none = [i for i, p in enumerate(params) if p.grad is None]
    print(f"CHAIN OK P={P} W={W} mode={mode} variant={'+'.join(sorted(variant))} loss={float(loss):.6f} finite={fin} params_without_grad={none}", flush=True)


if __name__ == "__main__":
    main()
