"""GPU test of the posing drop-in boundary: FusedGaussianConverter(gaussians, camera, iteration) against the oracle chain
(1-NN skinning weights -> pose oracle), then the deformed set drives the drop-in rasterizer exactly as
gaussian_renderer/__init__.py:100-129 does."""
import os
import types

import numpy as np
import pytest
import torch

import synth
from oracle import pose_oracle

pytestmark = pytest.mark.gpu


def test_converter_smpl_nn_matches_oracle_and_renders(built_lib):
    from avatar_pose.modules import FusedGaussianConverter, SMPLNNFused
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    dev = torch.device("cuda:0")
    P = 5000
    g = synth.canonical_gaussians(P, seed=0)
    rng = np.random.default_rng(4)
    verts = g["xyz"][rng.choice(P, 700, replace=False)] + rng.normal(0, 0.01, size=(700, 3)).astype(np.float32)
    vw = synth.skinning_weights(verts)
    gm = types.SimpleNamespace(active_sh_degree=3, max_sh_degree=3)
    for k in ("xyz", "rotation", "scaling", "opacity", "features_dc", "features_rest"):
        setattr(gm, "_" + k, torch.from_numpy(g[k]).to(dev).requires_grad_(True))
    ps, cam = synth.pose(2), synth.camera(W=128, H=128, fx=134.0, fy=134.0)
    camera = types.SimpleNamespace(bone_transforms=torch.from_numpy(ps["bone_transforms"]).to(dev),
                                   camera_center=torch.from_numpy(cam["camera_center"]).to(dev))
    conv = FusedGaussianConverter(SMPLNNFused(verts, vw)).to(dev)
    deformed, loss_reg, colors = conv(gm, camera, iteration=0)
    # oracle chain
    d2 = ((g["xyz"][:, None, :].astype(np.float64) - verts[None].astype(np.float64)) ** 2).sum(-1)
    w_ref = torch.from_numpy(vw[d2.argmin(1)]).double()
    T = lambda a: torch.from_numpy(a).double()  # fp64 truth (fp32 CPU matmuls are not trustworthy on every host)
    feat = torch.cat([T(g["features_dc"]), T(g["features_rest"])], 1)
    o = pose_oracle.pose_forward(T(g["xyz"]), T(g["rotation"]), T(g["scaling"]), T(g["opacity"]), feat, w_ref,
                                 T(ps["bone_transforms"]), T(cam["camera_center"]), 3)
    assert loss_reg == {}
    assert float((deformed.get_xyz.detach().cpu().double() - o["means3D"]).abs().max()) <= 1e-5
    assert float((colors.detach().cpu().double() - o["colors"]).abs().max()) <= 1e-5
    assert float((deformed.get_opacity.detach().cpu().double() - o["opacity"]).abs().max()) <= 1e-6
    cov = deformed.get_covariance(1.0).detach().cpu().double()
    assert float((cov - o["cov3D"]).abs().max()) <= 1e-5 * float(o["cov3D"].abs().max())
    assert torch.allclose(deformed.get_covariance(0.5), deformed.get_covariance(1.0) * 0.25)
    assert float((deformed.fwd_transform.cpu().double() - o["T_fwd"]).abs().max()) <= 1e-5
    # render exactly like gaussian_renderer/__init__.py
    d = lambda a: torch.as_tensor(np.ascontiguousarray(a)).to(dev)
    rs = GaussianRasterizationSettings(128, 128, cam["tanfovx"], cam["tanfovy"], torch.zeros(3, device=dev), 1.0,
                                       d(cam["world_view_transform"]), d(cam["full_proj_transform"]), 3, d(cam["camera_center"]), False, False)
    screenspace = torch.zeros_like(deformed.get_xyz, requires_grad=True)
    img, radii = GaussianRasterizer(rs)(means3D=deformed.get_xyz, means2D=screenspace, opacities=deformed.get_opacity,
                                        colors_precomp=colors, cov3D_precomp=deformed.get_covariance(1.0))
    img.sum().backward()
    assert (radii > 0).sum() > 100 and torch.isfinite(img).all()
    assert gm._xyz.grad is not None and torch.isfinite(gm._xyz.grad).all() and float(gm._xyz.grad.abs().max()) > 0
    assert gm._features_rest.grad is not None and screenspace.grad is not None


def test_converter_full_chain_trains_every_parameter(built_lib):
    """non-rigid (hash grid + MLP) -> skinning field -> fused posing -> rasterizer -> loss.backward(): every trainable
    tensor of the chain (Gaussians, hash table, both MLPs) receives a finite, non-zero gradient; a small step down the
    gradient (max |delta| 1e-4 per tensor) lowers an L2 image loss against a perturbed target."""
    from avatar_pose.modules import FusedGaussianConverter
    from avatar_pose.nonrigid import HashGridMLPFused
    from avatar_pose.skinning import SkinningFieldFused
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    P = 4000
    g = synth.canonical_gaussians(P, seed=3)
    gm = types.SimpleNamespace(active_sh_degree=3, max_sh_degree=3)
    for k in ("xyz", "rotation", "scaling", "opacity", "features_dc", "features_rest"):
        setattr(gm, "_" + k, torch.from_numpy(g[k]).to(dev).requires_grad_(True))
    lo, hi = (g["xyz"].min(0) - 0.05).tolist(), (g["xyz"].max(0) + 0.05).tolist()
    nr = HashGridMLPFused(n_cond=144, aabb_min=lo, aabb_max=hi, delay=10)
    sk = SkinningFieldFused(aabb_min=lo, aabb_max=hi)
    with torch.no_grad():
        nr.table.mul_(300.0)
        nr.lin3.weight.mul_(0.05)
        sk.lin4.weight.mul_(4.0)
    enc = lambda rots, Jtrs: torch.full((1, 144), 0.1, device=dev)
    conv = FusedGaussianConverter(sk, nr, enc).to(dev)
    ps, cam = synth.pose(2), synth.camera(W=128, H=128, fx=134.0, fy=134.0)
    camera = types.SimpleNamespace(bone_transforms=torch.from_numpy(ps["bone_transforms"]).to(dev),
                                   camera_center=torch.from_numpy(cam["camera_center"]).to(dev), rots=None, Jtrs=None)
    d = lambda a: torch.as_tensor(np.ascontiguousarray(a)).to(dev)
    rs = GaussianRasterizationSettings(128, 128, cam["tanfovx"], cam["tanfovy"], torch.zeros(3, device=dev), 1.0,
                                       d(cam["world_view_transform"]), d(cam["full_proj_transform"]), 3, d(cam["camera_center"]), False, False)

    def render(iteration):
        deformed, loss_reg, colors = conv(gm, camera, iteration)
        img, _ = GaussianRasterizer(rs)(means3D=deformed.get_xyz, means2D=torch.zeros_like(deformed.get_xyz, requires_grad=True),
                                        opacities=deformed.get_opacity, colors_precomp=colors,
                                        cov3D_precomp=deformed.get_covariance(1.0))
        return img, loss_reg

    with torch.no_grad():
        target = render(100)[0] * 0.8 + 0.05
        assert render(0)[1] == {}  # before `delay`: identity, no regulariser
    params = [p for p in conv.parameters()] + [getattr(gm, "_" + k) for k in ("xyz", "rotation", "scaling", "opacity", "features_dc", "features_rest")]
    img, loss_reg = render(100)
    assert set(loss_reg) == {"nr_xyz", "nr_scale", "nr_rot"}
    loss0 = ((img - target) ** 2).mean()
    (loss0 + 0.01 * sum(loss_reg.values())).backward()
    for name, p in list(conv.named_parameters()) + [("xyz", gm._xyz), ("rot", gm._rotation), ("scale", gm._scaling)]:
        assert p.grad is not None and torch.isfinite(p.grad).all() and float(p.grad.abs().max()) > 0, name
    with torch.no_grad():
        for p in params:
            p -= p.grad * (1e-4 / float(p.grad.abs().max()))
        loss1 = ((render(100)[0] - target) ** 2).mean()
    assert float(loss1) < float(loss0.detach()), (float(loss0.detach()), float(loss1))


def test_converter_default_config_mlp_texture_and_mask(built_lib):
    """The repository's default configuration end to end: non-rigid (hash grid + MLP with 16 extra feature outputs) ->
    skinning field -> fused posing -> colour-MLP texture (`mlp`, 32-d per-Gaussian features, use_sh false) -> ONE
    rasterisation returning colour + opacity (return_opacity) -> L1 + mask loss.  The colours equal the oracle's ColorMLP
    on the module's own intermediate tensors; every trainable tensor of the chain receives a finite, non-zero gradient."""
    from avatar_pose.modules import FusedGaussianConverter
    from avatar_pose.nonrigid import HashGridMLPFused
    from avatar_pose.skinning import SkinningFieldFused
    from avatar_pose.texture import ColorMLPFused
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    from oracle import color_oracle
    dev = torch.device("cuda:0")
    torch.manual_seed(1)
    P = 4000
    g = synth.canonical_gaussians(P, seed=3)
    gm = types.SimpleNamespace(active_sh_degree=0, max_sh_degree=0)
    for k in ("xyz", "rotation", "scaling", "opacity"):
        setattr(gm, "_" + k, torch.from_numpy(g[k]).to(dev).requires_grad_(True))
    gm._features_dc = (0.5 * torch.randn(P, 1, 1, device=dev)).requires_grad_(True)     # gaussian_model.py:182: [P, feature_dim, 1]
    gm._features_rest = (0.5 * torch.randn(P, 31, 1, device=dev)).requires_grad_(True)
    lo, hi = (g["xyz"].min(0) - 0.05).tolist(), (g["xyz"].max(0) + 0.05).tolist()
    nr = HashGridMLPFused(n_cond=144, feature_dim=16, aabb_min=lo, aabb_max=hi, delay=10)
    sk = SkinningFieldFused(aabb_min=lo, aabb_max=hi)
    tex = ColorMLPFused(n_frames=8)
    with torch.no_grad():
        nr.table.mul_(300.0)
        nr.lin3.weight.mul_(0.05)
        nr.lin3.weight[10:].mul_(100.0)  # visible non-rigid features
        sk.lin4.weight.mul_(4.0)
        for l in range(3):
            getattr(tex, f"lin{l}").weight.mul_(3.0)
    enc = lambda rots, Jtrs: torch.full((1, 144), 0.1, device=dev)
    conv = FusedGaussianConverter(sk, nr, enc, texture=tex).to(dev).eval()
    ps, cam = synth.pose(2), synth.camera(W=128, H=128, fx=134.0, fy=134.0)
    camera = types.SimpleNamespace(bone_transforms=torch.from_numpy(ps["bone_transforms"]).to(dev),
                                   camera_center=torch.from_numpy(cam["camera_center"]).to(dev), rots=None, Jtrs=None, latent_idx=5)
    d = lambda a: torch.as_tensor(np.ascontiguousarray(a)).to(dev)
    rs = GaussianRasterizationSettings(128, 128, cam["tanfovx"], cam["tanfovy"], torch.zeros(3, device=dev), 1.0,
                                       d(cam["world_view_transform"]), d(cam["full_proj_transform"]), 0, d(cam["camera_center"]), False, False)
    deformed, loss_reg, colors = conv(gm, camera, 100)
    assert colors.shape == (P, 3) and deformed.non_rigid_feature.shape == (P, 16)
    assert float(deformed.non_rigid_feature.abs().max()) > 1e-3
    # oracle ColorMLP on the chain's own intermediates (fp64)
    D = lambda t: t.detach().double().cpu()
    truth = color_oracle.color_forward(D(deformed.get_features).squeeze(-1), D(deformed.get_xyz), D(deformed.fwd_transform).reshape(P, 16),
                                       D(camera.camera_center), D(deformed.non_rigid_feature), D(tex.latent.weight[5]),
                                       [D(getattr(tex, f"lin{l}").weight) for l in range(3)],
                                       [D(getattr(tex, f"lin{l}").bias) for l in range(3)], 3, True, None)
    assert float((D(colors) - truth).abs().max()) <= 3e-5
    assert float(D(colors).std()) > 0.01
    screenspace = torch.zeros_like(deformed.get_xyz, requires_grad=True)
    img, radii, opim = GaussianRasterizer(rs)(means3D=deformed.get_xyz, means2D=screenspace, opacities=deformed.get_opacity,
                                              colors_precomp=colors, cov3D_precomp=deformed.get_covariance(1.0), return_opacity=True)
    target = torch.rand(3, 128, 128, device=dev)
    mask = (torch.rand(1, 128, 128, device=dev) > 0.5).float()
    loss = (img - target).abs().mean() + 0.1 * (opim - mask).abs().mean() + 0.01 * sum(loss_reg.values())
    loss.backward()
    named = list(conv.named_parameters()) + [(k, getattr(gm, "_" + k)) for k in ("xyz", "rotation", "scaling", "opacity", "features_dc", "features_rest")]
    for name, p in named:
        if name == "texture.latent.weight":
            assert p.grad is not None and float(p.grad[5].abs().max()) > 0 and float(p.grad[:5].abs().max()) == 0
            continue
        assert p.grad is not None and torch.isfinite(p.grad).all() and float(p.grad.abs().max()) > 0, name
    assert float(screenspace.grad.abs().max()) > 0


def test_smpl_nn_kernel_exact_nearest_vertex(built_lib):
    """smpl_nn_forward (rigid.py:41-47: knn_points K = 1 over the 6890 SMPL vertices) at the bench size: the chosen vertex is
    the fp64 argmin, except where two vertices are equidistant within FP32 rounding (then the chosen one is as near as the
    best within 1e-6 relative); duplicated vertices resolve to the lowest index; the gathered weights are that vertex's."""
    from avatar_pose.modules import SMPLNNFused
    dev = torch.device("cuda:0")
    P, V = 50000, 6890
    g = synth.canonical_gaussians(P, seed=0)
    rng = np.random.default_rng(9)
    verts = g["xyz"][rng.choice(P, V, replace=False)] + rng.normal(0, 0.004, size=(V, 3)).astype(np.float32)
    verts[4000] = verts[17]  # an exact duplicate: index 17 must win
    vw = synth.skinning_weights(verts)
    m = SMPLNNFused(verts, vw).to(dev)
    w, idx = m.query_weights(torch.from_numpy(g["xyz"]).to(dev), return_index=True)
    idx = idx.cpu().numpy().astype(np.int64)
    x64, v64 = g["xyz"].astype(np.float64), verts.astype(np.float64)
    best = np.empty(P, np.int64)
    for s in range(0, P, 5000):
        best[s:s + 5000] = ((x64[s:s + 5000, None, :] - v64[None]) ** 2).sum(-1).argmin(1)
    d_got = ((x64 - v64[idx]) ** 2).sum(1)
    d_best = ((x64 - v64[best]) ** 2).sum(1)
    differ = idx != best
    assert differ.mean() < 1e-3, f"{differ.sum()} indices differ from the fp64 argmin"
    assert np.all(d_got[differ] <= d_best[differ] * (1 + 1e-6) + 1e-14)
    assert not np.any(idx == 4000)
    assert torch.equal(w.cpu(), torch.from_numpy(vw)[torch.from_numpy(idx)])


@pytest.mark.parametrize("variant,mode", [("nr", "eager"), ("sk", "eager"), ("nr+sk+tex+enc", "eager"), ("nr+sk+tex+enc", "graph")])
def test_chain_fwd_bwd_at_bench_size(built_lib, variant, mode):
    """fwd+bwd of the converter chain + rasterizer at the bench size (50 000 Gaussians, 512x512), stage subsets included: the
    tcgen05 kernels run right before the blend kernels here and leave BF16 patterns in shared memory -- the backward blend once
    evaluated unstaged shared-memory slots for a strip that had run out of entries, and 0 * NaN reached its atomics (an illegal
    address at this size only).  Separate process: a device fault must not take the rest of the suite down.
    `graph`: the whole default-configuration step is captured in one CUDA graph (no host <-> device copy may occur inside it)."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "profiles", "default_chain_run.py"), "50000", "512", "3", mode, variant],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "CHAIN OK" in r.stdout and "finite=True" in r.stdout, (r.stdout[-500:], r.stderr[-1500:])
