"""CPU tests of the posing oracle: eval_sh pinned against the reference's own utils/sh_utils.py (importable on CPU in the
build container; skipped where /root/reference is absent), camera conventions pinned against utils/graphics_utils.py,
and closed-form checks of LBS / covariance / quaternion helpers."""
import importlib.util
import math
import os

import numpy as np
import pytest
import torch

import synth
from oracle import pose_oracle as po

REF = "/root/reference"


def _ref_module(rel, name):
    path = os.path.join(REF, rel)
    if not os.path.exists(path):
        pytest.skip("reference tree not present on this machine")
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.mark.parametrize("deg", [0, 1, 2, 3])
def test_eval_sh_equals_reference(deg):
    ref = _ref_module("utils/sh_utils.py", "ref_sh_utils")
    g = torch.Generator().manual_seed(deg)
    sh = torch.randn(500, 3, 16, generator=g)
    d = torch.nn.functional.normalize(torch.randn(500, 3, generator=g), dim=1)
    assert torch.equal(po.eval_sh(deg, sh, d), ref.eval_sh(deg, sh, d))


def test_camera_matches_reference_graphics_utils():
    ref = _ref_module("utils/graphics_utils.py", "ref_graphics_utils")
    cam = synth.camera(W=512, H=384, fx=537.0, fy=500.0, azimuth_deg=33.0)
    P = ref.getProjectionMatrix(znear=0.01, zfar=100.0, fovX=cam["FoVx"], fovY=cam["FoVy"]).transpose(0, 1)
    wvt = torch.from_numpy(cam["world_view_transform"])
    full = wvt.unsqueeze(0).bmm(P.unsqueeze(0)).squeeze(0)  # scene/cameras.py:37-39
    assert torch.allclose(full, torch.from_numpy(cam["full_proj_transform"]), atol=1e-6)
    assert abs(ref.focal2fov(537.0, 512) - cam["FoVx"]) < 1e-12
    # the pelvis projects to the image centre, in front of the camera
    J, _ = synth.skeleton()
    p = np.append(J[0], 1.0).astype(np.float32) @ cam["full_proj_transform"]
    assert abs(p[0] / p[3]) < 1e-4 and abs(p[1] / p[3]) < 1e-4 and p[3] > 3.0


def test_build_rotation_orthonormal_and_quat_product():
    g = torch.Generator().manual_seed(0)
    q1, q2 = torch.randn(64, 4, generator=g), torch.randn(64, 4, generator=g)
    R1, R2 = po.build_rotation(q1), po.build_rotation(q2)
    I = torch.eye(3).expand(64, 3, 3)
    assert torch.allclose(R1 @ R1.transpose(1, 2), I, atol=1e-5)
    assert torch.allclose(po.build_rotation(po.quaternion_multiply(q1, q2)), R1 @ R2, atol=1e-5)
    assert torch.allclose(po.build_rotation(po.quaternion_multiply_general_utils(q1, q2)), R2 @ R1, atol=1e-5)


def test_lbs_identity_and_rigid_motion():
    g = synth.canonical_gaussians(500, seed=1)
    w = torch.from_numpy(synth.skinning_weights(g["xyz"]))
    assert torch.allclose(w.sum(1), torch.ones(500), atol=1e-6) and (w >= 0).all() and ((w > 0).sum(1) <= 4).all()
    xyz, rot = torch.from_numpy(g["xyz"]), torch.from_numpy(g["rotation"])
    eye = torch.eye(4).repeat(24, 1, 1)
    T, xb, Rb = po.lbs(xyz, rot, w, eye)
    assert torch.allclose(xb, xyz, atol=1e-6) and torch.allclose(Rb, po.build_rotation(rot), atol=1e-6)
    # one global rigid motion applied to every bone moves every Gaussian rigidly, whatever the weights
    a = 0.7
    M = torch.tensor([[math.cos(a), -math.sin(a), 0, 0.3], [math.sin(a), math.cos(a), 0, -0.2], [0, 0, 1, 0.1], [0, 0, 0, 1.0]])
    T, xb, Rb = po.lbs(xyz, rot, w, M.repeat(24, 1, 1))
    assert torch.allclose(xb, xyz @ M[:3, :3].T + M[:3, 3], atol=1e-5)
    cov = po.covariance(Rb, torch.from_numpy(g["scaling"]))
    cov0 = po.covariance(po.build_rotation(rot), torch.from_numpy(g["scaling"]))
    full = lambda c: torch.stack([c[:, 0], c[:, 1], c[:, 2], c[:, 1], c[:, 3], c[:, 4], c[:, 2], c[:, 4], c[:, 5]], 1).view(-1, 3, 3)
    assert torch.allclose(full(cov), M[:3, :3] @ full(cov0) @ M[:3, :3].T, atol=1e-6, rtol=1e-4)


def test_bone_transforms_rest_pose_is_identity():
    B = synth.bone_transforms_from_rotmats(np.stack([np.eye(3)] * 24))
    assert np.allclose(B, np.eye(4)[None].repeat(24, 0), atol=1e-6)
    seq = synth.pose_sequence(n_frames=8)
    assert len(seq) == 8 and seq[0]["bone_transforms"].shape == (24, 4, 4)


def test_pose_chain_matches_reference_code():
    """The posing oracle against outputs of the reference's OWN code (SMPLNN.forward, build_rotation,
    build_covariance_from_scaling_rotation, SH2RGB.forward -- lifted out of their files with `ast` and executed unmodified:
    tests/golden/make_pose_chain_golden.py -> tests/golden/ref_pose_chain.npz).  fp64 oracle vs the fp32 reference run:
    <= 2e-6 relative (fp32 rounding of the reference's matmuls); fp32 oracle: same bound."""
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_pose_chain.npz"))
    for dt in (torch.float64, torch.float32):
        T = lambda k: torch.from_numpy(z[k]).to(dt)
        feat = torch.cat([T("features_dc"), T("features_rest")], 1)
        o = po.pose_forward(T("xyz"), T("rotation"), T("scaling"), T("opacity"), feat, T("weights"), T("bone_transforms"),
                            T("camera_center"), 3, 1.0, True)
        for k, tol in (("means3D", 2e-6), ("T_fwd", 2e-6), ("rotation_bar", 2e-6), ("cov3D", 5e-6), ("colors", 5e-6)):
            ref = torch.from_numpy(z[k]).double()
            err = float((o[k].double().reshape(ref.shape) - ref).abs().max() / ref.abs().max())
            assert err <= tol, (str(dt), k, err)
        o07 = po.pose_forward(T("xyz"), T("rotation"), T("scaling"), T("opacity"), feat, T("weights"), T("bone_transforms"),
                              T("camera_center"), 3, 0.7, True)
        ref = torch.from_numpy(z["cov3D_mod07"]).double()
        assert float((o07["cov3D"].double() - ref).abs().max() / ref.abs().max()) <= 5e-6
