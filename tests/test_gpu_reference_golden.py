"""The CUDA path against outputs of the REFERENCE'S OWN CODE (tests/golden/ref_*.npz: reference functions / classes lifted out
of their files with `ast` and executed unmodified on the CPU by tests/golden/make_*_golden.py) on the same inputs -- no
oracle in between.  Tolerances: fp32 elementwise stages (posing) 2e-5 relative; tensor-core MLP stages with the 2-way BF16
split: 3e-5 abs on [0,1]-valued outputs (skinning weights, colours after the sigmoid)."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _rel(a, b):
    b = torch.as_tensor(b).double()
    return float((a.detach().double().cpu().reshape(b.shape) - b).abs().max() / b.abs().max())


def test_fused_pose_matches_reference_chain(built_lib):
    """pose_fwd_kernel vs SMPLNN.forward + build_covariance_from_scaling_rotation + SH2RGB.forward (ref_pose_chain.npz)."""
    from avatar_pose import fused_pose
    dev = torch.device("cuda:0")
    z = np.load(os.path.join(GOLD, "ref_pose_chain.npz"))
    t = lambda k: torch.from_numpy(z[k]).to(dev)
    means, cov, col, op, T_fwd, Rbar = fused_pose(t("xyz"), t("rotation"), t("scaling"), t("opacity"), t("features_dc"), t("features_rest"),
                                                  t("weights"), t("bone_transforms"), t("camera_center"), 3, 1.0, True, return_aux=True)
    assert _rel(means, z["means3D"]) <= 2e-5
    assert _rel(T_fwd, z["T_fwd"]) <= 2e-5
    assert _rel(Rbar, z["rotation_bar"]) <= 2e-5
    assert _rel(cov, z["cov3D"]) <= 2e-5
    assert _rel(col, z["colors"]) <= 2e-5
    cov07 = fused_pose(t("xyz"), t("rotation"), t("scaling"), t("opacity"), t("features_dc"), t("features_rest"), t("weights"),
                       t("bone_transforms"), t("camera_center"), 3, 0.7, True)[1]
    assert _rel(cov07, z["cov3D_mod07"]) <= 2e-5


def test_skinning_field_matches_reference_code(built_lib):
    """sk_fwd_kernel (+ pose_fwd_kernel behind it) vs the reference's SkinningField path (ref_skinning.npz)."""
    from avatar_pose import fused_pose
    from avatar_pose.skinning import SkinningFieldFused
    dev = torch.device("cuda:0")
    z = np.load(os.path.join(GOLD, "ref_skinning.npz"))
    t = lambda k: torch.from_numpy(z[k]).to(dev)
    m = SkinningFieldFused(d_out=25, aabb_min=z["aabb_min"].tolist(), aabb_max=z["aabb_max"].tolist())
    with torch.no_grad():
        for l in range(5):
            getattr(m, f"lin{l}").weight.copy_(torch.from_numpy(z[f"W{l}"]))
            getattr(m, f"lin{l}").bias.copy_(torch.from_numpy(z[f"b{l}"]))
    m = m.to(dev)
    w, logits = m.query_weights(t("xyz"), return_logits=True)
    scale = float(np.abs(z["logits"]).max())
    assert float((logits.cpu() - torch.from_numpy(z["logits"])).abs().max()) <= 3e-5 * scale + 1e-6
    assert float((w.cpu() - torch.from_numpy(z["weights"])).abs().max()) <= 3e-5
    P = z["xyz"].shape[0]
    zeros = lambda *s: torch.zeros(*s, device=dev)
    means, _, _, _, T_fwd, Rbar = fused_pose(t("xyz"), t("rotation"), zeros(P, 3), zeros(P, 1), zeros(P, 1, 3), zeros(P, 15, 3), w,
                                             t("bone_transforms"), zeros(3), 3, 1.0, True, return_aux=True)
    assert _rel(means, z["means3D"]) <= 1e-4 and _rel(T_fwd, z["T_fwd"]) <= 1e-4 and _rel(Rbar, z["rotation_bar"]) <= 1e-4


def test_color_mlp_matches_reference_class(built_lib):
    """color_fwd_kernel vs the reference's ColorMLP + VanillaCondMLP (ref_color_mlp.npz)."""
    from avatar_pose.texture import ColorMLPFused
    dev = torch.device("cuda:0")
    z = np.load(os.path.join(GOLD, "ref_color_mlp.npz"))
    t = lambda k: torch.from_numpy(z[k]).to(dev)
    m = ColorMLPFused()
    with torch.no_grad():
        for l in range(3):
            getattr(m, f"lin{l}").weight.copy_(torch.from_numpy(z[f"W{l}"]))
            getattr(m, f"lin{l}").bias.copy_(torch.from_numpy(z[f"b{l}"]))
    m = m.to(dev)
    col, inp = m.colors(t("features"), t("xyz"), t("T_fwd"), t("camera_center"), t("non_rigid_feature"), t("latent"), None, return_inputs=True)
    assert float((inp.cpu() - torch.from_numpy(z["inputs"][:, :63])).abs().max()) <= 2e-6
    assert float((col.cpu() - torch.from_numpy(z["colors"])).abs().max()) <= 3e-5
