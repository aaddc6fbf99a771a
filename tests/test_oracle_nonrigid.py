"""CPU test: the non-rigid oracle (oracle/nonrigid_oracle.py: conditional MLP + delta application) and the pose-encoder
oracle, chained, against outputs of the reference's OWN HashGridwithMLP.forward run with its own HierarchicalPoseEncoder,
VanillaCondMLP and AABB (tests/golden/make_nonrigid_golden.py -> tests/golden/ref_nonrigid.npz).  What stays outside the pin:
the hash-grid encoding itself (tiny-cuda-nn, not installed: the fixture feeds a seeded feature matrix in its place) and the
sign convention of pytorch3d's quaternion_multiply, which standardises the product to a non-negative real part -- the oracle
and the kernel return the raw Hamilton product, the same rotation (build_rotation is even in the quaternion)."""
import os

import numpy as np
import torch

from oracle import encoder_oracle as eo
from oracle import nonrigid_oracle as no

HERE = os.path.dirname(os.path.abspath(__file__))


def test_nonrigid_oracle_matches_reference_forward():
    z = np.load(os.path.join(HERE, "golden", "ref_nonrigid.npz"))
    T = lambda k: torch.from_numpy(z[k])
    cond = eo.encode(T("rots")[0], T("Jtrs")[0], T("enc_W0"), T("enc_b0"), T("enc_Wa"), T("enc_ba"), T("enc_Wb"), T("enc_bb"))
    assert float((cond - T("cond")[0]).abs().max()) <= 1e-6
    W, b = [T(f"W{l}") for l in range(4)], [T(f"b{l}") for l in range(4)]
    o = no.nonrigid_forward(T("xyz"), T("scaling"), T("rotation"), T("aabb_min"), T("aabb_max"), None, None, cond, W, b,
                            features=T("hash_feat"))
    assert float((o["xyz"] - T("xyz_out")).abs().max()) <= 1e-6
    assert float((o["scaling"] - T("scaling_out")).abs().max()) <= 1e-6
    assert float((o["deltas"][:, 10:] - T("non_rigid_feature")).abs().max()) <= 1e-6
    rot = o["rotation"]
    std = torch.where(rot[:, :1] < 0, -rot, rot)  # pytorch3d standardises; same rotation
    assert float((std - T("rotation_out")).abs().max()) <= 2e-6
    assert bool((rot[:, 0] < 0).any()), "the fixture should exercise the sign convention"
    d = o["deltas"]
    loss = torch.stack([d[:, :3].norm(p=2, dim=1).mean(), d[:, 3:6].norm(p=1, dim=1).mean(), d[:, 7:10].norm(p=1, dim=1).mean()])
    assert float((loss - T("loss")).abs().max()) <= 1e-6  # non_rigid.py:288-297 (what FusedGaussianConverter returns as loss_reg)


def test_skinning_oracle_matches_reference_code():
    """Skinning-field oracle (MLP -> hierarchical softmax / softmax -> LBS weights) and the posing oracle's LBS against the
    reference's OWN hierarchical_softmax, SkinningField.softmax / get_forward_transform / forward, VanillaCondMLP, AABB and
    build_rotation (tests/golden/make_skinning_golden.py -> tests/golden/ref_skinning.npz)."""
    from oracle import pose_oracle as po
    z = np.load(os.path.join(HERE, "golden", "ref_skinning.npz"))
    T = lambda k: torch.from_numpy(z[k])
    W, b = [T(f"W{l}") for l in range(5)], [T(f"b{l}") for l in range(5)]
    w, logits = no.skinning_weights_forward(T("xyz"), T("aabb_min"), T("aabb_max"), W, b)
    assert float((logits - T("logits")).abs().max()) <= 2e-5 * float(T("logits").abs().max())
    assert float((w - T("weights")).abs().max()) <= 2e-6
    assert float((w.sum(1) - 1).abs().max()) <= 1e-5 and float(T("weights").max()) > 0.5
    w24 = torch.softmax(T("logits")[:, :24], -1)
    assert float((w24 - T("weights_softmax24")).abs().max()) <= 1e-6
    T_fwd, x_bar, R_bar = po.lbs(T("xyz"), T("rotation"), T("weights"), T("bone_transforms"))
    assert float((T_fwd.reshape(-1, 4, 4) - T("T_fwd")).abs().max()) <= 2e-6
    assert float((x_bar - T("means3D")).abs().max()) <= 2e-6
    assert float((R_bar - T("rotation_bar")).abs().max()) <= 2e-6
