"""GPU parity of the non-rigid deformation forward (hash grid + tensor-core MLP, nonrigid_forward through
avatar_pose.nonrigid.HashGridMLPFused) against the torch oracle in fp64.  Tolerances: hash-grid features <= 3e-4 of
max|table| (the finest level has scale 2047: an FP32 position carries ~1e-4 cells of rounding, which the fp64 oracle does
not have); MLP deltas, evaluated by the oracle on the SAME features, <= 2e-5 of max|delta| + 1e-6 (2-way BF16 split on
the tensor cores, FP32 accumulate); end to end <= 1e-3 of max|delta|."""
import numpy as np
import pytest
import torch

import synth
from oracle import nonrigid_oracle as nro

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("P,feature_dim", [(5000, 0), (777, 16), (128, 0)])
def test_nonrigid_forward(built_lib, P, feature_dim):
    from avatar_pose.nonrigid import HashGridMLPFused
    dev = torch.device("cuda:0")
    torch.manual_seed(3)
    g = synth.canonical_gaussians(P, seed=0)
    xyz, sc, rot = (torch.from_numpy(g[k]) for k in ("xyz", "scaling", "rotation"))
    lo, hi = xyz.min(0).values - 0.05, xyz.max(0).values + 0.05
    m = HashGridMLPFused(n_cond=144, feature_dim=feature_dim, aabb_min=lo.tolist(), aabb_max=hi.tolist())
    with torch.no_grad():  # a trained-like state: table and last layer large enough to matter
        m.table.mul_(3000.0)
        m.lin3.weight.mul_(0.3)
    cond = torch.randn(144) * 0.5
    md = m.to(dev)
    oxyz, osc, orot, deltas, feats = md.deform(xyz.to(dev), sc.to(dev), rot.to(dev), cond.to(dev), return_features=True)
    torch.cuda.synchronize()
    d64 = lambda t: t.detach().cpu().double()
    W = [d64(l.weight) for l in (m.lin0, m.lin1, m.lin2, m.lin3)]
    b = [d64(l.bias) for l in (m.lin0, m.lin1, m.lin2, m.lin3)]
    ref = nro.nonrigid_forward(xyz.double(), sc.double(), rot.double(), d64(m.aabb_min), d64(m.aabb_max), d64(m.table),
                               m.levels, cond.double(), W, b)
    assert float((feats.cpu().double() - ref["features"]).abs().max()) <= 3e-4 * float(m.table.detach().abs().max())
    scale = float(ref["deltas"].abs().max())
    assert scale > 1e-3
    assert float((deltas.cpu().double() - ref["deltas"]).abs().max()) <= 1e-3 * scale  # end to end, incl. FP32 encoding
    ref = nro.nonrigid_forward(xyz.double(), sc.double(), rot.double(), d64(m.aabb_min), d64(m.aabb_max), d64(m.table),
                               m.levels, cond.double(), W, b, features=feats.cpu().double())  # the MLP on its own
    err = float((deltas.cpu().double() - ref["deltas"]).abs().max())
    assert err <= 2e-5 * scale + 1e-6, (err, scale)
    assert float((oxyz.cpu().double() - ref["xyz"]).abs().max()) <= 2e-6 + 2e-5 * scale
    assert float((osc.cpu().double() - ref["scaling"]).abs().max()) <= 2e-6 + 2e-5 * scale
    assert float((orot.cpu().double() - ref["rotation"]).abs().max()) <= 2e-6 + 2e-5 * scale


def test_nonrigid_delay_is_identity(built_lib):
    from avatar_pose.nonrigid import HashGridMLPFused
    dev = torch.device("cuda:0")
    m = HashGridMLPFused(delay=3000).to(dev)
    x = torch.randn(10, 3, device=dev)
    out = m.deform(x, x, torch.randn(10, 4, device=dev), torch.zeros(144, device=dev), iteration=100)
    assert out[0] is x and out[3] is None


@pytest.mark.parametrize("P,d_out", [(6000, 25), (300, 24)])
def test_skinning_field_weights(built_lib, P, d_out):
    """Skinning MLP + hierarchical softmax on the tensor cores vs the fp64 oracle: logits <= 2e-5 of max|logit| + 1e-6,
    weights <= 2e-5 abs, rows sum to 1; then the weights drive the fused posing kernel."""
    from avatar_pose.skinning import SkinningFieldFused
    from avatar_pose import fused_pose
    dev = torch.device("cuda:0")
    torch.manual_seed(5)
    g = synth.canonical_gaussians(P, seed=1)
    xyz = torch.from_numpy(g["xyz"])
    lo, hi = xyz.min(0).values - 0.05, xyz.max(0).values + 0.05
    m = SkinningFieldFused(d_out=d_out, aabb_min=lo.tolist(), aabb_max=hi.tolist())
    with torch.no_grad():
        m.lin4.weight.mul_(8.0)  # spread the logits as a trained field would
    md = m.to(dev)
    wts, logits = md.query_weights(xyz.to(dev), return_logits=True)
    torch.cuda.synchronize()
    d64 = lambda t: t.detach().cpu().double()
    W = [d64(getattr(m, f"lin{l}").weight) for l in range(5)]
    b = [d64(getattr(m, f"lin{l}").bias) for l in range(5)]
    ref_w, ref_logits = nro.skinning_weights_forward(xyz.double(), d64(m.aabb_min), d64(m.aabb_max), W, b)
    scale = float(ref_logits.abs().max())
    assert scale > 0.5
    assert float((logits.cpu().double() - ref_logits).abs().max()) <= 2e-5 * scale + 1e-6
    assert float((wts.cpu().double() - ref_w).abs().max()) <= 2e-5
    assert float((wts.sum(1) - 1).abs().max()) <= 1e-5
    ps = synth.pose(1)
    t = lambda k: torch.from_numpy(g[k]).to(dev)
    means = fused_pose(t("xyz"), t("rotation"), t("scaling"), t("opacity"), t("features_dc"), t("features_rest"), wts,
                       torch.from_numpy(ps["bone_transforms"]).to(dev), torch.zeros(3, device=dev), 3)[0]
    assert torch.isfinite(means).all()
