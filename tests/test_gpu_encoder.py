"""GPU parity of the one-launch hierarchical pose encoder (gs_encoder.cu) against the golden outputs of the reference's own
class (tests/golden/ref_pose_encoder.npz) and the oracle; <= 2e-6 abs (FP32 dot products of length <= 288 in a different
summation order).  Gradients of the one-launch backward kernel against fp64 autograd of the oracle."""
import os

import numpy as np
import pytest
import torch

from oracle import encoder_oracle as eo

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _load(tag, dev):
    from avatar_pose.pose_encoder import HierarchicalPoseEncoderFused
    z = np.load(os.path.join(HERE, "golden", "ref_pose_encoder.npz"))
    T = lambda k: torch.from_numpy(z[f"{tag}_{k}"])
    m = HierarchicalPoseEncoderFused(rel_joints=(tag == "rel"))
    sd = {"layer_0.weight": T("W0"), "layer_0.bias": T("b0")}
    for j in range(24):
        sd[f"layers.{j}.0.weight"], sd[f"layers.{j}.0.bias"] = T("Wa")[j], T("ba")[j]
        sd[f"layers.{j}.2.weight"], sd[f"layers.{j}.2.bias"] = T("Wb")[j], T("bb")[j]
    m.load_reference_state_dict(sd)
    return m.to(dev), T


@pytest.mark.parametrize("tag", ["abs", "rel"])
def test_pose_encoder_matches_reference_golden(built_lib, tag):
    dev = torch.device("cuda:0")
    m, T = _load(tag, dev)
    with torch.no_grad():
        y = m(T("rots").to(dev), T("Jtrs").to(dev))
    assert y.shape == (1, 144)
    assert float((y.cpu() - T("out")).abs().max()) <= 2e-6 * max(1.0, float(T("out").abs().max()))


@pytest.mark.parametrize("tag", ["abs", "rel"])
def test_pose_encoder_gradients(built_lib, tag):
    """pose_encoder_backward (one launch) against fp64 autograd of the oracle, <= 1e-4 of each gradient's max; in `rel` mode the
    reference forms the relative joints under no_grad, so Jtrs receives no gradient."""
    dev = torch.device("cuda:0")
    m, T = _load(tag, dev)
    rots, Jtrs = T("rots").to(dev).requires_grad_(True), T("Jtrs").to(dev).requires_grad_(True)
    y = m(rots, Jtrs)
    g = torch.randn(1, 144, generator=torch.Generator().manual_seed(2))
    y.backward(g.to(dev))
    D = lambda k: T(k).double().requires_grad_(True)
    leaves = {k: D(k) for k in ("rots", "Jtrs", "W0", "b0", "Wa", "ba", "Wb", "bb")}
    ref = eo.encode(leaves["rots"][0], leaves["Jtrs"][0], leaves["W0"], leaves["b0"], leaves["Wa"], leaves["ba"], leaves["Wb"], leaves["bb"],
                    rel_joints=(tag == "rel"))
    ref.backward(g.double()[0])
    rel = lambda a, b: float((a.double().cpu() - b).abs().max() / b.abs().max().clamp_min(1e-30))
    assert rel(rots.grad, leaves["rots"].grad) <= 1e-4
    if tag == "rel":
        assert Jtrs.grad is None and leaves["Jtrs"].grad is None
    else:
        assert rel(Jtrs.grad, leaves["Jtrs"].grad) <= 1e-4
    for k, p in (("W0", m.W0), ("b0", m.b0), ("Wa", m.Wa), ("ba", m.ba), ("Wb", m.Wb), ("bb", m.bb)):
        assert rel(p.grad, leaves[k].grad) <= 1e-4, k
