"""CPU tests of the colour-MLP oracle (oracle/color_oracle.py): `sh_bases` pinned against the committed outputs of the
reference's own eval_sh_bases (tests/golden/ref_sh_bases.npz, made by tests/golden/make_sh_bases_golden.py) and, where the
reference tree is present, against the live function; the composition against hand-built cases; and the two identities
the CUDA kernel relies on (latent code folded into the first bias, zero padding of the input row)."""
import importlib.util
import os

import numpy as np
import pytest
import torch

from oracle import color_oracle as co

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.parametrize("deg", [0, 1, 2, 3])
def test_sh_bases_golden(deg):
    z = np.load(os.path.join(HERE, "golden", "ref_sh_bases.npz"))
    got = co.sh_bases(deg, torch.from_numpy(z["dirs"])).numpy()
    assert np.array_equal(got.view(np.uint32), z[f"bases_deg{deg}"].view(np.uint32))  # same expressions: bit-exact


def test_sh_bases_live_reference():
    path = "/root/reference/utils/sh_utils.py"
    if not os.path.exists(path):
        pytest.skip("reference tree not present on this machine")
    spec = importlib.util.spec_from_file_location("ref_sh_utils", path)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    d = torch.nn.functional.normalize(torch.randn(500, 3, generator=torch.Generator().manual_seed(3)))
    assert torch.equal(ref.eval_sh_bases(3, d), co.sh_bases(3, d))


def test_compose_input_layout_and_view_dir():
    c = co.synthetic_case(64, seed=1)
    x = co.compose_input(c["features"], c["xyz"], c["T_fwd"], c["camera_center"], c["non_rigid_feature"], c["latent"], 3, True)
    assert x.shape == (64, 32 + 15 + 16 + 16)
    assert torch.equal(x[:, :32], c["features"]) and torch.equal(x[:, 47:63], c["non_rigid_feature"])
    assert torch.equal(x[:, 63:], c["latent"].reshape(1, -1).expand(64, -1))
    # canonical view direction of Gaussian 0 by hand: R^T (xyz - cam), normalised, SH basis without the DC term
    R = c["T_fwd"][0].reshape(4, 4)[:3, :3]
    d = R.t() @ (c["xyz"][0] - c["camera_center"])
    d = d / (d.norm() + 1e-12)
    assert torch.allclose(x[0, 32:47], co.sh_bases(3, d)[1:], atol=1e-6)
    # world-space direction when cano_view_dir is off: independent of T_fwd
    y = co.compose_input(c["features"], c["xyz"], None, c["camera_center"], c["non_rigid_feature"], c["latent"], 3, False)
    d = c["xyz"][0] - c["camera_center"]
    assert torch.allclose(y[0, 32:47], co.sh_bases(3, d / (d.norm() + 1e-12))[1:], atol=1e-6)


def test_latent_fold_and_padding_identities():
    """What gs_color.cu does to the first layer: W0 . [x | latent] + b0 == W0[:, :63] . x + (b0 + W0[:, 63:] . latent), and
    zero-padded K columns contribute nothing."""
    c = co.synthetic_case(200, seed=2, dtype=torch.float64)
    x = co.compose_input(c["features"], c["xyz"], c["T_fwd"], c["camera_center"], c["non_rigid_feature"], c["latent"], 3, True)
    W0, b0 = c["Ws"][0], c["bs"][0]
    full = x @ W0.t() + b0
    xp = torch.cat([x[:, :63], torch.zeros(200, 1, dtype=torch.float64)], 1)
    W0p = torch.cat([W0[:, :63], torch.zeros(64, 1, dtype=torch.float64)], 1)
    folded = xp @ W0p.t() + (b0 + W0[:, 63:] @ c["latent"])
    assert torch.allclose(full, folded, atol=1e-12)


def test_color_range_and_dtype_agreement():
    c32, c64 = co.synthetic_case(500, seed=3), co.synthetic_case(500, seed=3, dtype=torch.float64)
    args = lambda c: (c["features"], c["xyz"], c["T_fwd"], c["camera_center"], c["non_rigid_feature"], c["latent"], c["Ws"], c["bs"], 3, True, c["view_noise"])
    a, b = co.color_forward(*args(c32)), co.color_forward(*args(c64))
    assert a.shape == (500, 3) and float(a.min()) > 0 and float(a.max()) < 1
    assert float((a.double() - b).abs().max()) < 5e-6


def test_color_oracle_matches_reference_colormlp():
    """The whole colour-MLP oracle against the reference's OWN ColorMLP + VanillaCondMLP classes (lifted out of their files
    with `ast` and executed unmodified: tests/golden/make_color_mlp_golden.py -> tests/golden/ref_color_mlp.npz)."""
    z = np.load(os.path.join(HERE, "golden", "ref_color_mlp.npz"))
    T = lambda k: torch.from_numpy(z[k])
    Ws, bs = [T(f"W{l}") for l in range(3)], [T(f"b{l}") for l in range(3)]
    x = co.compose_input(T("features"), T("xyz"), T("T_fwd"), T("camera_center"), T("non_rigid_feature"), T("latent"), 3, True)
    assert x.shape == (400, 79)
    assert float((x - T("inputs")).abs().max()) <= 1e-6
    col = co.color_forward(T("features"), T("xyz"), T("T_fwd"), T("camera_center"), T("non_rigid_feature"), T("latent"), Ws, bs, 3, True)
    assert float((col - T("colors")).abs().max()) <= 1e-6
    assert float(T("colors").std()) > 0.05
