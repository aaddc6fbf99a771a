"""A/B against the REFERENCE'S OWN CUDA rasterizer (the unmodified diff-gaussian-rasterization sources rebuilt for sm_100
into the git-ignored baseline/_ref/, which travels to the GPU box with the snapshot) on IDENTICAL posed inputs, at the
bench size: BASELINE.json configs[1]/[2] = 50 000 Gaussians, 512x512 -- random-init and a trained-like opaque scene.

north_star's contract, checked against the reference itself (not against this repository's oracle):
  bit-exact      sorted 64-bit keys, point list, tile ranges, per-pixel contributor counts (n_contrib), radii
  <= 1e-5 abs    image, final_T
  <= 1e-4 rel    every gradient (the reference's atomics are order-nondeterministic)
The reference's private byte buffers are decoded with tests/golden/make_ref_cuda_golden.py:decode.  Mismatch COUNTS are
printed and written to gpurun_out/ref_parity_<scene>.json.  The default blend uses CUDA's expf -- the instruction sequence
nvcc emits for the reference's `exp(power)` (cuda_rasterizer/forward.cu:335-351) -- so that the skip / termination decisions
are the reference's, bit for bit."""
import importlib.util
import json
import os
import sys

import numpy as np
import pytest
import torch

import scenes
from test_gpu_rasterizer import run_cuda, rel_err

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = os.path.join(ROOT, "baseline", "_ref", "diff_gaussian_rasterization")


def _gen():
    spec = importlib.util.spec_from_file_location("make_ref_cuda_golden", os.path.join(ROOT, "tests", "golden", "make_ref_cuda_golden.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture(scope="module")
def ref_mod():
    if not (os.path.isdir(REF_DIR) and any(f.endswith(".so") for f in os.listdir(REF_DIR))):
        pytest.skip("baseline/_ref (the reference rasterizer rebuilt for sm_100) is absent")
    gen = _gen()
    return gen, gen.load_ref()


def trained_like_scene(P=50000, W=512, H=512):
    """Opaque surface Gaussians (the state after training): early termination everywhere, T crossing 1e-4."""
    return scenes.avatar_scene(P=P, W=W, H=H, random_init=False)


SCENES = {"cfgA_random_init_50k_512": lambda: scenes.avatar_scene(P=50000, W=512, H=512, random_init=True),
          "cfgA_trained_like_50k_512": trained_like_scene,
          "sh_scale_rot_20k_400x304": lambda: scenes.generic_scene(P=20000, W=400, H=304, seed=11)}


@pytest.mark.parametrize("name", list(SCENES))
def test_against_reference_cuda_build(built_lib, ref_mod, name):
    gen, ref = ref_mod
    sc = SCENES[name]()
    deg, mod = 3, 1.0
    r = gen.run(ref, sc, deg, mod)                 # the reference: forward, decode of its buffers, backward with the L1 gradient
    ours = run_cuda(sc, sh_degree=deg, scale_modifier=mod, dL=r["dL_dpix"])   # this library, the SAME tensors and upstream gradient
    H, W = sc["H"], sc["W"]
    dbg = ours["dbg"]
    N = int(r["N"])
    rep = {"scene": name, "P": int(sc["means3D"].shape[0]), "W": W, "H": H, "N_ref": N, "N_ours": int(ours["N"])}
    rep["radii_mismatch"] = int((ours["radii"] != r["radii"]).sum())
    if ours["N"] == N:
        rep["keys_mismatch"] = int((dbg["keys"].view(np.uint64) != r["keys"]).sum())
        rep["point_list_mismatch"] = int((dbg["point_list"].view(np.uint32) != r["point_list"]).sum())
    else:
        rep["keys_mismatch"] = rep["point_list_mismatch"] = -1
    rep["ranges_mismatch"] = int((dbg["ranges"].view(np.uint32).reshape(-1, 2) != r["ranges"]).any(1).sum())
    nc_o, nc_r = dbg["n_contrib"].view(np.uint32).reshape(H, W), r["n_contrib"].reshape(H, W)
    rep["n_contrib_mismatch_pixels"] = int((nc_o != nc_r).sum())
    rep["E_fwd"] = int(nc_r.astype(np.int64).sum())
    rep["final_T_max_abs"] = float(np.abs(dbg["final_T"].reshape(H, W) - r["final_T"].reshape(H, W)).max())
    rep["final_T_bit_mismatch"] = int((dbg["final_T"].view(np.uint32).reshape(-1) != r["final_T"].view(np.uint32).reshape(-1)).sum())
    rep["image_max_abs"] = float(np.abs(ours["img"] - r["color"]).max())
    rep["image_bit_mismatch"] = int((ours["img"].view(np.uint32) != r["color"].view(np.uint32)).sum())
    g = ours["grads"]
    pairs = {"means3D": "grad_means3D", "means2D": "grad_means2D", "opacities": "grad_opacities", "colors_precomp": "grad_colors_precomp",
             "cov3D_precomp": "grad_cov3D_precomp", "shs": "grad_shs", "scales": "grad_scales", "rotations": "grad_rotations"}
    rep["grad_rel"] = {k: rel_err(g[k], r[v]) for k, v in pairs.items() if g.get(k) is not None and v in r}
    print("\nREF-PARITY " + json.dumps(rep))
    out_dir = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out_dir, exist_ok=True)
    with open(os.path.join(out_dir, f"ref_parity_{name}.json"), "w") as f:
        json.dump(rep, f, indent=1)
    assert rep["N_ours"] == rep["N_ref"]
    assert rep["radii_mismatch"] == 0
    assert rep["keys_mismatch"] == 0, "sorted keys differ from the reference build"
    assert rep["point_list_mismatch"] == 0
    assert rep["ranges_mismatch"] == 0
    assert rep["n_contrib_mismatch_pixels"] == 0, "per-pixel contributor counts differ from the reference build"
    assert rep["image_max_abs"] <= 1e-5 and rep["final_T_max_abs"] <= 1e-5
    for k, e in rep["grad_rel"].items():
        assert e <= 1e-4, (k, e)
