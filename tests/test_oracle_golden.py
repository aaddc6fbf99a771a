"""CPU test: pins the oracle (oracle/gs_oracle.c) against outputs of the REFERENCE's own CUDA rasterizer (sm_100
rebuild run on a B200; fixtures tests/golden/ref_cuda_*.npz, generator tests/golden/make_ref_cuda_golden.py).

Bit-exact: radii, tiles_touched, depth bits, pixel means, conic, sorted keys, point list, tile ranges, n_contrib.
Image / final_T <= 1e-5 (north_star; the only arithmetic difference is exp(): the oracle's FMA polynomial vs CUDA's
MUFU-based expf).  Gradients <= 1e-4 relative."""
import os

import numpy as np
import pytest

import scenes
from oracle import gs_oracle

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = {"avatar_3000_128": (lambda: scenes.avatar_scene(P=3000, W=128, H=128), 3, 1.0),
         "avatar_1500_200x120": (lambda: scenes.avatar_scene(P=1500, W=200, H=120), 3, 1.0),
         "generic_sh_2000_160x96": (lambda: scenes.generic_scene(), 3, 0.9)}
GRADS = {"means3D": "dL_dmeans3D", "means2D": "dL_dmeans2D", "colors_precomp": "dL_dcolors", "opacities": "dL_dopacity",
         "cov3D_precomp": "dL_dcov3D", "shs": "dL_dsh", "scales": "dL_dscales", "rotations": "dL_drotations"}


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_matches_reference_cuda(name):
    mk, deg, mod = CASES[name]
    sc = mk()
    G = np.load(os.path.join(GOLD, f"ref_cuda_{name}.npz"))
    fw = scenes.oracle_forward(sc, deg, mod)
    H, W = sc["H"], sc["W"]
    assert fw["N"] == int(G["N"])
    assert np.array_equal(fw["radii"], G["radii"])
    assert np.array_equal(fw["tiles_touched"], G["tiles_touched"])
    vis = fw["radii"] > 0
    for k in ("depths", "means2D", "conic_opacity"):
        assert np.array_equal(bits(fw[k][vis]), bits(G[k][vis])), k
    if sc.get("scales") is not None:
        assert np.array_equal(bits(fw["cov3D"][vis]), bits(G["cov3D"][vis])), "cov3D from scale/rotation"
        assert np.abs(fw["rgb"][vis] - G["rgb"][vis]).max() <= 2e-7  # SH colour: tolerance only
    assert np.array_equal(fw["keys"], G["keys"])
    assert np.array_equal(fw["point_list"], G["point_list"])
    assert np.array_equal(fw["ranges"], G["ranges"])
    assert np.array_equal(fw["n_contrib"], G["n_contrib"].reshape(H, W))
    assert np.abs(fw["final_T"] - G["final_T"].reshape(H, W)).max() <= 1e-5
    assert np.abs(fw["color"] - G["color"]).max() <= 1e-5
    bw = gs_oracle.backward(fw, G["dL_dpix"])
    for k, v in GRADS.items():
        if "grad_" + k in G:
            ref = G["grad_" + k].astype(np.float64)
            err = np.abs(bw[v].astype(np.float64) - ref).max() / max(np.abs(ref).max(), 1e-30)
            assert err <= 1e-4, (k, err)


CFGA = {"cfgA_random_init_50k_512": lambda: scenes.avatar_scene(P=50000, W=512, H=512, random_init=True),
        "cfgA_trained_like_50k_512": lambda: scenes.avatar_scene(P=50000, W=512, H=512, random_init=False)}


@pytest.mark.parametrize("name", sorted(CFGA))
def test_oracle_matches_reference_cuda_at_bench_size(name):
    """The oracle pinned AT SCALE: BASELINE.json configs[1]/[2] (50 000 Gaussians, 512x512), random-init and trained-like, against
    the reference's CUDA build (compact fixtures of tests/golden/make_ref_cuda_golden.py `cfgA`: SHA-256 of the integer artefacts,
    the contributor counts, an image window + 16x16 block means, gradient sums + leading rows)."""
    import hashlib
    sha = lambda a: hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()
    sc = CFGA[name]()
    G = np.load(os.path.join(GOLD, f"ref_cuda_{name}.npz"))
    fw = scenes.oracle_forward(sc, 3, 1.0)
    H, W = sc["H"], sc["W"]
    assert fw["N"] == int(G["N"])
    assert sha(fw["radii"].astype(np.int32)) == str(G["sha_radii"])
    assert sha(fw["tiles_touched"].astype(np.uint32)) == str(G["sha_tiles_touched"])
    assert sha(fw["keys"].astype(np.uint64)) == str(G["sha_keys"])
    assert sha(fw["point_list"].astype(np.uint32)) == str(G["sha_point_list"])
    assert sha(fw["ranges"].astype(np.uint32)) == str(G["sha_ranges"])
    nc = fw["n_contrib"].reshape(H, W)
    assert int((nc != G["n_contrib"]).sum()) == 0, "per-pixel contributor counts differ from the reference build"
    assert sha(nc.astype(np.uint32)) == str(G["sha_n_contrib"])
    crop = (slice(192, 320), slice(192, 320))
    assert np.abs(fw["color"][:, crop[0], crop[1]] - G["color_crop"]).max() <= 1e-5
    assert np.abs(fw["final_T"].reshape(H, W)[crop] - G["final_T_crop"]).max() <= 1e-5
    blk = fw["color"].astype(np.float64).reshape(3, H // 16, 16, W // 16, 16).mean((2, 4))
    assert np.abs(blk - G["color_block_mean"]).max() <= 1e-5
    blkT = fw["final_T"].astype(np.float64).reshape(H // 16, 16, W // 16, 16).mean((1, 3))
    assert np.abs(blkT - G["final_T_block_mean"]).max() <= 1e-5
    bw = gs_oracle.backward(fw, scenes.l1_grad(fw["color"]))
    for k, v in GRADS.items():
        if "grad_" + k + "_head" in G:
            ours = bw[v].astype(np.float64)
            scale = max(float(G["grad_" + k + "_max_abs"]), 1e-30)
            head = G["grad_" + k + "_head"].astype(np.float64)
            assert np.abs(ours.reshape(ours.shape[0], -1)[:head.shape[0]] - head).max() / scale <= 1e-4, k
            assert abs(np.abs(ours).sum() - float(G["grad_" + k + "_abs_sum"])) <= 1e-4 * float(G["grad_" + k + "_abs_sum"]) + 1e-12, k
