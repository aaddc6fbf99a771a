"""GPU parity tests of the tile rasterizer: the CUDA path (through the drop-in package and the C ABI) against the CPU
oracle on identical seeded inputs.  Bit-exact: radii, tiles_touched, sort keys, point list, tile ranges, n_contrib (and,
because the FP32 operation order is shared by specification, final_T and the image).  Gradients: <= 1e-4 relative
(north_star tolerance; the sums are order-dependent).

The tests of this file run the blend in ORACLE EXP MODE (gs_set_exp_mode(1): the polynomial exp that oracle/gs_oracle.c
reproduces bit for bit); the default mode -- CUDA's expf, the reference's instruction sequence -- is compared with the
reference's own CUDA build in tests/test_gpu_vs_reference_cuda.py and with the oracle in test_default_exp_mode_vs_oracle."""
import os

import numpy as np
import pytest
import torch

import scenes

pytestmark = pytest.mark.gpu


def _dev():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch.device("cuda:0")


@pytest.fixture(autouse=True)
def oracle_exp_mode(built_lib):
    from diff_gaussian_rasterization import _C
    prev = _C.set_exp_mode(1)
    yield
    _C.set_exp_mode(prev)


def run_cuda(sc, sh_degree=3, scale_modifier=1.0, dL=None, debug=False):
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer, _C
    dev = _dev()
    cam = sc["cam"]
    d = lambda a: None if a is None else torch.as_tensor(np.ascontiguousarray(a)).to(dev)
    rs = GaussianRasterizationSettings(sc["H"], sc["W"], cam["tanfovx"], cam["tanfovy"], d(sc["bg"]), scale_modifier,
                                       d(cam["world_view_transform"]), d(cam["full_proj_transform"]), sh_degree,
                                       d(cam["camera_center"]), False, debug)
    names = ["means3D", "opacities", "colors_precomp", "shs", "scales", "rotations", "cov3D_precomp"]
    t = {k: (d(sc.get(k)).requires_grad_(True) if sc.get(k) is not None else None) for k in names}
    means2D = torch.zeros_like(t["means3D"], requires_grad=True)
    img, radii = GaussianRasterizer(rs)(means3D=t["means3D"], means2D=means2D, opacities=t["opacities"], shs=t["shs"],
                                        colors_precomp=t["colors_precomp"], scales=t["scales"],
                                        rotations=t["rotations"], cov3D_precomp=t["cov3D_precomp"])
    fn = img.grad_fn
    saved = fn.saved_tensors
    geom, lst, imgbuf = saved[7], saved[8], saved[9]
    N = fn.num_rendered
    P = sc["means3D"].shape[0]
    dbg = {k: v.cpu().numpy() for k, v in _C.debug_state(P, sc["W"], sc["H"], N, radii, geom, lst, imgbuf).items()}
    out = dict(img=img.detach().cpu().numpy(), radii=radii.cpu().numpy(), N=N, dbg=dbg)
    if dL is not None:
        img.backward(d(dL))
        out["E_bwd"] = _C.read_counters(geom)[0]  # counted by the backward blend
        out["grads"] = {k: (v.grad.cpu().numpy() if v is not None and v.grad is not None else None) for k, v in t.items()}
        out["grads"]["means2D"] = means2D.grad.cpu().numpy()
    torch.cuda.synchronize()
    return out


def check_forward_exact(fw, out):
    H, W = fw["H"], fw["W"]
    dbg = out["dbg"]
    assert out["N"] == fw["N"]
    assert np.array_equal(out["radii"], fw["radii"])
    assert np.array_equal(dbg["tiles_touched"].view(np.uint32), fw["tiles_touched"])
    vis = fw["radii"] > 0
    assert np.array_equal(dbg["depths"].view(np.uint32)[vis], fw["depths"].view(np.uint32)[vis])
    assert np.array_equal(dbg["means2D"].reshape(-1, 2).view(np.uint32)[vis], fw["means2D"].view(np.uint32)[vis])
    assert np.array_equal(dbg["conic_opacity"].reshape(-1, 4).view(np.uint32)[vis], fw["conic_opacity"].view(np.uint32)[vis])
    assert np.array_equal(dbg["keys"].view(np.uint64), fw["keys"]), "sorted keys differ"
    assert np.array_equal(dbg["point_list"].view(np.uint32), fw["point_list"]), "point list differs"
    assert np.array_equal(dbg["ranges"].view(np.uint32).reshape(-1, 2), fw["ranges"]), "tile ranges differ"
    assert np.array_equal(dbg["n_contrib"].view(np.uint32).reshape(H, W), fw["n_contrib"]), "n_contrib differs"
    assert np.array_equal(dbg["final_T"].reshape(H, W).view(np.uint32), fw["final_T"].view(np.uint32)), "final_T differs"
    assert np.abs(out["img"] - fw["color"]).max() <= 1e-5
    assert np.array_equal(out["img"].view(np.uint32), fw["color"].view(np.uint32)), "image not bit-identical"


def rel_err(a, b):
    """max |a-b| / max|b| (gradient tolerance of north_star: 1e-4 relative)."""
    scale = max(float(np.abs(b).max()), 1e-30)
    return float(np.abs(a.astype(np.float64) - b.astype(np.float64)).max()) / scale


@pytest.mark.parametrize("P,W,H", [(3000, 128, 128), (20000, 256, 256), (1500, 200, 120), (3001, 128, 96)])
def test_avatar_forward_backward(built_lib, P, W, H):
    from oracle import gs_oracle
    sc = scenes.avatar_scene(P=P, W=W, H=H)
    fw = scenes.oracle_forward(sc)
    assert fw["N"] > 0
    dL = scenes.l1_grad(fw["color"])
    out = run_cuda(sc, dL=dL)
    check_forward_exact(fw, out)
    bw = gs_oracle.backward(fw, dL)
    g = out["grads"]
    assert rel_err(g["means3D"], bw["dL_dmeans3D"]) <= 1e-4
    assert rel_err(g["means2D"], bw["dL_dmeans2D"]) <= 1e-4
    assert rel_err(g["colors_precomp"], bw["dL_dcolors"]) <= 1e-4
    assert rel_err(g["opacities"], bw["dL_dopacity"]) <= 1e-4
    assert rel_err(g["cov3D_precomp"], bw["dL_dcov3D"]) <= 1e-4
    assert np.all(g["means2D"][:, 2] == 0)


def test_sh_scale_rot_path(built_lib):
    """In-rasterizer SH colours + covariance from scales/rotations, non-zero background, W/H not multiples of 16,
    points behind the camera."""
    from oracle import gs_oracle
    sc = scenes.generic_scene()
    fw = scenes.oracle_forward(sc, sh_degree=3, scale_modifier=0.9)
    assert (fw["radii"] == 0).any() and (fw["radii"] > 0).any()
    dL = scenes.l1_grad(fw["color"])
    out = run_cuda(sc, sh_degree=3, scale_modifier=0.9, dL=dL)
    check_forward_exact(fw, out)
    vis = fw["radii"] > 0
    assert np.array_equal(out["dbg"]["rgb"].reshape(-1, 3).view(np.uint32)[vis], fw["rgb"].view(np.uint32)[vis])
    assert np.array_equal(out["dbg"]["cov3D"].reshape(-1, 6).view(np.uint32)[vis], fw["cov3D"].view(np.uint32)[vis])
    assert np.array_equal(out["dbg"]["clamped"].reshape(-1, 3)[vis], fw["clamped"][vis])
    bw = gs_oracle.backward(fw, dL)
    g = out["grads"]
    assert rel_err(g["means3D"], bw["dL_dmeans3D"]) <= 1e-4
    assert rel_err(g["shs"], bw["dL_dsh"]) <= 1e-4
    assert rel_err(g["scales"], bw["dL_dscales"]) <= 1e-4
    assert rel_err(g["rotations"], bw["dL_drotations"]) <= 1e-4
    assert rel_err(g["opacities"], bw["dL_dopacity"]) <= 1e-4


@pytest.mark.parametrize("deg", [0, 1, 2])
def test_lower_sh_degrees(built_lib, deg):
    sc = scenes.generic_scene(P=800, W=96, H=64, seed=5)
    fw = scenes.oracle_forward(sc, sh_degree=deg)
    out = run_cuda(sc, sh_degree=deg)
    check_forward_exact(fw, out)


def test_bench_config_50k_512(built_lib):
    """BASELINE.json configs[1]/[2] at full size (50k random-init Gaussians, 512x512): every integer artefact and the image
    bit-exact vs the oracle, gradients <= 1e-4 relative."""
    from oracle import gs_oracle
    sc = scenes.avatar_scene(P=50000, W=512, H=512, random_init=True)
    fw = scenes.oracle_forward(sc)
    dL = scenes.l1_grad(fw["color"])
    out = run_cuda(sc, dL=dL)
    check_forward_exact(fw, out)
    bw = gs_oracle.backward(fw, dL)
    g = out["grads"]
    for k, v in (("means3D", "dL_dmeans3D"), ("means2D", "dL_dmeans2D"), ("colors_precomp", "dL_dcolors"),
                 ("opacities", "dL_dopacity"), ("cov3D_precomp", "dL_dcov3D")):
        assert rel_err(g[k], bw[v]) <= 1e-4, k


def test_many_gaussians_small_grid_unfused_placement(built_lib):
    """130k Gaussians on a 128x128 image: more than GS_BIN_FUSED_MAX_BLOCKS blocks of 1024, so direct placement runs as
    two kernels (count, place) instead of the co-resident fused grid; integer artefacts and image bit-exact."""
    sc = scenes.avatar_scene(P=130000, W=128, H=128)
    fw = scenes.oracle_forward(sc)
    assert fw["N"] > 100000
    check_forward_exact(fw, run_cuda(sc))


def test_trained_like_opaque_scene(built_lib):
    """High-opacity Gaussians: pixels terminate early (T < 1e-4), exercising the done / early-exit paths of the pipelined
    forward and the checkpoint restarts of the segmented backward."""
    from oracle import gs_oracle
    sc = scenes.avatar_scene(P=30000, W=256, H=256)
    sc["opacities"] = np.clip(sc["opacities"] * 0 + 0.97, 0, 1).astype(np.float32)
    sc["cov3D_precomp"] = (sc["cov3D_precomp"] * 9.0).astype(np.float32)  # 3x larger splats: deep, opaque lists
    fw = scenes.oracle_forward(sc)
    assert (fw["final_T"] < 1e-3).sum() > 1000 and (fw["ranges"][:, 1] - fw["ranges"][:, 0]).max() > 4 * 256
    dL = scenes.l1_grad(fw["color"])
    out = run_cuda(sc, dL=dL)
    check_forward_exact(fw, out)
    bw = gs_oracle.backward(fw, dL)
    g = out["grads"]
    for k, v in (("means3D", "dL_dmeans3D"), ("means2D", "dL_dmeans2D"), ("colors_precomp", "dL_dcolors"),
                 ("opacities", "dL_dopacity"), ("cov3D_precomp", "dL_dcov3D")):
        assert rel_err(g[k], bw[v]) <= 1e-4, k


def test_stress_1m_1080p_properties(built_lib):
    """BASELINE.json configs[4] size (1M Gaussians, 1920x1080): size-independent properties instead of the oracle --
    sortedness of the (tile | depth) keys, ranges partition the list, sum(tiles_touched) = N, contributor counts within the
    tile lists, run-to-run bit-identical forward, finite gradients."""
    from diff_gaussian_rasterization import _C
    rng = np.random.default_rng(0)
    P, W, H = 1_000_000, 1920, 1080
    import synth
    cam = synth.camera(W=W, H=H, fx=1600.0, fy=1600.0, dist=3.5)
    means = (rng.normal(0, 1, size=(P, 3)) * np.array([1.2, 0.7, 0.5]) + np.array([0, 0.9, 0])).astype(np.float32)
    s = np.exp(rng.uniform(-6.0, -4.3, size=(P, 3))).astype(np.float32)
    q = rng.normal(size=(P, 4)).astype(np.float32)
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    sc = dict(cam=cam, W=W, H=H, bg=np.array([0.1, 0.2, 0.3], np.float32), means3D=means, scales=s, rotations=q,
              opacities=rng.uniform(0.05, 0.9, size=(P, 1)).astype(np.float32),
              colors_precomp=rng.uniform(0, 1, size=(P, 3)).astype(np.float32))
    dL = (rng.normal(size=(3, H, W)) / (3 * H * W)).astype(np.float32)
    out = run_cuda(sc, dL=dL)
    N, dbg = out["N"], out["dbg"]
    assert N > 2_000_000
    keys = dbg["keys"].view(np.uint64)
    assert np.all(keys[1:] >= keys[:-1]), "keys not sorted"
    assert int(dbg["tiles_touched"].view(np.uint32).astype(np.int64).sum()) == N
    ranges = dbg["ranges"].view(np.uint32).reshape(-1, 2).astype(np.int64)
    ne = ranges[ranges[:, 1] > ranges[:, 0]]
    assert ne[0, 0] == 0 and ne[-1, 1] == N and np.all(ne[1:, 0] == ne[:-1, 1]), "ranges do not partition the list"
    tiles_of_keys = (keys >> np.uint64(32)).astype(np.int64)
    tix = np.nonzero(ranges[:, 1] > ranges[:, 0])[0]
    assert np.all(tiles_of_keys[ne[:, 0]] == tix) and np.all(tiles_of_keys[ne[:, 1] - 1] == tix)
    gx = (W + 15) // 16
    ncon = dbg["n_contrib"].view(np.uint32).reshape(H, W).astype(np.int64)
    lens = (ranges[:, 1] - ranges[:, 0]).reshape(-1, gx)
    assert np.all(ncon <= np.kron(lens, np.ones((16, 16), np.int64))[:H, :W])
    assert np.isfinite(out["img"]).all() and out["img"].min() >= 0
    for k, v in out["grads"].items():
        if v is not None:
            assert np.isfinite(v).all(), k
    out2 = run_cuda(sc)
    assert np.array_equal(out2["img"].view(np.uint32), out["img"].view(np.uint32)), "forward not run-to-run identical"
    assert out2["N"] == N and np.array_equal(out2["dbg"]["point_list"], dbg["point_list"])


def test_graph_capture_and_capacity_overflow_recapture(built_lib):
    """CUDA-graph replay of forward+backward equals the eager result; a pair capacity that is too small for the frame is
    detected after the replay, the step is re-captured with a larger capacity and re-run (avatar_pose.graphs.GraphedStep)."""
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer, _C
    from avatar_pose.graphs import GraphedStep
    dev = _dev()
    sc = scenes.avatar_scene(P=20000, W=256, H=256)
    fw = scenes.oracle_forward(sc)
    cam = sc["cam"]
    d = lambda a: torch.as_tensor(np.ascontiguousarray(a)).to(dev)
    rs = GaussianRasterizationSettings(256, 256, cam["tanfovx"], cam["tanfovy"], d(sc["bg"]), 1.0, d(cam["world_view_transform"]),
                                       d(cam["full_proj_transform"]), 3, d(cam["camera_center"]), False, False)
    means = d(sc["means3D"]).requires_grad_(True)
    cols, cov, op = d(sc["colors_precomp"]).requires_grad_(True), d(sc["cov3D_precomp"]), d(sc["opacities"])

    def step():
        means.grad = None
        cols.grad = None
        img, radii = GaussianRasterizer(rs)(means3D=means, means2D=torch.zeros_like(means, requires_grad=True), opacities=op,
                                            colors_precomp=cols, cov3D_precomp=cov)
        img.mean().backward()
        return img

    g = GraphedStep(step)
    img = g.replay(check=True).clone()
    grad = means.grad.clone()
    assert np.array_equal(img.detach().cpu().numpy().view(np.uint32), fw["color"].view(np.uint32))
    # now force an overflow: shrink the capacity the NEXT capture will use, re-capture, replay with the check
    st = _C._dev_state(dev)
    assert fw["N"] > 8192
    st.capacity = 4096
    regrows = _C.stats["capacity_regrows"]
    g.capture()
    img2 = g.replay(check=True)
    assert _C.stats["capacity_regrows"] > regrows and st.capacity >= fw["N"]
    assert np.array_equal(img2.detach().cpu().numpy().view(np.uint32), fw["color"].view(np.uint32))
    assert torch.allclose(means.grad, grad, rtol=1e-4, atol=1e-9)


def test_default_exp_mode_vs_oracle(built_lib):
    """Default mode (CUDA expf, the reference's instruction sequence) against the oracle (polynomial exp, <= 1 ulp from expf):
    every artefact that does not depend on exp stays bit-exact; n_contrib may differ only where an alpha sits within an ulp of
    1/255 (or T of 1e-4): at most a handful of pixels; image <= 1e-5; gradients <= 1e-4."""
    from diff_gaussian_rasterization import _C
    from oracle import gs_oracle
    sc = scenes.avatar_scene(P=20000, W=256, H=256)
    fw = scenes.oracle_forward(sc)
    dL = scenes.l1_grad(fw["color"])
    prev = _C.set_exp_mode(0)
    try:
        out = run_cuda(sc, dL=dL)
    finally:
        _C.set_exp_mode(prev)
    dbg = out["dbg"]
    assert out["N"] == fw["N"] and np.array_equal(out["radii"], fw["radii"])
    assert np.array_equal(dbg["keys"].view(np.uint64), fw["keys"]) and np.array_equal(dbg["point_list"].view(np.uint32), fw["point_list"])
    assert np.array_equal(dbg["ranges"].view(np.uint32).reshape(-1, 2), fw["ranges"])
    nmis = int((dbg["n_contrib"].view(np.uint32).reshape(256, 256) != fw["n_contrib"]).sum())
    assert nmis <= 4, f"{nmis} n_contrib pixels differ between expf and the polynomial exp"
    assert np.abs(out["img"] - fw["color"]).max() <= 1e-5
    assert np.abs(dbg["final_T"].reshape(256, 256) - fw["final_T"]).max() <= 1e-5
    bw = gs_oracle.backward(fw, dL)
    for k, v in (("means3D", "dL_dmeans3D"), ("means2D", "dL_dmeans2D"), ("colors_precomp", "dL_dcolors"),
                 ("opacities", "dL_dopacity"), ("cov3D_precomp", "dL_dcov3D")):
        assert rel_err(out["grads"][k], bw[v]) <= 1e-4, k


@pytest.mark.parametrize("opacity", [0.05, 0.6])
def test_deep_lists_chunked_chain(built_lib, opacity):
    """Tile lists of several thousand entries (many 512-entry forward chunks per tile, hand-off of the chain state between
    CTAs): translucent (no pixel terminates: every chunk runs) and opaque enough that pixels terminate inside a middle chunk
    (later chunks receive "done" and only pass the state on).  Everything bit-exact vs the oracle; E_bwd counter exact."""
    from diff_gaussian_rasterization import _C
    from oracle import gs_oracle
    rng = np.random.default_rng(5)
    P, W, H = 24000, 64, 48
    import synth
    cam = synth.camera(W=W, H=H, fx=70.0, fy=70.0, dist=3.5)
    means = (rng.normal(0, 1, size=(P, 3)) * np.array([0.5, 0.4, 0.5]) + np.array([0, 0.9, 0])).astype(np.float32)
    s = np.exp(rng.uniform(-3.2, -2.2, size=(P, 3))).astype(np.float32)
    q = rng.normal(size=(P, 4)).astype(np.float32)
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    sc = dict(cam=cam, W=W, H=H, bg=np.array([0.1, 0.2, 0.3], np.float32), means3D=means, scales=s, rotations=q,
              opacities=np.full((P, 1), opacity, np.float32), colors_precomp=rng.uniform(0, 1, size=(P, 3)).astype(np.float32))
    fw = scenes.oracle_forward(sc)
    lens = fw["ranges"][:, 1] - fw["ranges"][:, 0]
    assert lens.max() > 4 * 512, f"deepest list {lens.max()}"
    dL = scenes.l1_grad(fw["color"])
    out = run_cuda(sc, dL=dL)
    check_forward_exact(fw, out)
    assert out["E_bwd"] == fw["evals_blended"], (out["E_bwd"], fw["evals_blended"])
    if opacity > 0.5:
        assert (fw["n_contrib"].max(axis=None) < lens.max()), "expected early termination"
    bw = gs_oracle.backward(fw, dL)
    g = out["grads"]
    for k, v in (("means3D", "dL_dmeans3D"), ("means2D", "dL_dmeans2D"), ("colors_precomp", "dL_dcolors"),
                 ("opacities", "dL_dopacity"), ("scales", "dL_dscales"), ("rotations", "dL_drotations")):
        assert rel_err(g[k], bw[v]) <= 1e-4, k


def test_mark_visible(built_lib):
    """GaussianRasterizer.markVisible (rasterize_points.cu:198-217 -> checkFrustum, rasterizer_impl.cu:54-66: p_view.z > 0.2)
    against the oracle, bit-exact, on a cloud with points in front of, behind and right at the near plane; and the same mask
    through the reference build when baseline/_ref is on the box."""
    import os
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    from oracle import gs_oracle
    dev = _dev()
    sc = scenes.generic_scene(P=50001, W=64, H=64, seed=21, spread=2.5)
    cam = sc["cam"]
    means = sc["means3D"].copy()
    # a band of points whose view depth lies within a few ulp of the 0.2 threshold (the decision must be the reference's FMA order)
    V = cam["world_view_transform"]
    zaxis = V[:3, 2] / np.linalg.norm(V[:3, 2])
    depth = means @ V[:3, 2] + V[3, 2]
    k = 4000
    means[:k] = means[:k] - (depth[:k, None] - 0.2) * zaxis[None, :] / np.linalg.norm(V[:3, 2])
    means[:k] += (np.random.default_rng(0).integers(-3, 4, size=(k, 1)) * 1e-8).astype(np.float32)
    d = lambda a: torch.as_tensor(np.ascontiguousarray(a, dtype=np.float32)).to(dev)
    rs = GaussianRasterizationSettings(sc["H"], sc["W"], cam["tanfovx"], cam["tanfovy"], d(sc["bg"]), 1.0, d(cam["world_view_transform"]),
                                       d(cam["full_proj_transform"]), 3, d(cam["camera_center"]), False, False)
    mask = GaussianRasterizer(rs).markVisible(d(means))
    assert mask.dtype == torch.bool and mask.shape == (means.shape[0],)
    want = gs_oracle.mark_visible(means.astype(np.float32), cam["world_view_transform"], cam["full_proj_transform"]).astype(bool)
    got = mask.cpu().numpy()
    assert 0 < want.sum() < want.size
    assert np.array_equal(got, want), f"{int((got != want).sum())} visibility decisions differ from the oracle"
    assert GaussianRasterizer(rs).markVisible(torch.zeros(0, 3, device=dev)).numel() == 0
    ref_dir = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baseline", "_ref", "diff_gaussian_rasterization")
    if os.path.isdir(ref_dir) and any(f.endswith(".so") for f in os.listdir(ref_dir)):
        from test_gpu_vs_reference_cuda import _gen
        ref = _gen().load_ref()
        rrs = ref.GaussianRasterizationSettings(sc["H"], sc["W"], cam["tanfovx"], cam["tanfovy"], d(sc["bg"]), 1.0, d(cam["world_view_transform"]),
                                                d(cam["full_proj_transform"]), 3, d(cam["camera_center"]), False, False)
        rmask = ref.GaussianRasterizer(rrs).markVisible(d(means)).cpu().numpy()
        assert np.array_equal(got, rmask), f"{int((got != rmask).sum())} visibility decisions differ from the reference build"


@pytest.mark.parametrize("env", [{"GS_FWD_ORDER": "0"}, {"GS_FWD_ORDER": "1"}, {"GS_FWD_KERNEL": "block"}, {"GS_SORT_FUSED": "0", "GS_PREWAIT": "0"}],
                         ids=["order0", "order1", "block_kernel", "sort_chain_of_four_launches"])
def test_alternative_forward_variants(built_lib, env):
    """The forward blend has two kernels (chunked, the default for tile grids <= 2048 tiles; per-block with producer / consumer
    warps, the default above: GS_FWD_KERNEL=chunk|block forces one) and the chunked kernel three orders of its work items
    (GS_FWD_ORDER: 2 = mixed, the default; 0 = longest remaining chain first; 1 = level by level); the depth sort runs as one
    launch (default for small inputs) or as a chain of four (GS_SORT_FUSED=0, here with the prologues behind the wait,
    GS_PREWAIT=0).  All of them only SCHEDULE the
    same per-pixel operation sequence: images, n_contrib and gradients must not depend on the choice.  The variables are read once
    per process, so the bit-exact tests of the deep-list, bench-size, trained-like and A/B-against-the-reference scenes run in
    a child process."""
    import subprocess
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(here, "test_gpu_rasterizer.py"), os.path.join(here, "test_gpu_vs_reference_cuda.py"),
                        os.path.join(here, "test_gpu_opacity_pass.py"), "-q", "-m", "gpu", "-x", "-k",
                        "deep_lists or bench_config or trained_like or unfused or reference_cuda_build or opacity"],
                       env=dict(os.environ, **env), capture_output=True, text=True, timeout=1200)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-1000:]
