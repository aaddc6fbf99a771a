"""GPU parity tests of the tile rasterizer: the CUDA path (through the drop-in package and the C ABI) against the CPU
oracle on identical seeded inputs.  Bit-exact: radii, tiles_touched, sort keys, point list, tile ranges, n_contrib (and,
because the FP32 operation order is shared by specification, final_T and the image).  Gradients: <= 1e-4 relative
(north_star tolerance; the sums are order-dependent)."""
import numpy as np
import pytest
import torch

import scenes

pytestmark = pytest.mark.gpu


def _dev():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch.device("cuda:0")


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


@pytest.mark.parametrize("P,W,H", [(3000, 128, 128), (20000, 256, 256), (1500, 200, 120)])
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
