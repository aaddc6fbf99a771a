"""Fused opacity pass.  The reference's render() rasterises the same geometry twice when
return_opacity is set -- once with the colours, once with colors_precomp = ones, keeping channel 0
(gaussian_renderer/__init__.py:121-142) -- and autograd sums the two backward calls.  Here one forward blend and one
backward blend do both.  Oracle = the CPU restatement run TWICE exactly like the reference (second call with ones):
opacity image bit-exact against the second call's channel 0, colour image bit-exact against the first, gradients
<= 1e-4 relative against the SUM of the two oracle backward calls (tolerance of north_star; sums are order-dependent).
Also checked: the fused pass against two calls of this library's own plain API (the reference's exact call sequence)."""
import numpy as np
import pytest
import torch

import scenes
import synth

pytestmark = pytest.mark.gpu

GEOM = ("means3D", "opacities", "cov3D_precomp", "scales", "rotations")


def _dev():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch.device("cuda:0")


@pytest.fixture(autouse=True)
def oracle_exp_mode(built_lib):
    """Bit-exact comparisons with the CPU oracle: the blend runs with the polynomial exp the oracle reproduces."""
    from diff_gaussian_rasterization import _C
    prev = _C.set_exp_mode(1)
    yield
    _C.set_exp_mode(prev)


def rel_err(a, b):
    scale = max(float(np.abs(b).max()), 1e-30)
    return float(np.abs(a.astype(np.float64) - b.astype(np.float64)).max()) / scale


def _settings(sc, dev, sh_degree=3):
    from diff_gaussian_rasterization import GaussianRasterizationSettings
    cam = sc["cam"]
    d = lambda a: torch.as_tensor(np.ascontiguousarray(a)).to(dev)
    return GaussianRasterizationSettings(sc["H"], sc["W"], cam["tanfovx"], cam["tanfovy"], d(sc["bg"]), 1.0,
                                         d(cam["world_view_transform"]), d(cam["full_proj_transform"]), sh_degree,
                                         d(cam["camera_center"]), False, False)


def _leaves(sc, dev):
    d = lambda a: torch.as_tensor(np.ascontiguousarray(a)).to(dev)
    names = ["means3D", "opacities", "colors_precomp", "shs", "scales", "rotations", "cov3D_precomp"]
    t = {k: (d(sc[k]).requires_grad_(True) if sc.get(k) is not None else None) for k in names}
    t["means2D"] = torch.zeros_like(t["means3D"], requires_grad=True)
    return t


def _grads(t):
    return {k: (v.grad.detach().cpu().numpy() if v is not None and v.grad is not None else None) for k, v in t.items()}


def run_fused(sc, dL_img, dL_op):
    from diff_gaussian_rasterization import GaussianRasterizer
    dev = _dev()
    t = _leaves(sc, dev)
    img, radii, op = GaussianRasterizer(_settings(sc, dev))(
        means3D=t["means3D"], means2D=t["means2D"], opacities=t["opacities"], shs=t["shs"],
        colors_precomp=t["colors_precomp"], scales=t["scales"], rotations=t["rotations"],
        cov3D_precomp=t["cov3D_precomp"], return_opacity=True)
    assert op.shape == (1, sc["H"], sc["W"])
    d = lambda a: torch.as_tensor(np.ascontiguousarray(a)).to(dev)
    torch.autograd.backward([img, op], [d(dL_img), d(dL_op)])
    torch.cuda.synchronize()
    return img.detach().cpu().numpy(), op.detach().cpu().numpy(), radii.cpu().numpy(), _grads(t)


def run_two_calls(sc, dL_img, dL_op):
    """The reference's call sequence through the plain drop-in API."""
    from diff_gaussian_rasterization import GaussianRasterizer
    dev = _dev()
    t = _leaves(sc, dev)
    r = GaussianRasterizer(_settings(sc, dev))
    img, radii = r(means3D=t["means3D"], means2D=t["means2D"], opacities=t["opacities"], shs=t["shs"],
                   colors_precomp=t["colors_precomp"], scales=t["scales"], rotations=t["rotations"],
                   cov3D_precomp=t["cov3D_precomp"])
    op, _ = r(means3D=t["means3D"], means2D=t["means2D"], opacities=t["opacities"], shs=None,
              colors_precomp=torch.ones(t["means3D"].shape[0], 3, device=dev), scales=t["scales"],
              rotations=t["rotations"], cov3D_precomp=t["cov3D_precomp"])
    op = op[:1]
    d = lambda a: torch.as_tensor(np.ascontiguousarray(a)).to(dev)
    torch.autograd.backward([img, op], [d(dL_img), d(dL_op)])
    torch.cuda.synchronize()
    return img.detach().cpu().numpy(), op.detach().cpu().numpy(), radii.cpu().numpy(), _grads(t)


def oracle_two_calls(sc, dL_img, dL_op, sh_degree=3):
    from oracle import gs_oracle
    fw1 = scenes.oracle_forward(sc, sh_degree=sh_degree)
    sc2 = dict(sc)
    sc2["shs"] = None
    sc2["colors_precomp"] = np.ones((sc["means3D"].shape[0], 3), np.float32)
    fw2 = scenes.oracle_forward(sc2, sh_degree=sh_degree)
    bw1 = gs_oracle.backward(fw1, dL_img)
    dL2 = np.zeros_like(dL_img)
    dL2[0] = dL_op[0]  # `[:1]`: channels 1 and 2 of the second image receive no gradient
    bw2 = gs_oracle.backward(fw2, dL2)
    return fw1, fw2, bw1, bw2


def _upstream(fw1, fw2, seed=2):
    dL_img = scenes.l1_grad(fw1["color"], seed)
    H, W = fw1["H"], fw1["W"]
    mask = (synth.l1_target(H, W, seed + 5)[:1] > 0.5).astype(np.float32)  # mask loss, train.py:143-153 (l1 variant)
    dL_op = (0.1 * np.sign(fw2["color"][:1] - mask) / (H * W)).astype(np.float32)
    return dL_img, dL_op


@pytest.mark.parametrize("P,W,H,bg", [(3000, 128, 128, (0, 0, 0)), (20000, 256, 256, (1, 1, 1)), (1500, 200, 120, (0.3, 0.6, 0.1))])
def test_fused_opacity_vs_oracle_two_calls(built_lib, P, W, H, bg):
    sc = scenes.avatar_scene(P=P, W=W, H=H)
    sc["bg"] = np.asarray(bg, np.float32)
    fw1 = scenes.oracle_forward(sc)
    sc2 = dict(sc, colors_precomp=np.ones((P, 3), np.float32))
    fw2 = scenes.oracle_forward(sc2)
    dL_img, dL_op = _upstream(fw1, fw2)
    fw1, fw2, bw1, bw2 = oracle_two_calls(sc, dL_img, dL_op)
    img, op, radii, g = run_fused(sc, dL_img, dL_op)
    assert np.array_equal(radii, fw1["radii"])
    assert np.array_equal(img.view(np.uint32), fw1["color"].view(np.uint32)), "colour image not bit-identical"
    assert np.array_equal(op.view(np.uint32), fw2["color"][:1].view(np.uint32)), "opacity image not bit-identical"
    assert float(op.max()) > 0.5  # the avatar covers pixels
    assert rel_err(g["means3D"], bw1["dL_dmeans3D"] + bw2["dL_dmeans3D"]) <= 1e-4
    assert rel_err(g["means2D"], bw1["dL_dmeans2D"] + bw2["dL_dmeans2D"]) <= 1e-4
    assert rel_err(g["opacities"], bw1["dL_dopacity"] + bw2["dL_dopacity"]) <= 1e-4
    assert rel_err(g["cov3D_precomp"], bw1["dL_dcov3D"] + bw2["dL_dcov3D"]) <= 1e-4
    assert rel_err(g["colors_precomp"], bw1["dL_dcolors"]) <= 1e-4  # the second call's colours are constants


def test_fused_opacity_vs_two_library_calls_sh_path(built_lib):
    """SH + scale/rotation path, non-zero background, sizes that are not multiples of 16."""
    sc = scenes.generic_scene()
    fw1 = scenes.oracle_forward(sc)
    sc2 = dict(sc, shs=None, colors_precomp=np.ones((sc["means3D"].shape[0], 3), np.float32))
    fw2 = scenes.oracle_forward(sc2)
    dL_img, dL_op = _upstream(fw1, fw2)
    img_f, op_f, radii_f, gf = run_fused(sc, dL_img, dL_op)
    img_t, op_t, radii_t, gt = run_two_calls(sc, dL_img, dL_op)
    assert np.array_equal(radii_f, radii_t)
    assert np.array_equal(img_f.view(np.uint32), img_t.view(np.uint32))
    assert np.array_equal(op_f.view(np.uint32), op_t.view(np.uint32))
    assert np.array_equal(op_f.view(np.uint32), fw2["color"][:1].view(np.uint32))
    for k in ("means3D", "means2D", "opacities", "scales", "rotations", "shs"):
        assert rel_err(gf[k], gt[k]) <= 1e-4, k


def test_fused_opacity_only_one_upstream(built_lib):
    """Only the opacity image carries a gradient (autograd hands None for the colour image)."""
    from diff_gaussian_rasterization import GaussianRasterizer
    from oracle import gs_oracle
    dev = _dev()
    sc = scenes.avatar_scene(P=3000, W=128, H=128)
    sc2 = dict(sc, colors_precomp=np.ones((3000, 3), np.float32))
    fw2 = scenes.oracle_forward(sc2)
    t = _leaves(sc, dev)
    img, radii, op = GaussianRasterizer(_settings(sc, dev))(
        means3D=t["means3D"], means2D=t["means2D"], opacities=t["opacities"], colors_precomp=t["colors_precomp"],
        cov3D_precomp=t["cov3D_precomp"], return_opacity=True)
    op.sum().backward()
    torch.cuda.synchronize()
    dL2 = np.zeros((3, 128, 128), np.float32)
    dL2[0] = 1.0
    bw2 = gs_oracle.backward(fw2, dL2)
    g = _grads(t)
    assert rel_err(g["means3D"], bw2["dL_dmeans3D"]) <= 1e-4
    assert rel_err(g["opacities"], bw2["dL_dopacity"]) <= 1e-4
    assert float(np.abs(g["colors_precomp"]).max()) == 0.0


def test_fused_opacity_empty_and_invisible(built_lib):
    """P == 0 short-circuits to zeros like the plain path (rasterize_points.cu:81); a scene entirely behind the camera gives
    opacity == bg[0] * 1 everywhere (T stays 1) and zero gradients."""
    from diff_gaussian_rasterization import GaussianRasterizer
    dev = _dev()
    sc = scenes.avatar_scene(P=500, W=64, H=48)
    rs = _settings(sc, dev)
    z = lambda *s: torch.zeros(*s, device=dev)
    img, radii, op = GaussianRasterizer(rs)(means3D=z(0, 3), means2D=z(0, 3), opacities=z(0, 1), colors_precomp=z(0, 3),
                                            cov3D_precomp=z(0, 6), return_opacity=True)
    assert img.shape == (3, 48, 64) and op.shape == (1, 48, 64) and radii.numel() == 0
    assert float(img.abs().max()) == 0.0 and float(op.abs().max()) == 0.0
    sc["bg"] = np.asarray((0.25, 0.5, 0.75), np.float32)
    sc["means3D"] = sc["means3D"] + np.asarray([0, 0, 100.0], np.float32)  # far behind the camera
    t = _leaves(sc, dev)
    img, radii, op = GaussianRasterizer(_settings(sc, dev))(means3D=t["means3D"], means2D=t["means2D"], opacities=t["opacities"],
                                                            colors_precomp=t["colors_precomp"], cov3D_precomp=t["cov3D_precomp"],
                                                            return_opacity=True)
    (img.sum() + op.sum()).backward()
    torch.cuda.synchronize()
    assert int((radii > 0).sum()) == 0
    assert torch.all(op == 0.25) and torch.all(img[2] == 0.75)
    assert float(t["means3D"].grad.abs().max()) == 0.0
