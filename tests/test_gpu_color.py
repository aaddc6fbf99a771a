"""GPU parity of the colour-MLP texture (gs_color.cu, tcgen05 tensor cores) against the fp64 oracle restating
ColorMLP.forward (models/texture/texture.py:74-125).  Tolerances: the composed input row (features | SH basis of the
canonical view direction | non-rigid feature) <= 2e-6 abs (FP32 elementwise math); colours <= 3e-5 abs (three layers
with the 2-way BF16 split, ~2^-16 relative per product, then a sigmoid with slope <= 1/4)."""
import numpy as np
import pytest
import torch

from oracle import color_oracle as co

pytestmark = pytest.mark.gpu

COL_TOL = 3e-5
IN_TOL = 2e-6


def _dev():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch.device("cuda:0")


def _module(c, dev, **kw):
    from avatar_pose.texture import ColorMLPFused
    m = ColorMLPFused(**kw).to(dev)
    with torch.no_grad():
        for l in range(3):
            getattr(m, f"lin{l}").weight.copy_(c["Ws"][l])
            getattr(m, f"lin{l}").bias.copy_(c["bs"][l])
    return m


def _truth(c, cano=True, noise=True):
    c64 = {k: (v.double() if torch.is_tensor(v) else v) for k, v in c.items()}
    c64["Ws"], c64["bs"] = [w.double() for w in c["Ws"]], [b.double() for b in c["bs"]]
    return co.color_forward(c64["features"], c64["xyz"], c64["T_fwd"] if cano else None, c64["camera_center"],
                            c64["non_rigid_feature"], c64["latent"], c64["Ws"], c64["bs"], c["sh_degree"], cano,
                            c64["view_noise"] if (noise and cano) else None), c64


@pytest.mark.parametrize("P", [1, 127, 128, 1000, 50000])
def test_color_forward_default_config(built_lib, P):
    dev = _dev()
    c = co.synthetic_case(P, seed=P)
    m = _module(c, dev)
    d = lambda t: None if t is None else t.to(dev)
    col, inp = m.colors(d(c["features"]), d(c["xyz"]), d(c["T_fwd"]), d(c["camera_center"]), d(c["non_rigid_feature"]),
                        d(c["latent"]), d(c["view_noise"]), return_inputs=True)
    truth, c64 = _truth(c)
    x = co.compose_input(c64["features"], c64["xyz"], c64["T_fwd"], c64["camera_center"], c64["non_rigid_feature"], None, 3, True, c64["view_noise"])
    assert inp.shape == (P, 63)
    assert float((inp.double().cpu() - x).abs().max()) <= IN_TOL
    assert float((col.double().cpu() - truth).abs().max()) <= COL_TOL
    assert torch.isfinite(col).all()


@pytest.mark.parametrize("cano,latent_dim,non_rigid_dim,sh_degree", [(False, 16, 16, 3), (True, 0, 16, 3), (True, 16, 0, 2), (True, 16, 16, 0)])
def test_color_forward_variants(built_lib, cano, latent_dim, non_rigid_dim, sh_degree):
    dev = _dev()
    P = 3000
    c = co.synthetic_case(P, seed=7, sh_degree=sh_degree, non_rigid_dim=non_rigid_dim, latent_dim=latent_dim)
    m = _module(c, dev, sh_degree=sh_degree, non_rigid_dim=non_rigid_dim, latent_dim=latent_dim, cano_view_dir=cano)
    d = lambda t: None if t is None else t.to(dev)
    col = m.colors(d(c["features"]), d(c["xyz"]), d(c["T_fwd"]) if cano else None, d(c["camera_center"]),
                   d(c["non_rigid_feature"]), d(c["latent"]), None)
    truth, _ = _truth(c, cano=cano, noise=False)
    assert float((col.double().cpu() - truth).abs().max()) <= COL_TOL


@pytest.mark.parametrize("P", [2000, 60000])  # 60 000 = 469 tiles: more than the 444 forward / 148 backward persistent CTAs
def test_color_autograd_matches_oracle_autograd(built_lib, P):
    """Forward on the tensor cores, backward = the hand-written color_backward kernel: gradients against fp64 autograd of the
    oracle (which takes the LeakyReLU branch of every hidden unit from the implementation's saved activations: see
    tests/test_gpu_nonrigid.py), <= 1e-4 of each gradient's max."""
    dev = _dev()
    c = co.synthetic_case(P, seed=9)
    m = _module(c, dev)
    d = lambda t: t.to(dev).requires_grad_(True)
    feats, xyz, T, nrf, lat = d(c["features"]), d(c["xyz"]), d(c["T_fwd"]), d(c["non_rigid_feature"]), d(c["latent"])
    col = m.colors_autograd(feats, xyz, T, c["camera_center"].to(dev), nrf, lat, None)
    g = torch.randn(P, 3, generator=torch.Generator().manual_seed(1))
    col.backward(g.to(dev))
    sv = m.saved_buffer(P, dev)
    with torch.no_grad():
        m.colors(feats, xyz, T, c["camera_center"].to(dev), nrf, lat, None, _saved=sv)
    x = sv.view(torch.bfloat16)
    tiles = (P + 127) // 128
    x = x[:tiles * 3 * 2 * 8192].view(tiles, 3, 2, 8192)  # [tile][input, lin0 act, lin1 act][hi, lo][128 x 64 operand layout]
    masks = [(x[:, 1 + l, 0].float().view(tiles, 8, 128, 8).permute(0, 2, 1, 3).reshape(tiles * 128, 64)[:P] > 0).cpu() for l in range(2)]
    leaves = {k: c[k].double().requires_grad_(True) for k in ("features", "xyz", "T_fwd", "non_rigid_feature", "latent")}
    Ws, bs = [w.double().requires_grad_(True) for w in c["Ws"]], [b.double().requires_grad_(True) for b in c["bs"]]
    ref = co.color_forward(leaves["features"], leaves["xyz"], leaves["T_fwd"], c["camera_center"].double(),
                           leaves["non_rigid_feature"], leaves["latent"], Ws, bs, 3, True, None, masks=masks)
    ref.backward(g.double())
    rel = lambda a, b: float((a.double().cpu() - b).abs().max() / b.abs().max().clamp_min(1e-30))
    assert rel(feats.grad, leaves["features"].grad) <= 1e-4
    assert rel(xyz.grad, leaves["xyz"].grad) <= 1e-4
    assert rel(T.grad.reshape(P, -1), leaves["T_fwd"].grad.reshape(P, -1)) <= 1e-4
    assert rel(nrf.grad, leaves["non_rigid_feature"].grad) <= 1e-4
    assert rel(lat.grad, leaves["latent"].grad) <= 1e-4
    for l in range(3):
        assert rel(getattr(m, f"lin{l}").weight.grad, Ws[l].grad) <= 1e-4
        assert rel(getattr(m, f"lin{l}").bias.grad, bs[l].grad) <= 1e-4


def test_color_module_forward_duck_typed(built_lib):
    """ColorMLP.forward(gaussians, camera) boundary: reads get_features / get_xyz / fwd_transform / non_rigid_feature and
    camera_center; colours go straight into the rasterizer as colors_precomp."""
    dev = _dev()
    P = 1500
    c = co.synthetic_case(P, seed=4)
    m = _module(c, dev, n_frames=4).eval()
    with torch.no_grad():
        m.latent.weight[2].copy_(c["latent"])

    class G:
        get_features = c["features"].to(dev).unsqueeze(-1)
        get_xyz = c["xyz"].to(dev)
        fwd_transform = c["T_fwd"].to(dev).reshape(P, 4, 4)
        non_rigid_feature = c["non_rigid_feature"].to(dev)

    class Cam:
        camera_center = c["camera_center"].to(dev)
        latent_idx = 2

    col = m(G, Cam)
    truth, _ = _truth(c, noise=False)
    assert float((col.double().cpu() - truth).abs().max()) <= COL_TOL
