"""Generates tests/golden/ref_cuda_*.npz: outputs of the REFERENCE's own CUDA rasterizer (rebuilt for sm_100 into the
git-ignored baseline/_ref/) on the seeded scenes of tests/scenes.py.  Run on a GPU box:

    gpurun -- 'python tests/golden/make_ref_cuda_golden.py gpurun_out/golden'   # then copy the .npz files to tests/golden/
    gpurun -- 'python tests/golden/make_ref_cuda_golden.py gpurun_out/golden cfgA'   # the two bench-size scenes, compact form

The fixtures pin the CPU oracle (tests/test_oracle_golden.py, CPU-only): integer artefacts (radii, sorted keys, point
list, tile ranges, n_contrib) must match the reference build bit for bit, images <= 1e-5, gradients <= 1e-4 relative.
The reference's private byte buffers are decoded with the layout of cuda_rasterizer/rasterizer_impl.cu:155-194
(128-byte aligned bump allocation on the absolute address).
"""
import importlib.util
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (os.path.join(ROOT, "3dgs-avatar-release_b200"), ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)
import scenes  # noqa: E402


def load_ref():
    d = os.path.join(ROOT, "baseline", "_ref", "diff_gaussian_rasterization")
    spec = importlib.util.spec_from_file_location("ref_dgr", os.path.join(d, "__init__.py"), submodule_search_locations=[d])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["ref_dgr"] = mod
    spec.loader.exec_module(mod)
    return mod


class Chunk:
    def __init__(self, t):
        self.t, self.base, self.off = t, t.data_ptr(), 0

    def take(self, count, itemsize, dtype):
        addr = (self.base + self.off + 127) & ~127
        start = addr - self.base
        self.off = start + count * itemsize
        return self.t[start:start + count * itemsize].cpu().numpy().view(dtype)


def decode(geom, binning, img, P, N, HW):
    g = Chunk(geom)
    out = dict(depths=g.take(P, 4, np.float32))
    g.take(P * 3, 1, np.uint8)
    g.take(P, 4, np.int32)
    out["means2D"] = g.take(P * 2, 4, np.float32).reshape(P, 2)
    out["cov3D"] = g.take(P * 6, 4, np.float32).reshape(P, 6)
    out["conic_opacity"] = g.take(P * 4, 4, np.float32).reshape(P, 4)
    out["rgb"] = g.take(P * 3, 4, np.float32).reshape(P, 3)
    out["tiles_touched"] = g.take(P, 4, np.uint32)
    b = Chunk(binning)
    out["point_list"] = b.take(N, 4, np.uint32)
    b.take(N, 4, np.uint32)
    out["keys"] = b.take(N, 8, np.uint64)
    i = Chunk(img)
    out["final_T"] = i.take(HW, 4, np.float32)
    out["n_contrib"] = i.take(HW, 4, np.uint32)
    out["ranges"] = i.take(HW * 2, 4, np.uint32).reshape(HW, 2)
    return out


def run(ref, sc, sh_degree=3, scale_modifier=1.0):
    dev = torch.device("cuda:0")
    cam = sc["cam"]
    d = lambda a: None if a is None else torch.as_tensor(np.ascontiguousarray(a)).to(dev)
    rs = ref.GaussianRasterizationSettings(sc["H"], sc["W"], cam["tanfovx"], cam["tanfovy"], d(sc["bg"]), scale_modifier,
                                           d(cam["world_view_transform"]), d(cam["full_proj_transform"]), sh_degree,
                                           d(cam["camera_center"]), False, False)
    names = ["means3D", "opacities", "colors_precomp", "shs", "scales", "rotations", "cov3D_precomp"]
    t = {k: (d(sc.get(k)).requires_grad_(True) if sc.get(k) is not None else None) for k in names}
    means2D = torch.zeros_like(t["means3D"], requires_grad=True)
    img, radii = ref.GaussianRasterizer(rs)(means3D=t["means3D"], means2D=means2D, opacities=t["opacities"], shs=t["shs"],
                                            colors_precomp=t["colors_precomp"], scales=t["scales"], rotations=t["rotations"],
                                            cov3D_precomp=t["cov3D_precomp"])
    fn = img.grad_fn
    sv = fn.saved_tensors
    N = fn.num_rendered
    P, HW = sc["means3D"].shape[0], sc["H"] * sc["W"]
    out = decode(sv[7], sv[8], sv[9], P, N, HW)
    tiles = ((sc["W"] + 15) // 16) * ((sc["H"] + 15) // 16)
    out["ranges"] = out["ranges"][:tiles]
    out.update(N=np.int64(N), color=img.detach().cpu().numpy(), radii=radii.cpu().numpy())
    dL = scenes.l1_grad(out["color"])
    img.backward(d(dL))
    out["dL_dpix"] = dL
    for k, v in t.items():
        if v is not None and v.grad is not None:
            out["grad_" + k] = v.grad.cpu().numpy()
    out["grad_means2D"] = means2D.grad.cpu().numpy()
    return out


def main(outdir):
    os.makedirs(outdir, exist_ok=True)
    ref = load_ref()
    cases = {"avatar_3000_128": (scenes.avatar_scene(P=3000, W=128, H=128), 3, 1.0),
             "avatar_1500_200x120": (scenes.avatar_scene(P=1500, W=200, H=120), 3, 1.0),
             "generic_sh_2000_160x96": (scenes.generic_scene(), 3, 0.9)}
    for name, (sc, deg, mod) in cases.items():
        out = run(ref, sc, deg, mod)
        vis = out["radii"] > 0
        for k in ("depths", "means2D", "conic_opacity", "rgb", "cov3D"):  # unwritten entries of culled Gaussians are garbage
            out[k] = np.where(vis.reshape(-1, *([1] * (out[k].ndim - 1))), out[k], 0).astype(out[k].dtype)
        np.savez_compressed(os.path.join(outdir, f"ref_cuda_{name}.npz"), **out)
        print(name, "N", int(out["N"]), "visible", int(vis.sum()), "img mean", float(out["color"].mean()))


CFGA = {"cfgA_random_init_50k_512": lambda: scenes.avatar_scene(P=50000, W=512, H=512, random_init=True),
        "cfgA_trained_like_50k_512": lambda: scenes.avatar_scene(P=50000, W=512, H=512, random_init=False)}
CROP = (slice(192, 320), slice(192, 320))  # full-resolution window over the torso


def sha(a):
    import hashlib
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def compact(out, H, W):
    """The bench-size fixture (BASELINE.json configs[1]/[2]: 50 000 Gaussians, 512x512) in a few hundred KB: SHA-256 of the
    integer artefacts (bit-exactness needs no more), the contributor counts themselves (uint16, compresses well), a full
    resolution window + 16x16 block means of the image and of final_T, and per-tensor sums + the first rows of every gradient."""
    nc = out["n_contrib"].reshape(H, W)
    assert int(nc.max()) < 65536
    col, fT = out["color"], out["final_T"].reshape(H, W)
    c = dict(N=out["N"], sha_keys=sha(out["keys"]), sha_point_list=sha(out["point_list"]), sha_ranges=sha(out["ranges"]),
             sha_radii=sha(out["radii"].astype(np.int32)), sha_tiles_touched=sha(out["tiles_touched"]), sha_n_contrib=sha(nc.astype(np.uint32)),
             n_contrib=nc.astype(np.uint16), color_crop=col[:, CROP[0], CROP[1]].copy(), final_T_crop=fT[CROP[0], CROP[1]].copy(),
             color_block_mean=col.astype(np.float64).reshape(3, H // 16, 16, W // 16, 16).mean((2, 4)),
             final_T_block_mean=fT.astype(np.float64).reshape(H // 16, 16, W // 16, 16).mean((1, 3)))
    for k, v in out.items():
        if k.startswith("grad_"):
            v64 = v.astype(np.float64)
            c[k + "_sum"], c[k + "_abs_sum"], c[k + "_max_abs"] = v64.sum(), np.abs(v64).sum(), np.abs(v64).max()
            c[k + "_head"] = v.reshape(v.shape[0], -1)[:2048].copy()
    return c


def main_cfgA(outdir):
    os.makedirs(outdir, exist_ok=True)
    ref = load_ref()
    for name, mk in CFGA.items():
        sc = mk()
        out = run(ref, sc, 3, 1.0)
        np.savez_compressed(os.path.join(outdir, f"ref_cuda_{name}.npz"), **compact(out, sc["H"], sc["W"]))
        print(name, "N", int(out["N"]), "E_fwd", int(out["n_contrib"].astype(np.int64).sum()))


if __name__ == "__main__":
    d = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden")
    if len(sys.argv) > 2 and sys.argv[2] == "cfgA":
        main_cfgA(d)
    else:
        main(d)
