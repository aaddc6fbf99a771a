"""numpy/ctypes front-end of the C rasterizer oracle (oracle/gs_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's CPU legs.
The product package never imports this module.

The call structure mirrors CudaRasterizer::Rasterizer::forward / backward
(submodules/diff-gaussian-rasterization/cuda_rasterizer/rasterizer_impl.cu:198-336, 340-434).
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force=False):
    so = os.path.join(_HERE, "_build", "libgs_oracle.so")
    src = os.path.join(_HERE, "gs_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = ctypes.CDLL(build())
        _LIB.gso_bin.restype = ctypes.c_int64
        _LIB.gso_exp.restype = ctypes.c_float
        _LIB.gso_exp.argtypes = [ctypes.c_float]
        _LIB.gso_higher_msb.restype = ctypes.c_uint32
    return _LIB


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def _f32(a):
    if a is None:
        return None
    a = np.ascontiguousarray(np.asarray(a, dtype=np.float32))
    return a if a.size else None


def exp(x):
    return float(lib().gso_exp(ctypes.c_float(float(x))))


def set_num_threads(n):
    lib().gso_set_num_threads(int(n))


def num_threads():
    return int(lib().gso_num_threads())


def forward(means3D, opacities, W, H, tanfovx, tanfovy, view, proj, campos, bg, colors_precomp=None,
            shs=None, sh_degree=0, scales=None, rotations=None, scale_modifier=1.0, cov3D_precomp=None):
    """Returns a dict with every intermediate the parity tests compare (all numpy)."""
    L = lib()
    means3D = _f32(means3D).reshape(-1, 3)
    P = means3D.shape[0]
    opacities, view, proj, campos, bg = _f32(opacities), _f32(view), _f32(proj), _f32(campos), _f32(bg)
    colors_precomp, shs, scales, rotations, cov3D_precomp = map(_f32, (colors_precomp, shs, scales, rotations, cov3D_precomp))
    M = 0 if shs is None else shs.shape[1]
    radii = np.zeros(P, np.int32)
    means2D = np.zeros((P, 2), np.float32)
    depths = np.zeros(P, np.float32)
    cov3D = np.zeros((P, 6), np.float32)
    rgb = np.zeros((P, 3), np.float32)
    conic_opacity = np.zeros((P, 4), np.float32)
    tiles_touched = np.zeros(P, np.uint32)
    clamped = np.zeros((P, 3), np.uint8)
    f = ctypes.c_float
    L.gso_preprocess(P, int(sh_degree), M, _p(means3D), _p(scales), f(scale_modifier), _p(rotations), _p(opacities),
                     _p(shs), _p(cov3D_precomp), _p(colors_precomp), _p(view), _p(proj), _p(campos), int(W), int(H),
                     f(tanfovx), f(tanfovy), _p(radii), _p(means2D), _p(depths), _p(cov3D), _p(rgb),
                     _p(conic_opacity), _p(tiles_touched), _p(clamped))
    N = int(tiles_touched.astype(np.int64).sum())
    gx, gy = (W + 15) // 16, (H + 15) // 16
    keys_unsorted = np.zeros(max(N, 1), np.uint64)
    keys = np.zeros(max(N, 1), np.uint64)
    point_list = np.zeros(max(N, 1), np.uint32)
    ranges = np.zeros((gx * gy, 2), np.uint32)
    N2 = L.gso_bin(P, _p(means2D), _p(depths), _p(radii), int(W), int(H), ctypes.c_int64(N), _p(keys_unsorted),
                   _p(keys), _p(point_list), _p(ranges))
    assert N2 == N, (N2, N)
    feats = colors_precomp if colors_precomp is not None else rgb
    final_T = np.zeros(H * W, np.float32)
    n_contrib = np.zeros(H * W, np.uint32)
    out = np.zeros((3, H, W), np.float32)
    evals = np.zeros(2, np.uint64)
    L.gso_render(int(W), int(H), _p(ranges), _p(point_list), _p(means2D), _p(feats), _p(conic_opacity), _p(bg),
                 _p(final_T), _p(n_contrib), _p(out), _p(evals))
    return dict(P=P, M=M, W=W, H=H, N=N, color=out, radii=radii, means2D=means2D, depths=depths, cov3D=cov3D, rgb=rgb,
                conic_opacity=conic_opacity, tiles_touched=tiles_touched, clamped=clamped,
                keys_unsorted=keys_unsorted[:N], keys=keys[:N], point_list=point_list[:N], ranges=ranges,
                final_T=final_T.reshape(H, W), n_contrib=n_contrib.reshape(H, W), evals_visited=int(evals[0]),
                evals_blended=int(evals[1]),
                inputs=dict(means3D=means3D, opacities=opacities, view=view, proj=proj, campos=campos, bg=bg,
                            colors_precomp=colors_precomp, shs=shs, sh_degree=int(sh_degree), scales=scales,
                            rotations=rotations, scale_modifier=float(scale_modifier), cov3D_precomp=cov3D_precomp,
                            tanfovx=float(tanfovx), tanfovy=float(tanfovy)))


def backward(fw, dL_dpix):
    """fw = dict returned by forward(); dL_dpix [3,H,W].  Returns the 8 gradients of
    rasterize_points.cu:117-196 (plus dL_dconic)."""
    L = lib()
    i = fw["inputs"]
    P, M, W, H = fw["P"], fw["M"], fw["W"], fw["H"]
    dL_dpix = _f32(dL_dpix)
    feats = i["colors_precomp"] if i["colors_precomp"] is not None else fw["rgb"]
    dmean2D = np.zeros((P, 3), np.float32)
    dconic = np.zeros((P, 4), np.float32)
    dopac = np.zeros((P, 1), np.float32)
    dcolors = np.zeros((P, 3), np.float32)
    L.gso_render_backward(P, int(W), int(H), _p(fw["ranges"]), _p(fw["point_list"] if fw["N"] else np.zeros(1, np.uint32)),
                          _p(i["bg"]), _p(fw["means2D"]), _p(fw["conic_opacity"]), _p(feats),
                          _p(np.ascontiguousarray(fw["final_T"])), _p(np.ascontiguousarray(fw["n_contrib"])), _p(dL_dpix),
                          _p(dmean2D), _p(dconic), _p(dopac), _p(dcolors))
    dmean3D = np.zeros((P, 3), np.float32)
    dcov3D = np.zeros((P, 6), np.float32)
    dsh = np.zeros((P, M, 3), np.float32)
    dscale = np.zeros((P, 3), np.float32)
    drot = np.zeros((P, 4), np.float32)
    cov = i["cov3D_precomp"] if i["cov3D_precomp"] is not None else fw["cov3D"]
    f = ctypes.c_float
    L.gso_preprocess_backward(P, i["sh_degree"], M, _p(i["means3D"]), _p(fw["radii"]), _p(i["shs"]), _p(fw["clamped"]),
                              _p(i["scales"]), _p(i["rotations"]), f(i["scale_modifier"]), _p(cov), _p(i["view"]),
                              _p(i["proj"]), _p(i["campos"]), int(W), int(H), f(i["tanfovx"]), f(i["tanfovy"]),
                              _p(dmean2D), _p(dconic), _p(dmean3D), _p(dcolors), _p(dcov3D), _p(dsh), _p(dscale), _p(drot))
    return dict(dL_dmeans2D=dmean2D, dL_dcolors=dcolors, dL_dopacity=dopac, dL_dmeans3D=dmean3D, dL_dcov3D=dcov3D,
                dL_dsh=dsh, dL_dscales=dscale, dL_drotations=drot, dL_dconic=dconic)


def mark_visible(means3D, view, proj):
    means3D = _f32(means3D).reshape(-1, 3)
    out = np.zeros(means3D.shape[0], np.uint8)
    lib().gso_mark_visible(means3D.shape[0], _p(means3D), _p(_f32(view)), _p(_f32(proj)), _p(out))
    return out.astype(bool)
