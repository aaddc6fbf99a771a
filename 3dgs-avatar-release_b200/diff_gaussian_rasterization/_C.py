"""`_C`: the extension module behind `diff_gaussian_rasterization`.  The three entry points are a thin PyTorch C++ extension
(`_C_ext`, csrc/ext/ext.cpp: pybind11 like the reference's module) over the C-ABI library libgsplat_b200.so
(include/gsplat_b200.h); the debug / bench accessors below bind the same library with ctypes.

It exports the reference's three entry points with the reference's signatures
(submodules/diff-gaussian-rasterization/ext.cpp:15-19, rasterize_points.h:18-67):

    rasterize_gaussians(bg, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
                        projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos, prefiltered,
                        debug) -> (num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer)
    rasterize_gaussians_backward(bg, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp,
                        viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, sh, degree, campos, geomBuffer, R,
                        binningBuffer, imageBuffer, debug) -> (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D,
                        dL_dcov3D, dL_dsh, dL_dscales, dL_drotations)
    mark_visible(means3D, viewmatrix, projmatrix) -> bool[P]

so the reference's own `diff_gaussian_rasterization/__init__.py` could drive it unchanged.  torch is used for device
memory and the current stream only.  There is NO CPU fallback: a missing library or a non-CUDA tensor raises.
"""
import ctypes
import os
import time

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.environ.get("GS_LIB") or os.path.join(os.path.dirname(_HERE), "lib", "libgsplat_b200.so")  # GS_LIB: experiment builds

c_float_p = ctypes.c_void_p


class GsForwardArgs(ctypes.Structure):
    _fields_ = [("P", ctypes.c_int32), ("D", ctypes.c_int32), ("M", ctypes.c_int32), ("width", ctypes.c_int32),
                ("height", ctypes.c_int32), ("tan_fovx", ctypes.c_float), ("tan_fovy", ctypes.c_float),
                ("scale_modifier", ctypes.c_float), ("prefiltered", ctypes.c_int32), ("debug", ctypes.c_int32),
                ("background", c_float_p), ("means3D", c_float_p), ("shs", c_float_p), ("colors_precomp", c_float_p),
                ("opacities", c_float_p), ("scales", c_float_p), ("rotations", c_float_p),
                ("cov3D_precomp", c_float_p), ("viewmatrix", c_float_p), ("projmatrix", c_float_p),
                ("campos", c_float_p), ("out_color", c_float_p), ("radii", c_float_p), ("out_opacity", c_float_p)]


class GsBackwardArgs(ctypes.Structure):
    _fields_ = [("P", ctypes.c_int32), ("D", ctypes.c_int32), ("M", ctypes.c_int32), ("width", ctypes.c_int32),
                ("height", ctypes.c_int32), ("tan_fovx", ctypes.c_float), ("tan_fovy", ctypes.c_float),
                ("scale_modifier", ctypes.c_float), ("debug", ctypes.c_int32),
                ("background", c_float_p), ("means3D", c_float_p), ("shs", c_float_p), ("colors_precomp", c_float_p),
                ("scales", c_float_p), ("rotations", c_float_p), ("cov3D_precomp", c_float_p),
                ("viewmatrix", c_float_p), ("projmatrix", c_float_p), ("campos", c_float_p), ("radii", c_float_p),
                ("dL_dpix", c_float_p), ("dL_dmean2D", c_float_p), ("dL_dconic", c_float_p),
                ("dL_dopacity", c_float_p), ("dL_dcolor", c_float_p), ("dL_dmean3D", c_float_p),
                ("dL_dcov3D", c_float_p), ("dL_dsh", c_float_p), ("dL_dscale", c_float_p), ("dL_drot", c_float_p),
                ("dL_dopacity_pix", c_float_p)]


_lib = None


def lib():
    """Loads libgsplat_b200.so; raises (never falls back) when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            raise ImportError(
                f"{_LIB_PATH} is missing: build it with `python __graft_entry__.py build` "
                "(nvcc -gencode arch=compute_100a,code=sm_100a). There is no CPU fallback.")
        L = ctypes.CDLL(_LIB_PATH)
        sz = ctypes.c_size_t
        L.gs_geom_bytes.restype = sz
        L.gs_geom_bytes.argtypes = [ctypes.c_int32]
        L.gs_image_bytes.restype = sz
        L.gs_image_bytes.argtypes = [ctypes.c_int32, ctypes.c_int32]
        L.gs_list_bytes.restype = sz
        L.gs_list_bytes.argtypes = [ctypes.c_int64, ctypes.c_int32, ctypes.c_int32]
        L.gs_list_capacity.restype = ctypes.c_int64
        L.gs_list_capacity.argtypes = [sz, ctypes.c_int32, ctypes.c_int32]
        L.gs_scratch_bytes.restype = sz
        L.gs_scratch_bytes.argtypes = [ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_int64]
        vp = ctypes.c_void_p
        L.gs_forward.restype = ctypes.c_int
        L.gs_forward.argtypes = [ctypes.POINTER(GsForwardArgs), vp, sz, vp, sz, vp, sz, vp, sz, ctypes.c_int64, vp, vp]
        L.gs_backward.restype = ctypes.c_int
        L.gs_backward.argtypes = [ctypes.POINTER(GsBackwardArgs), vp, sz, vp, sz, vp, sz, ctypes.c_int64, vp]
        L.gs_mark_visible.restype = ctypes.c_int
        L.gs_mark_visible.argtypes = [ctypes.c_int32, vp, vp, vp, vp, vp]
        L.gs_read_num_rendered.restype = ctypes.c_int
        L.gs_read_num_rendered.argtypes = [vp, vp, vp]
        L.gs_debug_read.restype = ctypes.c_int
        L.gs_debug_read.argtypes = [ctypes.c_int32, ctypes.POINTER(GsForwardArgs), vp, vp, vp, ctypes.c_int64, vp, sz, vp]
        L.gs_set_exp_mode.restype = ctypes.c_int
        L.gs_set_exp_mode.argtypes = [ctypes.c_int]
        L.gs_read_counters.restype = ctypes.c_int
        L.gs_read_counters.argtypes = [vp, vp, vp]
        L.gs_last_cuda_error.restype = ctypes.c_int
        L.gs_version.restype = ctypes.c_char_p
        _lib = L
    return _lib


class GsError(RuntimeError):
    pass


_STATUS = {-1: "bad argument", -2: "workspace too small", -3: "CUDA error", -4: "For non-RGB, provide precomputed Gaussian colors!",
           -5: "size exceeds the limit of the built-in static scratch"}


def _check(status, what):
    if status != 0:
        extra = f" (cudaError {lib().gs_last_cuda_error()})" if status == -3 else ""
        raise GsError(f"{what}: {_STATUS.get(status, status)}{extra}")


def _ptr(t, name, dtype=torch.float32):
    """Device pointer of a tensor, or None for the empty placeholder tensors (`torch.Tensor([])`) with which the
    Python layer marks absent arguments (reference: __init__.py:197-207; null pointer in C++)."""
    if t is None or t.numel() == 0:
        return None
    if not t.is_cuda:
        raise GsError(f"{name} must be a CUDA tensor (there is no CPU path)")
    if t.dtype != dtype:
        raise GsError(f"{name} must be {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise GsError(f"{name} must be contiguous")
    return t.data_ptr()


def _c(t):
    return t if (t is None or t.is_contiguous()) else t.contiguous()


# ---- the three entry points: the thin PyTorch C++ extension (csrc/ext/ext.cpp -> _C_ext*.so, built by build.py) ----
try:
    from . import _C_ext as _ext
except ImportError as _ex:  # never falls back to anything else
    raise ImportError(f"diff_gaussian_rasterization._C_ext is missing ({_ex}): build it with `python __graft_entry__.py build` "
                      "(g++ against torch's headers, linked to lib/libgsplat_b200.so). There is no CPU fallback.") from None


def _call(fn, *args, **kw):
    try:
        return fn(*args, **kw)
    except RuntimeError as ex:  # std::runtime_error / TORCH_CHECK of the extension
        raise GsError(str(ex)) from None


def rasterize_gaussians(bg, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
                        projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos, prefiltered, debug):
    return _call(_ext.rasterize_gaussians, bg, means3D, colors, opacity, scales, rotations, float(scale_modifier), cov3D_precomp,
                 viewmatrix, projmatrix, float(tan_fovx), float(tan_fovy), int(image_height), int(image_width), sh, int(degree),
                 campos, bool(prefiltered), bool(debug))


def rasterize_gaussians_opacity(bg, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                                viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree,
                                campos, prefiltered, debug):
    """rasterize_gaussians + the opacity image of the reference's second rasterizer call (colours = 1, channel 0 kept:
    gaussian_renderer/__init__.py:131-142) from the same forward blend.
    -> (num_rendered, color, opacity_image [1,H,W], radii, geomBuffer, binningBuffer, imgBuffer)"""
    return _call(_ext.rasterize_gaussians_opacity, bg, means3D, colors, opacity, scales, rotations, float(scale_modifier),
                 cov3D_precomp, viewmatrix, projmatrix, float(tan_fovx), float(tan_fovy), int(image_height), int(image_width), sh,
                 int(degree), campos, bool(prefiltered), bool(debug))


def rasterize_gaussians_backward(bg, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp,
                                 viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, sh, degree, campos,
                                 geomBuffer, R, binningBuffer, imageBuffer, debug, dL_dout_opacity=None):
    """dL_dout_opacity ([1,H,W], only after rasterize_gaussians_opacity): upstream gradient of the fused opacity image;
    the returned geometry gradients are then the sum of the reference's two backward calls."""
    return _call(_ext.rasterize_gaussians_backward, bg, means3D, radii, colors, scales, rotations, float(scale_modifier),
                 cov3D_precomp, viewmatrix, projmatrix, float(tan_fovx), float(tan_fovy), dL_dout_color, sh, int(degree), campos,
                 geomBuffer, int(R), binningBuffer, imageBuffer, bool(debug), dL_dout_opacity)


def mark_visible(means3D, viewmatrix, projmatrix):
    return _call(_ext.mark_visible, means3D, viewmatrix, projmatrix)


class _DevStateProxy:
    """Pair-capacity high-water mark of a device (kept inside the extension)."""

    def __init__(self, index):
        self.index = index

    capacity = property(lambda s: _ext.get_capacity(s.index), lambda s, v: _ext.set_capacity(s.index, int(v)))


def _dev_index(device):
    return device.index if (device is not None and device.index is not None) else torch.cuda.current_device()


def _dev_state(device):
    return _DevStateProxy(_dev_index(device))


class _Stats:
    """forward_calls / backward_calls / capacity_regrows counters of the extension (read-only mapping)."""
    _keys = ("forward_calls", "backward_calls", "capacity_regrows")

    def __getitem__(self, k):
        return _ext.counters()[self._keys.index(k)]

    def keys(self):
        return self._keys


stats = _Stats()


def set_exp_mode(mode):
    """0: CUDA expf = the reference's instruction sequence (default); 1: the polynomial the CPU oracle reproduces bit for bit
    (include/gsplat_b200.h: gs_set_exp_mode).  Returns the previous mode."""
    return int(lib().gs_set_exp_mode(int(mode)))


def read_counters(geomBuffer):
    """(E_bwd, forward work items) of the forward call that filled `geomBuffer` (synchronises the current stream)."""
    buf = (ctypes.c_uint64 * 2)()
    with torch.cuda.device(geomBuffer.device):
        _check(lib().gs_read_counters(geomBuffer.data_ptr(), ctypes.cast(buf, ctypes.c_void_p),
                                      torch.cuda.current_stream(geomBuffer.device).cuda_stream), "gs_read_counters")
    return int(buf[0]), int(buf[1])


def release_cached():
    """Drops the per-device scratch caches (sort scratch, graph keep-alives, pinned words).  For an orderly shutdown: call
    after the CUDA graphs that captured rasterize_gaussians have been released."""
    _ext.release_cached()


def graph_overflow(device=None):
    """After replaying a CUDA graph that contains rasterize_gaussians (and synchronising): returns (N, overflowed).
    overflowed means the captured pair capacity was too small, the frame is invalid, and the graph must be re-captured
    (the capacity high-water mark has already been raised)."""
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    return _call(_ext.graph_overflow, _dev_index(dev))


_DEBUG_SPECS = {  # which -> (name, dtype, elems as fn(P, W, H, N))
    0: ("ranges", torch.int32, lambda P, W, H, N: ((W + 15) // 16) * ((H + 15) // 16) * 2),
    1: ("n_contrib", torch.int32, lambda P, W, H, N: W * H),
    2: ("final_T", torch.float32, lambda P, W, H, N: W * H),
    3: ("point_list", torch.int32, lambda P, W, H, N: N),
    4: ("keys", torch.int64, lambda P, W, H, N: N),
    5: ("means2D", torch.float32, lambda P, W, H, N: 2 * P),
    6: ("depths", torch.float32, lambda P, W, H, N: P),
    7: ("conic_opacity", torch.float32, lambda P, W, H, N: 4 * P),
    8: ("tiles_touched", torch.int32, lambda P, W, H, N: P),
    9: ("rgb", torch.float32, lambda P, W, H, N: 3 * P),
    10: ("cov3D", torch.float32, lambda P, W, H, N: 6 * P),
    11: ("clamped", torch.uint8, lambda P, W, H, N: 3 * P),
}


def debug_state(P, W, H, N, radii, geomBuffer, binningBuffer, imgBuffer, names=None):
    """Copies the private workspace arrays the bit-exact parity tests compare into fresh tensors (dict by name)."""
    L = lib()
    dev = geomBuffer.device
    out = {}
    a = GsForwardArgs()
    a.P, a.width, a.height = P, W, H
    a.radii = radii.data_ptr()
    stream = torch.cuda.current_stream(dev).cuda_stream
    for which, (name, dtype, n) in _DEBUG_SPECS.items():
        if names is not None and name not in names:
            continue
        cnt = int(n(P, W, H, N))
        t = torch.zeros(max(cnt, 1), dtype=dtype, device=dev)
        status = L.gs_debug_read(which, ctypes.byref(a), geomBuffer.data_ptr(), imgBuffer.data_ptr(),
                                 binningBuffer.data_ptr(), N, t.data_ptr(), t.numel() * t.element_size(), stream)
        _check(status, f"gs_debug_read({name})")
        out[name] = t[:cnt]
    return out
