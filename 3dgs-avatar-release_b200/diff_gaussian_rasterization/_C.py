"""`_C`: the extension module behind `diff_gaussian_rasterization` -- here a ctypes binding of the C-ABI library
libgsplat_b200.so (include/gsplat_b200.h) instead of the reference's pybind11 module.

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


_STATUS = {-1: "bad argument", -2: "workspace too small", -3: "CUDA error", -4: "For non-RGB, provide precomputed Gaussian colors!"}


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


class _DeviceState:
    """Per-device scratch cache: the sort scratch (not needed by backward), the pair-capacity high-water mark and the
    pinned word the pair count is copied into."""

    def __init__(self):
        self.scratch = None
        self.graph_keepalive = []
        self.capacity = 0
        self.pinned = torch.empty(2, dtype=torch.int64).pin_memory()
        self.pinned_np = self.pinned.numpy()


_state = {}
stats = {"forward_calls": 0, "backward_calls": 0, "capacity_regrows": 0}
_GROW = 1.25


def _dev_state(device):
    key = device.index if device.index is not None else torch.cuda.current_device()
    st = _state.get(key)
    if st is None:
        st = _state[key] = _DeviceState()
    return st


def rasterize_gaussians(bg, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
                        projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos, prefiltered,
                        debug):
    return _forward(bg, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
                    projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos, prefiltered, debug, None)


def rasterize_gaussians_opacity(bg, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                                viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree,
                                campos, prefiltered, debug):
    """rasterize_gaussians + the opacity image of the reference's second rasterizer call (colours = 1, channel 0 kept:
    gaussian_renderer/__init__.py:131-142) from the same forward blend.
    -> (num_rendered, color, opacity_image [1,H,W], radii, geomBuffer, binningBuffer, imgBuffer)"""
    H, W = int(image_height), int(image_width)
    op_img = torch.zeros(1, H, W, dtype=torch.float32, device=means3D.device)
    n, color, radii, geom, lst, img = _forward(bg, means3D, colors, opacity, scales, rotations, scale_modifier,
                                               cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, H, W, sh,
                                               degree, campos, prefiltered, debug, op_img)
    return n, color, op_img, radii, geom, lst, img


def _forward(bg, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
             projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos, prefiltered, debug,
             out_opacity):
    if means3D.dim() != 2 or means3D.size(1) != 3:
        raise GsError("means3D must have dimensions (num_points, 3)")  # rasterize_points.cu:57-59
    if not means3D.is_cuda:
        raise GsError("means3D must be a CUDA tensor (there is no CPU path)")
    L = lib()
    P, H, W = int(means3D.size(0)), int(image_height), int(image_width)
    dev = means3D.device
    with torch.cuda.device(dev):
        u8 = dict(dtype=torch.uint8, device=dev)
        if P == 0:  # rasterize_points.cu:81
            z = torch.empty(0, **u8)
            return 0, torch.zeros(3, H, W, dtype=torch.float32, device=dev), torch.zeros(0, dtype=torch.int32, device=dev), z, z.clone(), z.clone()
        means3D, colors, opacity, scales, rotations, cov3D_precomp, sh = map(_c, (means3D, colors, opacity, scales, rotations, cov3D_precomp, sh))
        bg, viewmatrix, projmatrix, campos = map(_c, (bg, viewmatrix, projmatrix, campos))
        out_color = torch.empty(3, H, W, dtype=torch.float32, device=dev)
        radii = torch.empty(P, dtype=torch.int32, device=dev)
        M = int(sh.size(1)) if (sh is not None and sh.numel() != 0) else 0
        a = GsForwardArgs(P, int(degree), M, W, H, float(tan_fovx), float(tan_fovy), float(scale_modifier),
                          int(bool(prefiltered)), int(bool(debug)), _ptr(bg, "bg"), _ptr(means3D, "means3D"),
                          _ptr(sh, "sh"), _ptr(colors, "colors_precomp"), _ptr(opacity, "opacities"),
                          _ptr(scales, "scales"), _ptr(rotations, "rotations"), _ptr(cov3D_precomp, "cov3D_precomp"),
                          _ptr(viewmatrix, "viewmatrix"), _ptr(projmatrix, "projmatrix"), _ptr(campos, "campos"),
                          out_color.data_ptr(), radii.data_ptr(),
                          out_opacity.data_ptr() if out_opacity is not None else None)
        st = _dev_state(dev)
        stream = torch.cuda.current_stream(dev).cuda_stream
        geom = torch.empty(L.gs_geom_bytes(P), **u8)
        img = torch.empty(L.gs_image_bytes(W, H), **u8)
        if st.capacity == 0:
            st.capacity = max(4 * P, 1 << 18)
        while True:
            # capacities live on the 4096 lattice so that backward can recover them from the buffer size
            cap = st.capacity = (st.capacity + 4095) // 4096 * 4096
            lst = torch.empty(L.gs_list_bytes(cap, W, H), **u8)
            need = L.gs_scratch_bytes(P, W, H, cap)
            if st.scratch is None or st.scratch.numel() < need or st.scratch.device != dev:
                st.scratch = torch.empty(int(need * 1.1) + 1024, **u8)
            st.pinned_np[:] = -1
            status = L.gs_forward(ctypes.byref(a), geom.data_ptr(), geom.numel(), img.data_ptr(), img.numel(),
                                  lst.data_ptr(), lst.numel(), st.scratch.data_ptr(), st.scratch.numel(), cap,
                                  st.pinned.data_ptr(), stream)
            _check(status, "gs_forward")
            stats["forward_calls"] += 1
            if torch.cuda.is_current_stream_capturing():
                # CUDA-graph capture: no host round trip.  The pair count lands in pinned memory on every replay;
                # the caller checks it with graph_overflow() after a replay (capacity is fixed inside a graph).
                st.graph_capacity = cap
                st.graph_keepalive.append(st.scratch)  # a later regrow must not free what the graph points at
                return -1, out_color, radii, geom, lst, img
            # The pair count N is produced on the device; the async copy into pinned memory lands while the blend
            # kernel is still running, so this wait overlaps GPU work (the reference blocks the pipeline instead,
            # rasterizer_impl.cu:281).
            t0 = None
            pn = st.pinned_np
            while pn[0] == -1 or pn[1] == -1:
                if t0 is None:
                    t0 = time.perf_counter()
                elif time.perf_counter() - t0 > 30.0:
                    torch.cuda.synchronize(dev)
                    if pn[0] == -1:
                        raise GsError("gs_forward: the pair count never arrived (kernel failure?)")
            N, flags = int(pn[0]), int(pn[1])
            if (flags >> 32) != 0:
                raise GsError(f"gs_forward: internal look-back timeout (error flag {flags >> 32})")
            if N <= cap:
                break
            st.capacity = int(N * _GROW) + 4096  # overflow: the image is invalid -> regrow and rerun
            stats["capacity_regrows"] += 1
        return N, out_color, radii, geom, lst, img


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
    for st in _state.values():
        st.scratch = None
        st.graph_keepalive.clear()
    _state.clear()


def graph_overflow(device=None):
    """After replaying a CUDA graph that contains rasterize_gaussians (and synchronising): returns (N, overflowed).
    overflowed means the captured pair capacity was too small, the frame is invalid, and the graph must be re-captured
    (the capacity high-water mark has already been raised)."""
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    st = _dev_state(dev)
    N, flags = int(st.pinned_np[0]), int(st.pinned_np[1])
    if (flags >> 32) != 0:
        raise GsError(f"gs_forward: internal look-back timeout (error flag {flags >> 32})")
    over = N > getattr(st, "graph_capacity", st.capacity)
    if over:
        st.capacity = int(N * _GROW) + 4096
        stats["capacity_regrows"] += 1
    return N, over


def rasterize_gaussians_backward(bg, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp,
                                 viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, sh, degree, campos,
                                 geomBuffer, R, binningBuffer, imageBuffer, debug, dL_dout_opacity=None):
    """dL_dout_opacity ([1,H,W], only after rasterize_gaussians_opacity): upstream gradient of the fused opacity image;
    the returned geometry gradients are then the sum of the reference's two backward calls."""
    L = lib()
    P = int(means3D.size(0))
    H, W = int(dL_dout_color.size(1)), int(dL_dout_color.size(2))
    dev = means3D.device
    M = int(sh.size(1)) if (sh is not None and sh.numel() != 0) else 0
    with torch.cuda.device(dev):
        # one slab carved into the 8 gradient tensors of rasterize_points.cu:151-159 (dL_dconic stays inside the library).  NOT
        # zero-filled: gs_backward writes every element of every output (culled Gaussians get explicit zeros)
        sizes = [3, 3, 3, 1, 6, 3 * M, 3, 4]  # means3D, means2D, colors, opacity, cov3D, sh, scales, rot
        pad4 = lambda n: (n + 3) // 4 * 4  # every part 16-byte aligned for any P
        flat = torch.empty(sum(pad4(P * s) for s in sizes), dtype=torch.float32, device=dev)
        parts, off = [], 0
        for s in sizes:
            parts.append(flat[off:off + P * s])
            off += pad4(P * s)
        dmeans3D, dmeans2D, dcolors, dopacity, dcov3D, dsh, dscales, drot = parts
        dmeans3D, dmeans2D, dcolors = dmeans3D.view(P, 3), dmeans2D.view(P, 3), dcolors.view(P, 3)
        dopacity, dcov3D = dopacity.view(P, 1), dcov3D.view(P, 6)
        dsh, dscales, drot = dsh.view(P, M, 3), dscales.view(P, 3), drot.view(P, 4)
        if P == 0:
            return dmeans2D, dcolors, dopacity, dmeans3D, dcov3D, dsh, dscales, drot
        means3D, colors, scales, rotations, cov3D_precomp, sh = map(_c, (means3D, colors, scales, rotations, cov3D_precomp, sh))
        bg, viewmatrix, projmatrix, campos, dL = map(_c, (bg, viewmatrix, projmatrix, campos, dL_dout_color))
        dLo = _c(dL_dout_opacity)
        a = GsBackwardArgs(P, int(degree), M, W, H, float(tan_fovx), float(tan_fovy), float(scale_modifier),
                           int(bool(debug)), _ptr(bg, "bg"), _ptr(means3D, "means3D"), _ptr(sh, "sh"),
                           _ptr(colors, "colors_precomp"), _ptr(scales, "scales"), _ptr(rotations, "rotations"),
                           _ptr(cov3D_precomp, "cov3D_precomp"), _ptr(viewmatrix, "viewmatrix"),
                           _ptr(projmatrix, "projmatrix"), _ptr(campos, "campos"), _ptr(radii, "radii", torch.int32),
                           _ptr(dL, "dL_dout_color"), dmeans2D.data_ptr(), None, dopacity.data_ptr(),
                           dcolors.data_ptr(), dmeans3D.data_ptr(), dcov3D.data_ptr(),
                           dsh.data_ptr() if M else None, dscales.data_ptr(), drot.data_ptr(),
                           _ptr(dLo, "dL_dout_opacity"))
        stream = torch.cuda.current_stream(dev).cuda_stream
        cap = L.gs_list_capacity(binningBuffer.numel(), W, H)
        status = L.gs_backward(ctypes.byref(a), geomBuffer.data_ptr(), geomBuffer.numel(), imageBuffer.data_ptr(),
                               imageBuffer.numel(), binningBuffer.data_ptr(), binningBuffer.numel(), cap, stream)
        _check(status, "gs_backward")
        stats["backward_calls"] += 1
    return dmeans2D, dcolors, dopacity, dmeans3D, dcov3D, dsh, dscales, drot


def mark_visible(means3D, viewmatrix, projmatrix):
    L = lib()
    P = int(means3D.size(0))
    dev = means3D.device
    if not means3D.is_cuda:
        raise GsError("means3D must be a CUDA tensor (there is no CPU path)")
    with torch.cuda.device(dev):
        present = torch.zeros(P, dtype=torch.bool, device=dev)
        if P:
            means3D, viewmatrix, projmatrix = map(_c, (means3D, viewmatrix, projmatrix))
            status = L.gs_mark_visible(P, _ptr(means3D, "means3D"), _ptr(viewmatrix, "viewmatrix"),
                                       _ptr(projmatrix, "projmatrix"), present.data_ptr(),
                                       torch.cuda.current_stream(dev).cuda_stream)
            _check(status, "gs_mark_visible")
    return present


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
