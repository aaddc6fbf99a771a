"""Per-work-unit %globaltimer trace of blend_fwd_kernel at cfg-A (opt-in build: -DGS_FWD_TRACE into lib/libgsplat_b200_trace.so).
    python profiles/fwd_trace.py            (on a GPU box; prints where the forward blend's time goes)"""
import ctypes, os, subprocess, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "3dgs-avatar-release_b200")
sys.path[:0] = [PKG, ROOT, os.path.join(ROOT, "tests")]
lib = os.path.join(PKG, "lib", "libgsplat_b200_trace.so")
if not os.path.exists(lib):
    import glob
    subprocess.check_call(["/usr/local/cuda/bin/nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC",
                           "-shared", "-DGS_FWD_TRACE"] + sorted(glob.glob(os.path.join(PKG, "csrc", "*.cu"))) + ["-o", lib])
import torch
from diff_gaussian_rasterization import _C
_C._LIB_PATH = lib
import scenes
from test_gpu_rasterizer import run_cuda
sc = scenes.avatar_scene(P=50000, W=512, H=512, random_init=True)
for _ in range(3):
    out = run_cuda(sc)
L = _C.lib()
n = 16384 * 4
buf = (ctypes.c_uint64 * n)()
L.gs_debug_fwd_trace.argtypes = [ctypes.c_void_p, ctypes.c_int]
assert L.gs_debug_fwd_trace(buf, n) == 0
t = np.frombuffer(buf, dtype=np.uint64).reshape(-1, 4).astype(np.int64)
ranges = out["dbg"]["ranges"].view(np.uint32).reshape(-1, 2).astype(np.int64)
lens = ranges[:, 1] - ranges[:, 0]
n_items = int(sum((l + 511) // 512 for l in lens if l > 0))
u = t[:n_items * 8]
t0 = u[:, 0].min()
u = u - t0
print("units", len(u), "span_us", (u[:, 3].max()) / 1e3, "first start", u[:, 0].min() / 1e3, "last start", u[:, 0].max() / 1e3)
prod, wait, chain = u[:, 1] - u[:, 0], u[:, 2] - u[:, 1], u[:, 3] - u[:, 2]
for name, a in (("rows", prod), ("wait", wait), ("chain", chain)):
    print(f"{name:6s} mean {a.mean()/1e3:7.2f} us  p50 {np.median(a)/1e3:7.2f}  p90 {np.percentile(a,90)/1e3:7.2f}  max {a.max()/1e3:7.2f}  sum {a.sum()/1e3:9.1f}")
# deepest tile: its units are the first of every remaining level; print unit timelines of the 12 slowest-finishing units
order = np.argsort(u[:, 3])[::-1][:16]
for i in order:
    print("unit", i, "start %.1f rows_done %.1f state %.1f chain_done %.1f" % tuple(u[i] / 1e3))
json.dump({"units": int(len(u)), "span_us": float(u[:, 3].max() / 1e3), "rows_mean_us": float(prod.mean() / 1e3), "wait_mean_us": float(wait.mean() / 1e3),
           "chain_mean_us": float(chain.mean() / 1e3), "chain_max_us": float(chain.max() / 1e3)}, open(os.path.join(ROOT, "gpurun_out", "fwd_trace.json"), "w"))
