"""Critical path of blend_fwd_kernel's deepest tile at cfg-A from the per-unit %globaltimer trace (-DGS_FWD_TRACE build of the
library, built here INTO lib/libgsplat_b200.so: run on a scratch GPU box only).
    python profiles/fwd_critical_path.py"""
import ctypes, glob, json, os, subprocess, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "3dgs-avatar-release_b200")
sys.path[:0] = [PKG, ROOT, os.path.join(ROOT, "tests")]
lib = os.path.join(PKG, "lib", "libgsplat_b200.so")
subprocess.check_call(["/usr/local/cuda/bin/nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC",
                       "-shared", "-DGS_FWD_TRACE"] + sorted(glob.glob(os.path.join(PKG, "csrc", "*.cu"))) + ["-o", lib])
import torch
from diff_gaussian_rasterization import _C
import scenes
from test_gpu_rasterizer import run_cuda
CH = 256
sc = scenes.avatar_scene(P=50000, W=512, H=512, random_init=True)
for _ in range(4):
    out = run_cuda(sc)
L = _C.lib()
n = 16384 * 4
buf = (ctypes.c_uint64 * n)()
L.gs_debug_fwd_trace.argtypes = [ctypes.c_void_p, ctypes.c_int]
assert L.gs_debug_fwd_trace(buf, n) == 0
t = np.frombuffer(buf, dtype=np.uint64).reshape(-1, 4).astype(np.int64)
ranges = out["dbg"]["ranges"].view(np.uint32).reshape(-1, 2).astype(np.int64)
lens = ranges[:, 1] - ranges[:, 0]
nch = (lens + CH - 1) // CH
n_items = int(nch.sum())
ge = np.array([(nch > k).sum() for k in range(int(nch.max()) + 2)])
base = np.concatenate([[0], np.cumsum(ge)])
u = t[:n_items * 8].copy()
t0 = u[:, 0].min()
u -= t0
deep = int(np.argmax(nch))
D = int(nch[deep])
last = np.argsort(u[:, 3])[::-1][:8]
print("last units to finish:", [(int(i), round(float(u[i, 0]) / 1e3, 1), round(float(u[i, 3]) / 1e3, 1)) for i in last], "of", n_items * 8, "(+", int((nch == 0).sum()), "empty tiles)")
print(f"items {n_items} units {n_items*8} deepest tile {deep}: {int(lens[deep])} entries = {D} chunks; kernel span {u[:,3].max()/1e3:.1f} us")
rows, wait, chain = u[:, 1] - u[:, 0], u[:, 2] - u[:, 1], u[:, 3] - u[:, 2]
for name, a in (("rows", rows), ("wait", wait), ("chain", chain)):
    print(f"{name:6s} mean {a.mean()/1e3:6.2f} us p50 {np.median(a)/1e3:6.2f} p90 {np.percentile(a,90)/1e3:6.2f} max {a.max()/1e3:6.2f} sum {a.sum()/1e3:9.1f}")
res = {}
for blk in range(8):
    tl = []
    for c in range(D):
        # pos(deepest) = 0 at every level if its chunk count is unique
        order = os.environ.get("GS_FWD_ORDER", "0")
        if order == "0":
            fi = n_items - base[D - c]
        elif order == "1":
            fi = base[c]
        else:  # mixed: bucket K = 2 c + (D - nch); the deepest tile is first in its bucket
            hist = np.bincount(nch, minlength=D + 1)
            valid = lambda K, n: n >= 1 and (K - D + n) >= 0 and (K - D + n) % 2 == 0 and (K - D + n) // 2 < n
            fi = sum(hist[n] for K in range(2 * c) for n in range(1, D + 1) if valid(K, n))
        tl.append(u[fi * 8 + blk] / 1e3)
    tl = np.array(tl)
    gaps = tl[1:, 2] - tl[:-1, 3]       # predecessor published -> this unit has the state (hand-off latency, or waiting for own rows)
    late_rows = np.maximum(tl[1:, 1] - tl[:-1, 3], 0)  # this unit's rows finished AFTER the predecessor's chain: rows on the critical path
    print(f"blk {blk}: start {tl[0,0]:.1f} end {tl[-1,3]:.1f} | chain sum {chain_sum:.1f}" if False else
          f"blk {blk}: first start {tl[0,0]:5.1f} last chain_done {tl[-1,3]:5.1f} | sum chain {(tl[:,3]-tl[:,2]).sum():5.1f} sum gap {gaps.sum():5.1f} (of which rows not ready {late_rows.sum():5.1f}) first rows {tl[0,1]-tl[0,0]:4.1f}")
    res[blk] = dict(end=float(tl[-1, 3]), chain=float((tl[:, 3] - tl[:, 2]).sum()), gap=float(gaps.sum()), late_rows=float(late_rows.sum()))
    if blk == 0:
        for c in range(D):
            print("   chunk %2d start %5.1f rows_done %5.1f state %5.1f chain_done %5.1f" % (c, *tl[c]))
json.dump(res, open(os.path.join(ROOT, "gpurun_out", "fwd_critical_path.json"), "w"), indent=1)
