"""Per-batch clock64 trace of the forward blend on the bench scene's deepest tile (library built with -DGS_FWD_TRACE).
Prints, in SM cycles: consumer wait / process / gap per batch and producer fill time.  python profiles/tools/fwd_trace.py"""
import ctypes, sys, os
import numpy as np
import torch
sys.path[:0] = [os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))]
import bench
from diff_gaussian_rasterization import _C

dev = torch.device("cuda:0")
g, wts, frames = bench.make_workload(50000, 512, 512, 1)
hp = bench.HotPath(dev, g, wts, 512, 512)
f = hp.frame_to_device(frames[0])
for _ in range(3):
    hp.render(f)
torch.cuda.synchronize()
L = _C.lib()
buf = (ctypes.c_longlong * (512 * 8))()
L.gs_debug_fwd_trace.argtypes = [ctypes.c_void_p, ctypes.c_int]
assert L.gs_debug_fwd_trace(buf, 512 * 8) == 0
t = np.array(buf[:], dtype=np.int64).reshape(512, 8)
good = (t[:, 2] > 0) & (t[:, 3] >= t[:, 2]) & (t[:, 4] >= t[:, 3]) & (t[:, 4] - t[:, 2] < 10_000_000) & (t[:, 1] > t[:, 0]) & (t[:, 1] - t[:, 0] < 10_000_000)
nb = int(np.argmin(good)) if not good.all() else len(good)  # the shared trace array is not cleared: stop at the first stale row
t = t[:nb]
wait, proc = t[:, 3] - t[:, 2], t[:, 4] - t[:, 3]
gap = t[1:, 2] - t[:-1, 4]
fill = t[:, 1] - t[:, 0]
print("batches", nb, "total cycles", int(t[-1, 4] - t[0, 2]), "per batch", float(t[-1, 4] - t[0, 2]) / nb)
for name, v in (("consumer wait", wait), ("consumer process", proc), ("consumer gap", gap), ("producer fill", fill)):
    print(f"{name:18s} mean {v.mean():8.1f}  p50 {np.percentile(v, 50):8.1f}  p90 {np.percentile(v, 90):8.1f}  max {v.max():8d}")
cnt = t[:, 5]
print("rows per batch: mean", cnt.mean(), "p50", np.percentile(cnt, 50), "max", cnt.max())
A = np.stack([cnt, np.ones_like(cnt)], 1).astype(float)
coef = np.linalg.lstsq(A, proc.astype(float), rcond=None)[0]
print("consumer process ~= %.1f cycles/row * rows + %.1f" % (coef[0], coef[1]))
lead = t[:, 2] - t[:, 1]  # > 0: the batch was ready before the consumer asked for it
print("producer lead (ready before asked) mean", lead.mean(), "frac ready", float((lead > 0).mean()))
