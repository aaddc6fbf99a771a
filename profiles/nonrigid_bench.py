"""Times the non-rigid forward (hash grid + tcgen05 MLP) at P Gaussians with CUDA events; prints one JSON line.
Run on a GPU box:  python profiles/nonrigid_bench.py [P]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "3dgs-avatar-release_b200"), ROOT]
import numpy as np, torch
import synth
from avatar_pose.nonrigid import HashGridMLPFused

P = int(sys.argv[1]) if len(sys.argv) > 1 else 50000
dev = torch.device("cuda:0")
g = synth.canonical_gaussians(P, seed=0)
xyz, sc, rot = (torch.from_numpy(g[k]).to(dev) for k in ("xyz", "scaling", "rotation"))
m = HashGridMLPFused(aabb_min=(xyz.min(0).values - 0.05).tolist(), aabb_max=(xyz.max(0).values + 0.05).tolist()).to(dev)
cond = torch.randn(144, device=dev)
for _ in range(5):
    m.deform(xyz, sc, rot, cond)
torch.cuda.synchronize()
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
ts = []
for _ in range(50):
    flush.zero_()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record(); m.deform(xyz, sc, rot, cond); e.record()
    torch.cuda.synchronize()
    ts.append(s.elapsed_time(e))
ms = float(np.median(ts))
flop_alg = 2.0 * (32 * 128 + 128 * 128 * 2 + 128 * 10) * P           # cond folded: 76 288 flop / Gaussian
flop_issued = 3.0 * 2.0 * (32 * 128 + 128 * 128 * 2 + 128 * 16) * P   # 3 BF16 MMAs per product, N padded to 16
print(json.dumps({"P": P, "ms_median": ms, "alg_tflops": flop_alg / ms / 1e9, "issued_bf16_tflops": flop_issued / ms / 1e9,
                  "note": "prep (weight split) + hash encode + 4 tcgen05 layers + delta application, L2 flushed"}))
