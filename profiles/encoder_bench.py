"""Times the one-launch hierarchical pose encoder against the same module evaluated with torch ops on the same GPU (what the
reference runs: ~150 small kernels per frame); CUDA events, 200 calls each.  python profiles/encoder_bench.py"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "3dgs-avatar-release_b200"), ROOT]
import torch
from avatar_pose.pose_encoder import HierarchicalPoseEncoderFused, _torch_encode

dev = torch.device("cuda:0")
m = HierarchicalPoseEncoderFused().to(dev)
rots, Jtrs = torch.randn(1, 24, 9, device=dev), torch.randn(1, 24, 3, device=dev)
P = [p.detach() for p in m.params()]


def timeit(fn, n=200):
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / n * 1e3


with torch.no_grad():
    us = timeit(lambda: m.encode(rots, Jtrs))
    us_torch = timeit(lambda: _torch_encode(rots[0], Jtrs[0], *P, False))
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        y = _torch_encode(rots[0], Jtrs[0], *P, False)
    us_torch_graph = timeit(g.replay)
    g2 = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g2):
        y2 = m.encode(rots, Jtrs)
    us_graph = timeit(g2.replay)
    err = float((y - y2.reshape(-1)).abs().max())
print(json.dumps({"fused_us_eager": us, "torch_us_eager": us_torch, "fused_us_in_graph": us_graph, "torch_us_in_graph": us_torch_graph,
                  "max_abs_diff": err}))
