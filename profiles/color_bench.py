"""Times the colour-MLP texture forward (gs_color.cu: compose_input + 3 tcgen05 layers + sigmoid) at P Gaussians with CUDA
events, against the same stage in plain PyTorch on the same GPU (the reference's ColorMLP is ~25 torch kernels); prints one
JSON line.  Run on a GPU box:  python profiles/color_bench.py [P]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "3dgs-avatar-release_b200"), ROOT]
import numpy as np, torch
from avatar_pose.texture import ColorMLPFused
from avatar_pose import _torch_ref

P = int(sys.argv[1]) if len(sys.argv) > 1 else 50000
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
r = lambda *s: torch.randn(*s, generator=g).to(dev)
feats, xyz, nrf, lat, cam = r(P, 32), 0.5 * r(P, 3), 0.5 * r(P, 16), 0.5 * r(16), torch.tensor([0.2, 0.9, 3.5], device=dev)
T = torch.eye(4, device=dev).repeat(P, 1, 1)
T[:, :3, :3] = torch.linalg.qr(r(P, 3, 3))[0]
m = ColorMLPFused().to(dev)


def timeit(fn, n=50):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    ts = []
    for _ in range(n):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    return float(np.median(ts))


ms = timeit(lambda: m.colors(feats, xyz, T, cam, nrf, lat))
params = [p.detach() for p in m.params()]
with torch.no_grad():
    ms_torch = timeit(lambda: _torch_ref.color_mlp(feats, xyz, T.reshape(P, 16), cam, nrf, lat, params[0::2], params[1::2], 3, True, None))
    err = float((m.colors(feats, xyz, T, cam, nrf, lat) - _torch_ref.color_mlp(feats, xyz, T.reshape(P, 16), cam, nrf, lat, params[0::2], params[1::2], 3, True, None)).abs().max())
flop_alg = 2.0 * (63 * 64 + 64 * 64 + 64 * 3) * P                 # latent folded
flop_issued = 3.0 * 2.0 * (64 * 64 + 64 * 64 + 64 * 16) * P       # 3 BF16 MMAs per product, K padded to 64, N to 16
bytes_alg = (32 + 3 + 16 + 16 + 3) * 4.0 * P                      # features, xyz, T_fwd, non-rigid feature in; colours out
print(json.dumps({"P": P, "ms_median": ms, "torch_ms_median": ms_torch, "speedup_vs_torch": ms_torch / ms, "max_abs_diff_vs_torch_fp32": err,
                  "alg_tflops": flop_alg / ms / 1e9, "issued_bf16_tflops": flop_issued / ms / 1e9, "alg_GBps": bytes_alg / ms / 1e6,
                  "note": "prep (weight split) + persistent tcgen05 kernel, L2 flushed; HBM-bound stage: 280 B/Gaussian"}))
