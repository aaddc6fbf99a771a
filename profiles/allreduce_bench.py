"""Times the hand-written slab all-reduce against NCCL on the same buffer size and prints the kernel's own phase timestamps
(a multi-rank barrier kernel cannot be replayed under ncu).  torchrun --nproc-per-node N profiles/allreduce_bench.py [P]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "3dgs-avatar-release_b200"), ROOT]
import numpy as np, torch, torch.distributed as dist
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
dev = torch.device("cuda", torch.cuda.current_device())
dist.init_process_group("nccl", device_id=dev)
from avatar_pose.dist import SymmetricGradSlab
P = int(sys.argv[1]) if len(sys.argv) > 1 else 50000
n = 59 * P
res = {}
for blocks, mc in ((32, True), (144, True)):
    slab = SymmetricGradSlab(n, blocks=blocks)
    x = torch.randn(n, device=dev)
    ref = x.clone()

    def timeit(fn, reps=30):
        ts = []
        for _ in range(5):
            fn()
        for _ in range(reps):
            torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(); fn(); e.record(); torch.cuda.synchronize()
            ts.append(s.elapsed_time(e) * 1e3)
        return float(np.median(ts))

    us = timeit(lambda: slab.all_reduce(n))
    st = slab.error_flag[2:10].view(torch.int64).cpu().numpy()
    res[f"native_{'nvls' if slab.multicast else 'p2p'}_blocks{slab.blocks}"] = {"us": us, "phases_us": [float(st[i + 1] - st[i]) / 1e3 for i in range(3)]}
    slab.check()
res["nccl_us"] = timeit(lambda: dist.all_reduce(ref))
if rank == 0:
    print(json.dumps({"world": world, "bytes": 4 * n, **res}))
dist.barrier(); torch.cuda.synchronize()
slab.close()
dist.destroy_process_group()
