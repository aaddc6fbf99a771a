import os, sys
ROOT = "/root/repo"
sys.path[:0] = [os.path.join(ROOT, "3dgs-avatar-release_b200"), ROOT]
import torch, torch.distributed as dist
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank); dev = torch.device("cuda", rank)
dist.init_process_group("nccl", device_id=dev)
from avatar_pose.dist import SymmetricGradSlab
slab = SymmetricGradSlab(59 * 3001)
pad = slab.hdl.get_signal_pad(rank, (2304,), dtype=torch.int32)
torch.cuda.synchronize()
nz = pad.nonzero().flatten().tolist()
print(f"rank {rank} pad nonzero after init: {nz[:20]} vals {[int(pad[i]) for i in nz[:20]]}", flush=True)
n = 59 * 3001
x = torch.full((n,), float(rank + 1), device=dev)
slab.buf[:n].copy_(x)
torch.cuda.synchronize(); dist.barrier()
slab.all_reduce()
torch.cuda.synchronize()
print(f"rank {rank} result head {slab.buf[:4].tolist()} tail {slab.buf[n-4:n].tolist()} mid {slab.buf[n//2:n//2+2].tolist()} err {int(slab.error_flag[0])} ctrl {slab.error_flag[12:16].tolist()}", flush=True)
nz = pad.nonzero().flatten().tolist()
print(f"rank {rank} pad nonzero after 1 call: {nz[:20]} vals {[int(pad[i]) for i in nz[:20]]}", flush=True)
bad = (slab.buf[:n] != 3.0).nonzero().flatten()
print(f"rank {rank} wrong elements: {bad.numel()} first {bad[:5].tolist()} last {bad[-5:].tolist() if bad.numel() else []}", flush=True)
dist.barrier(); slab.close(); dist.destroy_process_group()
