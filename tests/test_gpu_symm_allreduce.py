"""Two-GPU test of the hand-written NVLink all-reduce over the symmetric gradient slab (gs_allreduce.cu) against NCCL's
all_reduce on the same data, eager and inside a CUDA graph, plus the end-to-end property of SURVEY 8(e): the N-rank summed
gradient equals the sum of the single-GPU gradients.  Needs >= 2 GPUs (skipped on the one-GPU box; run with
`gpurun --gpus 2 -- python -m pytest tests/test_gpu_symm_allreduce.py -m gpu`)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))

WORKER = r'''
import os, sys
import numpy as np
import torch, torch.distributed as dist
sys.path[:0] = [os.path.join(ROOT, "3dgs-avatar-release_b200"), ROOT]
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dev = torch.device("cuda", rank)
dist.init_process_group("nccl", device_id=dev)
import avatar_pose, synth
from avatar_pose.dist import SymmetricGradSlab
P = 3001
slab = SymmetricGradSlab(59 * P)
n = 59 * P
for trial in range(3):
    g = torch.Generator(device="cpu").manual_seed(100 * trial + rank)
    x = torch.randn(n, generator=g).to(dev)
    ref = x.clone()
    dist.all_reduce(ref)
    slab.buf.zero_()
    slab.buf[:n].copy_(x)
    torch.cuda.synchronize(); dist.barrier()
    slab.all_reduce()
    torch.cuda.synchronize()
    slab.check()
    err = float((slab.buf[:n] - ref).abs().max() / ref.abs().max())
    assert err <= 1e-6, ("eager", trial, err)
    # every rank holds the same bits
    mine = slab.buf[:n].clone(); other = mine.clone()
    dist.broadcast(other, src=0)
    assert torch.equal(mine, other), "replicas differ"
# inside a CUDA graph, replayed with fresh data
static = torch.zeros(n, device=dev)
s = torch.cuda.Stream()
s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    slab.buf[:n].copy_(static); slab.all_reduce()
torch.cuda.current_stream().wait_stream(s)
torch.cuda.synchronize(); dist.barrier()
graph = torch.cuda.CUDAGraph()
with torch.cuda.graph(graph):
    slab.buf[:n].copy_(static)
    slab.all_reduce()
for trial in range(3):
    x = torch.randn(n, generator=torch.Generator().manual_seed(7 * trial + rank)).to(dev)
    ref = x.clone(); dist.all_reduce(ref)
    static.copy_(x)
    graph.replay()
    torch.cuda.synchronize(); slab.check()
    assert float((slab.buf[:n] - ref).abs().max() / ref.abs().max()) <= 1e-6, ("graph", trial)
# SURVEY 8(e): N-rank summed gradient == sum of the single-GPU gradients (pose backward writes into the symmetric slab)
gs = synth.canonical_gaussians(P, seed=0)
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
params = {k: t(gs[k]).requires_grad_(True) for k in ("xyz", "rotation", "scaling", "opacity", "features_dc", "features_rest")}
wts = t(synth.skinning_weights(gs["xyz"]))
cam = synth.camera(W=64, H=64, fx=67.0, fy=67.0)
def grads_for(r, into_slab):
    for v in params.values(): v.grad = None
    avatar_pose.set_grad_slab_buffer(slab.buf if into_slab else None)
    ps = synth.pose(1 + r)
    means, cov, col, op = avatar_pose.fused_pose(params["xyz"], params["rotation"], params["scaling"], params["opacity"],
                                                 params["features_dc"], params["features_rest"], wts, t(ps["bone_transforms"]),
                                                 t(cam["camera_center"]), 3)
    ((means ** 2).sum() + cov.sum() + (col * op).sum()).backward()
    return avatar_pose.last_grad_slab()
single = sum(grads_for(r, False).clone() for r in range(world))
g_mine = grads_for(rank, True)
assert g_mine.data_ptr() == slab.buf.data_ptr(), "pose backward must write into the symmetric slab"
torch.cuda.synchronize(); dist.barrier()
slab.all_reduce()
torch.cuda.synchronize(); slab.check()
err = float((slab.buf[:n] - single).abs().max() / single.abs().max())
assert err <= 1e-5, ("dp-sum", err)
assert params["xyz"].grad.data_ptr() == slab.buf.data_ptr()  # .grad views see the reduced values
print(f"rank {rank}: ok multicast={bool(slab.multicast)} blocks={slab.blocks}", flush=True)
dist.barrier(); torch.cuda.synchronize()
# orderly teardown (no os._exit: the interpreter's exit handlers must run): graph first, then the slab, then the group
graph.reset(); del graph
for v in params.values(): v.grad = None
avatar_pose.set_grad_slab_buffer(None)
del g_mine, single
slab.close()
torch.cuda.synchronize(); dist.barrier()
dist.destroy_process_group()
'''


def test_symmetric_slab_allreduce_two_gpus(built_lib, tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    root = os.path.dirname(HERE)
    script = tmp_path / "worker.py"
    script.write_text(f"ROOT = {root!r}\n" + WORKER)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", "29533", str(script)], capture_output=True, text=True, timeout=600)
    log = os.path.join(root, "gpurun_out")
    if os.path.isdir(log):
        with open(os.path.join(log, "symm_allreduce_worker.log"), "w") as f:
            f.write(r.stdout + "\n---- stderr ----\n" + r.stderr)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count(": ok") == 2, r.stdout[-2000:]
