"""World-size-2 gloo test (CPU) of the multi-GPU host logic: frame sharding covers every frame once; the all-reduced
gradient slab equals the sum of the single-rank slabs (gradients from the CPU posing oracle, one pose per rank); the
slab layout matches the .grad views."""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _rank_slab(rank, P):
    """Per-rank flat gradient slab from the oracle: pose `rank+1`, loss = sum of weighted outputs."""
    import synth
    from oracle import pose_oracle
    from avatar_pose import dist as apd
    g = synth.canonical_gaussians(P, seed=0)
    w = torch.from_numpy(synth.skinning_weights(g["xyz"]))
    ps = synth.pose(1 + rank)
    cam = synth.camera(azimuth_deg=40.0 * rank)
    names = ["xyz", "rotation", "scaling", "opacity", "features_dc", "features_rest"]
    leaf = {k: torch.from_numpy(g[k]).clone().requires_grad_(True) for k in names}
    feat = torch.cat([leaf["features_dc"], leaf["features_rest"]], 1)
    out = pose_oracle.pose_forward(leaf["xyz"], leaf["rotation"], leaf["scaling"], leaf["opacity"], feat, w,
                                   torch.from_numpy(ps["bone_transforms"]), torch.from_numpy(cam["camera_center"]), 3)
    gen = torch.Generator().manual_seed(11)
    loss = sum((out[k] * torch.randn(out[k].shape, generator=gen)).sum() for k in ("means3D", "cov3D", "colors", "opacity"))
    loss.backward()
    n_rest = leaf["features_rest"].shape[1]
    slab = torch.zeros(apd.slab_layout(P, n_rest)["_total"])
    views = apd.slab_views(slab, P, n_rest)
    for k in names:
        views[k].copy_(leaf[k].grad.reshape(views[k].shape))
    return slab, n_rest


def _worker(rank, world, port, P, q):
    for p in (os.path.join(ROOT, "3dgs-avatar-release_b200"), ROOT, HERE):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from avatar_pose import dist as apd
    torch.set_num_threads(2)
    frames = apd.shard_frames(7, rank, world)
    slab, n_rest = _rank_slab(rank, P)
    mine = slab.clone()
    apd.allreduce_gradients(slab)
    stats = (torch.full((P, 1), float(rank + 1)), torch.ones(P, 1), torch.full((P,), float(10 * (rank + 1))))
    apd.allreduce_densification_stats(*stats)
    q.put((rank, frames, mine.numpy(), slab.numpy(), [s.numpy() for s in stats]))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gradient_allreduce_and_sharding():
    world, P = 2, 400
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, P, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=240) for _ in range(world)], key=lambda t: t[0])
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    frames = sorted(res[0][1] + res[1][1])
    assert frames == list(range(7)) and res[0][1] == [0, 2, 4, 6] and res[1][1] == [1, 3, 5]
    total = res[0][2].astype(np.float64) + res[1][2].astype(np.float64)
    assert np.abs(res[0][2] - res[1][2]).max() > 0  # the ranks really had different poses
    for r in range(world):
        err = np.abs(res[r][3] - total).max() / max(np.abs(total).max(), 1e-30)
        assert err <= 1e-6, err  # N-rank summed gradient == sum of the single-rank gradients
        ga, dn, mr = res[r][4]
        assert np.all(ga == 3.0) and np.all(dn == 2.0) and np.all(mr == 20.0)


def test_slab_layout_matches_fused_pose_backward():
    sys.path.insert(0, os.path.join(ROOT, "3dgs-avatar-release_b200"))
    from avatar_pose import dist as apd
    lay = apd.slab_layout(10, 15)
    assert lay["_total"] == 10 * 59  # 59 floats per Gaussian
    assert [lay[k][0] for k in ("xyz", "rotation", "scaling", "opacity", "features_dc", "features_rest")] == [0, 30, 70, 100, 110, 140]
