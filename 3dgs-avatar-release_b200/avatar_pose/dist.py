"""Multi-GPU host logic of the hot path: one process per GPU, frames sharded round-robin, parameters
replicated, and -- for data-parallel training -- ONE all-reduce per step over the flat per-Gaussian gradient slab that
pose_backward writes.  No torch.cuda here: the same code runs under gloo on CPU in the tests."""
import torch
import torch.distributed as dist

SLAB_FIELDS = (("xyz", 3), ("rotation", 4), ("scaling", 3), ("opacity", 1), ("features_dc", 3), ("features_rest", None))


def shard_frames(n_frames, rank, world):
    """Frame f -> rank f mod world (static round-robin: frames have near-equal cost)."""
    return list(range(rank, n_frames, world))


def slab_layout(P, n_rest):
    """Attribute-major layout of the gradient slab: name -> (offset, width, shape). Matches avatar_pose._FusedPose.backward."""
    out, off = {}, 0
    for name, w in SLAB_FIELDS:
        w = 3 * n_rest if w is None else w
        shape = {"features_dc": (P, 1, 3), "features_rest": (P, n_rest, 3)}.get(name, (P, w))
        out[name] = (off, w, shape)
        off += P * w
    out["_total"] = off
    return out


def slab_views(slab, P, n_rest):
    lay = slab_layout(P, n_rest)
    assert slab.numel() == lay["_total"], "slab size does not match the layout"
    return {k: slab[v[0]:v[0] + P * v[1]].view(v[2]) for k, v in lay.items() if k != "_total"}


def allreduce_gradients(slab, average=False, group=None):
    """Sum (or mean) of the per-rank gradient slabs, in place: the only collective of a data-parallel step."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return slab
    dist.all_reduce(slab, op=dist.ReduceOp.SUM, group=group)
    if average:
        slab.div_(dist.get_world_size(group))
    return slab


def allreduce_densification_stats(grad_accum, denom, max_radii2D, group=None):
    """Keeps replicas identical: sums of the view-space gradient accumulators (scene/gaussian_model.py:464-466) and the
    max of the screen radii (train.py:219-220)."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return
    dist.all_reduce(grad_accum, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(denom, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(max_radii2D, op=dist.ReduceOp.MAX, group=group)
