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


class SymmetricGradSlab:
    """The flat gradient slab in symmetric memory + the hand-written NVLink all-reduce over it (`slab_allreduce`,
    include/gsplat_b200.h; gs_allreduce.cu).  Collective constructor (every rank of `group`, same `n`):

        slab = SymmetricGradSlab(59 * P)              # once, after init_process_group("nccl")
        avatar_pose.set_grad_slab_buffer(slab.buf)    # pose backward now writes its gradients straight into it
        loss.backward(); slab.all_reduce()            # one kernel; capturable in the step's CUDA graph

    `buf` is padded to a multiple of 4 * world floats (pad stays zero).  Construction raises where symmetric memory or the
    NVLS multicast object is unavailable (the caller then keeps using allreduce_gradients = NCCL).  A cross-rank barrier that
    never completes (a dead or very late peer) leaves the slab untouched and sets an error word; `all_reduce()` surfaces it
    lazily on the NEXT call (no synchronisation on the hot path) and `check()` after a synchronisation."""

    def __init__(self, n, group=None, blocks=32):
        import ctypes
        import torch.distributed._symmetric_memory as symm
        from diff_gaussian_rasterization import _C as _gs
        group = group if group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        dev = torch.device("cuda", torch.cuda.current_device())
        q = 4 * self.world
        self.n = (int(n) + q - 1) // q * q
        try:
            symm.enable_symm_mem_for_group(group.group_name)
        except Exception:
            pass
        self.buf = symm.empty(self.n, dtype=torch.float32, device=dev)
        self.buf.zero_()
        self.hdl = symm.rendezvous(self.buf, group)
        self.multicast = int(self.hdl.multicast_ptr) if getattr(self.hdl, "has_multicast_support", False) else 0
        if not self.multicast:
            raise RuntimeError("SymmetricGradSlab: no NVLS multicast support for this group (use allreduce_gradients = NCCL)")
        self.byte_offset = int(self.buf.data_ptr()) - int(self.hdl.buffer_ptrs[self.rank])
        pad = int(getattr(self.hdl, "signal_pad_size", 0) or 0)
        if pad and pad < (256 + self.world) * 4:  # the kernel keeps its barrier epochs in words [256, 256 + world) of the pad
            raise RuntimeError(f"SymmetricGradSlab: signal pad of {pad} bytes is too small for the in-kernel barrier")
        self.blocks = max(1, int(blocks))
        self.error_flag = torch.zeros(16, dtype=torch.int32, device=dev)  # [0] error code, [2:10] phase timestamps of block 0
        self._err_host = torch.zeros(1, dtype=torch.int32).pin_memory()  # error word as of the previous call (async copy)

        class Args(ctypes.Structure):
            _fields_ = [("n", ctypes.c_int64), ("rank", ctypes.c_int32), ("world", ctypes.c_int32), ("peer_ptrs", ctypes.c_void_p),
                        ("signal_pads", ctypes.c_void_p), ("multicast_ptr", ctypes.c_void_p), ("byte_offset", ctypes.c_int64),
                        ("blocks", ctypes.c_int32), ("pad_", ctypes.c_int32), ("error_flag", ctypes.c_void_p)]
        self._Args, self._gs, self._ctypes = Args, _gs, ctypes
        L = _gs.lib()
        L.slab_allreduce.restype = ctypes.c_int
        L.slab_allreduce.argtypes = [ctypes.POINTER(Args), ctypes.c_void_p]
        self._L = L
        torch.cuda.synchronize(dev)
        dist.barrier(group)

    def all_reduce(self, n=None):
        """In-place sum over the ranks of buf[:n] (default: the whole buffer), on the current stream."""
        capturing = torch.cuda.is_current_stream_capturing()
        if not capturing and int(self._err_host[0]) != 0:  # lazily: what an earlier call left behind, no synchronisation here
            raise RuntimeError(f"slab_allreduce: an earlier call's cross-rank barrier timed out (code {int(self._err_host[0])}); "
                               "the slab was left unreduced")
        a = self._Args()
        n = self.n if n is None else (int(n) + 3) // 4 * 4
        a.n, a.rank, a.world = n, self.rank, self.world
        a.peer_ptrs, a.signal_pads = int(self.hdl.buffer_ptrs_dev), int(self.hdl.signal_pad_ptrs_dev)
        a.multicast_ptr = self.multicast if self.multicast else None
        a.byte_offset, a.blocks, a.error_flag = self.byte_offset, self.blocks, self.error_flag.data_ptr()
        dev = self.buf.device
        with torch.cuda.device(dev):
            self._gs._check(self._L.slab_allreduce(self._ctypes.byref(a), torch.cuda.current_stream(dev).cuda_stream), "slab_allreduce")
            self._err_host.copy_(self.error_flag[:1], non_blocking=True)
        return self.buf

    def close(self):
        """Releases the symmetric buffer and its rendezvous handle (collective in effect: call on every rank, after the graphs
        that captured all_reduce() have been released and before destroy_process_group)."""
        self.hdl = None
        self.buf = None
        self.error_flag = None

    def check(self):
        """After a synchronisation: raises if an in-kernel barrier gave up (a peer never arrived)."""
        e = int(self.error_flag[0].item())
        if e:
            raise RuntimeError(f"slab_allreduce: cross-rank barrier timed out (code {e})")


def allreduce_densification_stats(grad_accum, denom, max_radii2D, group=None):
    """Keeps replicas identical: sums of the view-space gradient accumulators (scene/gaussian_model.py:464-466) and the
    max of the screen radii (train.py:219-220)."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return
    dist.all_reduce(grad_accum, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(denom, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(max_radii2D, op=dist.ReduceOp.MAX, group=group)
