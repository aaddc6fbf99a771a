"""Optimiser-side per-Gaussian work of a training iteration: host side of `adam_step` / `densify_stats`
(include/gsplat_b200.h).  Mirrors what GaussianModel.training_setup builds (scene/gaussian_model.py:201-221: one
torch.optim.Adam with six named parameter groups, lr per group, eps 1e-15) and what train.py:219-220,231-232 run per
iteration, as ONE kernel over a flat parameter slab instead of one foreach chain per group.

    flat = FlatGaussianParams(gaussians)          # re-homes _xyz ... _features_rest as views of one slab
    opt = FusedAdam(flat, lrs={"xyz": 1.6e-4, ...}, eps=1e-15)
    for g in opt.param_groups: ...                # `name` / `lr` like torch's: update_learning_rate() works unchanged
    loss.backward(); opt.step(); opt.zero_grad()

The slab layout is the flat gradient slab `pose_backward` writes (avatar_pose.last_grad_slab()), so after a fused_pose
backward the gradients are consumed in place; otherwise the `.grad` tensors are gathered with one cat.  No CPU fallback."""
import ctypes

import torch

from diff_gaussian_rasterization import _C as _gs

vp = ctypes.c_void_p
ADAM_MAX_GROUPS = 8
# slab order = gradient-slab order of pose_backward (avatar_pose/__init__.py): attribute-major
SLAB_ORDER = ("xyz", "rotation", "scaling", "opacity", "f_dc", "f_rest")
_ATTR = {"xyz": "_xyz", "rotation": "_rotation", "scaling": "_scaling", "opacity": "_opacity", "f_dc": "_features_dc",
         "f_rest": "_features_rest"}


class AdamArgs(ctypes.Structure):
    _fields_ = [("n", ctypes.c_int64), ("n_groups", ctypes.c_int32), ("pad_", ctypes.c_int32),
                ("group_end", ctypes.c_int64 * ADAM_MAX_GROUPS), ("lrs", vp), ("beta1", ctypes.c_float),
                ("beta2", ctypes.c_float), ("eps", ctypes.c_float), ("pad2_", ctypes.c_float), ("params", vp), ("grads", vp),
                ("exp_avg", vp), ("exp_avg_sq", vp), ("state", vp)]


_configured = False


def _lib():
    global _configured
    L = _gs.lib()
    if not _configured:
        L.adam_state_bytes.restype = ctypes.c_size_t
        L.adam_state_bytes.argtypes = []
        L.adam_step.restype = ctypes.c_int
        L.adam_step.argtypes = [ctypes.POINTER(AdamArgs), vp]
        L.densify_stats.restype = ctypes.c_int
        L.densify_stats.argtypes = [ctypes.c_int32, vp, vp, vp, vp, vp, vp]
        _configured = True
    return L


class FlatGaussianParams:
    """Moves the six per-Gaussian parameter tensors of a GaussianModel-like object into ONE flat FP32 slab (attribute-major,
    the gradient-slab layout) and re-binds the attributes as views of it, so one kernel can update all of them."""

    def __init__(self, gaussians):
        self.gaussians = gaussians
        self._bind([getattr(gaussians, _ATTR[k]) for k in SLAB_ORDER])

    def _bind(self, tensors):
        gaussians = self.gaussians
        dev = tensors[0].device
        if not tensors[0].is_cuda:
            raise _gs.GsError("FlatGaussianParams needs CUDA tensors (there is no CPU path)")
        self.sizes = [t.numel() for t in tensors]
        self.slab = torch.empty(sum(self.sizes), dtype=torch.float32, device=dev)
        self.views, off = {}, 0
        for k, t in zip(SLAB_ORDER, tensors):
            v = self.slab[off:off + t.numel()].view(t.shape)
            with torch.no_grad():
                v.copy_(t)
            v.requires_grad_(True)
            setattr(gaussians, _ATTR[k], v)
            self.views[k] = v
            off += t.numel()
        self.group_end = [sum(self.sizes[:i + 1]) for i in range(len(self.sizes))]


class FusedAdam:
    def __init__(self, flat, lrs, betas=(0.9, 0.999), eps=1e-15):
        """flat: FlatGaussianParams; lrs: {"xyz": .., "f_dc": .., "f_rest": .., "opacity": .., "scaling": .., "rotation": ..}."""
        self.flat, self.betas, self.eps = flat, betas, eps
        dev = flat.slab.device
        self.param_groups = [{"params": [flat.views[k]], "lr": float(lrs[k]), "name": k} for k in SLAB_ORDER]
        self.exp_avg = torch.zeros_like(flat.slab)
        self.exp_avg_sq = torch.zeros_like(flat.slab)
        self.state = torch.zeros(_lib().adam_state_bytes(), dtype=torch.uint8, device=dev)
        self._lrs_host = torch.empty(len(SLAB_ORDER), dtype=torch.float32).pin_memory()
        self._lrs_dev = torch.zeros(len(SLAB_ORDER), dtype=torch.float32, device=dev)
        self._lrs_last = None

    def _grad_slab(self):
        import avatar_pose
        slab = avatar_pose.last_grad_slab()
        grads = [self.flat.views[k].grad for k in SLAB_ORDER]
        if slab is not None and slab.numel() == self.flat.slab.numel() and all(g is not None for g in grads):
            off, in_place = 0, True
            for g, n in zip(grads, self.flat.sizes):
                in_place &= g.is_contiguous() and g.data_ptr() == slab.data_ptr() + 4 * off
                off += n
            if in_place:
                return slab
        z = lambda k: torch.zeros_like(self.flat.views[k]) if self.flat.views[k].grad is None else self.flat.views[k].grad
        return torch.cat([z(k).reshape(-1) for k in SLAB_ORDER])

    def push_lrs(self):
        """Copies the groups' learning rates to the device when they changed (the xyz rate does, every iteration)."""
        lrs = tuple(float(g["lr"]) for g in self.param_groups)
        if lrs != self._lrs_last:
            for i, v in enumerate(lrs):
                self._lrs_host[i] = v
            self._lrs_dev.copy_(self._lrs_host, non_blocking=True)
            self._lrs_last = lrs

    @torch.no_grad()
    def step(self, grad_slab=None):
        L = _lib()
        flat = self.flat
        dev = flat.slab.device
        grads = self._grad_slab() if grad_slab is None else grad_slab
        self.push_lrs()
        a = AdamArgs()
        a.n, a.n_groups = flat.slab.numel(), len(SLAB_ORDER)
        for i, e in enumerate(flat.group_end):
            a.group_end[i] = e
        a.lrs, a.beta1, a.beta2, a.eps = self._lrs_dev.data_ptr(), self.betas[0], self.betas[1], self.eps
        a.params, a.grads = flat.slab.data_ptr(), _gs._ptr(grads, "grads")
        a.exp_avg, a.exp_avg_sq, a.state = self.exp_avg.data_ptr(), self.exp_avg_sq.data_ptr(), self.state.data_ptr()
        with torch.cuda.device(dev):
            _gs._check(L.adam_step(ctypes.byref(a), torch.cuda.current_stream(dev).cuda_stream), "adam_step")

    @torch.no_grad()
    def resize(self, keep_mask=None, new_tensors=None):
        """The optimiser surgery of densify-and-prune with the reference's semantics (scene/gaussian_model.py: `_prune_optimizer`
        :251-270 keeps the rows of `keep_mask` in every parameter AND in exp_avg / exp_avg_sq; `cat_tensors_to_optimizer` :290-312
        appends `new_tensors` = {"xyz": [n,3], "f_dc": [n,1,3], "f_rest": [n,15,3], "opacity": [n,1], "scaling": [n,3],
        "rotation": [n,4]} with ZERO moments); the step counter and the learning rates are kept.  Re-homes the parameters in a new
        slab and re-binds the GaussianModel attributes; returns {name: new parameter tensor} like the reference's helpers."""
        flat = self.flat
        dev = flat.slab.device
        params, m1, m2, off = {}, {}, {}, 0
        for k, n in zip(SLAB_ORDER, flat.sizes):
            shape = flat.views[k].shape
            parts = [flat.views[k].detach(), self.exp_avg[off:off + n].view(shape), self.exp_avg_sq[off:off + n].view(shape)]
            if keep_mask is not None:
                parts = [t[keep_mask] for t in parts]
            if new_tensors is not None:
                ext = new_tensors[k].detach().to(dev, torch.float32)
                parts = [torch.cat([parts[0], ext]), torch.cat([parts[1], torch.zeros_like(ext)]), torch.cat([parts[2], torch.zeros_like(ext)])]
            params[k], m1[k], m2[k] = parts
            off += n
        flat._bind([params[k] for k in SLAB_ORDER])
        self.exp_avg = torch.cat([m1[k].reshape(-1) for k in SLAB_ORDER])
        self.exp_avg_sq = torch.cat([m2[k].reshape(-1) for k in SLAB_ORDER])
        for grp in self.param_groups:
            grp["params"] = [flat.views[grp["name"]]]
        return dict(flat.views)

    def zero_grad(self, set_to_none=True):
        for v in self.flat.views.values():
            if set_to_none:
                v.grad = None
            elif v.grad is not None:
                v.grad.zero_()

    def step_count(self):
        return int(self.state.view(torch.int64)[0].item())


@torch.no_grad()
def add_densification_stats(max_radii2D, xyz_gradient_accum, denom, radii, viewspace_grad):
    """train.py:219-220 + GaussianModel.add_densification_stats (scene/gaussian_model.py:464-466) in one launch;
    visibility_filter = radii > 0.  Updates the three accumulators in place."""
    L = _lib()
    P = radii.shape[0]
    dev = radii.device
    if not radii.is_cuda:
        raise _gs.GsError("add_densification_stats needs CUDA tensors (there is no CPU path)")
    p = _gs._ptr
    vg = viewspace_grad if viewspace_grad.is_contiguous() else viewspace_grad.contiguous()
    if vg.shape != (P, 3):
        raise _gs.GsError("viewspace_grad must be [P,3] (screenspace_points.grad)")
    with torch.cuda.device(dev):
        _gs._check(L.densify_stats(P, p(radii, "radii", torch.int32), p(vg, "viewspace_grad"), p(max_radii2D, "max_radii2D"),
                                   p(xyz_gradient_accum, "xyz_gradient_accum"), p(denom, "denom"),
                                   torch.cuda.current_stream(dev).cuda_stream), "densify_stats")
