#!/usr/bin/env python
"""bench.py -- 3DGS-Avatar hot path on B200: training step fwd+bwd (posing + rasterizer) at 512x512, 50k Gaussians.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one pass of the hot path over one frame (BASELINE.json config[2]): fused posing forward -> rasterizer
forward -> L1 loss -> rasterizer backward -> posing backward.  The forward-only render rate (config[1]) rides along as
`render_fps`.  N>1 (torchrun): every rank takes a different camera/pose (frame sharding, weak scaling); the training
step adds one NCCL all-reduce of the flat per-Gaussian gradient buffer.
`--impl reference`: the same workload through the CPU restatement of the reference (oracle/), rank 0 only.
Prints ONE JSON line (contract in the task statement).
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (os.path.join(ROOT, "3dgs-avatar-release_b200"), ROOT):
    if _p not in sys.path:
        sys.path.insert(0, _p)

if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
    os.environ["NCCL_DEBUG"] = "WARN"  # keep stdout to the one JSON line (NCCL prints its version banner there)

import numpy as np
import torch

P_DEFAULT, W_DEFAULT, H_DEFAULT = 50_000, 512, 512
METRIC = "fwd+bwd training iterations/s @512x512, 50k Gaussians (posing + rasterizer)"


def workload_config(args):
    """The `config` object of the JSON line: identical text in both arms (ours / --impl reference) for the same flags."""
    W, H, P, world = args.width, args.height, args.P, args.gpus
    return {"workload": f"training step fwd+bwd, {W}x{H}, {P} random-init Gaussians, SH deg 3, rigid LBS posing + tile "
                        "rasterizer (BASELINE.json configs[2]); one camera/pose per rank",
            "l2": "GPU arm: flushed between timed steps (256 MiB write, outside the per-step events); CPU arm: host caches, not flushed",
            "parallelism": f"dp{world} (frame-sharded; one all-reduce of the flat gradient slab per step)" if world > 1 else "single GPU"}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "reference-cuda"])
    ap.add_argument("--P", type=int, default=P_DEFAULT)
    ap.add_argument("--width", type=int, default=W_DEFAULT)
    ap.add_argument("--height", type=int, default=H_DEFAULT)
    ap.add_argument("--cpu-steps", type=int, default=3, help="steps of the cpu_baseline sample inside the ours arm")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--eager", action="store_true", help="time the eager Python path instead of CUDA-graph replay of the same step")
    ap.add_argument("--no-stress", action="store_true", help="skip the 1M-Gaussian / 1080p stress leg (BASELINE.json configs[3])")
    ap.add_argument("--profile-only", action="store_true", help="run warmup+steps training steps and exit (for ncu)")
    return ap.parse_args()


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            pk = json.load(f)
        return float(pk["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def make_workload(P, W, H, n_frames):
    """Synthetic inputs: random-init canonical Gaussians (config[2]), skinning weights, a pose and
    a ring camera per frame, an L1 target per frame.  All numpy float32."""
    import synth
    g = synth.canonical_gaussians(P, seed=0, random_init=True)
    wts = synth.skinning_weights(g["xyz"])
    frames = []
    for f in range(n_frames):
        cam = synth.camera(W=W, H=H, fx=537.0 * W / 512.0, fy=537.0 * H / 512.0, azimuth_deg=360.0 * f / max(n_frames, 1))
        frames.append(dict(cam=cam, pose=synth.pose(1 + f), target=synth.l1_target(H, W, seed=2 + f)))
    return g, wts, frames


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        sm = sorted(float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in self.rows if len(r) > 3 + i)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons,
                "samples": len(self.rows)}


# --------------------------------------------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------------------------------------------
class HotPath:
    """Device-resident canonical Gaussians + the per-frame hot path through the public API
    (avatar_pose.fused_pose + diff_gaussian_rasterization.GaussianRasterizer)."""

    def __init__(self, dev, g, wts, W, H):
        from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
        from avatar_pose import fused_pose
        self.dev, self.W, self.H = dev, W, H
        from avatar_pose.loss import l1_loss
        self.Settings, self.Rasterizer, self.fused_pose = GaussianRasterizationSettings, GaussianRasterizer, fused_pose
        self.l1_loss = l1_loss
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
        self.params = {k: t(g[k]).requires_grad_(True) for k in ("xyz", "rotation", "scaling", "opacity", "features_dc", "features_rest")}
        self.weights = t(wts)
        self.bg = torch.zeros(3, device=dev)
        self.P = g["xyz"].shape[0]
        # screenspace_points (gaussian_renderer/__init__.py:76-80) only receives a gradient: one persistent leaf instead of a
        # zeros_like fill per step; `one` = the upstream gradient of the scalar loss (no ones_like fill either)
        self.means2D = torch.zeros(self.P, 3, device=dev, requires_grad=True)
        self.one = torch.ones((), device=dev)

    def frame_to_device(self, fr):
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(self.dev)
        cam = fr["cam"]
        return dict(bones=t(fr["pose"]["bone_transforms"]), view=t(cam["world_view_transform"]),
                    proj=t(cam["full_proj_transform"]), campos=t(cam["camera_center"]), target=t(fr["target"]),
                    tanx=cam["tanfovx"], tany=cam["tanfovy"])

    def frame_to_pinned(self, fr):
        cam = fr["cam"]
        small = np.concatenate([fr["pose"]["bone_transforms"].reshape(-1), cam["world_view_transform"].reshape(-1),
                                cam["full_proj_transform"].reshape(-1), cam["camera_center"].reshape(-1)]).astype(np.float32)
        return dict(small=torch.from_numpy(small).pin_memory(), target=torch.from_numpy(fr["target"]).pin_memory(),
                    tanx=cam["tanfovx"], tany=cam["tanfovy"])

    def upload(self, pf):
        """H2D of one frame's inputs from pinned host memory (pose, camera, target image)."""
        s = pf["small"].to(self.dev, non_blocking=True)
        tgt = pf["target"].to(self.dev, non_blocking=True)
        return dict(bones=s[:384].view(24, 4, 4), view=s[384:400].view(4, 4), proj=s[400:416].view(4, 4), campos=s[416:419],
                    target=tgt, tanx=pf["tanx"], tany=pf["tany"])

    def static_frame(self, fr):
        """Per-frame inputs in fixed device storage (what a CUDA graph reads); `load_static` overwrites them."""
        cam = fr["cam"]
        small = torch.zeros(419, device=self.dev)
        target = torch.zeros(3, self.H, self.W, device=self.dev)
        f = dict(small=small, bones=small[:384].view(24, 4, 4), view=small[384:400].view(4, 4), proj=small[400:416].view(4, 4),
                 campos=small[416:419], target=target, tanx=cam["tanfovx"], tany=cam["tanfovy"])
        self.load_static(f, self.frame_to_pinned(fr))
        return f

    def load_static(self, f, pf):
        f["small"].copy_(pf["small"], non_blocking=True)
        f["target"].copy_(pf["target"], non_blocking=True)

    def forward(self, f):
        p = self.params
        means, cov, col, op = self.fused_pose(p["xyz"], p["rotation"], p["scaling"], p["opacity"], p["features_dc"],
                                              p["features_rest"], self.weights, f["bones"], f["campos"], 3)
        rs = self.Settings(self.H, self.W, f["tanx"], f["tany"], self.bg, 1.0, f["view"], f["proj"], 3, f["campos"], False, False)
        means2D = self.means2D if torch.is_grad_enabled() else means
        img, radii = self.Rasterizer(rs)(means3D=means, means2D=means2D, opacities=op, colors_precomp=col, cov3D_precomp=cov)
        return img, radii

    def train_step_mask(self, f, fused):
        """Training step with the mask loss of train.py:106-111,143-153 (return_opacity=True, the reference's default): the
        opacity image comes from the fused pass (fused=True) or from the reference's second rasterizer call with
        colors_precomp = ones (gaussian_renderer/__init__.py:131-142) through the plain API (fused=False)."""
        p = self.params
        for v in p.values():
            v.grad = None
        means, cov, col, op = self.fused_pose(p["xyz"], p["rotation"], p["scaling"], p["opacity"], p["features_dc"],
                                              p["features_rest"], self.weights, f["bones"], f["campos"], 3)
        rs = self.Settings(self.H, self.W, f["tanx"], f["tany"], self.bg, 1.0, f["view"], f["proj"], 3, f["campos"], False, False)
        means2D = torch.zeros_like(means, requires_grad=True)
        r = self.Rasterizer(rs)
        if fused:
            img, _, opim = r(means3D=means, means2D=means2D, opacities=op, colors_precomp=col, cov3D_precomp=cov, return_opacity=True)
        else:
            img, _ = r(means3D=means, means2D=means2D, opacities=op, colors_precomp=col, cov3D_precomp=cov)
            opim, _ = r(means3D=means, means2D=means2D, opacities=op,
                        colors_precomp=torch.ones(op.shape[0], 3, device=op.device), cov3D_precomp=cov)
            opim = opim[:1]
        loss = self.l1_loss(img, f["target"]) + 0.1 * self.l1_loss(opim, f["target"][:1])  # mask_loss_type 'l1'
        loss.backward()
        return loss

    def train_iteration(self, f, opt, stats, fused):
        """One whole training iteration around the hot path (train.py:100-232 minus logging/densify-and-prune): zero_grad,
        posing + rasterizer fwd, L1, backward, densification statistics (train.py:219-220), optimiser step (:231-232).
        fused=True: avatar_pose.optim (one densify_stats launch + one adam_step launch over the flat slab);
        fused=False: the reference's statements with torch ops + torch.optim.Adam on the same tensors."""
        opt.zero_grad(set_to_none=True)
        p = self.params
        means, cov, col, op = self.fused_pose(p["xyz"], p["rotation"], p["scaling"], p["opacity"], p["features_dc"],
                                              p["features_rest"], self.weights, f["bones"], f["campos"], 3)
        rs = self.Settings(self.H, self.W, f["tanx"], f["tany"], self.bg, 1.0, f["view"], f["proj"], 3, f["campos"], False, False)
        means2D = torch.zeros_like(means, requires_grad=True)
        img, radii = self.Rasterizer(rs)(means3D=means, means2D=means2D, opacities=op, colors_precomp=col, cov3D_precomp=cov)
        loss = self.l1_loss(img, f["target"])
        loss.backward()
        if fused:
            from avatar_pose.optim import add_densification_stats
            add_densification_stats(stats["max_radii2D"], stats["xyz_gradient_accum"], stats["denom"], radii, means2D.grad)
        else:
            with torch.no_grad():  # train.py:219-220, gaussian_model.py:464-466 with where() instead of boolean indexing
                vis = radii > 0    # (same values; masks of dynamic size cannot be captured in a CUDA graph)
                stats["max_radii2D"].copy_(torch.where(vis, torch.max(stats["max_radii2D"], radii.float()), stats["max_radii2D"]))
                vis1 = vis.unsqueeze(1)
                stats["xyz_gradient_accum"].add_(torch.where(vis1, torch.norm(means2D.grad[:, :2], dim=-1, keepdim=True), 0.0))
                stats["denom"].add_(vis1.float())
        opt.step()
        return loss

    def train_step(self, f):
        for v in self.params.values():
            v.grad = None
        self.means2D.grad = None
        img, _ = self.forward(f)
        loss = self.l1_loss(img, f["target"])  # l1_loss, train.py:121 (fused forward + gradient kernel)
        loss.backward(self.one)
        return loss

    @torch.no_grad()
    def render(self, f):
        return self.forward(f)[0]


def default_config_step(hp, g, frame, f, GraphedStep, flush, dev, eager, n_steps):
    """fwd+bwd of the reference's default model (configs/config.yaml: non_rigid hashgrid, rigid skinning_field, texture shallow_mlp,
    use_sh false / 32 features per Gaussian, return_opacity + L1 mask loss) through FusedGaussianConverter + the rasterizer."""
    import types
    from avatar_pose.modules import FusedGaussianConverter
    from avatar_pose.nonrigid import HashGridMLPFused
    from avatar_pose.skinning import SkinningFieldFused
    from avatar_pose.texture import ColorMLPFused
    from avatar_pose.pose_encoder import HierarchicalPoseEncoderFused
    P = hp.P
    torch.manual_seed(0)
    gm = types.SimpleNamespace(active_sh_degree=0, max_sh_degree=0)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    for k in ("xyz", "rotation", "scaling", "opacity"):
        setattr(gm, "_" + k, t(g[k]).requires_grad_(True))
    gm._features_dc = (0.5 * torch.randn(P, 1, 1, device=dev)).requires_grad_(True)   # gaussian_model.py:182: [P, feature_dim, 1]
    gm._features_rest = (0.5 * torch.randn(P, 31, 1, device=dev)).requires_grad_(True)
    lo, hi = (g["xyz"].min(0) - 0.05).tolist(), (g["xyz"].max(0) + 0.05).tolist()
    nr = HashGridMLPFused(n_cond=144, feature_dim=16, aabb_min=lo, aabb_max=hi, delay=0)
    with torch.no_grad():
        nr.lin3.weight.mul_(0.01)  # small initial deformation, like the reference's last_layer_init
    conv = FusedGaussianConverter(SkinningFieldFused(aabb_min=lo, aabb_max=hi), nr, HierarchicalPoseEncoderFused(),
                                  texture=ColorMLPFused(n_frames=8)).to(dev).eval()
    params = [p for p in conv.parameters()] + [getattr(gm, "_" + k) for k in ("xyz", "rotation", "scaling", "opacity", "features_dc", "features_rest")]
    rots, Jtrs = t(frame["pose"]["rots"]), t(frame["pose"]["Jtrs"])

    def step():
        for p in params:
            p.grad = None
        camera = types.SimpleNamespace(bone_transforms=f["bones"], camera_center=f["campos"], rots=rots, Jtrs=Jtrs, latent_idx=3)
        deformed, loss_reg, colors = conv(gm, camera, 100)
        rs = hp.Settings(hp.H, hp.W, f["tanx"], f["tany"], hp.bg, 1.0, f["view"], f["proj"], 0, f["campos"], False, False)
        means2D = torch.zeros_like(deformed.get_xyz, requires_grad=True)
        img, radii, opim = hp.Rasterizer(rs)(means3D=deformed.get_xyz, means2D=means2D, opacities=deformed.get_opacity,
                                             colors_precomp=colors, cov3D_precomp=deformed.get_covariance(1.0), return_opacity=True)
        loss = hp.l1_loss(img, f["target"]) + 0.1 * hp.l1_loss(opim, f["target"][:1])
        loss.backward()
        return loss

    fn = step if eager else GraphedStep(step).replay
    ms = float(np.mean(time_steps(fn, n_steps, 3, flush, dev)))
    loss = float(fn().item())
    grads_ok = all(p.grad is not None and bool(torch.isfinite(p.grad).all()) for p in params if p is not conv.texture.latent.weight)
    return {"ms_per_step": ms, "iter_s": 1e3 / ms, "loss": loss, "all_parameters_receive_finite_gradients": grads_ok,
            "what": "fwd+bwd of the reference's default configuration at cfg-A size: hash-grid non-rigid MLP -> skinning-field MLP -> fused "
                    "posing -> colour MLP -> rasterizer with fused opacity pass -> L1 + mask loss; the three MLP stages run hand-written "
                    "tcgen05 forward AND backward kernels, the pose encoder one launch forward and one backward (no torch / cuBLAS kernels on the "
                    "path apart from the mask-loss scaling and the latent embedding lookup)"}


def render_sequence(hp, args, rank, world, dev, GraphedStep, eager, n_frames=64):
    """SURVEY 8(d) metric 1 as the reference's render.py loop runs it: a 64-frame animation (slerp between seeded key poses, ring
    camera), frames sharded f mod N over the ranks, every frame end to end -- H2D of its pose + camera from pinned memory, posing
    + rasterizer forward, D2H of the image into pinned memory (double-buffered: the copy of frame k overlaps frame k+1).
    Returns (frames of this rank, device milliseconds for them)."""
    import synth
    W, H = hp.W, hp.H
    poses = synth.pose_sequence(n_frames, seed=11)
    mine = list(range(rank, n_frames, world))
    frames = [dict(cam=synth.camera(W=W, H=H, fx=537.0 * W / 512.0, fy=537.0 * H / 512.0, azimuth_deg=360.0 * f / n_frames),
                   pose=poses[f], target=np.zeros((3, H, W), np.float32)) for f in mine]
    smalls = torch.stack([hp.frame_to_pinned(fr)["small"] for fr in frames]).pin_memory()
    statics = [hp.static_frame(frames[0]), hp.static_frame(frames[0])]
    with torch.no_grad():  # untimed eager pass over the shard: sizes the pair capacity for the busiest frame before capture
        for i in range(len(mine)):
            statics[0]["small"].copy_(smalls[i], non_blocking=True)
            hp.render(statics[0])
    torch.cuda.synchronize(dev)
    fns = [(lambda s=s: hp.render(s)) if eager else GraphedStep(lambda s=s: hp.render(s)).replay for s in statics]
    pins = [torch.empty(3, H, W).pin_memory(), torch.empty(3, H, W).pin_memory()]
    copy_stream = torch.cuda.Stream(dev)
    main = torch.cuda.current_stream(dev)
    done = [None, None]
    rendered = [torch.cuda.Event(), torch.cuda.Event()]
    copied = [torch.cuda.Event(), torch.cuda.Event()]
    checksum = 0.0

    def run(timed):
        nonlocal checksum
        for i in range(len(mine)):
            slot = i & 1
            if done[slot] is not None:
                main.wait_event(done[slot])          # this slot's previous image has left the device
            statics[slot]["small"].copy_(smalls[i], non_blocking=True)
            img = fns[slot]()
            ev = rendered[slot]
            ev.record(main)
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(ev)
                pins[slot].copy_(img, non_blocking=True)
                d = copied[slot]
                d.record(copy_stream)
            done[slot] = d
            if i and done[slot ^ 1] is not None:
                done[slot ^ 1].synchronize()         # the caller holds frame i-1 in host memory
                if timed:
                    checksum += float(pins[slot ^ 1][0, H // 2, W // 2])
        for d in done:
            if d is not None:
                main.wait_event(d)

    run(False)
    torch.cuda.synchronize(dev)
    best = float("inf")
    for _ in range(5):  # a shard is 8 frames (~1.5 ms) at N = 8: one host hiccup would decide a single pass; best of five passes
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(main)
        run(True)
        e1.record(main)
        torch.cuda.synchronize(dev)
        best = min(best, e0.elapsed_time(e1))
    return len(mine), best, checksum


def stress_step(dev, GraphedStep, eager, flush, n_steps, P=1_000_000, W=1920, H=1080):
    """BASELINE.json configs[3]: 1M Gaussians at 1080p, fwd+bwd of posing + rasterizer.  The 50k avatar proxy is densified to
    1M points in the same capsules (scales shrunk by the cube root of the density ratio, times 14 so that a splat still covers
    a few tiles at 1080p from 3.5 m: N is of the order of 1e7 pairs, SURVEY 8's assumption for this configuration)."""
    import synth
    g = synth.canonical_gaussians(P, seed=5, random_init=True)
    g = dict(g)
    g["scaling"] = (g["scaling"] + np.log(14.0) - np.log(P / 50_000.0) / 3.0).astype(np.float32)
    wts = synth.skinning_weights(g["xyz"])
    hp = HotPath(dev, g, wts, W, H)
    fr = dict(cam=synth.camera(W=W, H=H, fx=537.0 * H / 512.0, fy=537.0 * H / 512.0), pose=synth.pose(1), target=synth.l1_target(H, W, seed=2))
    f = hp.static_frame(fr)
    fn = (lambda: hp.train_step(f)) if eager else GraphedStep(lambda: hp.train_step(f)).replay
    ms = float(np.mean(time_steps(fn, n_steps, 3, flush, dev)))
    ms_r = float(np.mean(time_steps(lambda: hp.render(f), n_steps, 3, flush, dev)))
    img, _ = hp.forward(f)
    N = int(img.grad_fn.num_rendered)
    return {"ms_per_step": ms, "iter_s": 1e3 / ms, "render_ms_eager": ms_r, "N_pairs": N, "P": P, "W": W, "H": H,
            "what": "BASELINE.json configs[3] (stress: 1M Gaussians, 1920x1080, SH deg 3): fwd+bwd training step, one GPU"}


def allreduce_check(symm_slab, n, dev):
    """The hand-written NVLS all-reduce against NCCL on the same slab contents (one seeded pattern per rank)."""
    import torch.distributed as dist
    rank = dist.get_rank()
    gen = torch.Generator(device=dev).manual_seed(1234 + rank)
    x = torch.randn(n, device=dev, generator=gen)
    ref = x.clone()
    dist.all_reduce(ref)
    symm_slab.buf[:n].copy_(x)
    torch.cuda.synchronize(dev)
    dist.barrier()
    symm_slab.all_reduce(n)
    torch.cuda.synchronize(dev)
    symm_slab.check()
    err = (symm_slab.buf[:n] - ref).abs().max() / ref.abs().max()
    dist.all_reduce(err, op=dist.ReduceOp.MAX)
    return float(err.item())


def time_steps(fn, steps, warmup, flush, dev, drain=None):
    """Per-step CUDA-event timing on the current stream with an L2 flush (outside the events) before every step.
    `drain` (pipelined end-to-end legs): completes whatever the last step left in flight, INSIDE its timed region."""
    for _ in range(warmup):
        fn()
    if drain is not None:
        drain()
    torch.cuda.synchronize(dev)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for i, (s, e) in enumerate(ev):
        if flush is not None:
            flush.zero_()
        s.record()
        fn()
        if drain is not None and i == steps - 1:
            drain()
        e.record()
    torch.cuda.synchronize(dev)
    return [s.elapsed_time(e) for s, e in ev]


def fp32_peak_tflops(dev):
    from diff_gaussian_rasterization import _C
    L = _C.lib()
    L.gs_fma_probe.restype = ctypes.c_int
    L.gs_fma_probe.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    out = torch.zeros(4, device=dev)
    blocks, threads, iters = 148 * 8, 256, 20000
    st = torch.cuda.current_stream(dev).cuda_stream
    best = 0.0
    for _ in range(4):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        L.gs_fma_probe(blocks, threads, iters, out.data_ptr(), st)
        e.record()
        torch.cuda.synchronize(dev)
        best = max(best, blocks * threads * 16.0 * iters / (s.elapsed_time(e) * 1e-3) / 1e12)
    return best


def ncu_traffic(stage):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the stage's kernel, from the newest committed
    `ncu --set full` summary under profiles/ that lists it (profiles/summarize.py); None when there is none."""
    import glob
    import re
    natural = lambda p: [int(t) if t.isdigit() else t for t in re.split(r"(\d+)", os.path.basename(p))]  # r1_v13 after r1_v7
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "*_kernels.json")) + glob.glob(os.path.join(ROOT, "profiles", "*", "*_kernels.json")),
                   key=natural, reverse=True)
    mult = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0}
    names = (stage + "_kernel", stage.replace("_fwd", "") + "_kernel")
    for path in files:
        try:
            with open(path) as f:
                ks = json.load(f)
            for k in ks:
                name = k["kernel"].replace("void ", "").strip()  # templated kernels are listed as "void name<...>"
                if name.startswith(names):
                    return (k.get("dram_read", 0.0) * mult.get(k.get("dram_read_unit", "byte"), 1.0) +
                            k.get("dram_write", 0.0) * mult.get(k.get("dram_write_unit", "byte"), 1.0))
        except Exception:
            continue
    return None


def stage_times(hp, f, reps, dev):
    """Average per-stage kernel durations (CUDA events on the launching stream, recorded inside the library)."""
    from diff_gaussian_rasterization import _C
    L = _C.lib()
    L.gs_profile_enable.argtypes = [ctypes.c_int]
    L.gs_profile_read.argtypes = [ctypes.c_void_p]
    L.gs_profile_enable(1)
    acc = np.zeros(7)
    buf = (ctypes.c_float * 7)()
    for _ in range(reps):
        hp.train_step(f)
        torch.cuda.synchronize(dev)
        L.gs_profile_read(ctypes.cast(buf, ctypes.c_void_p))
        acc += np.array(list(buf))
    L.gs_profile_enable(0)
    names = ["preprocess_fwd", "binning", "blend_fwd", "blend_bwd", "preprocess_bwd", "pose_fwd", "pose_bwd"]
    return {n: acc[i] / reps for i, n in enumerate(names)}


# --------------------------------------------------------------------------------------------------------------
# CPU arm: the reference's algorithm restated on the host (oracle/): torch-CPU posing chain + C rasterizer
# --------------------------------------------------------------------------------------------------------------
def cpu_step(g_t, wts_t, fr, W, H, backward=True):
    """One training step of the same workload on the CPU.  Returns (seconds, info)."""
    from oracle import gs_oracle, pose_oracle
    t0 = time.perf_counter()
    leaf = {k: v.detach().clone().requires_grad_(backward) for k, v in g_t.items()}
    feat = torch.cat([leaf["features_dc"], leaf["features_rest"]], 1)
    cam = fr["cam"]
    out = pose_oracle.pose_forward(leaf["xyz"], leaf["rotation"], leaf["scaling"], leaf["opacity"], feat, wts_t,
                                   torch.from_numpy(fr["pose"]["bone_transforms"]), torch.from_numpy(cam["camera_center"]), 3)
    fw = gs_oracle.forward(out["means3D"].detach().numpy(), out["opacity"].detach().numpy(), W, H, cam["tanfovx"], cam["tanfovy"],
                           cam["world_view_transform"], cam["full_proj_transform"], cam["camera_center"], np.zeros(3, np.float32),
                           colors_precomp=out["colors"].detach().numpy(), cov3D_precomp=out["cov3D"].detach().numpy())
    info = dict(N=fw["N"], E_fwd=fw["evals_visited"], E_bwd=fw["evals_blended"])
    if backward:
        dL = (np.sign(fw["color"] - fr["target"]) / fw["color"].size).astype(np.float32)
        bw = gs_oracle.backward(fw, dL)
        torch.autograd.backward([out["means3D"], out["cov3D"], out["colors"], out["opacity"]],
                                [torch.from_numpy(bw["dL_dmeans3D"]), torch.from_numpy(bw["dL_dcov3D"]),
                                 torch.from_numpy(bw["dL_dcolors"]), torch.from_numpy(bw["dL_dopacity"])])
    return time.perf_counter() - t0, info


def run_cpu(args, steps, warmup):
    from oracle import gs_oracle
    gs_oracle.build()
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    gs_oracle.set_num_threads(cores)
    g, wts, frames = make_workload(args.P, args.width, args.height, 1)
    g_t = {k: torch.from_numpy(g[k]) for k in ("xyz", "rotation", "scaling", "opacity", "features_dc", "features_rest")}
    wts_t = torch.from_numpy(wts)
    for _ in range(warmup):
        cpu_step(g_t, wts_t, frames[0], args.width, args.height)
    ts, info = [], {}
    for _ in range(steps):
        dt, info = cpu_step(g_t, wts_t, frames[0], args.width, args.height)
        ts.append(dt)
    return float(np.mean(ts)), cores, info


def reference_main(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps, warmup = args.steps, args.warmup  # one step of this arm = one full training step of the workload on the host (~0.15 s)
    sec, cores, info = run_cpu(args, steps, warmup)
    val = 1.0 / sec
    line = {"metric": METRIC, "value": val, "unit": "iter/s", "n_gpus": args.gpus, "steps": steps, "warmup": warmup,
            "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "impl": "reference",
            "config": workload_config(args),
            "workload_counts": {"N_pairs": info.get("N"), "E_fwd": info.get("E_fwd"), "E_bwd": info.get("E_bwd")},
            "cpu_baseline": {"value": val, "unit": "iter/s", "cores": cores, "kind": "port",
                             "sample": f"{steps} full training steps of the workload (oracle/: torch-CPU posing + OpenMP C rasterizer)"},
            "e2e": {"value": val, "unit": "iter/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "the reference's rasterizer has no CPU path; this arm times the CPU restatement (oracle port). The reference's own "
                    "CUDA rasterizer rebuilt for sm_100 is timed by the `ours` arm as `ref_cuda` when baseline/_ref is present."}
    print(json.dumps(line))


# --------------------------------------------------------------------------------------------------------------
# the reference's own CUDA rasterizer rebuilt for sm_100 (baseline/_ref, git-ignored) + the reference's PyTorch posing
# chain on the GPU: the GPU baseline of BASELINE.md 3.1.  Reported beside our number, never on our path.
# --------------------------------------------------------------------------------------------------------------
def load_ref_cuda():
    import importlib.util
    d = os.path.join(ROOT, "baseline", "_ref", "diff_gaussian_rasterization")
    init = os.path.join(d, "__init__.py")
    if not os.path.exists(init):
        return None
    try:
        spec = importlib.util.spec_from_file_location("ref_dgr", init, submodule_search_locations=[d])
        mod = importlib.util.module_from_spec(spec)
        sys.modules["ref_dgr"] = mod
        spec.loader.exec_module(mod)
        return mod
    except Exception as ex:  # a wrong-arch build, missing libs ...
        sys.stderr.write(f"[bench] baseline/_ref not loadable: {ex}\n")
        return None


def ref_cuda_times(ref, hp, f, steps, warmup, flush, dev):
    """fwd+bwd of the reference rasterizer behind the reference's PyTorch posing chain (oracle.pose_oracle is the
    line-by-line restatement of that chain; here it runs on the GPU as the reference would)."""
    from oracle import pose_oracle
    p = hp.params

    def step(backward=True):
        for v in p.values():
            v.grad = None
        feat = torch.cat([p["features_dc"], p["features_rest"]], 1)
        out = pose_oracle.pose_forward(p["xyz"], p["rotation"], p["scaling"], p["opacity"], feat, hp.weights, f["bones"], f["campos"], 3)
        rs = ref.GaussianRasterizationSettings(hp.H, hp.W, f["tanx"], f["tany"], hp.bg, 1.0, f["view"], f["proj"], 3, f["campos"], False, False)
        means2D = torch.zeros_like(out["means3D"], requires_grad=True)
        img, _ = ref.GaussianRasterizer(rs)(means3D=out["means3D"], means2D=means2D, opacities=out["opacity"],
                                            colors_precomp=out["colors"], cov3D_precomp=out["cov3D"])
        if backward:
            (img - f["target"]).abs().mean().backward()
        return img

    def fwd_only():
        with torch.no_grad():
            return step(False)

    def rast_only():
        with torch.no_grad():
            rs = ref.GaussianRasterizationSettings(hp.H, hp.W, f["tanx"], f["tany"], hp.bg, 1.0, f["view"], f["proj"], 3, f["campos"], False, False)
            return ref.GaussianRasterizer(rs)(means3D=rast_in[0], means2D=rast_in[0], opacities=rast_in[3],
                                              colors_precomp=rast_in[2], cov3D_precomp=rast_in[1])[0]

    with torch.no_grad():
        rast_in = [t.detach().clone() for t in hp.fused_pose(p["xyz"], p["rotation"], p["scaling"], p["opacity"], p["features_dc"],
                                                            p["features_rest"], hp.weights, f["bones"], f["campos"], 3)]
    t_train = float(np.mean(time_steps(step, steps, warmup, flush, dev)))
    t_render = float(np.mean(time_steps(fwd_only, steps, warmup, flush, dev)))
    t_rast = float(np.mean(time_steps(rast_only, steps, warmup, flush, dev)))
    img_ref = fwd_only()
    return t_train, t_render, t_rast, img_ref, rast_only()


def ours_main(args):
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py (ours) needs a CUDA device; there is no CPU fallback")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    # stdout carries exactly ONE JSON line: library banners (NCCL prints its version there) go to stderr
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    import __graft_entry__ as ge
    if rank == 0:
        ge.build()
    if world > 1:
        dist.barrier()
    from diff_gaussian_rasterization import _C
    import avatar_pose
    P, W, H = args.P, args.width, args.height
    g, wts, frames = make_workload(P, W, H, world)
    hp = HotPath(dev, g, wts, W, H)
    f_dev = hp.frame_to_device(frames[rank])
    f_pin = hp.frame_to_pinned(frames[rank])
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2, written between timed steps

    from avatar_pose.graphs import GraphedStep as _GraphedStep
    graphs_alive = []

    def GraphedStep(fn):
        g_ = _GraphedStep(fn)
        graphs_alive.append(g_)
        return g_

    f_static = hp.static_frame(frames[rank])
    eager = args.eager
    from avatar_pose import dist as apd

    # the one collective of the data-parallel step: the hand-written NVLink all-reduce over the symmetric gradient slab
    # (gs_allreduce.cu: NVLS multimem two-shot); NCCL only where symmetric memory / multicast cannot be set up
    symm_slab, collective = None, "none (single GPU)"
    if world > 1:
        collective = "nccl all_reduce of the flat gradient slab"
        # measured on this pool (profiles/r1_v12_allreduce_n*.json): 11.8 MB slab, N = 2: kernel 69 us vs NCCL 67 us; N = 8: kernel
        # 66 us vs NCCL 110 us -> the kernel from four ranks up, NCCL for two (GS_NCCL_ALLREDUCE=1 / =0 force either)
        want_native = os.environ.get("GS_NCCL_ALLREDUCE", "1" if world < 4 else "0") != "1"
        if want_native:
            try:
                symm_slab = apd.SymmetricGradSlab(59 * P)
                avatar_pose.set_grad_slab_buffer(symm_slab.buf)
                collective = ("slab_allreduce kernel over symmetric memory (" +
                              "NVLS multimem.ld_reduce / multimem.st two-shot)")
            except Exception as ex:
                symm_slab = None
                print(f"[bench] symmetric-memory all-reduce unavailable, using NCCL: {str(ex)[:200]}", file=sys.stderr)
        ok = torch.tensor([1.0 if symm_slab is not None else 0.0], device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)  # all ranks must take the same path
        if float(ok.item()) == 0.0 and symm_slab is not None:
            symm_slab = None
            avatar_pose.set_grad_slab_buffer(None)
            collective = "nccl all_reduce of the flat gradient slab"

    def step_with_allreduce(f):
        """One data-parallel training step: fwd + bwd, then ONE all-reduce of the flat per-Gaussian gradient slab."""
        loss = hp.train_step(f)
        if world > 1:
            if symm_slab is not None:
                symm_slab.all_reduce(avatar_pose.last_grad_slab().numel())
            else:
                apd.allreduce_gradients(avatar_pose.last_grad_slab())
        return loss

    # the NCCL all-reduce is captured inside the graph too (one launch per step for everything)
    g_train = None if eager else GraphedStep(lambda: step_with_allreduce(f_static))
    g_render = None if eager else GraphedStep(lambda: hp.render(f_static))

    def train_resident():
        return step_with_allreduce(f_dev) if eager else g_train.replay()

    loss_host = torch.empty(1).pin_memory()
    # e2e input pipeline: two static frames / two graphs in ping-pong; the H2D copy of step k+1's inputs (pose, camera,
    # target image, from pinned host memory) runs on a copy stream while step k computes, and step k+1 waits for it.
    e2e = {"k": 0, "ready": [None, None], "read": [None, None]}
    copy_stream = torch.cuda.Stream(dev)
    loss_pins = [torch.empty(1).pin_memory(), torch.empty(1).pin_memory()]
    if not eager:
        f_static2 = hp.static_frame(frames[rank])
        e2e_frames = [f_static, f_static2]

        def e2e_graph_fn(slot):
            # the D2H copy of the loss is a node of the step's graph (pinned destination): no separate stream operation per step
            loss = step_with_allreduce(e2e_frames[slot])
            loss_pins[slot].copy_(loss.detach().reshape(1), non_blocking=True)
            return loss
        e2e_graphs = [GraphedStep(lambda: e2e_graph_fn(0)), GraphedStep(lambda: e2e_graph_fn(1))]

        ready_evs = [torch.cuda.Event(), torch.cuda.Event()]

        def prefetch(slot):
            # overwrite the slot only after its previous reader (the graph of step k-1) has finished -- NOT after the step
            # just launched, which reads the other slot and is what this copy overlaps
            if e2e["read"][slot] is not None:
                copy_stream.wait_event(e2e["read"][slot])
            with torch.cuda.stream(copy_stream):
                hp.load_static(e2e_frames[slot], f_pin)
                ev = ready_evs[slot]
                ev.record(copy_stream)
            e2e["ready"][slot] = ev

    e2e.update(loss_ev=[None, None], last_loss=None)

    e2e_evs = [torch.cuda.Event(), torch.cuda.Event()]

    def train_e2e():
        """One end-to-end step.  Software-pipelined by one step, like a training loop that logs the previous iteration's
        loss: the D2H copy of THIS step's loss is enqueued right behind its graph, and the host then blocks on the PREVIOUS
        step's copy -- so the next graph launch is never held up by a host round trip, and every step's loss is still read."""
        if eager:
            loss = step_with_allreduce(hp.upload(f_pin))
            loss_host.copy_(loss.detach().reshape(1), non_blocking=True)
            torch.cuda.current_stream(dev).synchronize()
            return
        main = torch.cuda.current_stream(dev)
        slot = e2e["k"] & 1
        if e2e["ready"][slot] is None:
            prefetch(slot)  # very first step: nothing was prefetched yet
        main.wait_event(e2e["ready"][slot])
        e2e_graphs[slot].replay()                      # step + D2H of its loss into loss_pins[slot]
        ev = e2e_evs[slot]                             # one event: the slot's inputs have been read AND its loss is on the host
        ev.record(main)
        e2e["read"][slot] = ev
        e2e["loss_ev"][slot] = ev
        prefetch(slot ^ 1)  # inputs of the NEXT step: the copy runs while this step's kernels do
        if e2e["loss_ev"][slot ^ 1] is not None:
            e2e["loss_ev"][slot ^ 1].synchronize()
            e2e["last_loss"] = float(loss_pins[slot ^ 1][0])  # the caller reads the previous step's loss
        e2e["k"] += 1

    def train_drain():
        if not eager and e2e["loss_ev"][(e2e["k"] - 1) & 1] is not None:
            e2e["loss_ev"][(e2e["k"] - 1) & 1].synchronize()
            e2e["last_loss"] = float(loss_pins[(e2e["k"] - 1) & 1][0])

    img_host = torch.empty(3, H, W).pin_memory()
    img_pins = [img_host, torch.empty(3, H, W).pin_memory()]
    rnd = {"k": 0, "out_done": [None, None]}
    if not eager:
        g_render2 = GraphedStep(lambda: hp.render(f_static2))
        rnd_frames, rnd_graphs = [f_static, f_static2], [g_render, g_render2]

    def render_resident():
        return hp.render(f_dev) if eager else g_render.replay()

    def render_e2e():
        """One end-to-end frame: H2D of pose + camera, render, D2H of the image.  Two output buffers: the D2H copy of frame k
        runs on the copy stream while frame k+1 renders; the host blocks on frame k-1's copy."""
        if eager:
            img = hp.render(hp.upload(f_pin))
            img_host.copy_(img, non_blocking=True)
            torch.cuda.current_stream(dev).synchronize()
            return
        main = torch.cuda.current_stream(dev)
        slot = rnd["k"] & 1
        if rnd["out_done"][slot] is not None:
            main.wait_event(rnd["out_done"][slot])  # this slot's previous image has left the device
        rnd_frames[slot]["small"].copy_(f_pin["small"], non_blocking=True)
        img = rnd_graphs[slot].replay()
        done = torch.cuda.Event()
        done.record(main)
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(done)
            img_pins[slot].copy_(img, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(copy_stream)
        rnd["out_done"][slot] = ev
        if rnd["out_done"][slot ^ 1] is not None:
            rnd["out_done"][slot ^ 1].synchronize()  # the caller has frame k-1 in host memory
        rnd["k"] += 1

    def render_drain():
        if not eager and rnd["out_done"][(rnd["k"] - 1) & 1] is not None:
            torch.cuda.current_stream(dev).wait_event(rnd["out_done"][(rnd["k"] - 1) & 1])  # inside the last timed region

    def timed(fn, steps, warmup, drain=None):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)
        ts = time_steps(fn, steps, warmup, flush, dev, drain)
        torch.cuda.synchronize(dev)
        tot = torch.tensor([float(np.sum(ts))], device=dev)
        if world > 1:
            dist.barrier()
            dist.all_reduce(tot, op=dist.ReduceOp.MAX)
        return float(tot.item()) / steps, ts

    sampler = ClockSampler(local)
    if args.profile_only:
        for _ in range(args.warmup + args.steps):
            train_resident()
        torch.cuda.synchronize(dev)
        os.dup2(saved_stdout, 1)
        print(json.dumps({"profile_only": True, "steps": args.warmup + args.steps}), flush=True)
        finish(world, dev, graphs_alive, symm_slab, sampler)
        return
    if rank == 0:
        sampler.start()
    ms_train, ts_train = timed(train_resident, args.steps, args.warmup)
    ms_train_e2e, _ = timed(train_e2e, args.steps, args.warmup, train_drain)
    ms_render, _ = timed(render_resident, args.steps, args.warmup)
    ms_render_e2e, _ = timed(render_e2e, args.steps, args.warmup, render_drain)
    sampler.stop_flag = True
    # hot-cache numbers (no L2 flush), for information
    torch.cuda.synchronize(dev)
    hot = time_steps(train_resident, max(args.steps // 2, 5), 3, None, dev)
    ms_eager = float(np.mean(time_steps(lambda: hp.train_step(f_dev), max(args.steps // 2, 5), 3, flush, dev)))
    if not eager:
        torch.cuda.synchronize(dev)
        n_graph, over = _C.graph_overflow(dev)
        assert not over, f"the captured pair capacity overflowed (N={n_graph}): timed frames were invalid"
    # SURVEY 8(d) metric 2, second form: the step with the reference's default return_opacity=True (mask loss) -- fused
    # opacity pass vs the reference's two rasterizer calls through the same library
    mask_pass = None
    if world == 1:
        try:
            n_m = max(args.steps // 2, 5)
            if eager:
                fn_f, fn_t = (lambda: hp.train_step_mask(f_dev, True)), (lambda: hp.train_step_mask(f_dev, False))
            else:
                g_f = GraphedStep(lambda: hp.train_step_mask(f_static, True))
                g_t = GraphedStep(lambda: hp.train_step_mask(f_static, False))
                fn_f, fn_t = g_f.replay, g_t.replay
            ms_f = float(np.mean(time_steps(fn_f, n_m, 3, flush, dev)))
            ms_t = float(np.mean(time_steps(fn_t, n_m, 3, flush, dev)))
            lf, lt = float(fn_f().item()), float(fn_t().item())
            mask_pass = {"fused_ms_per_step": ms_f, "fused_iter_s": 1e3 / ms_f, "two_call_ms_per_step": ms_t,
                         "two_call_iter_s": 1e3 / ms_t, "loss_fused": lf, "loss_two_call": lt,
                         "what": "fwd+bwd step with return_opacity=True + L1 mask loss: ONE rasterisation carrying the opacity "
                                 "channel vs the reference's two rasterizer calls (both through this library)"}
        except Exception as ex:  # informational leg: never takes the headline down
            mask_pass = {"unavailable": str(ex)[:300]}
    # the whole iteration (hot path + densification statistics + Adam) in one graph
    full_iter = None
    if world == 1:
        try:
            import types as _types
            from avatar_pose.optim import FlatGaussianParams, FusedAdam
            lrs = {"xyz": 1.6e-4, "f_dc": 2.5e-3, "f_rest": 1.25e-4, "opacity": 5e-2, "scaling": 5e-3, "rotation": 1e-3}
            n_m = max(args.steps // 2, 5)
            res = {}
            for fused in (True, False):
                h2 = HotPath(dev, g, wts, W, H)
                names = {"xyz": "_xyz", "rotation": "_rotation", "scaling": "_scaling", "opacity": "_opacity",
                         "features_dc": "_features_dc", "features_rest": "_features_rest"}
                stats = {"max_radii2D": torch.zeros(P, device=dev), "xyz_gradient_accum": torch.zeros(P, 1, device=dev),
                         "denom": torch.zeros(P, 1, device=dev)}
                if fused:
                    gm = _types.SimpleNamespace(**{a: h2.params[k] for k, a in names.items()})
                    flat = FlatGaussianParams(gm)
                    h2.params = {k: getattr(gm, a) for k, a in names.items()}
                    opt = FusedAdam(flat, lrs, eps=1e-15)
                else:
                    opt = torch.optim.Adam([{"params": [h2.params["xyz"]], "lr": lrs["xyz"]}, {"params": [h2.params["features_dc"]], "lr": lrs["f_dc"]},
                                            {"params": [h2.params["features_rest"]], "lr": lrs["f_rest"]}, {"params": [h2.params["opacity"]], "lr": lrs["opacity"]},
                                            {"params": [h2.params["scaling"]], "lr": lrs["scaling"]}, {"params": [h2.params["rotation"]], "lr": lrs["rotation"]}],
                                           lr=0.0, eps=1e-15, capturable=not eager)
                fn = (lambda h2=h2, opt=opt, stats=stats, fused=fused: h2.train_iteration(f_static if not eager else f_dev, opt, stats, fused))
                if not eager:
                    fn = GraphedStep(fn).replay
                res["fused" if fused else "torch"] = float(np.mean(time_steps(fn, n_m, 3, flush, dev)))
            full_iter = {"fused_ms_per_iteration": res["fused"], "fused_iter_s": 1e3 / res["fused"],
                         "torch_optimizer_ms_per_iteration": res["torch"], "torch_optimizer_iter_s": 1e3 / res["torch"],
                         "what": "hot path fwd+bwd + densification statistics + Adam step: avatar_pose.optim (2 launches over the flat slab) "
                                 "vs torch.optim.Adam(capturable) + the reference's torch statements, both around this library's hot path"}
        except Exception as ex:  # informational leg
            full_iter = {"unavailable": str(ex)[:300]}
    # The reference's DEFAULT configuration (configs/config.yaml:1-9): hashgrid non-rigid + skinning_field + shallow colour MLP +
    # return_opacity mask loss -- every stage hand-written (tcgen05 forward and backward of the three MLPs), one training step
    default_step = None
    if world == 1:
        try:
            default_step = default_config_step(hp, g, frames[rank], f_static if not eager else f_dev, GraphedStep, flush, dev, eager,
                                               max(args.steps // 2, 5))
        except Exception as ex:  # informational leg
            default_step = {"unavailable": str(ex)[:300]}
    # metric 1 as a job: a 64-frame animation rendered end to end, frames sharded over the ranks (all ranks take part)
    render_seq = None
    try:
        if world > 1:
            dist.barrier()
        n_mine, ms_mine, chk = render_sequence(hp, args, rank, world, dev, GraphedStep, eager)
        tot = torch.tensor([ms_mine], device=dev)
        if world > 1:
            dist.all_reduce(tot, op=dist.ReduceOp.MAX)
        _, over = _C.graph_overflow(dev) if not eager else (0, False)
        render_seq = {"frames": 64, "ms_total_max_over_ranks": float(tot.item()), "frames_per_s": 64 * 1e3 / float(tot.item()),
                      "pair_capacity_overflow": bool(over),
                      "h2d_bytes_per_frame": 419 * 4, "d2h_bytes_per_frame": 3 * H * W * 4,
                      "what": "render.py's loop: 64-frame slerp animation + ring camera, frame f on rank f mod N, per frame H2D pose+camera -> "
                              "posing + rasterizer forward -> D2H image (pinned, double-buffered); no collective; best of five passes per rank, max over ranks"}
    except Exception as ex:  # informational leg
        if world > 1:
            raise
        render_seq = {"unavailable": str(ex)[:300]}
    stress = None
    if world == 1 and not args.no_stress:
        try:
            stress = stress_step(dev, GraphedStep, eager, flush, 5)
        except Exception as ex:  # informational leg
            stress = {"unavailable": str(ex)[:300]}
    ar_check = None
    if symm_slab is not None:
        torch.cuda.synchronize(dev)
        symm_slab.check()  # raises if an in-kernel cross-rank barrier gave up
        ar_check = {"max_rel_err_vs_nccl": allreduce_check(symm_slab, 59 * P, dev)}
    if rank != 0:
        finish(world, dev, graphs_alive, symm_slab, sampler)
        return

    # ---- rank 0: stage times, roofline, baselines ----
    hbm_peak, peak_kind = load_peaks()
    fp32_peak = fp32_peak_tflops(dev)
    st = stage_times(hp, f_dev, 10, dev)
    img, radii = hp.forward(f_dev)
    fn = img.grad_fn
    N = int(fn.num_rendered)
    dbg = _C.debug_state(P, W, H, N, radii, fn.saved_tensors[7], fn.saved_tensors[8], fn.saved_tensors[9], names=["n_contrib"])
    E_fwd = int(dbg["n_contrib"].to(torch.int64).sum().item())
    geom_buf = fn.saved_tensors[7]
    img.sum().backward()
    E_bwd = _C.read_counters(geom_buf)[0]  # blended (pixel, entry) pairs, counted on the device by the backward blend
    cpu = None
    if not args.no_cpu_baseline and world == 1:  # N > 1: the other ranks spin in a barrier on the same host cores -- the CPU sample would be garbage
        sec, cores, info = run_cpu(args, args.cpu_steps, 1)
        cpu = {"value": 1.0 / sec, "unit": "iter/s", "cores": cores, "kind": "port",
               "sample": f"{args.cpu_steps} full training steps of this workload on the host (oracle/: torch-CPU posing + OpenMP C rasterizer)"}
        # informational cross-check (the CPU arm poses with torch-CPU fp32, whose matmul precision is host dependent, and
        # blends with the polynomial exp; the parity tests, not the bench, are the gate)
        cpu["matches_gpu"] = {"N_pairs": bool(info["N"] == N), "E_fwd": bool(info["E_fwd"] == E_fwd), "E_bwd": bool(info["E_bwd"] == E_bwd)}
    G = 1e9
    work = {  # ALGORITHMIC work per launch
        "pose_fwd": ("hbm", 384.0 * P), "pose_bwd": ("hbm", 716.0 * P), "preprocess_fwd": ("hbm", 76.0 * P),
        "binning": ("hbm", 28.0 * P + (12.0 + 152.0 + 8.0) * N), "preprocess_bwd": ("hbm", 144.0 * P),
        "blend_fwd": ("fp32", 24.0 * E_fwd), "blend_bwd": ("fp32", 70.0 * E_bwd if E_bwd is not None else None)}
    stages = {}
    for k, ms in st.items():
        kind, amount = work[k]
        ent = {"ms": ms, "bound": kind}
        if amount is not None and ms > 0:
            if kind == "hbm":
                ent.update(achieved=amount / (ms * 1e-3) / G, peak=hbm_peak, unit="GB/s")
            else:
                ent.update(achieved=amount / (ms * 1e-3) / 1e12, peak=fp32_peak, unit="TFLOP/s")
            ent["frac"] = ent["achieved"] / ent["peak"]
        stages[k] = ent
    # SURVEY 8(d), end to end: the time the step's algorithmic work would take with every stage at ITS roofline, over the step
    ideal_ms = sum(1e3 * work[k][1] / ((hbm_peak * G) if work[k][0] == "hbm" else (fp32_peak * 1e12)) for k in st if work[k][1] is not None)
    roof_step = {"ideal_ms": ideal_ms, "measured_ms": ms_train, "frac": ideal_ms / ms_train,
                 "what": "sum over the seven stages of (algorithmic bytes or flops / the stage's measured peak) divided by the measured step"}
    dom = max(st, key=lambda k: st[k])
    roof = dict(stages[dom])
    roof.update(kernel=dom, traffic=ncu_traffic(dom), peak_source=(f"{peak_kind} MEASURED_PEAKS.json hbm_gbs" if roof["bound"] == "hbm" else
                                                        "FP32 FMA probe measured live in bench.py (gs_fma_probe)"))
    roof.pop("ms", None)
    ref = load_ref_cuda()
    ref_cuda = None
    if ref is not None:
        try:
            t_train, t_render, t_rast, img_ref, img_ref_same = ref_cuda_times(ref, hp, f_dev, max(args.steps // 2, 10), 5, flush, dev)
            with torch.no_grad():
                ours_img = hp.render(f_dev)
                err = float((ours_img - img_ref).abs().max().item())
                err_same = float((ours_img - img_ref_same).abs().max().item())
            ref_cuda = {"train_ms": t_train, "train_iter_s": 1e3 / t_train, "render_ms": t_render, "render_fps": 1e3 / t_render,
                        "rasterize_fwd_ms": t_rast, "max_abs_image_diff_vs_ours": err,
                        "max_abs_image_diff_same_posed_inputs": err_same,
                        "note": "max_abs_image_diff_vs_ours compares two different POSING arms (the reference's ~60-kernel PyTorch chain vs fused_pose: "
                                "posed means differ in the last bits, which moves alpha >= 1/255 decisions); with the SAME posed tensors fed to "
                                "both rasterizers the images agree to max_abs_image_diff_same_posed_inputs (tests/test_gpu_vs_reference_cuda.py)",
                        "what": "reference CUDA rasterizer rebuilt for sm_100 + the reference's PyTorch posing chain on the same GPU"}
        except Exception as ex:
            ref_cuda = {"unavailable": str(ex)[:200]}
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    blocks_c = (P + 1023) // 1024
    # depth sort (gs_sort.cu): ONE launch while its CTAs (1024 keys each) are co-resident, else one launch per 8-bit digit
    depth = 1 if (P <= (1 << 19) and blocks_c <= 148 and os.environ.get("GS_SORT_FUSED", "1") != "0") else 4
    if tiles <= 2048:  # direct placement (gs_sort.cu): count+place fused into one grid up to 120 blocks, else two kernels
        binning = depth + (1 if blocks_c <= 120 else 3)            # depth sort, placement (its last block plans) | count, place, plan
    else:
        tile_passes = (max(tiles.bit_length(), 1) + 7) // 8
        binning = depth + 1 + tile_passes + 1 + 1                   # depth sort, emit, tile passes, ranges, plan
    # pose_fwd | preprocess, binning, blend_fwd | l1 loss fwd + bwd | blend_bwd, preprocess_bwd | pose_bwd
    launches = args.steps * (1 + (1 + binning + 1) + 2 + 2 + 1 + (1 if symm_slab is not None else 0))  # + slab_allreduce
    h2d = int(f_pin["small"].numel() * 4 + f_pin["target"].numel() * 4)
    line = {"metric": METRIC, "value": world * 1e3 / ms_train, "unit": "iter/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_train, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": workload_config(args),
            "workload_counts": {"N_pairs": N, "E_fwd": E_fwd, "E_bwd": E_bwd},
            "launch": "eager Python calls" if eager else "one CUDA-graph replay per step (avatar_pose.graphs.GraphedStep); eager_ms_per_step is the same step launched op by op",
            "collective": collective,
            "e2e": {"value": world * 1e3 / ms_train_e2e, "unit": "iter/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                    "ms_per_step": ms_train_e2e,
                    "pipeline": "per step: H2D of the NEXT step's pose+camera+target from pinned memory on a copy stream (double-buffered "
                                "static inputs), graph replay whose last node is the D2H copy of the loss into pinned memory; the host "
                                "blocks on the PREVIOUS step's loss (one-step software pipeline, every loss is read)"},
            "render_fps": world * 1e3 / ms_render,
            "render_e2e": {"value": world * 1e3 / ms_render_e2e, "unit": "frames/s", "h2d_bytes_per_step": h2d - 3 * H * W * 4,
                           "d2h_bytes_per_step": 3 * H * W * 4},
            "eager_ms_per_step": ms_eager, "ms_per_step_hot_l2": float(np.mean(hot)), "ms_per_step_p10_p90": [float(np.percentile(ts_train, 10)), float(np.percentile(ts_train, 90))],
            "gpu_launches": launches, "roofline": roof, "roofline_step": roof_step, "stages": stages, "fp32_peak_tflops_measured": fp32_peak,
            "cpu_baseline": cpu, "ref_cuda": ref_cuda, "return_opacity_step": mask_pass, "full_iteration": full_iter, "default_config_step": default_step, "render_seq64": render_seq, "stress_1m": stress, "allreduce_check": ar_check, "clocks": sampler.summary(), "regrows": _C.stats["capacity_regrows"]}
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    print(json.dumps(line), flush=True)
    if world > 1:
        os.dup2(2, 1)
    finish(world, dev, graphs_alive, symm_slab, sampler)


def finish(world, dev, graphs, symm_slab, sampler):
    """Orderly end of a run, so that the interpreter exits normally (atexit handlers run): the captured CUDA graphs are
    released FIRST (they hold kernels of the process group's communicator and pointers into the symmetric slab), then the
    gradient-slab redirection, then the process group."""
    import gc
    import torch.distributed as dist
    import avatar_pose
    from diff_gaussian_rasterization import _C
    sampler.stop_flag = True
    if sampler.is_alive():
        sampler.join(timeout=10)
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()  # nobody tears down while a peer's graph still reads this rank's slab
        torch.cuda.synchronize(dev)
    for g_ in graphs:
        g_.release()
    graphs.clear()
    avatar_pose.set_grad_slab_buffer(None)
    if symm_slab is not None:
        symm_slab.close()
    _C.release_cached()
    gc.collect()
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    sys.stdout.flush()
    sys.stderr.flush()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        reference_main(a)
    else:
        ours_main(a)
