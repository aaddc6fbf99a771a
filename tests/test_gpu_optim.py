"""Fused Adam + densification statistics (gs_optim.cu) against the reference's own optimiser class: torch.optim.Adam on
the CPU, configured exactly as GaussianModel.training_setup does (scene/gaussian_model.py:201-221: six named groups,
lr per group, eps 1e-15), stepped on identical gradients with the xyz learning rate rescheduled every iteration
(:223-229).  Tolerance: 2e-6 relative to max|param| per group after 25 steps (same FP32 operation order as torch's
_single_tensor_adam; the bias corrections are evaluated in double on both sides)."""
import types

import numpy as np
import pytest
import torch

import synth

pytestmark = pytest.mark.gpu

LRS = {"xyz": 1.6e-4, "f_dc": 2.5e-3, "f_rest": 2.5e-3 / 20.0, "opacity": 5e-2, "scaling": 5e-3, "rotation": 1e-3}
ATTR = {"xyz": "_xyz", "rotation": "_rotation", "scaling": "_scaling", "opacity": "_opacity", "f_dc": "_features_dc", "f_rest": "_features_rest"}
KEY = {"xyz": "xyz", "rotation": "rotation", "scaling": "scaling", "opacity": "opacity", "f_dc": "features_dc", "f_rest": "features_rest"}


def _model(P, dev, seed=0):
    g = synth.canonical_gaussians(P, seed=seed)
    gm = types.SimpleNamespace()
    for k, a in ATTR.items():
        setattr(gm, a, torch.from_numpy(g[KEY[k]]).to(dev).requires_grad_(True))
    return g, gm


@pytest.mark.parametrize("P", [1001, 20000])
def test_fused_adam_matches_torch_adam(built_lib, P):
    from avatar_pose.optim import FlatGaussianParams, FusedAdam, SLAB_ORDER
    dev = torch.device("cuda:0")
    g, gm = _model(P, dev)
    flat = FlatGaussianParams(gm)
    assert gm._xyz.data_ptr() == flat.slab.data_ptr() and flat.slab.numel() == 59 * P
    opt = FusedAdam(flat, LRS, eps=1e-15)
    # the reference's optimiser on the CPU (same group order as training_setup)
    cpu = {k: torch.from_numpy(g[KEY[k]]).clone().requires_grad_(True) for k in ATTR}
    ref = torch.optim.Adam([{"params": [cpu[k]], "lr": LRS[k], "name": k} for k in ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")],
                           lr=0.0, eps=1e-15)
    gen = torch.Generator().manual_seed(5)
    for it in range(25):
        lr_xyz = LRS["xyz"] * (0.01 ** (it / 25.0))  # get_expon_lr_func-style decay, rescheduled every iteration
        for grp in opt.param_groups:
            if grp["name"] == "xyz":
                grp["lr"] = lr_xyz
        for grp in ref.param_groups:
            if grp["name"] == "xyz":
                grp["lr"] = lr_xyz
        for k in ATTR:
            gr = torch.randn(cpu[k].shape, generator=gen) * (10.0 ** float(torch.randint(-6, 0, (1,), generator=gen)))
            gr[torch.rand(P, generator=gen) < 0.3] = 0.0  # invisible Gaussians get exact zeros
            cpu[k].grad = gr.clone()
            getattr(gm, ATTR[k]).grad = gr.to(dev)
        opt.step()
        ref.step()
        opt.zero_grad()
    assert opt.step_count() == 25
    for k in SLAB_ORDER:
        a, b = getattr(gm, ATTR[k]).detach().cpu(), cpu[k].detach()
        err = float((a - b).abs().max() / b.abs().max())
        assert err <= 2e-6, (k, err)
        m = opt.exp_avg[flat.group_end[SLAB_ORDER.index(k)] - a.numel():flat.group_end[SLAB_ORDER.index(k)]].cpu().view(a.shape)
        mref = ref.state[cpu[k]]["exp_avg"]
        assert float((m - mref).abs().max()) <= 1e-6 * float(mref.abs().max())


def test_fused_adam_consumes_pose_gradient_slab_in_place(built_lib):
    """After a fused_pose backward the .grad tensors are views of the flat gradient slab: the step reads it in place
    (no gather) and gives the same update as torch.optim.Adam on copies."""
    import avatar_pose
    from avatar_pose.optim import FlatGaussianParams, FusedAdam
    dev = torch.device("cuda:0")
    P = 4000
    g, gm = _model(P, dev, seed=2)
    flat = FlatGaussianParams(gm)
    opt = FusedAdam(flat, LRS)
    clones = {k: getattr(gm, ATTR[k]).detach().clone().requires_grad_(True) for k in ATTR}
    ref = torch.optim.Adam([{"params": [clones[k]], "lr": LRS[k]} for k in ATTR], lr=0.0, eps=1e-15)
    wts = torch.from_numpy(synth.skinning_weights(g["xyz"])).to(dev)
    ps, cam = synth.pose(1), synth.camera(W=64, H=64, fx=67.0, fy=67.0)
    bones, campos = torch.from_numpy(ps["bone_transforms"]).to(dev), torch.from_numpy(cam["camera_center"]).to(dev)

    def loss_of(p):
        means, cov, col, op = avatar_pose.fused_pose(p["xyz"], p["rotation"], p["scaling"], p["opacity"], p["f_dc"], p["f_rest"], wts, bones, campos, 3)
        return (means ** 2).sum() + cov.sum() + (col * op).sum()

    loss_of({k: getattr(gm, ATTR[k]) for k in ATTR}).backward()
    slab = avatar_pose.last_grad_slab()
    assert opt._grad_slab().data_ptr() == slab.data_ptr(), "the gradient slab must be consumed in place"
    loss_of(clones).backward()
    opt.step()
    ref.step()
    for k in ATTR:
        a, b = getattr(gm, ATTR[k]).detach(), clones[k].detach()
        assert float((a - b).abs().max() / b.abs().max()) <= 2e-6, k


def test_densification_stats(built_lib):
    from avatar_pose.optim import add_densification_stats
    dev = torch.device("cuda:0")
    P = 5003
    gen = torch.Generator().manual_seed(3)
    radii = torch.randint(-1, 40, (P,), generator=gen, dtype=torch.int32)
    radii[radii < 5] = 0
    vgrad = torch.randn(P, 3, generator=gen) * 1e-3
    mr, acc, den = torch.rand(P, generator=gen) * 30, torch.rand(P, 1, generator=gen), torch.randint(0, 5, (P, 1), generator=gen).float()
    # reference statements on the CPU (train.py:219-220, gaussian_model.py:464-466)
    vis = radii > 0
    mr_ref, acc_ref, den_ref = mr.clone(), acc.clone(), den.clone()
    mr_ref[vis] = torch.max(mr_ref[vis], radii[vis])
    acc_ref[vis] += torch.norm(vgrad[vis, :2], dim=-1, keepdim=True)
    den_ref[vis] += 1
    d = lambda t: t.to(dev)
    mr_d, acc_d, den_d = d(mr), d(acc), d(den)
    add_densification_stats(mr_d, acc_d, den_d, d(radii), d(vgrad))
    assert torch.equal(mr_d.cpu(), mr_ref) and torch.equal(den_d.cpu(), den_ref)
    assert float((acc_d.cpu() - acc_ref).abs().max()) <= 1e-7
    assert torch.equal(acc_d.cpu()[~vis], acc[~vis])


def test_fused_adam_resize_keeps_moments_like_the_reference(built_lib):
    """Densify-and-prune surgery (scene/gaussian_model.py:251-312): pruning keeps the surviving rows of every parameter and of
    both Adam moments, densification appends rows with zero moments, the step count goes on.  FusedAdam.resize against
    torch.optim.Adam on the CPU with the reference's state surgery done by hand, stepped before and after."""
    from avatar_pose.optim import FlatGaussianParams, FusedAdam, SLAB_ORDER
    dev = torch.device("cuda:0")
    P, n_new = 3000, 700
    g, gm = _model(P, dev, seed=4)
    flat = FlatGaussianParams(gm)
    opt = FusedAdam(flat, LRS, eps=1e-15)
    cpu = {k: torch.from_numpy(g[KEY[k]]).clone().requires_grad_(True) for k in ATTR}
    order = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")
    ref = torch.optim.Adam([{"params": [cpu[k]], "lr": LRS[k], "name": k} for k in order], lr=0.0, eps=1e-15)
    gen = torch.Generator().manual_seed(9)

    def steps(n):
        for _ in range(n):
            for k in ATTR:
                gr = torch.randn(cpu[k].shape, generator=gen) * 1e-3
                cpu[k].grad = gr.clone()
                getattr(gm, ATTR[k]).grad = gr.to(dev)
            opt.step()
            ref.step()
            opt.zero_grad()

    steps(6)
    keep = torch.rand(P, generator=gen) > 0.25
    new = {k: torch.randn(n_new, *cpu[k].shape[1:], generator=gen) * 0.1 for k in ATTR}
    # the reference's surgery on torch.optim.Adam (gaussian_model.py:251-270 then :290-312)
    for grp in ref.param_groups:
        k = grp["name"]
        st = ref.state.pop(grp["params"][0])
        st["exp_avg"] = torch.cat([st["exp_avg"][keep], torch.zeros_like(new[k])])
        st["exp_avg_sq"] = torch.cat([st["exp_avg_sq"][keep], torch.zeros_like(new[k])])
        cpu[k] = torch.cat([grp["params"][0].detach()[keep], new[k]]).requires_grad_(True)
        grp["params"][0] = cpu[k]
        ref.state[cpu[k]] = st
    out = opt.resize(keep.to(dev), {k: v.to(dev) for k, v in new.items()})
    P2 = int(keep.sum()) + n_new
    assert gm._xyz.shape == (P2, 3) and out["xyz"] is gm._xyz and flat.slab.numel() == 59 * P2 and opt.step_count() == 6
    steps(5)
    assert opt.step_count() == 11
    for k in SLAB_ORDER:
        a, b = getattr(gm, ATTR[k]).detach().cpu(), cpu[k].detach()
        assert a.shape == b.shape
        assert float((a - b).abs().max() / b.abs().max()) <= 2e-6, k
