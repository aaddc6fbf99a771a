"""Generates tests/golden/ref_color_mlp.npz by running the reference's OWN ColorMLP (models/texture/texture.py:40-125) with
its OWN VanillaCondMLP (models/network_utils.py:182-249) and eval_sh_bases on the CPU, in the default configuration
(configs/texture/shallow_mlp.yaml).  The two modules cannot be imported here (tinycudann / pytorch3d at import time), so the
class definitions are lifted out of their files with `ast` and executed unmodified.
    python tests/golden/make_color_mlp_golden.py"""
import ast
import importlib.util
import os
import types

import numpy as np
import torch
import torch.nn as nn

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"


def lift_class(path, name):
    text = open(os.path.join(REF, path)).read()
    node = next(n for n in ast.parse(text).body if isinstance(n, ast.ClassDef) and n.name == name)
    return ast.get_source_segment(text, node)


spec = importlib.util.spec_from_file_location("ref_sh_utils", os.path.join(REF, "utils/sh_utils.py"))
sh_utils = importlib.util.module_from_spec(spec)
spec.loader.exec_module(sh_utils)
ns = {"torch": torch, "nn": nn, "np": np, "eval_sh": sh_utils.eval_sh, "eval_sh_bases": sh_utils.eval_sh_bases,
      "augm_rots": sh_utils.augm_rots}
exec(lift_class("models/network_utils.py", "VanillaCondMLP"), ns)
exec(lift_class("models/texture/texture.py", "ColorPrecompute"), ns)
exec(lift_class("models/texture/texture.py", "ColorMLP"), ns)


class Cfg(dict):  # attribute + .get access, like the OmegaConf node the reference passes in
    __getattr__ = dict.__getitem__


cfg = Cfg(name="mlp", feature_dim=32, use_xyz=False, use_cov=False, use_normal=False, sh_degree=3, non_rigid_dim=16, latent_dim=16,
          cano_view_dir=True, view_noise=45, mlp=Cfg(n_neurons=64, n_hidden_layers=2, skip_in=[], cond_in=[], multires=0))
torch.manual_seed(11)
frame_dict = {10 * i: i for i in range(6)}
model = ns["ColorMLP"](cfg, {"frame_dict": frame_dict}).eval()
with torch.no_grad():
    for l in range(3):
        getattr(model.mlp, f"lin{l}").weight.mul_(3.0)  # beyond the near-linear regime of the default initialisation
g = torch.Generator().manual_seed(5)
P = 400
q, _ = torch.linalg.qr(torch.randn(P, 3, 3, generator=g))
T_fwd = torch.eye(4).repeat(P, 1, 1)
T_fwd[:, :3, :3] = q
T_fwd[:, :3, 3] = 0.3 * torch.randn(P, 3, generator=g)
gaussians = types.SimpleNamespace(get_features=torch.randn(P, 32, 1, generator=g), get_xyz=0.5 * torch.randn(P, 3, generator=g),
                                  fwd_transform=T_fwd, non_rigid_feature=0.5 * torch.randn(P, 16, generator=g))
camera = types.SimpleNamespace(camera_center=torch.tensor([0.2, 0.9, 3.5]), frame_id=30)
with torch.no_grad():
    colors = model(gaussians, camera)
    inputs = model.compose_input(gaussians, camera)
out = dict(features=gaussians.get_features.squeeze(-1), xyz=gaussians.get_xyz, T_fwd=T_fwd.reshape(P, 16), camera_center=camera.camera_center,
           non_rigid_feature=gaussians.non_rigid_feature, latent=model.latent.weight[frame_dict[30]].detach(), colors=colors, inputs=inputs)
for l in range(3):
    out[f"W{l}"], out[f"b{l}"] = getattr(model.mlp, f"lin{l}").weight.detach(), getattr(model.mlp, f"lin{l}").bias.detach()
np.savez_compressed(os.path.join(HERE, "ref_color_mlp.npz"), **{k: v.numpy().astype(np.float32) for k, v in out.items()})
print({k: tuple(v.shape) for k, v in out.items()}, float(colors.std()))
