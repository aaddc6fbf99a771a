"""Generates tests/golden/ref_pose_chain.npz by running the reference's OWN posing code on the CPU:
    SMPLNN.forward                      models/deformer/rigid.py:49-71      (rigid LBS of means and rotations)
    build_rotation / build_scaling_rotation / strip_lowerdiag / strip_symmetric   utils/general_utils.py:73-108,194-207
    build_covariance_from_scaling_rotation   scene/gaussian_model.py:28-32  (covariance_activation)
    SH2RGB.forward                      models/texture/texture.py:21-38     (with the reference's eval_sh, importable)
The modules themselves cannot be imported here (pytorch3d / trimesh / igl / tinycudann at import time) and the helpers
hard-code device="cuda", so each function's source is lifted out of its file with `ast` and executed UNMODIFIED in a
namespace whose `torch` is a thin proxy that drops device="cuda" from torch.zeros / torch.ones.  Everything else -- every
arithmetic statement -- is the reference's.
    python tests/golden/make_pose_chain_golden.py"""
import ast
import importlib.util
import os
import sys
import textwrap
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [os.path.join(ROOT, "3dgs-avatar-release_b200"), ROOT]
REF = "/root/reference"


class TorchProxy:
    """torch, except that zeros/ones ignore device='cuda' (the only edit needed to run the reference helpers on the CPU)."""

    def __getattr__(self, name):
        return getattr(torch, name)

    @staticmethod
    def _cpu(fn):
        def wrapped(*a, **k):
            if str(k.get("device", "")).startswith("cuda"):
                k.pop("device")
            return fn(*a, **k)
        return wrapped

    zeros = property(lambda self: self._cpu(torch.zeros))
    ones = property(lambda self: self._cpu(torch.ones))


def lift(path, name, cls=None):
    """Source of function `name` (optionally a method of class `cls`, or nested anywhere) from file `path`, dedented."""
    text = open(os.path.join(REF, path)).read()
    tree = ast.parse(text)
    scope = tree
    if cls is not None:
        scope = next(n for n in ast.walk(tree) if isinstance(n, ast.ClassDef) and n.name == cls)
    node = next(n for n in ast.walk(scope) if isinstance(n, ast.FunctionDef) and n.name == name)
    lines = text.splitlines()[node.lineno - 1:node.end_lineno]
    return textwrap.dedent("\n".join(lines))


spec = importlib.util.spec_from_file_location("ref_sh_utils", os.path.join(REF, "utils/sh_utils.py"))
sh_utils = importlib.util.module_from_spec(spec)
spec.loader.exec_module(sh_utils)
ns = {"torch": TorchProxy(), "np": np, "eval_sh": sh_utils.eval_sh, "augm_rots": sh_utils.augm_rots}
for fn in ("strip_lowerdiag", "strip_symmetric", "build_rotation", "build_scaling_rotation"):
    exec(lift("utils/general_utils.py", fn), ns)
exec(lift("scene/gaussian_model.py", "build_covariance_from_scaling_rotation"), ns)
exec(lift("models/deformer/rigid.py", "forward", cls="SMPLNN").replace("def forward(", "def smplnn_forward(", 1), ns)
exec(lift("models/texture/texture.py", "forward", cls="SH2RGB").replace("def forward(", "def sh2rgb_forward(", 1), ns)

import synth  # noqa: E402  (same synthetic inputs as the tests)

P = 512
g = synth.canonical_gaussians(P, seed=4)
wts = torch.from_numpy(synth.skinning_weights(g["xyz"]))
ps, cam = synth.pose(3), synth.camera(W=128, H=128, fx=134.0, fy=134.0)
T = torch.from_numpy


class Gaussians:
    def __init__(self):
        self._xyz, self._rotation = T(g["xyz"]), T(g["rotation"])
        self._features = torch.cat([T(g["features_dc"]), T(g["features_rest"])], 1)
        self.max_sh_degree, self.active_sh_degree = 3, 3

    get_xyz = property(lambda s: s._xyz)
    get_features = property(lambda s: s._features)

    def clone(self):
        c = Gaussians.__new__(Gaussians)
        c.__dict__.update(self.__dict__)
        return c

    def set_fwd_transform(self, t):
        self.fwd_transform = t


class Cfg(dict):
    cano_view_dir = True


camera = types.SimpleNamespace(bone_transforms=T(ps["bone_transforms"]), camera_center=T(cam["camera_center"]))
with torch.no_grad():
    deformed = ns["smplnn_forward"](types.SimpleNamespace(query_weights=lambda xyz: wts), Gaussians(), 0, camera)
    cov = ns["build_covariance_from_scaling_rotation"](torch.exp(T(g["scaling"])), 1.0, deformed.rotation_precomp)
    cov_mod = ns["build_covariance_from_scaling_rotation"](torch.exp(T(g["scaling"])), 0.7, deformed.rotation_precomp)
    colors = ns["sh2rgb_forward"](types.SimpleNamespace(cfg=Cfg(), training=False), deformed, camera)
out = dict(xyz=g["xyz"], rotation=g["rotation"], scaling=g["scaling"], opacity=g["opacity"], features_dc=g["features_dc"],
           features_rest=g["features_rest"], weights=wts.numpy(), bone_transforms=ps["bone_transforms"],
           camera_center=cam["camera_center"], means3D=deformed._xyz.numpy(), T_fwd=deformed.fwd_transform.numpy(),
           rotation_bar=deformed.rotation_precomp.numpy(), cov3D=cov.numpy(), cov3D_mod07=cov_mod.numpy(), colors=colors.numpy())
np.savez_compressed(os.path.join(HERE, "ref_pose_chain.npz"), **{k: np.asarray(v, np.float32) for k, v in out.items()})
print({k: np.asarray(v).shape for k, v in out.items()})
