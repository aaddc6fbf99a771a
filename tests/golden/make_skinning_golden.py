"""Generates tests/golden/ref_skinning.npz by running the reference's OWN skinning-field code on the CPU: the module-level
hierarchical_softmax (models/deformer/rigid.py:85-129), SkinningField.softmax / get_forward_transform / forward (:160-171,
:189-196, :215-236) with its own VanillaCondMLP (models/network_utils.py:182-249), AABB (utils/dataset_utils.py:66-76) and
build_rotation (utils/general_utils.py:87-108) -- sources lifted with `ast`, executed unmodified (torch proxy drops
device='cuda' in torch.zeros, as in make_pose_chain_golden.py).  Non-distilled path, 25 logits (configs/rigid/skinning_field.yaml).
    python tests/golden/make_skinning_golden.py"""
import ast
import math
import os
import textwrap
import types

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"


class TorchProxy:
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
    text = open(os.path.join(REF, path)).read()
    tree = ast.parse(text)
    body = tree.body if cls is None else next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == cls).body
    node = next(n for n in body if isinstance(n, (ast.FunctionDef, ast.ClassDef)) and n.name == name)
    return textwrap.dedent("\n".join(text.splitlines()[node.lineno - 1:node.end_lineno]))


ns = {"torch": TorchProxy(), "nn": nn, "np": np, "F": F, "math": math}
exec(lift("models/network_utils.py", "VanillaCondMLP"), ns)
exec(lift("utils/dataset_utils.py", "AABB"), ns)
exec(lift("utils/general_utils.py", "build_rotation"), ns)
exec(lift("models/deformer/rigid.py", "hierarchical_softmax"), ns)
for m in ("softmax", "get_forward_transform", "forward"):
    exec(lift("models/deformer/rigid.py", m, cls="SkinningField").replace(f"def {m}(", f"def sf_{m}(", 1), ns)


class Cfg(dict):
    __getattr__ = dict.__getitem__


torch.manual_seed(31)
g = torch.Generator().manual_seed(8)
P = 350
net = ns["VanillaCondMLP"](3, 0, 25, Cfg(n_neurons=128, n_hidden_layers=4, skip_in=[], cond_in=[], multires=0))
with torch.no_grad():
    net.lin4.weight.mul_(6.0)  # logits of useful magnitude
aabb = ns["AABB"](np.array([1.1, 1.9, 0.6], np.float32), np.array([-1.0, -0.2, -0.5], np.float32))
self_ = types.SimpleNamespace(distill=False, lbs_network=net, aabb=aabb, training=False)
self_.softmax = lambda logit: ns["sf_softmax"](self_, logit)
self_.get_forward_transform = lambda xyz, tfs: ns["sf_get_forward_transform"](self_, xyz, tfs)


class Gaussians:
    def __init__(self):
        self._xyz = torch.rand(P, 3, generator=g) * torch.tensor([2.0, 2.0, 1.0]) + torch.tensor([-0.95, -0.15, -0.45])
        self._rotation = torch.randn(P, 4, generator=g)

    get_xyz = property(lambda s: s._xyz)

    def clone(self):
        c = Gaussians.__new__(Gaussians)
        c.__dict__.update(self.__dict__)
        return c

    def set_fwd_transform(self, t):
        self.fwd_transform = t


q, _ = torch.linalg.qr(torch.randn(24, 3, 3, generator=g))
tfs = torch.eye(4).repeat(24, 1, 1)
tfs[:, :3, :3] = q
tfs[:, :3, 3] = 0.2 * torch.randn(24, 3, generator=g)
gs = Gaussians()
with torch.no_grad():
    deformed = ns["sf_forward"](self_, gs, 0, types.SimpleNamespace(bone_transforms=tfs))
    logits = net(aabb.normalize(gs._xyz, sym=True))
    weights = ns["sf_softmax"](self_, logits)
    w24 = ns["sf_softmax"](self_, logits[:, :24])
out = dict(xyz=gs._xyz, rotation=gs._rotation, bone_transforms=tfs, aabb_min=aabb.coord_min, aabb_max=aabb.coord_max, logits=logits,
           weights=weights, weights_softmax24=w24, T_fwd=deformed.fwd_transform, means3D=deformed._xyz, rotation_bar=deformed.rotation_precomp)
for l in range(5):
    out[f"W{l}"], out[f"b{l}"] = getattr(net, f"lin{l}").weight.detach(), getattr(net, f"lin{l}").bias.detach()
np.savez_compressed(os.path.join(HERE, "ref_skinning.npz"), **{k: v.numpy().astype(np.float32) for k, v in out.items()})
print({k: tuple(v.shape) for k, v in out.items() if not k[0] in "Wb"}, float(weights.sum(1).min()), float(weights.max()))
