"""Generates tests/golden/ref_nonrigid.npz by running the reference's OWN HashGridwithMLP.forward
(models/deformer/non_rigid.py:226-300) with its OWN HierarchicalPoseEncoder, VanillaCondMLP (models/network_utils.py) and AABB
(utils/dataset_utils.py:66-76) on the CPU -- class / method sources lifted out of their files with `ast` and executed
unmodified.  Two pieces of that forward live in third-party packages that are not installed here and are stubbed, both
stated in the test that uses this file:
  * self.hashgrid (tiny-cuda-nn): replaced by a callable returning a SEEDED RANDOM feature matrix [P,32] (saved in the
    fixture), so everything downstream of the encoding is the reference's own arithmetic and the encoding stays unpinned;
  * tf.quaternion_multiply (pytorch3d 0.7.1 transforms): the published definition, quaternion_raw_multiply followed by
    standardize_quaternion (real part made non-negative), restated below.
    python tests/golden/make_nonrigid_golden.py"""
import ast
import os
import textwrap
import types

import numpy as np
import torch
import torch.nn as nn

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"


def lift(path, name, cls=None):
    text = open(os.path.join(REF, path)).read()
    tree = ast.parse(text)
    if cls is None:
        node = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == name)
        return ast.get_source_segment(text, node)
    scope = next(n for n in ast.walk(tree) if isinstance(n, ast.ClassDef) and n.name == cls)
    node = next(n for n in scope.body if isinstance(n, ast.FunctionDef) and n.name == name)
    return textwrap.dedent("\n".join(text.splitlines()[node.lineno - 1:node.end_lineno]))


def quaternion_multiply(a, b):  # pytorch3d.transforms.quaternion_multiply
    aw, ax, ay, az = torch.unbind(a, -1)
    bw, bx, by, bz = torch.unbind(b, -1)
    q = torch.stack((aw * bw - ax * bx - ay * by - az * bz, aw * bx + ax * bw + ay * bz - az * by,
                     aw * by - ax * bz + ay * bw + az * bx, aw * bz + ax * by - ay * bx + az * bw), -1)
    return torch.where(q[..., 0:1] < 0, -q, q)


import math  # noqa: E402

ns = {"torch": torch, "nn": nn, "np": np, "math": math, "tf": types.SimpleNamespace(quaternion_multiply=quaternion_multiply)}
exec(lift("models/network_utils.py", "HierarchicalPoseEncoder"), ns)
exec(lift("models/network_utils.py", "VanillaCondMLP"), ns)
exec(lift("utils/dataset_utils.py", "AABB"), ns)
exec(lift("models/deformer/non_rigid.py", "forward", cls="HashGridwithMLP").replace("def forward(", "def nonrigid_forward(", 1), ns)


class Cfg(dict):
    __getattr__ = dict.__getitem__


torch.manual_seed(21)
P, FEAT = 300, 16
g = torch.Generator().manual_seed(6)
enc = ns["HierarchicalPoseEncoder"](num_joints=24, rel_joints=False, dim_per_joint=6, out_dim=-1)
mlp = ns["VanillaCondMLP"](32, 144, 10 + FEAT, Cfg(n_neurons=128, n_hidden_layers=3, skip_in=[], cond_in=[0], multires=0, last_layer_init=False))
with torch.no_grad():
    mlp.lin3.weight.mul_(0.3)
hash_feat = 0.1 * torch.randn(P, 32, generator=g)
aabb = ns["AABB"](np.array([1.1, 1.9, 0.6], np.float32), np.array([-1.0, -0.2, -0.5], np.float32))
self_ = types.SimpleNamespace(delay=0, feature_dim=FEAT, latent_dim=0, pose_encoder=enc, aabb=aabb, hashgrid=lambda xyz_norm: hash_feat,
                              mlp=mlp, cfg=Cfg(scale_offset="logit", rot_offset="mult"))


class Gaussians:
    def __init__(self):
        self._xyz = torch.rand(P, 3, generator=g) * torch.tensor([2.0, 2.0, 1.0]) + torch.tensor([-0.95, -0.15, -0.45])
        self._scaling = torch.randn(P, 3, generator=g) - 4.0
        self._rotation = torch.randn(P, 4, generator=g)

    get_xyz = property(lambda s: s._xyz)

    def clone(self):
        c = Gaussians.__new__(Gaussians)
        c.__dict__.update(self.__dict__)
        return c


gs = Gaussians()
camera = types.SimpleNamespace(rots=torch.randn(1, 24, 9, generator=g), Jtrs=0.5 * torch.randn(1, 24, 3, generator=g), frame_id=0)
with torch.no_grad():
    deformed, loss_reg = ns["nonrigid_forward"](self_, gs, 100, camera, True)
    cond = enc(camera.rots, camera.Jtrs)
out = dict(xyz=gs._xyz, scaling=gs._scaling, rotation=gs._rotation, rots=camera.rots, Jtrs=camera.Jtrs, hash_feat=hash_feat, cond=cond,
           aabb_min=aabb.coord_min, aabb_max=aabb.coord_max, xyz_out=deformed._xyz, scaling_out=deformed._scaling,
           rotation_out=deformed._rotation, non_rigid_feature=deformed.non_rigid_feature,
           loss=torch.stack([loss_reg["nr_xyz"], loss_reg["nr_scale"], loss_reg["nr_rot"]]))
for l in range(4):
    out[f"W{l}"], out[f"b{l}"] = getattr(mlp, f"lin{l}").weight.detach(), getattr(mlp, f"lin{l}").bias.detach()
out["enc_W0"], out["enc_b0"] = enc.layer_0.weight.detach(), enc.layer_0.bias.detach()
out["enc_Wa"] = torch.stack([l[0].weight.detach() for l in enc.layers]); out["enc_ba"] = torch.stack([l[0].bias.detach() for l in enc.layers])
out["enc_Wb"] = torch.stack([l[2].weight.detach() for l in enc.layers]); out["enc_bb"] = torch.stack([l[2].bias.detach() for l in enc.layers])
np.savez_compressed(os.path.join(HERE, "ref_nonrigid.npz"), **{k: v.numpy().astype(np.float32) for k, v in out.items()})
print({k: tuple(v.shape) for k, v in out.items() if not k.startswith(("W", "b", "enc"))}, [float(v) for v in out["loss"]])
