"""Generates tests/golden/ref_pose_encoder.npz from the reference's OWN HierarchicalPoseEncoder.  The module that holds it
(models/network_utils.py) imports tinycudann at the top and cannot be imported in this container, so the class definition
is lifted out of the file with `ast` and executed as is (no edits) with torch / nn / np in scope.
    python tests/golden/make_pose_encoder_golden.py"""
import ast
import os

import numpy as np
import torch
import torch.nn as nn

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/models/network_utils.py"
text = open(SRC).read()
node = next(n for n in ast.parse(text).body if isinstance(n, ast.ClassDef) and n.name == "HierarchicalPoseEncoder")
ns = {"torch": torch, "nn": nn, "np": np}
exec(ast.get_source_segment(text, node), ns)
out = {}
for tag, rel in (("abs", False), ("rel", True)):
    torch.manual_seed(7)
    enc = ns["HierarchicalPoseEncoder"](num_joints=24, rel_joints=rel, dim_per_joint=6, out_dim=-1)
    g = torch.Generator().manual_seed(3)
    rots = torch.randn(1, 24, 9, generator=g)
    Jtrs = torch.randn(1, 24, 3, generator=g) * 0.5
    with torch.no_grad():
        y = enc(rots, Jtrs)
    out[f"{tag}_rots"], out[f"{tag}_Jtrs"], out[f"{tag}_out"] = rots.numpy(), Jtrs.numpy(), y.numpy()
    out[f"{tag}_W0"], out[f"{tag}_b0"] = enc.layer_0.weight.detach().numpy(), enc.layer_0.bias.detach().numpy()
    out[f"{tag}_Wa"] = np.stack([l[0].weight.detach().numpy() for l in enc.layers])
    out[f"{tag}_ba"] = np.stack([l[0].bias.detach().numpy() for l in enc.layers])
    out[f"{tag}_Wb"] = np.stack([l[2].weight.detach().numpy() for l in enc.layers])
    out[f"{tag}_bb"] = np.stack([l[2].bias.detach().numpy() for l in enc.layers])
np.savez_compressed(os.path.join(HERE, "ref_pose_encoder.npz"), **out)
print({k: v.shape for k, v in out.items() if k.startswith("abs")})
