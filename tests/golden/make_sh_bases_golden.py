"""Generates tests/golden/ref_sh_bases.npz by importing the reference's own utils/sh_utils.py (importable on CPU in the
build container): eval_sh_bases(deg, dirs) for deg 0..3 on seeded unit directions.  Pins oracle/color_oracle.sh_bases.
    python tests/golden/make_sh_bases_golden.py"""
import importlib.util
import os

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
spec = importlib.util.spec_from_file_location("ref_sh_utils", "/root/reference/utils/sh_utils.py")
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)
g = torch.Generator().manual_seed(11)
dirs = torch.nn.functional.normalize(torch.randn(256, 3, generator=g))
out = {"dirs": dirs.numpy()}
for deg in range(4):
    out[f"bases_deg{deg}"] = ref.eval_sh_bases(deg, dirs).numpy()
np.savez_compressed(os.path.join(HERE, "ref_sh_bases.npz"), **out)
print({k: v.shape for k, v in out.items()})
