"""A/B of forward-blend build variants at cfg-A: python profiles/fwd_variants.py lib1.so lib2.so ...  (each in a child process)"""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import os, sys, json, ctypes
import numpy as np, torch
ROOT = sys.argv[1]
sys.path[:0] = [os.path.join(ROOT, "3dgs-avatar-release_b200"), ROOT, os.path.join(ROOT, "tests")]
import scenes
from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer, _C
dev = torch.device("cuda:0")
sc = scenes.avatar_scene(P=50000, W=512, H=512, random_init=True)
cam = sc["cam"]
d = lambda a: torch.as_tensor(np.ascontiguousarray(a)).to(dev)
rs = GaussianRasterizationSettings(512, 512, cam["tanfovx"], cam["tanfovy"], d(sc["bg"]), 1.0, d(cam["world_view_transform"]), d(cam["full_proj_transform"]), 3, d(cam["camera_center"]), False, False)
means = d(sc["means3D"]).requires_grad_(True); cols = d(sc["colors_precomp"]).requires_grad_(True); cov = d(sc["cov3D_precomp"]); op = d(sc["opacities"])
L = _C.lib()
L.gs_profile_enable.argtypes = [ctypes.c_int]; L.gs_profile_read.argtypes = [ctypes.c_void_p]
L.gs_profile_enable(1)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
acc = np.zeros(7); buf = (ctypes.c_float * 7)(); n = 0
for it in range(25):
    means.grad = None; cols.grad = None
    flush.zero_()
    img, radii = GaussianRasterizer(rs)(means3D=means, means2D=torch.zeros_like(means, requires_grad=True), opacities=op, colors_precomp=cols, cov3D_precomp=cov)
    img.mean().backward()
    torch.cuda.synchronize()
    L.gs_profile_read(ctypes.cast(buf, ctypes.c_void_p))
    if it >= 5:
        acc += np.array(list(buf)); n += 1
names = ["preprocess_fwd", "binning", "blend_fwd", "blend_bwd", "preprocess_bwd"]
print(json.dumps({k: round(float(acc[i] / n) * 1e3, 1) for i, k in enumerate(names)}))
'''
for lib in sys.argv[1:]:
    env = dict(os.environ, GS_LIB=os.path.abspath(lib))
    r = subprocess.run([sys.executable, "-c", CHILD, ROOT], env=env, capture_output=True, text=True, timeout=600)
    print(os.path.basename(lib), (r.stdout.strip().splitlines() or ["?"])[-1], r.stderr[-300:] if r.returncode else "")
