"""Four forward+backward rasterizer calls at cfg-A (the command the ncu captures of the blend kernels wrap)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "3dgs-avatar-release_b200"), ROOT, os.path.join(ROOT, "tests")]
import scenes
from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
dev = torch.device("cuda:0")
sc = scenes.avatar_scene(P=50000, W=512, H=512, random_init=True)
cam = sc["cam"]
d = lambda a: torch.as_tensor(np.ascontiguousarray(a)).to(dev)
rs = GaussianRasterizationSettings(512, 512, cam["tanfovx"], cam["tanfovy"], d(sc["bg"]), 1.0, d(cam["world_view_transform"]), d(cam["full_proj_transform"]), 3, d(cam["camera_center"]), False, False)
means = d(sc["means3D"]).requires_grad_(True); cols = d(sc["colors_precomp"]).requires_grad_(True); cov = d(sc["cov3D_precomp"]); op = d(sc["opacities"])
for it in range(4):
    means.grad = None; cols.grad = None
    img, radii = GaussianRasterizer(rs)(means3D=means, means2D=torch.zeros_like(means, requires_grad=True), opacities=op, colors_precomp=cols, cov3D_precomp=cov)
    img.mean().backward()
torch.cuda.synchronize()
print("ok", float(img.mean()))
