"""CPU tests (no GPU): the C-ABI library loads and exports every symbol include/gsplat_b200.h declares; the ctypes
structures mirror the C structs; the product path refuses to run without CUDA (no silent CPU fallback)."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "gsplat_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = re.findall(r"^\s*(?:const\s+)?(?:int|size_t|void|char\*?)\s*\*?\s*(\w+)\s*\(", src, flags=re.M)
    return sorted(set(n for n in names if n.startswith(("gs_", "pose_"))))


def test_header_symbols_exported(built_lib):
    lib = ctypes.CDLL(built_lib)
    syms = declared_symbols()
    assert len(syms) >= 14, syms
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/gsplat_b200.h but not exported"
    lib.gs_version.restype = ctypes.c_char_p
    assert b"sm_100a" in lib.gs_version()


def test_workspace_sizes(built_lib):
    from diff_gaussian_rasterization import _C
    L = _C.lib()
    assert L.gs_geom_bytes(1000) >= 1000 * 48 + 256
    assert L.gs_image_bytes(512, 512) >= 512 * 512 * 8 + 1024 * 8
    assert L.gs_list_bytes(10, 512, 512) >= 40
    for cap in (4096, 262144, 1 << 22):
        assert L.gs_list_capacity(L.gs_list_bytes(cap, 512, 512), 512, 512) == cap
    a, b = L.gs_scratch_bytes(1000, 64, 64, 1 << 16), L.gs_scratch_bytes(1000, 64, 64, 1 << 20)
    assert b > a > 0


def test_struct_layout_matches_header(built_lib):
    from diff_gaussian_rasterization import _C
    import avatar_pose
    # 10 x 4-byte scalars + 14 pointers / 9 scalars (+pad) + 22 pointers / 5 scalars (+pad) + 15 pointers
    assert ctypes.sizeof(_C.GsForwardArgs) == 40 + 14 * 8
    assert ctypes.sizeof(_C.GsBackwardArgs) == 40 + 22 * 8
    assert ctypes.sizeof(avatar_pose.PoseArgs) == 24 + 16 * 8
    assert ctypes.sizeof(avatar_pose.PoseBackwardArgs) == ctypes.sizeof(avatar_pose.PoseArgs) + 12 * 8
    # and against the library itself
    from avatar_pose.nonrigid import NonRigidArgs, NonRigidBackwardArgs
    from avatar_pose.skinning import SkinArgs, SkinBackwardArgs
    from avatar_pose.texture import ColorArgs, ColorBackwardArgs
    from avatar_pose.optim import AdamArgs
    from avatar_pose.pose_encoder import PoseEncoderArgs, PoseEncoderBackwardArgs
    L = _C.lib()
    L.gs_sizeof.restype = ctypes.c_size_t
    L.gs_sizeof.argtypes = [ctypes.c_int32]
    for which, st in enumerate((_C.GsForwardArgs, _C.GsBackwardArgs, avatar_pose.PoseArgs, avatar_pose.PoseBackwardArgs,
                                NonRigidArgs, SkinArgs, ColorArgs, AdamArgs, PoseEncoderArgs)):
        assert L.gs_sizeof(which) == ctypes.sizeof(st), st.__name__
    for which, st in ((10, NonRigidBackwardArgs), (11, SkinBackwardArgs), (12, ColorBackwardArgs), (13, PoseEncoderBackwardArgs)):
        assert L.gs_sizeof(which) == ctypes.sizeof(st), st.__name__


def test_no_cpu_fallback(built_lib):
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer, _C
    rs = GaussianRasterizationSettings(16, 16, 1.0, 1.0, torch.zeros(3), 1.0, torch.eye(4), torch.eye(4), 0, torch.zeros(3), False, False)
    r = GaussianRasterizer(rs)
    m = torch.zeros(4, 3)
    with pytest.raises(_C.GsError):
        r(means3D=m, means2D=m, opacities=torch.ones(4, 1), colors_precomp=torch.ones(4, 3), cov3D_precomp=torch.ones(4, 6))
    from avatar_pose import fused_pose
    with pytest.raises(_C.GsError):
        fused_pose(m, torch.ones(4, 4), m, torch.ones(4, 1), torch.ones(4, 1, 3), torch.ones(4, 15, 3), torch.ones(4, 24),
                   torch.eye(4).repeat(24, 1, 1), torch.zeros(3))


def test_argument_validation_messages():
    """Same exceptions as the reference's GaussianRasterizer.forward (__init__.py:191-195)."""
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    rs = GaussianRasterizationSettings(16, 16, 1.0, 1.0, torch.zeros(3), 1.0, torch.eye(4), torch.eye(4), 0, torch.zeros(3), False, False)
    r = GaussianRasterizer(rs)
    m = torch.zeros(4, 3)
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(means3D=m, means2D=m, opacities=m[:, :1])
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(means3D=m, means2D=m, opacities=m[:, :1], shs=torch.zeros(4, 16, 3), colors_precomp=m)
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(means3D=m, means2D=m, opacities=m[:, :1], colors_precomp=m)
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(means3D=m, means2D=m, opacities=m[:, :1], colors_precomp=m, scales=m, rotations=torch.zeros(4, 4), cov3D_precomp=torch.zeros(4, 6))
    assert GaussianRasterizationSettings._fields == ("image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier",
                                                     "viewmatrix", "projmatrix", "sh_degree", "campos", "prefiltered", "debug")
