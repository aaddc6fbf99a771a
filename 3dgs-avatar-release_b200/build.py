"""Builds libgsplat_b200.so (the C-ABI library, include/gsplat_b200.h) in-tree with nvcc for sm_100a.

    python 3dgs-avatar-release_b200/build.py [--force]

nvcc cross-compiles without a GPU; the .so is git-ignored but travels to the GPU box with the snapshot.
"""
import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "lib", "libgsplat_b200.so")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC",
              "-shared"]


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def needs_build():
    if not os.path.exists(LIB):
        return True
    deps = sources() + glob.glob(os.path.join(CSRC, "*.cuh")) + [os.path.join(HERE, "..", "include", "gsplat_b200.h")]
    return any(os.path.getmtime(d) > os.path.getmtime(LIB) for d in deps)


EXT_SRC = os.path.join(CSRC, "ext", "ext.cpp")


def ext_path():
    import sysconfig
    return os.path.join(HERE, "diff_gaussian_rasterization", "_C_ext" + sysconfig.get_config_var("EXT_SUFFIX"))


def build_ext(force=False):
    """The thin PyTorch C++ extension (csrc/ext/ext.cpp -> diff_gaussian_rasterization/_C_ext*.so): g++ against torch's headers,
    linked to libgsplat_b200.so (rpath $ORIGIN/../lib).  In-tree, so that it travels to the GPU box with the snapshot."""
    out = ext_path()
    deps = [EXT_SRC, os.path.join(HERE, "..", "include", "gsplat_b200.h")]
    if not force and os.path.exists(out) and all(os.path.getmtime(d) <= os.path.getmtime(out) for d in deps):
        return out
    import sysconfig
    import torch
    from torch.utils import cpp_extension as ce
    inc = ce.include_paths(device_type="cuda") if "device_type" in ce.include_paths.__code__.co_varnames else ce.include_paths(cuda=True)
    inc += [sysconfig.get_paths()["include"]]
    tlib = os.path.join(os.path.dirname(torch.__file__), "lib")
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-DTORCH_EXTENSION_NAME=_C_ext", "-DTORCH_API_INCLUDE_EXTENSION_H",
           f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}", "-Wno-deprecated-declarations"]
    cmd += [f"-I{i}" for i in inc] + [EXT_SRC, "-o", out, f"-L{tlib}", "-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda", "-ltorch",
                                      "-ltorch_python", f"-L{os.path.dirname(LIB)}", "-lgsplat_b200", "-Wl,-rpath,$ORIGIN/../lib",
                                      f"-Wl,-rpath,{tlib}", "-L/usr/local/cuda/lib64", "-lcudart"]
    env = dict(os.environ)
    env.pop("CC", None), env.pop("CXX", None)
    subprocess.check_call(cmd, env=env)
    return out


def build(force=False, verbose=False):
    lib = build_lib(force, verbose)
    build_ext(force)
    return lib


def build_lib(force=False, verbose=False):
    if not force and not needs_build():
        return LIB
    os.makedirs(os.path.dirname(LIB), exist_ok=True)
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + sources() + ["-o", LIB]
    env = dict(os.environ)
    env.pop("CC", None), env.pop("CXX", None)
    subprocess.check_call(cmd, env=env)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
