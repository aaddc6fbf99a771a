#!/usr/bin/env bash
# Builds the UNMODIFIED reference rasterizer (submodules/diff-gaussian-rasterization) for sm_100 into the
# git-ignored baseline/_ref/ (GPU-side A/B baseline + generator of tests/golden/ref_cuda_*.npz).
# The only patch is `#include <cstdint>` in cuda_rasterizer/rasterizer_impl.h, without which gcc 13 fails
# (SURVEY.md 8(c)).  The build happens in a /tmp copy because /root/reference is read-only.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
SRC=/root/reference/submodules/diff-gaussian-rasterization
[ -d "$SRC" ] || { echo "no reference tree at $SRC (GPU box?): keeping prebuilt baseline/_ref"; exit 0; }
TMP=$(mktemp -d /tmp/dgr_ref.XXXXXX)
cp -r "$SRC"/. "$TMP"/
sed -i '0,/#include/s//#include <cstdint>\n#include/' "$TMP/cuda_rasterizer/rasterizer_impl.h"
(cd "$TMP" && TORCH_CUDA_ARCH_LIST=10.0 MAX_JOBS=8 python setup.py build_ext --inplace > "$TMP/build.log" 2>&1) || { tail -30 "$TMP/build.log"; exit 1; }
rm -rf "$HERE/_ref"
mkdir -p "$HERE/_ref"
cp -r "$TMP/diff_gaussian_rasterization" "$HERE/_ref/"
ls -la "$HERE/_ref/diff_gaussian_rasterization"
rm -rf "$TMP"
