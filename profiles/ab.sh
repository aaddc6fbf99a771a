#!/usr/bin/env bash
# A/B of prebuilt library variants in ONE box: ab.sh TAG lib1.so lib2.so ...   (each is copied over lib/libgsplat_b200.so)
TAG=$1; shift; mkdir -p gpurun_out
for L in "$@"; do
  cp "$L" 3dgs-avatar-release_b200/lib/libgsplat_b200.so
  n=$(basename "$L" .so)
  timeout 300 python bench.py --steps 60 --warmup 10 > gpurun_out/${TAG}_${n}.json 2> gpurun_out/${TAG}_${n}.err || echo "bench $n failed"
done
