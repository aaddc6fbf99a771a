#!/usr/bin/env bash
# ncu --set full capture of kernels matching a regex during a short profile-only bench run.  Usage: ncu_one.sh TAG REGEX [count]
TAG=$1; KRE=$2; CNT=${3:-2}
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"${KRE}" -s 4 -c ${CNT} -o gpurun_out/${TAG}_full -f \
  python bench.py --profile-only --steps 3 --warmup 3 > gpurun_out/${TAG}_ncu2.log 2>&1
tail -3 gpurun_out/${TAG}_ncu2.log
