#!/usr/bin/env bash
# ncu launch list only (per-kernel durations of a short profile-only bench run).  Usage: launchlist.sh TAG
TAG=${1:-ll}; mkdir -p gpurun_out
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/${TAG}_launches.csv \
  python bench.py --profile-only --steps 2 --warmup 2 > gpurun_out/${TAG}_ncu1.log 2>&1; echo "ncu exit $?"
