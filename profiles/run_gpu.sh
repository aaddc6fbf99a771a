#!/usr/bin/env bash
# One gpurun call: GPU tests, bench, ncu launch list + full captures of the named kernels.  Usage: run_gpu.sh TAG [kernel-regex]
TAG=${1:-rX}; KRE=${2:-blend_}
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${TAG}_pytest.log
tail -5 gpurun_out/${TAG}_pytest.log
timeout 600 python bench.py --steps 100 --warmup 10 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"
tail -c 600 gpurun_out/${TAG}_bench.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/${TAG}_launches.csv \
  python bench.py --profile-only --steps 2 --warmup 2 > gpurun_out/${TAG}_ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"${KRE}" -s 30 -c 16 -o gpurun_out/${TAG}_full -f \
  python bench.py --profile-only --steps 3 --warmup 3 > gpurun_out/${TAG}_ncu2.log 2>&1
ls -la gpurun_out | tail
