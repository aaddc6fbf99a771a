#!/usr/bin/env bash
# quick GPU check: tests + bench only.  Usage: quick.sh TAG
TAG=${1:-q}; mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; tail -3 gpurun_out/${TAG}_pytest.log
timeout 600 python bench.py --steps 100 --warmup 10 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"; tail -c 400 gpurun_out/${TAG}_bench.err
