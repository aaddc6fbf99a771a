TAG=r1_v7; mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_opacity_pass.py -m gpu -x -q > gpurun_out/${TAG}_pytest_op.log 2>&1; tail -15 gpurun_out/${TAG}_pytest_op.log
timeout 500 python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest.log 2>&1; tail -5 gpurun_out/${TAG}_pytest.log
timeout 600 python bench.py --steps 100 --warmup 10 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"; tail -c 400 gpurun_out/${TAG}_bench.err
