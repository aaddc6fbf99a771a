TAG=r1_v10; mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_optim.py -m gpu -q > gpurun_out/${TAG}_pytest_optim.log 2>&1; tail -25 gpurun_out/${TAG}_pytest_optim.log
