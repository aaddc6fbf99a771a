mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_symm_allreduce.py tests/test_gpu_pose.py tests/test_gpu_rasterizer.py tests/test_gpu_optim.py -m gpu -x -q > gpurun_out/r1_v12_pytest_ar.log 2>&1; tail -3 gpurun_out/r1_v12_pytest_ar.log | cut -c1-300
grep -n "Error\|Traceback\|assert\|File \"/tmp\|ok multicast" gpurun_out/symm_allreduce_worker.log | head -20 | cut -c1-300
