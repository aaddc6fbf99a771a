TAG=r1_v9; mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_converter.py tests/test_gpu_color.py -m gpu -x -q > gpurun_out/${TAG}_pytest_conv.log 2>&1; tail -15 gpurun_out/${TAG}_pytest_conv.log
timeout 120 python profiles/color_bench.py 50000 > gpurun_out/${TAG}_color_50k.json 2>gpurun_out/${TAG}_color.err; cat gpurun_out/${TAG}_color_50k.json; tail -3 gpurun_out/${TAG}_color.err
timeout 120 python profiles/color_bench.py 1000000 > gpurun_out/${TAG}_color_1m.json 2>>gpurun_out/${TAG}_color.err; cat gpurun_out/${TAG}_color_1m.json
