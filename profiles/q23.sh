mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_symm_allreduce.py -m gpu -x -q 2>&1 | tail -2
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 profiles/allreduce_bench.py > gpurun_out/r1_v12_allreduce_n2.json 2> gpurun_out/r1_v12_allreduce_n2.err; echo "exit $?"; cat gpurun_out/r1_v12_allreduce_n2.json; tail -1 gpurun_out/r1_v12_allreduce_n2.err | cut -c1-300
