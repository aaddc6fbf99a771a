mkdir -p gpurun_out
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29522 profiles/allreduce_bench.py > gpurun_out/r1_v12_allreduce_n8.json 2> gpurun_out/r1_v12_allreduce_n8.err; echo "exit $?"; cat gpurun_out/r1_v12_allreduce_n8.json; tail -2 gpurun_out/r1_v12_allreduce_n8.err | cut -c1-300
