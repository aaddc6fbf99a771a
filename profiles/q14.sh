mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r1_v8_bench_n2.json 2> gpurun_out/r1_v8_bench_n2.err; echo "n2 exit $?"
tail -c 600 gpurun_out/r1_v8_bench_n2.json; echo; tail -c 300 gpurun_out/r1_v8_bench_n2.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/r1_v8_ref_n2.json 2> gpurun_out/r1_v8_ref_n2.err; echo "ref n2 exit $?"; tail -c 400 gpurun_out/r1_v8_ref_n2.json
