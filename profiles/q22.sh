mkdir -p gpurun_out
for mode in 0 1; do
GS_NCCL_ALLREDUCE=$mode timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2951$mode bench.py --gpus 2 --steps 100 --warmup 10 --no-cpu-baseline > gpurun_out/r1_v12_bench_n2_nccl$mode.json 2> gpurun_out/r1_v12_bench_n2_nccl$mode.err; echo "n2 mode $mode exit $?"
done
python - <<'PY'
import json
for m in (0,1):
    d=json.load(open(f'gpurun_out/r1_v12_bench_n2_nccl{m}.json')); print(m, d["value"], d["ms_per_step"], d["e2e"]["value"], d["config"]["collective"])
PY
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'slab_allreduce|pose_bwd' -c 6 --csv --log-file gpurun_out/r1_v12_ar_ncu.csv --target-processes all python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --profile-only --steps 2 --warmup 2 > gpurun_out/r1_v12_ar_ncu.log 2>&1; grep -c slab gpurun_out/r1_v12_ar_ncu.csv; grep "slab_allreduce\|pose_bwd" gpurun_out/r1_v12_ar_ncu.csv | cut -d, -f5,12- | head -8
