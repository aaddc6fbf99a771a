mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r1_v11_pytest.log 2>&1; tail -3 gpurun_out/r1_v11_pytest.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r1_v11_bench_n2.json 2> gpurun_out/r1_v11_bench_n2.err; echo "n2 exit $?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r1_v11_bench_n2.json')); print(d["value"], d["n_gpus"], d["ms_per_step"], d["e2e"]["value"], d["render_fps"], d["render_e2e"]["value"])
PY
