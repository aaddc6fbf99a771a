TAG=r1_v10; mkdir -p gpurun_out
timeout 600 python bench.py --steps 100 --warmup 10 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"; tail -c 300 gpurun_out/${TAG}_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r1_v10_bench.json'))
print(d["value"], d["ms_per_step"], d["render_fps"], d["e2e"]["value"], d["render_e2e"]["value"], d["eager_ms_per_step"])
PY
