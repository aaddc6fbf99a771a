mkdir -p gpurun_out
timeout 900 python bench.py --P 1000000 --width 1920 --height 1080 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r1_v11_bench_1m.json 2> gpurun_out/r1_v11_bench_1m.err; echo "exit $?"; tail -c 400 gpurun_out/r1_v11_bench_1m.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r1_v11_bench_1m.json'))
print(d["value"], d["ms_per_step"], d["render_fps"], d["config"]["N_pairs"], d["config"]["E_fwd"])
for k,v in d["stages"].items(): print(k, round(v["ms"]*1e3,1), v.get("achieved"), v.get("unit"), v.get("frac"))
print(d.get("return_opacity_step")); print(d.get("full_iteration")); print(d.get("ref_cuda"))
PY
