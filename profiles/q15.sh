TAG=r1_v9; mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; tail -4 gpurun_out/${TAG}_pytest.log
timeout 600 python bench.py --steps 100 --warmup 10 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"; tail -c 300 gpurun_out/${TAG}_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r1_v9_bench.json'))
print(d["value"], d["ms_per_step"], d["render_fps"], {k:round(v["ms"]*1e3,1) for k,v in d["stages"].items()})
print(d.get("return_opacity_step")); print(d.get("full_iteration")); print(d["e2e"]["value"])
PY
