bash profiles/ab.sh r1_ab variants/base.so variants/bwd3.so variants/base.so
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r1_ab_*.json')):
    d=json.load(open(f)); print(f, round(d["value"]), d["ms_per_step"], {k:round(v["ms"]*1e3,1) for k,v in d["stages"].items()})
PY
