bash profiles/run_gpu.sh r1_v7 'blend_|pose_|preprocess|bin_place|onesweep'
timeout 300 ncu --set full --clock-control none --import-source on -k regex:'color_fwd|adam_step|nr_fwd|sk_fwd' -c 6 -o gpurun_out/r1_v7_mlp_full -f \
  python profiles/color_bench.py 50000 > gpurun_out/r1_v7_ncu3.log 2>&1
tail -3 gpurun_out/r1_v7_ncu3.log
python - <<'PY'
import json
d=json.load(open('gpurun_out/r1_v7_bench.json'))
for k in ("value","ms_per_step","render_fps","return_opacity_step","full_iteration","clocks","roofline"): print(k, d.get(k))
print(d["e2e"])
PY
