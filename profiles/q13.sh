timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum --clock-control none -k regex:'blend_' -s 12 -c 9 --csv --log-file gpurun_out/r1_v8_blend.csv python bench.py --profile-only --steps 2 --warmup 2 > gpurun_out/r1_v8_ncu.log 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/r1_v8_blend.csv')) if len(r)>10]
h=rows[0]; ki=h.index("Kernel Name"); mi=h.index("Metric Name"); vi=h.index("Metric Value"); ii=h.index("ID")
cur={}
for r in rows[1:]:
    cur.setdefault((r[ii], r[ki][:40]), {})[r[mi]]=r[vi]
for k,v in cur.items(): print(k, v)
PY
