"""Turns a gpurun_out/<TAG>_full.ncu-rep + <TAG>_launches.csv into the committed summaries under profiles/:
<OUT>_kernels.json (per kernel: duration, dram bytes, issue/warp activity, registers, top stall reasons) and
<OUT>_launches.csv (the ncu launch list).   python profiles/summarize.py TAG OUT"""
import csv, json, subprocess, sys, shutil, os
tag, out = sys.argv[1], sys.argv[2]
here = os.path.dirname(os.path.abspath(__file__)); root = os.path.dirname(here)
rep = os.path.join(root, "gpurun_out", f"{tag}_full.ncu-rep")
txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines())); hdr = rows[0]
keep = {"gpu__time_duration.sum": "duration_us", "dram__bytes_read.sum": "dram_read", "dram__bytes_write.sum": "dram_write",
        "launch__grid_size": "grid", "launch__block_size": "block", "launch__registers_per_thread": "regs",
        "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_active_pct", "sm__warps_active.avg.pct_of_peak_sustained_active": "warps_active_pct",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm_throughput_pct", "sm__cycles_active.avg": "sm_cycles_active_avg",
        "sm__cycles_elapsed.avg": "sm_cycles_elapsed_avg", "smsp__inst_executed.sum": "warp_insts",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active": "tensor_pipe_active_pct",
        "lts__t_sectors_srcunit_tex_op_red.sum": "l2_red_sectors", "smsp__inst_executed_op_global_red.sum": "red_insts"}
units = {h: u for h, u in zip(hdr, rows[1])}
res = []
for r in rows[2:]:
    d = dict(zip(hdr, r))
    e = {"kernel": d["Kernel Name"].split("(")[0].replace("<unnamed>::", "")}
    for k, n in keep.items():
        if k in d and d[k] not in ("", "n/a"):
            try: e[n] = float(d[k])
            except ValueError: e[n] = d[k]
            if k.startswith("dram__bytes"): e[n + "_unit"] = units.get(k, "")
    st = {k: float(v) for k, v in d.items() if "issue_stalled" in k and k.endswith("_per_issue_active.ratio") and v not in ("", "n/a")}
    e["top_stalls_per_issue"] = {k.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""): round(v, 2)
                                 for k, v in sorted(st.items(), key=lambda x: -x[1])[:4]}
    res.append(e)
json.dump(res, open(os.path.join(here, f"{out}_kernels.json"), "w"), indent=1)
src = os.path.join(root, "gpurun_out", f"{tag}_launches.csv")
if os.path.exists(src): shutil.copy(src, os.path.join(here, f"{out}_launches.csv"))
for e in res: print(e["kernel"], e.get("duration_us"), e.get("dram_read"), e.get("dram_read_unit"), e.get("dram_write"), e.get("dram_write_unit"))
