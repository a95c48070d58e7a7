#!/usr/bin/env python
"""Extracts the per-launch counters bench.py's roofline needs from an .ncu-rep (ncu --set full) into profiles/r02_ncu_metrics.json.
Usage: ncu_metrics.py report.ncu-rep KEY WORKLOAD UNITS_PER_LAUNCH SOURCE_TXT  (KEY = kernel name bench.py uses, e.g. render_pass_kernel)"""
import csv, io, json, os, subprocess, sys
rep, key, workload, units, source = sys.argv[1], sys.argv[2], sys.argv[3], float(sys.argv[4]), sys.argv[5]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw))); hdr = rows[0]; d = dict(zip(hdr, rows[2]))
num = lambda k: float(d[k].replace(",", "")) if d.get(k) not in (None, "") else None
u = dict(zip(hdr, rows[1]))
def bytes_of(k):
    v = num(k); unit = u.get(k, "")
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1) if v is not None else None
m = {"workload": workload, "units_per_launch": units, "kernel_name": d["Kernel Name"], "duration_under_ncu": num("gpu__time_duration.sum"), "duration_unit": u.get("gpu__time_duration.sum", ""),
     "registers": num("launch__registers_per_thread"), "grid": num("launch__grid_size"),
     "inst_executed": num("smsp__inst_executed.sum"), "thread_inst_executed": num("smsp__thread_inst_executed.sum"),
     "lanes_per_inst": num("smsp__thread_inst_executed_per_inst_executed.ratio"),
     "issue_active_pct": num("smsp__issue_active.avg.pct_of_peak_sustained_active"),
     "dram_bytes": (bytes_of("dram__bytes_read.sum") or 0) + (bytes_of("dram__bytes_write.sum") or 0),
     "lts_bytes": (num("lts__t_sectors.sum") or 0) * 32.0 or bytes_of("lts__t_bytes.sum"),
     "l1_bytes": (num("l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum") or 0) * 32.0, "l2_hit_pct": num("lts__t_sector_hit_rate.pct"), "l1_hit_pct": num("l1tex__t_sector_hit_rate.pct"),
     "long_scoreboard": num("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio"),
     "warps_active_pct": num("sm__warps_active.avg.pct_of_peak_sustained_active"), "source": source,
     "captured": subprocess.run(["git", "-C", ROOT, "log", "-1", "--format=%cs, commit %h"], capture_output=True, text=True).stdout.strip()}
if m["thread_inst_executed"] is None and m["lanes_per_inst"]:
    m["thread_inst_executed"] = m["inst_executed"] * m["lanes_per_inst"]
path = os.path.join(ROOT, "profiles", "r02_ncu_metrics.json")
allm = json.load(open(path)) if os.path.exists(path) else {"_comment": "per-launch ncu counters (ncu --set full --clock-control none) of the dominant kernels at the benchmarked launch size; written by scripts/ncu_metrics.py, read by bench.py"}
allm[f"{key}:{workload}"] = m
json.dump(allm, open(path, "w"), indent=1)
print(json.dumps(m, indent=1))
