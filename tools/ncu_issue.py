#!/usr/bin/env python3
"""Summarise an ncu report of the rollout kernel into profiles/<name>_issue.json, the file bench.py quotes under
roofline.issue (issue-slot utilisation and lane efficiency: what SURVEY 8d asks for beside the HBM figure).

  python tools/ncu_issue.py gpurun_out/r2c_rollout.ncu-rep profiles/r2_rollout_issue.json "command that was profiled"

Run here (the CPU box has ncu); the capture itself comes from a gpurun call."""
import csv
import io
import json
import subprocess
import sys
import time

rep, out, cmd = sys.argv[1], sys.argv[2], sys.argv[3]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, vals = rows[0], rows[2]
m = dict(zip(hdr, vals))


def f(k):
    return float(m[k].replace(",", ""))


stalls = {k.split("issue_stalled_")[1].split("_per_issue")[0]: f(k) for k in hdr if "issue_stalled" in k and "per_issue_active" in k}
total = sum(v for k, v in stalls.items() if k != "selected")
head = subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
doc = {
    "kernel": m["Kernel Name"], "grid": int(f("launch__grid_size")), "registers": int(f("launch__registers_per_thread")),
    "kernel_ms": f("gpu__time_duration.sum"), "issue_slots_busy": f("smsp__issue_active.avg.pct_of_peak_sustained_active") / 100,
    "threads_per_warp_inst": f("smsp__thread_inst_executed_per_inst_executed.ratio"),
    "warps_active_frac": f("sm__warps_active.avg.pct_of_peak_sustained_active") / 100,
    "warps_per_sm": f("sm__warps_active.avg.pct_of_peak_sustained_active") / 100 * 64,
    "stall_no_instruction_per_issue": stalls["no_instruction"], "stall_no_instruction_share": stalls["no_instruction"] / total,
    "stalls_per_issue": {k: round(v, 3) for k, v in sorted(stalls.items(), key=lambda kv: -kv[1])[:6]},
    "warp_instructions": int(f("smsp__inst_executed.sum")),
    "dram_bytes": int((f("dram__bytes_read.sum") + f("dram__bytes_write.sum")) * {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1}[rows[1][hdr.index("dram__bytes_read.sum")]]),
    "l1_hit_rate": f("l1tex__t_sector_hit_rate.pct") / 100, "l2_hit_rate": f("lts__t_sector_hit_rate.pct") / 100,
    "captured": time.strftime("%Y-%m-%dT%H:%M:%SZ", time.gmtime()), "commit": head, "report": rep, "command": cmd,
    "how": "ncu --set full --clock-control none, one launch; summarised by tools/ncu_issue.py",
}
json.dump(doc, open(out, "w"), indent=1)
print(json.dumps(doc, indent=1))
