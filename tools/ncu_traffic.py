#!/usr/bin/env python
"""Writes profiles/mh_sweeps_ncu.json from one `ncu --set full` capture of the MH kernel: the numbers bench.py quotes
next to its live timing (DRAM traffic per launch, issue-slot utilisation, warp instructions per launch).
usage: python tools/ncu_traffic.py gpurun_out/x.ncu-rep "<what was launched>" [out.json]"""
import csv, io, json, subprocess, sys

rep, what = sys.argv[1], sys.argv[2]
out_path = sys.argv[3] if len(sys.argv) > 3 else "profiles/mh_sweeps_ncu.json"
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out[out.find('"ID"'):])))
hdr, units, r = rows[0], rows[1], rows[2]
def get(name):
    i = hdr.index(name)
    v = float(r[i].replace(",", ""))
    u = units[i]
    scale = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1.0, "ms": 1e-3, "us": 1e-6, "ns": 1e-9, "s": 1.0}.get(u, 1.0)
    return v * scale
d = {
    "source": rep.split("/")[-1], "kernel": r[hdr.index("Kernel Name")], "launch": what,
    "dram_bytes_per_launch": get("dram__bytes_read.sum") + get("dram__bytes_write.sum"),
    "duration_s_under_ncu": get("gpu__time_duration.sum"),
    "warp_instructions_per_launch": get("smsp__inst_executed.sum"),
    "issue_slot_utilisation_pct": get("smsp__issue_active.avg.pct_of_peak_sustained_active"),
    "fp64_pipe_active_pct": get("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"),
    "l1_hit_pct": get("l1tex__t_sector_hit_rate.pct"), "l2_hit_pct": get("lts__t_sector_hit_rate.pct"),
    "stall_cycles_per_issue": {k: get("smsp__average_warps_issue_stalled_%s_per_issue_active.ratio" % k)
                               for k in ("no_instruction", "long_scoreboard", "wait", "barrier", "branch_resolving", "short_scoreboard", "selected")},
    "registers_per_thread": get("launch__registers_per_thread"),
    "active_lanes_per_instruction": get("smsp__thread_inst_executed_per_inst_executed.ratio"),
    "eligible_warps_per_cycle": get("smsp__warps_eligible.avg.per_cycle_active"),
}
json.dump(d, open(out_path, "w"), indent=1)
print(json.dumps(d, indent=1))
