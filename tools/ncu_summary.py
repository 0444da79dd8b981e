#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU needed): key raw metrics per captured launch.
usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep [extra_regex]"""
import csv, io, re, subprocess, sys

KEYS = [
    r"^gpu__time_duration\.sum$", r"^dram__bytes_(read|write)\.sum$", r"^launch__(registers_per_thread|grid_size|block_size|occupancy_limit_\w+|waves_per_multiprocessor|shared_mem_per_block_\w+)$",
    r"^sm__warps_active\.avg\.pct_of_peak_sustained_active$", r"^sm__cycles_elapsed\.(avg|max)$", r"^sm__cycles_active\.avg$",
    r"^smsp__issue_active\.avg\.pct_of_peak_sustained_active$", r"^smsp__inst_executed\.sum$", r"^smsp__thread_inst_executed\.sum$",
    r"^smsp__thread_inst_executed_per_inst_executed\.ratio$", r"^sm__inst_executed_pipe_(fp64|alu|fma|xu|lsu)\.avg\.pct_of_peak_sustained_active$",
    r"^sm__pipe_fp64_cycles_active\.avg\.pct_of_peak_sustained_active$", r"^lts__t_bytes\.sum$", r"^lts__t_sector_hit_rate\.pct$",
    r"^l1tex__t_sector_hit_rate\.pct$", r"^l1tex__t_bytes\.sum$", r"^smsp__average_warp_latency_per_inst_issued\.ratio$",
    r"^smsp__average_warps_issue_stalled_\w+_per_issue_active\.ratio$", r"^sm__throughput\.avg\.pct_of_peak_sustained_elapsed$",
    r"^smsp__warps_eligible\.avg\.per_cycle_active$", r"^smsp__warps_active\.avg\.per_cycle_active$",
    r"^l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum$", r"^smsp__inst_executed_op_(shared|global|local)_(ld|st)\.sum$",
    r"^sm__sass_thread_inst_executed_op_(dfma|dadd|dmul)_pred_on\.sum$", r"^smsp__sass_average_branch_targets_threads_uniform\.pct$",
]

def main():
    rep = sys.argv[1]
    extra = sys.argv[2:] 
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    start = out.find('"ID"')
    rows = list(csv.reader(io.StringIO(out[start:])))
    hdr, units = rows[0], rows[1]
    pats = [re.compile(k) for k in KEYS + extra]
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")]
        print("== launch id %s: %s  grid %s block %s" % (r[0], name[:80], r[hdr.index("Grid Size")], r[hdr.index("Block Size")]))
        for h, u, v in zip(hdr, units, r):
            if any(p.search(h) for p in pats):
                print("  %-90s %-14s %s" % (h, u, v))

if __name__ == "__main__":
    main()
