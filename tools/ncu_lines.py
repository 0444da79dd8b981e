"""Hot source lines of an `ncu --page source --print-source cuda,sass --csv` dump: share of executed warp instructions,
share of stall samples, active threads per instruction.  usage: python tools/ncu_lines.py dump.csv [top]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 50
hdr = None
agg = {}
fname = ""
for r in rows:
    if len(r) >= 2 and r[0] == "File Path": fname = r[1].split("/")[-1]
    if len(r) > 5 and r[0] == "Line No": hdr = r; iA = 2; iS = hdr.index("# Samples"); iI = hdr.index("Instructions Executed"); iT = hdr.index("Thread Instructions Executed"); continue
    if hdr is None or len(r) <= iT or r[iA] != "-": continue
    try:
        key = (fname, int(r[0]), r[1])
        v = agg.setdefault(key, [0, 0, 0])
        v[0] += int(r[iS]); v[1] += int(r[iI]); v[2] += int(r[iT])
    except ValueError:
        pass
totI = sum(v[1] for v in agg.values()); totS = sum(v[0] for v in agg.values())
print("total warp inst", totI, "samples", totS)
for (f, ln, src), v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print("%-18s %4d inst %5.1f%% samp %5.1f%% thr/inst %4.1f | %s" % (f[:18], ln, 100 * v[1] / totI, 100 * v[0] / max(totS, 1), v[2] / max(v[1], 1), src.strip()[:95]))
