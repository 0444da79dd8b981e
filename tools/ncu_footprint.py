"""Static SASS footprint of a kernel from an `ncu --page source --print-source cuda,sass --csv` dump: instructions per source
region, and how many of them are hot (executed by at least `frac` of the reference count = the most executed instruction / 2).
usage: python tools/ncu_footprint.py dump.csv"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
fname = ""; cur = None; seen = set(); per = {}
iI = None
for r in rows:
    if len(r) >= 2 and r[0] == "File Path": fname = r[1].split("/")[-1]
    if len(r) > 5 and r[0] == "Line No": iI = r.index("Instructions Executed"); continue
    if iI is None or len(r) <= iI: continue
    if r[0] != "": 
        try: cur = (fname, int(r[0]))
        except ValueError: cur = None
        continue
    if cur is None or not r[2].startswith("0x"): continue
    if r[2] in seen: continue
    seen.add(r[2])
    try: ex = int(r[iI])
    except ValueError: ex = 0
    per.setdefault(cur, []).append(ex)
allx = sorted((e for v in per.values() for e in v), reverse=True)
ref = allx[len(allx) // 50]  # robust "once per step" count
print("static SASS instructions:", len(allx), "(%.0f kB)" % (len(allx) * 16 / 1024), " reference exec count:", ref)
for frac in (0.5, 0.1, 0.01):
    n = sum(1 for e in allx if e >= frac * ref)
    print("  executed >= %4.0f%% of steps: %5d instr (%.0f kB)" % (100 * frac, n, n * 16 / 1024))
# by file/100-line bucket
bk = {}
for (f, ln), v in per.items():
    b = (f, ln // 50 * 50)
    t = bk.setdefault(b, [0, 0])
    t[0] += len(v); t[1] += sum(1 for e in v if e >= 0.1 * ref)
for b in sorted(bk): print("  %-18s lines %4d-%4d: %5d static, %5d hot" % (b[0][:18], b[1], b[1] + 49, bk[b][0], bk[b][1]))
