"""Static size of the hot code per source line from an `ncu --page source --print-source cuda,sass --csv` dump:
for every source line, SASS instructions executed by >= frac of the per-step reference count.
usage: python tools/ncu_hotsize.py dump.csv [frac=0.1] [top=50]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
frac = float(sys.argv[2]) if len(sys.argv) > 2 else 0.1
top = int(sys.argv[3]) if len(sys.argv) > 3 else 50
fname = ""; cur = None; seen = set(); per = {}; src = {}
iI = None
for r in rows:
    if len(r) >= 2 and r[0] == "File Path": fname = r[1].split("/")[-1]
    if len(r) > 5 and r[0] == "Line No": iI = r.index("Instructions Executed"); continue
    if iI is None or len(r) <= iI: continue
    if r[0] != "":
        try: cur = (fname, int(r[0])); src[cur] = r[1]
        except ValueError: cur = None
        continue
    if cur is None or not r[2].startswith("0x") or r[2] in seen: continue
    seen.add(r[2])
    try: ex = int(r[iI])
    except ValueError: ex = 0
    per.setdefault(cur, []).append(ex)
allx = sorted((e for v in per.values() for e in v), reverse=True)
ref = allx[len(allx) // 50]
hot = {k: sum(1 for e in v if e >= frac * ref) for k, v in per.items()}
tot = sum(hot.values())
print("hot (>= %.0f%% of %d) static instructions: %d = %.1f kB" % (100 * frac, ref, tot, tot * 16 / 1024))
for k, n in sorted(hot.items(), key=lambda kv: -kv[1])[:top]:
    print("%-18s %5d  %4d hot / %4d static | %s" % (k[0], k[1], n, len(per[k]), src.get(k, "")[:110]))
