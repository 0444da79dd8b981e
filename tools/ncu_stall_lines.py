"""Source lines ranked by one stall reason's samples (ncu --page source --print-source cuda,sass --csv dump).
usage: python tools/ncu_stall_lines.py dump.csv stall_long_sb [top]"""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1]))); name = sys.argv[2]; top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
fname = ""; cur = None; hdr = None; agg = collections.Counter(); src = {}; seen = set(); ops = collections.defaultdict(collections.Counter)
for r in rows:
    if len(r) >= 2 and r[0] == "File Path": fname = r[1].split("/")[-1]
    if len(r) > 5 and r[0] == "Line No": hdr = r; continue
    if hdr is None or len(r) < len(hdr): continue
    if r[0] != "":
        try: cur = (fname, int(r[0])); src[cur] = r[1]
        except ValueError: cur = None
        continue
    if cur is None or not r[2].startswith("0x") or r[2] in seen: continue
    seen.add(r[2])
    try: v = int(r[hdr.index(name)])
    except ValueError: v = 0
    agg[cur] += v; ops[cur][r[3].split()[0] if r[3] else "?"] += v
tot = sum(agg.values())
print(name, "samples:", tot)
for k, v in agg.most_common(top):
    print("%-16s %5d %5.1f%% %-28s| %s" % (k[0], k[1], 100 * v / max(tot, 1), ",".join(o for o, _ in ops[k].most_common(3)), src[k][:100]))
