"""Per source line of a line range: total stall samples and the top stall reasons (ncu --page source --print-source cuda,sass --csv dump).
usage: python tools/ncu_func_lines.py dump.csv file.cuh first_line last_line"""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1]))); want = sys.argv[2]; lo, hi = int(sys.argv[3]), int(sys.argv[4])
fname = ""; cur = None; hdr = None; seen = set()
samples = collections.Counter(); execd = collections.Counter(); reasons = collections.defaultdict(collections.Counter); src = {}
for r in rows:
    if len(r) >= 2 and r[0] == "File Path": fname = r[1].split("/")[-1]
    if len(r) > 5 and r[0] == "Line No": hdr = r; stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]; continue
    if hdr is None or len(r) < len(hdr): continue
    if r[0] != "":
        try: cur = (fname, int(r[0])); src[cur] = r[1]
        except ValueError: cur = None
        continue
    if cur is None or not r[2].startswith("0x") or r[2] in seen: continue
    seen.add(r[2])
    def iv(x):
        try: return int(x)
        except ValueError: return 0
    samples[cur] += iv(r[hdr.index("# Samples")]); execd[cur] += iv(r[hdr.index("Instructions Executed")])
    for i, h in stall_cols: reasons[cur][h[6:]] += iv(r[i])
tot_all = sum(samples.values()); tot = 0
for k in sorted(samples):
    if k[0] != want or not (lo <= k[1] <= hi): continue
    tot += samples[k]
print("samples in range: %d of %d (%.1f%%)" % (tot, tot_all, 100.0 * tot / max(tot_all, 1)))
for k in sorted(samples):
    if k[0] != want or not (lo <= k[1] <= hi) or samples[k] == 0: continue
    top = ", ".join("%s %d" % (n, v) for n, v in reasons[k].most_common(3) if v)
    print("%5d %6d smp %8d inst | %-42s | %s" % (k[1], samples[k], execd[k], top, src[k].strip()[:90]))
