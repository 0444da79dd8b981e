"""Executed warp instructions and stall samples per FUNCTION (source line -> enclosing __device__/__global__ function) from an
`ncu --page source --print-source cuda,sass --csv` dump.  usage: python tools/ncu_func_share.py dump.csv"""
import csv, collections, sys, os, re
rows = list(csv.reader(open(sys.argv[1])))
fname = ""; cur = None; hdr = None; seen = set()
inst = collections.Counter(); smp = collections.Counter(); thr = collections.Counter()
for r in rows:
    if len(r) >= 2 and r[0] == "File Path": fname = r[1].split("/")[-1]
    if len(r) > 5 and r[0] == "Line No": hdr = r; continue
    if hdr is None or len(r) < len(hdr): continue
    if r[0] != "":
        try: cur = (fname, int(r[0]))
        except ValueError: cur = None
        continue
    if cur is None or not r[2].startswith("0x") or r[2] in seen: continue
    seen.add(r[2])
    def iv(x):
        try: return int(x)
        except ValueError: return 0
    inst[cur] += iv(r[hdr.index("Instructions Executed")]); smp[cur] += iv(r[hdr.index("# Samples")]); thr[cur] += iv(r[hdr.index("Thread Instructions Executed")])
src = {}
root = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "ssps_b200", "csrc")
def func_of(k):
    f, ln = k
    if f not in src:
        p = os.path.join(root, f)
        src[f] = open(p).read().splitlines() if os.path.exists(p) else None
    L = src[f]
    if L is None: return f
    i = min(ln, len(L)) - 1
    while i > 0 and not re.search(r"^\s*(template.*)?(__device__|__global__)", L[i]) and "__device__" not in L[i]: i -= 1
    m = re.search(r"(\w+)\s*\(", L[i].replace("__launch_bounds__(SSPS_BLOCK, SSPS_MIN_CTAS)", "").split("__device__")[-1].split("__global__")[-1])
    return "%s:%s" % (f, m.group(1) if m else i)
fi = collections.Counter(); fs = collections.Counter(); ft = collections.Counter()
for k in inst: fn = func_of(k); fi[fn] += inst[k]; fs[fn] += smp[k]; ft[fn] += thr[k]
ti, ts = sum(fi.values()), sum(fs.values())
print("total warp instructions %d, samples %d" % (ti, ts))
for fn, c in fi.most_common(30):
    print("%-44s inst %5.1f%%  samples %5.1f%%  active lanes %.1f" % (fn, 100.0 * c / ti, 100.0 * fs[fn] / max(ts, 1), ft[fn] / max(c, 1)))
