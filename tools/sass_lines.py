"""Static SASS size per source line / function of the MH kernel, without a GPU: nvdisasm -g of the cubin in a built .so.
usage: python tools/sass_lines.py variants/x.so [kernel-substring]   (needs -lineinfo, which every build has)"""
import re, subprocess, sys, tempfile, os, collections
so = os.path.abspath(sys.argv[1]); kern = sys.argv[2] if len(sys.argv) > 2 else "mh_sweeps_kernelI12PhiloxSourceLb0"
d = tempfile.mkdtemp(); subprocess.run(["cuobjdump", "-xelf", "all", so], cwd=d, capture_output=True)
cub = [os.path.join(d, f) for f in os.listdir(d) if f.endswith(".cubin")][0]
out = subprocess.run(["nvdisasm", "-g", "-c", cub], capture_output=True, text=True).stdout.splitlines()
inside = False; cur = None; per = collections.Counter(); n = 0
for ln in out:
    if ln.startswith("//---") and ".text." in ln: inside = kern in ln; continue
    if not inside: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m: cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
    if re.match(r"\s+/\*[0-9a-f]{4,6}\*/", ln): per[cur] += 1; n += 1
print("static instructions:", n, "(%.1f kB)" % (n * 16 / 1024))
src = {}
for f in set(k[0] for k in per if k):
    p = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "ssps_b200", "csrc", f)
    if os.path.exists(p): src[f] = open(p).read().splitlines()
# per function (by scanning back for a __device__ / __global__ line)
fn = collections.Counter()
for (k, c) in per.items():
    if not k or k[0] not in src: fn[k[0] if k else "?"] += c; continue
    L = src[k[0]]; i = min(k[1], len(L)) - 1
    while i > 0 and not re.search(r"__device__|__global__", L[i]): i -= 1
    m = re.search(r"(\w+)\s*\(", L[i][L[i].find("__"):].replace("__device__", "").replace("__forceinline__", "").replace("__noinline__", "").replace("__global__", "").replace("__launch_bounds__(SSPS_BLOCK, SSPS_MIN_CTAS)", ""))
    fn[(k[0], m.group(1) if m else str(i))] += c
for k, c in fn.most_common(40): print("%6d  %s" % (c, k))
if len(sys.argv) > 3:      # per-line listing of one function: python tools/sass_lines.py x.so kernel func
    want = sys.argv[3]
    for (k, c) in sorted(per.items(), key=lambda kv: (kv[0] or ("", 0))):
        if not k or k[0] not in src: continue
        L = src[k[0]]; i = min(k[1], len(L)) - 1; j = i
        while j > 0 and not re.search(r"__device__|__global__", L[j]): j -= 1
        if want in L[j]: print("%5d %4d | %s" % (k[1], c, L[i].strip()[:120]))
