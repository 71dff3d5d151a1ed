"""Static instruction count of one kernel per source line, without a GPU: which lines carry the code footprint
(k_obs is sensitive to it: DESIGN.md section 3). Reads `nvdisasm -g -c` of the cubin extracted from the library.
usage: python tools/static_lines.py [library.so] [kernel name fragment, default k_obsILb0] [top N]"""
import collections, glob, os, re, subprocess, sys, tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpudrive_lab_b200", "libgpudrive_b200.so")
kern = sys.argv[2] if len(sys.argv) > 2 else "k_obsILb0"
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
with tempfile.TemporaryDirectory() as d:
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=d, capture_output=True)
    cubin = max(glob.glob(os.path.join(d, "*.cubin")), key=os.path.getsize)
    txt = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.split("\n")
start = [i for i, l in enumerate(txt) if l.lstrip().startswith(".section") and kern in l and ".text." in l][0]
end = next((i for i in range(start + 1, len(txt)) if txt[i].lstrip().startswith(".section")), len(txt))
cur, cnt = None, collections.Counter()
for l in txt[start:end]:
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
    elif re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", l):
        cnt[cur] += 1
byfile = collections.Counter()
for (f, _), c in cnt.items():
    byfile[f] += c
print("instructions:", sum(cnt.values()), dict(byfile))
for (f, ln), c in cnt.most_common(top):
    print(f"{c:5d}  {f}:{ln}")
