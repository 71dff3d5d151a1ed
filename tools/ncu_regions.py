"""Aggregate an ncu source-page csv of k_obs (`ncu -i X.ncu-rep --page source --csv --print-source sass,cuda`) by code
region. Region boundaries are found by anchor strings in the CURRENT source files, so the tool follows edits.
usage: python tools/ncu_regions.py source.csv [agents per launch]"""
import csv, os, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "gpudrive_lab_b200", "csrc")
rows = list(csv.reader(open(sys.argv[1])))
agents = float(sys.argv[2]) if len(sys.argv) > 2 else 524288.0
cur = None; hdr = None; data = {}
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if r[0] == "Line No": hdr = r; continue
    if hdr is None or len(r) < 8 or r[2] != "-": continue
    si = hdr.index("# Samples"); ie = hdr.index("Instructions Executed")
    k = (cur, int(r[0])); v = data.get(k, (0, 0))
    data[k] = (v[0] + int(r[si] or 0), v[1] + int(r[ie] or 0))

M, O = "kernels_obs_main.cuh", "kernels_obs.cuh"
# (region name, file, anchor that starts it); a region ends where the next one of the same file starts
ANCHORS = [
    ("staging/self/abs", M, "k_obs(const SceneDev"),
    ("partner rows", M, "---- partner observations"),
    ("lidar setup + agents", M, "---- lidar set-up"),
    ("group list", M, "---- road groups: kNN candidates + lidar"),
    ("band loop control", M, "// ---- bands of increasing distance"),
    ("knn sweep", M, "if (do_n) {"),
    ("lidar per entity", M, "if (do_l) {"),
    ("lidar write-out", M, "---- lidar write-out"),
    ("select", M, "---- road-map rows"),
    ("emission", M, "// One chunk of 32 candidates"),
    ("best_t/ray_bits/open_for", O, "__device__ __forceinline__ float best_t("),
    ("ray_box fast", O, "// Clip one ray against one box in the box frame"),
    ("ray_box exact", O, "// The same clip with the reference's own arithmetic"),
    ("fast_atan2/ray_range", O, "// atan2 with |error| < 2e-4 rad"),
    ("cull/consider", O, "// Which rays can touch this box?"),
    ("plane_mask", O, "// z-slab membership of the three lidar planes"),
    ("commit", O, "// lidar clip stage: every lane brings one box"),
    ("expand", O, "// The owner of pair j"),
    ("open_rays", O, "// Rays of each plane with no hit closer than"),
    ("flush_rows", O, "// Emit `m` rows held one per lane"),
]
starts = {}
for f in (M, O):
    src = open(os.path.join(CSRC, f)).read().split("\n")
    for name, ff, anchor in ANCHORS:
        if ff != f: continue
        ln = next((i + 1 for i, line in enumerate(src) if anchor in line), None)
        if ln is None:
            print(f"(anchor for '{name}' not found in {f})")
            continue
        starts[(f, name)] = ln
regions = []
for f in (M, O):
    names = sorted([(ln, name) for (ff, name), ln in starts.items() if ff == f])
    for i, (ln, name) in enumerate(names):
        end = names[i + 1][0] - 1 if i + 1 < len(names) else 10 ** 9
        regions.append((name, f, ln, end))

ts = sum(v[0] for v in data.values()) or 1; ti = sum(v[1] for v in data.values()) or 1
used = set()
print(f"warp instructions per agent (sum over source lines): {ti / agents:.0f}")
for name, f, a, b in regions:
    ks = [k for k in data if k[0] == f and a <= k[1] <= b]
    used |= set(ks)
    s = sum(data[k][0] for k in ks); i = sum(data[k][1] for k in ks)
    print(f"{name:26s} {100 * s / ts:5.1f}% smp {100 * i / ti:5.1f}% inst {i / agents:7.0f} inst/agent")
rest = {}
for k, v in data.items():
    if k not in used:
        o = rest.get(k[0], (0, 0)); rest[k[0]] = (o[0] + v[0], o[1] + v[1])
for f, v in sorted(rest.items(), key=lambda x: -x[1][1]):
    print(f"{f:26s} {100 * v[0] / ts:5.1f}% smp {100 * v[1] / ti:5.1f}% inst {v[1] / agents:7.0f} inst/agent")
