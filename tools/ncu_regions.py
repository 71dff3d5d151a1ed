"""Aggregate an ncu source-page csv of k_obs by code region (kernel experiments)."""
import csv, sys
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
regions = [("staging/self/abs", M, 1, 73), ("partner rows", M, 74, 105), ("lidar setup + agents", M, 106, 153),
           ("group list", M, 154, 204), ("band loop control", M, 205, 237), ("knn sweep", M, 238, 256),
           ("lidar per entity", M, 257, 280), ("lidar write-out", M, 281, 311), ("select", M, 312, 473),
           ("emission", M, 474, 520), ("best_t/ray_bits/open_for", O, 76, 89), ("ray_box fast", O, 90, 120),
           ("ray_box exact", O, 121, 151), ("fast_atan2/ray_range", O, 152, 176), ("consider", O, 177, 202),
           ("plane_mask", O, 203, 217), ("expand", O, 218, 277), ("open_rays", O, 278, 292), ("flush_rows", O, 293, 320)]
ts = sum(v[0] for v in data.values()); ti = sum(v[1] for v in data.values())
used = set()
print(f"warp instructions per agent: {ti / agents:.0f}")
for name, f, a, b in regions:
    ks = [k for k in data if k[0] == f and a <= k[1] <= b]
    used |= set(ks)
    s = sum(data[k][0] for k in ks); i = sum(data[k][1] for k in ks)
    print(f"{name:24s} {100 * s / ts:5.1f}% smp {100 * i / ti:5.1f}% inst {i / agents:7.0f} inst/agent")
rest = {}
for k, v in data.items():
    if k not in used:
        o = rest.get(k[0], (0, 0)); rest[k[0]] = (o[0] + v[0], o[1] + v[1])
for f, v in sorted(rest.items(), key=lambda x: -x[1][1]):
    print(f"{f:24s} {100 * v[0] / ts:5.1f}% smp {100 * v[1] / ti:5.1f}% inst {v[1] / agents:7.0f} inst/agent")
