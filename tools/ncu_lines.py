"""Aggregate an `ncu --page source --print-source sass,cuda --csv` dump per CUDA source line."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur = None; data = []; hdr = None
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if r[0] == "Line No": hdr = r; continue
    if hdr is None or r[0] in ("Function Name",): continue
    if len(r) < 8 or r[2] != "-": continue      # per-line summary rows have "-" as address
    try:
        si = hdr.index("# Samples"); ie = hdr.index("Instructions Executed")
        data.append((int(r[si] or 0), int(r[ie] or 0), cur, r[0], r[1].strip()[:100]))
    except ValueError:
        pass
tot = sum(d[0] for d in data) or 1; toti = sum(d[1] for d in data) or 1
print(f"total samples {tot}, warp instructions {toti}")
for d in sorted(data, reverse=True)[:top]:
    print(f"{100*d[0]/tot:5.1f}% smp {100*d[1]/toti:5.1f}% inst  {d[2]}:{d[3]}  {d[4]}")
