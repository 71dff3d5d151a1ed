# executed warp instructions and duration of k_obs / k_lidar_fix for each variants/lib_<name>.so given
mkdir -p gpurun_out
for v in "$@"; do
  GD_B200_LIB=$PWD/variants/lib_$v.so timeout 600 ncu --metrics smsp__inst_executed.sum,gpu__time_duration.sum,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:"k_obs|k_lidar_fix" -s 6 -c 2 --csv --log-file gpurun_out/inst_$v.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
  python - "$v" <<'PY'
import csv, sys
v = sys.argv[1]
rows = [r for r in csv.reader(open(f"gpurun_out/inst_{v}.csv")) if len(r) > 10 and r[0].isdigit()]
out = {}
for r in rows:
    out.setdefault(r[4].split("(")[0], {})[r[-3]] = r[-1]
for k, m in out.items():
    print(v, k, {a: m[a] for a in sorted(m)})
PY
done
