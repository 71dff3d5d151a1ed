# kernel experiments: bench each variants/lib_*.so given on the command line (names), short runs
mkdir -p gpurun_out
for v in "$@"; do
  GD_B200_LIB=$PWD/variants/lib_$v.so timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/var_$v.json 2> gpurun_out/var_$v.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/var_$v.json"))
    print("$v", "ms/step", round(d["ms_per_step"], 3), "k_obs ms", round(d["roofline"]["avg_launch_ms"], 3), "e2e", round(d["e2e"]["value"] / 1e6, 2))
except Exception as e:
    print("$v", "failed", e)
PY
done
