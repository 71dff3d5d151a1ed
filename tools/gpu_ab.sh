# A/B of library variants at the full shard (8192 worlds) and at the 8-GPU shard size (1024 worlds) on one box
# usage: bash tools/gpu_ab.sh <variant> <variant> ...
mkdir -p gpurun_out
for pass in 1 2; do
for v in "$@"; do
  for w in 0 1024; do
    GD_B200_LIB=$PWD/variants/lib_$v.so timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --worlds $w > gpurun_out/ab_${v}_$w.json 2> gpurun_out/ab_${v}_$w.err
    python - <<PY
import json
try:
    d = json.load(open("gpurun_out/ab_${v}_$w.json"))
    print("$v", "worlds", d["config"]["total_worlds"], "ms/step", round(d["ms_per_step"], 4), "k_obs ms", round(d["roofline"]["avg_launch_ms"], 4), "e2e", round(d["e2e"]["value"] / 1e6, 2))
except Exception as e:
    print("$v", $w, "failed", e)
PY
  done
done
done
