# does L1 sharing of the road table between co-resident CTAs matter? same workload with 1..64 distinct scenes
mkdir -p gpurun_out
for u in 64 1 2 8; do
  timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --unique-scenes $u > gpurun_out/uq_$u.json 2> gpurun_out/uq_$u.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/uq_$u.json"))
    print("unique", $u, "ms/step", round(d["ms_per_step"], 4), "k_obs ms", round(d["roofline"]["avg_launch_ms"], 4))
except Exception as e:
    print($u, "failed", e)
PY
done
