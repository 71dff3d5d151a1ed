# benches of the other configurations + unique-scene sweep + k_agents ncu capture
mkdir -p gpurun_out
for c in 2 4 5; do
  timeout 900 python bench.py --config $c --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_cfg$c.json 2> gpurun_out/bench_cfg$c.err; echo "config $c rc=$?"; tail -2 gpurun_out/bench_cfg$c.err
done
for u in 64 512; do
  timeout 1200 python bench.py --unique-scenes $u --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_u$u.json 2> gpurun_out/bench_u$u.err; echo "unique $u rc=$?"; tail -2 gpurun_out/bench_u$u.err
done
python - <<'PY'
import json
for n in ("cfg2", "cfg4", "cfg5", "u64", "u512"):
    try:
        d = json.load(open(f"gpurun_out/bench_{n}.json"))
        r = d["roofline"]
        print(n, "ms/step", round(d["ms_per_step"], 3), "k_obs", round(r["avg_launch_ms"], 3), "k_agents", round(r["k_agents_avg_ms"], 3), "frac", round(r["frac"], 4),
              "value", round(d["value"] / 1e6, 2), "e2e", round(d["e2e"]["value"] / 1e6, 2), "sat/s", round(d["sat_tests_per_s"] / 1e9, 3), "stats", {k: d["episode_stats"][k] for k in ("episodes_reset", "dones", "goals_reached")})
    except Exception as e:
        print(n, "failed", e)
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_agents -s 4 -c 1 -o gpurun_out/k_agents_cfg3 -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_agents3.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_agents -s 4 -c 1 -o gpurun_out/k_agents_cfg5 -f python bench.py --config 5 --worlds 1024 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_agents5.log 2>&1
ls -la gpurun_out/*.ncu-rep
