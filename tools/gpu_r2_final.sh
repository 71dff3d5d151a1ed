# round-2 record run on one GPU: full GPU test suite, bench lines of every configuration, ncu launch list and
# full captures of k_obs / k_agents of the default bench command (each ncu pass only after the plain run exited 0)
set -x
O=gpurun_out/final; mkdir -p $O
timeout 2400 python -m pytest tests -m gpu -x -q --durations=6 > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -5 $O/pytest_gpu.log
timeout 900 python bench.py > $O/bench_r02_n1.json 2> $O/bench_r02_n1.err; echo "bench rc=$?"
for c in 2 4 5; do timeout 900 python bench.py --config $c --no-cpu-baseline > $O/bench_r02_cfg${c}_n1.json 2> $O/bench_cfg$c.err; echo "cfg$c rc=$?"; done
timeout 900 python bench.py --config 4 --separate-policy-obs --no-cpu-baseline > $O/bench_r02_cfg4_separate_n1.json 2> /dev/null
timeout 900 python bench.py --unique-scenes 512 --no-cpu-baseline > $O/bench_r02_u512_n1.json 2> $O/bench_u512.err; echo "u512 rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $O/ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_obs -s 4 -c 1 -o $O/r02_k_obs -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $O/ncu_obs.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_agents -s 4 -c 1 -o $O/r02_k_agents -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $O/ncu_agents.log 2>&1
if [ -n "$WITH_U512_NCU" ]; then timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_obs -s 4 -c 1 -o $O/r02_k_obs_u512 -f python bench.py --unique-scenes 512 --steps 3 --warmup 3 --no-cpu-baseline > $O/ncu_obs512.log 2>&1; fi
# the reference's CPU backend on this box's host cores for the other configurations (config 3 is in the default line)
for c in 2 4 5; do timeout 600 python bench.py --impl reference --config $c --steps 5 --warmup 1 --budget 15 > $O/bench_r02_ref_cfg$c.json 2> $O/ref_cfg$c.err; echo "ref cfg$c rc=$?"; done
ls -la $O
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/final/bench_r02_*.json")):
    try:
        d = json.load(open(f))
        if d.get("impl") == "reference":
            print(f.split("/")[-1], "reference", d["value"], d["cpu_baseline"]["cores"], "cores"); continue
        r = d["roofline"]
        print(f.split("/")[-1], "ms/step", round(d["ms_per_step"], 3), "k_obs", round(r["avg_launch_ms"], 3), "k_agents", round(r["k_agents_avg_ms"], 3), "frac", round(r["frac"], 4), "value", f"{d['value']:.3e}", "e2e", f"{d['e2e']['value']:.3e}", "cpu", (d.get("cpu_baseline") or {}).get("value"))
    except Exception as e:
        print(f, "failed", e)
PY
