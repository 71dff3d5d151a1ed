# closing run of round 2 on one GPU with the final kernels: GPU test suite, smoke, default bench, config 4 / 5 lines,
# launch list and one full ncu capture of k_obs (each ncu pass only after the plain run exited 0)
set -x
O=gpurun_out/final; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -3 $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 600 python bench.py > $O/bench_r02_n1.json 2> $O/bench_r02_n1.err; echo "bench rc=$?"
for c in 2 4 5; do timeout 600 python bench.py --config $c --no-cpu-baseline > $O/bench_r02_cfg${c}_n1.json 2> $O/bench_cfg$c.err; echo "cfg$c rc=$?"; done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $O/ncu_launch.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_obs -s 4 -c 1 -o $O/r02_k_obs -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $O/ncu_obs.log 2>&1
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/final/bench_r02_*n1.json")):
    try:
        d = json.load(open(f)); r = d["roofline"]
        print(f.split("/")[-1], "ms/step", round(d["ms_per_step"], 3), "k_obs", round(r["avg_launch_ms"], 3), "frac", round(r["frac"], 4), "value", f"{d['value']:.3e}", "e2e", f"{d['e2e']['value']:.3e}", "cpu", (d.get("cpu_baseline") or {}).get("value"))
    except Exception as e:
        print(f, "failed", e)
PY
