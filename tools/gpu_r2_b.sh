# kernel iteration: parity subset (lidar / kNN), lidar mismatch diagnosis on the bench config, work counters, bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py tests/test_gpu_bench_config.py -m gpu -x -q -k "delta_lidar or dense_map or knn_index or golden_rollout or expert_replay_linear or restatement or config3 or ragged or agent_stop" --durations=5 > gpurun_out/pytest_quick.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_quick.log
tail -25 gpurun_out/pytest_quick.log
timeout 300 python tools/diag_lidar.py 12 2>&1 | grep "^step\|mismatching" | cut -c1-170
bash tools/gpu_counters.sh 2>&1 | tail -18
timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_q.json"))
print("ms/step", round(d["ms_per_step"], 3), "k_obs ms", round(d["roofline"]["avg_launch_ms"], 3), "k_agents", round(d["roofline"]["k_agents_avg_ms"], 3), "frac", round(d["roofline"]["frac"], 4), "value", round(d["value"] / 1e6, 2), "e2e", round(d["e2e"]["value"] / 1e6, 2))
PY
