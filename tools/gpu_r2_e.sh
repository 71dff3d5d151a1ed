mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "fused_policy or env or expert_replay_linear or dense_map or policy_observation or golden_rollout" 2>&1 | tail -8 | cut -c1-220
bash tools/gpu_variants.sh prev cur
for extra in "" "--separate-policy-obs"; do
  timeout 900 python bench.py --config 4 --steps 30 --warmup 5 --no-cpu-baseline $extra > gpurun_out/bench_cfg4.json 2> gpurun_out/bench_cfg4.err; tail -1 gpurun_out/bench_cfg4.err
  python -c "
import json; d=json.load(open('gpurun_out/bench_cfg4.json')); r=d['roofline']; print('cfg4 $extra ms/step', round(d['ms_per_step'],3), 'k_obs', round(r['avg_launch_ms'],3), 'frac', round(r['frac'],3), 'value', round(d['value']/1e6,2), 'e2e', round(d['e2e']['value']/1e6,2), d['episode_stats']['episodes_reset'])"
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches1024.csv python bench.py --worlds 1024 --steps 3 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
tail -6 gpurun_out/launches1024.csv | cut -d, -f5,14-
