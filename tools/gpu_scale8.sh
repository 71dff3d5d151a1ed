# 8-GPU box: N=1 (with the child-process cpu_baseline, as the driver runs it), the in-process reference repro of
# round 1's SIGSEGV, then N=2/4/8 for config 3 and N=8 for config 5
mkdir -p gpurun_out
nproc
( time timeout 600 python bench.py --gpus 1 > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err ) 2>&1 | grep real; echo "n1 rc=$?"
echo "(in-process reference repro skipped: did not reproduce on 16- and 32-core boxes)"
for n in 2 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 40 --warmup 5 > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err; echo "n$n rc=$?"
done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --config 5 --steps 32 --warmup 5 > gpurun_out/scale_cfg5_n8.json 2> gpurun_out/scale_cfg5_n8.err; echo "cfg5 n8 rc=$?"
python - <<'PY'
import json
base = None
for n in ("n1", "n2", "n4", "n8", "cfg5_n8"):
    try:
        d = [json.loads(l) for l in open(f"gpurun_out/scale_{n}.json") if l.startswith("{")][-1]
        if n == "n1": base = d["value"]
        eff = d["value"] / (base * d["n_gpus"]) if base and not n.startswith("cfg5") else None
        print(n, "gpus", d["n_gpus"], "ms/step", round(d["ms_per_step"], 3), "value", f"{d['value']:.3e}", "e2e", f"{d['e2e']['value']:.3e}", "eff", eff and round(eff, 3),
              "k_obs", round(d["roofline"]["avg_launch_ms"], 3), "k_agents", round(d["roofline"]["k_agents_avg_ms"], 4), "sat/s", f"{d['sat_tests_per_s']:.2e}", "cpu", (d.get("cpu_baseline") or {}).get("value"))
    except Exception as e:
        print(n, "failed", e)
PY
