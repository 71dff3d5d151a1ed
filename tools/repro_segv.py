"""Round-1 SCALE N=1 SIGSEGV (bench.py --gpus 1 on the 32-core box): the reference Manager (ExecMode::CPU, one pinned
worker per core, 2 x cores worlds) was run INSIDE the CUDA process after the engine. This reproduces that combination
with faulthandler on, for a root cause. usage: python tools/repro_segv.py [worlds]"""
import faulthandler, os, sys
faulthandler.enable(all_threads=True)
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from gpudrive_lab_b200 import madrona_gpudrive as mg, workload as wl
W = 2048
scenes = wl.world_scenes(wl.make_scenes(3), W)
sim = mg.SimManager(mg.madrona.ExecMode.CUDA, 0, scenes, wl.engine_params(wl.params_dict(3)), max_agents=128)
for _ in range(5):
    sim.step()
print("engine ok; cores", len(os.sched_getaffinity(0)), flush=True)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
for worlds in (32, n, 2 * n):
    v, ms, steps, desc, _ = bench.run_reference_sample(3, 6, 1, worlds=worlds, budget_s=10.0)
    print("in-process reference ok:", worlds, "worlds", round(v), "agent-steps/s", flush=True)
