"""Kernel experiment: per-agent work counters of k_obs (needs a library built with -DGD_COUNTERS)."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gpudrive_lab_b200 import madrona_gpudrive as mg, _capi, workload as wl
W = int(sys.argv[1]) if len(sys.argv) > 1 else 256
scenes = wl.world_scenes(wl.make_scenes(3), W)
sim = mg.SimManager(mg.madrona.ExecMode.CUDA, 0, scenes, wl.engine_params(wl.params_dict(3)), max_agents=128)
lib = _capi.load()
out = (C.c_ulonglong * 16)()
act = sim.action_tensor().to_torch()
for t in range(10):
    act[:, :, 0].uniform_(-0.5, 1.5); act[:, :, 1].uniform_(-0.2, 0.2); act[:, :, 2].uniform_(-0.05, 0.05)
    sim.step()
lib.gd_debug_counters(out, 1)
sim.step()
lib.gd_debug_counters(out, 0)
n = out[0]
names = ["agents", "groups listed", "groups processed", " with lidar", " with knn", "(box, ray) pairs clipped", "clip-stage calls",
         "-", "in-radius candidates", "clip rounds", "agents over the candidate cap", "k_lidar_fix: road groups reached",
         "k_lidar_fix: boxes listed for the exact clip", "-", "k_lidar_fix: rays re-traced"]
for i, nm in enumerate(names):
    print(f"{nm:28s} {out[i]:12d}  per agent {out[i] / max(n, 1):9.2f}")
