"""Kernel diagnosis: list every lidar ray on which the engine and the reference Manager differ on the benchmarked
configuration (tests/test_gpu_bench_config.py's run), with the agent pose, for offline analysis."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
from gpudrive_lab_b200 import madrona_gpudrive as mg, workload as wl
from oracle import refmgr
from helpers import close
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 12
W = 1024
check = [0, 63, 64, 65, 130, 197, 264, 331, 398, 465, 532, 599, 666, 733, 800, 867, 934, 1001, 1023]
pd = wl.params_dict(3)
scenes = wl.world_scenes(wl.make_scenes(3), W)
sim = mg.SimManager(mg.madrona.ExecMode.CUDA, 0, scenes, wl.engine_params(pd), max_agents=128)
ref = refmgr.RefManager([scenes[w] for w in check], refmgr.RefParams(**pd))
idx = torch.as_tensor(check, device="cuda")
n = 0
for t in range(steps):
    a = wl.random_actions(t, W, 128, 2, seed=77)
    ref.tensor("action")[...] = a[check]
    sim.action_tensor().to_torch().copy_(torch.from_numpy(a))
    ref.step(); sim.step()
    lr = ref.tensor("lidar"); le = sim.lidar_tensor().to_torch().index_select(0, idx).cpu().numpy()
    ab = ref.tensor("absolute_self_observation")
    ok = close(le[..., 0], lr[..., 0], 1.0) & (le[..., 1] == lr[..., 1]) & close(le[..., 2:4], lr[..., 2:4], 1.0).all(-1)
    for (w, i, p, k) in np.argwhere(~ok):
        if i >= 64: continue
        n += 1
        print(f"step {t+1} world {check[w]} (scene {check[w] % 64}) agent {i} plane {p} ray {k}: ref {lr[w,i,p,k]} eng {le[w,i,p,k]} pose {ab[w,i,:8]}")
print(n, "mismatching rays")
