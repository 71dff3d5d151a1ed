"""Diagnostic: per-step max differences engine vs reference (run on the GPU box)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
from helpers import *
from gpudrive_lab_b200 import synth, madrona_gpudrive as mg
from oracle.refmgr import RefManager

model = int(sys.argv[1]) if len(sys.argv) > 1 else 2
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
scenes = synth.write_scenarios("/tmp/synth_diag", 2)
pd = param_dict(dynamicsModel=model, enableLidar=True)
ref = RefManager(scenes, ref_params(pd))
sim = mg.SimManager(mg.madrona.ExecMode.CUDA, 0, scenes, engine_params(pd))
act_e = sim.action_tensor().to_torch(); act_r = ref.tensor("action")
for t in range(steps):
    a = random_actions(t, 2, 128, model)
    act_r[...] = a; act_e.copy_(torch.from_numpy(a)); torch.cuda.synchronize()
    ref.step(); sim.step()
    r = snapshot_ref(ref); e = snapshot_engine(sim)
    A = 64
    ab_r, ab_e = r["absolute_self_observation"][:, :A], e["absolute_self_observation"][:, :A]
    eq = (ab_r == ab_e)
    print(f"step {t+1}: abs bit-equal frac pos={eq[...,0:2].mean():.4f} quat={eq[...,[3,6]].mean():.4f} yaw={eq[...,7].mean():.4f} "
          f"max|dpos|={np.abs(ab_r[...,0:2]-ab_e[...,0:2]).max():.3e} max|dyaw|={np.abs(ab_r[...,7]-ab_e[...,7]).max():.3e}", end=" ")
    so_r, so_e = r["self_observation"][:, :A], e["self_observation"][:, :A]
    print(f"speed eq={(so_r[...,0]==so_e[...,0]).mean():.4f}", end=" ")
    lr, le = r["lidar"][:, :A], e["lidar"][:, :A]
    d = np.abs(lr[..., 0] - le[..., 0]); rel = d / np.maximum(np.abs(lr[..., 0]), 1.0)
    print(f"lidar bit-eq={(lr==le).all(-1).mean():.5f} >1e-5: {(rel>1e-5).sum()} max rel={rel.max():.2e} type diff={(lr[...,1]!=le[...,1]).sum()}", end=" ")
    rm_r, rm_e = r["agent_roadmap"][:, :A], e["agent_roadmap"][:, :A]
    pr, pe = r["partner_observations"][:, :A], e["partner_observations"][:, :A]
    print(f"partner bit-eq={(pr==pe).mean():.5f} max|d|={np.abs(pr-pe).max():.2e}")
