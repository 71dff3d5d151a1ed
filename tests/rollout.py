"""Lock-step rollout of the CUDA engine (through the C ABI / SimManager stand-in) against the
reference Manager (oracle/_ref) on identical scenes and seeded actions."""
from __future__ import annotations

import numpy as np

from helpers import (compare_step, engine_params, random_actions, ref_params, report, snapshot_engine,
                     snapshot_ref)


def run_parity(scenes, pd, steps, S=128, seed=42, check_every=1, names=None, tag=None, allowed=None):
    import torch
    from gpudrive_lab_b200 import madrona_gpudrive as mg
    from oracle.refmgr import RefManager

    model = pd["dynamicsModel"]
    lidar = bool(pd["enableLidar"])
    ref = RefManager(scenes, ref_params(pd))
    sim = mg.SimManager(mg.madrona.ExecMode.CUDA, 0, scenes, engine_params(pd), max_agents=S)
    act_e = sim.action_tensor().to_torch()
    act_r = ref.tensor("action")
    W = len(scenes)
    stats = {}
    kw = {} if names is None else {"names": names}
    compare_step(snapshot_ref(ref, **kw), snapshot_engine(sim, **kw), pd["observationRadius"], lidar,
                 tag="after construction", stats=stats)
    for t in range(steps):
        a = random_actions(t, W, S, model, seed)
        act_r[...] = a
        act_e.copy_(torch.from_numpy(a))
        torch.cuda.synchronize()
        ref.step()
        sim.step()
        if (t + 1) % check_every == 0 or t == steps - 1:
            compare_step(snapshot_ref(ref, **kw), snapshot_engine(sim, **kw), pd["observationRadius"], lidar,
                         tag=f"step {t + 1}", stats=stats)
        # the reference clips actions in place (dynamics.hpp:55-56); so must we
        if model == 1:
            assert np.array_equal(act_r[..., :2], act_e.cpu().numpy()[..., :2]), "clipped action write-back differs"
    if tag is None:
        import inspect
        tag = inspect.stack()[1].function
    report(tag, stats, allowed)
    return ref, sim, stats
