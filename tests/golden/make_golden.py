"""Generates the golden fixtures in this directory from the UNMODIFIED reference (oracle/_ref, the
reference Manager in ExecMode::CPU compiled from /root/reference). Run in the build container:

    python tests/golden/make_golden.py

Outputs (all committed):
  womd_test_scene.json.gz   the reference's own test scenario (tests/test.json, data, not source)
  synth_a.json.gz, synth_b.json.gz   two small scenarios from gpudrive_lab_b200.synth
  loader_*.npz              static tensors the reference builds from each scenario (rows a16/a17)
  rollout_*.npz             exported tensors of seeded rollouts at selected steps (rows a1-a15)

The GPU box has no /root/reference: tests read these files only.
"""
from __future__ import annotations

import gzip
import json
import os
import shutil
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from helpers import param_dict, random_actions, ref_params  # noqa: E402

S = 128
SNAP = ["done", "info", "steps_remaining", "reward", "self_observation", "absolute_self_observation",
        "partner_observations", "agent_roadmap", "lidar", "action"]
LOADER = ["shape", "map_observation", "expert_trajectory", "metadata", "controlled_state", "response_type",
          "world_means", "map_name", "scenario_id"]

# name -> (scenes, params, steps, snapshot steps). Step 0 = state after construction.
ROLLOUTS = {
    "delta_lidar": (["synth_a", "synth_b"], dict(dynamicsModel=2, enableLidar=True, collisionBehaviour=2,
                                                  maxNumControlledAgents=16), 30, [0, 1, 12, 30]),
    "bicycle_stop": (["synth_a", "synth_b"], dict(dynamicsModel=1, collisionBehaviour=0,
                                                   maxNumControlledAgents=16), 91, [0, 2, 40, 91]),
    "classic_removed_linear": (["synth_b", "synth_a"], dict(dynamicsModel=0, collisionBehaviour=1,
                                                            roadObservationAlgorithm=1, rewardType=0,
                                                            maxNumControlledAgents=16), 40, [0, 7, 40]),
    "expert_womd": (["womd_test_scene"], dict(dynamicsModel=1, collisionBehaviour=0, maxNumControlledAgents=0,
                                              IgnoreNonVehicles=True, isStaticAgentControlled=False,
                                              polylineReductionThreshold=0.5, observationRadius=10.0, rewardType=0,
                                              distanceToGoalThreshold=1.0), 91, [0, 1, 45, 91]),
    "state_womd": (["womd_test_scene"], dict(dynamicsModel=3, collisionBehaviour=2, maxNumControlledAgents=4,
                                             enableLidar=True, observationRadius=60.0), 6, [0, 3, 6]),
}
LOADERS = {
    "synth_a": dict(), "synth_b": dict(polylineReductionThreshold=0.0),
    "womd_test_scene": dict(polylineReductionThreshold=0.0, initOnlyValidAgentsAtFirstStep=False),
    "womd_test_scene@reduced": dict(polylineReductionThreshold=0.5, IgnoreNonVehicles=True,
                                    isStaticAgentControlled=False, maxNumControlledAgents=2),
}


def write_scene_files():
    from gpudrive_lab_b200 import synth
    shutil.copyfileobj(open("/root/reference/tests/test.json", "rb"),
                       gzip.GzipFile(os.path.join(HERE, "womd_test_scene.json.gz"), "wb", 9, mtime=0))
    for name, idx, kw in (("synth_a", 3, dict(num_agents=8, extent=100.0, spacing=2.0)),
                          ("synth_b", 5, dict(num_agents=12, extent=120.0, spacing=1.5))):
        b = json.dumps(synth.make_scenario(idx, **kw), separators=(",", ":")).encode()
        with gzip.GzipFile(os.path.join(HERE, name + ".json.gz"), "wb", 9, mtime=0) as f:
            f.write(b)


def unpack(names, out_dir):
    paths = []
    for n in names:
        p = os.path.join(out_dir, n + ".json")
        if not os.path.exists(p):
            with gzip.open(os.path.join(HERE, n + ".json.gz"), "rb") as f, open(p, "wb") as g:
                shutil.copyfileobj(f, g)
        paths.append(p)
    return paths


def main():
    from oracle.refmgr import RefManager
    write_scene_files()
    tmp = tempfile.mkdtemp()
    for key, kw in LOADERS.items():
        scene = key.split("@")[0]
        pd = param_dict(**kw)
        ref = RefManager(unpack([scene], tmp), ref_params(pd))
        A, R = (int(x) for x in ref.tensor("shape")[0])
        out = {"params": json.dumps(pd)}
        for n in LOADER:
            t = ref.tensor(n)[0]
            if n == "map_observation":
                t = t[:R]
            elif n in ("expert_trajectory", "metadata", "controlled_state", "response_type"):
                t = t[:A]
            out[n] = np.array(t, copy=True)
        np.savez_compressed(os.path.join(HERE, f"loader_{key.replace('@', '_')}.npz"), **out)
        ref.close()
        print("loader", key, A, R)
    for key, (scenes, kw, steps, snaps) in ROLLOUTS.items():
        pd = param_dict(**kw)
        ref = RefManager(unpack(scenes, tmp), ref_params(pd))
        W = len(scenes)
        shape = ref.tensor("shape").copy()
        Amax = int(shape[:, 0].max())
        out = {"params": json.dumps(pd), "scenes": json.dumps(scenes), "shape": shape,
               "steps": np.array(steps), "snaps": np.array(snaps)}

        def snap(t):
            for n in SNAP:
                if n == "lidar" and not pd["enableLidar"]:
                    continue
                out[f"{n}@{t}"] = ref.tensor(n)[:, :Amax].copy()
        if 0 in snaps:
            snap(0)
        for t in range(steps):
            ref.tensor("action")[...] = random_actions(t, W, S, pd["dynamicsModel"], seed=7)
            ref.step()
            if t + 1 in snaps:
                snap(t + 1)
        np.savez_compressed(os.path.join(HERE, f"rollout_{key}.npz"), **out)
        ref.close()
        print("rollout", key, shape.tolist())
    shutil.rmtree(tmp)


if __name__ == "__main__":
    main()
