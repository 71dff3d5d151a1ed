"""The benchmark workloads of BASELINE.json (configs 2-5) as data: one place that says which scenes, which
Parameters and which actions a configuration means, shared by bench.py, the parity tests that check the very
configuration that is timed, and tools/. Host side only (numpy); nothing here touches the oracle.

SURVEY.md section 8(d): synthetic Waymo-shaped scenes (gpudrive_lab_b200.synth), 64 controlled vehicles per
world (128 in the collision-stress config), S = 128 agent slots like the reference's kMaxAgentCount.
"""
from __future__ import annotations

import os
import tempfile

import numpy as np

SLOTS = 128               # consts::kMaxAgentCount of the reference -> identical tensor shapes
K_ROAD = 200
LIDAR = 50

# enum values of src/init.hpp:76-109
BICYCLE, DELTA_LOCAL, CLASSIC, STATE = 1, 2, 0, 3
STOP, REMOVE, IGNORE = 0, 1, 2

CONFIGS = {
    2: dict(name="config 2", worlds=1024, agents=64, unique_scenes=64, lidar=False, dynamics=BICYCLE,
            collision=IGNORE, road_algorithm=0, scene_kw={},
            text="1024 worlds x 64 agents (synthetic Waymo-shaped scenes, ~2050 road entities), InvertibleBicycle "
                 "dynamics, self+partner+roadmap(K=200) observations, no lidar, collisions Ignore"),
    3: dict(name="config 3", worlds=8192, agents=64, unique_scenes=64, lidar=True, dynamics=DELTA_LOCAL,
            collision=IGNORE, road_algorithm=0, scene_kw={},
            text="8192 worlds x 64 agents (synthetic Waymo-shaped scenes, ~2050 road entities), DeltaLocal "
                 "dynamics, self+partner+roadmap(K=200)+lidar(3x50) observations, collisions Ignore"),
    4: dict(name="config 4", worlds=4096, agents=64, unique_scenes=512, lidar=False, dynamics=CLASSIC,
            collision=IGNORE, road_algorithm=1, scene_kw={},
            text="4096 worlds x 64 agents, PPO-style rollout with the env defaults (Classic dynamics, discrete "
                 "action table, linear road algorithm, collisions Ignore): step + device-side auto-reset of finished "
                 "worlds from a 512-scenario device pool + policy-observation read-out, 91-step episodes"),
    5: dict(name="config 5", worlds=8192, agents=128, unique_scenes=8, lidar=False, dynamics=BICYCLE,
            collision=STOP, road_algorithm=0, scene_kw=dict(stress=True, edge_spacing=0.25), threshold=0.0,
            # AgentStop removes a collided agent one step later and these scenes collide within a few steps: the
            # benchmark resets every world each 8 steps (inside the timed region) so that the agents it times are alive
            episode_steps=8,
            text="collision stress: 8192 worlds x 128 agents packed into a double intersection, ~4000 road-edge "
                 "segments per world (polyline reduction off), InvertibleBicycle dynamics, AgentStop, "
                 "self+partner+roadmap observations"),
}


def params_dict(config: int = 3) -> dict:
    """Parameters (src/init.hpp:111-127) of a configuration, as the keyword dict both sides are built from."""
    c = CONFIGS[config]
    # GD_BENCH_NO_LIDAR / GD_BENCH_NO_CLASSIC: kernel experiments only (time split of k_obs); a bench line is only
    # valid without them
    lidar = c["lidar"] and not os.environ.get("GD_BENCH_NO_LIDAR")
    return dict(polylineReductionThreshold=c.get("threshold", 0.1), observationRadius=50.0, rewardType=1,
                distanceToGoalThreshold=2.0, distanceToExpertThreshold=0.0, collisionBehaviour=c["collision"],
                maxNumControlledAgents=c["agents"], IgnoreNonVehicles=False,
                roadObservationAlgorithm=c["road_algorithm"], initOnlyValidAgentsAtFirstStep=True,
                isStaticAgentControlled=True, enableLidar=bool(lidar),
                disableClassicalObs=bool(os.environ.get("GD_BENCH_NO_CLASSIC")), dynamicsModel=c["dynamics"],
                readFromTracksToPredict=False)


def engine_params(d: dict):
    """madrona_gpudrive.Parameters from a keyword dict (reward fields live in rewardParams, src/init.hpp:83-88)."""
    from . import madrona_gpudrive as mg
    p = mg.Parameters()
    for k, v in d.items():
        if k == "rewardType":
            p.rewardParams.rewardType = v
        elif k in ("distanceToGoalThreshold", "distanceToExpertThreshold"):
            setattr(p.rewardParams, k, v)
        else:
            setattr(p, k, v)
    return p


def scene_dir() -> str:
    d = os.environ.get("GD_BENCH_SCENES", os.path.join(tempfile.gettempdir(), "gd_bench_scenes"))
    os.makedirs(d, exist_ok=True)
    return d


def make_scenes(config: int = 3, count: int | None = None) -> list[str]:
    """The distinct scenario files of a configuration (written once, seed = 0x5EED0000 + index). More than 1 024 of
    them (the distinct-scene measurements) are kept as binary scenario-pool files of 256 scenes each - generated,
    converted on all host cores (gd_pool_convert) and the JSON deleted chunk by chunk, so 8 192 scenes take
    3 GB of scratch instead of 8.6 GB and load in seconds instead of minutes; the references "<pool>#<i>" build
    bit-identical tensors (tests/test_cpu_pool.py)."""
    from . import synth
    c = CONFIGS[config]
    count = count if count is not None else c["unique_scenes"]
    kw = dict(num_agents=c["agents"], **c["scene_kw"])
    if count <= 1024:
        return synth.write_scenarios(scene_dir(), count, **kw)
    import shutil
    from . import scene_host
    thr = c.get("threshold", 0.1)
    tag = "_".join(f"{k}{v}" for k, v in sorted(kw.items()))
    refs = []
    for start in range(0, count, 256):
        n = min(256, count - start)
        pool = os.path.join(scene_dir(), f"pool_{tag}_thr{thr}_{start:06d}_{n}.gdpool")
        if not os.path.exists(pool):
            tmp = os.path.join(scene_dir(), f"chunk_{os.getpid()}_{start}")
            paths = synth.write_scenarios(tmp, n, start_index=start, **kw)
            scene_host.convert_pool(paths, pool + ".tmp", thr)
            os.replace(pool + ".tmp", pool)
            shutil.rmtree(tmp, ignore_errors=True)
        refs += [f"{pool}#{i}" for i in range(n)]
    return refs


def world_scenes(unique: list[str], total_worlds: int, rank: int = 0, world_size: int = 1) -> list[str]:
    """Scenario of every world of a rank: global world i runs unique[i % len(unique)] (shard.scenes_for_rank)."""
    from . import shard
    return shard.scenes_for_rank(unique, total_worlds, rank, world_size)


def random_actions(step: int, W: int, S: int, model: int, seed: int = 42) -> np.ndarray:
    """Seeded per (seed, step) action block [W,S,10] (ranges mirror tests/bicyclemodel.cpp:73-75 and
    SURVEY.md 8(d) configs 2-3)."""
    rng = np.random.Generator(np.random.PCG64([seed, step]))
    a = np.zeros((W, S, 10), np.float32)
    if model in (0, 1):      # Classic / InvertibleBicycle: acceleration, steering, head angle
        a[..., 0] = rng.uniform(-3.0, 2.0, (W, S))
        a[..., 1] = rng.uniform(-0.7, 0.7, (W, S))
        a[..., 2] = rng.uniform(-0.3, 0.3, (W, S))
    elif model == 2:         # DeltaLocal: dx, dy, dyaw
        a[..., 0] = rng.uniform(-0.5, 1.5, (W, S))
        a[..., 1] = rng.uniform(-0.2, 0.2, (W, S))
        a[..., 2] = rng.uniform(-0.05, 0.05, (W, S))
    else:                    # State: position(3), yaw, velocity(6)
        a[..., 0] = rng.uniform(-100, 100, (W, S))
        a[..., 1] = rng.uniform(-100, 100, (W, S))
        a[..., 2] = 1.0
        a[..., 3] = rng.uniform(-3, 3, (W, S))
        a[..., 4] = rng.uniform(-10, 10, (W, S))
        a[..., 5] = rng.uniform(-10, 10, (W, S))
    return a


def algorithmic_bytes(S: int, R: float, A: float, lidar: bool = True, policy_obs_floats: int = 0):
    """compulsory HBM bytes per agent-step, (whole step, the k_obs share)."""
    partner = 36 * (S - 1)
    roadmap = 36 * K_ROAD
    lid = 48 * LIDAR if lidar else 0
    tables = (36 * R + 64 * A) / A
    step = partner + roadmap + lid + 88 + 32 + 104 + 40 + tables + 4 * policy_obs_floats
    obs = partner + roadmap + lid + 88 + 36 + 4 + tables + 4 * policy_obs_floats
    return step, obs
