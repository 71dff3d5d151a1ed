"""Known-answer and golden-rollout checks shared by the CPU suite (run on the restatement oracle) and the
GPU suite (run on the CUDA engine through the C ABI). `make(scenes, param_dict)` returns an object with
tensor(name) -> numpy [W,S,...], set_actions(ndarray), step(), reset(list)."""
import math

import numpy as np

from golden_util import Rollout, unpack
from helpers import compare_step, engine_params, param_dict, random_actions


def _minstd_uniform(seed):
    """std::default_random_engine(seed) (libstdc++: minstd_rand0) feeding
    std::uniform_real_distribution<float> via generate_canonical<float, 24>."""
    state = [seed % 2147483647 or 1]

    def draw(lo, hi):
        state[0] = (state[0] * 16807) % 2147483647
        c = np.float32(np.float32(state[0] - 1) / np.float32(2147483646.0))
        if c >= 1.0:
            c = np.nextafter(np.float32(1), np.float32(0))
        return np.float32(c * np.float32(hi - lo) + np.float32(lo))
    return draw


def _closed_form(x, y, th, v0, acc, steer, L, dt=0.1):
    """StepBicycleModel, tests/bicyclemodel.cpp:84-100"""
    v = v0 + 0.5 * acc * dt
    beta = math.atan(math.tan(steer) * 0.5)
    w = v * math.cos(beta) * math.tan(steer) / L
    xn, yn = x + v * math.cos(th + beta) * dt, y + v * math.sin(th + beta) * dt
    tn = math.fmod(th + w * dt, 2 * math.pi)
    tn = tn - 2 * math.pi if tn > math.pi else (tn + 2 * math.pi if tn < -math.pi else tn)
    return xn, yn, tn, abs(v0 + acc * dt)


def run_classic_bicycle_closed_form(make, tmp_path):
    """tests/bicyclemodel.cpp:18-242: agent 0 of test.json, Classic model, 10 seeded random steps vs the
    closed form, EPSILON 1e-3 (tests/test_utils.hpp:11)."""
    from gpudrive_lab_b200 import scene_host
    pd = param_dict(polylineReductionThreshold=0.0, observationRadius=100.0, collisionBehaviour=2,
                    initOnlyValidAgentsAtFirstStep=False, dynamicsModel=0, maxNumControlledAgents=10000,
                    isStaticAgentControlled=False, rewardType=0, distanceToGoalThreshold=0.0)
    scenes = unpack(["womd_test_scene"], tmp_path)
    mgr = make(scenes, pd)
    length = scene_host.scene_tensors(scenes[0], engine_params(pd))["agent_static"][0, 0]
    draw = _minstd_uniform(42)
    for _ in range(10):
        ab, so = mgr.tensor("absolute_self_observation")[0, 0], mgr.tensor("self_observation")[0, 0]
        prev = (float(ab[0]), float(ab[1]), float(ab[7]), float(so[0]))
        acc, steer = float(draw(-3.0, 2.0)), float(draw(-0.7, 0.7))
        a = np.zeros((1, 128, 10), np.float32); a[0, 0, 0] = acc; a[0, 0, 1] = steer
        mgr.set_actions(a)
        exp = _closed_form(*prev, acc, steer, float(length))
        mgr.step()
        ab, so = mgr.tensor("absolute_self_observation")[0, 0], mgr.tensor("self_observation")[0, 0]
        assert abs(ab[0] - exp[0]) < 1e-3 and abs(ab[1] - exp[1]) < 1e-3 and abs(ab[7] - exp[2]) < 1e-3
        assert abs(so[0] - exp[3]) < 1e-3


def run_forward_inverse_dynamics(make, model, tmp_path):
    """tests/test_waymax_model.py:5-58 and tests/test_delta_model.py:5-60: feeding the loader's inverse
    action for agent 1 reproduces the expert state at t = 1."""
    from gpudrive_lab_b200 import scene_host
    pd = param_dict(polylineReductionThreshold=0.5, observationRadius=10.0, collisionBehaviour=0, rewardType=0,
                    distanceToGoalThreshold=1.0, distanceToExpertThreshold=1.0, maxNumControlledAgents=2,
                    IgnoreNonVehicles=True, dynamicsModel=model, isStaticAgentControlled=False)
    scenes = unpack(["womd_test_scene"], tmp_path)
    traj = scene_host.scene_tensors(scenes[0], engine_params(pd))["traj"][1]
    pos, vel, head, inv = traj[:182].reshape(91, 2), traj[182:364].reshape(91, 2), traj[364:455], traj[546:].reshape(91, 10)
    mgr = make(scenes, pd)
    ab, so = mgr.tensor("absolute_self_observation")[0, 1], mgr.tensor("self_observation")[0, 1]
    assert np.allclose(ab[:2], pos[0], atol=1e-2) and abs(ab[7] - head[0]) < 1e-2
    assert abs(so[0] - np.linalg.norm(vel[0])) < 1e-2
    a = np.zeros((1, 128, 10), np.float32); a[0, 1, :3] = inv[0, :3]
    mgr.set_actions(a)
    mgr.step()
    ab, so = mgr.tensor("absolute_self_observation")[0, 1], mgr.tensor("self_observation")[0, 1]
    assert np.allclose(ab[:2], pos[1], atol=2e-2), (ab[:2], pos[1])
    # heading bound: 3e-3 in the reference's tests. The unmodified reference itself (oracle/_ref) lands
    # 8.9e-3 away for InvertibleBicycle on this scene (1.564241 vs 1.573158), so its own bound does not hold
    # for that model; the value is pinned instead, the bound kept for DeltaLocal.
    if model == 1:
        assert ab[7] == np.float32(1.564241)
    else:
        assert abs(ab[7] - head[1]) < 3e-3
    assert abs(so[0] - np.linalg.norm(vel[1])) < 1e-3


def run_expert_replay(make, tmp_path):
    """tests/test_expert.py:5-60"""
    pd = param_dict(polylineReductionThreshold=0.5, observationRadius=10.0, collisionBehaviour=0, rewardType=0,
                    distanceToGoalThreshold=1.0, distanceToExpertThreshold=1.0, maxNumControlledAgents=0,
                    IgnoreNonVehicles=True, isStaticAgentControlled=False, dynamicsModel=0)
    mgr = make(unpack(["womd_test_scene"], tmp_path), pd)
    A = int(mgr.tensor("shape")[0, 0])
    for _ in range(200):
        if mgr.tensor("done")[0, :A].all():
            break
        mgr.step()
    info = mgr.tensor("info")[0, :A]
    veh = info[info[:, 4] == 7]
    assert veh[:, 3].sum() == A
    assert veh[:, :3].sum() == 0


# ---------------------------------------------------------------- golden rollouts from the reference
def _snapshot(mgr, like):
    out = {"shape": mgr.tensor("shape")}
    for n in like:
        if n != "shape":
            out[n] = mgr.tensor(n)
    return out


def run_golden_rollout(make, key, tmp_path, require_exact):
    g = Rollout(key)
    pd = g.params
    mgr = make(unpack(g.scenes, tmp_path), pd)
    assert np.array_equal(mgr.tensor("shape"), g.shape)
    lidar = bool(pd["enableLidar"])
    stats = {}
    exact = True
    for t in range(g.steps + 1):
        if t in g.snaps:
            gold = g.snapshot(t)
            mine = g.crop(_snapshot(mgr, gold.keys()))
            compare_step(gold, mine, pd["observationRadius"], lidar, tag=f"{key} step {t}", stats=stats)
            for n in ("absolute_self_observation", "self_observation", "reward", "partner_observations"):
                for w in range(len(g.scenes)):
                    A = int(g.shape[w, 0])
                    exact &= np.array_equal(gold[n][w, :A], mine[n][w, :A])
            if pd["dynamicsModel"] == 1:      # clipped write-back into the action tensor (live slots)
                for w in range(len(g.scenes)):
                    A = int(g.shape[w, 0])
                    assert np.array_equal(gold["action"][w, :A, :2], mine["action"][w, :A, :2])
        if t < g.steps:
            mgr.set_actions(random_actions(t, len(g.scenes), 128, pd["dynamicsModel"], seed=7))
            mgr.step()
    if require_exact:
        # same compiler, same libm, contraction off on both sides: the restatement is bit-identical here
        assert exact, "within tolerance but no longer bit-identical to the reference build"
        assert stats.get("lidar_mismatch", 0) == 0
    return stats
