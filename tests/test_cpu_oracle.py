"""Pins the CPU restatement oracle (oracle/step_oracle.cpp):
  * against the reference's own known-answer tests (SAT literals, ego-frame transform, the Classic
    bicycle closed form, forward o inverse dynamics, expert replay => all goals, zero collisions),
  * against golden rollouts recorded from the unmodified reference Manager (tests/golden/rollout_*.npz),
  * live against oracle/_ref where it is built.
No GPU involved."""
import numpy as np
import pytest

import kat

from golden_util import ROLLOUT_KEYS, Rollout, unpack
from helpers import (TENSORS_CMP, compare_step, engine_params, param_dict, random_actions, ref_params,
                     snapshot_ref)

PI_F = np.float32(3.14159265358979323846)


def to_radians(deg):          # madrona::math::toRadians, math.inl:7-11
    return np.float32(np.float32(PI_F / np.float32(180.0)) * np.float32(deg))


# ---------------------------------------------------------------- reference unit tests restated
def test_sat_axis_aligned_cases():
    """tests/CollisionDetectionTests.cpp:11-70"""
    from oracle.step_oracle import obb_collided
    assert obb_collided([0, 0, 0, 1, 1], [1, 1, 0, 1, 1])                    # Colliding
    assert not obb_collided([0, 0, 0, .5, .5], [2, 2, 0, .5, .5])           # NotColliding
    assert obb_collided([0, 0, 0, .5, .5], [1, 1, 0, .5, .5])               # PointIntersection: touching counts
    assert obb_collided([0, 0, 0, 1, 1], [0, 0, 0, .5, .5])                 # OneInsideOther


def test_sat_exhaustive_rotations():
    """tests/CollisionDetectionTests.cpp:72-85"""
    from oracle.step_oracle import obb_collided
    for deg in range(0, 360, 15):
        assert obb_collided([0, 0, 0, 1, 1], [0.5, 0.5, to_radians(deg), 1, 1]), deg


def test_egocentric_observation():
    """tests/EgocentricRoadObservationTests.cpp:9-21: heading is exactly toRadians(180)."""
    from oracle.step_oracle import observation_of
    out = observation_of([3, 0, to_radians(90)], [3, 3, to_radians(270)])
    assert out[0] - 3 < 1e-6 and out[1] - 0 < 1e-6
    assert out[2] == to_radians(180)


def _make(scenes, pd):
    from oracle.step_oracle import OracleManager
    return OracleManager(scenes, engine_params(pd))


def test_classic_bicycle_closed_form(tmp_path):
    kat.run_classic_bicycle_closed_form(_make, tmp_path)


@pytest.mark.parametrize("model", [1, 2])
def test_forward_inverse_dynamics(model, tmp_path):
    kat.run_forward_inverse_dynamics(_make, model, tmp_path)


def test_expert_replay_reaches_goals_without_collisions(tmp_path):
    kat.run_expert_replay(_make, tmp_path)


@pytest.mark.parametrize("key", ROLLOUT_KEYS)
def test_oracle_matches_golden_rollout(key, tmp_path):
    kat.run_golden_rollout(_make, key, tmp_path, require_exact=True)


def test_oracle_matches_reference_live(tmp_path):
    from oracle import refmgr
    if not refmgr.available():
        pytest.skip("oracle/_ref not built")
    from oracle.step_oracle import OracleManager
    scenes = unpack(["synth_b", "synth_a"], tmp_path)
    pd = param_dict(dynamicsModel=2, enableLidar=True, collisionBehaviour=0, maxNumControlledAgents=6)
    ref = refmgr.RefManager(scenes, ref_params(pd))
    mgr = OracleManager(scenes, engine_params(pd))
    stats = {}
    for t in range(25):
        a = random_actions(t, 2, 128, 2, seed=11)
        ref.tensor("action")[...] = a
        mgr.set_actions(a)
        ref.step(); mgr.step()
        if t == 9:
            ref.reset([1]); mgr.reset([1])
        compare_step(snapshot_ref(ref), kat._snapshot(mgr, TENSORS_CMP), 50.0, True, tag=f"step {t + 1}",
                     stats=stats)
    ref.close()


def test_bev_matches_reference_live(tmp_path):
    """The restated bird's-eye-view rasteriser (src/sim.cpp:462-555, src/rasterizer.hpp:12-78) is
    bit-identical to the reference's on a synthetic and on the reference's own test scene."""
    from oracle import refmgr
    if not refmgr.available():
        pytest.skip("oracle/_ref not built")
    from oracle.step_oracle import OracleManager
    scenes = unpack(["synth_b", "womd_test_scene"], tmp_path)
    pd = param_dict(dynamicsModel=1, maxNumControlledAgents=6, observationRadius=40.0)
    ref = refmgr.RefManager(scenes, ref_params(pd))
    mgr = OracleManager(scenes, engine_params(pd))
    mgr.enable_bev()
    A = ref.tensor("shape")[:, 0]
    for t in range(4):
        for w in range(2):
            r, o = ref.tensor("bev_observation")[w, :A[w]], mgr.tensor("bev_observation")[w, :A[w]]
            assert (r != 0).sum() > 1000 and np.array_equal(r, o), (t, w)
        a = random_actions(t, 2, 128, 1, seed=2)
        ref.tensor("action")[...] = a
        mgr.set_actions(a)
        ref.step(); mgr.step()
    ref.close()


def test_ragged_and_empty_scenes_match_reference(tmp_path):
    """Edge cases: a single agent, a map of two entities, a map with no road at all - lidar, k-nearest padding and
    partner rows of the restatement against the unmodified reference."""
    from helpers import ragged_scenes
    from oracle import refmgr
    if not refmgr.available():
        pytest.skip("oracle/_ref not built")
    scenes = ragged_scenes(tmp_path)
    for model in (0, 2):
        pd = param_dict(dynamicsModel=model, enableLidar=True)
        ref = refmgr.RefManager(scenes, ref_params(pd))
        orc = _make(scenes, pd)
        assert ref.tensor("shape")[:, 1].tolist()[2] == 0
        stats = {}
        for t in range(6):
            a = random_actions(t, len(scenes), 128, model)
            ref.tensor("action")[...] = a
            orc.set_actions(a)
            ref.step(); orc.step()
            compare_step(snapshot_ref(ref), kat._snapshot(orc, TENSORS_CMP), 50.0, True, tag=f"model {model} step {t + 1}",
                         stats=stats)
