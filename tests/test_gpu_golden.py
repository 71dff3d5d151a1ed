"""The CUDA engine, through the C ABI, against (1) the reference's own known-answer tests, (2) the golden
rollouts recorded from the unmodified reference (tests/golden/), (3) the CPU restatement oracle.
Nothing here reads /root/reference."""
import numpy as np
import pytest

import kat
from golden_util import ROLLOUT_KEYS, unpack
from helpers import TENSORS_CMP, EngineAdapter, compare_step, engine_params, param_dict, random_actions, report

pytestmark = pytest.mark.gpu


def _make(scenes, pd):
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device; the engine has no fallback"
    return EngineAdapter(scenes, pd)


def test_classic_bicycle_closed_form(tmp_path):
    kat.run_classic_bicycle_closed_form(_make, tmp_path)


@pytest.mark.parametrize("model", [1, 2])
def test_forward_inverse_dynamics(model, tmp_path):
    kat.run_forward_inverse_dynamics(_make, model, tmp_path)


def test_expert_replay_reaches_goals_without_collisions(tmp_path):
    kat.run_expert_replay(_make, tmp_path)


@pytest.mark.parametrize("key", ROLLOUT_KEYS)
def test_engine_matches_golden_rollout(key, tmp_path):
    stats = kat.run_golden_rollout(_make, key, tmp_path, require_exact=False)
    report("golden_" + key, stats)


def test_engine_matches_restatement_oracle(tmp_path):
    """Engine vs oracle/step_oracle.cpp (the "port" oracle), with a mid-episode partial reset."""
    from oracle.step_oracle import OracleManager
    scenes = unpack(["synth_b", "synth_a", "synth_b"], tmp_path)
    pd = param_dict(dynamicsModel=2, enableLidar=True, collisionBehaviour=0, maxNumControlledAgents=6)
    orc = OracleManager(scenes, engine_params(pd))
    eng = _make(scenes, pd)
    stats = {}
    for t in range(30):
        a = random_actions(t, 3, 128, 2, seed=11)
        orc.set_actions(a); eng.set_actions(a)
        orc.step(); eng.step()
        if t == 9:
            orc.reset([1]); eng.reset([1])
        compare_step(kat._snapshot(orc, TENSORS_CMP), kat._snapshot(eng, TENSORS_CMP), 50.0, True,
                     tag=f"step {t + 1}", stats=stats)
    report("restatement_oracle_30_steps", stats)


def test_knn_index_sets_bit_exact(tmp_path):
    """kNN road-entity index SETS per agent are identical to the oracle's heap selection."""
    from oracle.step_oracle import OracleManager
    scenes = unpack(["synth_b", "womd_test_scene"], tmp_path)
    pd = param_dict(dynamicsModel=1, observationRadius=40.0, maxNumControlledAgents=32)
    orc = OracleManager(scenes, engine_params(pd))
    eng = _make(scenes, pd)
    shape = orc.tensor("shape")
    ties = 0
    for t in range(12):
        a = random_actions(t, 2, 128, 1, seed=5)
        orc.set_actions(a); eng.set_actions(a)
        orc.step(); eng.step()
        ko, ke = orc.tensor("knn_index"), eng.tensor("knn_index")
        for w in range(2):
            for i in range(int(shape[w, 0])):
                so, se = set(ko[w, i][ko[w, i] >= 0].tolist()), set(ke[w, i][ke[w, i] >= 0].tolist())
                # bit-exact: same compiler-independent key arithmetic on both sides, ties broken by road index
                assert so == se, (t, w, i, sorted(so ^ se))
                # every selected road entity appears once
                kk = ke[w, i][ke[w, i] >= 0]
                assert len(set(kk.tolist())) == len(kk)
    assert ties == 0


def test_bev_grid_bit_exact(tmp_path):
    """collectBevObservationsSystem (a14): the 200 x 200 grid of every live agent equals the oracle's,
    which is bit-identical to the reference's (tests/test_cpu_oracle.py::test_bev_matches_reference_live)."""
    from oracle.step_oracle import OracleManager
    scenes = unpack(["synth_b", "womd_test_scene", "synth_a"], tmp_path)
    pd = param_dict(dynamicsModel=1, maxNumControlledAgents=6, observationRadius=40.0)
    orc = OracleManager(scenes, engine_params(pd))
    orc.enable_bev()
    eng = EngineAdapter(scenes, pd, enable_bev=True)
    shape = orc.tensor("shape")
    for t in range(8):
        go, ge = orc.tensor("bev_observation"), eng.tensor("bev_observation")
        assert ge.shape == (3, 128, 200, 200, 1)
        for w in range(3):
            A = int(shape[w, 0])
            assert (go[w, :A] != 0).sum() > 1000
            assert np.array_equal(go[w, :A], ge[w, :A]), (t, w, int((go[w, :A] != ge[w, :A]).sum()))
            assert not ge[w, A:].any()
        a = random_actions(t, 3, 128, 1, seed=9)
        orc.set_actions(a); eng.set_actions(a)
        orc.step(); eng.step()
        if t == 4:
            orc.reset([0, 2]); eng.reset([0, 2])


def test_dense_map_overflow_paths(tmp_path):
    """The reference's own scene at polylineReductionThreshold = 0 (9 899 road entities, 310 groups) with a
    100 m radius: more in-radius entities than the candidate buffer holds and more groups than one list
    pass, so the slow selection path and the multi-pass group loop are what produce the rows."""
    from oracle.step_oracle import OracleManager
    scenes = unpack(["womd_test_scene"], tmp_path)
    pd = param_dict(polylineReductionThreshold=0.0, observationRadius=100.0, dynamicsModel=1, enableLidar=True,
                    initOnlyValidAgentsAtFirstStep=False, maxNumControlledAgents=8)
    orc = OracleManager(scenes, engine_params(pd))
    eng = _make(scenes, pd)
    assert int(orc.tensor("shape")[0, 1]) == 9899
    stats = {}
    for t in range(3):
        a = random_actions(t, 1, 128, 1, seed=21)
        orc.set_actions(a); eng.set_actions(a)
        orc.step(); eng.step()
        compare_step(kat._snapshot(orc, TENSORS_CMP), kat._snapshot(eng, TENSORS_CMP), 100.0, True,
                     tag=f"dense step {t + 1}", stats=stats)
    report("dense_map_9899_entities", stats)
    # linear algorithm on the same map: first K in road-index order
    pd2 = dict(pd, roadObservationAlgorithm=1, enableLidar=False)
    orc2 = OracleManager(scenes, engine_params(pd2)); eng2 = _make(scenes, pd2)
    orc2.step(); eng2.step()
    ko, ke = orc2.tensor("knn_index")[0], eng2.tensor("knn_index")[0]
    for i in range(int(orc2.tensor("shape")[0, 0])):
        assert ko[i].tolist() == ke[i].tolist(), i          # same indices in the same (road-index) order
