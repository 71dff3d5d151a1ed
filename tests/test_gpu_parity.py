"""-m gpu: parity of the CUDA step engine against the reference's own CPU Manager (oracle/_ref)
on identical synthetic scenes and actions, through the C ABI. Bit-exact: done / info / collision
flags / ids / partner ordering / roadmap index sets. 1e-5 relative: positions, headings, speeds,
observations, lidar depths, rewards (rules in helpers.compare_step)."""
import numpy as np
import pytest

from helpers import param_dict
from rollout import run_parity

pytestmark = pytest.mark.gpu


def _need():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("CUDA device required for -m gpu tests (no CPU fallback exists)")
    from oracle import refmgr
    if not refmgr.available():
        pytest.fail("oracle/_ref/libgd_ref.so missing: run __graft_entry__.build() where /root/reference exists")


def _scenes(synth_dir, n, **kw):
    from gpudrive_lab_b200 import synth
    return synth.write_scenarios(synth_dir, n, **kw)


def test_bicycle_ignore_full_episode(synth_dir):
    """config 2 shape: InvertibleBicycle, self + partner + roadmap, 91 steps."""
    _need()
    scenes = _scenes(synth_dir, 2)
    pd = param_dict(dynamicsModel=1, collisionBehaviour=2)
    run_parity(scenes, pd, 91)          # (helpers.report: every exemption counter must be 0)


def test_delta_lidar_full_episode(synth_dir):
    """config 3 shape: DeltaLocal, full observations + lidar, 91 steps."""
    _need()
    scenes = _scenes(synth_dir, 2)
    pd = param_dict(dynamicsModel=2, enableLidar=True)
    run_parity(scenes, pd, 91)


def test_agent_stop_collisions(synth_dir):
    """AgentStop: collided agents freeze, then get removed; flags must stay bit-exact."""
    _need()
    scenes = _scenes(synth_dir, 2, stress=True)
    pd = param_dict(dynamicsModel=1, collisionBehaviour=0, enableLidar=True)
    ref, sim, _ = run_parity(scenes, pd, 91)
    info = ref.tensor("info")
    assert info[..., :3].sum() > 0, "the stress scene is expected to produce collisions"


def test_agent_removed_classic(synth_dir):
    _need()
    scenes = _scenes(synth_dir, 1, stress=True)
    pd = param_dict(dynamicsModel=0, collisionBehaviour=1)
    run_parity(scenes, pd, 60)


def test_expert_replay_linear_roadmap(synth_dir):
    """No controlled agents (pure expert replay, tests/test_expert.py) + the env-default 'linear'
    road algorithm (AllEntitiesWithRadiusFiltering), whose row ORDER is part of the contract: the first K
    in-radius entities in road-index order (src/sim.cpp:261-279)."""
    _need()
    from helpers import close, POS_SCALE
    scenes = _scenes(synth_dir, 2)
    pd = param_dict(dynamicsModel=1, maxNumControlledAgents=0, roadObservationAlgorithm=1,
                    collisionBehaviour=0, isStaticAgentControlled=False)
    ref, sim, _ = run_parity(scenes, pd, 91)
    rr = ref.tensor("agent_roadmap")
    re_ = sim.agent_roadmap_tensor().to_torch().cpu().numpy()
    shape = ref.tensor("shape")
    rows = 0
    for w in range(2):
        A = int(shape[w, 0])
        r, e = rr[w, :A], re_[w, :A]
        # row for row: type / id / mapType and the scale exact, ego-frame position and heading 1e-5
        assert np.array_equal(r[..., 6:9], e[..., 6:9]), "linear road-map rows are not in the reference's order"
        assert np.array_equal(r[..., 2:5], e[..., 2:5])
        assert close(e[..., 0:2], r[..., 0:2], POS_SCALE).all()
        rows += int((r[..., 6] != 0).sum())
    assert rows > 400


def test_state_model_and_distance_reward(synth_dir):
    _need()
    scenes = _scenes(synth_dir, 1)
    pd = param_dict(dynamicsModel=3, rewardType=0, observationRadius=30.0)
    run_parity(scenes, pd, 10)


def test_reset_and_partial_reset(synth_dir):
    _need()
    import torch
    scenes = _scenes(synth_dir, 3)
    pd = param_dict(dynamicsModel=1, enableLidar=True)
    from helpers import compare_step, snapshot_engine, snapshot_ref
    ref, sim, _ = run_parity(scenes, pd, 12)
    ref.reset([1]); sim.reset([1])
    compare_step(snapshot_ref(ref), snapshot_engine(sim), 50.0, True, tag="after reset([1])")
    assert int(sim.steps_remaining_tensor().to_torch()[1, 0, 0]) == 91
    assert int(sim.steps_remaining_tensor().to_torch()[0, 0, 0]) == 91 - 12
    ref.step(); sim.step()
    compare_step(snapshot_ref(ref), snapshot_engine(sim), 50.0, True, tag="step after reset")


def test_set_maps_and_delete_agents(synth_dir):
    _need()
    from helpers import compare_step, snapshot_engine, snapshot_ref
    scenes = _scenes(synth_dir, 4)
    pd = param_dict(dynamicsModel=2)
    ref, sim, _ = run_parity(scenes[:2], pd, 5)
    ref.set_maps(scenes[2:4]); sim.set_maps(scenes[2:4])
    compare_step(snapshot_ref(ref), snapshot_engine(sim), 50.0, False, tag="after set_maps")
    ids = [100, 103, 117]
    ref.delete_agents(1, ids); sim.deleteAgents({1: ids})
    compare_step(snapshot_ref(ref), snapshot_engine(sim), 50.0, False, tag="after deleteAgents")
    assert int(sim.shape_tensor().to_torch()[1, 0]) == 61
    ref.step(); sim.step()
    compare_step(snapshot_ref(ref), snapshot_engine(sim), 50.0, False, tag="step after deleteAgents")


def test_collision_stress_128_agents(synth_dir):
    """config 5 shape: every one of the 128 slots live, agents packed into a double intersection, road edges
    sampled at 0.25 m (about 4 000 edge segments), AgentStop. Collision / done / info flags bit-exact."""
    _need()
    import os
    d = os.path.join(synth_dir, "stress128")
    scenes = _scenes(d, 1, num_agents=128, stress=True, edge_spacing=0.25)
    pd = param_dict(dynamicsModel=1, collisionBehaviour=0, maxNumControlledAgents=128,
                    polylineReductionThreshold=0.0)
    ref, sim, _ = run_parity(scenes, pd, 40, names=["shape", "done", "info", "steps_remaining", "controlled_state",
                                                    "response_type", "self_observation",
                                                    "absolute_self_observation", "reward", "partner_observations"])
    shape = ref.tensor("shape")[0]
    assert shape[0] == 128 and shape[1] >= 4000
    info = ref.tensor("info")[0]
    assert (info[:, :3].sum(1) > 0).mean() >= 0.2, "expected >= 20 % of the agents to have collided"


def test_auto_reset_rollout_from_device_pool(synth_dir):
    """config 4 shape: episodes end, finished worlds are re-bound to other pooled scenarios on the device
    (gd_bind_worlds) and keep stepping; the reference does the same through set_maps. Tensors equal after
    every re-bind and at the end."""
    _need()
    import os
    import torch
    from gpudrive_lab_b200 import madrona_gpudrive as mg
    from helpers import (TENSORS_CMP, compare_step, engine_params, random_actions, ref_params, snapshot_engine,
                         snapshot_ref)
    from oracle.refmgr import RefManager
    d = os.path.join(synth_dir, "pool")
    pool = _scenes(d, 4, num_agents=12)
    pd = param_dict(dynamicsModel=2, maxNumControlledAgents=12, distanceToGoalThreshold=3.0)
    cur = [0, 1]
    ref = RefManager([pool[i] for i in cur], ref_params(pd))
    sim = mg.SimManager(mg.madrona.ExecMode.CUDA, 0, [pool[i] for i in cur], engine_params(pd))
    ids = [sim.pool_add(p) for p in pool]
    act = sim.action_tensor().to_torch()
    stats = {}
    for t in range(120):
        a = random_actions(t, 2, 128, 2, seed=13)
        ref.tensor("action")[...] = a
        act.copy_(torch.from_numpy(a)); torch.cuda.synchronize()
        ref.step(); sim.step()
        if t % 10 == 9:
            compare_step(snapshot_ref(ref), snapshot_engine(sim), 50.0, False, tag=f"step {t + 1}", stats=stats)
        if t in (45, 91):       # world (t == 45 ? 1 : 0) switches to the next pooled scenario
            w = 1 if t == 45 else 0
            cur[w] = (cur[w] + 2) % 4
            ref.set_maps([pool[i] for i in cur])              # reference: re-parses and rebuilds every world
            ref.reset([w])
            sim.bind_worlds([w], [ids[cur[w]]])               # engine: re-bind + reset on the device
            # the reference's set_maps re-creates every world (fresh episodes everywhere, src/mgr.cpp:590-654);
            # mirror that on the other world for a like-for-like state
            sim.reset([1 - w]); ref.reset([1 - w])
            compare_step(snapshot_ref(ref, TENSORS_CMP + ["map_observation", "expert_trajectory"]),
                         snapshot_engine(sim, TENSORS_CMP + ["map_observation", "expert_trajectory"]),
                         50.0, False, tag=f"after re-bind at {t + 1}", stats=stats)


def test_ragged_and_empty_scenes(tmp_path):
    """One agent; a map of two entities; a map with no road at all (empty road table, empty collidable table,
    no lidar hit on the road planes, all-padding road-map rows)."""
    from helpers import ragged_scenes
    scenes = ragged_scenes(tmp_path)
    for model in (0, 2):
        run_parity(scenes, param_dict(dynamicsModel=model, enableLidar=True), 12)
