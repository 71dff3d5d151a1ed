"""-m gpu: engine behaviour added in round 2 - device-side episode statistics (gd_stats), the stream contract
(no explicit synchronisation between a torch write of the actions and step()), one CUDA graph per step, and
the device-side auto-reset from the scenario pool."""
import numpy as np
import pytest

from golden_util import unpack
from helpers import engine_params, param_dict, random_actions, ref_params

pytestmark = pytest.mark.gpu


def _sim(scenes, pd, S=128):
    import torch
    assert torch.cuda.is_available()
    from gpudrive_lab_b200 import madrona_gpudrive as mg
    return mg.SimManager(mg.madrona.ExecMode.CUDA, 0, scenes, engine_params(pd), max_agents=S)


@pytest.mark.parametrize("behaviour", [0, 2])      # AgentStop, Ignore
def test_episode_statistics_match_reference_info_and_done(synth_dir, behaviour):
    """Every gd_stats counter equals the sum the reference's own info / done tensors give over a collision
    stress rollout (src/sim.cpp:597-626,713-724): flags raised (0 -> 1) per (agent, step), goals reached,
    agents done, live and controlled agent-steps. These are what the job's one NCCL reduction carries."""
    import torch
    from gpudrive_lab_b200 import synth
    from oracle.refmgr import RefManager
    scenes = synth.write_scenarios(synth_dir, 2, stress=True)
    pd = param_dict(dynamicsModel=1, collisionBehaviour=behaviour, distanceToGoalThreshold=3.0)
    ref = RefManager(scenes, ref_params(pd))
    sim = _sim(scenes, pd)
    act = sim.action_tensor().to_torch()
    steps = 91
    exp = dict(collided_road=0, collided_vehicle=0, collided_nonvehicle=0, goals_reached=0, dones=0)
    info0, done0 = ref.tensor("info").copy(), ref.tensor("done").copy()
    live = int(ref.tensor("shape")[:, 0].sum())
    ctrl = int(ref.tensor("controlled_state").sum())
    assert sim.stats()["steps"] == 0 and sim.stats()["dones"] == 0 and sim.stats()["episodes_reset"] == 2
    for t in range(steps):
        a = random_actions(t, 2, 128, 1, seed=31)
        ref.tensor("action")[...] = a
        act.copy_(torch.from_numpy(a))
        ref.step(); sim.step()
        info1, done1 = ref.tensor("info").copy(), ref.tensor("done").copy()
        for k, name in enumerate(("collided_road", "collided_vehicle", "collided_nonvehicle")):
            if behaviour == 2:      # Ignore: movementSystem clears the flags of a collided agent, the pass raises them again
                exp[name] += int((info1[..., k] == 1).sum())
            else:
                exp[name] += int(((info1[..., k] == 1) & (info0[..., k] == 0)).sum())
        exp["goals_reached"] += int(((info1[..., 3] == 1) & (info0[..., 3] == 0)).sum())
        exp["dones"] += int(((done1 == 1) & (done0 == 0)).sum())
        info0, done0 = info1, done1
    st = sim.stats()
    assert np.array_equal(sim.info_tensor().to_torch().cpu().numpy(), info1)
    assert st["steps"] == steps and st["agent_steps"] == live * steps and st["controlled_steps"] == ctrl * steps
    for k, v in exp.items():
        assert st[k] == v, (k, st[k], v)
    assert exp["collided_road"] + exp["collided_vehicle"] > 0 and exp["dones"] > 0
    assert st["sat_tests"] > 0
    # counters are per engine and keep counting across a reset; resets are counted as episodes
    sim.reset([0, 1])
    assert sim.stats()["episodes_reset"] == 4 and sim.stats()["dones"] == exp["dones"]


def test_step_is_ordered_after_default_stream_action_writes(synth_dir):
    """The drop-in call pattern of gpudrive/env/env_torch.py:606-660: `action_tensor().to_torch()[:, :, :3].copy_(a)`
    on torch's default stream, then sim.step(), with NO torch.cuda.synchronize() in between. The engine stream is
    a blocking one (as Madrona's), so the step must see the new actions: results equal those of a second engine
    that synchronises explicitly. Large enough (512 worlds) for the copy to be in flight at the launch."""
    import torch
    from gpudrive_lab_b200 import synth
    scenes_u = synth.write_scenarios(synth_dir, 2)
    scenes = [scenes_u[i % 2] for i in range(512)]
    pd = param_dict(dynamicsModel=2)
    a, b = _sim(scenes, pd), _sim(scenes, pd)
    act_a, act_b = a.action_tensor().to_torch(), b.action_tensor().to_torch()
    for t in range(25):
        h = torch.from_numpy(random_actions(t, 512, 128, 2, seed=3)[..., :3].copy()).pin_memory()
        act_a[:, :, :3].copy_(h, non_blocking=True)          # in flight when step() launches
        a.step()
        act_b[:, :, :3].copy_(h)
        torch.cuda.synchronize()
        b.step()
        torch.cuda.synchronize()
    for name in ("absolute_self_observation_tensor", "self_observation_tensor", "reward_tensor", "done_tensor"):
        assert torch.equal(getattr(a, name)().to_torch(), getattr(b, name)().to_torch()), name


def test_graph_step_equals_plain_launches(tmp_path):
    """GD_OPT_GRAPH: a step launched as one CUDA graph leaves the same tensors as the two plain launches, with
    and without per-kernel profiling events, and across a change of the launch parameters (knn index option)."""
    import torch
    scenes = unpack(["synth_a", "synth_b", "synth_a"], tmp_path)
    pd = param_dict(dynamicsModel=2, enableLidar=True, collisionBehaviour=0, maxNumControlledAgents=8)
    g, p = _sim(scenes, pd), _sim(scenes, pd)
    p.set_option(2, 0)
    for t in range(20):
        act = torch.from_numpy(random_actions(t, 3, 128, 2, seed=5))
        g.action_tensor().to_torch().copy_(act); p.action_tensor().to_torch().copy_(act)
        if t == 5:
            g.set_option(1, 1); p.set_option(1, 1)
        if t == 10:
            g.set_option(0, 0); p.set_option(0, 0)
        if t == 14:
            g.reset([1]); p.reset([1])
        g.step(); p.step()
        for name in ("absolute_self_observation_tensor", "partner_observations_tensor", "agent_roadmap_tensor",
                     "lidar_tensor", "info_tensor", "done_tensor", "reward_tensor"):
            assert torch.equal(getattr(g, name)().to_torch(), getattr(p, name)().to_torch()), (t, name)
    kg, kp = g.kernel_times(), p.kernel_times()
    assert kg["steps"] == kp["steps"] == 15 and kg["k_obs_ms"] > 0 and kg["k_agents_ms"] > 0
    assert g.stats() == p.stats()


def test_device_auto_reset_equals_host_driven_rebind(synth_dir):
    """gd_auto_reset (config 4): finished worlds move on to the next pooled scenario on the device. Equal, tensor
    for tensor and step for step, to the host-driven loop `read done -> bind_worlds(finished, next scene)`."""
    import os
    import torch
    from gpudrive_lab_b200 import synth
    pool = synth.write_scenarios(os.path.join(synth_dir, "pool_ar"), 5, num_agents=12)
    pd = param_dict(dynamicsModel=2, maxNumControlledAgents=12, distanceToGoalThreshold=4.0)
    worlds = [pool[i % 5] for i in range(7)]
    a, b = _sim(worlds, pd), _sim(worlds, pd)
    ids_a, ids_b = [a.pool_add(p) for p in pool], [b.pool_add(p) for p in pool]
    a.auto_reset_pool(ids_a)
    cursor = [w % 5 for w in range(7)]
    names = ("shape_tensor", "absolute_self_observation_tensor", "self_observation_tensor", "partner_observations_tensor",
             "agent_roadmap_tensor", "done_tensor", "info_tensor", "reward_tensor", "steps_remaining_tensor",
             "expert_trajectory_tensor", "map_observation_tensor", "controlled_state_tensor", "metadata_tensor")
    resets = 0
    for t in range(200):
        act = torch.from_numpy(random_actions(t, 7, 128, 2, seed=17))
        a.action_tensor().to_torch().copy_(act); b.action_tensor().to_torch().copy_(act)
        a.step(); a.auto_reset()
        b.step()
        done = b.done_tensor().to_torch()[..., 0].cpu().numpy()
        ctrl = b.controlled_state_tensor().to_torch()[..., 0].cpu().numpy()
        fin = [w for w in range(7) if ctrl[w].sum() > 0 and (done[w][ctrl[w] == 1] == 1).all()]
        if fin:
            for w in fin:
                cursor[w] = (cursor[w] + 1 + w % 4) % 5
            b.bind_worlds(fin, [ids_b[cursor[w]] for w in fin])
            resets += len(fin)
        a.sync()
        for n in names:
            assert torch.equal(getattr(a, n)().to_torch(), getattr(b, n)().to_torch()), (t, n)
    assert resets >= 7, "every world is expected to finish at least one episode in 200 steps"
    sa, sb = a.stats(), b.stats()
    assert sa["episodes_reset"] == sb["episodes_reset"] == 7 + resets
    for k in ("agent_steps", "controlled_steps", "dones", "goals_reached", "collided_vehicle"):
        assert sa[k] == sb[k], k
    # the host mirror follows the device binding: deleteAgents after an auto-reset acts on the current scene
    aid = int(a.absolute_self_observation_tensor().to_torch()[3, 1, 13].item())
    n0 = int(a.shape_tensor().to_torch()[3, 0])
    a.deleteAgents({3: [aid]}); b.deleteAgents({3: [aid]})
    assert int(a.shape_tensor().to_torch()[3, 0]) == n0 - 1
    for n in names:
        assert torch.equal(getattr(a, n)().to_torch(), getattr(b, n)().to_torch()), n


@pytest.mark.parametrize("algorithm", [0, 1])      # k-nearest, linear (the env default)
def test_fused_policy_observation_equals_separate_kernel(synth_dir, algorithm):
    """GD_OPT_FUSED_POLICY_OBS: the policy observation k_obs writes as a second output, from the
    rows it holds on chip, equals bit for bit what the separate gd_policy_obs pass computes from the exported
    tensors (itself pinned against the reference's Python classes) - every slot, padding slots included, across
    steps, a partial reset, set_maps, deleteAgents and a device-side re-bind."""
    import os
    import torch
    from gpudrive_lab_b200 import synth
    pool = synth.write_scenarios(os.path.join(synth_dir, "pool_fp"), 4, num_agents=12)
    big = synth.write_scenarios(synth_dir, 1)          # 64 agents, ~2 000 road entities: rows fill all K = 200
    scenes = [pool[0], big[0], pool[1]]
    pd = param_dict(dynamicsModel=2, roadObservationAlgorithm=algorithm, enableLidar=bool(algorithm == 0),
                    maxNumControlledAgents=64, distanceToGoalThreshold=3.0)
    f, s = _sim(scenes, pd), _sim(scenes, pd)
    f.set_option(4, 1)                                 # fused: k_obs<true> + k_bind keep the buffer current

    def check(tag):
        a = f.policy_observation_tensor(normalize=True, torch_cuda_arithmetic=True).to_torch()
        b = s.policy_observation_tensor(normalize=True, torch_cuda_arithmetic=True).to_torch()
        assert a.shape == b.shape == (3, 128, 6 + 127 * 6 + 200 * 13)
        if not torch.equal(a.view(torch.int32), b.view(torch.int32)):
            bad = (a.view(torch.int32) != b.view(torch.int32)).nonzero()[:6].tolist()
            raise AssertionError(f"{tag}: fused policy observation differs at {bad}: "
                                 f"{[float(a[tuple(i)]) for i in bad]} vs {[float(b[tuple(i)]) for i in bad]}")
        for n in ("agent_roadmap_tensor", "partner_observations_tensor", "self_observation_tensor"):
            assert torch.equal(getattr(f, n)().to_torch(), getattr(s, n)().to_torch()), (tag, n)

    check("after enabling")
    with pytest.raises(RuntimeError):
        f.policy_observation_tensor(normalize=True, torch_cuda_arithmetic=False)      # the buffer holds the CUDA form
    launches = f.launch_count()
    f.policy_observation_tensor(normalize=True, torch_cuda_arithmetic=True)
    assert f.launch_count() == launches, "the fused read-out must not launch anything"
    for t in range(14):
        act = torch.from_numpy(random_actions(t, 3, 128, 2, seed=29))
        f.action_tensor().to_torch().copy_(act); s.action_tensor().to_torch().copy_(act)
        f.step(); s.step()
        check(f"step {t + 1}")
        if t == 4:
            f.reset([1]); s.reset([1]); check("partial reset")
        if t == 7:
            f.set_maps([pool[2], big[0], pool[3]]); s.set_maps([pool[2], big[0], pool[3]]); check("set_maps")
        if t == 10:
            aid = int(f.absolute_self_observation_tensor().to_torch()[0, 2, 13].item())
            f.deleteAgents({0: [aid]}); s.deleteAgents({0: [aid]}); check("deleteAgents")
    ida, idb = f.pool_add(pool[0]), s.pool_add(pool[0])
    f.bind_worlds([1], [ida]); s.bind_worlds([1], [idb])      # a 64-agent world becomes a 12-agent one: 52 new padding slots
    check("bind_worlds")
    f.step(); s.step(); check("step after bind_worlds")
    f.set_option(4, 0)
    f.step(); s.step()
    a = f.policy_observation_tensor(normalize=True).to_torch(); b = s.policy_observation_tensor(normalize=True).to_torch()
    assert torch.equal(a, b)
