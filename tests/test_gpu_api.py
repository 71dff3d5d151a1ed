"""C-ABI entry points that have no reference equivalent (host-buffer step, async step, device scenario
pool, statistics, kernel timing) behave like the reference-shaped calls they wrap."""
import numpy as np
import pytest

from golden_util import unpack
from helpers import engine_params, param_dict, random_actions

pytestmark = pytest.mark.gpu


def _sim(scenes, pd, S=128):
    import torch
    assert torch.cuda.is_available()
    from gpudrive_lab_b200 import madrona_gpudrive as mg
    return mg.SimManager(mg.madrona.ExecMode.CUDA, 0, scenes, engine_params(pd), max_agents=S)


def test_step_host_equals_step(tmp_path):
    import torch
    scenes = unpack(["synth_a", "synth_b", "synth_a"], tmp_path)
    pd = param_dict(dynamicsModel=2, enableLidar=True, collisionBehaviour=0, maxNumControlledAgents=8)
    a, b = _sim(scenes, pd), _sim(scenes, pd)
    W, S = 3, 128
    h_rew = torch.empty(W, S, 1).pin_memory()
    h_done = torch.empty(W, S, 1, dtype=torch.int32).pin_memory()
    h_info = torch.empty(W, S, 5, dtype=torch.int32).pin_memory()
    act_b = b.action_tensor().to_torch()
    for t in range(12):
        act = random_actions(t, W, S, 2, seed=3)
        h_act = torch.from_numpy(np.ascontiguousarray(act[..., :3])).pin_memory()
        a.step_host(h_act, h_rew, h_done, h_info)
        act_b[:, :, :3].copy_(torch.from_numpy(act[..., :3]))
        torch.cuda.synchronize()
        b.step()
        assert torch.equal(h_rew, b.reward_tensor().to_torch().cpu())
        assert torch.equal(h_done, b.done_tensor().to_torch().cpu())
        assert torch.equal(h_info, b.info_tensor().to_torch().cpu())
        for name in ("self_observation_tensor", "partner_observations_tensor", "agent_roadmap_tensor", "lidar_tensor"):
            assert torch.equal(getattr(a, name)().to_torch(), getattr(b, name)().to_torch()), name
    # action columns beyond those passed stay untouched (zero after reset)
    assert float(a.action_tensor().to_torch()[:, :, 3:].abs().sum()) == 0.0


def test_async_step_and_kernel_times(tmp_path):
    import torch
    scenes = unpack(["synth_b"], tmp_path)
    pd = param_dict(dynamicsModel=1)
    a, b = _sim(scenes, pd), _sim(scenes, pd)
    a.set_option(1, 1)
    n0 = a.launch_count()
    for _ in range(5):
        a.step_async()
    a.sync()
    for _ in range(5):
        b.step()
    assert a.launch_count() - n0 == 10            # two kernels per step
    kt = a.kernel_times()
    assert kt["steps"] == 5 and kt["k_obs_ms"] > 0 and kt["k_agents_ms"] > 0
    assert torch.equal(a.absolute_self_observation_tensor().to_torch(), b.absolute_self_observation_tensor().to_torch())
    st = a.stats()
    assert st["steps"] == 5 and st["agent_steps"] == 5 * int(a.shape_tensor().to_torch()[0, 0])


def test_device_scenario_pool_rebind(tmp_path):
    """Re-binding a world to a pooled scenario gives the same tensors as constructing it there."""
    import torch
    sa, sb = unpack(["synth_a", "synth_b"], tmp_path)
    pd = param_dict(dynamicsModel=2, maxNumControlledAgents=8)
    sim = _sim([sa, sa], pd)
    ida, idb = sim.pool_add(sa), sim.pool_add(sb)
    assert ida != idb
    for _ in range(3):
        sim.step()
    sim.bind_worlds([1], [idb])          # world 1 now runs scenario b, freshly reset
    ref = _sim([sa, sb], pd)
    for name in ("shape_tensor", "map_observation_tensor", "expert_trajectory_tensor", "controlled_state_tensor",
                 "self_observation_tensor", "agent_roadmap_tensor", "partner_observations_tensor", "done_tensor",
                 "steps_remaining_tensor", "info_tensor", "metadata_tensor", "world_means_tensor"):
        x, y = getattr(sim, name)().to_torch()[1], getattr(ref, name)().to_torch()[1]
        assert torch.equal(x, y), name
    assert int(sim.steps_remaining_tensor().to_torch()[0, 0, 0]) == 91 - 3      # world 0 untouched


def test_scene_cache_is_bounded_under_set_maps_churn(tmp_path):
    """An env that resamples with set_maps over a large dataset (swap_data_batch) must not grow the device scenario
    cache for ever: scenes no world is bound to are evicted beyond max(W, 64) and their ids / arena blocks reused,
    while ids the caller holds (pool_add) stay valid, and a scene that comes back after eviction is rebuilt
    identically."""
    import torch
    from gpudrive_lab_b200 import synth
    paths = []
    for i in range(0, 150, 15):       # (batches below write_scenarios' fork-pool threshold: this process has threads)
        paths += synth.write_scenarios(str(tmp_path / "many"), 15, start_index=i, num_agents=4, extent=80.0, spacing=2.0)
    pd = param_dict(dynamicsModel=2, maxNumControlledAgents=4)
    sim = _sim(paths[:2], pd, S=8)
    pinned = sim.pool_add(paths[2])                    # id 2, held by the caller across the churn
    assert pinned == 2
    for i in range(4, 150, 2):                         # 73 calls, 146 further distinct scenes
        sim.set_maps(paths[i:i + 2])
        sim.step()
    fresh_id = sim.pool_add(paths[3])                  # never seen before: takes a recycled id, not id >= 149
    assert fresh_id < 80, fresh_id
    names = ("shape_tensor", "map_observation_tensor", "expert_trajectory_tensor", "self_observation_tensor",
             "agent_roadmap_tensor", "partner_observations_tensor", "absolute_self_observation_tensor", "done_tensor")
    # scenes evicted long ago come back bit-identical to a fresh engine; the held id still binds to its scene
    sim.set_maps(paths[:2])
    ref = _sim(paths[:2], pd, S=8)
    for name in names:
        assert torch.equal(getattr(sim, name)().to_torch(), getattr(ref, name)().to_torch()), name
    sim.bind_worlds([0], [pinned])
    ref2 = _sim([paths[2], paths[1]], pd, S=8)
    for name in names:
        assert torch.equal(getattr(sim, name)().to_torch(), getattr(ref2, name)().to_torch()), name
    # and the most recent scenes were still cached (binding them again needs no upload: same behaviour either way)
    sim.set_maps(paths[148:150])
    ref3 = _sim(paths[148:150], pd, S=8)
    for _ in range(3):
        sim.step(); ref3.step()
    for name in names:
        assert torch.equal(getattr(sim, name)().to_torch(), getattr(ref3, name)().to_torch()), name


def test_errors_are_reported_not_fatal(tmp_path):
    scenes = unpack(["synth_a"], tmp_path)
    sim = _sim(scenes, param_dict())
    with pytest.raises(RuntimeError):
        sim.set_maps(scenes + scenes)                 # length must equal num_worlds (src/mgr.cpp:592)
    with pytest.raises(RuntimeError):
        sim.bev_observation_tensor()
    with pytest.raises(RuntimeError):
        sim.reset([5])
    sim.step()                                        # the handle is still usable


def test_policy_observation_matches_reference_python(tmp_path):
    """gd_policy_obs on the golden input tensors is bit-identical to the output of the
    reference's own Python post-processing, and on live engine tensors to its numpy restatement."""
    import os
    import torch
    from golden_util import GOLDEN
    from policy_obs_ref import policy_obs
    g = np.load(os.path.join(GOLDEN, "policy_obs.npz"))
    scenes = unpack(["synth_a", "synth_b"], tmp_path)
    sim = _sim(scenes, param_dict(dynamicsModel=2, maxNumControlledAgents=12))
    A = g["self_observation"].shape[1]
    so, po, rm = (sim.self_observation_tensor().to_torch(), sim.partner_observations_tensor().to_torch(),
                  sim.agent_roadmap_tensor().to_torch())
    so[:, :A].copy_(torch.from_numpy(g["self_observation"]))
    po[:, :A].copy_(torch.from_numpy(g["partner_observations"]))
    rm[:, :A].copy_(torch.from_numpy(g["agent_roadmap"]))
    torch.cuda.synchronize()
    for norm, key in ((True, "obs_norm"), (False, "obs_raw")):
        out = sim.policy_observation_tensor(normalize=norm).to_torch().cpu().numpy()
        assert out.shape == (2, 128, 3368)
        assert np.array_equal(out[:, :A].view(np.uint32), g[key].view(np.uint32)), key
    # live tensors after a few steps, every slot (padding slots included, like the Python path)
    for _ in range(3):
        sim.step()
    out = sim.policy_observation_tensor().to_torch().cpu().numpy()
    exp = policy_obs(so.cpu().numpy(), po.cpu().numpy(), rm.cpu().numpy(), True)
    assert np.array_equal(out.view(np.uint32), exp.view(np.uint32))


def test_binary_scenario_pool_equals_json(tmp_path):
    """Worlds created from / re-mapped to binary pool entries are bit-identical to JSON ones."""
    import torch
    from gpudrive_lab_b200 import scene_host
    sa, sb = unpack(["synth_a", "synth_b"], tmp_path)
    pd = param_dict(dynamicsModel=2, enableLidar=True)
    refs = scene_host.convert_pool([sa, sb], str(tmp_path / "s.gdpool"), pd["polylineReductionThreshold"])
    a, b = _sim([sa, sb], pd), _sim(refs, pd)
    names = ("shape_tensor", "map_observation_tensor", "expert_trajectory_tensor", "self_observation_tensor",
             "agent_roadmap_tensor", "partner_observations_tensor", "lidar_tensor", "absolute_self_observation_tensor",
             "metadata_tensor", "world_means_tensor", "map_name_tensor", "scenario_id_tensor")
    for _ in range(3):
        a.step(); b.step()
    for n in names:
        assert torch.equal(getattr(a, n)().to_torch(), getattr(b, n)().to_torch()), n
    a.set_maps([sb, sa]); b.set_maps([refs[1], refs[0]])
    a.step(); b.step()
    for n in names:
        assert torch.equal(getattr(a, n)().to_torch(), getattr(b, n)().to_torch()), n
    b.deleteAgents({0: [int(b.absolute_self_observation_tensor().to_torch()[0, 1, 13].item())]})
    a.deleteAgents({0: [int(a.absolute_self_observation_tensor().to_torch()[0, 1, 13].item())]})
    assert torch.equal(a.shape_tensor().to_torch(), b.shape_tensor().to_torch())


def test_shaped_reward_and_expert_actions_match_reference_python(tmp_path):
    """gd_shaped_reward / gd_expert_actions: on the golden inputs they reproduce what the
    reference's own get_rewards / get_expert_actions returned (tests/golden/make_env_post_golden.py), and on
    live tensors they equal the numpy restatement."""
    import os
    import torch
    import env_post_ref as R
    G = np.load(os.path.join(os.path.dirname(__file__), "golden", "env_post.npz"))
    sa, sb = unpack(["synth_a", "synth_b"], tmp_path)

    def bits(a):
        return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)
    kw = dict(collision_weight=-0.75, goal_achieved_weight=1.25, off_road_weight=-0.5, log_distance_weight=0.02)
    modes = ("sparse_on_goal_achieved", "weighted_combination", "reward_conditioned", "distance_to_logs")
    for model in range(4):
        sim = _sim([sa, sb], param_dict(dynamicsModel=model, collisionBehaviour=0, maxNumControlledAgents=12))
        A = G["info"].shape[1]
        # golden inputs written over the exports (slots >= A keep the engine's values)
        sim.info_tensor().to_torch()[:, :A] = torch.from_numpy(G["info"]).cuda()
        sim.reward_tensor().to_torch()[:, :A] = torch.from_numpy(G["reward"]).cuda()
        sim.expert_trajectory_tensor().to_torch()[:, :A] = torch.from_numpy(G["expert_trajectory"]).cuda()
        sim.absolute_self_observation_tensor().to_torch()[:, :A] = torch.from_numpy(G["absolute_self_observation"]).cuda()
        acts = sim.expert_actions_tensor().to_torch().cpu().numpy()
        assert np.array_equal(bits(acts[:, :A]), bits(G["expert_actions_%d" % model])), model
        if model:
            continue
        w = torch.zeros((2, 128, 3), dtype=torch.float32, device="cuda")
        w[:, :A] = torch.from_numpy(G["agent_weights"]).cuda()
        wts = torch.from_numpy(G["world_time_steps"]).cuda()
        for mode in modes:
            r = sim.shaped_reward_tensor(mode, agent_weights=w, world_time_steps=wts, **kw).to_torch().cpu().numpy()
            want = G["reward_" + mode]
            if mode == "distance_to_logs":
                assert np.allclose(r[:, :A], want, rtol=1e-6, atol=1e-9), mode
            else:
                assert np.array_equal(bits(r[:, :A]), bits(want)), mode
    # live tensors, default time step (episode length - steps remaining)
    sim = _sim([sa, sb], param_dict(dynamicsModel=0, collisionBehaviour=0, maxNumControlledAgents=12))
    for t in range(12):
        sim.action_tensor().to_torch().copy_(torch.from_numpy(random_actions(t, 2, 128, 0, seed=23)).cuda())
        sim.step()
    snap = {n: getattr(sim, n + "_tensor")().to_torch().cpu().numpy()
            for n in ("info", "reward", "expert_trajectory", "absolute_self_observation")}
    for mode in ("weighted_combination", "distance_to_logs"):
        r = sim.shaped_reward_tensor(mode, **kw).to_torch().cpu().numpy()
        want = R.shaped_reward(snap["info"], snap["reward"], snap["expert_trajectory"], snap["absolute_self_observation"],
                               mode, world_time_steps=np.array([12, 12]), **kw)
        assert np.allclose(r, want, rtol=1e-6, atol=1e-9), mode


def test_headless_cli_runs(tmp_path):
    """The C++ driver steps real worlds through the C ABI and reports the reference's two FPS lines."""
    import re
    import subprocess
    from gpudrive_lab_b200 import build
    exe = build.HEADLESS if __import__("os").path.exists(build.HEADLESS) else build.build_headless()
    sa, sb = unpack(["synth_a", "synth_b"], tmp_path)
    r = subprocess.run([exe, "CUDA", "20", "--rand-actions", "--worlds", "6", "--scenes", sa, sb],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    fps = float(re.search(r"^FPS ([0-9.]+)", r.stdout, re.M).group(1))
    afps = float(re.search(r"^Agent-Normalized FPS ([0-9.]+)", r.stdout, re.M).group(1))
    m = re.search(r"steps (\d+) agent_steps (\d+)", r.stdout)
    assert fps > 0 and afps > fps and int(m.group(1)) == 20 and int(m.group(2)) > 0
