"""-m gpu: the reference's own `GPUDriveTorchEnv` (gpudrive/env/env_torch.py), UNMODIFIED, running on the B200
engine through `gpudrive_lab_b200.install()` - north_star: "gpudrive/env's GpuDriveTorchEnv ... run unchanged".

The env package is pure Python; __graft_entry__.build() installs a copy of it into the git-ignored
baseline/_ref/ (it travels to the GPU box like oracle/_ref). gymnasium, mediapy and the visualisation /
VBD stacks are not in this image and are not on the path under test: they are stubbed exactly as
tests/golden/make_env_post_golden.py stubs them. What runs for real is the env's call pattern against the
simulator: keyword construction with batch_render_view_width=None (base_env.py:160-191), sim.reset(list)
(env_torch.py:421-425), in-place `action_tensor().to_torch()[:, :, :3].copy_()` + sim.step() (:606-660),
get_obs / get_rewards / get_dones / get_infos (:426-604,1172-1216), deleteAgents (:1339) and set_maps via
swap_data_batch (:1351-1370).
"""
import importlib.abc
import importlib.machinery
import os
import sys
import types
from unittest import mock

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_PY = os.path.join(ROOT, "baseline", "_ref")


class _Box:
    def __init__(self, low=None, high=None, shape=None, dtype=None):
        self.low, self.high, self.shape, self.dtype = low, high, tuple(shape or ()), dtype


class _Discrete:
    def __init__(self, n):
        self.n = int(n)


class _Tuple:
    def __init__(self, spaces):
        self.spaces = tuple(spaces)


class _Stub(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    """Import stand-ins for packages that are absent here and unused on the path under test."""
    PREFIX = ("gpudrive.visualize", "matplotlib", "mediapy", "jax", "pygame", "seaborn", "huggingface_hub", "box",
              "wandb", "gpudrive.integrations", "imageio", "PIL", "cv2")

    def find_spec(self, name, path, target=None):
        if any(name == p or name.startswith(p + ".") for p in self.PREFIX):
            return importlib.machinery.ModuleSpec(name, self, is_package=True)

    def create_module(self, spec):
        m = mock.MagicMock(); m.__path__ = []; m.__name__ = spec.name; m.__spec__ = spec
        return m

    def exec_module(self, module):
        pass


def import_reference_env():
    """Import gpudrive.env.env_torch from baseline/_ref over `gpudrive_lab_b200.install()`. Returns the env
    classes and the installed simulator module (whose SimManager attribute the parity run below swaps for the
    reference-backed one and back: the env looks it up at construction time)."""
    if not os.path.isdir(os.path.join(REF_PY, "gpudrive", "env")):
        pytest.fail("baseline/_ref/gpudrive is missing: __graft_entry__.build() installs it where /root/reference exists")
    import gpudrive_lab_b200
    simulator = gpudrive_lab_b200.install()
    if "gymnasium" not in sys.modules:
        gym = types.ModuleType("gymnasium"); gym.Env = type("Env", (), {})
        spaces = types.ModuleType("gymnasium.spaces")
        spaces.Box, spaces.Discrete, spaces.Tuple = _Box, _Discrete, _Tuple
        gym.spaces = spaces
        sys.modules["gymnasium"] = gym; sys.modules["gymnasium.spaces"] = spaces
    if not any(isinstance(f, _Stub) for f in sys.meta_path):
        sys.meta_path.insert(0, _Stub())
    if REF_PY not in sys.path:
        sys.path.insert(0, REF_PY)
    from gpudrive.env.config import EnvConfig
    from gpudrive.env.dataset import SceneDataLoader
    from gpudrive.env.env_torch import GPUDriveTorchEnv
    return GPUDriveTorchEnv, EnvConfig, SceneDataLoader, simulator


def make_dataset(out_dir, count, num_agents=12):
    """Synthetic scenes under the file prefix SceneDataLoader filters on ("tfrecord")."""
    import shutil
    from gpudrive_lab_b200 import synth
    src = synth.write_scenarios(os.path.join(out_dir, "src"), count, num_agents=num_agents)
    d = os.path.join(out_dir, "data")
    os.makedirs(d, exist_ok=True)
    for i, p in enumerate(src):
        shutil.copy(p, os.path.join(d, f"tfrecord-{i:05d}.json"))
    return d


def run_env_rollout(GPUDriveTorchEnv, EnvConfig, SceneDataLoader, data_dir, device, steps=20, seed=0):
    """reset -> `steps` x (step_dynamics, get_obs, get_rewards, get_dones, get_infos) -> deleteAgents ->
    swap_data_batch -> reset, exactly through the env's public methods. Returns what it observed."""
    import torch
    torch.manual_seed(seed)       # remove_agents_by_id samples with torch.randperm
    cfg = EnvConfig(dynamics_model="classic", reward_type="weighted_combination", collision_weight=-0.5,
                    goal_achieved_weight=1.0, off_road_weight=-0.5, dist_to_goal_threshold=2.0)
    loader = SceneDataLoader(root=data_dir, batch_size=4, dataset_size=8, sample_with_replacement=False)
    env = GPUDriveTorchEnv(config=cfg, data_loader=loader, max_cont_agents=64, device=device)
    out = {"obs": [], "rew": [], "done": [], "info": []}
    obs = env.reset()
    out["obs"].append(obs.cpu().numpy())
    assert obs.shape == (4, 128, 6 + 127 * 6 + 200 * 13)
    g = torch.Generator().manual_seed(seed)
    n_act = env.action_space.n
    assert n_act == 7 * 13
    for t in range(steps):
        actions = torch.randint(0, n_act, (4, 128), generator=g).to(device)
        env.step_dynamics(actions)
        out["obs"].append(env.get_obs().cpu().numpy())
        out["rew"].append(env.get_rewards(collision_weight=-0.5, goal_achieved_weight=1.0, off_road_weight=-0.5).cpu().numpy())
        out["done"].append(env.get_dones().cpu().numpy())
        info = env.get_infos()
        out["info"].append(np.stack([info.off_road.cpu().numpy(), info.collided.cpu().numpy(),
                                     info.goal_achieved.cpu().numpy()], -1))
    out["mask0"] = env.cont_agent_mask.cpu().numpy().copy()
    env.remove_agents_by_id(0.25, remove_controlled_agents=True)        # -> sim.deleteAgents
    out["mask_after_delete"] = env.cont_agent_mask.cpu().numpy().copy()
    out["obs_after_delete"] = env.get_obs().cpu().numpy()
    env.swap_data_batch()                                              # -> sim.set_maps(next batch)
    out["obs_after_swap"] = env.reset().cpu().numpy()
    out["mask_after_swap"] = env.cont_agent_mask.cpu().numpy().copy()
    return env, out


def test_reference_env_runs_unchanged_on_the_engine(synth_dir):
    import torch
    assert torch.cuda.is_available()
    GPUDriveTorchEnv, EnvConfig, SceneDataLoader, simulator = import_reference_env()
    data = make_dataset(os.path.join(synth_dir, "env_ds"), 8)
    env, out = run_env_rollout(GPUDriveTorchEnv, EnvConfig, SceneDataLoader, data, "cuda")
    sim = env.sim
    assert isinstance(sim, simulator.SimManager)
    # the rollout did something: agents moved, rewards / dones / infos have the env's shapes
    assert out["obs"][0].shape == out["obs"][-1].shape and not np.array_equal(out["obs"][0], out["obs"][-1])
    assert out["rew"][-1].shape == (4, 128) and out["done"][-1].shape == (4, 128) and out["info"][-1].shape == (4, 128, 3)
    per_world = out["mask0"].sum(1)
    assert per_world.min() >= 10
    assert out["mask_after_delete"].sum() == sum(int(n) - max(1, int(0.25 * int(n))) for n in per_world)
    assert out["mask_after_swap"].sum() >= 40                         # fresh scenes, nothing deleted
    assert not np.array_equal(out["obs_after_swap"], out["obs"][0])   # the second batch of scenes
    # get_obs() of the reference env == the engine's device-side policy observation, bit for bit
    want = env.get_obs()
    got = sim.policy_observation_tensor(normalize=True, torch_cuda_arithmetic=True).to_torch()   # env device = "cuda"
    assert got.shape == want.shape
    assert torch.equal(got.view(torch.int32), want.contiguous().view(torch.int32))
    # get_rewards() == gd_shaped_reward
    r_env = env.get_rewards(collision_weight=-0.5, goal_achieved_weight=1.0, off_road_weight=-0.5)
    r_dev = sim.shaped_reward_tensor("weighted_combination", collision_weight=-0.5, goal_achieved_weight=1.0,
                                     off_road_weight=-0.5).to_torch()
    assert torch.equal(r_env, r_dev)
    # and 20 more steps on the swapped batch, then a partial reset by index
    g = torch.Generator().manual_seed(1)
    for t in range(20):
        env.step_dynamics(torch.randint(0, env.action_space.n, (4, 128), generator=g).cuda())
    env.reset(env_idx_list=[1, 3])
    t_left = sim.steps_remaining_tensor().to_torch()[:, 0, 0].tolist()
    assert t_left == [91 - 20, 91, 91 - 20, 91]
    env.close()

    # ---- the same rollout of the same env on the reference's own simulator (oracle/_ref, CPU): every
    # observation / reward / done / info the env returned must agree. The env default is the "linear" road
    # algorithm, whose row order is part of the contract, so observations are compared element for element.
    from ref_sim import RefSimManager
    engine_cls = simulator.SimManager
    simulator.SimManager = RefSimManager
    try:
        env_r, ref = run_env_rollout(GPUDriveTorchEnv, EnvConfig, SceneDataLoader, data, "cpu")
    finally:
        simulator.SimManager = engine_cls
    assert isinstance(env_r.sim, RefSimManager)
    for k in ("mask0", "mask_after_delete", "mask_after_swap"):
        assert np.array_equal(out[k], ref[k]), k
    for t in range(len(ref["done"])):
        assert np.array_equal(out["done"][t], ref["done"][t]), ("done", t)
        assert np.array_equal(out["info"][t], ref["info"][t]), ("info", t)
        assert np.array_equal(out["rew"][t], ref["rew"][t]), ("reward", t)
    worst = 0.0
    for name, a_list, b_list in (("obs", out["obs"], ref["obs"]),
                                 ("obs_after_delete", [out["obs_after_delete"]], [ref["obs_after_delete"]]),
                                 ("obs_after_swap", [out["obs_after_swap"]], [ref["obs_after_swap"]])):
        for t, (a, b) in enumerate(zip(a_list, b_list)):
            # normalised observations (positions / 1000, headings / 2 pi wrap at +-0.5): 1e-5 absolute
            d = np.abs(a.astype(np.float64) - b.astype(np.float64))
            d = np.minimum(d, np.abs(d - 1.0))
            worst = max(worst, float(d.max()))
            assert d.max() <= 1e-5, (name, t, float(d.max()), np.argwhere(d > 1e-5)[:4].tolist())
    print(f"env rollout on the engine vs on the reference simulator: max |obs difference| = {worst:.2e}")
