"""Golden vectors for gd_shaped_reward / gd_expert_actions: the reference's own GPUDriveTorchEnv.get_rewards and
get_expert_actions (gpudrive/env/env_torch.py:469-604, 1445-1509), called unbound on the CPU over tensors of
the unmodified reference Manager. Run in the build container:

    python tests/golden/make_env_post_golden.py

gymnasium and the visualisation stack are not installed here; they are stubbed (the two functions use neither).
Writes env_post.npz: the input tensors (cropped to 16 agent slots) and the expected outputs."""
import importlib.abc
import importlib.machinery
import os
import sys
import tempfile
import types
from unittest import mock

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, "/root/reference")

import gpudrive_lab_b200  # noqa: E402

gpudrive_lab_b200.install()
gym = types.ModuleType("gymnasium"); gym.Env = type("Env", (), {})
spaces = types.ModuleType("gymnasium.spaces")
for n in ("Box", "Discrete", "Tuple"):
    setattr(spaces, n, mock.MagicMock())
gym.spaces = spaces
sys.modules["gymnasium"] = gym; sys.modules["gymnasium.spaces"] = spaces


class _Stub(importlib.abc.MetaPathFinder, importlib.abc.Loader):
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


sys.meta_path.insert(0, _Stub())
from gpudrive.env.env_torch import GPUDriveTorchEnv  # noqa: E402

from golden_util import unpack  # noqa: E402
from helpers import param_dict, random_actions, ref_params  # noqa: E402


class _CpuTensor(torch.Tensor):
    """The reference moves some clones to its default device "cuda"; there is none here."""
    def to(self, *args, **kwargs):
        args = tuple("cpu" if (isinstance(a, str) and a.startswith("cuda")) else a for a in args)
        return super().to(*args, **kwargs)


class _T:
    def __init__(self, a):
        self.a = a

    def to_torch(self):
        return torch.from_numpy(np.array(self.a, copy=True)).as_subclass(_CpuTensor)


class _Sim:
    def __init__(self, t):
        self.t = t

    def info_tensor(self): return _T(self.t["info"])
    def reward_tensor(self): return _T(self.t["reward"])
    def expert_trajectory_tensor(self): return _T(self.t["expert_trajectory"])
    def absolute_self_observation_tensor(self): return _T(self.t["absolute_self_observation"])


def main():
    from oracle.refmgr import RefManager
    tmp = tempfile.mkdtemp()
    scenes = unpack(["synth_a", "synth_b"], tmp)
    # AgentStop + random actions: some agents collide / go off road, so the Info flags are not all zero
    pd = param_dict(dynamicsModel=0, collisionBehaviour=0, maxNumControlledAgents=12)
    ref = RefManager(scenes, ref_params(pd))
    steps = 25
    for t in range(steps):
        ref.tensor("action")[...] = random_actions(t, 2, 128, 0, seed=23)
        ref.step()
    A = 16
    t = {k: ref.tensor(k)[:, :A].copy() for k in ("info", "reward", "expert_trajectory", "absolute_self_observation")}
    W = 2
    out = dict(t)
    rng = np.random.default_rng(5)
    weights = rng.uniform(-1, 1, (W, A, 3)).astype(np.float32)
    wts = np.array([steps, 7], np.int32)
    out["agent_weights"] = weights; out["world_time_steps"] = wts
    out["info_sum"] = np.array([int(t["info"][..., :4].sum())])
    for mode in ("sparse_on_goal_achieved", "weighted_combination", "reward_conditioned", "distance_to_logs"):
        me = types.SimpleNamespace(sim=_Sim(t), config=types.SimpleNamespace(reward_type=mode), num_worlds=W,
                                   max_agent_count=A, backend="torch", device="cpu",
                                   reward_weights_tensor=torch.from_numpy(weights))
        r = GPUDriveTorchEnv.get_rewards(me, collision_weight=-0.75, goal_achieved_weight=1.25, off_road_weight=-0.5,
                                         world_time_steps=torch.from_numpy(wts).long(), log_distance_weight=0.02)
        out["reward_" + mode] = r.numpy().astype(np.float32)
    for name, model in (("classic", 0), ("bicycle", 1), ("delta_local", 2), ("state", 3)):
        me = types.SimpleNamespace(sim=_Sim(t), config=types.SimpleNamespace(dynamics_model=name), num_worlds=W,
                                   max_agent_count=A, backend="torch", device="cpu")
        acts = GPUDriveTorchEnv.get_expert_actions(me)[0]
        out["expert_actions_%d" % model] = acts.numpy().astype(np.float32)
    np.savez_compressed(os.path.join(HERE, "env_post.npz"), **out)
    print({k: v.shape for k, v in out.items()}, "info flags set:", out["info_sum"])


if __name__ == "__main__":
    main()
