"""Drop-in stand-in for the reference's nanobind module `madrona_gpudrive`
(src/bindings.cpp:16-150) backed by the B200 engine through its C ABI.

    import gpudrive_lab_b200.madrona_gpudrive as madrona_gpudrive        # or
    gpudrive_lab_b200.install()   # registers it as sys.modules["madrona_gpudrive"]

Same names, argument meaning and tensor layouts as the reference. Differences, all explicit:
  * ExecMode.CPU raises: this engine has no CPU path (the reference's CPU mode is oracle/_ref).
  * enable_batch_renderer=True, rgb_tensor(), depth_tensor() raise: no graphics stack on this box.
  * bev_observation_tensor() needs SimManager(..., enable_bev=True): the grid is 160 KB per agent slot,
    so it is opt-in here (the reference always computes it).
  * agent_roadmap rows: with AllEntitiesWithRadiusFiltering (the env default, "linear") they come in
    road-entity index order exactly like the reference's (src/sim.cpp:261-279). With
    KNearestEntitiesWithRadiusFiltering the reference leaves them in its internal heap order, which
    already differs between its debug and release builds; here they come by distance band, then by
    position in the spatially sorted road table. The set of rows per agent is identical;
    knn_index_tensor() gives the road-entity index of every row.
  * Streams: the engine launches on a blocking CUDA stream, so tensor writes made with torch on its
    default stream are ordered before step() / reset() without an explicit synchronize, as with the
    reference.
"""
from __future__ import annotations

import ctypes as C
import enum
import types

from . import _capi

# module attributes, src/bindings.cpp:23-28
kMaxAgentCount = 128
kMaxRoadEntityCount = 10000
kMaxAgentMapObservationsCount = 200
episodeLen = 91
numLidarSamples = 50
vehicleScale = 0.7


class RewardType(enum.IntEnum):          # src/init.hpp:76-81
    DistanceBased = 0
    OnGoalAchieved = 1
    Dense = 2


class CollisionBehaviour(enum.IntEnum):  # src/init.hpp:90-95
    AgentStop = 0
    AgentRemoved = 1
    Ignore = 2


class DynamicsModel(enum.IntEnum):       # src/init.hpp:97-103
    Classic = 0
    InvertibleBicycle = 1
    DeltaLocal = 2
    State = 3


class FindRoadObservationsWith(enum.IntEnum):   # src/init.hpp:105-109
    KNearestEntitiesWithRadiusFiltering = 0
    AllEntitiesWithRadiusFiltering = 1


class EntityType(enum.IntEnum):          # src/types.hpp:24-38, python names src/bindings.cpp:76-88
    _None = 0
    RoadEdge = 1
    RoadLine = 2
    RoadLane = 3
    CrossWalk = 4
    SpeedBump = 5
    StopSign = 6
    Vehicle = 7
    Pedestrian = 8
    Cyclist = 9
    Padding = 10
    NumTypes = 11


class RewardParams:                      # src/init.hpp:83-88
    def __init__(self):
        self.rewardType = RewardType.DistanceBased
        self.distanceToGoalThreshold = 0.0
        self.distanceToExpertThreshold = 0.0


class Parameters:                        # src/init.hpp:111-127 (same defaults)
    def __init__(self):
        self.polylineReductionThreshold = 0.0
        self.observationRadius = 0.0
        self.rewardParams = RewardParams()
        self.collisionBehaviour = CollisionBehaviour.AgentStop
        self.maxNumControlledAgents = 10000
        self.IgnoreNonVehicles = False
        self.roadObservationAlgorithm = FindRoadObservationsWith.KNearestEntitiesWithRadiusFiltering
        self.initOnlyValidAgentsAtFirstStep = True
        self.isStaticAgentControlled = False
        self.enableLidar = False
        self.disableClassicalObs = False
        self.dynamicsModel = DynamicsModel.Classic
        self.readFromTracksToPredict = False

    def _to_c(self) -> _capi.GdParams:
        rp = self.rewardParams
        return _capi.GdParams(
            float(self.polylineReductionThreshold), float(self.observationRadius), int(rp.rewardType),
            float(rp.distanceToGoalThreshold), float(rp.distanceToExpertThreshold),
            int(self.collisionBehaviour), int(self.maxNumControlledAgents), int(bool(self.IgnoreNonVehicles)),
            int(self.roadObservationAlgorithm), int(bool(self.initOnlyValidAgentsAtFirstStep)),
            int(bool(self.isStaticAgentControlled)), int(bool(self.enableLidar)),
            int(bool(self.disableClassicalObs)), int(self.dynamicsModel), int(bool(self.readFromTracksToPredict)))


class _ExecMode(enum.IntEnum):           # madrona::ExecMode
    CPU = 0
    CUDA = 1


class Tensor:
    """madrona.Tensor (external/madrona/src/python/bindings.cpp:316-352): zero-copy view of an
    export buffer that stays valid (and aliased) across step()/reset()."""

    def __init__(self, desc: _capi.GdTensorDesc, owner):
        self._ptr = desc.ptr
        self._shape = tuple(int(desc.shape[i]) for i in range(desc.ndim))
        self._dtype = int(desc.dtype)
        self._device = int(desc.device)
        self._owner = owner          # keeps the SimManager (and its device memory) alive

    @property
    def __cuda_array_interface__(self):
        return {"shape": self._shape, "typestr": "<f4" if self._dtype == 0 else "<i4",
                "data": (self._ptr, False), "version": 2, "strides": None}

    def to_torch(self):
        import torch
        t = torch.as_tensor(self, device=torch.device("cuda", self._device))
        t._gd_owner = self._owner
        return t

    def to_jax(self):
        raise RuntimeError("to_jax(): jax is not installed in this image")

    def device_ptr(self) -> int:
        return self._ptr

    @property
    def shape(self):
        return self._shape


madrona = types.ModuleType(__name__ + ".madrona")
madrona.ExecMode = _ExecMode
madrona.Tensor = Tensor


class SimManager:
    """src/bindings.cpp:91-149. num_worlds = len(scenes); the constructor performs the first reset."""

    def __init__(self, exec_mode, gpu_id, scenes, params, enable_batch_renderer=False,
                 batch_render_view_width=64, batch_render_view_height=64, max_agents=None, enable_bev=False):
        if int(exec_mode) != int(_ExecMode.CUDA):
            raise RuntimeError("gpudrive_lab_b200 implements ExecMode.CUDA only; there is no CPU fallback")
        if enable_batch_renderer:
            raise RuntimeError("batch renderer: unavailable on this box (no graphics stack)")
        lib = _capi.load()
        self._lib = lib
        self._h = C.c_void_p()
        cfg = _capi.GdConfig(int(gpu_id), int(max_agents if max_agents is not None else kMaxAgentCount),
                             params._to_c(), 1 if enable_bev else 0)
        arr = (C.c_char_p * len(scenes))(*[str(s).encode() for s in scenes])
        _capi.check(lib.gd_create(C.byref(cfg), arr, len(scenes), C.byref(self._h)))
        self.num_worlds = len(scenes)

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            self._lib.gd_destroy(h)

    # ---- control
    def step(self):
        _capi.check(self._lib.gd_step(self._h))

    def step_async(self):
        _capi.check(self._lib.gd_step_async(self._h))

    def sync(self):
        _capi.check(self._lib.gd_sync(self._h))

    def reset(self, worlds):
        worlds = [int(w) for w in worlds]
        a = (C.c_int32 * max(len(worlds), 1))(*worlds)
        _capi.check(self._lib.gd_reset(self._h, a, len(worlds)))

    def set_maps(self, scenes):
        arr = (C.c_char_p * len(scenes))(*[str(s).encode() for s in scenes])
        _capi.check(self._lib.gd_set_maps(self._h, arr, len(scenes)))

    def deleteAgents(self, agents_to_delete: dict):
        for world, ids in agents_to_delete.items():
            ids = [int(i) for i in ids]
            a = (C.c_int32 * max(len(ids), 1))(*ids)
            _capi.check(self._lib.gd_delete_agents(self._h, int(world), a, len(ids)))

    def step_host(self, actions, reward=None, done=None, info=None):
        """Host-buffer step (gd_step_host): `actions` is a CPU torch/numpy array [W,S,cols]."""
        cols = int(actions.shape[-1])
        ptr = lambda x: None if x is None else (x.data_ptr() if hasattr(x, "data_ptr") else x.ctypes.data)
        _capi.check(self._lib.gd_step_host(self._h, ptr(actions), cols, ptr(reward), ptr(done), ptr(info)))

    def pool_add(self, path) -> int:
        out = C.c_int32()
        _capi.check(self._lib.gd_pool_add(self._h, str(path).encode(), C.byref(out)))
        return out.value

    def bind_worlds(self, worlds, scene_ids):
        n = len(worlds)
        a = (C.c_int32 * max(n, 1))(*[int(w) for w in worlds])
        b = (C.c_int32 * max(n, 1))(*[int(s) for s in scene_ids])
        _capi.check(self._lib.gd_bind_worlds(self._h, a, b, n))

    def auto_reset_pool(self, scene_ids):
        """Register the pooled scene ids (pool_add) the device-side auto-reset cycles through."""
        n = len(scene_ids)
        a = (C.c_int32 * max(n, 1))(*[int(s) for s in scene_ids])
        _capi.check(self._lib.gd_auto_reset_pool(self._h, a, n))

    def auto_reset(self):
        """After a step: every world whose controlled agents are all done moves on to its next pooled scenario,
        is reset and has its observations recomputed - on the device, stream-ordered, no host round trip."""
        _capi.check(self._lib.gd_auto_reset(self._h))

    def stats(self) -> dict:
        s = _capi.GdStats()
        _capi.check(self._lib.gd_stats(self._h, C.byref(s)))
        return {n: int(getattr(s, n)) for n, _ in s._fields_}

    def set_option(self, option: int, value: int):
        _capi.check(self._lib.gd_set_option(self._h, int(option), int(value)))

    def kernel_times(self) -> dict:
        a, o, n = C.c_double(), C.c_double(), C.c_int64()
        _capi.check(self._lib.gd_kernel_times(self._h, C.byref(a), C.byref(o), C.byref(n)))
        return {"k_agents_ms": a.value, "k_obs_ms": o.value, "steps": n.value}

    def policy_observation_tensor(self, normalize=True, torch_cuda_arithmetic=False) -> Tensor:
        """Flat policy observation [W, S, 6 + (S-1)*6 + 200*13] (GPUDriveTorchEnv.get_obs defaults), computed on
        the device from the current exports; recomputed at every call. torch divides a tensor by a Python scalar
        exactly on the CPU and by multiplying with the float reciprocal on CUDA: torch_cuda_arithmetic selects
        which of the two get_obs() results is reproduced bit for bit (default: the CPU one, pinned by goldens)."""
        d = _capi.GdTensorDesc()
        mode = 0 if not normalize else (2 if torch_cuda_arithmetic else 1)
        _capi.check(self._lib.gd_policy_obs(self._h, mode, C.byref(d)))
        return Tensor(d, self)

    _REWARD_MODES = {"sparse_on_goal_achieved": 0, "weighted_combination": 1, "reward_conditioned": 2,
                     "distance_to_logs": 3}

    def shaped_reward_tensor(self, reward_type="weighted_combination", collision_weight=-0.5,
                             goal_achieved_weight=1.0, off_road_weight=-0.5, log_distance_weight=0.01,
                             agent_weights=None, world_time_steps=None) -> Tensor:
        """GPUDriveTorchEnv.get_rewards (gpudrive/env/env_torch.py:469-604) in one device pass: f32 [W, S].
        agent_weights: CUDA float32 tensor [W, S, 3] for "reward_conditioned"; world_time_steps: CUDA int32
        tensor [W] for "distance_to_logs" (default: episode length - steps remaining)."""
        def ptr(t, dtype):
            if t is None:
                return None
            import torch
            assert t.is_cuda and t.dtype == dtype and t.is_contiguous(), "need a contiguous CUDA tensor"
            return t.data_ptr()
        import torch
        cfg = _capi.GdRewardShaping(self._REWARD_MODES[reward_type], collision_weight, goal_achieved_weight,
                                    off_road_weight, log_distance_weight, ptr(agent_weights, torch.float32),
                                    ptr(world_time_steps, torch.int32))
        d = _capi.GdTensorDesc()
        _capi.check(self._lib.gd_shaped_reward(self._h, C.byref(cfg), C.byref(d)))
        return Tensor(d, self)

    def expert_actions_tensor(self) -> Tensor:
        """GPUDriveTorchEnv.get_expert_actions()[0] (env_torch.py:1445-1509): f32 [W, S, 91, 3] clamped inferred
        actions, or [W, S, 91, 10] for the State dynamics model."""
        d = _capi.GdTensorDesc()
        _capi.check(self._lib.gd_expert_actions(self._h, C.byref(d)))
        return Tensor(d, self)

    def launch_count(self) -> int:
        return int(self._lib.gd_launch_count(self._h))

    def cuda_stream(self) -> int:
        return int(self._lib.gd_stream(self._h))

    # ---- tensors
    def _tensor(self, name) -> Tensor:
        d = _capi.GdTensorDesc()
        _capi.check(self._lib.gd_tensor(self._h, _capi.TENSOR_IDS[name], C.byref(d)))
        return Tensor(d, self)

    def action_tensor(self): return self._tensor("action")
    def reward_tensor(self): return self._tensor("reward")
    def done_tensor(self): return self._tensor("done")
    def self_observation_tensor(self): return self._tensor("self_observation")
    def map_observation_tensor(self): return self._tensor("map_observation")
    def partner_observations_tensor(self): return self._tensor("partner_observations")
    def lidar_tensor(self): return self._tensor("lidar")
    def steps_remaining_tensor(self): return self._tensor("steps_remaining")
    def shape_tensor(self): return self._tensor("shape")
    def controlled_state_tensor(self): return self._tensor("controlled_state")
    def agent_roadmap_tensor(self): return self._tensor("agent_roadmap")
    def absolute_self_observation_tensor(self): return self._tensor("absolute_self_observation")
    def bev_observation_tensor(self): return self._tensor("bev_observation")
    def valid_state_tensor(self): return self._tensor("valid_state")
    def info_tensor(self): return self._tensor("info")
    def response_type_tensor(self): return self._tensor("response_type")
    def expert_trajectory_tensor(self): return self._tensor("expert_trajectory")
    def world_means_tensor(self): return self._tensor("world_means")
    def metadata_tensor(self): return self._tensor("metadata")
    def map_name_tensor(self): return self._tensor("map_name")
    def scenario_id_tensor(self): return self._tensor("scenario_id")
    def deleted_agents_tensor(self): return self._tensor("deleted_agents")
    def knn_index_tensor(self): return self._tensor("knn_index")

    def rgb_tensor(self):
        raise RuntimeError("rgb_tensor: batch renderer unavailable on this box (no graphics stack)")

    def depth_tensor(self):
        raise RuntimeError("depth_tensor: batch renderer unavailable on this box (no graphics stack)")
