"""ctypes front-end for oracle/_ref/libgd_ref.so — the UNMODIFIED reference
`Manager` (src/mgr.hpp:28-100) built in ExecMode::CPU by oracle/ref/Makefile.

TEST INFRASTRUCTURE. Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this module. It is the checker
and the timed CPU baseline, never part of the product path.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libgd_ref.so")


class _Params(C.Structure):
    # mirrors GdRefParams in oracle/ref/ref_capi.cpp (== Parameters, src/init.hpp:111-127)
    _fields_ = [
        ("polylineReductionThreshold", C.c_float),
        ("observationRadius", C.c_float),
        ("rewardType", C.c_int32),
        ("distanceToGoalThreshold", C.c_float),
        ("distanceToExpertThreshold", C.c_float),
        ("collisionBehaviour", C.c_int32),
        ("maxNumControlledAgents", C.c_uint32),
        ("ignoreNonVehicles", C.c_int32),
        ("roadObservationAlgorithm", C.c_int32),
        ("initOnlyValidAgentsAtFirstStep", C.c_int32),
        ("isStaticAgentControlled", C.c_int32),
        ("enableLidar", C.c_int32),
        ("disableClassicalObs", C.c_int32),
        ("dynamicsModel", C.c_int32),
        ("readFromTracksToPredict", C.c_int32),
    ]


class _Tensor(C.Structure):
    _fields_ = [("ptr", C.c_void_p), ("dtype", C.c_int32), ("ndim", C.c_int32),
                ("shape", C.c_int64 * 8)]


@dataclass
class RefParams:
    """Field names and defaults follow src/init.hpp:111-127."""
    polylineReductionThreshold: float = 0.0
    observationRadius: float = 0.0
    rewardType: int = 0
    distanceToGoalThreshold: float = 0.0
    distanceToExpertThreshold: float = 0.0
    collisionBehaviour: int = 0          # AgentStop
    maxNumControlledAgents: int = 10000
    IgnoreNonVehicles: bool = False
    roadObservationAlgorithm: int = 0    # KNearestEntitiesWithRadiusFiltering
    initOnlyValidAgentsAtFirstStep: bool = True
    isStaticAgentControlled: bool = False
    enableLidar: bool = False
    disableClassicalObs: bool = False
    dynamicsModel: int = 0               # Classic
    readFromTracksToPredict: bool = False


_lib = None


def available() -> bool:
    return os.path.exists(LIB_PATH)


def _load():
    global _lib
    if _lib is None:
        if "GD_REF_ASSETS" not in os.environ:
            ref_assets = "/root/reference/assets/"
            os.environ["GD_REF_ASSETS"] = (
                ref_assets if os.path.isdir(ref_assets)
                else os.path.join(_HERE, "ref", "assets") + "/")
        lib = C.CDLL(LIB_PATH)
        lib.gd_ref_create.restype = C.c_void_p
        lib.gd_ref_create.argtypes = [C.POINTER(C.c_char_p), C.c_int, C.POINTER(_Params)]
        lib.gd_ref_destroy.argtypes = [C.c_void_p]
        lib.gd_ref_step.argtypes = [C.c_void_p]
        lib.gd_ref_reset.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.c_int]
        lib.gd_ref_set_maps.argtypes = [C.c_void_p, C.POINTER(C.c_char_p), C.c_int]
        lib.gd_ref_delete_agents.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_int32), C.c_int]
        lib.gd_ref_tensor.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(_Tensor)]
        lib.gd_ref_tensor.restype = C.c_int
        lib.gd_ref_consts.argtypes = [C.POINTER(C.c_int32)]
        _lib = lib
    return _lib


def consts():
    out = (C.c_int32 * 4)()
    _load().gd_ref_consts(out)
    return {"kMaxAgentCount": out[0], "kMaxRoadEntityCount": out[1],
            "kMaxAgentMapObservationsCount": out[2], "numLidarSamples": out[3]}


TENSOR_NAMES = [
    "action", "reward", "done", "self_observation", "map_observation",
    "partner_observations", "agent_roadmap", "lidar", "bev_observation",
    "steps_remaining", "shape", "controlled_state", "absolute_self_observation",
    "info", "response_type", "expert_trajectory", "world_means", "metadata",
    "deleted_agents", "map_name", "scenario_id",
]


class RefManager:
    """The reference SimManager surface (src/bindings.cpp:91-149) over numpy views."""

    def __init__(self, scenes, params: RefParams):
        lib = _load()
        p = _Params(
            params.polylineReductionThreshold, params.observationRadius, params.rewardType,
            params.distanceToGoalThreshold, params.distanceToExpertThreshold,
            params.collisionBehaviour, params.maxNumControlledAgents,
            int(params.IgnoreNonVehicles), params.roadObservationAlgorithm,
            int(params.initOnlyValidAgentsAtFirstStep), int(params.isStaticAgentControlled),
            int(params.enableLidar), int(params.disableClassicalObs), params.dynamicsModel,
            int(params.readFromTracksToPredict))
        arr = (C.c_char_p * len(scenes))(*[s.encode() for s in scenes])
        self._h = lib.gd_ref_create(arr, len(scenes), C.byref(p))
        self.num_worlds = len(scenes)
        self._views = {}

    def close(self):
        if self._h:
            _load().gd_ref_destroy(self._h)
            self._h = None
            self._views = {}

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def step(self):
        _load().gd_ref_step(self._h)

    def reset(self, worlds):
        a = (C.c_int32 * len(worlds))(*worlds)
        _load().gd_ref_reset(self._h, a, len(worlds))

    def set_maps(self, scenes):
        arr = (C.c_char_p * len(scenes))(*[s.encode() for s in scenes])
        _load().gd_ref_set_maps(self._h, arr, len(scenes))

    def delete_agents(self, world, ids):
        a = (C.c_int32 * len(ids))(*ids)
        _load().gd_ref_delete_agents(self._h, world, a, len(ids))

    def tensor(self, name) -> np.ndarray:
        """Zero-copy numpy view of the export buffer (stable across steps)."""
        if name in self._views:
            return self._views[name]
        t = _Tensor()
        rc = _load().gd_ref_tensor(self._h, name.encode(), C.byref(t))
        if rc != 0:
            raise KeyError(name)
        shape = tuple(t.shape[i] for i in range(t.ndim))
        n = int(np.prod(shape))
        ctype = C.c_float if t.dtype == 0 else C.c_int32
        buf = (ctype * n).from_address(t.ptr)
        v = np.frombuffer(buf, dtype=np.float32 if t.dtype == 0 else np.int32).reshape(shape)
        self._views[name] = v
        return v
