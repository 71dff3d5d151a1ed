"""ctypes binding of libgpudrive_b200.so (include/gpudrive_b200.h). No torch import here."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GD_B200_LIB") or os.path.join(_HERE, "libgpudrive_b200.so")   # override: kernel experiments


class GdParams(C.Structure):
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


class GdConfig(C.Structure):
    _fields_ = [("gpu_id", C.c_int32), ("max_agents", C.c_int32), ("params", GdParams), ("flags", C.c_int32)]


class GdTensorDesc(C.Structure):
    _fields_ = [("ptr", C.c_void_p), ("dtype", C.c_int32), ("ndim", C.c_int32),
                ("shape", C.c_int64 * 8), ("device", C.c_int32)]


class GdRewardShaping(C.Structure):
    _fields_ = [("mode", C.c_int32), ("collision_weight", C.c_float), ("goal_achieved_weight", C.c_float),
                ("off_road_weight", C.c_float), ("log_distance_weight", C.c_float),
                ("agent_weights", C.c_void_p), ("world_time_steps", C.c_void_p)]


class GdStats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in (
        "steps", "agent_steps", "controlled_steps", "episodes_reset", "collided_road",
        "collided_vehicle", "collided_nonvehicle", "goals_reached", "dones", "sat_tests")]


class GdSceneInfo(C.Structure):
    _fields_ = [("num_agents", C.c_int32), ("num_roads", C.c_int32), ("num_controlled", C.c_int32),
                ("mean", C.c_float * 3)]


# every symbol include/gpudrive_b200.h declares
EXPORTED_SYMBOLS = [
    "gd_create", "gd_destroy", "gd_step", "gd_step_async", "gd_sync", "gd_reset", "gd_set_maps",
    "gd_delete_agents", "gd_tensor", "gd_step_host", "gd_pool_add", "gd_bind_worlds", "gd_stats",
    "gd_num_worlds", "gd_max_agents", "gd_launch_count", "gd_stream", "gd_last_error",
    "gd_scene_info", "gd_scene_tensors", "gd_scene_road_pose", "gd_policy_obs", "gd_set_option", "gd_kernel_times",
    "gd_pool_convert", "gd_pool_count", "gd_shaped_reward", "gd_expert_actions", "gd_auto_reset_pool", "gd_auto_reset",
    "gd_scatter_actions_noalloc",
]

TENSOR_IDS = {
    "action": 0, "reward": 1, "done": 2, "self_observation": 3, "map_observation": 4,
    "partner_observations": 5, "agent_roadmap": 6, "lidar": 7, "bev_observation": 8,
    "steps_remaining": 9, "shape": 10, "controlled_state": 11, "absolute_self_observation": 12,
    "valid_state": 13, "info": 14, "response_type": 15, "expert_trajectory": 16,
    "world_means": 17, "metadata": 18, "deleted_agents": 19, "map_name": 20, "scenario_id": 21,
    "knn_index": 22,
}

_lib = None


def load():
    """Load the native library; raises (loudly) if it has not been built. There is no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`. "
            "gpudrive_lab_b200 has no CPU or PyTorch fallback.")
    lib = C.CDLL(LIB_PATH)
    H = C.c_void_p
    lib.gd_create.argtypes = [C.POINTER(GdConfig), C.POINTER(C.c_char_p), C.c_int32, C.POINTER(H)]
    lib.gd_destroy.argtypes = [H]
    lib.gd_step.argtypes = [H]
    lib.gd_step_async.argtypes = [H]
    lib.gd_sync.argtypes = [H]
    lib.gd_reset.argtypes = [H, C.POINTER(C.c_int32), C.c_int32]
    lib.gd_set_maps.argtypes = [H, C.POINTER(C.c_char_p), C.c_int32]
    lib.gd_delete_agents.argtypes = [H, C.c_int32, C.POINTER(C.c_int32), C.c_int32]
    lib.gd_tensor.argtypes = [H, C.c_int32, C.POINTER(GdTensorDesc)]
    lib.gd_step_host.argtypes = [H, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.gd_pool_add.argtypes = [H, C.c_char_p, C.POINTER(C.c_int32)]
    lib.gd_bind_worlds.argtypes = [H, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.c_int32]
    lib.gd_auto_reset_pool.argtypes = [H, C.POINTER(C.c_int32), C.c_int32]
    lib.gd_auto_reset.argtypes = [H]
    lib.gd_scatter_actions_noalloc.argtypes = [C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_uint64]
    lib.gd_shaped_reward.argtypes = [H, C.POINTER(GdRewardShaping), C.POINTER(GdTensorDesc)]
    lib.gd_expert_actions.argtypes = [H, C.POINTER(GdTensorDesc)]
    lib.gd_pool_convert.argtypes = [C.POINTER(C.c_char_p), C.c_int32, C.c_float, C.c_char_p, C.c_int32]
    lib.gd_pool_count.argtypes = [C.c_char_p, C.POINTER(C.c_int32), C.POINTER(C.c_float)]
    lib.gd_stats.argtypes = [H, C.POINTER(GdStats)]
    lib.gd_set_option.argtypes = [H, C.c_int32, C.c_int32]
    lib.gd_kernel_times.argtypes = [H, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_int64)]
    lib.gd_num_worlds.argtypes = [H]
    lib.gd_max_agents.argtypes = [H]
    lib.gd_launch_count.argtypes = [H]
    lib.gd_launch_count.restype = C.c_uint64
    lib.gd_stream.argtypes = [H]
    lib.gd_stream.restype = C.c_uint64
    lib.gd_last_error.restype = C.c_char_p
    lib.gd_scene_info.argtypes = [C.c_char_p, C.POINTER(GdParams), C.c_int32, C.POINTER(GdSceneInfo)]
    lib.gd_scene_tensors.argtypes = [C.c_char_p, C.POINTER(GdParams), C.c_int32, C.c_void_p, C.c_int32,
                                     C.c_void_p, C.c_void_p, C.c_int32]
    lib.gd_policy_obs.argtypes = [H, C.c_int32, C.POINTER(GdTensorDesc)]
    lib.gd_scene_road_pose.argtypes = [C.c_char_p, C.POINTER(GdParams), C.c_int32, C.c_void_p, C.c_int32]
    _lib = lib
    return lib


def check(rc: int):
    if rc != 0:
        raise RuntimeError(load().gd_last_error().decode(errors="replace"))
