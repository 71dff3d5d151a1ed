"""Host-only view of the scenario loader (gd_scene_info / gd_scene_tensors): parses a scenario
and builds its static tensors on the CPU, with no GPU involved. Used by the CPU test-suite to
check the loader against the reference and by tools that size a scenario pool."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _capi

AGENT_STATIC_COLS = ["length", "width", "height", "scale_d0", "scale_d1", "scale_d2", "goal_x", "goal_y",
                     "type", "id", "response_type", "controlled", "is_sdc", "is_ooi", "is_ttp", "difficulty"]


def scene_info(path: str, params, max_agents: int = 128) -> dict:
    lib = _capi.load()
    p = params._to_c()
    out = _capi.GdSceneInfo()
    _capi.check(lib.gd_scene_info(str(path).encode(), C.byref(p), max_agents, C.byref(out)))
    return {"num_agents": out.num_agents, "num_roads": out.num_roads,
            "num_controlled": out.num_controlled, "mean": tuple(out.mean)}


def scene_tensors(path: str, params, max_agents: int = 128) -> dict:
    info = scene_info(path, params, max_agents)
    lib = _capi.load()
    p = params._to_c()
    R, A = info["num_roads"], info["num_agents"]
    map_obs = np.zeros((max(R, 1), 9), np.float32)
    traj = np.zeros((max(A, 1), 1456), np.float32)
    stat = np.zeros((max(A, 1), 16), np.float32)
    _capi.check(lib.gd_scene_tensors(str(path).encode(), C.byref(p), max_agents, map_obs.ctypes.data, R,
                                     traj.ctypes.data, stat.ctypes.data, A))
    pose = np.zeros((max(R, 1), 4), np.float32)
    _capi.check(lib.gd_scene_road_pose(str(path).encode(), C.byref(p), max_agents, pose.ctypes.data, R))
    info.update(map_obs=map_obs[:R], traj=traj[:A], agent_static=stat[:A], road_pose=pose[:R])
    return info


def convert_pool(json_paths, out_path: str, polyline_reduction_threshold: float, threads: int = 0) -> list[str]:
    """Parse JSON scenarios once, in parallel, into one binary scenario-pool file (gd_pool_convert).
    Returns the scene references ("<out_path>#<i>") that every scene-path argument accepts afterwards."""
    lib = _capi.load()
    arr = (C.c_char_p * len(json_paths))(*[str(p).encode() for p in json_paths])
    _capi.check(lib.gd_pool_convert(arr, len(json_paths), float(polyline_reduction_threshold), str(out_path).encode(), threads))
    return [f"{out_path}#{i}" for i in range(len(json_paths))]


def pool_scenes(pool_path: str) -> list[str]:
    """Scene references of an existing pool file."""
    lib = _capi.load()
    n, thr = C.c_int32(0), C.c_float(0)
    _capi.check(lib.gd_pool_count(str(pool_path).encode(), C.byref(n), C.byref(thr)))
    return [f"{pool_path}#{i}" for i in range(n.value)]
