"""TEST INFRASTRUCTURE ONLY: ctypes front end of oracle/step_oracle.cpp (the CPU restatement of the
reference's step). Imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg; never by
the product package.

`OracleManager(scenes, params_dict)` mirrors the slice of Manager the parity tests drive: step(), reset(),
tensor(name) -> numpy [W, S, ...] in the reference's layouts. Entity tables come from the host scenario
loader (gpudrive_lab_b200.scene_host), which is pinned against the reference separately."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_build", "libgd_oracle.so")

_F32 = {"action": (0, 10), "reward": (1, 1), "self_observation": (2, 8), "partner_observations": (3, None),
        "agent_roadmap": (4, 200 * 9), "lidar": (5, 3 * 50 * 4), "absolute_self_observation": (6, 14), "bev_observation": (7, 200 * 200)}
_I32 = {"done": (16, 1), "steps_remaining": (17, 1), "info": (18, 5), "controlled_state": (19, 1),
        "response_type": (20, 1), "knn_index": (21, 200)}
_SHAPES = {"partner_observations": lambda S: (S, S - 1, 9), "agent_roadmap": lambda S: (S, 200, 9),
           "lidar": lambda S: (S, 3, 50, 4), "knn_index": lambda S: (S, 200),
           "bev_observation": lambda S: (S, 200, 200, 1)}

_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(HERE, "step_oracle.cpp")
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", HERE])
    return LIB


def load():
    global _lib
    if _lib is None:
        build()
        lib = C.CDLL(LIB)
        lib.gdo_create.restype = C.c_void_p
        lib.gdo_create.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        lib.gdo_destroy.argtypes = [C.c_void_p]
        lib.gdo_step.argtypes = [C.c_void_p]
        lib.gdo_reset.argtypes = [C.c_void_p, C.c_int]
        lib.gdo_enable_bev.argtypes = [C.c_void_p]
        lib.gdo_tensor.restype = C.c_void_p
        lib.gdo_tensor.argtypes = [C.c_void_p, C.c_int]
        lib.gdo_obb_collided.argtypes = [C.c_void_p, C.c_void_p]
        lib.gdo_observation_of.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        lib.gdo_forward_model.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib = lib
    return _lib


def roads_table(scene: dict) -> np.ndarray:
    """[R,12] = x,y,z, qw,qz, d0,d1,d2, yaw, type, id, mapType from the host loader's tensors."""
    m, p = scene["map_obs"], scene["road_pose"]
    out = np.zeros((len(m), 12), np.float32)
    out[:, 0:2] = m[:, 0:2]; out[:, 2] = p[:, 0]; out[:, 3] = p[:, 1]; out[:, 4] = p[:, 2]
    out[:, 5:8] = m[:, 2:5]; out[:, 8] = m[:, 5]; out[:, 9] = m[:, 6]; out[:, 10] = m[:, 7]; out[:, 11] = m[:, 8]
    return out


class OracleWorld:
    def __init__(self, scene: dict, cparams, S: int):
        lib = load()
        self.S, self.A, self.R = S, int(scene["num_agents"]), int(scene["num_roads"])
        stat = np.ascontiguousarray(scene["agent_static"], np.float32)
        traj = np.ascontiguousarray(scene["traj"], np.float32)
        roads = np.ascontiguousarray(roads_table(scene))
        self._h = lib.gdo_create(C.addressof(cparams), S, self.A, self.R, stat.ctypes.data, traj.ctypes.data,
                                 roads.ctypes.data)
        self._lib = lib

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            self._lib.gdo_destroy(h)

    def view(self, name) -> np.ndarray:
        S = self.S
        if name in _F32:
            tid, _ = _F32[name]; ct, dt = C.c_float, np.float32
        else:
            tid, _ = _I32[name]; ct, dt = C.c_int32, np.int32
        shape = _SHAPES[name](S) if name in _SHAPES else (S, (_F32.get(name) or _I32.get(name))[1])
        n = int(np.prod(shape))
        ptr = self._lib.gdo_tensor(self._h, tid)
        return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(ct)), shape=(n,)).view(dt).reshape(shape)


class OracleManager:
    """W independent OracleWorlds behind the Manager-shaped surface the tests use."""

    def __init__(self, scenes, params, S: int = 128):
        # params: gpudrive_lab_b200.madrona_gpudrive.Parameters (host loader + oracle read the same struct)
        from gpudrive_lab_b200 import scene_host
        self.S = S
        self._cp = params._to_c()
        cache = {}
        self.worlds = []
        for s in scenes:
            if s not in cache:
                cache[s] = scene_host.scene_tensors(s, params, S)
            self.worlds.append(OracleWorld(cache[s], self._cp, S))
        self.num_worlds = len(self.worlds)

    def enable_bev(self):
        for w in self.worlds:
            w._lib.gdo_enable_bev(w._h)

    def step(self):
        for w in self.worlds:
            self._lib_step(w)

    @staticmethod
    def _lib_step(w):
        w._lib.gdo_step(w._h)

    def reset(self, worlds):
        sel = set(int(x) for x in worlds)
        for i, w in enumerate(self.worlds):
            w._lib.gdo_reset(w._h, 1 if i in sel else 0)

    def set_actions(self, actions: np.ndarray):
        for i, w in enumerate(self.worlds):
            w.view("action")[...] = actions[i]

    def tensor(self, name) -> np.ndarray:
        if name == "shape":
            return np.array([[w.A, w.R] for w in self.worlds], np.int32)
        return np.stack([w.view(name) for w in self.worlds])


def obb_collided(a, b) -> bool:
    a = np.asarray(a, np.float32); b = np.asarray(b, np.float32)
    return bool(load().gdo_obb_collided(a.ctypes.data, b.ctypes.data))


def observation_of(ego, ent):
    ego = np.asarray(ego, np.float32); ent = np.asarray(ent, np.float32)
    out = np.zeros(3, np.float32)
    load().gdo_observation_of(ego.ctypes.data, ent.ctypes.data, out.ctypes.data)
    return out


def forward_model(model: int, state, action):
    state = np.asarray(state, np.float32)
    act = np.zeros(10, np.float32); act[:len(action)] = action
    out = np.zeros(8, np.float32)
    load().gdo_forward_model(int(model), state.ctypes.data, act.ctypes.data, out.ctypes.data)
    return out
