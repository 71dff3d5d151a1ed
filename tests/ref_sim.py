"""A `madrona_gpudrive.SimManager` look-alike over the UNMODIFIED reference Manager (oracle/_ref, ExecMode::CPU):
test infrastructure that lets the reference's own Python env run on the reference's own simulator on a box
where the reference's nanobind module cannot be built, so that the same env rollout on the B200 engine can be
compared with it (tests/test_gpu_env.py)."""
from __future__ import annotations

import torch

from oracle import refmgr


class _HostTensor:
    def __init__(self, a):
        self._a = a

    def to_torch(self):
        return torch.from_numpy(self._a)          # zero-copy alias of the reference's export buffer


class RefSimManager:
    def __init__(self, exec_mode, gpu_id, scenes, params, enable_batch_renderer=False,
                 batch_render_view_width=64, batch_render_view_height=64):
        pc = params._to_c()
        d = {f[0]: getattr(pc, f[0]) for f in pc._fields_}
        d["IgnoreNonVehicles"] = d.pop("ignoreNonVehicles")
        for k in ("IgnoreNonVehicles", "initOnlyValidAgentsAtFirstStep", "isStaticAgentControlled", "enableLidar",
                  "disableClassicalObs", "readFromTracksToPredict"):
            d[k] = bool(d[k])
        self._m = refmgr.RefManager(list(scenes), refmgr.RefParams(**d))

    def step(self):
        self._m.step()

    def reset(self, worlds):
        self._m.reset(list(worlds))

    def set_maps(self, scenes):
        self._m.set_maps(list(scenes))

    def deleteAgents(self, agents_to_delete):
        for w, ids in agents_to_delete.items():
            self._m.delete_agents(int(w), [int(i) for i in ids])

    def __getattr__(self, name):
        if name.endswith("_tensor"):
            return lambda: _HostTensor(self._m.tensor(name[:-7]))
        raise AttributeError(name)

    def close(self):
        self._m.close()
