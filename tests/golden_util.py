"""Access to the committed golden fixtures (tests/golden/, generated from the unmodified reference by
tests/golden/make_golden.py)."""
from __future__ import annotations

import gzip
import json
import os
import shutil

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
ROLLOUT_KEYS = ["delta_lidar", "bicycle_stop", "classic_removed_linear", "expert_womd", "state_womd"]
LOADER_KEYS = ["synth_a", "synth_b", "womd_test_scene", "womd_test_scene_reduced"]


def unpack(names, out_dir):
    """gunzip the named scenarios into out_dir, return their paths."""
    paths = []
    for n in names:
        p = os.path.join(str(out_dir), n + ".json")
        if not os.path.exists(p):
            with gzip.open(os.path.join(GOLDEN, n + ".json.gz"), "rb") as f, open(p, "wb") as g:
                shutil.copyfileobj(f, g)
        paths.append(p)
    return paths


class Rollout:
    def __init__(self, key):
        z = np.load(os.path.join(GOLDEN, f"rollout_{key}.npz"))
        self.z = z
        self.params = json.loads(str(z["params"]))
        self.scenes = json.loads(str(z["scenes"]))
        self.shape = z["shape"]
        self.steps = int(z["steps"])
        self.snaps = [int(x) for x in z["snaps"]]
        self.amax = int(self.shape[:, 0].max())

    def snapshot(self, t) -> dict:
        out = {"shape": self.shape}
        for k in self.z.files:
            if k.endswith(f"@{t}"):
                out[k.split("@")[0]] = self.z[k]
        return out

    def crop(self, snap: dict) -> dict:
        """Cut a full [W,S,...] snapshot to the agent range the fixture stores."""
        return {k: (v if k == "shape" else v[:, :self.amax]) for k, v in snap.items()}


def loader(key) -> dict:
    z = np.load(os.path.join(GOLDEN, f"loader_{key}.npz"))
    d = {k: z[k] for k in z.files}
    d["params"] = json.loads(str(d["params"]))
    return d
