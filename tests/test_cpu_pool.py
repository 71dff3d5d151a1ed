"""Binary scenario pool: a pool entry must build exactly the tensors of the JSON it was
converted from (which tests/test_cpu_loader.py pins to the reference), for every parameter set, and bad
input must be reported, not crash. Host only."""
import os
import time

import numpy as np
import pytest

from golden_util import unpack
from helpers import engine_params, param_dict


def _same(a, b):
    assert (a["num_agents"], a["num_roads"], a["num_controlled"]) == (b["num_agents"], b["num_roads"], b["num_controlled"])
    assert a["mean"] == b["mean"]
    for k in ("map_obs", "traj", "agent_static", "road_pose"):
        assert np.array_equal(a[k].view(np.uint32), b[k].view(np.uint32)), k


def test_pool_entries_build_identical_tensors(tmp_path):
    from gpudrive_lab_b200 import scene_host
    scenes = unpack(["womd_test_scene", "synth_a", "synth_b"], tmp_path)
    pool = str(tmp_path / "scenes.gdpool")
    refs = scene_host.convert_pool(scenes, pool, 0.1, threads=2)
    assert refs == scene_host.pool_scenes(pool) and len(refs) == 3
    for pd in (param_dict(), param_dict(maxNumControlledAgents=4, initOnlyValidAgentsAtFirstStep=False),
               param_dict(IgnoreNonVehicles=True, isStaticAgentControlled=True)):
        for js, ref in zip(scenes, refs):
            _same(scene_host.scene_tensors(js, engine_params(pd), 128), scene_host.scene_tensors(ref, engine_params(pd), 128))
    # a pool can be converted from pool references too, and one thread gives the same file as many
    pool2 = str(tmp_path / "again.gdpool")
    scene_host.convert_pool(refs, pool2, 0.1, threads=1)
    assert open(pool, "rb").read() == open(pool2, "rb").read()


def test_pool_errors_are_reported(tmp_path):
    from gpudrive_lab_b200 import scene_host
    scenes = unpack(["synth_a"], tmp_path)
    pool = str(tmp_path / "one.gdpool")
    scene_host.convert_pool(scenes, pool, 0.1)
    p = engine_params(param_dict())
    with pytest.raises(RuntimeError, match="does not exist"):
        scene_host.scene_info(pool + "#1", p)
    with pytest.raises(RuntimeError, match="polylineReductionThreshold"):
        scene_host.scene_info(pool + "#0", engine_params(param_dict(polylineReductionThreshold=0.5)))
    bad = tmp_path / "bad.gdpool"
    bad.write_bytes(b"not a pool at all")
    with pytest.raises(RuntimeError, match="not a scenario pool"):
        scene_host.scene_info(str(bad) + "#0", p)
    trunc = tmp_path / "trunc.gdpool"
    trunc.write_bytes(open(pool, "rb").read()[:4000])
    with pytest.raises(RuntimeError, match="truncated"):
        scene_host.scene_info(str(trunc) + "#0", p)
    with pytest.raises(RuntimeError):
        scene_host.convert_pool([str(tmp_path / "missing.json")], str(tmp_path / "x.gdpool"), 0.1)


def test_pool_read_is_faster_than_json(tmp_path):
    """The point of the format: reading an entry back costs a fraction of parsing its JSON."""
    from gpudrive_lab_b200 import scene_host
    scenes = unpack(["womd_test_scene"], tmp_path)
    refs = scene_host.convert_pool(scenes, str(tmp_path / "p.gdpool"), 0.1)
    p = engine_params(param_dict())

    def best(path, n=5):
        ts = []
        for _ in range(n):
            t0 = time.perf_counter(); scene_host.scene_info(path, p); ts.append(time.perf_counter() - t0)
        return min(ts)
    tj, tb = best(scenes[0]), best(refs[0])
    print(f"scene_info: json {tj * 1e3:.2f} ms, pool {tb * 1e3:.2f} ms ({os.path.getsize(scenes[0]) / 1e6:.2f} MB json)")
    assert tb < tj
