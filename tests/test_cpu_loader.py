"""Host scenario loader (JSON -> entity tables) against the reference:
bit-exact with the golden tensors generated from the unmodified reference Manager, and, where
/root/reference and oracle/_ref are present, live on the reference's real example scenes."""
import glob
import os

import numpy as np
import pytest

from golden_util import LOADER_KEYS, loader, unpack
from helpers import REF_EXAMPLES, engine_params, param_dict, ref_params


def _check(scene_path, pd, gold):
    from gpudrive_lab_b200 import scene_host
    t = scene_host.scene_tensors(scene_path, engine_params(pd), 128)
    A, R = int(gold["shape"][0]), int(gold["shape"][1])
    assert (t["num_agents"], t["num_roads"]) == (A, R)
    assert np.array_equal(t["map_obs"].view(np.uint32), gold["map_observation"][:R].view(np.uint32)), "map_observation"
    assert np.array_equal(t["traj"].view(np.uint32), gold["expert_trajectory"][:A].view(np.uint32)), "expert_trajectory"
    st = t["agent_static"]
    assert np.array_equal(st[:, 12:16].astype(np.int32), gold["metadata"][:A]), "metadata"
    assert np.array_equal(st[:, 11].astype(np.int32), gold["controlled_state"][:A, 0]), "controlled"
    assert np.array_equal(st[:, 10].astype(np.int32), gold["response_type"][:A, 0]), "response type"
    assert np.array_equal(np.asarray(t["mean"], np.float32).view(np.uint32),
                          gold["world_means"].astype(np.float32).view(np.uint32)), "world mean"
    assert t["num_controlled"] == int(gold["controlled_state"][:A].sum())


@pytest.mark.parametrize("key", LOADER_KEYS)
def test_loader_matches_golden(key, tmp_path):
    g = loader(key)
    scene = key.replace("_reduced", "")
    _check(unpack([scene], tmp_path)[0], g["params"], g)


def test_reference_observation_test_restated(tmp_path):
    """tests/observationTest.cpp:16-138 restated over created entities: every road segment's row is the
    mean-centred midpoint of its two polyline points, with the type of its road (threshold 0)."""
    import gzip, json
    from gpudrive_lab_b200 import scene_host
    path = unpack(["womd_test_scene"], tmp_path)[0]
    raw = json.load(open(path))
    pd = param_dict(polylineReductionThreshold=0.0, initOnlyValidAgentsAtFirstStep=False)
    t = scene_host.scene_tensors(path, engine_params(pd), 128)
    # calc_mean: running mean over objects' first positions then road points (json_serialization.hpp:307-336)
    mx, my = np.float32(t["mean"][0]), np.float32(t["mean"][1])
    types = {"road_edge": 1, "road_line": 2, "lane": 3, "crosswalk": 4, "speed_bump": 5, "stop_sign": 6}
    rows = t["map_obs"]
    i = 0
    for road in raw["roads"]:
        ty = types.get(road["type"])
        if ty is None:
            continue                          # "driveway" etc. create no entity
        g = road["geometry"]
        if ty <= 3:
            for a, b in zip(g[:-1], g[1:]):
                ex = (a["x"] + b["x"]) / 2 - mx
                ey = (a["y"] + b["y"]) / 2 - my
                assert abs(rows[i, 0] - ex) < 1e-3 and abs(rows[i, 1] - ey) < 1e-3 and rows[i, 6] == ty, (i, rows[i])
                i += 1
        elif ty in (4, 5):
            cx = sum(p["x"] for p in g[:4]) / 4 - mx
            cy = sum(p["y"] for p in g[:4]) / 4 - my
            assert abs(rows[i, 0] - cx) < 1e-3 and abs(rows[i, 1] - cy) < 1e-3 and rows[i, 6] == ty
            i += 1
        else:
            assert abs(rows[i, 0] - (g[0]["x"] - mx)) < 1e-3 and abs(rows[i, 1] - (g[0]["y"] - my)) < 1e-3
            assert rows[i, 6] == ty
            i += 1
    assert i == t["num_roads"] == 9899


@pytest.mark.skipif(not os.path.isdir(REF_EXAMPLES), reason="reference checkout not present")
@pytest.mark.parametrize("kw", [dict(), dict(polylineReductionThreshold=0.0, dynamicsModel=2),
                                dict(IgnoreNonVehicles=True, isStaticAgentControlled=False, maxNumControlledAgents=3,
                                     dynamicsModel=0)])
def test_loader_matches_reference_live(kw):
    from oracle import refmgr
    if not refmgr.available():
        pytest.skip("oracle/_ref not built")
    pd = param_dict(**kw)
    for path in sorted(glob.glob(os.path.join(REF_EXAMPLES, "*.json"))):
        ref = refmgr.RefManager([path], ref_params(pd))
        gold = {n: ref.tensor(n)[0].copy() for n in ("shape", "map_observation", "expert_trajectory", "metadata",
                                                      "controlled_state", "response_type", "world_means")}
        ref.close()
        _check(path, pd, gold)
