"""The C-ABI library loads on a machine without a GPU, exports every symbol include/gpudrive_b200.h
declares, and fails loudly (error code + message, no fallback) when asked to compute without a device."""
import ctypes as C
import os
import re

import pytest

from helpers import ROOT, engine_params, param_dict


def _declared():
    hdr = open(os.path.join(ROOT, "include", "gpudrive_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(gd_[a-z_0-9]+)\s*\(", hdr)))


def test_header_symbols_exported():
    from gpudrive_lab_b200 import _capi
    lib = _capi.load()
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/gpudrive_b200.h but not exported"
    assert sorted(_capi.EXPORTED_SYMBOLS) == names


def test_struct_layouts_match_header():
    from gpudrive_lab_b200 import _capi
    assert C.sizeof(_capi.GdParams) == 15 * 4
    assert C.sizeof(_capi.GdConfig) == 8 + 15 * 4 + 4
    assert C.sizeof(_capi.GdTensorDesc) == 8 + 4 + 4 + 8 * 8 + 8      # device + padding
    assert C.sizeof(_capi.GdStats) == 10 * 8
    assert len(_capi.TENSOR_IDS) == 23


def test_no_cpu_fallback(tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("needs a machine without a GPU")
    from gpudrive_lab_b200 import madrona_gpudrive as mg
    from golden_util import unpack
    scenes = unpack(["synth_a"], tmp_path)
    with pytest.raises(RuntimeError, match="(?i)cuda|device|driver"):
        mg.SimManager(mg.madrona.ExecMode.CUDA, 0, scenes, engine_params(param_dict()))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        mg.SimManager(mg.madrona.ExecMode.CPU, 0, scenes, engine_params(param_dict()))


def test_renderer_items_unavailable(tmp_path):
    from gpudrive_lab_b200 import madrona_gpudrive as mg
    from golden_util import unpack
    with pytest.raises(RuntimeError, match="unavailable on this box"):
        mg.SimManager(mg.madrona.ExecMode.CUDA, 0, unpack(["synth_a"], tmp_path), engine_params(param_dict()),
                      enable_batch_renderer=True)


def test_module_surface_matches_reference_bindings():
    """Names the reference's nanobind module exports (src/bindings.cpp:16-150)."""
    from gpudrive_lab_b200 import madrona_gpudrive as mg
    for n in ("kMaxAgentCount", "kMaxRoadEntityCount", "kMaxAgentMapObservationsCount", "episodeLen",
              "numLidarSamples", "vehicleScale", "RewardType", "CollisionBehaviour", "DynamicsModel",
              "FindRoadObservationsWith", "EntityType", "RewardParams", "Parameters", "SimManager", "madrona"):
        assert hasattr(mg, n), n
    for n in ("step", "reset", "set_maps", "deleteAgents", "action_tensor", "reward_tensor", "done_tensor",
              "self_observation_tensor", "map_observation_tensor", "partner_observations_tensor", "lidar_tensor",
              "steps_remaining_tensor", "shape_tensor", "controlled_state_tensor", "agent_roadmap_tensor",
              "absolute_self_observation_tensor", "bev_observation_tensor", "valid_state_tensor", "info_tensor",
              "response_type_tensor", "expert_trajectory_tensor", "world_means_tensor", "metadata_tensor",
              "map_name_tensor", "scenario_id_tensor", "deleted_agents_tensor", "rgb_tensor", "depth_tensor"):
        assert callable(getattr(mg.SimManager, n)), n
    assert mg.kMaxAgentCount == 128 and mg.episodeLen == 91 and mg.kMaxAgentMapObservationsCount == 200
    assert int(mg.EntityType.Vehicle) == 7 and int(mg.DynamicsModel.DeltaLocal) == 2


def test_install_registers_module():
    import sys
    import gpudrive_lab_b200
    gpudrive_lab_b200.install()
    import madrona_gpudrive
    assert madrona_gpudrive.SimManager is gpudrive_lab_b200.madrona_gpudrive.SimManager
    sys.modules.pop("madrona_gpudrive", None)


def test_headless_cli_builds_and_refuses_cpu():
    """tools/gd_headless.cpp (counterpart of src/headless.cpp) links against the C ABI alone."""
    import subprocess
    from gpudrive_lab_b200 import build
    exe = build.build_headless()
    r = subprocess.run([exe, "CPU", "3"], capture_output=True, text=True)
    assert r.returncode != 0 and "no CPU backend" in r.stderr
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode != 0 and "NUM_STEPS" in r.stderr
