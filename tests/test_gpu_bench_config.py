"""-m gpu: parity ON the benchmarked configurations, not on toy world counts.

Each test builds the engine exactly as bench.py does (gpudrive_lab_b200.workload: same scenes, same
Parameters, S = 128, worlds cycled over the scenario pool), steps it with host-generated seeded actions for
every world, and compares a spread of worlds - first, last, the pool wrap-around (63 / 64 / 65), and enough
others to cover >= 16 distinct scenes - against the unmodified reference Manager (oracle/_ref) built on
exactly those worlds' scenes and action slices, with the normal compare_step rules. A wrong world -> scene
indirection, a 32-bit offset overflow or a launch trimmed to the wrong slot count at w >> 4 fails here.
"""
import numpy as np
import pytest

from helpers import compare_step, ref_params, report, snapshot_ref, TENSORS_CMP

pytestmark = pytest.mark.gpu


def _sub_snapshot(sim, worlds, names):
    import torch
    idx = torch.as_tensor(worlds, device="cuda")
    getters = {"agent_roadmap": sim.agent_roadmap_tensor, "partner_observations": sim.partner_observations_tensor,
               "self_observation": sim.self_observation_tensor,
               "absolute_self_observation": sim.absolute_self_observation_tensor,
               "controlled_state": sim.controlled_state_tensor, "response_type": sim.response_type_tensor,
               "steps_remaining": sim.steps_remaining_tensor, "shape": sim.shape_tensor, "done": sim.done_tensor,
               "info": sim.info_tensor, "reward": sim.reward_tensor, "lidar": sim.lidar_tensor}
    return {n: getters[n]().to_torch().index_select(0, idx).cpu().numpy() for n in names}


def _run(config, total_worlds, check_worlds, steps, names=TENSORS_CMP):
    import torch
    assert torch.cuda.is_available(), "CUDA device required (no CPU fallback exists)"
    from gpudrive_lab_b200 import madrona_gpudrive as mg
    from gpudrive_lab_b200 import workload as wl
    from oracle import refmgr
    assert refmgr.available(), "oracle/_ref/libgd_ref.so missing: run __graft_entry__.build() where /root/reference exists"
    c = wl.CONFIGS[config]
    pd = wl.params_dict(config)
    unique = wl.make_scenes(config)
    scenes = wl.world_scenes(unique, total_worlds)
    sim = mg.SimManager(mg.madrona.ExecMode.CUDA, 0, scenes, wl.engine_params(pd), max_agents=wl.SLOTS)
    sim.set_option(0, 0)        # as the benchmark runs it
    ref = refmgr.RefManager([scenes[w] for w in check_worlds], ref_params(pd))
    assert len({scenes[w] for w in check_worlds}) >= min(16, len(unique))
    act_e = sim.action_tensor().to_torch()
    act_r = ref.tensor("action")
    model, lidar, radius = pd["dynamicsModel"], bool(pd["enableLidar"]), pd["observationRadius"]
    stats = {}
    compare_step(snapshot_ref(ref, names), _sub_snapshot(sim, check_worlds, names), radius, lidar,
                 tag=f"{c['name']} after construction", stats=stats)
    for t in range(steps):
        a = wl.random_actions(t, total_worlds, wl.SLOTS, model, seed=77)
        act_r[...] = a[check_worlds]
        act_e.copy_(torch.from_numpy(a))          # default stream; the engine's blocking stream orders the step after it
        ref.step()
        sim.step()
        compare_step(snapshot_ref(ref, names), _sub_snapshot(sim, check_worlds, names), radius, lidar,
                     tag=f"{c['name']} step {t + 1}", stats=stats)
    report(f"bench_{c['name'].replace(' ', '')}_{total_worlds}worlds_{len(check_worlds)}checked", stats)
    return sim, ref, stats


# scene of world w = w % 64: these 19 worlds cover 17 distinct scenes, both ends and the pool wrap-around
SPREAD_1024 = [0, 63, 64, 65, 130, 197, 264, 331, 398, 465, 532, 599, 666, 733, 800, 867, 934, 1001, 1023]


def test_config3_bench_engine_matches_reference():
    """config 3 (the headline workload): DeltaLocal, full observations + 3 x 50 lidar, 1 024 worlds over the
    64-scene pool - one rank's share of the 8-GPU run - 12 steps."""
    _run(3, 1024, SPREAD_1024, 12)


def test_config2_bench_engine_matches_reference():
    """config 2: InvertibleBicycle, no lidar, all 1 024 worlds of the configuration."""
    _run(2, 1024, SPREAD_1024, 12)


def test_config5_bench_engine_matches_reference():
    """config 5 (collision stress): 128 live agents per world, ~4 000 road-edge segments, AgentStop; 256 worlds
    over the 8-scene pool, every scene checked. Collision / done / info flags bit-exact."""
    names = ["shape", "done", "info", "steps_remaining", "controlled_state", "response_type", "self_observation",
             "absolute_self_observation", "reward", "partner_observations", "agent_roadmap"]
    sim, ref, _ = _run(5, 256, [0, 7, 8, 9, 74, 131, 196, 253, 254, 255], 16, names=names)
    shape = ref.tensor("shape")
    assert (shape[:, 0] == 128).all() and (shape[:, 1] >= 4000).all()
    info = ref.tensor("info")
    assert info[..., :3].sum() > 0, "the stress scenes are expected to produce collisions within 16 steps"
    st = sim.stats()
    assert st["sat_tests"] > 0 and st["collided_road"] + st["collided_vehicle"] > 0
