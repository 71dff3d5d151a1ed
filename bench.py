#!/usr/bin/env python
"""Benchmark of the GPUDrive step path (BASELINE.json: controlled agent-steps/sec, 8192 worlds x
64 agents, DeltaLocal dynamics, full observations + 3x50-ray lidar, 1/2/4/8 B200).

    python bench.py --gpus N --steps K --warmup W          # this engine (CUDA)
    python bench.py --impl reference ...                    # the reference's own CPU Manager

One "step" = one SimManager.step() over every world of the rank (actions already in HBM for
`value`; through gd_step_host with pinned host buffers for `e2e`). Worlds are sharded across
ranks (8192 / N each, strong scaling, no collective inside the step); the only collective is the
end-of-run reduction of the timing / statistics vector.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

TOTAL_WORLDS = 8192
# dram__bytes_read.sum + dram__bytes_write.sum of one k_obs launch of the default workload, from the committed
# ncu capture profiles/r01_v13_k_obs_raw.txt (88.9 MB read + 7.421 GB written)
K_OBS_DRAM_BYTES_PER_LAUNCH = 7.509e9
AGENTS = 64
SLOTS = 128               # consts::kMaxAgentCount of the reference -> identical tensor shapes
UNIQUE_SCENES = 64        # distinct synthetic scenarios, cycled over the worlds
K_ROAD = 200
LIDAR = 50


def params_dict():
    # config 3 of BASELINE.json; collision behaviour / road algorithm as SURVEY.md 8(d) config 2-3
    # GD_BENCH_NO_LIDAR / GD_BENCH_NO_CLASSIC: kernel experiments only (time split of k_obs); the bench line
    # is only valid without them
    return dict(polylineReductionThreshold=0.1, observationRadius=50.0, rewardType=1,
                distanceToGoalThreshold=2.0, distanceToExpertThreshold=0.0, collisionBehaviour=2,
                maxNumControlledAgents=AGENTS, IgnoreNonVehicles=False, roadObservationAlgorithm=0,
                initOnlyValidAgentsAtFirstStep=True, isStaticAgentControlled=True,
                enableLidar=not os.environ.get("GD_BENCH_NO_LIDAR"),
                disableClassicalObs=bool(os.environ.get("GD_BENCH_NO_CLASSIC")), dynamicsModel=2, readFromTracksToPredict=False)


def algorithmic_bytes(S, R, A, lidar=True):
    """compulsory HBM bytes per agent-step, whole step and the k_obs share."""
    partner = 36 * (S - 1)
    roadmap = 36 * K_ROAD
    lid = 48 * LIDAR if lidar else 0
    tables = (36 * R + 64 * A) / A
    step = partner + roadmap + lid + 88 + 32 + 104 + 40 + tables
    obs = partner + roadmap + lid + 88 + 36 + 4 + tables
    return step, obs


def scene_dir():
    d = os.environ.get("GD_BENCH_SCENES", os.path.join(tempfile.gettempdir(), "gd_bench_scenes"))
    os.makedirs(d, exist_ok=True)
    return d


def make_scenes(count):
    from gpudrive_lab_b200 import synth
    return synth.write_scenarios(scene_dir(), count, num_agents=AGENTS)


class ClockSampler:
    """nvidia-smi sampled DURING the timed region (B200_PROFILING.md clocks line)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


# --------------------------------------------------------------------------------------------
# reference arm / CPU baseline: the UNMODIFIED reference Manager, ExecMode::CPU (oracle/_ref)
# --------------------------------------------------------------------------------------------

def run_reference_sample(steps, warmup, worlds=None, budget_s=20.0):
    """Times Manager::step() of the reference on a bounded sample of the same workload.
    Returns (controlled agent-steps/s, ms per step, description dict)."""
    import numpy as np
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from helpers import random_actions, ref_params
    from oracle import refmgr
    if not refmgr.available():
        raise RuntimeError("oracle/_ref/libgd_ref.so is missing (built by __graft_entry__.build() where "
                           "/root/reference exists)")
    cores = len(os.sched_getaffinity(0))
    W = worlds if worlds is not None else max(2 * cores, 8)
    scenes_u = make_scenes(min(UNIQUE_SCENES, W))
    scenes = [scenes_u[i % len(scenes_u)] for i in range(W)]
    pd = params_dict()
    t0 = time.time()
    mgr = refmgr.RefManager(scenes, ref_params(pd))
    build_s = time.time() - t0
    ctrl = int(mgr.tensor("controlled_state")[:, :, 0].sum())
    act = mgr.tensor("action")
    for i in range(warmup):
        act[...] = random_actions(i, W, SLOTS, pd["dynamicsModel"])
        mgr.step()
    # bound the sample: keep the whole call within `budget_s` of CPU work
    t0 = time.time()
    act[...] = random_actions(warmup, W, SLOTS, pd["dynamicsModel"])
    mgr.step()
    one = time.time() - t0
    steps = max(2, min(steps, int(budget_s / max(one, 1e-3))))
    acts = [random_actions(warmup + 1 + i, W, SLOTS, pd["dynamicsModel"]) for i in range(steps)]
    t0 = time.perf_counter()
    for i in range(steps):
        act[...] = acts[i]
        mgr.step()
    dt = time.perf_counter() - t0
    value = ctrl * steps / dt
    desc = {"kind": "reference", "cores": cores, "value": value, "unit": "controlled agent-steps/s",
            "sample": f"{W} worlds x {AGENTS} agents x {steps} steps of the same workload through the reference's "
                      f"Manager::step() (ExecMode::CPU, one pinned worker per core; includes its per-step export "
                      f"copies and the BEV pass it cannot disable); construction {build_s:.1f}s not timed"}
    mgr.close()
    return value, dt / steps * 1e3, steps, desc


def main_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    value, ms, steps, desc = run_reference_sample(args.steps, max(args.warmup, 1), budget_s=60.0)
    line = {
        "impl": "reference", "metric": "controlled agent-steps/sec", "value": value, "unit": "agent-steps/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": max(args.warmup, 1), "ms_per_step": ms,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.gpus),
        "cpu_baseline": desc,
        "e2e": {"value": value, "unit": "agent-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


def workload_config(n_gpus):
    return {"workload": "config 3: 8192 worlds x 64 agents (synthetic Waymo-shaped scenes, ~2050 road entities), "
                        "DeltaLocal dynamics, self+partner+roadmap(K=200)+lidar(3x50) observations, collisions Ignore",
            "total_worlds": TOTAL_WORLDS, "worlds_per_gpu": TOTAL_WORLDS // n_gpus, "agents_per_world": AGENTS,
            "agent_slots": SLOTS, "unique_scenes": UNIQUE_SCENES, "parallelism": f"worlds sharded over {n_gpus} GPU(s)",
            "l2": "per-step export traffic (>= 1 GB per GPU) exceeds the 126 MB L2; no explicit flush needed"}


# --------------------------------------------------------------------------------------------
# engine arm
# --------------------------------------------------------------------------------------------

def main_engine(args):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from helpers import engine_params
    from gpudrive_lab_b200 import madrona_gpudrive as mg

    rank = int(os.environ.get("RANK", "0"))
    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world_size != args.gpus:
        if world_size == 1 and args.gpus > 1:
            raise SystemExit("--gpus N > 1 must be launched with torch.distributed.run (one rank per GPU)")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the engine has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world_size > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    from gpudrive_lab_b200 import shard
    total_worlds = args.worlds or TOTAL_WORLDS
    W = len(shard.world_block(total_worlds, rank, world_size))
    pd = params_dict()
    if rank == 0:
        scenes_u = make_scenes(UNIQUE_SCENES)
    if world_size > 1:
        dist.barrier()
    scenes_u = make_scenes(UNIQUE_SCENES)
    # contiguous block of worlds per rank; world i uses scenario (global index) % UNIQUE_SCENES
    scenes = shard.scenes_for_rank(scenes_u, total_worlds, rank, world_size)
    t0 = time.time()
    sim = mg.SimManager(mg.madrona.ExecMode.CUDA, local_rank, scenes, engine_params(pd), max_agents=SLOTS)
    build_s = time.time() - t0
    sim.set_option(0, 0)        # knn index side buffer off (debug aid, not part of the reference's tensors)
    stream = torch.cuda.ExternalStream(sim.cuda_stream(), device=local_rank)
    shape = sim.shape_tensor().to_torch()
    ctrl_t = sim.controlled_state_tensor().to_torch()
    n_ctrl = int(ctrl_t.sum().item())
    n_live = int(shape[:, 0].sum().item())
    R_mean = float(shape[:, 1].float().mean().item())
    action = sim.action_tensor().to_torch()

    # action ring, resident in HBM before the timed region (DeltaLocal: dx, dy, dyaw)
    g = torch.Generator(device="cuda"); g.manual_seed(1234 + rank)
    ring = []
    for _ in range(8):
        a = torch.empty(W, SLOTS, 3, device="cuda")
        a[..., 0].uniform_(-0.5, 1.5, generator=g); a[..., 1].uniform_(-0.2, 0.2, generator=g)
        a[..., 2].uniform_(-0.05, 0.05, generator=g)
        ring.append(a)
    torch.cuda.synchronize()

    def barrier():
        if world_size > 1:
            dist.barrier()
        torch.cuda.synchronize()

    all_worlds = list(range(W))
    step_no = [0]

    def one_step(k):
        if step_no[0] == 91:       # episode boundary: reset inside the timed region (time, not a step)
            sim.reset(all_worlds)
            step_no[0] = 0
        with torch.cuda.stream(stream):
            action[:, :, :3].copy_(ring[k % 8])
        sim.step_async()
        step_no[0] += 1

    for k in range(args.warmup):
        one_step(k)
    sim.sync()

    # ---- timed region: exactly K steps, device time on the engine's stream, max over ranks
    sampler = ClockSampler(local_rank)
    sim.set_option(1, 1)
    launches0 = sim.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    sampler.start()
    ev0.record(stream)
    for k in range(args.steps):
        one_step(args.warmup + k)
    ev1.record(stream)
    sim.sync()
    barrier()
    clocks = sampler.stop()
    ms_total = ev0.elapsed_time(ev1)
    launches = sim.launch_count() - launches0 + args.steps     # + the action copy kernel per step
    kt = sim.kernel_times()
    sim.set_option(1, 0)

    # ---- e2e: the same K steps through gd_step_host with pinned host buffers
    e2e_steps = max(3, min(args.steps, 20))
    h_act = torch.empty(W, SLOTS, 3).pin_memory()
    h_act.copy_(ring[0].cpu())
    h_rew = torch.empty(W, SLOTS, 1).pin_memory()
    h_done = torch.empty(W, SLOTS, 1, dtype=torch.int32).pin_memory()
    h_info = torch.empty(W, SLOTS, 5, dtype=torch.int32).pin_memory()
    if step_no[0] + e2e_steps + 1 > 91:
        sim.reset(all_worlds); step_no[0] = 0
    sim.step_host(h_act, h_rew, h_done, h_info); step_no[0] += 1
    barrier()
    t0 = time.perf_counter()
    for k in range(e2e_steps):
        sim.step_host(h_act, h_rew, h_done, h_info)
        _ = float(h_rew[0, 0, 0])
    barrier()
    e2e_s = time.perf_counter() - t0
    step_no[0] += e2e_steps

    # ---- reduce over ranks: max time, summed agent counts
    ms_total, (tot_ctrl, tot_live) = shard.reduce_timing(ms_total, [n_ctrl, n_live], device="cuda")
    e2e_s, _ = shard.reduce_timing(e2e_s, [0.0], device="cuda")
    episode_stats = shard.reduce_stats(sim.stats(), device="cuda")     # the one NCCL exchange of the job

    if rank == 0:
        value = tot_ctrl * args.steps / (ms_total * 1e-3)
        e2e_value = tot_ctrl * e2e_steps / e2e_s
        peak, peak_src = measured_peak_gbs()
        b_step, b_obs = algorithmic_bytes(SLOTS, R_mean, AGENTS, True)
        obs_ms = kt["k_obs_ms"] / max(kt["steps"], 1)
        agents_ms = kt["k_agents_ms"] / max(kt["steps"], 1)
        achieved = b_obs * n_live / (obs_ms * 1e-3) / 1e9 if obs_ms > 0 else 0.0
        line = {
            "metric": "controlled agent-steps/sec", "value": value, "unit": "agent-steps/s",
            "n_gpus": world_size, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(world_size),
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "agent-steps/s", "steps": e2e_steps,
                    "h2d_bytes_per_step": total_worlds * SLOTS * 3 * 4,
                    "d2h_bytes_per_step": total_worlds * SLOTS * (4 + 4 + 20),
                    "api": "gd_step_host (pinned host actions in, reward/done/info out)"},
            "gpu_launches": launches,
            "roofline": {"bound": "hbm", "kernel": "k_obs", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "peak_source": peak_src,
                         "traffic": K_OBS_DRAM_BYTES_PER_LAUNCH if (world_size == 1 and total_worlds == TOTAL_WORLDS) else None,
                         "traffic_source": "profiles/r01_v13_k_obs_raw.txt (ncu --set full: dram__bytes_read.sum + dram__bytes_write.sum)",
                         "bytes_per_agent_step": b_obs, "agent_steps_per_launch": n_live,
                         "avg_launch_ms": obs_ms, "k_agents_avg_ms": agents_ms,
                         "step_bytes_per_agent_step": b_step,
                         "step_frac_of_hbm_roofline": (b_step * tot_live * args.steps / (ms_total * 1e-3) / 1e9) / (peak * world_size)},
            "agent_normalized_fps": tot_live * args.steps / (ms_total * 1e-3),
            "episode_stats": episode_stats,
            "build_s": build_s,
        }
        if world_size == 1 and not args.no_cpu_baseline:
            try:
                _, _, _, desc = run_reference_sample(20, 1, budget_s=15.0)
                line["cpu_baseline"] = desc
            except Exception as e:      # the baseline is reported, never required for the GPU number
                line["cpu_baseline"] = {"value": None, "unit": "controlled agent-steps/s", "cores": None,
                                        "kind": "reference", "sample": f"unavailable: {e}"}
        print(json.dumps(line))
    if world_size > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="engine", choices=["engine", "reference"])
    ap.add_argument("--worlds", type=int, default=0, help="override the total world count (debug)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "engine":
        args.warmup = 3
    return main_reference(args) if args.impl == "reference" else main_engine(args)


if __name__ == "__main__":
    sys.exit(main())
