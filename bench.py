#!/usr/bin/env python
"""Benchmark of the GPUDrive step path (BASELINE.json: controlled agent-steps/sec, 8192 worlds x
64 agents, DeltaLocal dynamics, full observations + 3x50-ray lidar, 1/2/4/8 B200).

    python bench.py --gpus N --steps K --warmup W          # this engine (CUDA), config 3
    python bench.py --config {2,3,4,5} ...                  # the other BASELINE.json configurations
    python bench.py --unique-scenes 8192 ...                # distinct scenes per job (default 64, cycled)
    python bench.py --impl reference ...                    # the reference's own CPU Manager

One "step" = one SimManager.step() over every world of the rank (actions already in HBM for
`value`; through gd_step_host with pinned host buffers for `e2e`). Worlds are sharded across
ranks (total / N each, strong scaling, no collective inside the step); the only collective is the
end-of-run reduction of the timing / statistics vector. The workloads themselves (scenes, Parameters,
actions) are defined in gpudrive_lab_b200/workload.py, which the parity tests share.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from gpudrive_lab_b200 import workload as wl  # noqa: E402

SLOTS = wl.SLOTS


def k_obs_traffic(config, unique_scenes):
    """dram__bytes_read.sum + dram__bytes_write.sum of one k_obs launch from the committed `ncu --set full`
    capture of this workload on one GPU (profiles/k_obs_traffic.json), or None."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "k_obs_traffic.json")))
        e = t.get(f"config{config}_u{unique_scenes}")
        return (float(e["bytes"]), e["source"]) if e else (None, None)
    except Exception:
        return None, None


class ClockSampler:
    """nvidia-smi sampled DURING the timed region (B200_PROFILING.md clocks line). It is started ahead of the
    region (nvidia-smi needs ~0.1 s to produce its first line, an 8-GPU step only 1 ms) and every line is
    time-stamped on receipt; stop(t0, t1) keeps the samples taken inside [t0, t1], or, if the region was shorter
    than one sampling period, the ones nearest to it."""
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
                                          "-lms", "20", "-i", str(self.gpu)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def wait_ready(self, timeout=2.0):
        t_end = time.time() + timeout
        while self.proc and not self.lines and time.time() < t_end:
            time.sleep(0.01)

    def stop(self, t0=None, t1=None):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.05)
        self.proc.terminate()
        lines = list(self.lines)
        window = "timed region"
        if t0 is not None:
            inside = [(t, ln) for t, ln in lines if t0 <= t <= t1 + 0.02]
            if not inside and lines:        # region shorter than a sampling period: the sample(s) nearest to it
                mid = 0.5 * (t0 + t1)
                inside = sorted(lines, key=lambda x: abs(x[0] - mid))[:2]
                window = "nearest to the timed region"
            lines = inside
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for _, ln in lines:
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
                "samples": len(sm), "window": window, "reasons": sorted(reasons)}


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


def workload_config(config, n_gpus, total_worlds, unique_scenes, extra=None):
    c = wl.CONFIGS[config]
    d = {"workload": f"{c['name']}: {c['text']}",
         "total_worlds": total_worlds, "worlds_per_gpu": total_worlds // max(n_gpus, 1),
         "agents_per_world": c["agents"], "agent_slots": SLOTS, "unique_scenes": unique_scenes,
         "parallelism": f"worlds sharded over {n_gpus} GPU(s)",
         "l2": "per-step export traffic (>= 1 GB per GPU) exceeds the 126 MB L2; no explicit flush needed"}
    if extra:
        d.update(extra)
    return d


# --------------------------------------------------------------------------------------------
# reference arm / CPU baseline: the UNMODIFIED reference Manager, ExecMode::CPU (oracle/_ref)
# --------------------------------------------------------------------------------------------

def run_reference_sample(config, steps, warmup, worlds=None, budget_s=20.0):
    """Times Manager::step() of the reference on a bounded sample of the same workload.
    Returns (controlled agent-steps/s, ms per step, steps, description dict, worlds)."""
    from oracle import refmgr
    if not refmgr.available():
        raise RuntimeError("oracle/_ref/libgd_ref.so is missing (built by __graft_entry__.build() where "
                           "/root/reference exists)")
    c = wl.CONFIGS[config]
    cores = len(os.sched_getaffinity(0))
    W = worlds if worlds is not None else max(2 * cores, 8)
    scenes_u = wl.make_scenes(config, min(c["unique_scenes"], W))
    scenes = [scenes_u[i % len(scenes_u)] for i in range(W)]
    pd = wl.params_dict(config)
    t0 = time.time()
    mgr = refmgr.RefManager(scenes, refmgr.RefParams(**pd))
    build_s = time.time() - t0
    ctrl = int(mgr.tensor("controlled_state")[:, :, 0].sum())
    act = mgr.tensor("action")
    model = pd["dynamicsModel"]
    for i in range(warmup):
        act[...] = wl.random_actions(i, W, SLOTS, model)
        mgr.step()
    # bound the sample: keep the whole call within `budget_s` of CPU work
    t0 = time.time()
    act[...] = wl.random_actions(warmup, W, SLOTS, model)
    mgr.step()
    one = time.time() - t0
    steps = max(2, min(steps, int(budget_s / max(one, 1e-3))))
    acts = [wl.random_actions(warmup + 1 + i, W, SLOTS, model) for i in range(steps)]
    t0 = time.perf_counter()
    for i in range(steps):
        act[...] = acts[i]
        mgr.step()
    dt = time.perf_counter() - t0
    value = ctrl * steps / dt
    desc = {"kind": "reference", "cores": cores, "value": value, "unit": "controlled agent-steps/s",
            "sample": f"{W} worlds x {c['agents']} agents x {steps} steps of the same workload ({c['name']}) through the "
                      f"reference's Manager::step() (ExecMode::CPU, one pinned worker per core; includes its per-step "
                      f"export copies and the BEV pass it cannot disable); construction {build_s:.1f}s not timed"}
    mgr.close()
    return value, dt / steps * 1e3, steps, desc, W


def main_reference(args):
    """The reference arm. Runs on rank 0 only, in a process that never touches CUDA."""
    import faulthandler
    faulthandler.enable()
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    c = wl.CONFIGS[args.config]
    value, ms, steps, desc, W = run_reference_sample(args.config, args.steps, max(args.warmup, 1),
                                                     worlds=args.worlds or None, budget_s=args.budget)
    line = {
        "impl": "reference", "metric": "controlled agent-steps/sec", "value": value, "unit": "agent-steps/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": max(args.warmup, 1), "ms_per_step": ms,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        # the engine arm's workload; this line stepped a bounded sample of it: `sample_worlds` worlds
        "config": workload_config(args.config, args.gpus, c["worlds"], args.unique_scenes or c["unique_scenes"],
                                  {"sample_worlds": W, "sample_steps": steps,
                                   "parallelism": f"bounded sample on {desc['cores']} host cores, one world per worker"}),
        "cpu_baseline": desc,
        "e2e": {"value": value, "unit": "agent-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    line["config"]["worlds_per_gpu"] = None     # a CPU run: `sample_worlds` is what this line actually stepped
    print(json.dumps(line), flush=True)
    return 0


def cpu_baseline_child(config):
    """The cpu_baseline leg in a child process: a crash on the CPU side (the reference is C++ driven through
    ctypes; round 1 lost a GPU result to a SIGSEGV in it) cannot take the GPU number with it."""
    cmd = [sys.executable, os.path.abspath(__file__), "--impl", "reference", "--config", str(config),
           "--steps", "20", "--warmup", "1", "--budget", "15"]
    env = dict(os.environ)
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE"):
        env.pop(k, None)
    try:
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=env)
    except subprocess.TimeoutExpired:
        return {"value": None, "unit": "controlled agent-steps/s", "cores": None, "kind": "reference",
                "sample": "unavailable: the reference child process timed out after 900 s", "rc": None}
    for ln in reversed(r.stdout.strip().splitlines()):
        try:
            d = json.loads(ln)
            desc = d["cpu_baseline"]
            desc["rc"] = r.returncode
            return desc
        except Exception:
            continue
    return {"value": None, "unit": "controlled agent-steps/s", "cores": None, "kind": "reference", "rc": r.returncode,
            "sample": "unavailable: the reference child process printed no result; stderr tail: " + r.stderr[-600:]}


# --------------------------------------------------------------------------------------------
# engine arm
# --------------------------------------------------------------------------------------------

def main_engine(args):
    import torch
    import torch.distributed as dist
    from gpudrive_lab_b200 import madrona_gpudrive as mg
    from gpudrive_lab_b200 import shard

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

    config = args.config
    c = wl.CONFIGS[config]
    total_worlds = args.worlds or c["worlds"]
    unique = args.unique_scenes or c["unique_scenes"]
    W = len(shard.world_block(total_worlds, rank, world_size))
    pd = wl.params_dict(config)
    model = pd["dynamicsModel"]
    if rank == 0:
        wl.make_scenes(config, unique)
    if world_size > 1:
        dist.barrier()
    scenes_u = wl.make_scenes(config, unique)
    # contiguous block of worlds per rank; world i uses scenario (global index) % unique
    scenes = wl.world_scenes(scenes_u, total_worlds, rank, world_size)
    t0 = time.time()
    sim = mg.SimManager(mg.madrona.ExecMode.CUDA, local_rank, scenes, wl.engine_params(pd), max_agents=SLOTS)
    build_s = time.time() - t0
    sim.set_option(0, 0)        # knn index side buffer off (debug aid, not part of the reference's tensors)
    auto_reset = config == 4    # PPO-style rollout: device-side auto-reset + policy-observation read-out per step
    if auto_reset:
        sim.auto_reset_pool([sim.pool_add(p) for p in scenes_u])
        sim.set_option(3, 1)    # the policy observation is consumed on the device: no host wait per step
        if not args.separate_policy_obs:
            sim.set_option(4, 1)   # k_obs writes the policy observation itself (GD_OPT_FUSED_POLICY_OBS)
    stream = torch.cuda.ExternalStream(sim.cuda_stream(), device=local_rank)
    shape = sim.shape_tensor().to_torch()
    ctrl_t = sim.controlled_state_tensor().to_torch()
    n_ctrl = int(ctrl_t.sum().item())
    n_live = int(shape[:, 0].sum().item())
    R_mean = float(shape[:, 1].float().mean().item())
    action = sim.action_tensor().to_torch()

    # action ring, resident in HBM before the timed region
    g = torch.Generator(device="cuda"); g.manual_seed(1234 + rank)
    ring = []
    for _ in range(8):
        a = torch.empty(W, SLOTS, 3, device="cuda")
        if auto_reset:          # the env's discrete action table (gpudrive/env/config.py:64-70), random policy stand-in
            acc = torch.round(torch.linspace(-4.0, 4.0, 7), decimals=3).cuda()
            steer = torch.round(torch.linspace(-torch.pi, torch.pi, 13), decimals=3).cuda()
            a[..., 0] = acc[torch.randint(0, 7, (W, SLOTS), device="cuda", generator=g)]
            a[..., 1] = steer[torch.randint(0, 13, (W, SLOTS), device="cuda", generator=g)]
            a[..., 2] = 0.0
        elif model == wl.DELTA_LOCAL:
            a[..., 0].uniform_(-0.5, 1.5, generator=g); a[..., 1].uniform_(-0.2, 0.2, generator=g)
            a[..., 2].uniform_(-0.05, 0.05, generator=g)
        else:
            a[..., 0].uniform_(-3.0, 2.0, generator=g); a[..., 1].uniform_(-0.7, 0.7, generator=g)
            a[..., 2].uniform_(-0.3, 0.3, generator=g)
        ring.append(a)
    torch.cuda.synchronize()

    def barrier():
        if world_size > 1:
            dist.barrier()
        torch.cuda.synchronize()

    all_worlds = list(range(W))
    step_no = [0]
    episode_steps = c.get("episode_steps", 91)

    def one_step(k):
        if not auto_reset and step_no[0] == episode_steps:   # episode boundary: reset inside the timed region (time, not a step)
            sim.reset(all_worlds)
            step_no[0] = 0
        with torch.cuda.stream(stream):
            action[:, :, :3].copy_(ring[k % 8])
        sim.step_async()
        if auto_reset:
            sim.auto_reset()
            sim.policy_observation_tensor(torch_cuda_arithmetic=True)     # the rollout's read-out, on the device
        step_no[0] += 1

    sampler = ClockSampler(local_rank)
    sampler.start()                       # ahead of the warm-up: nvidia-smi needs ~0.1 s to produce its first line
    for k in range(args.warmup):
        one_step(k)
    sim.sync()

    # ---- timed region: exactly K steps, device time on the engine's stream, max over ranks
    sampler.wait_ready()
    sim.set_option(1, 1)
    launches0 = sim.launch_count()
    stats0 = sim.stats()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_wall0 = time.time()
    ev0.record(stream)
    for k in range(args.steps):
        one_step(args.warmup + k)
    ev1.record(stream)
    sim.sync()
    t_wall1 = time.time()
    barrier()
    clocks = sampler.stop(t_wall0, t_wall1)
    ms_total = ev0.elapsed_time(ev1)
    launches = sim.launch_count() - launches0 + args.steps     # + the action copy kernel per step
    kt = sim.kernel_times()
    stats1 = sim.stats()
    sim.set_option(1, 0)

    # ---- e2e: the same steps through gd_step_host with pinned host buffers
    e2e_steps = max(3, min(args.steps, 20))
    h_act = torch.empty(W, SLOTS, 3).pin_memory()
    h_act.copy_(ring[0].cpu())
    h_rew = torch.empty(W, SLOTS, 1).pin_memory()
    h_done = torch.empty(W, SLOTS, 1, dtype=torch.int32).pin_memory()
    h_info = torch.empty(W, SLOTS, 5, dtype=torch.int32).pin_memory()
    if not auto_reset:
        sim.reset(all_worlds); step_no[0] = 0

    def e2e_step():
        if not auto_reset and step_no[0] == episode_steps:
            sim.reset(all_worlds); step_no[0] = 0
        step_no[0] += 1
        sim.step_host(h_act, h_rew, h_done, h_info)
        if auto_reset:
            sim.auto_reset()
            sim.policy_observation_tensor(torch_cuda_arithmetic=True)
        return float(h_rew[0, 0, 0])

    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for k in range(e2e_steps):
        e2e_step()
    sim.sync()
    barrier()
    e2e_s = time.perf_counter() - t0

    # ---- reduce over ranks: max time, summed agent counts
    sat = float(stats1["sat_tests"] - stats0["sat_tests"])
    ms_total, (tot_ctrl, tot_live, tot_sat) = shard.reduce_timing(ms_total, [n_ctrl, n_live, sat], device="cuda")
    e2e_s, _ = shard.reduce_timing(e2e_s, [0.0], device="cuda")
    episode_stats = shard.reduce_stats(sim.stats(), device="cuda")     # the one NCCL exchange of the job

    if rank == 0:
        value = tot_ctrl * args.steps / (ms_total * 1e-3)
        e2e_value = tot_ctrl * e2e_steps / e2e_s
        peak, peak_src = measured_peak_gbs()
        fused_pol = auto_reset and not args.separate_policy_obs        # k_obs then also writes 3 368 floats per agent
        b_step, b_obs = wl.algorithmic_bytes(SLOTS, R_mean, c["agents"], bool(pd["enableLidar"]),
                                             policy_obs_floats=(6 + (SLOTS - 1) * 6 + wl.K_ROAD * 13) if fused_pol else 0)
        obs_ms = kt["k_obs_ms"] / max(kt["steps"], 1)
        agents_ms = kt["k_agents_ms"] / max(kt["steps"], 1)
        achieved = b_obs * n_live / (obs_ms * 1e-3) / 1e9 if obs_ms > 0 else 0.0
        traffic, traffic_src = (None, None)
        if world_size == 1 and total_worlds == c["worlds"]:
            traffic, traffic_src = k_obs_traffic(config, unique)
        line = {
            "metric": "controlled agent-steps/sec", "value": value, "unit": "agent-steps/s",
            "n_gpus": world_size, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(config, world_size, total_worlds, unique),
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "agent-steps/s", "steps": e2e_steps,
                    "h2d_bytes_per_step": total_worlds * SLOTS * 3 * 4,
                    "d2h_bytes_per_step": total_worlds * SLOTS * (4 + 4 + 20),
                    "api": "gd_step_host (pinned host actions in, reward/done/info out; observations stay on the device)"
                           + (" + gd_auto_reset + gd_policy_obs" if auto_reset else "")},
            "gpu_launches": launches,
            "roofline": {"bound": "hbm", "kernel": "k_obs", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "peak_source": peak_src,
                         "traffic": traffic, "traffic_source": traffic_src,
                         "bytes_per_agent_step": b_obs, "agent_steps_per_launch": n_live,
                         "avg_launch_ms": obs_ms, "k_agents_avg_ms": agents_ms,
                         "step_bytes_per_agent_step": b_step,
                         "step_frac_of_hbm_roofline": (b_step * tot_live * args.steps / (ms_total * 1e-3) / 1e9) / (peak * world_size)},
            "agent_normalized_fps": tot_live * args.steps / (ms_total * 1e-3),
            "sat_tests_per_s": tot_sat / (ms_total * 1e-3),
            "episode_stats": episode_stats,
            "build_s": build_s,
        }
        if world_size == 1 and not args.no_cpu_baseline:
            # in a child process, after every GPU number is in hand; the GPU line is printed whatever happens there
            try:
                line["cpu_baseline"] = cpu_baseline_child(config)
            except Exception as e:
                line["cpu_baseline"] = {"value": None, "unit": "controlled agent-steps/s", "cores": None,
                                        "kind": "reference", "sample": f"unavailable: {e}"}
        print(json.dumps(line), flush=True)
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
    ap.add_argument("--config", type=int, default=3, choices=sorted(wl.CONFIGS), help="BASELINE.json configuration")
    ap.add_argument("--worlds", type=int, default=0, help="override the total world count (debug)")
    ap.add_argument("--unique-scenes", type=int, default=0, help="distinct scenarios per job (default: the config's)")
    ap.add_argument("--budget", type=float, default=60.0, help="reference arm: seconds of CPU work to sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--separate-policy-obs", action="store_true",
                    help="config 4: compute the policy observation in its own pass instead of as k_obs' second output")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "engine":
        args.warmup = 3
    return main_reference(args) if args.impl == "reference" else main_engine(args)


if __name__ == "__main__":
    sys.exit(main())
