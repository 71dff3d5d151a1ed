"""World sharding across the GPUs of one box. Worlds are independent, so each rank
owns a contiguous block of them behind its own SimManager (the reference API already takes gpu_id,
src/bindings.cpp:93-101). No data-path collective exists; the only exchange is the reduction of the small
episode-statistics vector (gd_stats) for reporting, done here with torch.distributed (NCCL on GPUs, gloo in
the CPU tests)."""
from __future__ import annotations

from typing import Sequence

STAT_FIELDS = ("steps", "agent_steps", "controlled_steps", "episodes_reset", "collided_road",
               "collided_vehicle", "collided_nonvehicle", "goals_reached", "dones", "sat_tests")


def world_block(total_worlds: int, rank: int, world_size: int) -> range:
    """Contiguous block of global world indices owned by `rank`; the first total % world_size ranks hold one
    extra world, so the blocks tile [0, total) exactly."""
    if not (0 <= rank < world_size):
        raise ValueError(f"rank {rank} outside [0, {world_size})")
    base, extra = divmod(total_worlds, world_size)
    start = rank * base + min(rank, extra)
    return range(start, start + base + (1 if rank < extra else 0))


def scenes_for_rank(scene_pool: Sequence[str], total_worlds: int, rank: int, world_size: int) -> list[str]:
    """Scenario path of every world of this rank: global world i runs scene_pool[i % len(scene_pool)], so the
    assignment does not depend on how many GPUs the job uses."""
    return [scene_pool[i % len(scene_pool)] for i in world_block(total_worlds, rank, world_size)]


def reduce_stats(local: dict, device=None, group=None) -> dict:
    """Sum the per-rank episode statistics over all ranks (every field is a counter except `steps`, which is
    the same on every rank and is reduced with MAX). Returns the global dict on every rank. Without an
    initialised process group it returns `local` unchanged."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return dict(local)
    vec = torch.tensor([int(local.get(k, 0)) for k in STAT_FIELDS], dtype=torch.int64, device=device)
    mx = vec[:1].clone()
    dist.all_reduce(vec, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(mx, op=dist.ReduceOp.MAX, group=group)
    out = {k: int(v) for k, v in zip(STAT_FIELDS, vec.tolist())}
    out["steps"] = int(mx.item())
    return out


def reduce_timing(ms_local: float, counters: Sequence[float], device=None, group=None):
    """Max of a device-measured duration over ranks and the sum of the given counters."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return float(ms_local), [float(c) for c in counters]
    t = torch.tensor([ms_local], dtype=torch.float64, device=device)
    c = torch.tensor(list(counters), dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    dist.all_reduce(c, op=dist.ReduceOp.SUM, group=group)
    return float(t.item()), c.tolist()
