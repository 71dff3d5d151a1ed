"""Multi-GPU host logic on CPU: world blocks tile the job for any N, the scene assignment is independent of
N, and the episode-statistics reduction works over a real world_size-2 process group (gloo)."""
import os
import socket

import pytest
import torch
import torch.multiprocessing as mp

from gpudrive_lab_b200 import shard


@pytest.mark.parametrize("total,n", [(8192, 1), (8192, 2), (8192, 8), (10, 4), (3, 8), (0, 2)])
def test_world_blocks_tile(total, n):
    seen = []
    for r in range(n):
        b = shard.world_block(total, r, n)
        seen.extend(b)
        assert len(b) in (total // n, total // n + 1)
    assert seen == list(range(total))
    with pytest.raises(ValueError):
        shard.world_block(total, n, n)


def test_scene_assignment_independent_of_gpu_count():
    pool = [f"s{i}" for i in range(7)]
    one = shard.scenes_for_rank(pool, 100, 0, 1)
    for n in (2, 4, 8):
        cat = sum((shard.scenes_for_rank(pool, 100, r, n) for r in range(n)), [])
        assert cat == one


def test_reduce_without_process_group_is_identity():
    d = {"steps": 3, "agent_steps": 10}
    assert shard.reduce_stats(d) == d
    assert shard.reduce_timing(1.5, [2, 3]) == (1.5, [2.0, 3.0])


def _worker(rank, world_size, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    try:
        block = shard.world_block(10, rank, world_size)
        local = {"steps": 91, "agent_steps": 64 * 91 * len(block), "controlled_steps": 60 * 91 * len(block),
                 "episodes_reset": len(block), "collided_road": rank + 1, "collided_vehicle": 0,
                 "collided_nonvehicle": 0, "goals_reached": 5 * rank, "dones": 7}
        g = shard.reduce_stats(local)
        ms, cnt = shard.reduce_timing(10.0 + rank, [len(block), 1.0])
        q.put((rank, g, ms, cnt))
    finally:
        dist.destroy_process_group()


def test_stats_reduction_world_size_2():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    res = sorted((q.get() for _ in range(2)), key=lambda x: x[0])
    for rank, g, ms, cnt in res:
        assert g["steps"] == 91
        assert g["agent_steps"] == 64 * 91 * 10 and g["controlled_steps"] == 60 * 91 * 10
        assert g["episodes_reset"] == 10 and g["collided_road"] == 3 and g["goals_reached"] == 5 and g["dones"] == 14
        assert ms == 11.0 and cnt == [10.0, 2.0]
