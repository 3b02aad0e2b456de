"""N>1 host logic on CPU: frame-pair sharding + max-over-ranks timing with world_size 2 over gloo (no GPU)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from stereo_vision_b200 import shard


def test_shard_range_partitions_any_total():
    for total in (0, 1, 7, 8, 256, 257):
        for world in (1, 2, 3, 8):
            spans = [shard.shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard.shard_range(8, 2, 2)


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    a, b = shard.shard_range(256, rank, world)
    ms = shard.max_over_ranks(10.0 + 5.0 * rank)           # the slower rank defines the step time
    got = shard.gather_per_rank([float(i) for i in range(a, min(a + 3, b))])
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, a, b, ms, got))


def test_two_ranks_gloo():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=120) for _ in ps)
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, a0, b0, ms0, g0), (r1, a1, b1, ms1, g1) = res
    assert (a0, b0, a1, b1) == (0, 128, 128, 256)
    assert ms0 == ms1 == 15.0
    assert g0 == g1 == [[0.0, 1.0, 2.0], [128.0, 129.0, 130.0]]
