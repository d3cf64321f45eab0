"""CPU: multi-GPU host logic (frame sharding + result gather) with a world_size-2 gloo group."""
import os
import socket

import numpy as np
import pytest
import torch.multiprocessing as mp

from mantis_b200.sharding import array_to_results, frame_shard, gather_results, results_to_array


def test_frame_shard_covers_range():
    for n in (0, 1, 7, 4096, 4097):
        for world in (1, 2, 3, 8):
            spans = [frame_shard(n, world, r) for r in range(world)]
            assert spans[0][0] == 0
            assert sum(c for _, c in spans) == n
            for (f0, c0), (f1, _) in zip(spans, spans[1:]):
                assert f0 + c0 == f1
            assert max(c for _, c in spans) - min(c for _, c in spans) <= 1
    with pytest.raises(ValueError):
        frame_shard(10, 2, 2)


def test_result_array_round_trip():
    from mantis_b200.binding import Result

    res = (Result * 3)()
    for i in range(3):
        res[i].pos[0] = 1.5 * i
        res[i].status = i
        res[i].n_particles = 700 + i
    back = array_to_results(results_to_array(res))
    assert [back[i].pos[0] for i in range(3)] == [0.0, 1.5, 3.0]
    assert [back[i].n_particles for i in range(3)] == [700, 701, 702]


def _worker(rank, world, port, n_total, q):
    import torch.distributed as dist
    from mantis_b200.binding import Result

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    first, count = frame_shard(n_total, world, rank)
    res = (Result * count)()
    for i in range(count):  # stand-in for this rank's mantis_localize_batch() output
        res[i].pos[0] = float(first + i)
        res[i].status = (first + i) % 4
    full = gather_results(res, dist, n_total)
    if rank == 0:
        q.put(([full[i].pos[0] for i in range(n_total)], [full[i].status for i in range(n_total)]))
    else:
        assert full is None
    dist.barrier()
    dist.destroy_process_group()


def test_gather_results_gloo_world2():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    n_total, world = 11, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_total, q)) for r in range(world)]
    for p in procs:
        p.start()
    pos, status = q.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert pos == [float(i) for i in range(n_total)]
    assert status == [i % 4 for i in range(n_total)]
