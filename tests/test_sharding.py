"""The N > 1 path on the host: batch slicing and the max-over-ranks timing, over gloo with
world_size 2 (no GPU, no collective on the data path — SURVEY.md §8(e))."""
import os
import socket

import numpy as np
import pytest

from cantera_b200 import sharding


def test_slices_cover_batch_exactly():
    for n in (0, 1, 7, 1 << 20, (1 << 22) + 3):
        for world in (1, 2, 3, 8):
            b = [sharding.slice_bounds(n, r, world) for r in range(world)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(world - 1))
            sizes = [e - s for s, e in b]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sharding.slice_bounds(10, 2, 2)


def test_permuted_shares_are_disjoint_and_complete():
    n, world = 1000, 8
    parts = [sharding.permuted_indices(n, r, world) for r in range(world)]
    allidx = np.concatenate(parts)
    assert np.array_equal(np.sort(allidx), np.arange(n))


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n = 1001
    b, e = sharding.slice_bounds(n, rank, world)
    idx = sharding.permuted_indices(n, rank, world)
    # every rank "processes" its slice; the timed region is bracketed by a barrier on both sides
    dist.barrier()
    local_ms = 10.0 + 5.0 * rank
    dist.barrier()
    ms = sharding.timed_max(local_ms)
    units = sharding.total_units(e - b)
    q.put((rank, b, e, len(idx), ms, units))
    dist.destroy_process_group()


def test_two_ranks_gloo():
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, b0, e0, n0, ms0, u0), (r1, b1, e1, n1, ms1, u1) = out
    assert (b0, e0, b1, e1) == (0, 501, 501, 1001)
    assert n0 + n1 == 1001
    assert ms0 == ms1 == 15.0          # max over ranks, identical on both
    assert u0 == u1 == 1001            # whole-job units
