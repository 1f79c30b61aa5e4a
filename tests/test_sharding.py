"""CPU tests of the multi-GPU host logic: chunk assignment, chunk windows, and the per-iteration all-reduce of the tallies
over a world of 2 gloo ranks."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from smcsmc_b200.sharding import chunks_for_rank, chunk_windows, allreduce_tallies


def test_chunks_for_rank_partition():
    for n in (0, 1, 7, 16, 30):
        for world in (1, 2, 3, 8):
            parts = [chunks_for_rank(n, r, world) for r in range(world)]
            assert sorted(sum(parts, [])) == list(range(n))
            sizes = [len(p) for p in parts]
            assert max(sizes) - min(sizes) <= 1
            for p in parts:
                assert p == list(range(p[0], p[0] + len(p))) if p else True
    with pytest.raises(ValueError):
        chunks_for_rank(4, 2, 2)


def test_chunk_windows():
    w = chunk_windows(1e8, 16)
    assert w[0] == (1, 6.25e6) and w[15][0] == 1 + int(15 * 6.25e6)
    assert sum(x[1] for x in w) == 1e8


def _fake_chunk_result(c, E):
    rng = np.random.default_rng(1000 + c)
    return rng.random((6, E)), -1000.0 * (c + 1)


def _worker(rank, world, port, E, nchunks, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    t, lnz = np.zeros((6, E)), 0.0
    for c in chunks_for_rank(nchunks, rank, world):
        a, b = _fake_chunk_result(c, E)
        t += a
        lnz += b
    tt, ll = allreduce_tallies(t, lnz)
    out[rank] = (tt, ll)
    dist.destroy_process_group()


def test_allreduce_tallies_world2():
    E, nchunks, world = 5, 7, 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, 29512, E, nchunks, out), nprocs=world, join=True)
    want_t = sum(_fake_chunk_result(c, E)[0] for c in range(nchunks))
    want_l = sum(_fake_chunk_result(c, E)[1] for c in range(nchunks))
    for r in range(world):
        np.testing.assert_allclose(out[r][0], want_t, rtol=1e-14)
        assert out[r][1] == pytest.approx(want_l)


def test_allreduce_identity_without_process_group():
    t = np.arange(12.0).reshape(6, 2)
    tt, l = allreduce_tallies(t, -3.5)
    assert np.array_equal(tt, t) and l == -3.5
