"""N > 1 path on CPU: world_size-2 gloo.  The data path has no collective; what crosses ranks is the all-reduce of the
per-variant tally vector (u64) and, per refinement iteration, of tkpg (f64).  Checks shard bounds, the Reducer, and that the
solve driven by all-reduced partial estimates equals the single-rank solve."""
from __future__ import annotations

import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from speq_b200 import driver


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_shard_bounds_cover_everything():
    for n in (0, 1, 7, 100, 10_000_001):
        for w in (1, 2, 3, 8):
            b = [driver.shard_bounds(n, r, w) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1
    assert list(driver.batch_ranges(3, 11, 4)) == [(3, 7), (7, 11)]


def _worker(rank: int, world: int, port: int, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        red = driver.Reducer(dist.group.WORLD)
        V = 5
        rng = np.random.default_rng(100 + rank)
        ints = rng.integers(0, 2 ** 40, V + 4).astype(np.uint64)       # unique[V], total, ambiguous, lookups, reads
        s_int = red.sum_u64(ints)
        f = rng.random(V)
        s_f = red.sum_f64(f)
        # a fake per-rank "refine": tkpg_r(p) = w_r * p / sum(p) * total_r  -- linear in the rank's share
        total_r = float(1000 * (rank + 1))

        def estimate(p):
            part = total_r * p / p.sum()
            return red.sum_f64(part)

        out = estimate(np.arange(1.0, V + 1))
        q.put((rank, ints, s_int, f, s_f, out))
    finally:
        dist.destroy_process_group()


def test_gloo_world2_reduction():
    world = 2
    port = _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in range(world)], key=lambda x: x[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    ints = [r[1] for r in res]
    want_i = ints[0] + ints[1]
    want_f = res[0][3] + res[1][3]
    p = np.arange(1.0, 6.0)
    want_e = 3000.0 * p / p.sum()
    for r in res:
        assert np.array_equal(r[2], want_i) and r[2].dtype == np.uint64      # integer tallies: exact, rank-count independent
        assert np.allclose(r[4], want_f, rtol=1e-15)
        assert np.allclose(r[5], want_e, rtol=1e-15)


def test_bench_strong_scaling_shards_partition_the_read_set():
    """bench.py shards the chunks of ONE global read set over the ranks (strong scaling): at every N the ranks' chunk ranges are
    disjoint, ordered and cover all chunks, so the all-reduced tallies do not depend on N."""
    import importlib.util
    from pathlib import Path
    spec = importlib.util.spec_from_file_location("bench", Path(__file__).resolve().parent.parent / "bench.py")
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    for n_chunks in (1, 7, 200, 201):
        for world in (1, 2, 4, 8):
            parts = [bench.shard_chunks(n_chunks, r, world) for r in range(world)]
            assert parts[0][0] == 0 and parts[-1][1] == n_chunks
            assert all(a[1] == b[0] for a, b in zip(parts[:-1], parts[1:]))
            assert max(b - a for a, b in parts) - min(b - a for a, b in parts) <= 1
