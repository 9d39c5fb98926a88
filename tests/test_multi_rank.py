"""world_size-2 gloo tests (CPU) of the N>1 host path: shard arithmetic, the single gather,
and global ordering.  The per-rank 'solver' here is the oracle (test infrastructure) so the
gathered bytes can be checked against a one-process solve of the whole batch."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import garb200_pkg  # spawned ranks re-import this module without conftest.py

garb200_pkg.load()
from aligator_swk_b200 import sharding  # noqa: E402
from aligator_swk_b200.synth import generate_problems  # noqa: E402


def test_shard_bounds_cover_and_balance():
    for total in (0, 1, 2, 7, 16, 16384, 16385):
        for world in (1, 2, 3, 4, 8):
            spans = [sharding.shard_bounds(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sharding.shard_bounds(4, 2, 2)


def test_trajectory_sizes_and_carve():
    sizes = sharding.trajectory_sizes(3, 5, 4, 2, 1)
    assert sizes == [3 * 6 * 4, 3 * 6 * 2, 3 * 6 * 1, 3 * (4 + 5 * 4)]
    buf = torch.arange(sum(sizes), dtype=torch.float64)
    xs, us, vs, lb = sharding.carve(buf, sizes)
    assert xs[0] == 0 and us[0] == sizes[0] and lb[-1] == sum(sizes) - 1
    xs[0] = -1.0
    assert buf[0] == -1.0          # views, not copies
    with pytest.raises(ValueError):
        sharding.carve(buf[:10], sizes)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, total, dims, tmp):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for p in (root, os.path.join(root, "oracle")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import garb200_pkg
    garb200_pkg.load()
    import pyoracle
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        nx, nu, nc, T = dims
        lo, hi = sharding.shard_bounds(total, rank, world)
        pp = generate_problems(total, T, nx, nu, nc, 0, seed=7, dist="explicit")   # same bytes on all ranks
        sizes = sharding.trajectory_sizes(hi - lo, T, nx, nu, nc)
        send = torch.empty(sum(sizes), dtype=torch.float64)
        xs, us, vs, lb = sharding.carve(send, sizes)
        for i, b in enumerate(range(lo, hi)):
            sol = pyoracle.prox_solve(pp, 1e-8, 1e-8, b=b)
            xs[i * (T + 1) * nx:(i + 1) * (T + 1) * nx] = torch.from_numpy(np.ascontiguousarray(sol["xs"]).ravel())
            us[i * (T + 1) * nu:(i + 1) * (T + 1) * nu] = torch.from_numpy(np.ascontiguousarray(sol["us"]).ravel())
            if nc:
                vs[i * (T + 1) * nc:(i + 1) * (T + 1) * nc] = torch.from_numpy(np.ascontiguousarray(sol["vs"]).ravel())
            lb[i * (nx + T * nx):(i + 1) * (nx + T * nx)] = torch.from_numpy(np.ascontiguousarray(sol["lbdas"]).ravel())
        recv = sharding.gather_to_root(dist, send, rank, world)
        # max-over-ranks timing reduction used by bench.py
        t = torch.tensor([float(rank + 1)], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        assert t.item() == world
        if rank == 0:
            assert len(recv) == world
            g = sharding.assemble(recv, sizes)
            np.save(os.path.join(tmp, "xs.npy"), g[0].numpy())
            np.save(os.path.join(tmp, "lb.npy"), g[3].numpy())
        else:
            assert recv is None
    finally:
        dist.destroy_process_group()


def test_gloo_world2_gather_matches_single_process(tmp_path, oracle):
    nx, nu, nc, T, total = 4, 2, 0, 6, 6
    port = _free_port()
    mp.spawn(_worker, args=(2, port, total, (nx, nu, nc, T), str(tmp_path)), nprocs=2, join=True)
    pp = generate_problems(total, T, nx, nu, nc, 0, seed=7, dist="explicit")
    xs = np.load(tmp_path / "xs.npy").reshape(total, T + 1, nx)
    lb = np.load(tmp_path / "lb.npy").reshape(total, nx + T * nx)
    for b in range(total):
        ref = oracle.prox_solve(pp, 1e-8, 1e-8, b=b)
        assert np.array_equal(xs[b], np.asarray(ref["xs"]).reshape(T + 1, nx)), b
        assert np.array_equal(lb[b], np.asarray(ref["lbdas"]).ravel()), b
