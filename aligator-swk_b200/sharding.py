"""Batch sharding across ranks (one process per GPU).

The path shards over independent LQ problems: rank r owns the contiguous slice
``shard_bounds(total, r, world)`` of the batch, runs the whole backward/forward sweep on it
with no data-path collective, and a single gather brings the packed trajectories to rank 0
(BASELINE.json north_star).  The reference has no multi-process layer (its parallelism is
OpenMP legs inside one problem, /root/reference/include/aligator/gar/parallel-solver.hxx:84-133);
this module is the only place that touches ``torch.distributed`` and works on any backend
(nccl on the GPU box, gloo in tests/test_multi_rank.py).
"""
from __future__ import annotations


def shard_bounds(total: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous [lo, hi) slice of ``total`` problems for ``rank``; the first ``total % world``
    ranks get one extra problem.  Empty slices are legal (total < world)."""
    if world <= 0 or not (0 <= rank < world) or total < 0:
        raise ValueError(f"bad shard request total={total} rank={rank} world={world}")
    q, r = divmod(total, world)
    lo = rank * q + min(rank, r)
    return lo, lo + q + (1 if rank < r else 0)


def trajectory_sizes(batch: int, horizon: int, nx: int, nu: int, nc: int, nc0: int | None = None) -> list[int]:
    """Element counts of the packed (xs, us, vs, lbdas) arrays of garb200_forward (garb200.h)."""
    c0 = nx if nc0 is None else nc0
    return [batch * (horizon + 1) * nx, batch * (horizon + 1) * nu, batch * (horizon + 1) * nc,
            batch * (c0 + horizon * nx)]


def carve(buf, sizes):
    """Views of one flat buffer as the four packed arrays (no copies)."""
    out, off = [], 0
    for sz in sizes:
        out.append(buf[off:off + sz])
        off += sz
    if off > buf.numel():
        raise ValueError("buffer too small for the packed trajectories")
    return out


def gather_to_root(dist, send, rank: int, world: int, recv=None):
    """One gather of every rank's flat trajectory buffer to rank 0.  ``recv`` (rank 0 only)
    may be a preallocated list of ``world`` tensors shaped like ``send``; returns it on rank 0,
    None elsewhere.  All ranks must hold the same number of problems (weak scaling)."""
    import torch
    if world == 1 or dist is None:
        return [send]
    if rank == 0 and recv is None:
        recv = [torch.empty_like(send) for _ in range(world)]
    dist.gather(send, recv if rank == 0 else None, dst=0)
    return recv if rank == 0 else None


def assemble(recv, sizes):
    """Rank-0 side: list of per-rank flat buffers -> four arrays in global problem order
    (rank-major, which is the order of ``shard_bounds``)."""
    import torch
    parts = [carve(r, sizes) for r in recv]
    return [torch.cat([p[i] for p in parts]) for i in range(4)]
