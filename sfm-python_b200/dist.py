"""Point-sharded multi-GPU bundle adjustment: one process per GPU (torchrun), NCCL inside libb200ba.

Sharding: observations are point-major (``sfm_lite/graph.py:186-191``), so a contiguous
point range owns a contiguous observation range; each rank eliminates its own point blocks, the packed
buffer ``[S blocks | b | g_c | diag U | cost]`` is all-reduced ONCE per linearisation (ncclAllReduce over
NVLink, issued by the library on its own stream), and every rank runs the same reduced-camera solve on
the same bits, so accept/reject decisions stay in lock-step without further synchronisation.
``torch.distributed`` is only the bootstrap plumbing (rank/world, broadcast of the NCCL unique id,
gathering the point parameters at the end).
"""
from __future__ import annotations

import numpy as np

from .bundle_adjustment import BAResult, Engine, MESSAGES


def plan_shards(point_indices, n_pts, world):
    """Contiguous point ranges [(lo, hi)] * world, balanced by observation count."""
    deg = np.bincount(np.asarray(point_indices, dtype=np.int64), minlength=n_pts)
    return plan_shards_from_degrees(deg, world)


def plan_shards_from_degrees(deg, world):
    csum = np.concatenate([[0], np.cumsum(deg)])
    n = int(csum[-1])
    bounds = [0]
    for r in range(1, world):
        target = (n * r) // world
        p = int(np.searchsorted(csum, target, side="left"))
        p = min(max(p, bounds[-1]), len(deg))
        bounds.append(p)
    bounds.append(len(deg))
    return [(bounds[r], bounds[r + 1]) for r in range(world)]


def local_problem(camera_indices, point_indices, points_2d, lo, hi):
    """Observations of points [lo, hi) with shard-local point ids; also returns their global positions."""
    pi = np.asarray(point_indices, dtype=np.int64)
    sel = np.flatnonzero((pi >= lo) & (pi < hi))
    return np.asarray(camera_indices)[sel], pi[sel] - lo, np.asarray(points_2d)[sel], sel


def split_x(x, n_cam, lo, hi):
    """Shard-local parameter vector: all cameras (replicated) + points [lo, hi)."""
    x = np.asarray(x, dtype=np.float64)
    return np.concatenate([x[:6 * n_cam], x[6 * n_cam + 3 * lo:6 * n_cam + 3 * hi]])


class DistributedBA:
    """Collective bundle adjustment over the default ``torch.distributed`` process group."""

    def __init__(self, device=None):
        import torch
        import torch.distributed as td

        if not td.is_initialized():
            raise RuntimeError("torch.distributed is not initialised (launch with torchrun)")
        self.td = td
        self.rank, self.world = td.get_rank(), td.get_world_size()
        if device is None:
            import os
            device = int(os.environ.get("LOCAL_RANK", self.rank))
        torch.cuda.set_device(device)
        self.engine = Engine(device)
        box = [Engine.nccl_unique_id() if self.rank == 0 else None]
        td.broadcast_object_list(box, src=0)
        self.engine.comm_init(box[0], self.rank, self.world)
        self.range = (0, 0)
        self.n_cam = self.n_pts = 0

    def set_local_problem(self, n_cam, n_pts_global, point_range, camera_indices, local_point_indices, points_2d, K):
        """Collective. The caller already holds only its shard (``synthetic.make_problem(point_range=...)``)."""
        lo, hi = point_range
        self.engine.set_problem(n_cam, hi - lo, camera_indices, local_point_indices, points_2d, K)
        self.range, self.n_cam, self.n_pts = (lo, hi), n_cam, n_pts_global
        return self

    def set_problem(self, n_cam, n_pts, camera_indices, point_indices, points_2d, K):
        """Collective. Every rank passes the same full arrays (the drop-in case); each keeps its shard."""
        ranges = plan_shards(point_indices, n_pts, self.world)
        lo, hi = ranges[self.rank]
        ci, pil, uv, _ = local_problem(camera_indices, point_indices, points_2d, lo, hi)
        return self.set_local_problem(n_cam, n_pts, (lo, hi), ci, pil, uv, K)

    def set_fixed_cameras(self, indices=()):
        """Hold these cameras fixed on every rank (``fixed_camera_indices`` of ``bundle_adjustment.py:6``)."""
        self.engine.set_fixed_cameras(indices)
        return self

    def solve_local(self, x_local, **kw):
        """Collective LM; ``x_local`` = cameras + this shard's points. Returns (x_local, Result)."""
        return self.engine.solve(x_local, **kw)

    def solve(self, x0, **kw):
        """Collective LM on the full parameter vector; every rank returns the full solution."""
        lo, hi = self.range
        xl, r = self.solve_local(split_x(x0, self.n_cam, lo, hi), **kw)
        parts = [None] * self.world
        self.td.all_gather_object(parts, (lo, hi, xl[6 * self.n_cam:]))
        x = np.empty(6 * self.n_cam + 3 * self.n_pts)
        x[:6 * self.n_cam] = xl[:6 * self.n_cam]
        for plo, phi, xp in parts:
            x[6 * self.n_cam + 3 * plo:6 * self.n_cam + 3 * phi] = xp
        return BAResult(x=x, cost=r.cost, optimality=r.optimality, nfev=int(r.nfev), njev=int(r.njev),
                        status=int(r.status), message=MESSAGES.get(int(r.status), ""), success=int(r.status) > 0,
                        rms=r.rms, n_iter=int(r.n_iter), initial_cost=r.initial_cost, pcg_iters=int(r.pcg_iters),
                        solve_ms=r.solve_ms)
