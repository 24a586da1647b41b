"""CPU, world_size 2 over gloo: the sharding the multi-GPU path relies on -- per-shard point elimination followed by a
SUM all-reduce of [S | b | cost] equals the unsharded reduced camera system."""
import os
import sys

import numpy as np
import torch.multiprocessing as mp

from conftest import ROOT


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import torch
    import torch.distributed as td
    import ba_oracle as O
    import b200ba
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    td.init_process_group("gloo", rank=rank, world_size=world)
    pb = b200ba.synthetic.make_problem(7, 120, 480, seed=5, window=6)       # every rank holds the full problem
    lo, hi = b200ba.dist.plan_shards(pb.point_indices, pb.n_pts, world)[rank]
    ci, pil, uv, _ = b200ba.dist.local_problem(pb.camera_indices, pb.point_indices, pb.points_2d, lo, hi)
    xl = b200ba.dist.split_x(pb.x0, pb.n_cam, lo, hi)
    lam = 1e-3
    s = O.schur_system(xl, pb.n_cam, hi - lo, ci, pil, uv, pb.K, lam)
    packed = torch.from_numpy(np.concatenate([s["S"].ravel(), s["b"], [s["cost"]]]))
    td.all_reduce(packed, op=td.ReduceOp.SUM)
    # same generator, shard-only: bit-identical to the slice of the full problem
    sh = b200ba.synthetic.make_problem(7, 120, 480, seed=5, window=6, point_range=(lo, hi))
    same = bool(np.array_equal(sh.points_2d, uv) and np.array_equal(sh.x0, xl))
    if rank == 0:
        ref = O.schur_system(pb.x0, *pb.args(), lam)
        n = 6 * pb.n_cam
        S = packed[:n * n].numpy().reshape(n, n)
        out.put((float(np.abs(S - ref["S"]).max() / np.abs(ref["S"]).max()),
                 float(np.abs(packed[n * n:n * n + n].numpy() - ref["b"]).max() / np.abs(ref["b"]).max()),
                 float(abs(packed[-1].item() - ref["cost"]) / ref["cost"]), same))
    td.destroy_process_group()


def test_sharded_schur_allreduce_equals_unsharded():
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    res = out.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    eS, eb, ec, same = res
    assert eS < 1e-12 and eb < 1e-12 and ec < 1e-13 and same
