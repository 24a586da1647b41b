"""Setup / solve wall-time breakdown of the sharded path (torchrun, one rank per GPU): rank 0 prints the phases of
Engine.set_problem (B200BA_SETUP_PROF) on its shard of the workload.

usage: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
           tools/e2e_prof_multi.py [--workload D]
"""
import argparse
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
if rank == 0:
    os.environ.setdefault("B200BA_SETUP_PROF", "1")
import torch  # noqa: E402
import torch.distributed as td  # noqa: E402

import b200ba  # noqa: E402
from b200ba import dist as bd  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="D")
ap.add_argument("--iters", type=int, default=5)
a = ap.parse_args()
torch.cuda.set_device(local)
td.init_process_group("nccl", device_id=torch.device("cuda", local))
n_cam, n_pts, n_obs = b200ba.synthetic.CONFIGS[a.workload]
deg = b200ba.synthetic.point_degrees(n_pts, n_obs, min(16, n_cam), 0)
lo, hi = bd.plan_shards_from_degrees(deg, world)[rank]
pb = b200ba.synthetic.make_problem(n_cam, n_pts, n_obs, seed=0, point_range=(lo, hi))
pin = lambda arr: torch.from_numpy(np.ascontiguousarray(arr)).pin_memory().numpy()
ci, pi, uv, x0 = pin(pb.camera_indices), pin(pb.point_indices - lo), pin(pb.points_2d), pin(pb.x0)
dba = bd.DistributedBA(local)
eng = dba.engine
for rep in range(4):
    torch.cuda.synchronize(); td.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    eng.set_problem(n_cam, hi - lo, ci, pi, uv, pb.K)
    t1 = time.perf_counter()
    x, r = eng.solve(x0, ftol=1e-10, fixed_iters=a.iters, solver="pcg")
    t2 = time.perf_counter()
    print(f"rank {rank} rep {rep}: set_problem {1e3 * (t1 - t0):.1f} ms, solve {1e3 * (t2 - t1):.1f} ms "
          f"(device solve_ms {r.solve_ms:.1f}, njev {r.njev})", file=sys.stderr, flush=True)
td.destroy_process_group()
