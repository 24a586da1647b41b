"""Linearisation step + exchange time at N GPUs (torchrun), for NCCL environment experiments.
usage: torchrun --nproc-per-node N tools/xchg_bench.py [D]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as td  # noqa: E402

import b200ba  # noqa: E402
from b200ba import dist as bd  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "D"
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
td.init_process_group("nccl", device_id=torch.device("cuda", lr))
n_cam, n_pts, n_obs = b200ba.synthetic.CONFIGS[wl]
deg = b200ba.synthetic.point_degrees(n_pts, n_obs, min(16, n_cam), 0)
lo, hi = bd.plan_shards_from_degrees(deg, world)[rank]
pb = b200ba.synthetic.make_problem(n_cam, n_pts, n_obs, seed=0, point_range=(lo, hi))
dba = bd.DistributedBA(lr)
eng = dba.engine
eng.set_problem(n_cam, hi - lo, pb.camera_indices, pb.point_indices - lo, pb.points_2d, pb.K)
x = torch.from_numpy(pb.x0).cuda()
for _ in range(5):
    eng.linearize_dev(x.data_ptr(), 1e-4)
torch.cuda.synchronize(); td.barrier(); torch.cuda.synchronize()
t = time.perf_counter()
K = 50
for _ in range(K):
    eng.linearize_dev(x.data_ptr(), 1e-4)
torch.cuda.synchronize(); td.barrier()
dt = (time.perf_counter() - t) / K * 1e3
eng.set_profiling(True); eng.reset_stats()
for _ in range(10):
    eng.linearize_dev(x.data_ptr(), 1e-4)
st = eng.stats()
if rank == 0:
    print(f"[{world} GPUs] NCCL_ALGO={os.environ.get('NCCL_ALGO')} NCCL_PROTO={os.environ.get('NCCL_PROTO')} "
          f"step {dt:.3f} ms  lin {st['ms_lin_fused'] / 10:.3f} reduce {st['ms_slot_reduce'] / 10:.3f} exchange {st['ms_allreduce'] / 10:.3f}",
          flush=True)
td.destroy_process_group()
