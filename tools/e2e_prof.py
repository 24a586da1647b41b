"""Wall-time breakdown of the host-buffer path: Engine.set_problem (setup phases via B200BA_SETUP_PROF) and Engine.solve.

usage: python tools/e2e_prof.py [--workload D] [--iters 5]
"""
import argparse
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("B200BA_SETUP_PROF", "1")
import torch  # noqa: E402

import b200ba  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="D")
ap.add_argument("--iters", type=int, default=5)
ap.add_argument("--pageable", action="store_true", help="hand over plain numpy arrays (what sfm.py does) instead of page-locked ones")
ap.add_argument("--shard", type=int, default=1, help="take the first of this many point shards (one GPU, no exchange)")
a = ap.parse_args()
n_cam, n_pts, n_obs = b200ba.synthetic.CONFIGS[a.workload]
if a.shard > 1:
    from b200ba import dist as bd
    deg = b200ba.synthetic.point_degrees(n_pts, n_obs, min(16, n_cam), 0)
    lo, hi = bd.plan_shards_from_degrees(deg, a.shard)[0]
    pb = b200ba.synthetic.make_problem(n_cam, n_pts, n_obs, seed=0, point_range=(lo, hi))
    pb.point_indices = pb.point_indices - lo
    n_pts = hi - lo
else:
    pb = b200ba.synthetic.make_problem(n_cam, n_pts, n_obs, seed=0)
pin = (lambda arr: np.ascontiguousarray(arr).copy()) if a.pageable else (lambda arr: torch.from_numpy(np.ascontiguousarray(arr)).pin_memory().numpy())
ci, pi, uv, x0 = pin(pb.camera_indices), pin(pb.point_indices), pin(pb.points_2d), pin(pb.x0)
eng = b200ba.Engine(0)
solver = "pcg" if n_cam > 64 else "cholesky"
for rep in range(3):
    t0 = time.perf_counter()
    eng.set_problem(n_cam, n_pts, ci, pi, uv, pb.K)
    t1 = time.perf_counter()
    x, r = eng.solve(x0, ftol=1e-10, fixed_iters=a.iters, solver=solver)
    t2 = time.perf_counter()
    print(f"rep {rep}: set_problem {1e3 * (t1 - t0):.1f} ms, solve {1e3 * (t2 - t1):.1f} ms (device solve_ms {r.solve_ms:.1f}, "
          f"njev {r.njev})", file=sys.stderr)
