"""Dense multi-CTA Cholesky (k_chol_mc, FP64 tensor cores) against the PCG paths: time to ftol = 1e-10 and per-solve kernel
time on chains of 128 .. 1000 cameras (config C = 1000 / 500k / 2M).  python tools/exp_dense.py [quick]"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ba
eng = b200ba.Engine(0)
shapes = [(128, 12800, 51200), (256, 25600, 102400), (512, 51200, 204800), (1000, 500000, 2000000)]
if len(sys.argv) > 1 and sys.argv[1] == "quick":
    shapes = shapes[-1:]
for n_cam, n_pts, n_obs in shapes:
    pb = b200ba.synthetic.make_problem(n_cam, n_pts, n_obs, seed=0)
    eng.set_problem(*pb.args())
    row = {}
    for solver in ("cholesky", "pcg"):
        eng.solve(pb.x0, ftol=1e-10, solver=solver, max_nfev=4)          # warm-up (allocations, attributes)
        x, r = eng.solve(pb.x0, ftol=1e-10, solver=solver)
        row[solver] = (r.solve_ms, r.n_iter, r.cost, r.status, r.pcg_iters)
    eng.set_profiling(True)
    eng.reset_stats()
    eng.solve(pb.x0, ftol=1e-10, solver="cholesky", fixed_iters=3)
    st = eng.stats()
    eng.set_profiling(False)
    n = 6 * n_cam
    ms = st["ms_chol_mc"] / 3
    flops = n ** 3 / 3.0
    print(f"C={n_cam} n={n}: cholesky {row['cholesky'][0]:.2f} ms / {row['cholesky'][1]} LM it (cost {row['cholesky'][2]:.6f}, status {row['cholesky'][3]}) | "
          f"pcg {row['pcg'][0]:.2f} ms / {row['pcg'][1]} LM it / {row['pcg'][4]} PCG it (cost {row['pcg'][2]:.6f}) | "
          f"k_chol_mc {ms:.3f} ms per solve = {flops / (ms * 1e-3) / 1e12 if ms > 0 else 0:.2f} TFLOP/s (n^3/3)", flush=True)
