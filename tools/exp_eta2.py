"""Forcing term of the inexact reduced-camera solve against LM iterations / time to ftol at config D."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ba
pb = b200ba.synthetic.make_config(sys.argv[1] if len(sys.argv) > 1 else "D")
eng = b200ba.Engine(0)
eng.set_problem(*pb.args())
for rtol in (0.0, 0.1, 0.03, 0.01, 0.003, 0.001):
    x, r = eng.solve(pb.x0, ftol=1e-10, solver="pcg", pcg_rtol=rtol, pcg_max_iter=2000)
    print(f"pcg_rtol {rtol or 'auto':>6}: {r.solve_ms:8.1f} ms  {r.n_iter:3d} LM it  {r.pcg_iters:6d} PCG it  cost {r.cost:.6f} status {r.status}", flush=True)
