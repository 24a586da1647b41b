"""Scratch experiment: LM convergence vs PCG tolerance on a synthetic config (run on the GPU box)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import b200ba

wl = sys.argv[1] if len(sys.argv) > 1 else "C"
pb = b200ba.synthetic.make_config(wl)
eng = b200ba.Engine(0)
eng.set_problem(*pb.args())
for rtol in [float(a) for a in sys.argv[2:]] or [1e-6, 1e-3, 1e-2, 1e-1]:
    t = time.time()
    x, r = eng.solve(pb.x0, ftol=1e-10, solver="pcg", pcg_rtol=rtol, pcg_max_iter=2000)
    print(f"{wl} rtol {rtol:g}: cost {r.cost:.9f} status {r.status} lm_iters {r.n_iter} pcg_total {r.pcg_iters} "
          f"solve_ms {r.solve_ms:.1f} wall {time.time()-t:.2f}s opt {r.optimality:.2e}", flush=True)
