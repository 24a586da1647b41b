import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ba
wl = sys.argv[1]
pb = b200ba.synthetic.make_config(wl)
eng = b200ba.Engine(0)
eng.set_problem(*pb.args())
for eta in (0.0, 0.3, 0.1):
    x, r = eng.solve(pb.x0, ftol=1e-10, solver="pcg", pcg_rtol=eta, max_nfev=100)
    print(f"{wl} pcg_rtol {eta or 'auto'}: cost {r.cost:.6f} status {r.status} lm_iters {r.n_iter} pcg {r.pcg_iters} solve_ms {r.solve_ms:.1f}", flush=True)
