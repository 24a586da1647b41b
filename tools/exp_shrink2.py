import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ba
eng = b200ba.Engine(0)
for shape, seed in (((300, 30000, 120000), 3), ((128, 12800, 51200), 0), ((256, 25600, 102400), 0), ((400, 40000, 160000), 5), ((600, 60000, 240000), 6), ((160, 6000, 24000), 12), ((200, 20000, 80000), 4)):
    pb = b200ba.synthetic.make_problem(*shape, seed=seed)
    eng.set_problem(*pb.args())
    for f in ("0.3333", "0.1"):
        os.environ["B200BA_LAMBDA_SHRINK"] = f
        x, r = eng.solve(pb.x0, ftol=1e-10, solver="pcg", max_nfev=200)
        print(f"{shape} shrink {f:>6}: {r.solve_ms:8.2f} ms {r.n_iter:3d} LM it (nfev {r.nfev}) {r.pcg_iters:5d} PCG it cost {r.cost:.9f} status {r.status}", flush=True)
