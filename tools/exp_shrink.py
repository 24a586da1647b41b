"""lambda shrink factor on well-predicted steps against LM iterations / time to ftol (B200BA_LAMBDA_SHRINK)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import b200ba
eng = b200ba.Engine(0)
gold = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
for wl in ("B", "C", "D"):
    pb = b200ba.synthetic.make_config(wl)
    eng.set_problem(*pb.args())
    for f in ("0.3333", "0.1"):
        os.environ["B200BA_LAMBDA_SHRINK"] = f
        x, r = eng.solve(pb.x0, ftol=1e-10)
        print(f"{wl} shrink {f:>6}: {r.solve_ms:8.2f} ms {r.n_iter:3d} LM it (nfev {r.nfev}) {r.pcg_iters:5d} PCG it cost {r.cost:.9f} status {r.status}", flush=True)
for k in range(9):
    d = np.load(os.path.join(gold, f"sceaux_ba_{k}.npz"))
    eng.set_problem(int(d["n_cam"]), int(d["n_pts"]), d["ci"], d["pi"], d["uv"], d["K"])
    out = []
    for f in ("0.3333", "0.1"):
        os.environ["B200BA_LAMBDA_SHRINK"] = f
        x, r = eng.solve(d["x0"], ftol=float(d["ftol"]))
        out.append(f"{f}: {r.n_iter} it cost {r.cost:.9f} st {r.status}")
    print(f"sceaux {k}: " + " | ".join(out) + f" | scipy {float(d['scipy_cost']):.9f}", flush=True)
