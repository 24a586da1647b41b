import os, sys
sys.path.insert(0, "/root/repo")
import b200ba
eng = b200ba.Engine(0)
for shape in ((49, 7776, 31843), (24, 2400, 9600), (64, 6400, 25600), (9, 1400, 3000)):
    pb = b200ba.synthetic.make_problem(*shape, seed=0, window=min(16, shape[0]))
    eng.set_problem(*pb.args())
    for mc in ("0", "1"):
        os.environ["B200BA_CHOL_MC"] = mc
        for solver in ("cholesky", "pcg"):
            eng.solve(pb.x0, ftol=1e-10, solver=solver)
            x, r = eng.solve(pb.x0, ftol=1e-10, solver=solver)
            print(shape, "mc", mc, solver, f"{r.solve_ms:.3f} ms {r.n_iter} it {r.solve_ms / r.n_iter:.3f} ms/it cost {r.cost:.6f}", flush=True)
