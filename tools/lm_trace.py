import sys, os
sys.path.insert(0, "/root/repo")
import b200ba
pb = b200ba.synthetic.make_config(sys.argv[1] if len(sys.argv) > 1 else "C")
eng = b200ba.Engine(0)
eng.set_problem(*pb.args())
x, r = eng.solve(pb.x0, ftol=1e-10, solver="pcg", verbose=2)
print(r.n_iter, r.pcg_iters, r.solve_ms)
