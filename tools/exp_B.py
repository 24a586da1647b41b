import sys; import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ba
pb=b200ba.synthetic.make_config('B'); eng=b200ba.Engine(0); eng.set_problem(*pb.args())
for solver in ('cholesky','pcg'):
    for kw in (dict(fixed_iters=10), dict()):
        eng.solve(pb.x0, ftol=1e-10, solver=solver, **kw)
        x,r=eng.solve(pb.x0, ftol=1e-10, solver=solver, **kw)
        print(solver, kw, 'cost %.9f status %d iters %d pcg %d solve_ms %.3f'%(r.cost,r.status,r.n_iter,r.pcg_iters,r.solve_ms), flush=True)
