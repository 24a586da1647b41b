"""Fixed-iteration LM solve for an ncu launch list (gpu__time_duration per kernel): python tools/tl_launches.py [C|D] [iters]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ba  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "D"
pb = b200ba.synthetic.make_config(wl)
eng = b200ba.Engine(0)
eng.set_problem(*pb.args())
x, r = eng.solve(pb.x0, ftol=1e-10, solver="pcg", fixed_iters=int(sys.argv[2]) if len(sys.argv) > 2 else 10)
print(r.solve_ms, r.pcg_iters)
