"""Per-phase device milliseconds of LM iterations (profiling mode): python tools/lm_phases.py [C|D] [iters]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ba  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "D"
it = int(sys.argv[2]) if len(sys.argv) > 2 else 10
pb = b200ba.synthetic.make_config(wl)
eng = b200ba.Engine(0)
eng.set_problem(*pb.args())
eng.solve(pb.x0, fixed_iters=3, solver="pcg")
eng.set_profiling(True)
eng.reset_stats()
x, r = eng.solve(pb.x0, fixed_iters=it, solver="pcg")
st = eng.stats()
print(wl, {k: round(st[k] / it, 3) for k in ("ms_lin_fused", "ms_slot_reduce", "ms_camera_solve", "ms_backsub_trial")},
      "pcg iters", r.pcg_iters, "tl builds", st["n_tl_builds"])
