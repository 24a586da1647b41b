"""Where the time of the SceauxCastle replay (BASELINE configs[0]) goes: set_problem vs solve, per call."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ba
gold = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
eng = b200ba.Engine(0)
for rep in range(2):
    for k in range(9):
        d = np.load(os.path.join(gold, f"sceaux_ba_{k}.npz"))
        args = (int(d["n_cam"]), int(d["n_pts"]), d["ci"], d["pi"], d["uv"], d["K"])
        t0 = time.perf_counter()
        key = b200ba.bundle_adjustment._problem_key(*args)
        t1 = time.perf_counter()
        eng.set_problem(*args)
        t2 = time.perf_counter()
        x, r = eng.solve(d["x0"], ftol=float(d["ftol"]))
        t3 = time.perf_counter()
        if rep:
            print(f"call {k}: C {args[0]} N {len(args[2])}: key {1e3 * (t1 - t0):.3f} set_problem {1e3 * (t2 - t1):.3f} solve {1e3 * (t3 - t2):.3f} ms "
                  f"(device {r.solve_ms:.3f} ms, {r.n_iter} LM it -> {r.solve_ms / r.n_iter:.3f} ms/it)", flush=True)
