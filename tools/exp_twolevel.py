"""Two-level PCG preconditioner vs block-Jacobi: LM iterations with a fixed count and to convergence (run on the GPU box).

usage: python tools/exp_twolevel.py [C|D] [fixed_iters]      env B200BA_TWOLEVEL=0|1|2, B200BA_AGG=<cameras per aggregate>
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ba  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "C"
fixed = int(sys.argv[2]) if len(sys.argv) > 2 else 10
pb = b200ba.synthetic.make_config(wl)
eng = b200ba.Engine(0)
eng.set_problem(*pb.args())
tag = f"TWOLEVEL={os.environ.get('B200BA_TWOLEVEL', '1')} AGG={os.environ.get('B200BA_AGG', 'auto')}"
for kw in (dict(fixed_iters=fixed), dict()):
    eng.reset_stats()
    t = time.time()
    x, r = eng.solve(pb.x0, ftol=1e-10, solver="pcg", verbose=2 if kw else 0, max_nfev=60, **kw)
    st = eng.stats()
    print(f"[{tag}] {wl} {kw or 'converge'}: cost {r.cost:.9f} status {r.status} lm_iters {r.n_iter} pcg_total {r.pcg_iters} "
          f"solve_ms {r.solve_ms:.1f} ({r.solve_ms / max(r.n_iter, 1):.2f} ms/iter) tl_builds {st['n_tl_builds']} "
          f"wall {time.time() - t:.2f}s opt {r.optimality:.2e}", flush=True)
    pr = eng.debug_fetch("tl_prof", 8)
    if len(pr) and pr[6] > 0:
        names = ("local", "sync", "coarse", "sync", "spmv", "reduce")
        print("   k_pcg2 us/iteration: " + ", ".join(f"{n} {pr[k] / pr[6] * 1e-3:.1f}" for k, n in enumerate(names)) +
              f"  (total {pr[:6].sum() / pr[6] * 1e-3:.1f}, {int(pr[6])} iterations since set_problem)", flush=True)
