"""Worker of tests/test_gpu_multi.py: launched by torch.distributed.run, one process per GPU.

Every rank builds the same seeded problem, keeps its point shard, and checks (a) the sharded reduced camera system
against the unsharded one computed on this rank's own GPU by a second, single-rank engine, (b) the sharded LM against the
unsharded LM.  B200BA_XCHG selects the exchange (0 = all-reduce, 1 = sparse all-gather exchange).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as td  # noqa: E402

sys.path.insert(0, os.path.join(ROOT, "oracle"))
import ba_oracle as O  # noqa: E402
import b200ba  # noqa: E402
from b200ba import dist as bd  # noqa: E402


def main():
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    td.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = td.get_rank(), td.get_world_size()
    n_cam, n_pts, n_obs = (int(v) for v in sys.argv[1:4])
    pb = b200ba.synthetic.make_problem(n_cam, n_pts, n_obs, seed=3)
    dba = bd.DistributedBA(local)
    dba.set_problem(n_cam, n_pts, pb.camera_indices, pb.point_indices, pb.points_2d, pb.K)
    lo, hi = dba.range
    lam = 1e-3
    ne = dba.engine.normal_equations(bd.split_x(pb.x0, n_cam, lo, hi), lam)
    solver = "pcg" if n_cam > 64 else "cholesky"
    res = dba.solve(pb.x0, ftol=1e-10, max_nfev=30, solver=solver)

    one = b200ba.Engine(local)
    one.set_problem(n_cam, n_pts, pb.camera_indices, pb.point_indices, pb.points_2d, pb.K)
    ref = one.normal_equations(pb.x0, lam)
    x1, r1 = one.solve(pb.x0, ftol=1e-10, max_nfev=30, solver=solver)

    sc = np.abs(ref["S"]).max()
    err_S = np.abs(ne["S"] - ref["S"]).max() / sc
    err_b = np.abs(ne["b"] - ref["b"]).max() / np.abs(ref["b"]).max()
    err_c = abs(ne["cost"] - ref["cost"]) / ref["cost"]
    err_lm = abs(res.cost - r1.cost) / r1.cost
    # every rank must hold the same reduced system bit for bit (replicated camera solve stays in lock-step)
    t = torch.from_numpy(np.concatenate([ne["S"].ravel(), ne["b"]])).cuda()
    t0 = t.clone()
    td.broadcast(t0, src=0)
    same = bool(torch.equal(t, t0))
    # and against the ORACLE (not only sharded-CUDA vs unsharded-CUDA): cost, camera gradient and b from the vectorised
    # numpy blocks; the dense reduced system itself on the small shape (the oracle's Schur loop is per point)
    nb = O.normal_blocks(pb.x0, *pb.args())
    err_oc = abs(ne["cost"] - nb["cost"]) / nb["cost"]
    err_og = np.abs(ne["grad"][:6 * n_cam] - nb["gc"].ravel()).max() / np.abs(nb["gc"]).max()
    err_oS = 0.0
    if n_cam <= 64:
        so = O.schur_system(pb.x0, *pb.args(), lam)
        err_oS = max(np.abs(ne["S"] - so["S"]).max() / np.abs(so["S"]).max(), np.abs(ne["b"] - so["b"]).max() / np.abs(so["b"]).max())
    chk = 0.5 * float(np.sum(O.residuals(res.x, *pb.args()) ** 2))
    err_ox = abs(chk - res.cost) / res.cost
    ok = err_S < 1e-11 and err_b < 1e-11 and err_c < 1e-13 and err_lm < 1e-8 and same
    ok = ok and err_oc < 1e-12 and err_og < 1e-9 and err_oS < 1e-9 and err_ox < 1e-9
    print(f"rank {rank}/{world} xchg={os.environ.get('B200BA_XCHG', 'auto')} S {err_S:.2e} b {err_b:.2e} cost {err_c:.2e} "
          f"lm {err_lm:.2e} ({res.cost:.6f} vs {r1.cost:.6f}) identical={same} | vs oracle: cost {err_oc:.2e} g_c {err_og:.2e} "
          f"S,b {err_oS:.2e} cost(x*) {err_ox:.2e} {'OK' if ok else 'FAIL'}", flush=True)
    td.barrier()
    td.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
