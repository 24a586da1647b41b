"""Per-item cycle breakdown of k_lin_flow (build with `make EXTRA=-DFLOW_PROF=1`)."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200ba
wl = sys.argv[1] if len(sys.argv) > 1 else "C"
n_cam, n_pts, n_obs = b200ba.synthetic.CONFIGS[wl]
pb = b200ba.synthetic.make_problem(n_cam, n_pts, n_obs, seed=0)
eng = b200ba.Engine(0)
eng.set_problem(*pb.args())
x = torch.from_numpy(pb.x0).cuda()
for _ in range(3):
    eng.linearize_dev(x.data_ptr(), 1e-4)
nsm = 148
p = eng.debug_fetch("flow_prof", 16 * 16 * nsm, dtype=np.int64).reshape(-1, 16)
names = ["A", "C", "B", "wait", "queue", "items", "total", "K"]
print("mean per warp:", {n: float(p[:, i].mean()) for i, n in enumerate(names)})
tot = p[:, 6].mean()
print("fractions of total:", {n: round(float(p[:, i].mean() / tot), 3) for i, n in enumerate(names[:5])})
print("A sub-phases (fraction of A): load+project %.3f scan %.3f factor %.3f" % tuple(float(p[:, i].mean() / p[:, 0].mean()) for i in (8, 9, 10)))
print("per item cycles: A+C+B", float((p[:, 0] + p[:, 1] + p[:, 2]).sum() / p[:, 5].sum()))
st = eng.stats()
print(st["n_tiles"], st["n_pair_slots"], st["n_self_slots"])
