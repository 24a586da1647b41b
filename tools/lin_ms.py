"""Device time of the linearisation kernels on one workload (events around each kernel): tools/ab_variants.sh runs this
once per library variant (B200BA_LIB)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200ba
wl = sys.argv[1] if len(sys.argv) > 1 else "D"
n_cam, n_pts, n_obs = b200ba.synthetic.CONFIGS[wl]
pb = b200ba.synthetic.make_problem(n_cam, n_pts, n_obs, seed=0)
eng = b200ba.Engine(0)
eng.set_problem(*pb.args())
x = torch.from_numpy(pb.x0).cuda()
for _ in range(5):
    eng.linearize_dev(x.data_ptr(), 1e-4)
eng.set_profiling(True)
eng.reset_stats()
reps = 20
for _ in range(reps):
    eng.linearize_dev(x.data_ptr(), 1e-4)
st = eng.stats()
eng.set_profiling(False)
ne = eng.normal_equations(pb.x0, 1e-4, want_S=False)
print(f"{os.environ.get('B200BA_LIB', 'default'):50s} {wl} lin_fused {st['ms_lin_fused'] / reps:.4f} ms  slot_reduce {st['ms_slot_reduce'] / reps:.4f} ms  "
      f"cost {ne['cost']:.10e} |b| {np.abs(ne['b']).sum():.12e}")
