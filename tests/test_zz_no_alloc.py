"""GPU: the allocation-free residual entry point (b200ba_residuals_noalloc, no reference counterpart -- added for callers
that cannot call cudaMalloc) against the existing b200ba_residuals, through the C ABI, two calls in a row.

Collected last on purpose.  Uses its own handle so it passes alone and after the others."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_noalloc_residuals_match_the_allocating_entry_twice_in_a_row():
    import torch

    import b200ba
    eng = b200ba.Engine(0)
    try:
        dev = torch.device("cuda", 0)
        # two different problems back to back: the second call must not see anything of the first in the shared scratch
        for n_cam, n_pts, n_obs, seed in ((40, 3000, 12000, 21), (9, 500, 2100, 22)):
            pb = b200ba.synthetic.make_problem(n_cam, n_pts, n_obs, seed=seed, window=min(8, n_cam))
            rng = np.random.default_rng(seed)
            perm = rng.permutation(pb.n_obs)                       # caller's order need not be point-major
            ci, pi, uv = pb.camera_indices[perm], pb.point_indices[perm], pb.points_2d[perm]
            eng.set_problem(n_cam, n_pts, ci, pi, uv, pb.K)
            r_ref, cost_ref = eng.residuals(pb.x0, want_cost=True)
            ci_d = torch.from_numpy(ci.astype(np.int32)).to(dev)
            pi_d = torch.from_numpy(pi.astype(np.int32)).to(dev)
            uv_d = torch.from_numpy(np.ascontiguousarray(uv, dtype=np.float64)).to(dev)
            x_d = torch.from_numpy(pb.x0).to(dev)
            r_d = torch.full((2 * pb.n_obs,), float("nan"), dtype=torch.float64, device=dev)
            c_d = torch.full((1,), float("nan"), dtype=torch.float64, device=dev)
            torch.cuda.synchronize()
            for _ in range(2):
                eng.residuals_noalloc(n_cam, n_pts, pb.n_obs, ci_d.data_ptr(), pi_d.data_ptr(), uv_d.data_ptr(), pb.K,
                                      x_d.data_ptr(), r_d.data_ptr(), c_d.data_ptr())
                r = r_d.cpu().numpy()
                assert np.abs(r - r_ref).max() <= 1e-9 * np.abs(r_ref).max()
                assert abs(float(c_d.item()) - cost_ref) <= 1e-12 * cost_ref
                r_d.fill_(float("nan"))
                c_d.fill_(float("nan"))
                torch.cuda.synchronize()
        # beyond the stated maximum: refused with a clear error, nothing launched
        with pytest.raises(ValueError, match="4096"):
            eng.residuals_noalloc(4097, 10, 0, 0, 0, 0, np.eye(3), x_d.data_ptr())
        # cost-only and residual-only calls
        eng.residuals_noalloc(n_cam, n_pts, pb.n_obs, ci_d.data_ptr(), pi_d.data_ptr(), uv_d.data_ptr(), pb.K, x_d.data_ptr(),
                              0, c_d.data_ptr())
        assert abs(float(c_d.item()) - cost_ref) <= 1e-12 * cost_ref
    finally:
        eng.close()
