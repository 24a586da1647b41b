"""GPU parity of the robust loss (scipy's loss= / f_scale= keywords of the call at sfm.py:266).

The oracle's restatement of scipy's rho / scaling is pinned against scipy's own functions on the CPU
(tests/test_oracle.py::test_robust_scaling_matches_scipy); goldens: oracle/make_golden.py robust."""
import numpy as np
import pytest

import ba_oracle as O
from conftest import golden_args, load_golden

pytestmark = pytest.mark.gpu
LOSSES = ["soft_l1", "huber", "cauchy", "arctan"]


def _outlier_problem(seed=5):
    import b200ba
    pb = b200ba.synthetic.make_problem(9, 300, 1300, seed=seed, window=8)
    rng = np.random.default_rng(seed)
    uv = pb.points_2d.copy()
    bad = rng.choice(pb.n_obs, size=pb.n_obs // 20, replace=False)
    uv[bad] += rng.uniform(-80, 80, size=(len(bad), 2))
    return pb, (pb.n_cam, pb.n_pts, pb.camera_indices, pb.point_indices, uv, pb.K)


@pytest.mark.parametrize("loss", LOSSES)
def test_scaled_normal_equations_match_the_oracle(engine, loss):
    """S, b, gradient and cost of one linearisation under the loss vs the dense oracle (scipy's scaling of J and f)."""
    pb, args = _outlier_problem()
    engine.set_problem(*args)
    engine.set_loss(loss, 1.5)
    lam = 1e-3
    ne = engine.normal_equations(pb.x0, lam)
    ref = O.schur_system(pb.x0, *args, lam, loss=loss, f_scale=1.5)
    g = np.concatenate([ref["gc"].ravel(), ref["gp"].ravel()])
    assert abs(ne["cost"] - ref["cost"]) <= 1e-12 * ref["cost"]
    assert np.abs(ne["S"] - ref["S"]).max() <= 1e-9 * np.abs(ref["S"]).max()
    assert np.abs(ne["b"] - ref["b"]).max() <= 1e-9 * np.abs(ref["b"]).max()
    assert np.abs(ne["grad"] - g).max() <= 1e-9 * np.abs(g).max()
    # residuals / Jacobian blocks stay the plain ones (scipy's fun and jac)
    r = engine.residuals(pb.x0)
    assert np.abs(r - O.residuals(pb.x0, *args)).max() <= 1e-9 * np.abs(r).max()
    d, pred, newc = engine.lm_step(pb.x0, lam, solver="cholesky")
    d_ref, pred_ref, _ = O.lm_step(pb.x0, *args, lam, loss=loss, f_scale=1.5)
    assert np.abs(d - d_ref).max() <= 1e-6 * np.abs(d_ref).max()
    assert abs(pred - pred_ref) <= 1e-6 * abs(pred_ref)
    new_ref = O.robust_cost(O.residuals(pb.x0 + d_ref, *args), loss, 1.5)
    assert abs(newc - new_ref) <= 1e-6 * new_ref
    engine.set_loss("linear")


@pytest.mark.parametrize("k", [0, 1, 2])
@pytest.mark.parametrize("loss", ["soft_l1", "huber", "cauchy"])
def test_robust_converged_cost_vs_scipy(engine, k, loss):
    """SceauxCastle calls 0-2 with 5 % outliers, started at the converged linear-loss solution (a robust cost is not
    convex: from the raw 292 px starts scipy and any LM end in different minima; goldens: oracle/make_golden.py robust).
    Through the drop-in least_squares: the robust cost is the oracle's cost at the returned x (1e-9), never above scipy's
    (one-sided 1e-6; scipy's TRF crawls on these gauge-free problems and mostly stops on its evaluation cap), two-sided
    1e-6 where scipy itself converged, and within 1e-5 of the oracle LM run with the same loss."""
    import b200ba
    g = load_golden("robust_scipy.npz")
    d = load_golden(f"sceaux_ba_{k}.npz")
    n_cam, n_pts, ci, pi, _, K = golden_args(d)
    uv, x_start, f_scale = g[f"uv_{k}"], g[f"x_start_{k}"], float(g["f_scale"])
    mask = b200ba.create_sparsity_matrix(n_cam, n_pts, len(ci), ci, pi)
    res = b200ba.least_squares(b200ba.compute_residuals, x_start, jac_sparsity=mask, verbose=0, x_scale='jac', ftol=1e-10,
                               method='trf', loss=loss, f_scale=f_scale, args=(n_cam, n_pts, ci, pi, uv, K))
    assert res.status > 0
    chk = O.robust_cost(O.residuals(res.x, n_cam, n_pts, ci, pi, uv, K), loss, f_scale)
    assert abs(chk - res.cost) <= 1e-9 * res.cost
    sc, lm = float(g[f"cost_{k}_{loss}"]), float(g[f"lm_cost_{k}_{loss}"])
    assert res.cost <= sc * (1 + 1e-6), (res.cost, sc, lm)
    if int(g[f"status_{k}_{loss}"]) > 0 and abs(sc - lm) <= 1e-6 * lm:
        assert abs(res.cost - sc) <= 1e-6 * sc
    assert abs(res.cost - lm) <= 1e-5 * lm, (res.cost, lm)


def test_loss_resets_with_the_problem_and_bad_arguments(engine):
    pb, args = _outlier_problem()
    engine.set_problem(*args)
    engine.set_loss("cauchy", 3.0)
    c_rob = engine.normal_equations(pb.x0, 0.0, want_S=False)["cost"]
    engine.set_problem(*args)
    c_lin = engine.normal_equations(pb.x0, 0.0, want_S=False)["cost"]
    r = O.residuals(pb.x0, *args)
    assert abs(c_lin - 0.5 * float(r @ r)) <= 1e-12 * c_lin
    assert abs(c_rob - O.robust_cost(r, "cauchy", 3.0)) <= 1e-12 * c_rob
    with pytest.raises(ValueError):
        engine.set_loss("tukey")
    with pytest.raises(ValueError):
        engine.set_loss("huber", -1.0)
