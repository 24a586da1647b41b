"""GPU parity, stage 3: converged cost / RMS of the LM solve vs scipy (reference kwargs) and the oracle LM."""
import numpy as np
import pytest

import ba_oracle as O
from conftest import golden_args, load_golden

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("k", range(9))
@pytest.mark.parametrize("solver", ["cholesky", "pcg"])
def test_sceaux_converged_cost(engine, k, solver):
    """Frozen SceauxCastle BA calls. Calls 0-7: two-sided 1e-6 vs scipy's converged cost (sfm.py:266-268 kwargs);
    call 8 (scipy stalls): one-sided vs scipy's capped cost and two-sided vs the oracle LM minimum."""
    d = load_golden(f"sceaux_ba_{k}.npz")
    args = golden_args(d)
    engine.set_problem(*args)
    x, res = engine.solve(d["x0"], ftol=float(d["ftol"]), solver=solver)
    assert res.status > 0
    cost_check = 0.5 * float(np.sum(O.residuals(x, *args) ** 2))
    assert abs(cost_check - res.cost) <= 1e-9 * res.cost
    sc, lm = float(d["scipy_cost"]), float(d["lm_cost"])
    assert res.cost <= sc * (1 + 1e-6)
    if int(d["scipy_status"]) > 0:
        assert abs(res.cost - sc) <= 1e-6 * sc
        assert abs(res.rms - float(d["scipy_rms"])) <= 1e-6 * float(d["scipy_rms"])
    assert abs(res.cost - lm) <= 1e-6 * lm


def test_true_reference_scipy_costs(engine):
    """Calls 0-2 were also solved by scipy with the REAL reference residual + sparsity (sceaux_scipy_ref.npz)."""
    ref = load_golden("sceaux_scipy_ref.npz")
    for k in (0, 1, 2):
        d = load_golden(f"sceaux_ba_{k}.npz")
        engine.set_problem(*golden_args(d))
        x, res = engine.solve(d["x0"], ftol=1e-10)
        assert abs(res.cost - float(ref[f"cost_{k}"])) <= 1e-6 * float(ref[f"cost_{k}"])
        assert abs(res.rms - float(ref[f"rms_{k}"])) <= 1e-6 * float(ref[f"rms_{k}"])


def test_least_squares_dropin_surface():
    """scipy-shaped call exactly as written at sfm.py:264-268."""
    import b200ba
    d = load_golden("sceaux_ba_0.npz")
    n_cam, n_pts, ci, pi, uv, K = golden_args(d)
    jac_sparsity = b200ba.create_sparsity_matrix(n_cam, n_pts, len(ci), ci, pi)
    result = b200ba.least_squares(b200ba.compute_residuals, d["x0"], jac_sparsity=jac_sparsity, verbose=0,
                                  x_scale='jac', ftol=1e-10, method='trf', args=(n_cam, n_pts, ci, pi, uv, K))
    assert result.x.shape == d["x0"].shape and result.success
    assert abs(result.cost - float(d["scipy_cost"])) <= 1e-6 * float(d["scipy_cost"])
    assert result.fun.shape == (2 * len(ci),)
    assert abs(0.5 * float(result.fun @ result.fun) - result.cost) <= 1e-9 * result.cost
    assert result.grad.shape == d["x0"].shape
    assert abs(np.abs(result.grad).max() - result.optimality) <= 1e-6 * max(1.0, result.optimality)


@pytest.mark.parametrize("solver", ["cholesky", "pcg"])
def test_config_B_converged_cost(engine, solver):
    """Config B (49 / 7776 / 31843): oracle LM minimum 4951.324381508 (two-sided 1e-6)."""
    import b200ba
    pb = b200ba.synthetic.make_config("B")
    engine.set_problem(*pb.args())
    x, res = engine.solve(pb.x0, ftol=1e-10, solver=solver)
    assert res.status > 0
    ref_cost = 4951.324381508434
    assert abs(res.cost - ref_cost) <= 1e-6 * ref_cost
    cost_check = 0.5 * float(np.sum(O.residuals(x, *pb.args()) ** 2))
    assert abs(cost_check - res.cost) <= 1e-9 * res.cost


def test_inexact_pcg_reaches_the_exact_minimum(engine):
    """Mid-size problem (PCG territory, 200 cameras): the default inexact-Newton forcing term must converge to the same
    cost as a tightly solved reduced system, within the 1e-6 bar of north_star."""
    import b200ba
    pb = b200ba.synthetic.make_problem(200, 20000, 80000, seed=4)
    engine.set_problem(*pb.args())
    x_a, a = engine.solve(pb.x0, ftol=1e-10, solver="pcg")
    x_b, b = engine.solve(pb.x0, ftol=1e-10, solver="pcg", pcg_rtol=1e-10, pcg_max_iter=5000)
    assert a.status > 0 and b.status > 0
    assert abs(a.cost - b.cost) <= 1e-6 * b.cost
    assert a.pcg_iters < b.pcg_iters
    cost_check = 0.5 * float(np.sum(O.residuals(x_a, *pb.args()) ** 2))
    assert abs(cost_check - a.cost) <= 1e-9 * a.cost
