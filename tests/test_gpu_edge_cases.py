"""GPU: edge cases of the BA core through the C ABI -- empty and ragged inputs, rank-deficient points, unobserved cameras,
bad indices, non-finite parameters.  The reference has no tests of its own for this path; the expected
behaviour is what its numpy / scipy code does on the same input (shapes, errors) or, where that is an exception deep inside
scipy, a clean error code here."""
import numpy as np
import pytest

import ba_oracle as O

pytestmark = pytest.mark.gpu


def _small(seed=2):
    import b200ba
    return b200ba.synthetic.make_problem(7, 120, 500, seed=seed, window=7)


def test_empty_problem_and_zero_observations(engine):
    """No observations: residual vector of length 0, cost 0, the solve returns x0 unchanged with a success status (scipy on
    an empty residual vector reports cost 0 as well)."""
    K = np.eye(3)
    engine.set_problem(2, 3, np.zeros(0, dtype=np.int64), np.zeros(0, dtype=np.int64), np.zeros((0, 2)), K)
    x0 = np.arange(12 + 9, dtype=np.float64)
    r = engine.residuals(x0)
    assert r.shape == (0,) and engine.cost(x0) == 0.0
    x, res = engine.solve(x0, ftol=1e-10)
    assert np.array_equal(x, x0) and res.cost == 0.0 and res.status > 0
    engine.set_problem(0, 0, np.zeros(0, dtype=np.int64), np.zeros(0, dtype=np.int64), np.zeros((0, 2)), K)
    assert engine.residuals(np.zeros(0)).shape == (0,)


def test_single_observation_points_and_unobserved_camera(engine):
    """Points seen once have a rank-2 3x3 block (they are frozen, like the oracle does), a camera without observations has a
    zero block (its parameters must come back bit-identical); the other unknowns still converge to the oracle LM cost."""
    pb = _small()
    keep = np.ones(pb.n_obs, dtype=bool)
    first = np.flatnonzero(np.r_[True, np.diff(pb.point_indices) != 0])
    for p in range(0, 20):                                   # points 0..19 keep only their first observation
        keep[first[p] + 1:first[p + 1]] = False
    keep &= pb.camera_indices != 3                           # camera 3 observes nothing
    ci, pi, uv = pb.camera_indices[keep], pb.point_indices[keep], pb.points_2d[keep]
    args = (pb.n_cam, pb.n_pts, ci, pi, uv, pb.K)
    engine.set_problem(*args)
    r = engine.residuals(pb.x0)
    assert np.abs(r - O.residuals(pb.x0, *args)).max() <= 1e-9 * np.abs(r).max()
    for solver in ("cholesky", "pcg"):
        x, res = engine.solve(pb.x0, ftol=1e-10, solver=solver)
        assert res.status > 0 and np.isfinite(res.cost)
        assert np.array_equal(x[18:24], pb.x0[18:24])                    # the unobserved camera did not move
        chk = 0.5 * float(np.sum(O.residuals(x, *args) ** 2))
        assert abs(chk - res.cost) <= 1e-9 * res.cost
        # the oracle's dense solve is singular with an unobserved camera: there it is held fixed explicitly
        ref = O.lm_solve(pb.x0, *args, ftol=1e-10, fixed_cameras=(3,))
        assert abs(res.cost - ref["cost"]) <= 1e-6 * ref["cost"]


def test_bad_inputs_raise_clean_errors(engine):
    import b200ba
    pb = _small()
    args = list(pb.args())
    bad = pb.camera_indices.copy()
    bad[5] = pb.n_cam                                        # camera index out of range
    with pytest.raises(ValueError):
        engine.set_problem(pb.n_cam, pb.n_pts, bad, pb.point_indices, pb.points_2d, pb.K)
    badp = pb.point_indices.copy()
    badp[7] = -1
    with pytest.raises(ValueError):
        engine.set_problem(pb.n_cam, pb.n_pts, pb.camera_indices, badp, pb.points_2d, pb.K)
    with pytest.raises(ValueError):                          # ragged: index arrays of different length
        engine.set_problem(pb.n_cam, pb.n_pts, pb.camera_indices[:-1], pb.point_indices, pb.points_2d, pb.K)
    with pytest.raises(ValueError):
        engine.set_problem(pb.n_cam, pb.n_pts, pb.camera_indices, pb.point_indices, pb.points_2d[:-1], pb.K)
    with pytest.raises(ValueError):
        engine.set_problem(pb.n_cam, pb.n_pts, pb.camera_indices, pb.point_indices, pb.points_2d, np.eye(4))
    # a failed set_problem leaves the handle without a problem, not with a stale one
    with pytest.raises(b200ba.BAError):
        engine.residuals(pb.x0)
    engine.set_problem(*args)
    with pytest.raises(ValueError):                          # parameter vector of the wrong length
        engine.residuals(pb.x0[:-1])
    x_nan = pb.x0.copy()
    x_nan[40] = np.nan
    with pytest.raises(b200ba.BAError):                      # non-finite cost at the start: NOT_FINITE, not a hang or garbage
        engine.solve(x_nan, ftol=1e-10)
    x, res = engine.solve(pb.x0, ftol=1e-10)                 # and the handle is usable afterwards
    assert res.status > 0


def test_int64_pixels_duplicates_and_max_nfev(engine):
    """int64 pixel arrays (graph.py:112-113), an observation listed twice (counts twice, as in the reference's residual
    vector), and the max_nfev cap reported as status 0 like scipy."""
    pb = _small(seed=4)
    uv = pb.points_2d.astype(np.int64)
    ci = np.r_[pb.camera_indices, pb.camera_indices[:3]]
    pi = np.r_[pb.point_indices, pb.point_indices[:3]]
    uv2 = np.r_[uv, uv[:3]]
    args = (pb.n_cam, pb.n_pts, ci, pi, uv2, pb.K)
    engine.set_problem(*args)
    r = engine.residuals(pb.x0)
    assert np.abs(r - O.residuals(pb.x0, *args)).max() <= 1e-9 * np.abs(r).max()
    ne = engine.normal_equations(pb.x0, 1e-3)
    ref = O.schur_system(pb.x0, *args, 1e-3)
    assert np.abs(ne["S"] - ref["S"]).max() <= 1e-9 * np.abs(ref["S"]).max()
    x, res = engine.solve(pb.x0, ftol=1e-15, xtol=1e-15, gtol=1e-15, max_nfev=3)
    assert res.status == 0 and res.nfev <= 4
