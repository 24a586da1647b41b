"""CPU: the numpy oracle pinned to outputs of the REAL reference (tests/golden, made by oracle/make_golden.py)."""
import os
import sys

import numpy as np
import pytest

import ba_oracle as O
from conftest import ROOT, golden_args, load_golden

FIX = ["ref_residual_small.npz", "ref_residual_int_unsorted.npz", "ref_residual_pnp9.npz", "ref_residual_B2000.npz",
       "ref_residual_angles_skewK.npz"]


@pytest.mark.parametrize("name", FIX)
def test_vectorised_residual_vs_reference(name):
    d = load_golden(name)
    r = O.residuals(d["x"], *golden_args(d))
    ref = d["r_ref"]
    assert np.abs(r - ref).max() <= 1e-11 * np.abs(ref).max()


@pytest.mark.parametrize("name", ["ref_residual_small.npz", "ref_residual_pnp9.npz", "ref_residual_int_unsorted.npz"])
def test_loop_port_vs_reference_bit_exact(name):
    """The per-observation port follows bundle_adjustment.py:20-42 operation by operation."""
    d = load_golden(name)
    r = O.residuals_loop(d["x"], *golden_args(d))
    assert np.array_equal(r, d["r_ref"])


@pytest.mark.parametrize("k", range(9))
def test_residual_on_sceaux_dumps(k):
    d = load_golden(f"sceaux_ba_{k}.npz")
    r = O.residuals(d["x0"], *golden_args(d))
    assert np.abs(r - d["r0_ref"]).max() <= 1e-11 * np.abs(d["r0_ref"]).max()
    assert d["uv"].dtype == np.int64 and np.all(np.diff(d["pi"]) >= 0)


@pytest.mark.parametrize("name", ["ref_residual_small.npz", "ref_residual_pnp9.npz"])
def test_sparsity_vs_reference(name):
    """create_sparsity_matrix (bundle_adjustment.py:6-17): same non-zero pattern, from the oracle and from the product's
    lazy mask."""
    import b200ba
    d = load_golden(name)
    n_cam, n_pts, ci, pi = int(d["n_cam"]), int(d["n_pts"]), d["ci"], d["pi"]
    for S in (O.sparsity(n_cam, n_pts, len(ci), ci, pi),
              b200ba.create_sparsity_matrix(n_cam, n_pts, len(ci), ci, pi).tocsr()):
        S = S.tocoo()
        order = np.lexsort((S.col, S.row))
        assert tuple(S.shape) == tuple(d["sp_shape"])
        assert np.array_equal(S.row[order], d["sp_row"]) and np.array_equal(S.col[order], d["sp_col"])
    assert b200ba.create_sparsity_matrix(n_cam, n_pts, len(ci), ci, pi).nnz == len(d["sp_row"])


def test_analytic_jacobian_vs_central_differences_of_reference_residual_model():
    d = load_golden("ref_residual_angles_skewK.npz")
    n_cam, n_pts, ci, pi, uv, K = golden_args(d)
    x = d["x"]
    J = O.dense_jacobian(x, n_cam, n_pts, ci, pi, K)
    r0 = O.residuals(x, n_cam, n_pts, ci, pi, uv, K)
    ok = np.abs(r0) < 1e4                       # skip rows whose depth is ~0 (ill-conditioned projection)
    rng = np.random.default_rng(0)
    for col in rng.choice(J.shape[1], 40, replace=False):
        h = 1e-6
        e = np.zeros_like(x)
        e[col] = h
        fd = (O.residuals(x + e, n_cam, n_pts, ci, pi, uv, K) - O.residuals(x - e, n_cam, n_pts, ci, pi, uv, K)) / (2 * h)
        scale = max(1.0, np.abs(J[ok, col]).max())
        assert np.abs(fd[ok] - J[ok, col]).max() <= 2e-5 * scale


def test_schur_step_equals_dense_normal_equations():
    import b200ba
    pb = b200ba.synthetic.make_problem(5, 40, 150, seed=1, window=5)
    lam = 1e-3
    delta, pred, s = O.lm_step(pb.x0, *pb.args(), lam)
    J = O.dense_jacobian(pb.x0, pb.n_cam, pb.n_pts, pb.camera_indices, pb.point_indices, pb.K)
    r = O.residuals(pb.x0, *pb.args())
    H = J.T @ J
    Hd = H + lam * np.diag(np.diag(H))
    ref = np.linalg.solve(Hd, -J.T @ r)
    assert np.abs(delta - ref).max() <= 1e-8 * np.abs(ref).max()
    assert abs(pred - (0.5 * r @ r - 0.5 * np.sum((r + J @ ref) ** 2))) <= 1e-8 * abs(pred)


@pytest.mark.parametrize("k", [0, 3])
def test_oracle_lm_reaches_scipy_cost_on_dumps(k):
    d = load_golden(f"sceaux_ba_{k}.npz")
    res = O.lm_solve(d["x0"], *golden_args(d), ftol=1e-10)
    assert abs(res["cost"] - float(d["scipy_cost"])) <= 1e-6 * float(d["scipy_cost"])
    assert abs(res["cost"] - float(d["lm_cost"])) <= 1e-9 * float(d["lm_cost"])


def test_true_reference_scipy_golden_consistent():
    """scipy + the real reference residual (calls 0-2) agrees with scipy + the vectorised twin to 1e-7."""
    ref = load_golden("sceaux_scipy_ref.npz")
    for k in (0, 1, 2):
        d = load_golden(f"sceaux_ba_{k}.npz")
        assert abs(float(ref[f"cost_{k}"]) - float(d["scipy_cost"])) <= 1e-7 * float(d["scipy_cost"])


# ------------------------------------------------------------------ next rows: reprojection error, fixed cameras
def test_reproj_error_oracle_vs_reference():
    """oracle.reproj_error / total_reproj_error against the REAL reference reproj_error (cv2_lite/solve_pnp.py:9-14) run
    per observation as in sfm.py:234-238 (ref_reproj_error.npz)."""
    d = load_golden("ref_reproj_error.npz")
    e, tot = O.total_reproj_error(d["K"], d["Hs"], d["X3d"], d["ci"], d["pi"], d["uv"])
    assert np.abs(e - d["per_obs"]).max() <= 1e-12 * d["per_obs"].max()
    assert abs(tot - float(d["errs"])) <= 1e-12 * float(d["errs"])
    m = d["ci"] == 4
    vec = O.reproj_error(d["X3d"][d["pi"][m]], d["uv"][m], d["K"], d["Hs"][4, :3, :3], d["Hs"][4, :3, 3])
    assert np.abs(vec - d["cam4_vec"]).max() <= 1e-12 * np.abs(d["cam4_vec"]).max()
    assert abs(O.calc_reproj_error(d["X3d"][d["pi"][m]], d["uv"][m], d["K"], d["Hs"][4, :3, :3], d["Hs"][4, :3, 3]) -
               float(d["cam4_mean"])) <= 1e-12 * float(d["cam4_mean"])


@pytest.mark.parametrize("k", [0, 2])
def test_oracle_lm_with_fixed_camera(k):
    """`fixed_camera_indices` (bundle_adjustment.py:6,13): the oracle LM with that camera's rows/columns deleted keeps it
    bit-fixed, reaches the frozen minimum and beats scipy run with the REAL reference mask (scipy stops on xtol far
    above the minimum once columns are masked, sceaux_fixed_ref.npz)."""
    g = load_golden("sceaux_fixed_ref.npz")
    d = load_golden(f"sceaux_ba_{k}.npz")
    fixed = tuple(int(c) for c in g[f"fixed_{k}"])
    lm = O.lm_solve(d["x0"], *golden_args(d), ftol=float(d["ftol"]), fixed_cameras=fixed)
    c = fixed[0]
    assert np.array_equal(lm["x"][6 * c:6 * c + 6], d["x0"][6 * c:6 * c + 6])
    assert abs(lm["cost"] - float(g[f"lm_cost_{k}"])) <= 1e-6 * float(g[f"lm_cost_{k}"])
    assert lm["cost"] <= float(g[f"cost_{k}"]) * (1 + 1e-6)
    assert float(g[f"moved_{k}"]) <= 1e-9     # scipy, too, left the camera where it was


def test_pose_conversion_oracle_vs_golden():
    """oracle.pack_cameras / unpack_cameras (sfm.py:252-257 / 273-278 through scipy Rotation) against the frozen vectors;
    guards against a scipy on the test machine that converts differently from the one the goldens were made with."""
    d = load_golden("ref_pose_conversions.npz")
    p = O.pack_cameras(d["H"])
    assert np.abs(p - d["params"]).max() <= 1e-12
    Hb = O.unpack_cameras(d["params"])
    assert np.abs(Hb[:, :3, :3] - d["R_back"]).max() <= 1e-14
    assert np.array_equal(Hb[:, :3, 3], d["params"][:, 3:])


def test_gather_observations_matches_reference_iteration():
    """b200ba.gather_observations against a literal restatement of Graph.pt3ds_pt2ds (graph.py:186-191) + the index
    remapping of sfm.py:248-256 on a small random track set (host logic only)."""
    import b200ba
    rng = np.random.default_rng(3)
    n_pts, cam_ids = 50, [2, 3, 5, 8, 9, 11]
    registered = [3, 9, 2, 11]
    tracks = [[(int(c), int(rng.integers(0, 999)), int(rng.integers(0, 2000)), int(rng.integers(0, 1500)))
               for c in rng.choice(cam_ids, size=int(rng.integers(2, 5)), replace=False)] for _ in range(n_pts)]
    data = [(idx, cam_id, np.array([x, y])) for idx in range(n_pts) for cam_id, _, x, y in tracks[idx] if cam_id in registered]
    pt_indices, cam_indices, pt2ds = map(np.array, zip(*data))                      # sfm.py:248-249
    camera_map = {c: i for i, c in enumerate(sorted(registered))}                     # node order = ascending id
    camera_indices = np.array([camera_map[c] for c in cam_indices])                   # sfm.py:256
    flat = [(idx, c, x, y) for idx in range(n_pts) for c, _, x, y in tracks[idx]]
    tp, tc, tx, ty = map(np.array, zip(*flat))
    ci, pi, uv = b200ba.gather_observations(tp, tc, np.stack([tx, ty], 1), registered)
    assert np.array_equal(ci, camera_indices) and np.array_equal(pi, pt_indices) and np.array_equal(uv, pt2ds)


def test_triangulation_oracle_vs_cv2_and_reference():
    """oracle.triangulate_points_linear2 / visibility_mask against the REAL cv2.triangulatePoints (the call at sfm.py:216)
    and the reference's triangulate_points_linear2 (ref_triangulate.npz), incl. the reference's docstring example."""
    d = load_golden("ref_triangulate.npz")
    M1, M2 = d["K"] @ d["H1"][:3], d["K"] @ d["H2"][:3]
    X = O.triangulate_points_linear2(M1, M2, d["pts1"].T, d["pts2"].T)
    scale = np.abs(d["X_cv2"]).max(axis=0)
    assert (np.abs(X - d["X_cv2"]).max(axis=0) <= 1e-9 * scale).all()
    assert (np.abs(X - d["X_ref_linear2"]).max(axis=0) <= 1e-9 * scale).all()
    assert np.array_equal(O.visibility_mask(d["H1"], d["H2"], X), d["mask"])
    Xd = O.triangulate_points_linear2(d["doc_M1"], d["doc_M2"], d["doc_pts1"].T, d["doc_pts2"].T)
    assert np.abs(Xd - d["doc_X_cv2"]).max() <= 1e-9 * np.abs(d["doc_X_cv2"]).max()


@pytest.mark.parametrize("name", ["doc9", "s60", "s800"])
def test_pnp_oracle_vs_cv2_and_reference(name):
    """oracle.solve_pnp (DLT + scipy refinement, solve_pnp.py:17-66) against the REAL cv2.solvePnP (sfm.py:207-208) and
    the reference's solve_pnp on its own 9-point example and seeded scenes (ref_pnp.npz)."""
    d = load_golden("ref_pnp.npz")
    rv, tv, cost = O.solve_pnp(d[f"{name}_X"], d[f"{name}_uv"], d["K"])
    got = np.hstack([rv, tv])
    assert np.abs(got - d[f"{name}_cv2"]).max() <= 1e-5 * np.abs(d[f"{name}_cv2"]).max()
    assert np.abs(got - d[f"{name}_ref"]).max() <= 1e-5 * np.abs(d[f"{name}_ref"]).max()
    assert cost <= float(d[f"{name}_cost_cv2"]) * (1 + 1e-9) and cost <= float(d[f"{name}_cost_ref"]) * (1 + 1e-9)


def test_configB_golden_is_consistent():
    """tests/golden/configB_scipy.npz (full scipy R2 solve of config B): the stored scipy solution reproduces the stored
    cost through the oracle residual, scipy stopped above the oracle LM minimum, and the arbiter could not go lower."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("b200ba_synthetic_t", os.path.join(ROOT, "sfm-python_b200", "synthetic.py"))
    syn = importlib.util.module_from_spec(spec)
    sys.modules["b200ba_synthetic_t"] = syn
    spec.loader.exec_module(syn)
    g = load_golden("configB_scipy.npz")
    pb = syn.make_config("B")
    assert abs(float(pb.x0.sum()) - float(g["x0_sum"])) <= 1e-9 * abs(float(g["x0_sum"]))
    r = O.residuals(g["scipy_x"], *pb.args())
    assert abs(0.5 * float(r @ r) - float(g["scipy_cost"])) <= 1e-12 * float(g["scipy_cost"])
    assert float(g["lm_cost"]) <= float(g["scipy_cost"])
    assert float(g["lm_cost"]) - float(g["arbiter_from_lm_cost"]) <= 1e-6 * float(g["lm_cost"])


@pytest.mark.parametrize("loss", ["soft_l1", "huber", "cauchy", "arctan"])
def test_robust_scaling_matches_scipy(loss):
    """The oracle's rho / scaling restatement against scipy's own construct_loss_function and
    scale_for_robust_loss_function (the code path behind loss= at sfm.py:266)."""
    from scipy.optimize._lsq.common import scale_for_robust_loss_function
    from scipy.optimize._lsq.least_squares import construct_loss_function
    rng = np.random.default_rng(3)
    f = np.concatenate([rng.standard_normal(200) * 4.0, [0.0, 1e-9, 2.5, -2.5, 1e4]])
    J = rng.standard_normal((len(f), 5))
    for C in (0.7, 2.5):
        lf = construct_loss_function(len(f), loss, C)
        Js, fs = scale_for_robust_loss_function(J.copy(), f.copy(), lf(f))
        js, r = O.robust_scale(f, loss, C, clamp=0.0)          # clamp=0: scipy's own rows
        # where rho' + 2 rho'' z cancels to rounding noise (huber's linear part, far outliers) scipy's row factor is
        # sqrt(noise or eps) ~ 1e-8: only the products J_s^T J_s (~ 0) and J_s^T f_s = J^T rho' f are meaningful there
        ok = js > 1e-6
        assert np.abs(Js - J * js[:, None])[ok].max() <= 1e-13 * np.abs(Js).max()
        assert np.abs(fs - r)[ok].max() <= 1e-12 * np.abs(fs[ok]).max()
        assert np.abs(Js - J * js[:, None]).max() <= 1e-7 * np.abs(J).max()
        g_s, g_o = Js * fs[:, None], (J * js[:, None]) * r[:, None]
        assert np.abs(g_s - g_o).max() <= 1e-12 * np.abs(g_s).max()
        assert abs(lf(f, cost_only=True) - O.robust_cost(f, loss, C)) <= 1e-13 * O.robust_cost(f, loss, C)
        # the default clamp (used by the LM loops) changes no row whose second-order weight is at least 5 % of rho' and
        # never changes the gradient
        jc, rc = O.robust_scale(f, loss, C)
        keep = js * js >= O.ROBUST_CLAMP * O.robust_rho((f / C) ** 2, loss)[1]
        assert np.array_equal(jc[keep], js[keep]) and np.abs(rc - r)[keep].max() == 0.0
        g_c = (J * jc[:, None]) * rc[:, None]
        assert np.abs(g_c - g_s).max() <= 1e-12 * np.abs(g_s).max()


def test_nonlinear_triangulation_oracle_vs_reference_golden():
    """The oracle's per-point minimiser of the reference's triangulation objective reaches a value not above the REAL
    reference function's (tests/golden/ref_triangulate.npz, cv2_lite/triangulate_points.py:122-184) and lands within the
    reference's own stopping accuracy of its points."""
    t = load_golden("ref_triangulate.npz")
    M1, M2 = t["K"] @ t["H1"][:3], t["K"] @ t["H2"][:3]
    vis = np.flatnonzero(t["mask"])[:120]
    p1, p2 = t["pts1"][vis], t["pts2"][vis]
    Xo = O.triangulate_points_nonlinear(M1, M2, p1.T, p2.T)
    Xr = t["X_ref_nonlinear"][:, :120]
    co = O.triangulation_objective(M1, M2, p1.T, p2.T, Xo)
    cr = O.triangulation_objective(M1, M2, p1.T, p2.T, Xr)
    assert float(co.sum()) <= float(cr.sum()) * (1 + 1e-12)
    assert np.abs(Xo - Xr).max() <= 1e-2 * np.abs(Xr).max()
