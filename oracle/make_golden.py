"""Generate tests/golden/*.npz from the REAL reference (run in the build container only).

TEST INFRASTRUCTURE.  ``/root/reference`` does not exist on the GPU box, so the
vectors produced here are committed and nothing at test time reads the reference.

    python oracle/make_golden.py residual      # seeded problems -> reference compute_residuals / sparsity
    python oracle/make_golden.py sceaux        # SceauxCastle --use_ba: frozen per-call BA problem dumps
    python oracle/make_golden.py scipy [k...]  # scipy (sfm.py:266-268 kwargs) on the dumps -> cost / RMS
    python oracle/make_golden.py synthB        # scipy vs oracle LM on a reduced synthetic problem

What is pinned:
* ``ref_residual_*.npz``  -- outputs of ``/root/reference/sfm_lite/bundle_adjustment.py``
  (``compute_residuals`` :35-42, ``create_sparsity_matrix`` :6-17), imported by path.
* ``sceaux_ba_k.npz``     -- arguments of the k-th ``least_squares`` call made by
  ``SFM._apply_bundle_adjustment`` (sfm.py:266-268) on the SceauxCastle images, captured by
  replacing ``sfm_lite.sfm.least_squares`` in a writable copy of the reference, plus the
  reference residual at ``x0`` and the converged cost of scipy run with the reference kwargs.
"""
from __future__ import annotations

import ast
import importlib.util
import os
import shutil
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GOLD = os.path.join(ROOT, "tests", "golden")
REF = "/root/reference"
sys.path.insert(0, HERE)
import ba_oracle as O  # noqa: E402


def load_by_path(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def ref_ba():
    return load_by_path("ref_ba", os.path.join(REF, "sfm_lite", "bundle_adjustment.py"))


def synthetic():
    return load_by_path("b200ba_synthetic", os.path.join(ROOT, "sfm-python_b200", "synthetic.py"))


# ---------------------------------------------------------------- residual fixtures
def gen_residual():
    ref = ref_ba()
    syn = synthetic()
    rng = np.random.default_rng(1234)
    os.makedirs(GOLD, exist_ok=True)

    def dump(name, x, n_cam, n_pts, ci, pi, uv, K, with_sparsity=False):
        r = ref.compute_residuals(x, n_cam, n_pts, ci, pi, uv, K)
        out = dict(x=x, n_cam=n_cam, n_pts=n_pts, ci=ci, pi=pi, uv=uv, K=K, r_ref=r)
        if with_sparsity:
            S = ref.create_sparsity_matrix(n_cam, n_pts, len(ci), ci, pi).tocoo()
            order = np.lexsort((S.col, S.row))
            out.update(sp_row=S.row[order].astype(np.int32), sp_col=S.col[order].astype(np.int32),
                       sp_shape=np.array(S.shape))
        np.savez_compressed(os.path.join(GOLD, name), **out)
        print(name, "N", len(ci), "max|r|", np.abs(r).max())

    # (1) small synthetic, float pixels, identity camera at index 0, with sparsity pattern
    pb = syn.make_problem(6, 80, 300, seed=3, window=6)
    dump("ref_residual_small.npz", pb.x0, pb.n_cam, pb.n_pts, pb.camera_indices, pb.point_indices,
         pb.points_2d, pb.K, with_sparsity=True)
    # (2) int64 truncated pixels (graph.py:112-113), unsorted observation order, identity camera moved to index 2
    pb = syn.make_problem(5, 60, 200, seed=4, window=5, int_pixels=True)
    perm = rng.permutation(pb.n_obs)
    cams = pb.x0[:30].reshape(5, 6).copy()
    cams[[0, 2]] = cams[[2, 0]]
    ci = pb.camera_indices.copy()
    ci = np.where(ci == 0, 2, np.where(ci == 2, 0, ci))
    x = np.concatenate([cams.ravel(), pb.x0[30:]])
    dump("ref_residual_int_unsorted.npz", x, 5, 60, ci[perm], pb.point_indices[perm], pb.points_2d[perm], pb.K)
    # (3) rotation-angle edge cases and a skewed K with a non-unit last row
    thetas = [0.0, 1e-12, 1e-9, 1e-7, 2.4e-5, 9.9e-5, 1.01e-4, 1e-3, 0.5, 0.99, 2.95, np.pi - 1e-9, np.pi, 3.3]
    n_cam = len(thetas)
    axes = rng.standard_normal((n_cam, 3))
    axes /= np.linalg.norm(axes, axis=1, keepdims=True)
    cams = np.concatenate([axes * np.array(thetas)[:, None], 0.3 * rng.standard_normal((n_cam, 3))], 1)
    n_pts = 40
    pts = rng.standard_normal((n_pts, 3)) * 2.0
    ci = np.repeat(np.arange(n_cam), n_pts)
    pi = np.tile(np.arange(n_pts), n_cam)
    Ksk = np.array([[2905.88, 3.5, 1416.0], [0.0, 2890.1, 1064.0], [1e-5, -2e-5, 1.02]])
    # move every point in front of every camera is impossible here; residuals are still well defined (z may be < 0)
    uv = rng.uniform(0, 2000, size=(len(ci), 2))
    x = np.concatenate([cams.ravel(), pts.ravel()])
    dump("ref_residual_angles_skewK.npz", x, n_cam, n_pts, ci, pi, uv, Ksk)
    # (4) one camera / nine points known-answer data of cv2_lite/solve_pnp.py:70-83 (parsed, not copied into the repo source)
    src = open(os.path.join(REF, "cv2_lite", "solve_pnp.py")).read()
    tree = ast.parse(src)
    lits = {}
    for node in ast.walk(tree):
        if isinstance(node, ast.Assign) and isinstance(node.targets[0], ast.Name) \
                and node.targets[0].id in ("point3ds", "point2ds", "K") and isinstance(node.value, ast.Call):
            lits[node.targets[0].id] = np.array(ast.literal_eval(node.value.args[0]), dtype=np.float64)
    p3, p2, Kp = lits["point3ds"], lits["point2ds"], lits["K"]
    import cv2
    ok, rvec, tvec = cv2.solvePnP(p3, p2, Kp, None)
    R, _ = cv2.Rodrigues(rvec)
    centre = (-R.T @ tvec).ravel()                      # BA convention: T is the camera centre
    x = np.concatenate([rvec.ravel(), centre, p3.ravel()])
    dump("ref_residual_pnp9.npz", x, 1, 9, np.zeros(9, dtype=np.int64), np.arange(9), p2, Kp, with_sparsity=True)
    # (5) a 2000-observation slice of config B at its start
    pb = syn.make_config("B")
    n = 2000
    dump("ref_residual_B2000.npz", pb.x0, pb.n_cam, pb.n_pts, pb.camera_indices[:n], pb.point_indices[:n],
         pb.points_2d[:n], pb.K)


# ---------------------------------------------------------------- SceauxCastle dumps
def gen_sceaux(workdir="/tmp/ref_copy", max_calls=11):
    if not os.path.isdir(workdir):
        shutil.copytree(REF, workdir)
    stub = os.path.join(workdir, "_stubs", "matplotlib")
    os.makedirs(stub, exist_ok=True)
    with open(os.path.join(stub, "__init__.py"), "w") as f:
        f.write("def use(*a, **k):\n    pass\n")
    open(os.path.join(stub, "pyplot.py"), "w").close()
    sys.path.insert(0, os.path.join(workdir, "_stubs"))
    sys.path.insert(0, workdir)
    os.chdir(workdir)
    import sfm_lite.sfm as S
    from sfm_lite import SFM, load_calibration_data

    calls = []

    class _Res:
        pass

    def spy(fun, x0, **kw):
        k = len(calls)
        n_cam, n_pts, ci, pi, uv, K = kw["args"]
        t = time.time()
        r0 = fun(x0, *kw["args"])                                  # the reference residual itself
        t_ref = time.time() - t
        rec = dict(x0=np.array(x0), n_cam=n_cam, n_pts=n_pts, ci=np.asarray(ci), pi=np.asarray(pi),
                   uv=np.asarray(uv), K=np.asarray(K), r0_ref=r0, ftol=kw["ftol"], t_ref_residual=t_ref)
        t = time.time()
        sc = O.scipy_solve(x0, n_cam, n_pts, ci, pi, uv, K, ftol=kw["ftol"], max_nfev=600)
        t_sc = time.time() - t
        lm = O.lm_solve(x0, n_cam, n_pts, ci, pi, uv, K, ftol=kw["ftol"])
        rec.update(scipy_cost=sc.cost, scipy_rms=O.rms(sc.fun), scipy_status=sc.status, scipy_nfev=sc.nfev,
                   scipy_max_nfev=600, scipy_t=t_sc, lm_cost=lm["cost"], lm_rms=lm["rms"], lm_nit=lm["nit"],
                   scipy_fun="oracle.residuals (vectorised twin; reference-residual runs are in sceaux_scipy_ref.npz)")
        print(f"[BA call {k}] C/P/N {n_cam}/{n_pts}/{len(ci)} rms0 {O.rms(r0):.1f} | scipy status {sc.status} nfev {sc.nfev} "
              f"cost {sc.cost:.6f} ({t_sc:.1f}s) | lm nit {lm['nit']} cost {lm['cost']:.6f}", flush=True)
        np.savez_compressed(os.path.join(GOLD, f"sceaux_ba_{k}.npz"), **rec)
        calls.append(rec)
        res = _Res()
        # continue the incremental run from scipy's answer where scipy converged (pure-reference trajectory
        # up to 1e-8), otherwise from the exact-Schur LM answer (scipy stalls from call 8 on)
        res.x = sc.x if sc.status > 0 else lm["x"]
        return res

    S.least_squares = spy
    K = load_calibration_data("./ImageDataset_SceauxCastle/images/K.txt")
    sfm = SFM("./ImageDataset_SceauxCastle/images", K)
    sfm.construct(min_matches=80, use_ba=True, ba_tol=1e-10, verbose=0)
    print("BA calls:", len(calls))


# ---------------------------------------------------------------- true-reference scipy on dumps
def gen_scipy(ks):
    """scipy + the REAL reference residual and sparsity (R1 of BASELINE.md) on selected dumps."""
    ref = ref_ba()
    from scipy.optimize import least_squares
    out_path = os.path.join(GOLD, "sceaux_scipy_ref.npz")
    out = dict(np.load(out_path)) if os.path.exists(out_path) else {}
    for k in ks:
        d = np.load(os.path.join(GOLD, f"sceaux_ba_{k}.npz"))
        args = (int(d["n_cam"]), int(d["n_pts"]), d["ci"], d["pi"], d["uv"], d["K"])
        t = time.time()
        S = ref.create_sparsity_matrix(args[0], args[1], len(d["ci"]), d["ci"], d["pi"])
        res = least_squares(ref.compute_residuals, d["x0"], jac_sparsity=S, verbose=0, x_scale="jac",
                            ftol=float(d["ftol"]), method="trf", args=args)
        dt = time.time() - t
        out[f"cost_{k}"] = res.cost
        out[f"rms_{k}"] = O.rms(res.fun)
        out[f"status_{k}"] = res.status
        out[f"nfev_{k}"] = res.nfev
        out[f"njev_{k}"] = res.njev
        out[f"t_{k}"] = dt
        print(f"[ref scipy {k}] cost {res.cost:.9f} rms {O.rms(res.fun):.6f} status {res.status} nfev {res.nfev} {dt:.1f}s",
              flush=True)
        np.savez_compressed(out_path, **out)


def gen_fixed(ks):
    """The `fixed_camera_indices` hook (bundle_adjustment.py:6,13): scipy with the REAL reference sparsity mask built
    for a fixed camera (its columns are absent, so scipy never moves it).  The fixed camera is the one with the lowest
    reprojection RMS at x0 (a camera that sits at the previous call's optimum; on call 0 the identity camera of
    sfm.py:176) -- fixing one of the two freshly re-registered cameras leaves no good minimum.
    Call 0 also runs the REAL reference residual; later calls use the vectorised twin.  The oracle LM with the same
    camera fixed gives the arbiter minimum."""
    ref = ref_ba()
    from scipy.optimize import least_squares
    out = {}
    for k in ks:
        d = np.load(os.path.join(GOLD, f"sceaux_ba_{k}.npz"))
        n_cam = int(d["n_cam"])
        args = (n_cam, int(d["n_pts"]), d["ci"], d["pi"], d["uv"], d["K"])
        r0 = O.residuals(d["x0"], *args).reshape(-1, 2)
        per_cam = [np.sqrt(np.mean(r0[d["ci"] == c] ** 2)) if np.any(d["ci"] == c) else np.inf for c in range(n_cam)]
        fixed = (int(np.argmin(per_cam)),)
        t = time.time()
        S = ref.create_sparsity_matrix(args[0], args[1], len(d["ci"]), d["ci"], d["pi"], fixed_camera_indices=fixed)
        fun = ref.compute_residuals if k == 0 else O.residuals
        res = least_squares(fun, d["x0"], jac_sparsity=S, verbose=0, x_scale="jac", ftol=float(d["ftol"]), method="trf",
                            args=args)
        moved = float(np.abs(res.x[6 * fixed[0]:6 * fixed[0] + 6] - d["x0"][6 * fixed[0]:6 * fixed[0] + 6]).max())
        lm = O.lm_solve(d["x0"], *args, ftol=float(d["ftol"]), fixed_cameras=fixed)
        out[f"fixed_{k}"] = np.array(fixed)
        out[f"cost_{k}"] = res.cost
        out[f"rms_{k}"] = O.rms(res.fun)
        out[f"status_{k}"] = res.status
        out[f"nfev_{k}"] = res.nfev
        out[f"moved_{k}"] = moved
        out[f"real_residual_{k}"] = int(k == 0)
        out[f"lm_cost_{k}"] = lm["cost"]
        out[f"lm_rms_{k}"] = lm["rms"]
        print(f"[fixed {k}] fixed {fixed} (rms at x0 per camera {np.round(per_cam, 2)}) scipy cost {res.cost:.9f} rms "
              f"{O.rms(res.fun):.6f} status {res.status} nfev {res.nfev} moved {moved:.1e} | oracle LM cost {lm['cost']:.9f} "
              f"status {lm['status']} nit {lm['nit']}  {time.time() - t:.1f}s", flush=True)
    np.savez_compressed(os.path.join(GOLD, "sceaux_fixed_ref.npz"), **out)


def gen_poses():
    """Pose <-> parameter conversions exactly as written at sfm.py:252-257 / 273-278 (scipy Rotation, the reference's
    own dependency; scipy version recorded): random rotations, theta in {0, 1e-12 .. pi - 1e-9}, slightly and clearly
    non-orthogonal matrices (cv2.recoverPose / cv2.Rodrigues outputs are orthogonal to ~1e-16)."""
    import scipy
    from scipy.spatial.transform import Rotation as Rot
    rng = np.random.default_rng(5)
    w = [rng.standard_normal(3) * s for s in (0.0, 1e-12, 1e-8, 1e-4, 9e-4, 1.1e-3, 0.3, 1.0, 2.5)]
    for th in (np.pi - 1e-9, np.pi - 1e-4, 3.0, np.pi):
        a = rng.standard_normal(3)
        w.append(a / np.linalg.norm(a) * th)
    w += [np.array([np.pi, 0, 0]), np.array([0, -np.pi + 1e-7, 0]), np.array([0, 0, 3.1])]
    w += list(rng.standard_normal((40, 3)) * 1.5)
    w = np.array(w)
    R = Rot.from_rotvec(w).as_matrix()
    Hs = np.tile(np.eye(4), (len(w), 1, 1))
    Hs[:, :3, :3] = R
    Hs[:, :3, 3] = rng.standard_normal((len(w), 3)) * 5
    noisy = Hs.copy()
    noisy[10:30, :3, :3] += 1e-9 * rng.standard_normal((20, 3, 3))      # fails the 1e-12 orthogonality test
    noisy[30:45, :3, :3] += 1e-3 * rng.standard_normal((15, 3, 3))
    allH = np.concatenate([Hs, noisy])
    params = np.stack([np.hstack([Rot.from_matrix(H[:3, :3]).as_rotvec(), H[:3, 3].flatten()]) for H in allH])  # sfm.py:252-255
    back = np.stack([Rot.from_rotvec(p[:3]).as_matrix() for p in params])                                        # sfm.py:277
    np.savez_compressed(os.path.join(GOLD, "ref_pose_conversions.npz"), H=allH, params=params, R_back=back,
                        scipy_version=scipy.__version__)
    print("pose golden:", allH.shape, "scipy", scipy.__version__, "max angle", np.linalg.norm(params[:, :3], axis=1).max())


def gen_pnp():
    """PnP as at sfm.py:207-208: the REAL cv2.solvePnP and the reference's cv2_lite.solve_pnp.solve_pnp on the
    reference's own 9-point example (solve_pnp.py:70-83) and on seeded scenes (60 and 800 points, 0.5 px noise)."""
    import cv2
    sys.path.insert(0, REF)
    import cv2_lite.solve_pnp as sp
    syn = synthetic()
    K = syn.K_SCEAUX
    rng = np.random.default_rng(31)
    cases = {}
    p3 = np.array([[-2.83510726, 0.35871423, 7.39620667], [-2.64956519, 1.26203440, 7.24983088],
                   [-2.67895385, -0.01053622, 7.470429], [-2.49225235, 0.81401947, 7.36628308],
                   [-2.31513876, -0.16642927, 7.48943918], [-2.40576114, -1.25461447, 8.02639394],
                   [-2.12475213, 0.15831901, 7.43070466], [-1.88875539, 1.02978915, 8.18975227],
                   [-1.70972510, -0.60470877, 8.44642688]])
    p2 = np.array([[232.86323547, 1230.31469727], [294.15100098, 1622.02612305], [312.4229126, 1076.39611816],
                   [377.7600708, 1423.75268555], [470.68951416, 1014.42590332], [477.48834229, 604.61303711],
                   [544.51257324, 1148.95483398], [672.24969482, 1470.36743164], [754.92327881, 871.83178711]])
    cases["doc9"] = (p3, p2)
    for name, n in (("s60", 60), ("s800", 800)):
        w = rng.standard_normal(3) * 0.2
        R = O.rodrigues(w[None])[0]
        t = np.array([0.3, -0.2, 0.5]) + 0.1 * rng.standard_normal(3)
        X = np.stack([rng.uniform(-3, 3, n), rng.uniform(-2, 2, n), rng.uniform(4, 12, n)], 1)
        p = (K @ (R @ X.T + t[:, None])).T
        cases[name] = (X, p[:, :2] / p[:, 2:3] + 0.5 * rng.standard_normal((n, 2)))
    out = {"K": K, "cv2_version": cv2.__version__}
    for name, (X, uv) in cases.items():
        ok, rv, tv = cv2.solvePnP(X, uv, K, np.zeros((1, 5)))
        ok2, rv2, tv2 = sp.solve_pnp(X, uv, K)
        cost_cv = 0.5 * float(np.sum(sp.reproj_error(X, uv, K, cv2.Rodrigues(rv)[0], tv.ravel()) ** 2))
        cost_ref = 0.5 * float(np.sum(sp.reproj_error(X, uv, K, cv2.Rodrigues(rv2)[0], tv2.ravel()) ** 2))
        out.update({f"{name}_X": X, f"{name}_uv": uv, f"{name}_cv2": np.hstack([rv.ravel(), tv.ravel()]),
                    f"{name}_ref": np.hstack([rv2.ravel(), tv2.ravel()]), f"{name}_cost_cv2": cost_cv, f"{name}_cost_ref": cost_ref})
        print(f"[pnp {name}] n {len(X)} cv2 {np.round(np.hstack([rv.ravel(), tv.ravel()]), 6)} cost {cost_cv:.9f} | reference "
              f"{np.round(np.hstack([rv2.ravel(), tv2.ravel()]), 6)} cost {cost_ref:.9f}")
    np.savez_compressed(os.path.join(GOLD, "ref_pnp.npz"), **out)


def gen_triangulate():
    """Two-view triangulation as at sfm.py:214-221: the REAL cv2.triangulatePoints (OpenCV, the call the reference makes)
    and the reference's own restatement triangulate_points_linear2 (cv2_lite/triangulate_points.py:66-94, loaded by path)
    on a seeded two-camera scene with pixel noise, plus the four-point example of the reference's docstring
    (triangulate_points.py:137-161) and the visibility mask."""
    import cv2
    tp = load_by_path("ref_tri", os.path.join(REF, "cv2_lite", "triangulate_points.py"))
    syn = synthetic()
    rng = np.random.default_rng(21)
    K = syn.K_SCEAUX
    H1 = np.eye(4)
    H2 = np.eye(4)
    H2[:3, :3] = O.rodrigues(np.array([[0.02, -0.15, 0.01]]))[0]
    H2[:3, 3] = [-1.0, 0.05, 0.1]
    n = 500
    X = np.stack([rng.uniform(-2, 2, n), rng.uniform(-1.5, 1.5, n), rng.uniform(-3, 9, n)], 1)   # some points behind the cameras
    X[np.abs(X[:, 2]) < 0.3, 2] = 4.0
    Xh = np.concatenate([X, np.ones((n, 1))], 1).T
    M1, M2 = K @ H1[:3], K @ H2[:3]
    def proj(M):
        p = M @ Xh
        return (p[:2] / p[2]).T + 0.5 * rng.standard_normal((n, 2))
    pts1, pts2 = proj(M1), proj(M2)
    cv = cv2.triangulatePoints(M1, M2, pts1.T, pts2.T)
    cv = cv / cv[-1]                                                                 # sfm.py:217
    ref2 = tp.triangulate_points_linear2(M1, M2, pts1.T, pts2.T)
    P1 = np.linalg.inv(H1) @ cv
    P2 = np.linalg.inv(H2) @ cv
    mask = (P1[2, :] > 0) & (P2[2, :] > 0)                                             # sfm.py:219-221
    dM1 = np.array([[2.26648588e+03, -1.20335940e+01, 2.30479750e+03, -8.10676693e+03],
                    [-5.27039708e+02, 2.80897851e+03, 1.18663534e+03, -8.67245706e+02],
                    [-3.31554792e-01, -8.05176652e-02, 9.39993790e-01, -2.70143322e-01]])
    dM2 = np.array([[2.57889420e+03, 2.21368015e+01, 1.94884818e+03, -5.51844918e+03],
                    [-2.65746745e+02, 2.88325548e+03, 1.09199408e+03, -4.01890574e+02],
                    [-1.92197430e-01, -2.02833813e-02, 9.81146642e-01, -7.57015427e-02]])
    d1 = np.array([[356.43856812, 1148.45336914], [360.64306641, 1452.07055664], [362.01452637, 1746.91137695],
                   [380.1892395, 1124.84753418]])
    d2 = np.array([[238.98179626, 1011.63299561], [251.32275391, 1309.41662598], [263.76693726, 1595.79040527],
                   [265.42037964, 989.33227539]])
    dcv = cv2.triangulatePoints(dM1, dM2, d1.T, d2.T)
    dcv = dcv / dcv[-1]
    # the reference's non-linear refinement (triangulate_points.py:122-184; one scipy call over all points, ~20 s) on the
    # visible points, and the value of its own objective there
    vis = np.flatnonzero(mask)
    nl = tp.triangulate_points(M1, M2, pts1[vis].T, pts2[vis].T)
    nl_cost = float(np.sum(O.triangulation_objective(M1, M2, pts1[vis].T, pts2[vis].T, nl)))
    np.savez_compressed(os.path.join(GOLD, "ref_triangulate.npz"), K=K, H1=H1, H2=H2, pts1=pts1, pts2=pts2, X_cv2=cv,
                        X_ref_linear2=ref2, mask=mask, X_ref_nonlinear=nl, ref_nonlinear_cost=nl_cost, doc_M1=dM1, doc_M2=dM2, doc_pts1=d1, doc_pts2=d2, doc_X_cv2=dcv,
                        cv2_version=cv2.__version__)
    print("triangulate golden: n", n, "visible", int(mask.sum()), "max |cv2 - linear2| rel",
          np.abs(cv - ref2).max() / np.abs(cv).max(), "cv2", cv2.__version__)


def gen_reproj():
    """Reprojection error in the pipeline's R X + t convention: the REAL reference functions reproj_error
    (cv2_lite/solve_pnp.py:9-14) and the loop body of sfm.py:234-238 (calc_reproj_error, sfm.py:38-41, restated in two
    lines here because sfm.py itself needs matplotlib/Tk to import) on seeded scenes."""
    sys.path.insert(0, REF)
    import cv2_lite.solve_pnp as sp
    syn = synthetic()
    rng = np.random.default_rng(11)
    pb = syn.make_problem(9, 300, 1200, seed=3, window=6)
    cams = pb.x_true[:54].reshape(9, 6)
    Hs = np.zeros((9, 4, 4))
    Hs[:, :3, :3] = O.rodrigues(cams[:, :3])
    Hs[:, :3, 3] = -np.einsum("nij,nj->ni", Hs[:, :3, :3], cams[:, 3:])     # R (X - T) = R X + t with t = -R T
    Hs[:, 3, 3] = 1.0
    Hs[:, :3, 3] += 1e-3 * rng.standard_normal((9, 3))
    X3d = pb.x_true[54:].reshape(-1, 3) + 1e-2 * rng.standard_normal((300, 3))
    ci, pi, uv = pb.camera_indices, pb.point_indices, np.floor(pb.points_2d)
    errs, per = 0.0, np.empty(len(ci))
    for n in range(len(ci)):                      # the loop of sfm.py:234-238, one observation per call
        e = sp.reproj_error(X3d[pi[n]].reshape(1, 3), uv[n].reshape(1, 2), pb.K, Hs[ci[n], :3, :3], Hs[ci[n], :3, 3])
        per[n] = np.linalg.norm(e, axis=1).mean()  # sfm.py:40
        errs += per[n]
    one = sp.reproj_error(X3d[pi[ci == 4]], uv[ci == 4], pb.K, Hs[4, :3, :3], Hs[4, :3, 3])
    np.savez_compressed(os.path.join(GOLD, "ref_reproj_error.npz"), Hs=Hs, X3d=X3d, ci=ci, pi=pi, uv=uv, K=pb.K, per_obs=per,
                        errs=errs, cam4_mean=float(np.linalg.norm(one, axis=1).mean()), cam4_vec=one)
    print("reproj golden: n", len(ci), "errs", errs, "mean", errs / len(ci), "camera 4 mean", np.linalg.norm(one, axis=1).mean())


def gen_synthB():
    """A reduced BAL-shaped problem where scipy (reference kwargs) can be run to the end."""
    syn = synthetic()
    pb = syn.make_problem(12, 400, 1500, seed=7, window=8)
    t = time.time()
    sc = O.scipy_solve(pb.x0, *pb.args(), ftol=1e-10)
    lm = O.lm_solve(pb.x0, *pb.args(), ftol=1e-10)
    print("scipy", sc.cost, sc.status, sc.nfev, time.time() - t, "lm", lm["cost"], lm["nit"])
    np.savez_compressed(os.path.join(GOLD, "synth_small_scipy.npz"), seed=7, n_cam=12, n_pts=400, n_obs=1500, window=8,
                        scipy_cost=sc.cost, scipy_rms=O.rms(sc.fun), scipy_status=sc.status, scipy_nfev=sc.nfev,
                        lm_cost=lm["cost"], lm_rms=lm["rms"], x0_sum=float(pb.x0.sum()))


def gen_configB():
    """BASELINE.md section 3 R2 on the repo's own config B (49 / 7776 / 31843): scipy TRF with the sfm.py:266-268 kwargs and
    the vectorised residual, run to the end (about 40 min of CPU), with a (seconds, cost) trace of every residual
    evaluation for time-to-cost; beside it the oracle LM minimum from the same start and the arbiter (scipy restarted
    from the oracle-LM point with ftol = xtol = 1e-15 must not lower the cost by 1e-6)."""
    syn = synthetic()
    pb = syn.make_config("B")
    args = pb.args()
    trace = []
    t0 = time.time()

    def fun(x, *a):
        r = O.residuals(x, *a)
        trace.append((time.time() - t0, 0.5 * float(r @ r)))
        return r

    sc = O.scipy_solve(pb.x0, *args, ftol=1e-10, fun=fun)
    secs = time.time() - t0
    tr = np.asarray(trace)
    reach = tr[tr[:, 1] <= sc.cost * (1 + 1e-6)]
    ttc = float(reach[0, 0]) if len(reach) else secs
    print("scipy R2: cost", sc.cost, "status", sc.status, "nfev", sc.nfev, "njev", sc.njev, "seconds", secs, "time-to-cost", ttc)
    t1 = time.time()
    lm = O.lm_solve(pb.x0, *args, ftol=1e-10)
    lm_secs = time.time() - t1
    arb = O.scipy_solve(lm["x"], *args, ftol=1e-15, xtol=1e-15, max_nfev=6)
    print("oracle LM: cost", lm["cost"], "iterations", lm["nit"], "seconds", lm_secs, "arbiter cost", arb.cost)
    # thin the trace: every 16th evaluation is plenty for a time-to-cost curve
    np.savez_compressed(os.path.join(GOLD, "configB_scipy.npz"), seed=0, x0_sum=float(pb.x0.sum()), scipy_cost=sc.cost,
                        scipy_rms=O.rms(sc.fun), scipy_status=sc.status, scipy_nfev=sc.nfev, scipy_njev=sc.njev,
                        scipy_optimality=sc.optimality, scipy_seconds=secs, scipy_time_to_cost=ttc, scipy_x=sc.x,
                        trace=tr[::16], lm_cost=lm["cost"], lm_rms=lm["rms"], lm_iterations=lm["nit"],
                        lm_seconds=lm_secs, arbiter_from_lm_cost=arb.cost, scipy_version=__import__("scipy").__version__)


def gen_robust(ks=(0, 1, 2)):
    """scipy's robust losses on the frozen SceauxCastle BA calls with 5 % outliers (uniform +-10 px) injected into the pixels (seeded).
    A robust cost is not convex, and from the raw starts of these calls (RMS 292 px) scipy's TRF and an LM loop end in
    different local minima, so the comparison starts where a pipeline would use a robust loss: at the converged
    linear-loss solution of the clean problem (oracle LM), then least_squares(..., loss=L, f_scale=2.0) with the
    sfm.py:266-268 kwargs on the oracle residual (bit-identical to the reference's; the REAL reference residual is
    evaluated at every scipy solution as a cross-check) next to the oracle LM with the same loss."""
    ref = ref_ba()
    out = {}
    f_scale = 2.0
    for k in ks:
        d = np.load(os.path.join(GOLD, f"sceaux_ba_{k}.npz"))
        n_cam, n_pts, ci, pi, K = int(d["n_cam"]), int(d["n_pts"]), d["ci"], d["pi"], d["K"]
        clean = O.lm_solve(d["x0"], n_cam, n_pts, ci, pi, d["uv"], K, ftol=1e-10)
        x_start = clean["x"]
        rng = np.random.default_rng([77, k])
        uv = d["uv"].astype(np.float64)
        bad = rng.choice(len(ci), size=max(len(ci) // 20, 1), replace=False)
        uv[bad] += rng.uniform(-10.0, 10.0, size=(len(bad), 2))
        out[f"uv_{k}"] = uv
        out[f"x_start_{k}"] = x_start
        args = (n_cam, n_pts, ci, pi, uv, K)
        for loss in ("soft_l1", "huber", "cauchy"):
            t = time.time()
            sc = O.scipy_solve(x_start, *args, ftol=1e-10, loss=loss, f_scale=f_scale, max_nfev=1500)
            r_real = ref.compute_residuals(sc.x, *args)
            assert abs(O.robust_cost(r_real, loss, f_scale) - sc.cost) <= 1e-12 * sc.cost
            lm = O.lm_solve(x_start, *args, ftol=1e-10, loss=loss, f_scale=f_scale, max_iter=300)
            print(k, loss, "scipy", sc.cost, sc.status, sc.nfev, "lm", lm["cost"], lm["status"], lm["nit"],
                  "rel gap", (sc.cost - lm["cost"]) / lm["cost"], time.time() - t, flush=True)
            out[f"cost_{k}_{loss}"] = sc.cost
            out[f"status_{k}_{loss}"] = sc.status
            out[f"nfev_{k}_{loss}"] = sc.nfev
            out[f"lm_cost_{k}_{loss}"] = lm["cost"]
    np.savez_compressed(os.path.join(GOLD, "robust_scipy.npz"), f_scale=f_scale, **out)


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "residual"
    if what == "residual":
        gen_residual()
    elif what == "sceaux":
        gen_sceaux()
    elif what == "scipy":
        gen_scipy([int(a) for a in sys.argv[2:]] or [0, 1, 2])
    elif what == "poses":
        gen_poses()
    elif what == "pnp":
        gen_pnp()
    elif what == "triangulate":
        gen_triangulate()
    elif what == "reproj":
        gen_reproj()
    elif what == "fixed":
        gen_fixed([int(a) for a in sys.argv[2:]] or [0, 1, 2])
    elif what == "synthB":
        gen_synthB()
    elif what == "robust":
        gen_robust()
    elif what == "configB":
        gen_configB()
