"""CPU oracle for the bundle-adjustment hot path.  TEST INFRASTRUCTURE ONLY.

This module is the checker, never the product: only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / ``--impl reference``
legs may import it.  The product (``sfm-python_b200``) never routes through it
and raises when its CUDA library is missing.

It restates, in numpy, the algorithm of the reference path

* ``sfm_lite/bundle_adjustment.py:20-32``  project_points   -> :func:`project_points_loop`
* ``sfm_lite/bundle_adjustment.py:35-42``  compute_residuals -> :func:`residuals_loop`,
  :func:`residuals` (vectorised twin, O1' in SURVEY.md section 8c)
* ``sfm_lite/bundle_adjustment.py:6-17``   create_sparsity_matrix -> :func:`sparsity`
* ``sfm_lite/sfm.py:266-268``              the scipy call -> :func:`scipy_solve`

and adds the analytic pieces the CUDA engine implements (Jacobian blocks,
normal equations, point-block Schur complement, a reference LM loop) so that
every kernel has a numpy twin to be compared with.

Parity pinning: the reference ships no tests or golden vectors for this path
(SURVEY.md section 4), so the oracle is pinned against OUTPUTS OF THE REFERENCE
ITSELF: ``oracle/make_golden.py`` imports ``/root/reference/sfm_lite/
bundle_adjustment.py`` by path, evaluates it on seeded problems and on the
frozen SceauxCastle BA dumps and commits the vectors under ``tests/golden/``;
``tests/test_oracle.py`` checks every function below against those vectors.
"""
from __future__ import annotations

import numpy as np

# --------------------------------------------------------------------------
# Rodrigues / rotation-vector helpers
# --------------------------------------------------------------------------

_SMALL = 1e-4  # series switch-over for a=(1-cos t)/t^2, b=(t-sin t)/t^3


def rodrigues(rvec):
    """Rotation matrices for an (n,3) array of rotation vectors.

    Same map as ``scipy.spatial.transform.Rotation.from_rotvec(...).as_matrix()``
    used at ``bundle_adjustment.py:25`` (agrees to <= 7e-16):
    R = I + s [w]x + a [w]x^2 with s = sin(t)/t, a = (1-cos t)/t^2.
    """
    w = np.atleast_2d(np.asarray(rvec, dtype=np.float64))
    t2 = np.einsum("ij,ij->i", w, w)
    t = np.sqrt(t2)
    small = t < _SMALL
    ts = np.where(small, 1.0, t)
    s = np.where(small, 1.0 - t2 / 6.0 + t2 * t2 / 120.0, np.sin(ts) / ts)
    # 1-cos t = 2 sin^2(t/2): avoids the cancellation of the naive form
    a = np.where(small, 0.5 - t2 / 24.0 + t2 * t2 / 720.0, 2.0 * np.sin(0.5 * ts) ** 2 / (ts * ts))
    Wx = skew(w)
    W2 = Wx @ Wx
    R = np.eye(3)[None] + s[:, None, None] * Wx + a[:, None, None] * W2
    return R


def skew(v):
    v = np.atleast_2d(np.asarray(v, dtype=np.float64))
    z = np.zeros(len(v))
    return np.stack(
        [
            np.stack([z, -v[:, 2], v[:, 1]], -1),
            np.stack([v[:, 2], z, -v[:, 0]], -1),
            np.stack([-v[:, 1], v[:, 0], z], -1),
        ],
        -2,
    )


def right_jacobian(rvec):
    """J_r(w) = I - a [w]x + b [w]x^2 (right Jacobian of SO(3)); d(R(w)Y)/dw = -R [Y]x J_r."""
    w = np.atleast_2d(np.asarray(rvec, dtype=np.float64))
    t2 = np.einsum("ij,ij->i", w, w)
    t = np.sqrt(t2)
    small = t < _SMALL
    ts = np.where(small, 1.0, t)
    a = np.where(small, 0.5 - t2 / 24.0 + t2 * t2 / 720.0, 2.0 * np.sin(0.5 * ts) ** 2 / (ts * ts))
    b = np.where(small, 1.0 / 6.0 - t2 / 120.0 + t2 * t2 / 5040.0, (ts - np.sin(ts)) / (ts * ts * ts))
    Wx = skew(w)
    return np.eye(3)[None] - a[:, None, None] * Wx + b[:, None, None] * (Wx @ Wx)


# --------------------------------------------------------------------------
# Residual (reference: bundle_adjustment.py:20-42)
# --------------------------------------------------------------------------

def unpack(x, n_cameras, n_points):
    """``x`` layout of sfm.py:255-257 / bundle_adjustment.py:38-39."""
    x = np.asarray(x, dtype=np.float64)
    cams = x[: 6 * n_cameras].reshape(n_cameras, 6)
    pts = x[6 * n_cameras: 6 * n_cameras + 3 * n_points].reshape(n_points, 3)
    return cams, pts


def project_points_loop(points, camera_params, K):
    """Per-observation restatement of bundle_adjustment.py:20-32 (one 3x4 matrix per observation).

    Deliberately keeps the reference's cost structure (a Python loop and a scipy
    ``Rotation`` per observation) because it doubles as the timed "port" CPU
    baseline of bench.py.
    """
    from scipy.spatial.transform import Rotation

    K = np.asarray(K, dtype=np.float64)
    out = np.empty((len(points), 2))
    for n in range(len(points)):
        cam = camera_params[n]
        R = Rotation.from_rotvec(cam[:3]).as_matrix()
        M = K @ np.column_stack((R, -(R @ cam[3:6])))       # K [R | -R T]
        ph = M @ np.array([points[n][0], points[n][1], points[n][2], 1.0])
        out[n, 0] = ph[0] / ph[2]
        out[n, 1] = ph[1] / ph[2]
    return out


def residuals_loop(x, n_cameras, n_points, camera_indices, point_indices, points_2d, K):
    """Faithful (slow) restatement of bundle_adjustment.py:35-42."""
    cams, pts = unpack(x, n_cameras, n_points)
    proj = project_points_loop(pts[point_indices], cams[camera_indices], K)
    return (proj - points_2d).ravel()


def residuals(x, n_cameras, n_points, camera_indices, point_indices, points_2d, K):
    """Vectorised twin of :func:`residuals_loop`: r = pi(K R(w)(X - T)) - uv, interleaved (2N,)."""
    cams, pts = unpack(x, n_cameras, n_points)
    K = np.asarray(K, dtype=np.float64)
    R = rodrigues(cams[:, :3])
    ci = np.asarray(camera_indices)
    pi = np.asarray(point_indices)
    Y = pts[pi] - cams[ci, 3:6]
    Z = np.einsum("nij,nj->ni", R[ci], Y)
    p = Z @ K.T
    proj = p[:, :2] / p[:, 2:3]
    return (proj - np.asarray(points_2d)).ravel()


def sparsity(n_cameras, n_points, n_obs, camera_indices, point3d_indices, fixed_camera_indices=()):
    """0/1 Jacobian structure of bundle_adjustment.py:6-17 as CSR (built without the Python loop)."""
    from scipy.sparse import csr_matrix

    ci = np.asarray(camera_indices, dtype=np.int64)
    pi = np.asarray(point3d_indices, dtype=np.int64)
    assert len(ci) == len(pi)
    obs = np.arange(n_obs, dtype=np.int64)
    free = ~np.isin(ci, np.asarray(list(fixed_camera_indices), dtype=np.int64))
    rows, cols = [], []
    for r in range(2):
        for k in range(6):
            rows.append(2 * obs[free] + r)
            cols.append(6 * ci[free] + k)
        for k in range(3):
            rows.append(2 * obs + r)
            cols.append(6 * n_cameras + 3 * pi + k)
    rows = np.concatenate(rows)
    cols = np.concatenate(cols)
    return csr_matrix((np.ones(len(rows), dtype=int), (rows, cols)),
                      shape=(2 * n_obs, 6 * n_cameras + 3 * n_points))


# --------------------------------------------------------------------------
# Analytic Jacobian blocks (replaces scipy's 2-point FD at sfm.py:266)
# --------------------------------------------------------------------------

def jacobian_blocks(x, n_cameras, n_points, camera_indices, point_indices, K):
    """Per-observation blocks Jc (N,2,6) [d/dw | d/dT] and Jp (N,2,3) = d/dX.

    Y = X - T, Z = R Y, p = K Z, (u,v) = p[:2]/p[2];
    A = [[1,0,-u],[0,1,-v]]/p2, B = A K;  J_X = B R;  J_T = -J_X;  J_w = -J_X [Y]x J_r(w).
    """
    cams, pts = unpack(x, n_cameras, n_points)
    K = np.asarray(K, dtype=np.float64)
    ci = np.asarray(camera_indices)
    pi = np.asarray(point_indices)
    R = rodrigues(cams[:, :3])[ci]
    Jr = right_jacobian(cams[:, :3])[ci]
    Y = pts[pi] - cams[ci, 3:6]
    Z = np.einsum("nij,nj->ni", R, Y)
    p = Z @ K.T
    u = p[:, 0] / p[:, 2]
    v = p[:, 1] / p[:, 2]
    n = len(ci)
    A = np.zeros((n, 2, 3))
    A[:, 0, 0] = 1.0
    A[:, 1, 1] = 1.0
    A[:, 0, 2] = -u
    A[:, 1, 2] = -v
    A /= p[:, 2][:, None, None]
    B = A @ K
    JX = B @ R
    Jw = -(JX @ skew(Y)) @ Jr
    Jc = np.concatenate([Jw, -JX], axis=2)
    return Jc, JX


def dense_jacobian(x, n_cameras, n_points, camera_indices, point_indices, K):
    Jc, Jp = jacobian_blocks(x, n_cameras, n_points, camera_indices, point_indices, K)
    n = len(camera_indices)
    J = np.zeros((2 * n, 6 * n_cameras + 3 * n_points))
    for i in range(n):
        c = int(camera_indices[i])
        p = int(point_indices[i])
        J[2 * i:2 * i + 2, 6 * c:6 * c + 6] = Jc[i]
        J[2 * i:2 * i + 2, 6 * n_cameras + 3 * p:6 * n_cameras + 3 * p + 3] = Jp[i]
    return J



# --------------------------------------------------------------------------
# Robust loss (scipy's loss= / f_scale= keywords of the call at sfm.py:266; the reference keeps 'linear')
# restated from scipy _lsq/least_squares.py (soft_l1 / huber / cauchy / arctan, construct_loss_function) and
# _lsq/common.py scale_for_robust_loss_function; pinned against scipy itself in tests/golden/robust_scipy.npz.
# --------------------------------------------------------------------------

def robust_rho(z, loss):
    """rho(z), rho'(z), rho''(z) for z = (f / f_scale)^2, elementwise."""
    z = np.asarray(z, dtype=np.float64)
    if loss == "linear":
        return z, np.ones_like(z), np.zeros_like(z)
    if loss == "soft_l1":
        t = 1.0 + z
        return 2.0 * (np.sqrt(t) - 1.0), t ** -0.5, -0.5 * t ** -1.5
    if loss == "huber":
        m = z <= 1.0
        zs = np.where(m, 1.0, z)
        return (np.where(m, z, 2.0 * np.sqrt(zs) - 1.0), np.where(m, 1.0, zs ** -0.5),
                np.where(m, 0.0, -0.5 * zs ** -1.5))
    if loss == "cauchy":
        t = 1.0 + z
        return np.log1p(z), 1.0 / t, -1.0 / t ** 2
    if loss == "arctan":
        t = 1.0 + z * z
        return np.arctan(z), 1.0 / t, -2.0 * z / t ** 2
    raise ValueError(loss)


def robust_cost(r, loss="linear", f_scale=1.0):
    """0.5 * f_scale^2 * sum(rho((r / f_scale)^2)): scipy's OptimizeResult.cost under a robust loss."""
    r = np.asarray(r, dtype=np.float64)
    if loss == "linear":
        return 0.5 * float(np.sum(r * r))
    return 0.5 * f_scale ** 2 * float(np.sum(robust_rho((r / f_scale) ** 2, loss)[0]))


ROBUST_CLAMP = 0.05


def robust_scale(r, loss, f_scale, clamp=ROBUST_CLAMP):
    """(row factors of J, scaled residual) for residual components r (any shape): scipy's scale_for_robust_loss_function
    with ONE deliberate difference -- J_scale = max(rho' + 2 rho'' z, clamp * rho') where scipy clamps at machine epsilon.
    scipy steps inside a trust region, so rows whose second-order weight cancels to ~0 (every Huber outlier, far
    soft_l1 / Cauchy outliers) are harmless there; in a Marquardt-damped Schur solve they leave point blocks numerically
    singular and the LM loop stalls far above scipy's cost (measured: Huber 392 vs 261 on SceauxCastle call 0).  The
    gradient J_s^T f_s = J^T rho' f is exact for any clamp, so the minimiser is unchanged; clamp=0 gives scipy's rows."""
    z = (r / f_scale) ** 2
    _, r1, r2 = robust_rho(z, loss)
    js = np.sqrt(np.maximum(np.maximum(r1 + 2.0 * r2 * z, clamp * r1), np.finfo(float).eps))
    return js, r * r1 / js

# --------------------------------------------------------------------------
# Normal equations + point-block Schur complement
# --------------------------------------------------------------------------

def normal_blocks(x, n_cameras, n_points, camera_indices, point_indices, points_2d, K, loss="linear", f_scale=1.0):
    """U (C,6,6), V (P,3,3), gc (C,6), gp (P,3), cost = 0.5 sum r^2, plus per-observation Jc, Jp, r."""
    ci = np.asarray(camera_indices)
    pi = np.asarray(point_indices)
    r = residuals(x, n_cameras, n_points, ci, pi, points_2d, K).reshape(-1, 2)
    Jc, Jp = jacobian_blocks(x, n_cameras, n_points, ci, pi, K)
    cost = robust_cost(r, loss, f_scale)
    if loss != "linear":
        js, r = robust_scale(r, loss, f_scale)
        Jc = Jc * js[:, :, None]
        Jp = Jp * js[:, :, None]
    U = np.zeros((n_cameras, 6, 6))
    V = np.zeros((n_points, 3, 3))
    gc = np.zeros((n_cameras, 6))
    gp = np.zeros((n_points, 3))
    np.add.at(U, ci, np.einsum("nri,nrj->nij", Jc, Jc))
    np.add.at(V, pi, np.einsum("nri,nrj->nij", Jp, Jp))
    np.add.at(gc, ci, np.einsum("nri,nr->ni", Jc, r))
    np.add.at(gp, pi, np.einsum("nri,nr->ni", Jp, r))
    return dict(U=U, V=V, gc=gc, gp=gp, cost=cost, Jc=Jc, Jp=Jp, r=r)


def schur_system(x, n_cameras, n_points, camera_indices, point_indices, points_2d, K, lam, loss="linear", f_scale=1.0):
    """Dense reduced camera system with Marquardt damping  (H + lam diag H) d = -g.

    S = U* - sum_p W_p V_p*^-1 W_p^T,  b = -gc + sum_p W_p V_p*^-1 gp  (SURVEY 8 a8).
    Returns S (6C,6C), b (6C,), and the pieces needed for back-substitution.
    """
    ci = np.asarray(camera_indices)
    pi = np.asarray(point_indices)
    nb = normal_blocks(x, n_cameras, n_points, ci, pi, points_2d, K, loss, f_scale)
    U, V, gc, gp, Jc, Jp = nb["U"], nb["V"], nb["gc"], nb["gp"], nb["Jc"], nb["Jp"]
    Us = U.copy()
    Vs = V.copy()
    i6 = np.arange(6)
    i3 = np.arange(3)
    Us[:, i6, i6] *= (1.0 + lam)
    Vs[:, i3, i3] *= (1.0 + lam)
    Vinv = np.zeros_like(Vs)
    ok = np.einsum("pii->p", Vs) > 0                          # points without observations stay frozen
    if ok.any():
        Vinv[ok] = np.linalg.inv(Vs[ok])
    W = np.einsum("nri,nrj->nij", Jc, Jp)                    # (N,6,3)
    n = 6 * n_cameras
    S = np.zeros((n, n))
    b = -gc.reshape(-1).copy()
    for c in range(n_cameras):
        S[6 * c:6 * c + 6, 6 * c:6 * c + 6] = Us[c]
    WVi = np.einsum("nij,njk->nik", W, Vinv[pi])             # (N,6,3)
    np.add.at(b.reshape(n_cameras, 6), ci, np.einsum("nij,nj->ni", WVi, gp[pi]))
    order = np.argsort(pi, kind="stable")
    starts = np.flatnonzero(np.r_[True, pi[order][1:] != pi[order][:-1], True])
    for a, e in zip(starts[:-1], starts[1:]):
        idx = order[a:e]
        for i in idx:
            for j in idx:
                S[6 * ci[i]:6 * ci[i] + 6, 6 * ci[j]:6 * ci[j] + 6] -= WVi[i] @ W[j].T
    return dict(S=S, b=b, Vinv=Vinv, W=W, **nb)


def lm_step(x, n_cameras, n_points, camera_indices, point_indices, points_2d, K, lam, fixed_cameras=(), loss="linear",
            f_scale=1.0):
    """One damped Gauss-Newton step via the Schur complement; returns (delta, predicted_reduction, sys).

    ``fixed_cameras``: the ``fixed_camera_indices`` hook of bundle_adjustment.py:6,13 -- those cameras' Jacobian
    columns are masked out in the reference, i.e. rows/columns of the reduced system are deleted (delta_c = 0)."""
    s = schur_system(x, n_cameras, n_points, camera_indices, point_indices, points_2d, K, lam, loss, f_scale)
    ci = np.asarray(camera_indices)
    pi = np.asarray(point_indices)
    for c in fixed_cameras:
        sl = slice(6 * c, 6 * c + 6)
        s["S"][sl, :] = 0.0
        s["S"][:, sl] = 0.0
        s["S"][sl, sl] = np.eye(6)
        s["b"][sl] = 0.0
        s["gc"][c] = 0.0
    dc = np.linalg.solve(s["S"], s["b"])
    Wt_dc = np.zeros((n_points, 3))
    np.add.at(Wt_dc, pi, np.einsum("nij,ni->nj", s["W"], dc.reshape(n_cameras, 6)[ci]))
    dp = np.einsum("pij,pj->pi", s["Vinv"], -s["gp"] - Wt_dc)
    delta = np.concatenate([dc, dp.reshape(-1)])
    g = np.concatenate([s["gc"].reshape(-1), s["gp"].reshape(-1)])
    d2 = np.concatenate([np.einsum("cii->ci", s["U"]).reshape(-1), np.einsum("pii->pi", s["V"]).reshape(-1)])
    pred = 0.5 * (lam * float(delta @ (d2 * delta)) - float(g @ delta))
    return delta, pred, s


def lm_solve(x0, n_cameras, n_points, camera_indices, point_indices, points_2d, K,
             ftol=1e-10, xtol=1e-8, gtol=1e-8, max_iter=100, lam0=1e-4, verbose=0, fixed_cameras=(), loss="linear",
             f_scale=1.0):
    """Reference LM loop (numpy twin of the engine's b200ba_solve).

    Marquardt damping lam*diag(H); accept iff the cost decreases; Nielsen update
    lam *= max(1/3, 1-(2 rho-1)^3) on accept (0.1 when rho > 0.9), lam *= nu, nu *= 2 on reject;
    scipy-compatible termination (scipy _lsq/common.py check_termination).
    """
    x = np.array(x0, dtype=np.float64)
    args = (n_cameras, n_points, camera_indices, point_indices, points_2d, K)
    r = residuals(x, *args)
    cost = robust_cost(r, loss, f_scale)
    lam, nu = lam0, 2.0
    status, nfev, it = 0, 1, 0
    while it < max_iter:
        it += 1
        delta, pred, s = lm_step(x, *args, lam, fixed_cameras, loss, f_scale)
        g = np.concatenate([s["gc"].reshape(-1), s["gp"].reshape(-1)])
        gnorm = float(np.max(np.abs(g)))
        if gnorm < gtol:
            status = 1
            break
        x_new = x + delta
        r_new = residuals(x_new, *args)
        nfev += 1
        cost_new = robust_cost(r_new, loss, f_scale)
        actual = cost - cost_new
        rho = actual / pred if pred > 0 else -1.0
        step_norm = float(np.linalg.norm(delta))
        if verbose:
            print(f"it {it:3d} cost {cost:.9e} -> {cost_new:.9e} rho {rho:+.3f} lam {lam:.2e} |d| {step_norm:.2e} |g| {gnorm:.2e}")
        if np.isfinite(cost_new) and actual > 0:
            ftol_ok = actual < ftol * cost and rho > 0.25
            xtol_ok = step_norm < xtol * (xtol + float(np.linalg.norm(x)))
            x, cost = x_new, cost_new
            lam *= 0.1 if (rho > 0.9 and loss == "linear") else max(1.0 / 3.0, 1.0 - (2.0 * rho - 1.0) ** 3)   # as csrc/ba_lm.inc
            nu = 2.0
            if ftol_ok and xtol_ok:
                status = 4
            elif ftol_ok:
                status = 2
            elif xtol_ok:
                status = 3
            if status:
                break
        else:
            xtol_ok = step_norm < xtol * (xtol + float(np.linalg.norm(x)))
            if xtol_ok:
                status = 3
                break
            lam *= nu
            nu *= 2.0
    r = residuals(x, *args)
    return dict(x=x, cost=robust_cost(r, loss, f_scale), fun=r, status=status, nit=it, nfev=nfev,
                rms=float(np.sqrt(np.mean(r * r))))


# --------------------------------------------------------------------------
# The reference solver call (sfm.py:266-268)
# --------------------------------------------------------------------------

def scipy_solve(x0, n_cameras, n_points, camera_indices, point_indices, points_2d, K,
                ftol=1e-10, verbose=0, max_nfev=None, fun=None, xtol=1e-8, fixed_cameras=(), **kw):
    """scipy.optimize.least_squares with exactly the keyword arguments of sfm.py:266-268.

    ``fun`` defaults to the vectorised residual (R2 of BASELINE.md); pass
    :func:`residuals_loop` for the faithful per-observation port (R1).
    """
    from scipy.optimize import least_squares

    fun = residuals if fun is None else fun
    S = sparsity(n_cameras, n_points, len(camera_indices), camera_indices, point_indices, fixed_cameras)
    return least_squares(fun, np.asarray(x0, dtype=np.float64), jac_sparsity=S, verbose=verbose,
                         x_scale="jac", ftol=ftol, xtol=xtol, method="trf", max_nfev=max_nfev,
                         args=(n_cameras, n_points, camera_indices, point_indices, points_2d, K), **kw)


# --------------------------------------------------------------------------
# Pose <-> parameter conversion at the boundary
# --------------------------------------------------------------------------

def pack_cameras(Hs):
    """sfm.py:252-255 per camera: (Rot.from_matrix(R).as_rotvec(), T) with R, T = RT_from_H(H) (transforms.py:12-15)."""
    from scipy.spatial.transform import Rotation as Rot
    Hs = np.asarray(Hs, dtype=np.float64)
    return np.stack([np.hstack([Rot.from_matrix(H[:3, :3]).as_rotvec(), H[:3, 3].flatten()]) for H in Hs])


def unpack_cameras(params):
    """sfm.py:273-278 per camera: H_from_RT(Rot.from_rotvec(w).as_matrix(), T) (transforms.py:5-9)."""
    from scipy.spatial.transform import Rotation as Rot
    out = []
    for p in np.asarray(params, dtype=np.float64).reshape(-1, 6):
        H = np.eye(4)
        H[:3, :3] = Rot.from_rotvec(p[:3]).as_matrix()
        H[:3, 3] = p[3:]
        out.append(H)
    return np.stack(out)


# --------------------------------------------------------------------------
# Reprojection error in the pipeline's R X + t convention (next row 8f-2)
# --------------------------------------------------------------------------

def reproj_error(point3ds, point2ds, K, R, tvec):
    """cv2_lite/solve_pnp.py:9-14: (K (R X^T + t))^T, divided by its third column, minus the observations -> (n, 2)."""
    p = (np.asarray(K) @ (np.asarray(R) @ np.asarray(point3ds).T + np.asarray(tvec).reshape(3, 1))).T
    p = p / p[:, 2].reshape(-1, 1)
    return p[:, :2] - point2ds


def calc_reproj_error(points3d, points2d, K, R, tvec):
    """sfm_lite/sfm.py:38-41: mean of the row norms of reproj_error."""
    return float(np.linalg.norm(reproj_error(points3d, points2d, K, R, tvec), axis=1).mean())


def total_reproj_error(K, Hs, X3d, camera_indices, point_indices, points_2d):
    """The loop of sfm.py:233-238, vectorised: per-observation error norms (n,) and their sum `errs`."""
    Hs = np.asarray(Hs, dtype=np.float64)
    ci = np.asarray(camera_indices)
    pi = np.asarray(point_indices)
    X = np.asarray(X3d, dtype=np.float64)[pi]
    Z = np.einsum("nij,nj->ni", Hs[ci, :3, :3], X) + Hs[ci, :3, 3]
    p = Z @ np.asarray(K, dtype=np.float64).T
    e = np.linalg.norm(p[:, :2] / p[:, 2:3] - np.asarray(points_2d, dtype=np.float64), axis=1)
    return e, float(e.sum())


# --------------------------------------------------------------------------
# Two-view triangulation (next row 8f-4; sfm.py:214-221)
# --------------------------------------------------------------------------

def triangulate_points_linear2(M1, M2, pts1, pts2):
    """cv2_lite/triangulate_points.py:66-94 (the DLT of cv2.triangulatePoints, sfm.py:216): per point the 4 x 4 system
    [x1 M1[2] - M1[0]; y1 M1[2] - M1[1]; x2 M2[2] - M2[0]; y2 M2[2] - M2[1]], last right singular vector, X /= X[3].
    pts are (2, N); returns (4, N).  Batched SVD instead of the reference's Python loop."""
    M1, M2 = np.asarray(M1, dtype=np.float64), np.asarray(M2, dtype=np.float64)
    p1, p2 = np.asarray(pts1, dtype=np.float64), np.asarray(pts2, dtype=np.float64)
    A = np.stack([p1[0][:, None] * M1[2] - M1[0], p1[1][:, None] * M1[2] - M1[1],
                  p2[0][:, None] * M2[2] - M2[0], p2[1][:, None] * M2[2] - M2[1]], axis=1)      # (N, 4, 4)
    Vt = np.linalg.svd(A)[2]
    X = Vt[:, -1, :].T
    return X / X[3]


def triangulation_objective(M1, M2, pts1, pts2, X3d):
    """The reference's objective of cv2_lite/triangulate_points.py:94-118 per point: its residual vector is the SQUARED
    reprojection error of each of the four image coordinates, which scipy squares again -- sum of the 4th powers."""
    Xh = np.vstack([np.asarray(X3d, dtype=np.float64)[:3], np.ones(np.asarray(X3d).shape[1])])
    a, b = np.asarray(M1) @ Xh, np.asarray(M2) @ Xh
    e = np.vstack([a[:2] / a[2] - np.asarray(pts1), b[:2] / b[2] - np.asarray(pts2)])
    return np.sum(e ** 4, axis=0)


def triangulate_points_nonlinear(M1, M2, pts1, pts2, max_iter=100):
    """cv2_lite/triangulate_points.py:122-184: DLT start, then minimise the objective above.  The reference hands all points
    to one scipy least_squares call with default tolerances (it stops ~1e-3 short of the minimiser); the points are
    independent, so this restatement runs a damped Gauss-Newton per point to the minimiser.  pts are (2, N); returns (4, N)."""
    M1, M2 = np.asarray(M1, dtype=np.float64), np.asarray(M2, dtype=np.float64)
    p1, p2 = np.asarray(pts1, dtype=np.float64), np.asarray(pts2, dtype=np.float64)
    X = triangulate_points_linear2(M1, M2, p1, p2)[:3].copy()

    def res(x, i):
        xh = np.append(x, 1.0)
        a, b = M1 @ xh, M2 @ xh
        e = np.array([a[0] / a[2] - p1[0, i], a[1] / a[2] - p1[1, i], b[0] / b[2] - p2[0, i], b[1] / b[2] - p2[1, i]])
        Je = np.stack([(M1[0, :3] - a[0] / a[2] * M1[2, :3]) / a[2], (M1[1, :3] - a[1] / a[2] * M1[2, :3]) / a[2],
                       (M2[0, :3] - b[0] / b[2] * M2[2, :3]) / b[2], (M2[1, :3] - b[1] / b[2] * M2[2, :3]) / b[2]])
        return e * e, 2.0 * e[:, None] * Je

    for i in range(X.shape[1]):
        x, lam = X[:, i].copy(), 1e-3
        r, J = res(x, i)
        c = float(r @ r)
        for _ in range(max_iter):
            Hm = J.T @ J
            try:
                dx = np.linalg.solve(Hm + lam * np.diag(np.diag(Hm)) + 1e-300 * np.eye(3), -(J.T @ r))
            except np.linalg.LinAlgError:
                break
            r2, J2 = res(x + dx, i)
            c2 = float(r2 @ r2)
            if np.isfinite(c2) and c2 < c:
                done = (c - c2) < 1e-16 * c or np.linalg.norm(dx) < 1e-14 * np.linalg.norm(x)
                x, r, J, c, lam = x + dx, r2, J2, c2, max(lam / 3.0, 1e-12)
                if done:
                    break
            else:
                lam *= 4.0
                if lam > 1e12:
                    break
        X[:, i] = x
    return np.vstack([X, np.ones(X.shape[1])])


def visibility_mask(H1, H2, X3d_H):
    """sfm.py:219-221: both cameras see the point in front of them."""
    P1 = np.linalg.inv(H1) @ X3d_H
    P2 = np.linalg.inv(H2) @ X3d_H
    return (P1[2, :] > 0) & (P2[2, :] > 0)


# --------------------------------------------------------------------------
# PnP (next row 8f-4; cv2.solvePnP at sfm.py:207-208, restated by cv2_lite/solve_pnp.py:17-66)
# --------------------------------------------------------------------------

def solve_pnp_linear(point3ds, point2ds, K):
    """cv2_lite/solve_pnp.py:17-42: DLT on normalised image points, rotation part orthogonalised by an SVD, sign from
    the determinant (t is NOT rescaled there -- a quirk of the reference that only matters for the starting point)."""
    n = len(point3ds)
    ph = np.hstack((point2ds, np.ones((n, 1))))
    nrm = (np.linalg.inv(K) @ ph.T)[:2].T
    A = np.zeros((2 * n, 12))
    X, Y, Z = point3ds[:, 0], point3ds[:, 1], point3ds[:, 2]
    u, v = nrm[:, 0], nrm[:, 1]
    o, z = np.ones(n), np.zeros(n)
    A[0::2] = np.stack([X, Y, Z, o, z, z, z, z, -u * X, -u * Y, -u * Z, -u], 1)
    A[1::2] = np.stack([z, z, z, z, X, Y, Z, o, -v * X, -v * Y, -v * Z, -v], 1)
    h = np.linalg.svd(A)[2][-1].reshape(3, 4)
    U, _, Vt = np.linalg.svd(h[:, :3])
    R, t = U @ Vt, h[:, 3]
    if np.linalg.det(R) < 0:
        R, t = -R, -t
    return R, t


def solve_pnp(point3ds, point2ds, K, x0=None):
    """cv2_lite/solve_pnp.py:45-66: DLT start, then scipy least_squares on reproj_error over (rvec, tvec).  Returns
    (rvec (3,), tvec (3,), cost).  rvec <-> R through scipy Rotation (the reference's matrix_to_rvec / rvec_to_matrix are
    plain Rodrigues maps)."""
    from scipy.optimize import least_squares
    from scipy.spatial.transform import Rotation as Rot
    point3ds, point2ds = np.asarray(point3ds, dtype=np.float64), np.asarray(point2ds, dtype=np.float64)
    if x0 is None:
        R, t = solve_pnp_linear(point3ds, point2ds, K)
        x0 = np.hstack((Rot.from_matrix(R).as_rotvec(), t))

    def fun(p):
        return reproj_error(point3ds, point2ds, K, Rot.from_rotvec(p[:3]).as_matrix(), p[3:6]).ravel()

    res = least_squares(fun, x0, xtol=1e-14, ftol=1e-14, gtol=1e-14)
    return res.x[:3], res.x[3:6], float(res.cost)


def rms(r):
    r = np.asarray(r)
    return float(np.sqrt(np.mean(r * r)))
