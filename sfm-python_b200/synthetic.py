"""Seeded synthetic BAL-shaped bundle-adjustment problems in the reference's camera model.

Generator specified in SURVEY.md section 8(d): shared intrinsics ``K`` of the
SceauxCastle dataset (``ImageDataset_SceauxCastle/images/K.txt``), 6-parameter
cameras ``[rvec | T]`` packed as in ``sfm_lite/sfm.py:255-257``, observations
``pi(K R(w)(X - T))`` -- the convention of ``sfm_lite/bundle_adjustment.py:25-30``
(``T`` acts as a camera centre) -- plus 0.5 px Gaussian noise, point-major order
(``point_indices`` non-decreasing as produced by ``sfm_lite/graph.py:186-191``),
camera 0 exactly ``w = 0, T = 0`` (``sfm.py:176``), sliding-window visibility so
the reduced camera system is block-banded, and ``sum(k_p) == n_obs`` exactly.

Everything is drawn per block of ``BLOCK`` points from ``default_rng([seed, block])``
so any contiguous point range (one GPU's shard) can be generated on its own and
is bit-identical to the same range of the full problem.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

K_SCEAUX = np.array([[2905.88, 0.0, 1416.0], [0.0, 2905.88, 1064.0], [0.0, 0.0, 1.0]])
BLOCK = 65536

CONFIGS = {
    # name: (n_cam, n_pts, n_obs)      BASELINE.json configs[1..3]
    "B": (49, 7776, 31843),
    "C": (1000, 500_000, 2_000_000),
    "D": (10_000, 4_000_000, 20_000_000),
}


@dataclass
class BAProblem:
    n_cam: int
    n_pts: int                  # global point count
    camera_indices: np.ndarray  # (n,) int64
    point_indices: np.ndarray   # (n,) int64, global ids, non-decreasing
    points_2d: np.ndarray       # (n,2) float64 or int64
    K: np.ndarray
    x0: np.ndarray              # (6C + 3P_local,) start
    x_true: np.ndarray          # (6C + 3P_local,) noise-free parameters
    point_range: tuple          # (lo, hi) of the points held in x0 / x_true
    n_obs_total: int

    @property
    def n_obs(self):
        return len(self.camera_indices)

    def args(self):
        """The ``args=`` tuple of sfm.py:268 (valid when the problem holds every point)."""
        return (self.n_cam, self.n_pts, self.camera_indices, self.point_indices, self.points_2d, self.K)


def _rodrigues(w):
    t2 = np.einsum("ij,ij->i", w, w)
    t = np.sqrt(t2)
    small = t < 1e-4
    ts = np.where(small, 1.0, t)
    s = np.where(small, 1.0 - t2 / 6.0, np.sin(ts) / ts)
    a = np.where(small, 0.5 - t2 / 24.0, 2.0 * np.sin(0.5 * ts) ** 2 / (ts * ts))
    z = np.zeros(len(w))
    Wx = np.stack([np.stack([z, -w[:, 2], w[:, 1]], -1),
                   np.stack([w[:, 2], z, -w[:, 0]], -1),
                   np.stack([-w[:, 1], w[:, 0], z], -1)], -2)
    return np.eye(3)[None] + s[:, None, None] * Wx + a[:, None, None] * (Wx @ Wx)


def camera_truth(n_cam, spacing=0.3):
    c = np.arange(n_cam, dtype=np.float64)
    T = np.stack([spacing * c, 0.2 * np.sin(0.37 * c), 0.2 * np.cos(0.23 * c) - 0.2], -1)
    w = 0.08 * np.stack([np.sin(0.31 * c), np.sin(0.17 * c + 1.0), np.sin(0.13 * c + 2.0)], -1)
    w[0] = 0.0
    T[0] = 0.0
    return np.concatenate([w, T], axis=1)


def point_degrees(n_pts, n_obs, window, seed):
    """Observations per point: >= 2, <= window, summing to n_obs exactly."""
    if n_obs < 2 * n_pts or n_obs > window * n_pts:
        raise ValueError("need 2*n_pts <= n_obs <= window*n_pts")
    rng = np.random.default_rng([seed, 0x6465])
    base, rem = divmod(n_obs, n_pts)
    k = np.full(n_pts, base, dtype=np.int64)
    if rem:
        k[rng.choice(n_pts, size=rem, replace=False)] += 1
    # move single observations between random point pairs to vary the degree, sum unchanged
    if base >= 3 and base + 2 <= window and n_pts >= 8:
        m = n_pts // 8
        perm = rng.permutation(n_pts)
        k[perm[:m]] -= 1
        k[perm[m:2 * m]] += 1
    assert int(k.sum()) == n_obs and k.min() >= 2 and k.max() <= window
    return k


def make_problem(n_cam, n_pts, n_obs, seed=0, window=16, point_range=None, int_pixels=False,
                 noise_px=0.5, cam_noise=2e-3, pt_noise=2e-2, K=None):
    """Build the problem (or only the shard of ``point_range``); see module docstring."""
    K = K_SCEAUX.copy() if K is None else np.asarray(K, dtype=np.float64)
    window = int(min(window, n_cam))
    cams = camera_truth(n_cam)
    R_all = _rodrigues(cams[:, :3])
    deg = point_degrees(n_pts, n_obs, window, seed)
    obs_off = np.concatenate([[0], np.cumsum(deg)])
    lo, hi = (0, n_pts) if point_range is None else point_range
    crng = np.random.default_rng([seed, 0x63616d])
    cam0 = cams + cam_noise * crng.standard_normal(cams.shape)
    cam0[0] = 0.0

    n_loc = int(obs_off[hi] - obs_off[lo])
    ci = np.empty(n_loc, dtype=np.int64)
    pi = np.empty(n_loc, dtype=np.int64)
    uv = np.empty((n_loc, 2), dtype=np.float64)
    X_true = np.empty((hi - lo, 3))
    X0 = np.empty((hi - lo, 3))
    n_start = max(n_cam - window + 1, 1)
    for b in range(lo // BLOCK, (hi + BLOCK - 1) // BLOCK if hi > lo else 0):
        p0, p1 = b * BLOCK, min((b + 1) * BLOCK, n_pts)
        rng = np.random.default_rng([seed, 1 + b])
        nb = p1 - p0
        keys = rng.random((nb, window))
        xyz = rng.random((nb, 3))
        noise = rng.standard_normal((nb, 3))
        # crop the block to the requested range (draws above cover the whole block)
        a0, a1 = max(p0, lo), min(p1, hi)
        sl = slice(a0 - p0, a1 - p0)
        pid = np.arange(a0, a1)
        kk = deg[a0:a1]
        start = (pid * n_start) // n_pts
        order = np.argsort(keys[sl], axis=1, kind="stable")          # draw order of distinct cameras
        centre = cams[np.minimum(start + window // 2, n_cam - 1), 3]
        X = np.stack([centre + 2.0 * xyz[sl, 0] - 1.0, 2.0 * xyz[sl, 1] - 1.0, 6.0 + 4.0 * xyz[sl, 2]], -1)
        X_true[a0 - lo:a1 - lo] = X
        X0[a0 - lo:a1 - lo] = X + pt_noise * noise[sl]
        kmax = int(kk.max()) if len(kk) else 0
        sel = np.arange(kmax)[None, :] < kk[:, None]
        cam_of = (start[:, None] + order[:, :kmax])[sel]
        pt_of = np.broadcast_to(pid[:, None], sel.shape)[sel]
        o0 = int(obs_off[a0] - obs_off[lo])
        o1 = int(obs_off[a1] - obs_off[lo])
        ci[o0:o1] = cam_of
        pi[o0:o1] = pt_of
        Y = X[pt_of - a0] - cams[cam_of, 3:6]
        Z = np.einsum("nij,nj->ni", R_all[cam_of], Y)
        assert Z[:, 2].min() > 0
        p = Z @ K.T
        orng = np.random.default_rng([seed, 0x6f6273, b])
        px_noise = noise_px * orng.standard_normal((nb * window, 2))
        first = int(obs_off[a0] - obs_off[p0])
        uv[o0:o1] = p[:, :2] / p[:, 2:3] + px_noise[first:first + (o1 - o0)]
    if int_pixels:
        uv = uv.astype(np.int64)                                      # graph.py:112-113 truncation
    x_true = np.concatenate([cams.ravel(), X_true.ravel()])
    x0 = np.concatenate([cam0.ravel(), X0.ravel()])
    return BAProblem(n_cam, n_pts, ci, pi, uv, K, x0, x_true, (lo, hi), n_obs)


def make_config(name, seed=0, **kw):
    n_cam, n_pts, n_obs = CONFIGS[name]
    return make_problem(n_cam, n_pts, n_obs, seed=seed, **kw)
