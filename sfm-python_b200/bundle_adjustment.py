"""Host-side mirror of the reference's bundle-adjustment call surface, backed by libb200ba.so.

Drop-in names (rebind the two imports of ``sfm_lite/sfm.py:8,20`` and
``SFM._apply_bundle_adjustment`` -- ``sfm.py:243-281`` -- runs unchanged):

* :func:`compute_residuals`       <-> ``sfm_lite/bundle_adjustment.py:35-42``
* :func:`create_sparsity_matrix`  <-> ``sfm_lite/bundle_adjustment.py:6-17``
* :func:`least_squares`           <-> ``scipy.optimize.least_squares`` as called at ``sfm.py:266-268``

plus :func:`bundle_adjust`, the direct entry.  All numerics run in the CUDA
library; this module only converts dtypes/layouts (``points_2d`` arrives as int64,
``graph.py:112-113``) and packs the result object.  No CPU fallback exists.
"""
from __future__ import annotations

import ctypes as C
import os
import weakref

import numpy as np

from . import _lib

LOSSES = {"linear": 0, "soft_l1": 1, "huber": 2, "cauchy": 3, "arctan": 4}

MESSAGES = {
    0: "The maximum number of function evaluations is exceeded.",
    1: "`gtol` termination condition is satisfied.",
    2: "`ftol` termination condition is satisfied.",
    3: "`xtol` termination condition is satisfied.",
    4: "Both `ftol` and `xtol` termination conditions are satisfied.",
}


class BAError(RuntimeError):
    pass


def _ptr(a):
    return C.c_void_p(a.ctypes.data) if a is not None else C.c_void_p(0)


def default_device():
    return int(os.environ.get("LOCAL_RANK", "0"))


class _PinnedPool:
    """Free list of page-locked host buffers (``b200ba_host_alloc``) handed out as numpy arrays that OWN them: when the
    last view of an array dies its buffer returns to the list, so a steady stream of solves allocates nothing and the
    result download runs at link speed.  Small results use ordinary memory."""

    MIN_BYTES = 1 << 20
    MAX_CACHED = 4

    def __init__(self, lib):
        self.lib, self.free = lib, {}

    def empty(self, n):
        nbytes = 8 * int(n)
        if nbytes < self.MIN_BYTES:
            return np.empty(int(n), dtype=np.float64)
        lst = self.free.get(nbytes)
        ptr = lst.pop() if lst else self.lib.b200ba_host_alloc(nbytes)
        if not ptr:
            return np.empty(int(n), dtype=np.float64)
        buf = (C.c_double * int(n)).from_address(ptr)
        weakref.finalize(buf, self._release, nbytes, ptr)
        return np.frombuffer(buf, dtype=np.float64)

    def _release(self, nbytes, ptr):
        lst = self.free.setdefault(nbytes, [])
        if len(lst) < self.MAX_CACHED:
            lst.append(ptr)
        else:
            self.lib.b200ba_host_free(C.c_void_p(ptr))


def _as_dtype(a, dtype):
    """``np.ascontiguousarray(a, dtype)``; a LARGE array of another dtype (the reference's int64 pixel array at the
    size of config D is 320 MB) is converted in chunks on a few threads -- numpy's cast loop releases the GIL and one
    thread runs it at 1-2 GB/s."""
    a = np.asarray(a)
    if a.dtype == dtype or a.size < (1 << 21) or not a.flags.c_contiguous:
        return np.ascontiguousarray(a, dtype=dtype)
    from concurrent.futures import ThreadPoolExecutor
    out = np.empty(a.shape, dtype=dtype)
    src, dst = a.reshape(-1), out.reshape(-1)
    n_thr = max(1, min(8, (os.cpu_count() or 2) // 2))
    step = -(-src.size // (4 * n_thr))
    with ThreadPoolExecutor(n_thr) as ex:
        list(ex.map(lambda k: np.copyto(dst[k:k + step], src[k:k + step], casting="unsafe"), range(0, src.size, step)))
    return out


class Engine:
    """One GPU's handle (``b200ba_handle``): problem upload, residual/Jacobian/normal equations, LM solve."""

    def __init__(self, device=None, devices=None):
        """``device``: one CUDA ordinal (default ``LOCAL_RANK`` or 0).  ``devices``: a list of ordinals -> ONE engine that
        shards every problem over those GPUs inside this process (``b200ba_create_multi``); ``set_problem`` / ``residuals`` /
        ``normal_equations`` / ``solve`` then take and return the full arrays exactly like the single-GPU engine."""
        self.lib = _lib.load()
        h = C.c_void_p()
        if devices is not None and len(devices) > 1:
            self.devices = tuple(int(d) for d in devices)
            self.device = self.devices[0]
            ids = (C.c_int * len(self.devices))(*self.devices)
            rc = self.lib.b200ba_create_multi(C.byref(h), ids, len(self.devices))
        else:
            self.device = default_device() if device is None else int(device)
            if devices is not None and len(devices) == 1:
                self.device = int(devices[0])
            self.devices = (self.device,)
            rc = self.lib.b200ba_create(C.byref(h), self.device)
        if rc != 0:
            raise BAError(f"b200ba_create(devices={list(self.devices)}) failed: {_lib.STATUS.get(rc, rc)} "
                          "(CUDA devices are required; there is no CPU fallback)")
        self.h = h
        self._pinned = _PinnedPool(self.lib)
        self.n_cam = self.n_pts = self.n_obs = 0
        self.fixed = ()
        self.loss = ("linear", 1.0)
        self.rank, self.nranks = 0, 1

    # -- lifetime
    def close(self):
        if getattr(self, "h", None):
            self.lib.b200ba_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            msg = self.lib.b200ba_last_error(self.h)
            msg = msg.decode() if msg else ""
            exc = ValueError if rc in (-1, -6) else BAError
            raise exc(f"libb200ba: {_lib.STATUS.get(rc, rc)}: {msg}")

    # -- multi-GPU
    @staticmethod
    def nccl_unique_id():
        buf = C.create_string_buffer(128)
        rc = _lib.load().b200ba_nccl_unique_id(buf)
        if rc != 0:
            raise BAError("ncclGetUniqueId failed")
        return bytes(buf.raw)

    def comm_init(self, unique_id: bytes, rank: int, nranks: int):
        buf = C.create_string_buffer(unique_id, 128)
        self._check(self.lib.b200ba_comm_init(self.h, buf, rank, nranks))
        self.rank, self.nranks = rank, nranks

    # -- problem
    def set_problem(self, n_cam, n_pts, camera_indices, point_indices, points_2d, K):
        ci = _as_dtype(camera_indices, np.int64).ravel()
        pi = _as_dtype(point_indices, np.int64).ravel()
        uv = _as_dtype(points_2d, np.float64)        # the reference's pixels are int64 (graph.py:112-113)
        if len(ci) != len(pi):
            raise ValueError("camera_indices and point_indices differ in length")
        if uv.size != 2 * len(ci):
            raise ValueError("points_2d must have shape (n_obs, 2)")
        Kc = np.ascontiguousarray(K, dtype=np.float64)
        if Kc.shape != (3, 3):
            raise ValueError("K must be 3x3")
        self._check(self.lib.b200ba_set_problem(self.h, int(n_cam), int(n_pts), len(ci), _ptr(ci), _ptr(pi), _ptr(uv),
                                                _ptr(Kc)))
        self.n_cam, self.n_pts, self.n_obs = int(n_cam), int(n_pts), len(ci)
        self.fixed = ()
        self.loss = ("linear", 1.0)
        return self

    @property
    def n_params(self):
        return 6 * self.n_cam + 3 * self.n_pts

    def _x(self, x):
        x = np.ascontiguousarray(x, dtype=np.float64).ravel()
        if x.size != self.n_params:
            raise ValueError(f"x has {x.size} entries, expected 6*{self.n_cam} + 3*{self.n_pts}")
        return x

    # -- hot-path pieces
    def residuals(self, x, want_cost=False):
        x = self._x(x)
        r = np.empty(2 * self.n_obs, dtype=np.float64)
        cost = C.c_double(0.0)
        self._check(self.lib.b200ba_residuals(self.h, _ptr(x), _ptr(r), C.byref(cost)))
        return (r, cost.value) if want_cost else r

    def cost(self, x):
        x = self._x(x)
        cost = C.c_double(0.0)
        self._check(self.lib.b200ba_residuals(self.h, _ptr(x), C.c_void_p(0), C.byref(cost)))
        return cost.value

    def jacobian_blocks(self, x):
        x = self._x(x)
        Jc = np.empty((self.n_obs, 2, 6))
        Jp = np.empty((self.n_obs, 2, 3))
        self._check(self.lib.b200ba_jacobian_blocks(self.h, _ptr(x), _ptr(Jc), _ptr(Jp)))
        return Jc, Jp

    def reproj_errors(self, Rt, X3d, want_errors=True):
        """Per-observation ``|| pi(K (R X + t)) - uv ||`` over the problem's observations (the loop of
        ``sfm.py:234-238``); ``Rt`` is ``(n_cam, 12)`` = R row-major then t, or ``(n_cam, 4, 4)`` / ``(n_cam, 3, 4)`` poses
        ``H``.  Returns ``(errors or None, sum)``."""
        Rt = pack_Rt(Rt)
        X = np.ascontiguousarray(X3d, dtype=np.float64).reshape(-1, 3)
        if Rt.shape[0] != self.n_cam or X.shape[0] != self.n_pts:
            raise ValueError(f"expected {self.n_cam} poses and {self.n_pts} points")
        err = np.empty(self.n_obs, dtype=np.float64) if want_errors else None
        tot = C.c_double(0.0)
        self._check(self.lib.b200ba_reproj_errors(self.h, _ptr(Rt), _ptr(X), _ptr(err), C.byref(tot)))
        return err, tot.value

    def pack_cameras(self, poses):
        """``(n, 4, 4)`` poses -> ``(n, 6)`` rows ``[rotvec | T]``: the per-camera conversion of ``sfm.py:252-257``
        (``Rot.from_matrix(R).as_rotvec()``, ``T = H[:3, 3]``) for all cameras in one call."""
        H = np.ascontiguousarray(poses, dtype=np.float64).reshape(-1, 16)
        out = np.empty((len(H), 6), dtype=np.float64)
        self._check(self.lib.b200ba_pack_cameras(self.h, len(H), _ptr(H), _ptr(out)))
        return out

    def unpack_cameras(self, params):
        """``(n, 6)`` rows ``[rotvec | T]`` -> ``(n, 4, 4)`` poses: ``sfm.py:273-278`` for all cameras in one call."""
        P = np.ascontiguousarray(params, dtype=np.float64).reshape(-1, 6)
        out = np.empty((len(P), 4, 4), dtype=np.float64)
        self._check(self.lib.b200ba_unpack_cameras(self.h, len(P), _ptr(P), _ptr(out)))
        return out

    def solve_pnp_batch(self, problems, K, guesses=None, max_iter=100):
        """Independent PnP problems ``[(points3d (n, 3), points2d (n, 2)), ...]`` in one launch; returns ``(m, 9)`` rows
        ``rvec(3) tvec(3) cost iterations status`` (``cv2.solvePnP`` of ``sfm.py:207-208``, ``R X + t`` convention)."""
        counts = np.asarray([len(p[0]) for p in problems], dtype=np.int64)
        X = np.ascontiguousarray(np.concatenate([np.asarray(p[0], dtype=np.float64).reshape(-1, 3) for p in problems]))
        uv = np.ascontiguousarray(np.concatenate([np.asarray(p[1], dtype=np.float64).reshape(-1, 2) for p in problems]))
        if len(X) != counts.sum() or len(uv) != counts.sum():
            raise ValueError("points3d and points2d differ in length")
        Kc = np.ascontiguousarray(K, dtype=np.float64)
        g = None if guesses is None else np.ascontiguousarray(guesses, dtype=np.float64).reshape(len(problems), 6)
        out = np.empty((len(problems), 9), dtype=np.float64)
        self._check(self.lib.b200ba_solve_pnp(self.h, len(problems), _ptr(counts), _ptr(X), _ptr(uv), _ptr(Kc), _ptr(g),
                                              int(max_iter), _ptr(out)))
        return out

    def triangulate(self, M1, M2, pts1, pts2, H1=None, H2=None):
        """DLT triangulation of ``sfm.py:216-217`` for ``(n, 2)`` point arrays; returns ``X (n, 4)`` rows ``[X Y Z 1]`` and,
        when the poses ``H1, H2`` are given, the visibility mask of ``sfm.py:219-221``."""
        M1 = np.ascontiguousarray(M1, dtype=np.float64)
        M2 = np.ascontiguousarray(M2, dtype=np.float64)
        p1 = np.ascontiguousarray(pts1, dtype=np.float64).reshape(-1, 2)
        p2 = np.ascontiguousarray(pts2, dtype=np.float64).reshape(-1, 2)
        if M1.shape != (3, 4) or M2.shape != (3, 4) or len(p1) != len(p2):
            raise ValueError("M1, M2 must be 3x4 and the point arrays of equal length")
        X = np.empty((len(p1), 4), dtype=np.float64)
        mask = np.empty(len(p1), dtype=np.uint8) if H1 is not None and H2 is not None else None
        h1 = np.ascontiguousarray(H1, dtype=np.float64) if mask is not None else None
        h2 = np.ascontiguousarray(H2, dtype=np.float64) if mask is not None else None
        self._check(self.lib.b200ba_triangulate(self.h, len(p1), _ptr(M1), _ptr(M2), _ptr(p1), _ptr(p2), _ptr(h1), _ptr(h2),
                                                _ptr(X), _ptr(mask)))
        return (X, mask.astype(bool)) if mask is not None else X

    def triangulate_nonlinear(self, M1, M2, pts1, pts2, max_iter=100):
        """``cv2_lite.triangulate_points.triangulate_points`` (:122-184) for ``(n, 2)`` point arrays: DLT + per-point
        refinement of the reference's objective; returns ``X (n, 4)`` rows ``[X Y Z 1]``."""
        M1 = np.ascontiguousarray(M1, dtype=np.float64)
        M2 = np.ascontiguousarray(M2, dtype=np.float64)
        p1 = np.ascontiguousarray(pts1, dtype=np.float64).reshape(-1, 2)
        p2 = np.ascontiguousarray(pts2, dtype=np.float64).reshape(-1, 2)
        if M1.shape != (3, 4) or M2.shape != (3, 4) or len(p1) != len(p2):
            raise ValueError("M1, M2 must be 3x4 and the point arrays of equal length")
        X = np.empty((len(p1), 4), dtype=np.float64)
        self._check(self.lib.b200ba_triangulate_nonlinear(self.h, len(p1), _ptr(M1), _ptr(M2), _ptr(p1), _ptr(p2),
                                                          int(max_iter), _ptr(X)))
        return X

    def normal_equations(self, x, lam, want_S=True):
        x = self._x(x)
        n = 6 * self.n_cam
        S = np.empty((n, n)) if want_S else None
        b = np.empty(n)
        g = np.empty(self.n_params)
        cost = C.c_double(0.0)
        self._check(self.lib.b200ba_normal_equations(self.h, _ptr(x), float(lam), _ptr(S), _ptr(b), _ptr(g),
                                                     C.byref(cost)))
        return dict(S=S, b=b, grad=g, cost=cost.value)

    def lm_step(self, x, lam, solver="auto"):
        x = self._x(x)
        d = np.empty(self.n_params)
        pred, newc = C.c_double(0.0), C.c_double(0.0)
        self._check(self.lib.b200ba_lm_step(self.h, _ptr(x), float(lam), _lib.SOLVERS[solver], _ptr(d), C.byref(pred),
                                            C.byref(newc)))
        return d, pred.value, newc.value

    def make_options(self, ftol=1e-8, xtol=1e-8, gtol=1e-8, max_nfev=None, verbose=0, solver="auto", lambda0=0.0,
                     pcg_rtol=0.0, pcg_max_iter=0, fixed_iters=0):
        o = _lib.Options()
        o.ftol, o.xtol, o.gtol = float(ftol), float(xtol), float(gtol)
        o.max_nfev = int(max_nfev) if max_nfev else 0
        o.verbose = int(verbose)
        o.solver = _lib.SOLVERS[solver]
        o.lambda0, o.pcg_rtol = float(lambda0), float(pcg_rtol)
        o.pcg_max_iter, o.fixed_iters = int(pcg_max_iter), int(fixed_iters)
        return o

    def solve(self, x0, **kw):
        """LM on host ``x0``; returns (x, Result)."""
        xin = self._x(x0)
        x = self._pinned.empty(xin.size)
        o = self.make_options(**kw)
        res = _lib.Result()
        self._check(self.lib.b200ba_solve_io(self.h, _ptr(xin), _ptr(x), C.byref(o), C.byref(res)))
        return x, res

    # -- device-pointer variants (torch tensors are only carriers of device memory)
    def solve_dev(self, x_dev_ptr, **kw):
        o = self.make_options(**kw)
        res = _lib.Result()
        self._check(self.lib.b200ba_solve_dev(self.h, C.c_void_p(int(x_dev_ptr)), C.byref(o), C.byref(res)))
        return res

    def linearize_dev(self, x_dev_ptr, lam):
        self._check(self.lib.b200ba_linearize_dev(self.h, C.c_void_p(int(x_dev_ptr) if x_dev_ptr else 0), float(lam)))

    def residuals_dev(self, x_dev_ptr, r_dev_ptr=0):
        cost = C.c_double(0.0)
        self._check(self.lib.b200ba_residuals_dev(self.h, C.c_void_p(int(x_dev_ptr)),
                                                  C.c_void_p(int(r_dev_ptr) if r_dev_ptr else 0), C.byref(cost)))
        return cost.value

    def residuals_noalloc(self, n_cam, n_pts, n_obs, cam_dev_ptr, pt_dev_ptr, uv_dev_ptr, K, x_dev_ptr, r_dev_ptr=0,
                          cost_dev_ptr=0):
        """``b200ba_residuals_noalloc``: residual + cost from DEVICE arrays (int32 indices) with the scratch in a static
        device array -- no allocation on the call path, no ``set_problem``; at most 4096 cameras (no reference
        counterpart; for callers that cannot call cudaMalloc)."""
        Kc = np.ascontiguousarray(K, dtype=np.float64)
        self._check(self.lib.b200ba_residuals_noalloc(
            self.h, int(n_cam), int(n_pts), int(n_obs), C.c_void_p(int(cam_dev_ptr) or None), C.c_void_p(int(pt_dev_ptr) or None),
            C.c_void_p(int(uv_dev_ptr) or None), _ptr(Kc), C.c_void_p(int(x_dev_ptr) or None),
            C.c_void_p(int(r_dev_ptr) or None), C.c_void_p(int(cost_dev_ptr) or None)))

    def jacobian_blocks_dev(self, x_dev_ptr, r_dev_ptr, jc_dev_ptr, jp_dev_ptr):
        self._check(self.lib.b200ba_jacobian_blocks_dev(self.h, C.c_void_p(int(x_dev_ptr)),
                                                        C.c_void_p(int(r_dev_ptr) if r_dev_ptr else 0),
                                                        C.c_void_p(int(jc_dev_ptr)), C.c_void_p(int(jp_dev_ptr))))

    # -- introspection
    def stream(self):
        return int(self.lib.b200ba_stream(self.h) or 0)

    def stats(self):
        s = _lib.Stats()
        self._check(self.lib.b200ba_get_stats(self.h, C.byref(s)))
        return {n: getattr(s, n) for n, _ in s._fields_}

    def reset_stats(self):
        self._check(self.lib.b200ba_reset_stats(self.h))

    def fp64_peak_tflops(self):
        v = C.c_double(0.0)
        self._check(self.lib.b200ba_fp64_peak(self.h, C.byref(v)))
        return v.value

    def set_profiling(self, on):
        self._check(self.lib.b200ba_set_profiling(self.h, int(bool(on))))

    def set_fixed_cameras(self, indices=()):
        """Hold these cameras fixed (``fixed_camera_indices`` of ``bundle_adjustment.py:6``); () frees all."""
        idx = np.ascontiguousarray(list(indices) if not isinstance(indices, np.ndarray) else indices, dtype=np.int64).ravel()
        self._check(self.lib.b200ba_set_fixed_cameras(self.h, _ptr(idx) if len(idx) else C.c_void_p(0), len(idx)))
        self.fixed = tuple(int(i) for i in idx)

    def set_loss(self, loss="linear", f_scale=1.0):
        """Robust loss of the next solves (scipy's ``loss=`` / ``f_scale=``); reset to linear by ``set_problem``."""
        if loss not in LOSSES:
            raise ValueError(f"`loss` must be one of {sorted(LOSSES)}")
        self._check(self.lib.b200ba_set_loss(self.h, LOSSES[loss], float(f_scale)))
        self.loss = (loss, float(f_scale))

    def set_two_level(self, mode):
        """PCG preconditioner: 0 block-Jacobi, 1 automatic (default), 2 two-level from the first solve."""
        self._check(self.lib.b200ba_set_two_level(self.h, int(mode)))

    def debug_fetch(self, name, count, dtype=np.float64):
        out = np.empty(int(count), dtype=dtype)
        n = self.lib.b200ba_debug_fetch(self.h, name.encode(), _ptr(out), out.nbytes)
        if n < 0:
            self._check(int(n))
        return out[:n]


def pack_Rt(poses):
    """``(n, 4, 4)`` / ``(n, 3, 4)`` poses ``H = [R | t]`` (``transforms.py:5-9``) or ``(n, 12)`` rows -> contiguous ``(n, 12)``
    ``[R row-major | t]``."""
    a = np.asarray(poses, dtype=np.float64)
    if a.ndim == 3 and a.shape[1:] in ((4, 4), (3, 4)):
        a = np.concatenate([a[:, :3, :3].reshape(-1, 9), a[:, :3, 3]], axis=1)
    if a.ndim != 2 or a.shape[1] != 12:
        raise ValueError("poses must be (n, 4, 4), (n, 3, 4) or (n, 12)")
    return np.ascontiguousarray(a)


# ---------------------------------------------------------------------------------------------
# Reference call surface
# ---------------------------------------------------------------------------------------------
_cache = {"key": None, "engine": None, "n_gpus": None, "refs": ()}


def default_n_gpus():
    """GPUs the drop-in call uses: ``B200BA_GPUS`` (default 1), so an unmodified ``sfm_lite/sfm.py`` reaches several GPUs."""
    try:
        return max(1, int(os.environ.get("B200BA_GPUS", "1")))
    except ValueError:
        return 1


_FULL_DIGEST_BYTES = 8 << 20     # above this a blake2b pass over the arrays costs more than the solve it would save


def _problem_key(n_cam, n_pts, ci, pi, uv, K):
    """Key of the cached device problem.  Up to 8 MB of index / pixel data: a content digest (blake2b over the bytes of
    the three arrays and K), so two different problems never share a key.  Above that the digest would dominate the call
    (blake2b runs at about 0.4 GB/s: 1.5 s for the 640 MB of config D, ten times the solve): the key is then the identity
    of the caller's arrays -- object, address, shape, strides, dtype -- plus a digest of 3 x 73 k sampled elements, and
    `engine_for` keeps weak references to the arrays so that a recycled id or address cannot match.  What this does not
    see is an in-place edit of unsampled elements between two calls; `b200ba.reset()` drops the cached problem."""
    import hashlib
    h = hashlib.blake2b(digest_size=16)
    big = [ci, pi, uv]
    total = sum(a.nbytes for a in big if isinstance(a, np.ndarray))
    if total <= _FULL_DIGEST_BYTES or not all(isinstance(a, np.ndarray) for a in big):
        for a, dt in ((ci, np.int64), (pi, np.int64), (uv, None), (K, np.float64)):
            a = np.ascontiguousarray(a) if dt is None else np.ascontiguousarray(a, dtype=dt)
            h.update(str((a.dtype.str, a.shape)).encode())
            h.update(memoryview(a).cast("B"))
        return (int(n_cam), int(n_pts), "digest", h.digest())
    ident = []
    for a in big:
        ident.append((id(a), a.__array_interface__["data"][0], a.shape, a.strides, a.dtype.str))
        flat = a.reshape(-1) if a.flags.c_contiguous else a.ravel()
        step = max(1, flat.size // 65536)
        for part in (flat[:4096], flat[-4096:], flat[::step]):
            h.update(np.ascontiguousarray(part).tobytes())
    h.update(np.ascontiguousarray(K, dtype=np.float64).tobytes())
    return (int(n_cam), int(n_pts), "identity", tuple(ident), h.digest())


def _key_alive(key, arrays):
    """An identity key only stands while the very arrays it was taken from are alive (ids and addresses get recycled)."""
    if key is None or key[2] != "identity":
        return True
    refs = _cache.get("refs") or ()
    return len(refs) == len(arrays) and all(r() is a for r, a in zip(refs, arrays))


def engine_for(n_cam, n_pts, camera_indices, point_indices, points_2d, K, n_gpus=None):
    """The engine holding this problem on the device(s) (uploaded once, reused by later calls).  ``n_gpus`` (default
    ``B200BA_GPUS`` or 1): the first n GPUs of the box inside this process."""
    import weakref
    n_gpus = default_n_gpus() if n_gpus is None else max(1, int(n_gpus))
    key = _problem_key(n_cam, n_pts, camera_indices, point_indices, points_2d, K) + (n_gpus,)
    arrays = (camera_indices, point_indices, points_2d)
    if _cache["engine"] is not None and _cache["n_gpus"] != n_gpus:
        _cache["engine"].close()
        _cache["engine"], _cache["key"] = None, None
    if _cache["key"] != key or _cache["engine"] is None or not _key_alive(_cache["key"], arrays):
        eng = _cache["engine"] or _new_engine(n_gpus)
        _cache["n_gpus"] = n_gpus
        _cache["key"] = None                       # a failed upload leaves no stale key behind
        _cache["engine"] = eng
        eng.set_problem(n_cam, n_pts, camera_indices, point_indices, points_2d, K)
        _cache["key"] = key
        _cache["refs"] = tuple(weakref.ref(a) for a in arrays) if key[2] == "identity" else ()
    return _cache["engine"]


def reset():
    """Drop the cached device problem (and its engine): the next call uploads again."""
    if _cache["engine"] is not None:
        _cache["engine"].close()
    _cache.update(key=None, engine=None, n_gpus=None, refs=())


def _new_engine(n_gpus=1):
    return Engine(devices=list(range(n_gpus))) if n_gpus > 1 else Engine()


def compute_residuals(x, n_cameras, n_points, camera_indices, point_indices, points_2d, K):
    """Same signature / output as ``sfm_lite/bundle_adjustment.py:35-42``: (2N,) float64, projected - observed."""
    return engine_for(n_cameras, n_points, camera_indices, point_indices, points_2d, K).residuals(x)


def _aux_engine():
    _reproj_engine["engine"] = _reproj_engine["engine"] or Engine()
    return _reproj_engine["engine"]


def pack_parameters(poses, X3d, engine=None):
    """``initial_guess`` of ``sfm.py:252-257``: ``[w_0 T_0 ... w_{C-1} T_{C-1} | X3d.ravel()]`` from the registered cameras'
    poses ``H`` (dense camera-index order) and the points."""
    cams = (engine or _aux_engine()).pack_cameras(poses)
    return np.concatenate([cams.ravel(), np.asarray(X3d, dtype=np.float64).ravel()])


def unpack_parameters(x, n_cam, engine=None):
    """Write-back of ``sfm.py:271-281``: ``result.x`` -> (poses ``(n_cam, 4, 4)``, ``X3d (n_points, 3)``)."""
    x = np.asarray(x, dtype=np.float64)
    Hs = (engine or _aux_engine()).unpack_cameras(x[:6 * n_cam])
    return Hs, x[6 * n_cam:].reshape(-1, 3)


def solve_pnp(point3ds, point2ds, K, dist_coeffs=None, engine=None, **kw):
    """``cv2.solvePnP(point3ds, point2ds, K, np.zeros((1, 5)))`` as called at ``sfm.py:207-208`` (and the reference's
    ``cv2_lite.solve_pnp.solve_pnp``): returns ``(success, rvec (3, 1), tvec (3, 1))``.  Distortion must be zero."""
    if dist_coeffs is not None and np.any(np.asarray(dist_coeffs) != 0):
        raise NotImplementedError("lens distortion is not part of the reference's pipeline (zeros at sfm.py:207)")
    if pnp_is_degenerate(point3ds):
        raise NotImplementedError("coplanar (or fewer than 6) object points: the 12-column DLT start is not unique there; "
                                  "cv2.solvePnP switches to a homography start, which this library does not implement")
    o = (engine or _aux_engine()).solve_pnp_batch([(point3ds, point2ds)], K)[0]
    return bool(o[8] >= 0 and np.isfinite(o[6])), o[:3].reshape(3, 1).copy(), o[3:6].reshape(3, 1).copy()


def pnp_is_degenerate(point3ds, rel_tol=1e-4):
    """True when the DLT start of :func:`solve_pnp` is not unique: fewer than 6 points, or object points that are
    (numerically) coplanar -- smallest singular value of the centred point cloud below ``rel_tol`` times the largest (the
    planarity test cv2.solvePnP applies before choosing its homography initialisation)."""
    X = np.asarray(point3ds, dtype=np.float64).reshape(-1, 3)
    if len(X) < 6:
        return True
    sv = np.linalg.svd(X - X.mean(axis=0), compute_uv=False)
    return bool(sv[2] <= rel_tol * max(sv[0], 1e-300))


def triangulate_points(M1, M2, pts1, pts2, engine=None):
    """``cv2.triangulatePoints(M1, M2, pts1, pts2)`` as called at ``sfm.py:216`` (``pts`` are ``(2, N)``), already divided by
    the last row (what ``sfm.py:217`` does next): returns ``(4, N)`` with a last row of ones."""
    p1, p2 = np.asarray(pts1, dtype=np.float64), np.asarray(pts2, dtype=np.float64)
    if p1.ndim != 2 or p1.shape[0] != 2 or p1.shape != p2.shape:
        raise ValueError("pts1, pts2 must be (2, N)")
    return np.ascontiguousarray((engine or _aux_engine()).triangulate(M1, M2, p1.T, p2.T).T)


def triangulate_points_nonlinear(M1, M2, pts1, pts2, engine=None, max_iter=100):
    """Same signature / layout as the reference's ``cv2_lite.triangulate_points.triangulate_points`` (:122-184): ``pts`` are
    ``(2, N)``, returns ``(4, N)`` with a last row of ones."""
    p1, p2 = np.asarray(pts1, dtype=np.float64), np.asarray(pts2, dtype=np.float64)
    if p1.ndim != 2 or p1.shape[0] != 2 or p1.shape != p2.shape:
        raise ValueError("pts1, pts2 must be (2, N)")
    return np.ascontiguousarray((engine or _aux_engine()).triangulate_nonlinear(M1, M2, p1.T, p2.T, max_iter).T)


def triangulate_visible(K, H1, H2, pts1, pts2, engine=None):
    """``sfm.py:214-223`` in one call: ``M = K @ H[:3]``, triangulation, visibility mask; ``pts`` are ``(N, 2)`` as returned by
    ``edge.pt2ds_pt2ds()``.  Returns ``(X3d[mask] (m, 3), mask (N,))``."""
    K = np.asarray(K, dtype=np.float64)
    H1, H2 = np.asarray(H1, dtype=np.float64), np.asarray(H2, dtype=np.float64)
    X, mask = (engine or _aux_engine()).triangulate(K @ H1[:3], K @ H2[:3], pts1, pts2, H1, H2)
    return X[mask, :3], mask


def gather_observations(track_point_idx, track_cam_id, track_xy, registered_cam_ids):
    """Vectorised ``Graph.pt3ds_pt2ds`` + the index remapping of ``sfm.py:248-256``: from flat track arrays (one row per
    ``(point, camera, x, y)`` track entry, point-major as ``graph.py:186-191`` iterates) keep the entries of registered
    cameras and map graph camera ids to dense ranks.  Returns ``(camera_indices, point_indices, points_2d)``."""
    pidx = np.asarray(track_point_idx, dtype=np.int64)
    cid = np.asarray(track_cam_id, dtype=np.int64)
    xy = np.asarray(track_xy)
    reg = np.asarray(sorted(int(c) for c in registered_cam_ids), dtype=np.int64)   # node order = ascending graph id
    pos = np.searchsorted(reg, cid)
    keep = (pos < len(reg)) & (reg[np.minimum(pos, len(reg) - 1)] == cid)
    return pos[keep], pidx[keep], xy[keep]


_reproj_engine = {"engine": None}


def calc_reproj_error(points3d, points2d, K, R, tvec):
    """Same signature / value as ``sfm_lite/sfm.py:38-41``: mean over the points of ``|| pi(K (R X + t)) - x ||``
    (``reproj_error``, ``cv2_lite/solve_pnp.py:9-14``) for ONE camera."""
    X = np.ascontiguousarray(points3d, dtype=np.float64).reshape(-1, 3)
    uv = np.ascontiguousarray(points2d, dtype=np.float64).reshape(-1, 2)
    n = len(X)
    if len(uv) != n:
        raise ValueError("points3d and points2d differ in length")
    eng = _aux_engine()
    eng.set_problem(1, n, np.zeros(n, dtype=np.int64), np.arange(n, dtype=np.int64), uv, K)
    Rt = np.concatenate([np.asarray(R, dtype=np.float64).reshape(9), np.asarray(tvec, dtype=np.float64).reshape(3)])[None]
    _, tot = eng.reproj_errors(Rt, X, want_errors=False)
    return tot / n


def total_reproj_error(K, poses, X3d, camera_indices, point_indices, points_2d, engine=None):
    """The whole loop of ``sfm.py:233-238`` in one call: returns ``(errs, n)`` with ``errs`` the sum over all observations
    of the per-observation error (each ``calc_reproj_error`` call there sees one point, so its mean is the norm) and
    ``n`` the observation count; ``poses`` are the registered cameras' ``H`` in dense camera-index order."""
    Rt = pack_Rt(poses)
    X = np.ascontiguousarray(X3d, dtype=np.float64).reshape(-1, 3)
    eng = engine or engine_for(len(Rt), len(X), camera_indices, point_indices, points_2d, K)
    _, tot = eng.reproj_errors(Rt, X, want_errors=False)
    return tot, eng.n_obs


class JacobianSparsity:
    """Lazy stand-in for the ``lil_matrix`` of ``bundle_adjustment.py:6-17``.

    The engine uses analytic blocks and never reads the mask; it is only materialised (vectorised, no
    Python loop) if a caller asks for it (``tocsr`` / ``tolil`` / ``toarray``)."""

    def __init__(self, n_cameras, n_points, n_obs, camera_indices, point3d_indices, fixed_camera_indices=()):
        if len(camera_indices) != len(point3d_indices):
            raise AssertionError("camera_indices and point3d_indices differ in length")
        self.n_cameras, self.n_points, self.n_obs = int(n_cameras), int(n_points), int(n_obs)
        self.camera_indices = np.asarray(camera_indices, dtype=np.int64)
        self.point_indices = np.asarray(point3d_indices, dtype=np.int64)
        self.fixed = np.asarray(list(fixed_camera_indices), dtype=np.int64)
        self.shape = (2 * self.n_obs, 6 * self.n_cameras + 3 * self.n_points)
        self.dtype = np.dtype(int)

    @property
    def nnz(self):
        free = ~np.isin(self.camera_indices, self.fixed)
        return int(12 * free.sum() + 6 * len(self.camera_indices))

    def tocsr(self):
        from scipy.sparse import csr_matrix
        ci, pi = self.camera_indices, self.point_indices
        obs = np.arange(len(ci), dtype=np.int64)
        free = ~np.isin(ci, self.fixed)
        rows, cols = [], []
        for r in range(2):
            for k in range(6):
                rows.append(2 * obs[free] + r)
                cols.append(6 * ci[free] + k)
            for k in range(3):
                rows.append(2 * obs + r)
                cols.append(6 * self.n_cameras + 3 * pi + k)
        rows, cols = np.concatenate(rows), np.concatenate(cols)
        return csr_matrix((np.ones(len(rows), dtype=int), (rows, cols)), shape=self.shape)

    def tolil(self):
        return self.tocsr().tolil()

    def tocoo(self):
        return self.tocsr().tocoo()

    def toarray(self):
        return self.tocsr().toarray()


def create_sparsity_matrix(n_cameras, n_points, n_obs, camera_indices, point3d_indices, fixed_camera_indices=()):
    """Same signature as ``bundle_adjustment.py:6-17``; returns a lazy mask (see :class:`JacobianSparsity`)."""
    return JacobianSparsity(n_cameras, n_points, n_obs, camera_indices, point3d_indices, fixed_camera_indices)


class BAResult(dict):
    """OptimizeResult-like: ``x cost optimality nfev njev status message success`` are set eagerly;
    ``fun`` (2N residuals) and ``grad`` (J^T r) are computed by the engine on first access."""

    def __getattr__(self, name):
        if name in self:
            return self[name]
        if name in ("fun", "grad") and "_engine" in self.__dict__:
            eng = self.__dict__["_engine"]
            if name == "fun":
                self["fun"] = eng.residuals(self["x"])
            else:
                self["grad"] = eng.normal_equations(self["x"], 0.0, want_S=False)["grad"]
            return self[name]
        raise AttributeError(name)

    __setattr__ = dict.__setitem__

    def __repr__(self):
        keys = [k for k in self if k != "x"]
        return "BAResult(" + ", ".join(f"{k}={self[k]!r}" for k in keys) + ")"


def _is_ba_residual(fun):
    name = getattr(fun, "__name__", "")
    return fun is compute_residuals or name in ("compute_residuals", "residuals", "residuals_loop")


def bundle_adjust(x0, n_cam, n_pts, camera_indices, point_indices, points_2d, K, ftol=1e-10, xtol=1e-8, gtol=1e-8,
                  max_nfev=None, verbose=0, solver="auto", engine=None, fixed_camera_indices=(), loss="linear", f_scale=1.0,
                  n_gpus=None, **engine_opts):
    """Global bundle adjustment of ``x0`` (layout of ``sfm.py:255-257``) on the GPU; returns :class:`BAResult`.

    ``n_gpus``: GPUs of this process to shard the points over (default ``B200BA_GPUS`` or 1).
    ``fixed_camera_indices``: cameras held fixed, the hook of ``bundle_adjustment.py:6,13``; ``loss`` / ``f_scale``: scipy's
    robust-loss keywords (``result.cost`` is then ``0.5 * f_scale**2 * sum(rho((fun / f_scale)**2))`` like scipy's)."""
    eng = engine or engine_for(n_cam, n_pts, camera_indices, point_indices, points_2d, K, n_gpus)
    fixed = tuple(sorted(int(i) for i in fixed_camera_indices))
    if fixed != tuple(sorted(getattr(eng, "fixed", ()))):
        eng.set_fixed_cameras(fixed)
    if (loss, float(f_scale)) != getattr(eng, "loss", ("linear", 1.0)):
        eng.set_loss(loss, f_scale)
    x, r = eng.solve(x0, ftol=ftol, xtol=xtol, gtol=gtol, max_nfev=max_nfev, verbose=verbose, solver=solver,
                     **engine_opts)
    res = BAResult(x=x, cost=r.cost, optimality=r.optimality, nfev=int(r.nfev), njev=int(r.njev), status=int(r.status),
                   message=MESSAGES.get(int(r.status), ""), success=int(r.status) > 0, rms=r.rms, n_iter=int(r.n_iter),
                   initial_cost=r.initial_cost, pcg_iters=int(r.pcg_iters), solve_ms=r.solve_ms,
                   lambda_final=r.lambda_final)
    res.__dict__["_engine"] = eng
    return res


def least_squares(fun, x0, jac="2-point", bounds=None, method="trf", ftol=1e-8, xtol=1e-8, gtol=1e-8, x_scale="jac",
                  loss="linear", f_scale=1.0, diff_step=None, tr_solver=None, tr_options=None, jac_sparsity=None,
                  max_nfev=None, verbose=0, args=(), kwargs=None, **extra):
    """``scipy.optimize.least_squares``-shaped entry for the call at ``sfm.py:266-268``.

    ``fun`` must be the BA residual (ours or the reference's) and ``args`` the 6-tuple of ``sfm.py:268``;
    ``jac_sparsity``, ``x_scale='jac'`` and ``method='trf'`` are accepted and mapped to analytic blocks,
    Marquardt scaling and LM.  ``ftol`` keeps the meaning of ``--ba_tol`` (``main.py:36``)."""
    if not _is_ba_residual(fun) or len(args) != 6:
        raise TypeError("b200ba.least_squares only solves the bundle-adjustment problem of sfm_lite: pass "
                        "compute_residuals and args=(n_cam, n_points, camera_indices, point_indices, points_2d, K)")
    n_cam, n_pts, ci, pi, uv, K = args
    solver = extra.pop("solver", "auto")
    n_gpus = extra.pop("n_gpus", None)
    # keywords that would change scipy's answer are honoured or refused, never ignored
    if kwargs:
        raise NotImplementedError("`kwargs` for the residual function are not supported (the BA residual takes none)")
    if callable(jac) or jac not in ("2-point", "3-point", "cs"):
        raise NotImplementedError("a user Jacobian is not supported: the engine always uses its analytic blocks")
    if method not in ("trf", "lm", "dogbox"):
        raise ValueError("`method` must be 'trf', 'dogbox' or 'lm'")
    if bounds is not None:
        lo, hi = bounds
        if np.any(np.asarray(lo, dtype=np.float64) > -np.inf) or np.any(np.asarray(hi, dtype=np.float64) < np.inf):
            raise NotImplementedError("bounds are not supported (the reference call at sfm.py:266 passes none)")
    if not (isinstance(x_scale, str) and x_scale == "jac"):
        raise NotImplementedError("only x_scale='jac' (the reference's setting; Marquardt scaling) is implemented")
    if tr_solver not in (None, "lsmr", "exact"):
        raise ValueError("`tr_solver` must be None, 'exact' or 'lsmr'")
    if extra:
        raise TypeError(f"unexpected keyword arguments {sorted(extra)}")
    if loss not in LOSSES:
        raise ValueError(f"`loss` must be one of {sorted(LOSSES)}")
    return bundle_adjust(x0, n_cam, n_pts, ci, pi, uv, K, ftol=ftol, xtol=xtol, gtol=gtol, max_nfev=max_nfev,
                         verbose=verbose, solver=solver, fixed_camera_indices=_fixed_from_mask(jac_sparsity, n_cam),
                         loss=loss, f_scale=f_scale, n_gpus=n_gpus)


def _fixed_from_mask(mask, n_cam):
    """Cameras whose six columns are absent from ``jac_sparsity``: the lazy mask carries the set; for a real scipy mask
    (``create_sparsity_matrix`` of the reference returns a ``lil_matrix``) the all-zero camera column groups are read."""
    if mask is None:
        return ()
    if hasattr(mask, "fixed"):
        return tuple(int(i) for i in mask.fixed)
    try:
        from scipy.sparse import issparse
        m = mask.tocsc() if issparse(mask) else None
    except Exception:
        m = None
    if m is None:
        m = np.asarray(mask)
        nnz_col = (m != 0).sum(axis=0)
    else:
        nnz_col = np.diff(m.indptr)
    cam_cols = np.asarray(nnz_col).ravel()[:6 * int(n_cam)].reshape(int(n_cam), 6)
    zero = (cam_cols == 0).all(axis=1)
    partial = (cam_cols == 0).any(axis=1) & ~zero
    if partial.any():
        raise NotImplementedError("jac_sparsity masks single camera parameters; only whole cameras can be held fixed")
    return tuple(int(i) for i in np.flatnonzero(zero))


class _Cv2Proxy:
    """Stands in for the ``cv2`` name inside ``sfm_lite.sfm``: ``solvePnP`` (``sfm.py:207-208``) and ``triangulatePoints``
    (``sfm.py:216``) go to the GPU library, every other attribute (SIFT, matcher, ``findFundamentalMat`` ...) to the real
    module."""

    def __init__(self, real):
        self._real = real

    def __getattr__(self, name):
        return getattr(self._real, name)

    def solvePnP(self, objectPoints, imagePoints, cameraMatrix, distCoeffs=None, *a, **kw):
        X = np.asarray(objectPoints, dtype=np.float64).reshape(-1, 3)
        if a or kw or pnp_is_degenerate(X) or (distCoeffs is not None and np.any(np.asarray(distCoeffs) != 0)):
            # planar / tiny point sets, distortion, extrinsic guesses and solver flags stay with the real cv2
            return self._real.solvePnP(objectPoints, imagePoints, cameraMatrix, distCoeffs, *a, **kw)
        return solve_pnp(X, np.asarray(imagePoints, dtype=np.float64).reshape(-1, 2), cameraMatrix, distCoeffs)

    @staticmethod
    def triangulatePoints(projMatr1, projMatr2, projPoints1, projPoints2):
        return triangulate_points(projMatr1, projMatr2, projPoints1, projPoints2)


def install(sfm_module, geometry=False):
    """Rebind the three names inside an imported ``sfm_lite.sfm`` module.  ``geometry=True`` also routes
    the module's ``cv2.solvePnP`` / ``cv2.triangulatePoints`` calls (``sfm.py:207-216``) to the library."""
    sfm_module.least_squares = least_squares
    sfm_module.compute_residuals = compute_residuals
    sfm_module.create_sparsity_matrix = create_sparsity_matrix
    if geometry and hasattr(sfm_module, "cv2") and not isinstance(sfm_module.cv2, _Cv2Proxy):
        sfm_module.cv2 = _Cv2Proxy(sfm_module.cv2)
    return sfm_module
