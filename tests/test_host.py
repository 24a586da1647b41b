"""CPU: host-side logic -- the C-ABI library loads and exports every declared symbol, generator and shard planner."""
import ctypes
import os
import re

import numpy as np
import pytest

from conftest import ROOT


def test_library_exports_every_header_symbol():
    import b200ba
    lib = b200ba.bundle_adjustment._lib.load()
    hdr = open(os.path.join(ROOT, "include", "b200ba.h")).read()
    names = set(re.findall(r"\b(b200ba_[a-z0-9_]+)\s*\(", hdr))
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), n
    assert names <= set(b200ba.bundle_adjustment._lib.SIGNATURES)
    assert lib.b200ba_version() >= 100


def test_struct_layouts_match_header():
    import b200ba
    L = b200ba.bundle_adjustment._lib
    assert ctypes.sizeof(L.Options) == 64 and ctypes.sizeof(L.Result) == 80 and ctypes.sizeof(L.Stats) == 192


def test_no_cpu_fallback_without_gpu():
    import torch
    import b200ba
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(b200ba.BAError):
        b200ba.Engine(0)
    with pytest.raises(TypeError):
        b200ba.least_squares(lambda x: x, np.zeros(3))


def test_synthetic_generator_contract():
    import b200ba
    syn = b200ba.synthetic
    pb = syn.make_config("B")
    assert (pb.n_cam, pb.n_pts, pb.n_obs) == (49, 7776, 31843)
    assert np.all(np.diff(pb.point_indices) >= 0)
    assert np.bincount(pb.point_indices, minlength=pb.n_pts).min() >= 2
    assert np.all(pb.x0[:6] == 0.0)                                  # camera 0 exactly identity (sfm.py:176)
    # no duplicated (camera, point) pair
    key = pb.point_indices * pb.n_cam + pb.camera_indices
    assert len(np.unique(key)) == len(key)
    # shard == slice of the full problem, bit for bit
    sh = syn.make_config("B", point_range=(1000, 3000))
    m = (pb.point_indices >= 1000) & (pb.point_indices < 3000)
    assert np.array_equal(sh.camera_indices, pb.camera_indices[m]) and np.array_equal(sh.points_2d, pb.points_2d[m])
    assert np.array_equal(sh.x0[6 * 49:], pb.x0[6 * 49 + 3000:6 * 49 + 9000])
    ip = syn.make_problem(5, 60, 200, seed=4, window=5, int_pixels=True)
    assert ip.points_2d.dtype == np.int64


def test_shard_plan_balanced_and_covering():
    import b200ba
    deg = b200ba.synthetic.point_degrees(4000, 20000, 16, 0)
    for world in (1, 2, 4, 8):
        ranges = b200ba.dist.plan_shards_from_degrees(deg, world)
        assert ranges[0][0] == 0 and ranges[-1][1] == 4000
        assert all(ranges[i][1] == ranges[i + 1][0] for i in range(world - 1))
        loads = [int(deg[a:b].sum()) for a, b in ranges]
        assert max(loads) - min(loads) <= int(deg.max()) * 2
    pi = np.repeat(np.arange(4000), deg)
    assert b200ba.dist.plan_shards(pi, 4000, 4) == b200ba.dist.plan_shards_from_degrees(deg, 4)


def test_local_problem_and_split_x():
    import b200ba
    pb = b200ba.synthetic.make_problem(6, 80, 300, seed=3, window=6)
    ci, pil, uv, sel = b200ba.dist.local_problem(pb.camera_indices, pb.point_indices, pb.points_2d, 20, 50)
    assert pil.min() == 0 and pil.max() == 29 and np.array_equal(pb.point_indices[sel] - 20, pil)
    xl = b200ba.dist.split_x(pb.x0, 6, 20, 50)
    assert len(xl) == 36 + 90 and np.array_equal(xl[36:], pb.x0[36 + 60:36 + 150])


def test_install_rebinds_names_and_cv2_proxy_delegates():
    """b200ba.install on a stand-in for the imported sfm_lite.sfm module: the three BA names are rebound; with
    geometry=True the module's cv2 name becomes a proxy that keeps every other cv2 attribute (no GPU call here)."""
    import types

    import b200ba
    fake_cv2 = types.SimpleNamespace(SIFT_create=lambda: "sift", NORM_L2=4, solvePnP=lambda *a, **k: "cv2", triangulatePoints=lambda *a: "cv2")
    mod = types.SimpleNamespace(least_squares=None, compute_residuals=None, create_sparsity_matrix=None, cv2=fake_cv2)
    b200ba.install(mod)
    assert mod.least_squares is b200ba.least_squares and mod.compute_residuals is b200ba.compute_residuals
    assert mod.create_sparsity_matrix is b200ba.create_sparsity_matrix and mod.cv2 is fake_cv2
    b200ba.install(mod, geometry=True)
    assert mod.cv2.SIFT_create() == "sift" and mod.cv2.NORM_L2 == 4
    assert mod.cv2.solvePnP is not fake_cv2.solvePnP and mod.cv2.triangulatePoints is not fake_cv2.triangulatePoints
    proxy = mod.cv2
    b200ba.install(mod, geometry=True)
    assert mod.cv2 is proxy            # idempotent
    # coplanar object points (cv2 starts from a homography there), fewer than 6 points, distortion: the real cv2 answers
    plane = np.c_[np.random.default_rng(0).uniform(-1, 1, (12, 2)), np.full(12, 3.0)]
    assert b200ba.bundle_adjustment.pnp_is_degenerate(plane) and b200ba.bundle_adjustment.pnp_is_degenerate(plane[:4])
    assert not b200ba.bundle_adjustment.pnp_is_degenerate(np.random.default_rng(1).uniform(-1, 1, (12, 3)))
    assert proxy.solvePnP(plane, np.zeros((12, 2)), np.eye(3), np.zeros((1, 5))) == "cv2"
    assert proxy.solvePnP(np.random.default_rng(1).uniform(-1, 1, (12, 3)), np.zeros((12, 2)), np.eye(3), np.ones((1, 5))) == "cv2"
    with pytest.raises(NotImplementedError):
        b200ba.solve_pnp(plane, np.zeros((12, 2)), np.eye(3))


def _ba_args(n=40):
    rng = np.random.default_rng(0)
    ci = rng.integers(0, 4, n)
    pi = np.sort(rng.integers(0, 10, n))
    return (4, 10, ci, pi, rng.uniform(0, 100, (n, 2)), np.eye(3))


def test_least_squares_refuses_what_it_would_otherwise_ignore():
    """Keywords that change scipy's answer are honoured or refused before any device work."""
    import b200ba
    args = _ba_args()
    x0 = np.zeros(6 * 4 + 3 * 10)
    f = b200ba.compute_residuals
    with pytest.raises(NotImplementedError):
        b200ba.least_squares(f, x0, args=args, bounds=(np.zeros_like(x0), np.ones_like(x0)))
    with pytest.raises(NotImplementedError):
        b200ba.least_squares(f, x0, args=args, jac=lambda x, *a: None)
    with pytest.raises(NotImplementedError):
        b200ba.least_squares(f, x0, args=args, x_scale=2.0)
    with pytest.raises(NotImplementedError):
        b200ba.least_squares(f, x0, args=args, kwargs={"a": 1})
    with pytest.raises(ValueError):
        b200ba.least_squares(f, x0, args=args, loss="tukey")
    with pytest.raises(ValueError):
        b200ba.least_squares(f, x0, args=args, method="newton")
    with pytest.raises(TypeError):
        b200ba.least_squares(f, x0, args=args, frobnicate=1)


def test_fixed_cameras_are_read_from_a_real_scipy_mask():
    """A mask built by the reference's create_sparsity_matrix(..., fixed_camera_indices) is a lil_matrix without the
    cameras' columns: the drop-in derives the fixed set from the all-zero column groups."""
    import ba_oracle as O
    import b200ba
    n_cam, n_pts, ci, pi, uv, K = _ba_args()
    B = b200ba.bundle_adjustment
    assert B._fixed_from_mask(O.sparsity(n_cam, n_pts, len(ci), ci, pi, (2,)).tolil(), n_cam) == (2,)
    assert B._fixed_from_mask(O.sparsity(n_cam, n_pts, len(ci), ci, pi).tolil(), n_cam) == ()
    assert B._fixed_from_mask(b200ba.create_sparsity_matrix(n_cam, n_pts, len(ci), ci, pi, (1, 3)), n_cam) == (1, 3)
    assert B._fixed_from_mask(None, n_cam) == ()
    half = O.sparsity(n_cam, n_pts, len(ci), ci, pi).tolil()
    half[:, 6] = 0
    with pytest.raises(NotImplementedError):
        B._fixed_from_mask(half, n_cam)


def test_problem_key_is_a_content_digest():
    """Two problems that differ only by swapped pixel rows (same sums) must not share a cached device problem."""
    import b200ba
    n_cam, n_pts, ci, pi, uv, K = _ba_args()
    key = b200ba.bundle_adjustment._problem_key
    uv2 = uv.copy()
    uv2[[0, 1]] = uv2[[1, 0]]
    assert key(n_cam, n_pts, ci, pi, uv, K) == key(n_cam, n_pts, ci.copy(), pi.copy(), uv.copy(), K)
    assert key(n_cam, n_pts, ci, pi, uv, K) != key(n_cam, n_pts, ci, pi, uv2, K)
    assert key(n_cam, n_pts, ci, pi, uv, K) != key(n_cam, n_pts, ci, pi, uv.astype(np.int64), K)


def test_problem_key_of_large_arrays_is_identity_plus_sample(monkeypatch):
    """Above the full-digest limit the key is the identity of the caller's arrays plus a sampled digest (a blake2b pass over
    config D's 640 MB takes ten times the solve): the same objects give the same key, copies do not, a change in a
    sampled element does, and the cache insists that the very objects are still alive."""
    import b200ba
    B = b200ba.bundle_adjustment
    monkeypatch.setattr(B, "_FULL_DIGEST_BYTES", 1024)
    n_cam, n_pts, ci, pi, uv, K = _ba_args()
    k1 = B._problem_key(n_cam, n_pts, ci, pi, uv, K)
    assert k1[2] == "identity" and k1 == B._problem_key(n_cam, n_pts, ci, pi, uv, K)
    assert k1 != B._problem_key(n_cam, n_pts, ci.copy(), pi, uv, K)
    uv[0, 0] += 1.0
    assert k1 != B._problem_key(n_cam, n_pts, ci, pi, uv, K)
    uv[0, 0] -= 1.0
    import weakref
    monkeypatch.setitem(B._cache, "refs", tuple(weakref.ref(a) for a in (ci, pi, uv)))
    assert B._key_alive(k1 + (1,), (ci, pi, uv))
    assert not B._key_alive(k1 + (1,), (ci.copy(), pi, uv))
    monkeypatch.setitem(B._cache, "refs", ())
    assert not B._key_alive(k1 + (1,), (ci, pi, uv))
    # lists cannot be identified: they fall back to the full digest
    assert B._problem_key(n_cam, n_pts, ci.tolist(), pi, uv, K)[2] == "digest"


def _import_real_sfm():
    """The reference's own sfm_lite.sfm, imported from /root/reference with matplotlib stubbed (absent here; the module
    only uses it for plots).  Exists in the build container only: the GPU box has no reference tree."""
    import sys
    import types
    ref = "/root/reference"
    if not os.path.isdir(os.path.join(ref, "sfm_lite")):
        pytest.skip("reference tree not present")
    if "matplotlib" not in sys.modules:
        try:
            import matplotlib  # noqa: F401
        except Exception:
            mpl = types.ModuleType("matplotlib")
            mpl.use = lambda *a, **k: None
            sys.modules["matplotlib"] = mpl
            sys.modules["matplotlib.pyplot"] = types.ModuleType("matplotlib.pyplot")
    if ref not in sys.path:
        sys.path.insert(0, ref)
    try:
        import sfm_lite.sfm as S
    except Exception as e:      # a missing third-party import of the reference is not this repo's failure
        pytest.skip(f"reference module does not import here: {e!r}")
    return S


def test_install_on_the_real_sfm_module():
    """b200ba.install on the REAL sfm_lite.sfm: the three names the BA call of sfm.py:264-268 resolves are ours, and
    SFM._apply_bundle_adjustment, run on a small stand-in graph, hands our least_squares exactly the reference's packing
    (x0 layout of sfm.py:255-257, dense camera ranks, point-major observation arrays, ftol = tol) and writes back what it
    returns (sfm.py:271-281)."""
    import types

    import b200ba
    S = _import_real_sfm()
    saved = (S.least_squares, S.compute_residuals, S.create_sparsity_matrix)
    try:
        b200ba.install(S)
        g = S.SFM._apply_bundle_adjustment.__globals__
        assert g["least_squares"] is b200ba.least_squares and g["compute_residuals"] is b200ba.compute_residuals
        assert g["create_sparsity_matrix"] is b200ba.create_sparsity_matrix
        # a stand-in for SFM.graph: 3 registered cameras (graph ids 5, 2, 9) + 1 unregistered, 4 points
        rng = np.random.default_rng(0)

        def pose(k):
            from scipy.spatial.transform import Rotation as Rot
            H = np.eye(4)
            H[:3, :3] = Rot.from_rotvec(0.1 * rng.standard_normal(3)).as_matrix()
            H[:3, 3] = rng.standard_normal(3)
            return H
        nodes = [types.SimpleNamespace(idx=i, registered=i != 7, H=pose(i)) for i in (5, 2, 7, 9)]
        obs = [(0, None, 5, None, np.array([10, 20])), (0, None, 2, None, np.array([11, 21])),
               (1, None, 2, None, np.array([12, 22])), (1, None, 9, None, np.array([13, 23])),
               (2, None, 5, None, np.array([14, 24])), (2, None, 9, None, np.array([15, 25])),
               (3, None, 5, None, np.array([16, 26])), (3, None, 2, None, np.array([17, 27]))]

        class Graph:
            def __init__(self):
                self.nodes = nodes
                self.X3d = rng.standard_normal((4, 3))

            def pt3ds_pt2ds(self):
                return iter(obs)

            def __getitem__(self, cam_id):
                return next(n for n in nodes if n.idx == cam_id)
        seen = {}

        def spy(fun, x0, **kw):
            seen.update(fun=fun, x0=np.array(x0), kw=kw)
            x = np.array(x0)
            x[-12:] += 1.0                   # "optimised" points
            return types.SimpleNamespace(x=x)
        S.least_squares = spy
        me = types.SimpleNamespace(graph=Graph(), K=np.eye(3))
        X_before = me.graph.X3d.copy()
        S.SFM._apply_bundle_adjustment(me, tol=1e-7, verbose=0)
        assert seen["fun"] is b200ba.compute_residuals
        n_cam, n_pts, ci, pi, uv, K = seen["kw"]["args"]
        assert (n_cam, n_pts) == (3, 4) and seen["kw"]["ftol"] == 1e-7 and seen["kw"]["x_scale"] == "jac"
        assert list(ci) == [0, 1, 1, 2, 0, 2, 0, 1] and list(pi) == [0, 0, 1, 1, 2, 2, 3, 3]       # dense ranks 5->0, 2->1, 9->2
        assert np.asarray(uv).shape == (8, 2) and len(seen["x0"]) == 6 * 3 + 3 * 4
        assert np.allclose(seen["x0"][3:6], nodes[0].H[:3, 3]) and np.allclose(seen["x0"][18:], X_before.ravel())
        mask = seen["kw"]["jac_sparsity"]
        assert isinstance(mask, b200ba.bundle_adjustment.JacobianSparsity) and mask.shape == (16, 30)
        assert np.allclose(me.graph.X3d, X_before + 1.0)                                            # write-back of sfm.py:281
    finally:
        S.least_squares, S.compute_residuals, S.create_sparsity_matrix = saved


def test_large_host_arrays_are_converted_in_threads_with_identical_values():
    """The reference's pixel array is int64 (graph.py:112-113): above 2 M elements the cast to float64 runs chunked on
    threads; values, shape and contiguity are those of np.ascontiguousarray."""
    import b200ba
    f = b200ba.bundle_adjustment._as_dtype
    uv = np.random.default_rng(3).integers(-5, 4000, size=(1_200_000, 2))
    out = f(uv, np.float64)
    assert out.dtype == np.float64 and out.shape == uv.shape and out.flags.c_contiguous
    assert np.array_equal(out, uv.astype(np.float64))
    assert f(uv.astype(np.float64), np.float64).dtype == np.float64
    idx = np.arange(3_000_000, dtype=np.int32)
    assert np.array_equal(f(idx, np.int64), idx.astype(np.int64)) and f(idx, np.int64).dtype == np.int64
    assert np.array_equal(f([[1, 2], [3, 4]], np.float64), np.array([[1.0, 2.0], [3.0, 4.0]]))
