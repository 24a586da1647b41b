"""GPU parity, stage 2: J^T J / J^T r blocks, point-block Schur complement, one LM step, vs the numpy oracle."""
import numpy as np
import pytest

import ba_oracle as O
from conftest import golden_args, load_golden

pytestmark = pytest.mark.gpu


def _small():
    import b200ba
    return b200ba.synthetic.make_problem(7, 120, 480, seed=5, window=6)


@pytest.mark.parametrize("lam", [0.0, 1e-4, 3.0])
def test_schur_system_vs_oracle(engine, lam):
    pb = _small()
    engine.set_problem(*pb.args())
    out = engine.normal_equations(pb.x0, lam)
    ref = O.schur_system(pb.x0, *pb.args(), lam)
    g_ref = np.concatenate([ref["gc"].ravel(), ref["gp"].ravel()])
    assert abs(out["cost"] - ref["cost"]) <= 1e-12 * ref["cost"]
    assert np.abs(out["grad"] - g_ref).max() <= 1e-10 * np.abs(g_ref).max()
    assert np.abs(out["b"] - ref["b"]).max() <= 1e-9 * np.abs(ref["b"]).max()
    S, Sr = out["S"], ref["S"]
    assert np.abs(S - S.T).max() <= 1e-9 * np.abs(Sr).max()
    assert np.abs(S - Sr).max() <= 1e-9 * np.abs(Sr).max()


def test_schur_system_unsorted_and_duplicate_camera(engine):
    """Unsorted observation order, a point seen twice by one camera, a point with one observation, an unobserved point
    and an unobserved camera."""
    pb = _small()
    rng = np.random.default_rng(0)
    ci = pb.camera_indices.copy()
    pi = pb.point_indices.copy()
    uv = pb.points_2d.copy()
    ci[1] = ci[0]                       # point 0 seen twice by the same camera
    keep = np.ones(len(ci), bool)
    p5 = np.flatnonzero(pi == 5)
    keep[p5[1:]] = False                # point 5: single observation
    keep[pi == 9] = False               # point 9: no observation
    keep[ci == 3] = False               # camera 3: no observation
    ci, pi, uv = ci[keep], pi[keep], uv[keep]
    perm = rng.permutation(len(ci))
    ci, pi, uv = ci[perm], pi[perm], uv[perm]
    args = (pb.n_cam, pb.n_pts, ci, pi, uv, pb.K)
    engine.set_problem(*args)
    r = engine.residuals(pb.x0)
    r_ref = O.residuals(pb.x0, *args)
    assert np.abs(r - r_ref).max() <= 1e-9 * np.abs(r_ref).max()
    lam = 1e-3
    out = engine.normal_equations(pb.x0, lam)
    ref = O.schur_system(pb.x0, *args, lam)
    Sr = ref["S"].copy()
    live = np.ones(6 * pb.n_cam, bool)
    live[18:24] = False
    assert np.abs(out["S"][np.ix_(live, live)] - Sr[np.ix_(live, live)]).max() <= 1e-9 * np.abs(Sr).max()
    assert np.allclose(out["S"][18:24, 18:24], np.eye(6))           # unobserved camera: unit block
    assert np.abs(out["b"] - ref["b"]).max() <= 1e-9 * np.abs(ref["b"]).max()


@pytest.mark.parametrize("solver", ["cholesky", "pcg"])
def test_lm_step_vs_oracle(engine, solver):
    pb = _small()
    engine.set_problem(*pb.args())
    lam = 1e-3
    d, pred, newc = engine.lm_step(pb.x0, lam, solver=solver)
    d_ref, pred_ref, _ = O.lm_step(pb.x0, *pb.args(), lam)
    assert np.abs(d - d_ref).max() <= 1e-6 * np.abs(d_ref).max()
    assert abs(pred - pred_ref) <= 1e-8 * abs(pred_ref)
    r_new = O.residuals(pb.x0 + d, *pb.args())
    assert abs(newc - 0.5 * float(r_new @ r_new)) <= 1e-9 * newc


def test_big_point_path(engine):
    """A point observed by more than 32 cameras takes the multi-pass tile path."""
    import b200ba
    pb = b200ba.synthetic.make_problem(40, 30, 30 * 36, seed=2, window=40)
    engine.set_problem(*pb.args())
    lam = 1e-4
    out = engine.normal_equations(pb.x0, lam)
    ref = O.schur_system(pb.x0, *pb.args(), lam)
    assert np.abs(out["S"] - ref["S"]).max() <= 1e-9 * np.abs(ref["S"]).max()
    d, pred, newc = engine.lm_step(pb.x0, lam, solver="cholesky")
    d_ref, pred_ref, _ = O.lm_step(pb.x0, *pb.args(), lam)
    assert np.abs(d - d_ref).max() <= 1e-6 * np.abs(d_ref).max()


@pytest.mark.parametrize("tile_obs", [32, 64, 640, 1312])
def test_fused_matches_unfused_and_oracle(tile_obs, monkeypatch):
    """k_lin_fused (records in shared memory) against the unfused kernels and the oracle, several tile sizes,
    including points seen twice by one camera and tiles that cut the camera windows."""
    import b200ba
    pb = b200ba.synthetic.make_problem(23, 900, 4000, seed=9, window=7)
    ci = pb.camera_indices.copy()
    ci[1] = ci[0]
    ci[401] = ci[400]
    args = (pb.n_cam, pb.n_pts, ci, pb.point_indices, pb.points_2d, pb.K)
    lam = 2e-4
    ref = O.schur_system(pb.x0, *args, lam)
    g_ref = np.concatenate([ref["gc"].ravel(), ref["gp"].ravel()])
    outs = {}
    for fused, flow in (("1", "1"), ("1", "0"), ("0", "0")):
        monkeypatch.setenv("B200BA_FUSED", fused)
        monkeypatch.setenv("B200BA_FLOW", flow)
        monkeypatch.setenv("B200BA_TILE_OBS", str(tile_obs))
        eng = b200ba.Engine(0)
        eng.set_problem(*args)
        out = eng.normal_equations(pb.x0, lam)
        st = eng.stats()
        assert (st["n_self_slots"] > 0) == (fused == "1")
        outs[fused + flow] = out
        assert abs(out["cost"] - ref["cost"]) <= 1e-12 * ref["cost"]
        assert np.abs(out["grad"] - g_ref).max() <= 1e-10 * np.abs(g_ref).max()
        assert np.abs(out["b"] - ref["b"]).max() <= 1e-9 * np.abs(ref["b"]).max()
        assert np.abs(out["S"] - ref["S"]).max() <= 1e-9 * np.abs(ref["S"]).max()
        eng.close()
    assert np.abs(outs["10"]["S"] - outs["00"]["S"]).max() <= 1e-11 * np.abs(ref["S"]).max()
    if tile_obs <= 512:   # the dataflow kernel ran: same slot partials as the barrier version, bit for bit
        assert np.array_equal(outs["11"]["S"], outs["10"]["S"]) and np.array_equal(outs["11"]["b"], outs["10"]["b"])


def test_fused_is_bit_reproducible():
    import b200ba
    pb = b200ba.synthetic.make_problem(40, 3000, 13000, seed=3, window=9)
    eng = b200ba.Engine(0)
    eng.set_problem(*pb.args())
    a = eng.normal_equations(pb.x0, 1e-4)
    b = eng.normal_equations(pb.x0, 1e-4)
    assert np.array_equal(a["S"], b["S"]) and np.array_equal(a["b"], b["b"]) and a["cost"] == b["cost"]
    eng.close()


def test_device_tile_construction_equals_the_greedy_loop(engine):
    """Warp tiles are built on the device by pointer doubling (ba_setup.cu greedy_bins); they must equal the sequential
    greedy rule 'close the tile when the next point would push it past 32 observations, a larger point is its own
    tile' on a ragged problem: degrees 1..70, runs of unobserved points, one camera."""
    rng = np.random.default_rng(12)
    n_pts = 5000
    deg = rng.integers(1, 9, n_pts)
    deg[rng.choice(n_pts, 40, replace=False)] = rng.integers(33, 71, 40)      # points with more than a warp of observations
    deg[rng.choice(n_pts, 300, replace=False)] = 0                              # unobserved points
    deg[:3] = 0
    deg[-2:] = 0
    pi = np.repeat(np.arange(n_pts), deg)
    n = len(pi)
    ci = rng.integers(0, 6, n)
    uv = rng.uniform(0, 2000, (n, 2))
    K = np.array([[2900.0, 0, 1400], [0, 2900, 1000], [0, 0, 1]])
    engine.set_problem(6, n_pts, ci, pi, uv, K)
    off = np.concatenate([[0], np.cumsum(deg)])
    wt, start = [0], 0
    for p in range(n_pts):                                   # the rule, literally
        e = off[p + 1]
        if e - start > 32:
            if off[p] > start:
                wt.append(off[p])
            start = wt[-1]
            if e - start > 32:
                wt.append(e)
                start = e
    if n > start:
        wt.append(n)
    got = engine.debug_fetch("wt_off", len(wt) + 8, dtype=np.int32)
    assert len(got) == len(wt) and np.array_equal(got, np.asarray(wt))
    # and the residual on this ragged layout still matches the oracle
    x = np.concatenate([rng.normal(0, 0.05, 36), rng.uniform(-1, 1, (n_pts, 2)).ravel()[:0]])
    cams = rng.normal(0, 0.05, (6, 6))
    pts = np.stack([rng.uniform(-1, 1, n_pts), rng.uniform(-1, 1, n_pts), rng.uniform(5, 9, n_pts)], 1)
    x = np.concatenate([cams.ravel(), pts.ravel()])
    r = engine.residuals(x)
    r_ref = O.residuals(x, 6, n_pts, ci, pi, uv, K)
    assert np.abs(r - r_ref).max() <= 1e-9 * np.abs(r_ref).max()
