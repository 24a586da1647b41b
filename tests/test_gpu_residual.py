"""GPU parity, stage 1: residual and analytic Jacobian blocks through the C ABI vs the reference's outputs."""
import numpy as np
import pytest

import ba_oracle as O
from conftest import golden_args, load_golden

pytestmark = pytest.mark.gpu

RES_FIXTURES = ["ref_residual_small.npz", "ref_residual_int_unsorted.npz", "ref_residual_pnp9.npz",
                "ref_residual_B2000.npz", "ref_residual_angles_skewK.npz"]


@pytest.mark.parametrize("name", RES_FIXTURES)
def test_residual_matches_reference_golden(engine, name):
    """bundle_adjustment.py:35-42 evaluated by the real reference (oracle/make_golden.py) vs b200ba_residuals; 1e-9 relative."""
    d = load_golden(name)
    args = golden_args(d)
    engine.set_problem(*args)
    r, cost = engine.residuals(d["x"], want_cost=True)
    ref = d["r_ref"]
    assert r.shape == ref.shape
    assert np.abs(r - ref).max() <= 1e-9 * np.abs(ref).max()
    # element-wise too, except where the depth is so small that the projection itself is ill-conditioned
    ok = np.abs(ref) < 1e4
    assert np.all(np.abs(r - ref)[ok] <= 1e-9 * np.maximum(1.0, np.abs(ref[ok])))
    assert abs(cost - 0.5 * float(ref @ ref)) <= 1e-9 * 0.5 * float(ref @ ref)


@pytest.mark.parametrize("k", range(9))
def test_residual_sceaux_dumps(engine, k):
    """Frozen SceauxCastle BA calls (sfm.py:266-268 arguments): residual at x0 vs the reference's own output."""
    d = load_golden(f"sceaux_ba_{k}.npz")
    args = golden_args(d)
    engine.set_problem(*args)
    r = engine.residuals(d["x0"])
    ref = d["r0_ref"]
    assert np.abs(r - ref).max() <= 1e-9 * np.abs(ref).max()


def test_module_level_compute_residuals_signature():
    """Same positional signature as the reference function; int64 pixels accepted."""
    import b200ba
    d = load_golden("ref_residual_int_unsorted.npz")
    assert d["uv"].dtype == np.int64
    r = b200ba.compute_residuals(d["x"], *golden_args(d))
    assert np.abs(r - d["r_ref"]).max() <= 1e-9 * np.abs(d["r_ref"]).max()


@pytest.mark.parametrize("name", ["ref_residual_small.npz", "ref_residual_int_unsorted.npz",
                                  "ref_residual_angles_skewK.npz"])
def test_jacobian_blocks_vs_oracle(engine, name):
    d = load_golden(name)
    args = golden_args(d)
    engine.set_problem(*args)
    Jc, Jp = engine.jacobian_blocks(d["x"])
    Jc_o, Jp_o = O.jacobian_blocks(d["x"], args[0], args[1], args[2], args[3], args[5])
    sc = np.abs(Jc_o).max(axis=(1, 2), keepdims=True)
    assert np.all(np.abs(Jc - Jc_o) <= 1e-9 * sc)
    sp = np.abs(Jp_o).max(axis=(1, 2), keepdims=True)
    assert np.all(np.abs(Jp - Jp_o) <= 1e-9 * sp)


def test_jacobian_vs_central_differences_of_gpu_residual(engine):
    """Analytic blocks vs central differences of the engine's own residual (theta = 0 camera included)."""
    d = load_golden("ref_residual_small.npz")
    args = golden_args(d)
    engine.set_problem(*args)
    x = d["x"].copy()
    Jc, Jp = engine.jacobian_blocks(x)
    ci, pi = args[2], args[3]
    h = 1e-6
    for col in list(range(0, 12)) + [6 * args[0] + k for k in range(6)]:
        e = np.zeros_like(x)
        e[col] = h
        fd = (engine.residuals(x + e) - engine.residuals(x - e)).reshape(-1, 2) / (2 * h)
        if col < 6 * args[0]:
            c, k = divmod(col, 6)
            sel = ci == c
            an = Jc[sel][:, :, k]
        else:
            p, k = divmod(col - 6 * args[0], 3)
            sel = pi == p
            an = Jp[sel][:, :, k]
        assert np.abs(fd[sel] - an).max() <= 1e-6 * max(1.0, np.abs(an).max())
        assert np.abs(fd[~sel]).max() == 0.0
