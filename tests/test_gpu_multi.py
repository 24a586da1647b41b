"""Multi-GPU parity: the point-sharded reduced camera system and LM equal the unsharded ones.

Needs >= 2 GPUs on the box; skipped otherwise (the host-side shard logic is covered on CPU by test_dist_gloo.py).
"""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.gpu
@pytest.mark.parametrize("xchg", ["0", "1", "p2p"])
@pytest.mark.parametrize("shape", [(40, 4000, 16000), (300, 30000, 120000)])
def test_sharded_equals_unsharded(xchg, shape):
    """xchg: 0 = all-reduce of the packed buffer, 1 = sparse exchange through ncclAllGather, p2p = sparse exchange through
    peer memory (CUDA-IPC mapped receive slots, k_p2p_push / k_p2p_wait) -- all three must equal the unsharded system."""
    n = _gpus()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    world = 4 if n >= 4 else 2
    env = dict(os.environ, B200BA_XCHG="1" if xchg == "p2p" else xchg, B200BA_P2P="1" if xchg == "p2p" else "0")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "tests", "_multi_worker.py")] + [str(v) for v in shape]
    p = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    assert p.stdout.count(" OK") == world, p.stdout


@pytest.mark.gpu
def test_single_process_multi_gpu_dropin():
    """Row b of the scope table: the drop-in call of sfm.py:264-268 reaching several GPUs from ONE process
    (b200ba_create_multi behind Engine(devices=...) / least_squares(..., n_gpus=) / B200BA_GPUS): same converged cost as
    one GPU to 1e-12, same residuals and gradient, point-major and shuffled observation order, fixed cameras, robust loss."""
    n = _gpus()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    import numpy as np

    import ba_oracle as O
    import b200ba
    world = 4 if n >= 4 else 2
    pb = b200ba.synthetic.make_problem(300, 30000, 120000, seed=3)
    args = pb.args()
    one = b200ba.Engine(0)
    one.set_problem(*args)
    x1, r1 = one.solve(pb.x0, ftol=1e-10, solver="pcg")
    res1 = one.residuals(pb.x0)
    ne1 = one.normal_equations(pb.x0, 1e-3)
    # through the scipy-shaped drop-in
    mask = b200ba.create_sparsity_matrix(pb.n_cam, pb.n_pts, pb.n_obs, pb.camera_indices, pb.point_indices)
    rm = b200ba.least_squares(b200ba.compute_residuals, pb.x0, jac_sparsity=mask, verbose=0, x_scale='jac', ftol=1e-10,
                              method='trf', args=args, solver="pcg", n_gpus=world)
    assert rm.status == r1.status and abs(rm.cost - r1.cost) <= 1e-12 * r1.cost, (rm.cost, r1.cost)
    assert np.abs(rm.x - x1).max() <= 1e-9 * np.abs(x1).max()
    chk = 0.5 * float(np.sum(O.residuals(rm.x, *args) ** 2))
    assert abs(chk - rm.cost) <= 1e-9 * rm.cost
    eng = b200ba.bundle_adjustment._cache["engine"]
    assert len(eng.devices) == world
    rs = eng.residuals(pb.x0)
    assert np.abs(rs - res1).max() <= 1e-12 * np.abs(res1).max()
    nem = eng.normal_equations(pb.x0, 1e-3)
    assert np.abs(nem["S"] - ne1["S"]).max() <= 1e-11 * np.abs(ne1["S"]).max()
    assert np.abs(nem["b"] - ne1["b"]).max() <= 1e-11 * np.abs(ne1["b"]).max()
    assert np.abs(nem["grad"] - ne1["grad"]).max() <= 1e-11 * np.abs(ne1["grad"]).max()
    # shuffled observation order (any order is accepted, sfm.py always passes point-major)
    perm = np.random.default_rng(0).permutation(pb.n_obs)
    eng.set_problem(pb.n_cam, pb.n_pts, pb.camera_indices[perm], pb.point_indices[perm], pb.points_2d[perm], pb.K)
    assert np.abs(eng.residuals(pb.x0).reshape(-1, 2) - res1.reshape(-1, 2)[perm]).max() <= 1e-12 * np.abs(res1).max()
    xs, rsv = eng.solve(pb.x0, ftol=1e-10, solver="pcg")
    assert abs(rsv.cost - r1.cost) <= 1e-9 * r1.cost
    # fixed cameras and a robust loss go to every device
    eng.set_problem(*args)
    eng.set_fixed_cameras([0, 7])
    eng.set_loss("huber", 2.0)
    one.set_fixed_cameras([0, 7])
    one.set_loss("huber", 2.0)
    xa, ra = eng.solve(pb.x0, ftol=1e-10, solver="pcg")
    xb, rb = one.solve(pb.x0, ftol=1e-10, solver="pcg")
    assert abs(ra.cost - rb.cost) <= 1e-12 * rb.cost
    assert np.array_equal(xa[:6], pb.x0[:6]) and np.array_equal(xa[42:48], pb.x0[42:48])
    with pytest.raises(ValueError):
        eng.jacobian_blocks(pb.x0)            # not available on a multi-device handle: refused, not silently wrong
    one.close()
    b200ba.bundle_adjustment._cache["engine"].close()
    b200ba.bundle_adjustment._cache.update(engine=None, key=None, n_gpus=None)
