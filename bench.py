#!/usr/bin/env python
"""bench.py -- BA observations/s through one linearisation (residual + Jacobian blocks + J^T J / J^T r + Schur).

Contract (task statement): `python bench.py --gpus N --steps K --warmup W` (torchrun for N > 1, one rank per
GPU); rank 0 prints ONE JSON line.  A "step" is one full linearisation of the point-sharded problem with the
parameters resident in HBM, including the NCCL all-reduce of the reduced camera system at N > 1.
Workload = BASELINE.json configs[3] (10k cameras / 4M points / 20M observations) at every N (strong scaling, as
north_star states); `--workload B|C|D` selects another config.  Extra keys: `lm_iters_per_sec` (full LM
iterations of the device-resident solve), `e2e` (host buffers through the public API), `roofline`,
`cpu_baseline`, per-kernel milliseconds.  `--impl reference` times the reference's algorithm (oracle port of
sfm_lite/bundle_adjustment.py + scipy with the sfm.py:266-268 kwargs) on the host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "BA obs/sec (residual+Jacobian+Schur)"
UNIT = "obs/s"
WORKLOADS = {
    "B": "synthetic BAL-shaped 49 cameras / 7776 points / 31843 observations",
    "C": "synthetic 1k cameras / 500k points / 2M observations",
    "D": "synthetic 10k cameras / 4M points / 20M observations, point-sharded",
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="D", choices=list(WORKLOADS))
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--e2e-steps", type=int, default=4)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-sample-obs", type=int, default=8000)
    ap.add_argument("--sweep", action="store_true",
                    help="config E: residual + materialised Jacobian blocks, N = 1M..--sweep-max observations, vs the HBM roofline")
    ap.add_argument("--sweep-max", type=int, default=100_000_000)
    ap.add_argument("--no-sweep", action="store_true", help="skip the config-E points (1M..100M) inside the default run")
    return ap.parse_args()


# ---------------------------------------------------------------- clocks
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu):
        self.gpu, self.rows, self.proc = gpu, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i] == "Active"})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ---------------------------------------------------------------- CPU baseline (reference algorithm, host cores)
def reference_sample(workload, seed, n_obs_target):
    """A bounded sample of the workload: the observations of its first points, cameras renumbered densely."""
    import b200ba
    n_cam, n_pts, n_obs = b200ba.synthetic.CONFIGS[workload]
    k = max(n_obs // n_pts, 2)
    p_hi = max(min(n_pts, n_obs_target // k), 8)
    pb = b200ba.synthetic.make_problem(n_cam, n_pts, n_obs, seed=seed, point_range=(0, p_hi))
    cams, ci = np.unique(pb.camera_indices, return_inverse=True)
    x0 = np.concatenate([pb.x0[:6 * n_cam].reshape(n_cam, 6)[cams].ravel(), pb.x0[6 * n_cam:]])
    return dict(x0=x0, n_cam=len(cams), n_pts=p_hi, ci=ci.astype(np.int64), pi=pb.point_indices.copy(),
                uv=pb.points_2d, K=pb.K)


def load_reference():
    """The reference's OWN hot-path module staged by build() under oracle/_ref (sfm_lite/bundle_adjustment.py), or None."""
    import importlib.util
    path = os.path.join(ROOT, "oracle", "_ref", "ref_bundle_adjustment.py")
    if not os.path.exists(path):
        return None
    spec = importlib.util.spec_from_file_location("ref_bundle_adjustment", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def load_reference_gn():
    import importlib.util
    path = os.path.join(ROOT, "oracle", "_ref", "ref_least_squares.py")
    if not os.path.exists(path):
        return None
    spec = importlib.util.spec_from_file_location("ref_least_squares", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def reference_step(s, max_nfev=2):
    """One timed call of the reference path exactly as sfm.py:264-268 writes it: create_sparsity_matrix (its lil_matrix
    loop included) + scipy least_squares with the reference's own compute_residuals when oracle/_ref is staged
    (kind "reference"), else the oracle's per-observation port (kind "port").  Returns (seconds, njev, n_obs, kind)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ba_oracle as O
    ref = load_reference()
    args = (s["n_cam"], s["n_pts"], s["ci"], s["pi"], s["uv"], s["K"])
    t = time.perf_counter()
    if ref is not None:
        from scipy.optimize import least_squares
        mask = ref.create_sparsity_matrix(s["n_cam"], s["n_pts"], len(s["ci"]), s["ci"], s["pi"])
        res = least_squares(ref.compute_residuals, s["x0"], jac_sparsity=mask, verbose=0, x_scale="jac", ftol=1e-10,
                            method="trf", max_nfev=max_nfev, args=args)
    else:
        res = O.scipy_solve(s["x0"], *args, ftol=1e-10, max_nfev=max_nfev, fun=O.residuals_loop)
    return time.perf_counter() - t, int(res.njev), len(s["ci"]), "reference" if ref is not None else "port"


def r3_toy_gauss_newton():
    """R3 of BASELINE.md section 3: cv2_lite.least_squares (cv2_lite/least_squares.py:17-26, undamped Gauss-Newton with a
    dense forward-difference Jacobian), ONE iteration on SceauxCastle call 0, timing only (it diverges from real starts)."""
    gn, ref = load_reference_gn(), load_reference()
    fn = os.path.join(ROOT, "tests", "golden", "sceaux_ba_0.npz")
    if not os.path.exists(fn):
        return None
    d = np.load(fn)
    args = (int(d["n_cam"]), int(d["n_pts"]), d["ci"], d["pi"], d["uv"], d["K"])
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ba_oracle as O
    fun = ref.compute_residuals if ref is not None else O.residuals_loop
    t = time.perf_counter()
    if gn is not None:
        out = gn.least_squares(fun, d["x0"], args=args, max_iter=1)
        x1 = out["x"]
    else:      # restatement of least_squares.py:17-26 for one iteration
        x = np.array(d["x0"], dtype=float)
        f0 = fun(x, *args)
        J = np.zeros((f0.size, x.size))
        for i in range(x.size):
            x1 = x.copy()
            x1[i] += 1e-8
            J[:, i] = (fun(x1, *args) - f0) / 1e-8
        x1 = x + np.linalg.lstsq(J, -f0, rcond=None)[0]
    dt = time.perf_counter() - t
    r0, r1 = O.residuals(d["x0"], *args), O.residuals(x1, *args)
    return {"seconds_per_iteration": dt, "n_params": int(len(d["x0"])), "n_obs": int(len(d["ci"])),
            "obs_iters_per_sec": len(d["ci"]) / dt, "rms_px_before": O.rms(r0), "rms_px_after": O.rms(r1),
            "kind": "reference" if (gn is not None and ref is not None) else "port",
            "what": "cv2_lite.least_squares max_iter=1 on SceauxCastle BA call 0 with the reference residual"}


def cpu_baseline(workload, seed, n_obs_target):
    s = reference_sample(workload, seed, n_obs_target)
    dt, njev, n, kind = reference_step(s)
    # R2 of SURVEY 8d: the same scipy call with the vectorised numpy residual (the best the scipy path can do on a CPU)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ba_oracle as O
    t = time.perf_counter()
    r2 = O.scipy_solve(s["x0"], s["n_cam"], s["n_pts"], s["ci"], s["pi"], s["uv"], s["K"], ftol=1e-10, max_nfev=2)
    dt2 = time.perf_counter() - t
    out = {"value": n * njev / dt, "unit": UNIT, "cores": 1, "kind": kind,
           "vectorised_residual_value": n * int(r2.njev) / dt2,
           "sample": f"first {s['n_pts']} points / {n} observations / {s['n_cam']} cameras of the workload: "
                     f"create_sparsity_matrix + scipy TRF max_nfev=2 ({njev} linearisations) with "
                     f"{'the reference file staged under oracle/_ref' if kind == 'reference' else 'the per-observation residual port'}"
                     f", {dt:.1f} s", "seconds": dt}
    try:
        out["R3_toy_gauss_newton"] = r3_toy_gauss_newton()
    except Exception as e:      # a timing for the record must never take the bench line down
        out["R3_toy_gauss_newton"] = {"error": repr(e)}
    g = os.path.join(ROOT, "tests", "golden", "configB_scipy.npz")
    if os.path.exists(g):
        d = np.load(g)
        out["R2_config_B_full_solve"] = {
            "seconds": float(d["scipy_seconds"]), "time_to_cost_seconds": float(d["scipy_time_to_cost"]),
            "nfev": int(d["scipy_nfev"]), "cost": float(d["scipy_cost"]), "rms_px": float(d["scipy_rms"]),
            "note": "full scipy TRF solve (sfm.py:266-268 kwargs, vectorised residual) of config B recorded by "
                    "oracle/make_golden.py configB on the build container's CPU; time_to_cost = first evaluation within 1e-6 "
                    "of its final cost"}
    return out


def run_reference(a, rank):
    if rank != 0:
        return
    s = reference_sample(a.workload, a.seed, a.cpu_sample_obs)
    kind = "port"
    for _ in range(a.warmup):
        reference_step(s)
    tot, lin, n = 0.0, 0, 0
    for _ in range(a.steps):
        dt, njev, n, kind = reference_step(s)
        tot += dt
        lin += njev
    v = n * lin / tot
    sample = (f"each step = create_sparsity_matrix + scipy TRF max_nfev=2 on the first {s['n_pts']} points / {n} observations / "
              f"{s['n_cam']} cameras of the workload ("
              + ("the reference's own sfm_lite/bundle_adjustment.py staged under oracle/_ref" if kind == "reference"
                 else "per-observation residual port") + ", single Python thread as the reference runs it)")
    print(json.dumps({"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus,
                      "steps": a.steps, "warmup": a.warmup, "ms_per_step": 1e3 * tot / a.steps,
                      "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
                      "data": "synthetic", "config": {"workload": WORKLOADS[a.workload], "sample": sample},
                      "cpu_baseline": {"value": v, "unit": UNIT, "cores": 1, "kind": kind, "sample": sample},
                      "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


# ---------------------------------------------------------------- config E: residual + Jacobian sweep vs the HBM roofline
NOMINAL_HBM_GBS = 8000.0    # north_star's "about 8 TB/s"; the measured copy peak is the other denominator


def sweep_problem(n_obs, seed):
    """Config-E problem with n_obs observations (1000 cameras, 4 observations per point).  Up to 20 M it is generated; above,
    the 10 M problem is replicated with shifted point ids (distinct memory, same geometry) to keep the run short."""
    import b200ba
    if n_obs <= 20_000_000:
        pb = b200ba.synthetic.make_problem(1000, n_obs // 4, n_obs, seed=seed)
        return 1000, n_obs // 4, pb.camera_indices, pb.point_indices, pb.points_2d, pb.K, pb.x0
    base = b200ba.synthetic.make_problem(1000, 2_500_000, 10_000_000, seed=seed)
    k = n_obs // 10_000_000
    ci = np.tile(base.camera_indices, k)
    pi = np.concatenate([base.point_indices + j * 2_500_000 for j in range(k)])
    uv = np.tile(base.points_2d, (k, 1))
    x0 = np.concatenate([base.x0[:6000]] + [base.x0[6000:]] * k)
    return 1000, 2_500_000 * k, ci, pi, uv, base.K, x0


def sweep_rows(eng, seed, sweep_max, peak):
    """BASELINE.json configs[4]: k_residual (40 B/obs) and k_jacobian (184 B/obs materialised: r 16 + Jc 96 + Jp 48 out,
    u, v, cam, pt 24 in) at N = 1M .. 100M observations; kernel time from CUDA events placed directly around each kernel
    on the library stream (profiling mode), best-effort L2-cold: every buffer of a pass exceeds L2 from 2 M observations."""
    import torch
    rows = []
    saved = {k: os.environ.get(k) for k in ("B200BA_SCHUR_STREAM", "B200BA_FUSED")}
    os.environ["B200BA_SCHUR_STREAM"] = "1"      # the sweep needs no Schur structures: cheapest setup path
    os.environ["B200BA_FUSED"] = "0"
    try:
        for n_obs in (1_000_000, 2_000_000, 5_000_000, 10_000_000, 20_000_000, 50_000_000, 100_000_000):
            if n_obs > sweep_max:
                break
            n_cam, n_pts, ci, pi, uv, K, x0 = sweep_problem(n_obs, seed)
            eng.set_problem(n_cam, n_pts, ci, pi, uv, K)
            del ci, pi, uv
            x = torch.from_numpy(x0).cuda()
            r = torch.empty(2 * n_obs, dtype=torch.float64, device="cuda")
            jc = torch.empty(12 * n_obs, dtype=torch.float64, device="cuda")
            jp = torch.empty(6 * n_obs, dtype=torch.float64, device="cuda")
            eng.set_profiling(True)
            for _ in range(3):
                eng.jacobian_blocks_dev(x.data_ptr(), r.data_ptr(), jc.data_ptr(), jp.data_ptr())
                eng.residuals_dev(x.data_ptr(), r.data_ptr())
            eng.reset_stats()
            reps = 10
            for _ in range(reps):
                eng.jacobian_blocks_dev(x.data_ptr(), r.data_ptr(), jc.data_ptr(), jp.data_ptr())
                eng.residuals_dev(x.data_ptr(), r.data_ptr())
            st = eng.stats()
            eng.set_profiling(False)
            t_j = st["ms_jacobian"] / reps * 1e-3
            t_r = st["ms_residual_kernel"] / reps * 1e-3
            bj = 184 * n_obs + 48 * n_cam + 24 * n_pts
            br = 40 * n_obs + 48 * n_cam + 24 * n_pts
            rows.append({"n_obs": n_obs, "jacobian_ms": t_j * 1e3, "jacobian_GBps": bj / t_j / 1e9,
                         "jacobian_frac": bj / t_j / 1e9 / peak, "jacobian_frac_of_nominal_8TBs": bj / t_j / 1e9 / NOMINAL_HBM_GBS,
                         "jacobian_obs_per_s": n_obs / t_j, "residual_ms": t_r * 1e3, "residual_GBps": br / t_r / 1e9,
                         "residual_frac": br / t_r / 1e9 / peak, "residual_frac_of_nominal_8TBs": br / t_r / 1e9 / NOMINAL_HBM_GBS,
                         "residual_with_prep_and_reduce_ms": st["ms_residual"] / reps})
            del x, r, jc, jp
            torch.cuda.empty_cache()
    finally:
        for k, v in saved.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    return rows


def run_sweep(a):
    import b200ba
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    eng = b200ba.Engine(0)
    rows = sweep_rows(eng, a.seed, a.sweep_max, peak)
    print(json.dumps({"metric": "residual+Jacobian sweep (config E)", "unit": "GB/s", "peak_hbm_gbs": peak,
                      "peak_nominal_gbs": NOMINAL_HBM_GBS,
                      "peak_source": "MEASURED_PEAKS.json hbm_gbs (measured)" if peaks else "fallback 6650 GB/s",
                      "bytes_per_obs": {"jacobian": 184, "residual": 40}, "dtype": "f64", "data": "synthetic", "rows": rows}))


# ---------------------------------------------------------------- GPU arm
def run_b200(a, rank, world, local_rank):
    import torch

    import b200ba
    from b200ba import dist as bd
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local_rank)
    td = None
    if world > 1:
        import torch.distributed as tdist
        tdist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        td = tdist

    n_cam, n_pts, n_obs = b200ba.synthetic.CONFIGS[a.workload]
    deg = b200ba.synthetic.point_degrees(n_pts, n_obs, min(16, n_cam), a.seed)
    lo, hi = bd.plan_shards_from_degrees(deg, world)[rank]
    pb = b200ba.synthetic.make_problem(n_cam, n_pts, n_obs, seed=a.seed, point_range=(lo, hi))
    ci, pil, uv = pb.camera_indices, pb.point_indices - lo, pb.points_2d

    def pin(arr):
        return torch.from_numpy(np.ascontiguousarray(arr)).pin_memory().numpy()
    ci_h, pi_h, uv_h, x0_h = pin(ci), pin(pil), pin(uv), pin(pb.x0)

    if world > 1:
        dba = bd.DistributedBA(local_rank)
        eng = dba.engine
    else:
        eng = b200ba.Engine(local_rank)
    eng.set_problem(n_cam, hi - lo, ci_h, pi_h, uv_h, pb.K)
    lam = 1e-4
    x_dev = torch.from_numpy(pb.x0).to(f"cuda:{local_rank}")
    stream = torch.cuda.ExternalStream(eng.stream(), device=local_rank)

    def barrier():
        torch.cuda.synchronize()
        if td:
            td.barrier()
        torch.cuda.synchronize()

    def timed(fn, reps):
        """reps calls of fn between device events on the library's stream; max over ranks, in ms."""
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(reps):
            fn()
        e1.record(stream)
        e1.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device=f"cuda:{local_rank}", dtype=torch.float64)
        barrier()
        if td:
            td.all_reduce(ms, op=td.ReduceOp.MAX)
        return float(ms.item())

    # ---- value: K linearisations, parameters resident in HBM
    lin = lambda: eng.linearize_dev(x_dev.data_ptr(), lam)
    for _ in range(max(a.warmup, 3)):
        lin()
    eng.reset_stats()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ms_total = timed(lin, a.steps)
    st = eng.stats()
    launches = int(st["launches"])
    # keep the kernels running for >= 1.5 s more so that nvidia-smi (100 ms period) sees the clocks UNDER THIS LOAD;
    # these extra steps are not part of the timed K
    for _ in range(min(3000, int(1500.0 / max(ms_total / a.steps, 1e-3)) + 1)):   # same count on every rank (collective)
        lin()
    torch.cuda.synchronize()
    clocks = sampler.stop() if rank == 0 else None
    ms_step = ms_total / a.steps
    value = n_obs / (ms_step * 1e-3)

    # ---- per-kernel device times (profiling mode: events around each kernel family, separate pass)
    eng.set_profiling(True)
    eng.reset_stats()
    reps = max(a.steps, 5)
    for _ in range(reps):
        lin()
    sp = eng.stats()
    eng.set_profiling(False)
    fused = int(st["n_self_slots"]) > 0
    kkeys = ("ms_lin_fused", "ms_slot_reduce", "ms_allreduce") if fused else \
            ("ms_linearize_points", "ms_camera_blocks", "ms_schur", "ms_allreduce")
    kern_ms = {k: sp[k] / reps for k in kkeys}
    fp64_peak = eng.fp64_peak_tflops()

    # ---- LM iterations/s: device-resident solve, exactly K iterations (tolerances off)
    solver = "cholesky" if n_cam <= 64 else "pcg"
    x_work = x_dev.clone()
    eng.solve_dev(x_work.data_ptr(), fixed_iters=max(a.warmup, 3), solver=solver)
    x_work.copy_(x_dev)
    barrier()
    res = eng.solve_dev(x_work.data_ptr(), fixed_iters=a.steps, solver=solver)
    lm_ms = torch.tensor([res.solve_ms], device=f"cuda:{local_rank}", dtype=torch.float64)
    if td:
        td.all_reduce(lm_ms, op=td.ReduceOp.MAX)
    lm_ms = float(lm_ms.item())
    # phase breakdown of one LM iteration (profiling mode: events around each phase, separate pass)
    x_work.copy_(x_dev)
    eng.set_profiling(True)
    eng.reset_stats()
    eng.solve_dev(x_work.data_ptr(), fixed_iters=3, solver=solver)
    sq = eng.stats()
    eng.set_profiling(False)
    lm_phase = {k: sq[k] / 3 for k in ("ms_lin_fused", "ms_slot_reduce", "ms_allreduce", "ms_camera_solve", "ms_backsub_trial")}
    lm = {"lm_iters_per_sec": a.steps / (lm_ms * 1e-3), "ms_per_lm_iter": lm_ms / a.steps, "lm_phase_ms": lm_phase,
          "lm_phase_note": "first 3 LM iterations only (block-Jacobi PCG converges in a few iterations there); the "
                           "camera solve of later iterations is ms_per_lm_iter minus the other phases",
          "pcg_iters_per_lm_iter": res.pcg_iters / max(res.n_iter, 1), "solver": solver,
          "preconditioner": "two-level (gauge modes per camera aggregate, lagged coarse inverse) once block-Jacobi "
                            "needs > 24 iterations" if n_cam >= 128 else "block-Jacobi",
          "cost_initial": res.initial_cost, "cost_after": res.cost}
    # time to convergence with the reference's tolerances (ftol = 1e-10 = --ba_tol default, xtol = gtol = 1e-8)
    x_work.copy_(x_dev)
    barrier()
    rc = eng.solve_dev(x_work.data_ptr(), ftol=1e-10, solver=solver, max_nfev=200)
    cv = torch.tensor([rc.solve_ms], device=f"cuda:{local_rank}", dtype=torch.float64)
    if td:
        td.all_reduce(cv, op=td.ReduceOp.MAX)
    lm["converge"] = {"ms": float(cv.item()), "lm_iters": int(rc.n_iter), "pcg_iters": int(rc.pcg_iters),
                      "status": int(rc.status), "cost": rc.cost, "rms_px": rc.rms}

    # ---- parity inside the bench (outside every timed region): the cost the LM loop reported at x0 against the ORACLE's
    #      residual on the same arrays (each rank its shard, summed), so a wrong kernel cannot produce a bench line
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ba_oracle as O
    r_or = O.residuals(pb.x0, n_cam, hi - lo, ci, pil, uv, pb.K)
    c_or = torch.tensor([0.5 * float(r_or @ r_or)], device=f"cuda:{local_rank}", dtype=torch.float64)
    del r_or
    if td:
        td.all_reduce(c_or)
    cost_oracle = float(c_or.item())
    cost_err = abs(res.initial_cost - cost_oracle) / cost_oracle
    if cost_err > 1e-9:
        raise RuntimeError(f"parity: cost at x0 {res.initial_cost!r} differs from the oracle's {cost_oracle!r} ({cost_err:.2e})")
    lm["cost_initial_oracle"] = cost_oracle
    lm["cost_initial_rel_err"] = cost_err

    # ---- multi-GPU parity (N > 1): the sharded reduced camera system / b / cost of a 300-camera problem against an
    #      unsharded engine on rank 0 and against the oracle's cost, max relative error -> "shard_parity"
    shard_parity = None
    if world > 1:
        sp = b200ba.synthetic.make_problem(300, 30000, 120000, seed=3)
        dba.set_problem(*sp.args())
        slo, shi = dba.range
        ne = dba.engine.normal_equations(bd.split_x(sp.x0, 300, slo, shi), 1e-3)
        if rank == 0:
            one = b200ba.Engine(local_rank)
            one.set_problem(*sp.args())
            ref = one.normal_equations(sp.x0, 1e-3)
            one.close()
            r_sp = O.residuals(sp.x0, *sp.args())
            shard_parity = max(np.abs(ne["S"] - ref["S"]).max() / np.abs(ref["S"]).max(),
                               np.abs(ne["b"] - ref["b"]).max() / np.abs(ref["b"]).max(),
                               abs(ne["cost"] - ref["cost"]) / ref["cost"],
                               abs(ne["cost"] - 0.5 * float(r_sp @ r_sp)) / ne["cost"])
        barrier()
        eng.set_problem(n_cam, hi - lo, ci_h, pi_h, uv_h, pb.K)

    # ---- e2e: the public host API with host buffers; every step uploads the problem and x0, solves, downloads x
    e2e_iters = 5
    def e2e_step():
        t0 = time.perf_counter()
        eng.set_problem(n_cam, hi - lo, ci_h, pi_h, uv_h, pb.K)
        t1 = time.perf_counter()
        x, r = eng.solve(x0_h, ftol=1e-10, fixed_iters=e2e_iters, solver=solver)
        if os.environ.get("B200BA_SETUP_PROF"):
            print(f"[e2e] set_problem {1e3 * (t1 - t0):.1f} ms, solve {1e3 * (time.perf_counter() - t1):.1f} ms "
                  f"(device {r.solve_ms:.1f} ms)", file=sys.stderr)
        return r
    e2e_step()
    barrier()
    t0 = time.perf_counter()
    njev, step_s, step_parts = 0, [], []
    for _ in range(a.e2e_steps):
        ts = time.perf_counter()
        eng.set_problem(n_cam, hi - lo, ci_h, pi_h, uv_h, pb.K)
        tm = time.perf_counter()
        x_e2e, r_e2e = eng.solve(x0_h, ftol=1e-10, fixed_iters=e2e_iters, solver=solver)
        del x_e2e
        te = time.perf_counter()
        njev += int(r_e2e.njev)
        step_s.append(te - ts)
        step_parts.append((round(1e3 * (tm - ts), 2), round(1e3 * (te - tm), 2), round(r_e2e.solve_ms, 2)))
    torch.cuda.synchronize()
    dt = torch.tensor([time.perf_counter() - t0] + step_s, device=f"cuda:{local_rank}", dtype=torch.float64)
    if td:
        td.all_reduce(dt, op=td.ReduceOp.MAX)
    step_s = [float(v) for v in dt[1:].tolist()]
    dt = float(dt[0].item())
    h2d = ci_h.nbytes + pi_h.nbytes + uv_h.nbytes + x0_h.nbytes
    e2e = {"value": n_obs * njev / dt, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(x0_h.nbytes),
           "what": f"Engine.set_problem + Engine.solve (host numpy buffers, pinned) with {e2e_iters} LM iterations per step; "
                   f"observations x linearisations / wall time of all {a.e2e_steps} steps (max over ranks)",
           "seconds_per_step": dt / a.e2e_steps, "step_seconds": [round(v, 5) for v in step_s],
           "step_ms_set_problem_solve_device_rank0": step_parts,
           "value_at_median_step": n_obs * (njev / a.e2e_steps) / float(np.median(step_s))}

    # ---- the HBM-bound kernels of the path at this workload's size (SURVEY 8d families E1 / E2): residual only
    #      (40 B/observation) and residual + materialised 2x6 / 2x3 blocks (184 B/observation), CUDA events around each
    hbm_kernels = None
    if world == 1:
        n_loc0 = len(ci)
        r_t = torch.empty(2 * n_loc0, dtype=torch.float64, device=f"cuda:{local_rank}")
        jc_t = torch.empty(12 * n_loc0, dtype=torch.float64, device=f"cuda:{local_rank}")
        jp_t = torch.empty(6 * n_loc0, dtype=torch.float64, device=f"cuda:{local_rank}")
        for _ in range(3):
            eng.jacobian_blocks_dev(x_dev.data_ptr(), r_t.data_ptr(), jc_t.data_ptr(), jp_t.data_ptr())
            eng.residuals_dev(x_dev.data_ptr(), r_t.data_ptr())
        eng.set_profiling(True)
        eng.reset_stats()
        for _ in range(10):
            eng.jacobian_blocks_dev(x_dev.data_ptr(), r_t.data_ptr(), jc_t.data_ptr(), jp_t.data_ptr())
            eng.residuals_dev(x_dev.data_ptr(), r_t.data_ptr())
        sk = eng.stats()
        eng.set_profiling(False)
        del r_t, jc_t, jp_t
        torch.cuda.empty_cache()
        pk = 6545.3
        try:
            pk = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", pk))
        except Exception:
            pass
        bj = 184 * n_loc0 + 48 * n_cam + 24 * (hi - lo)
        br = 40 * n_loc0 + 48 * n_cam + 24 * (hi - lo)
        tj, tr = sk["ms_jacobian"] / 10 * 1e-3, sk["ms_residual_kernel"] / 10 * 1e-3
        hbm_kernels = {"k_jacobian": {"bound": "hbm", "ms": tj * 1e3, "achieved": bj / tj / 1e9, "peak": pk, "unit": "GB/s",
                                      "frac": bj / tj / 1e9 / pk, "frac_of_nominal_8TBs": bj / tj / 1e9 / NOMINAL_HBM_GBS,
                                      "bytes_per_obs": 184},
                       "k_residual": {"bound": "hbm", "ms": tr * 1e3, "achieved": br / tr / 1e9, "peak": pk, "unit": "GB/s",
                                      "frac": br / tr / 1e9 / pk, "frac_of_nominal_8TBs": br / tr / 1e9 / NOMINAL_HBM_GBS,
                                      "bytes_per_obs": 40, "with_cam_prep_and_reduction_ms": sk["ms_residual"] / 10}}
    # ---- e2e as the drop-in caller sees it: PAGEABLE numpy arrays (what sfm.py hands over), upload + solve to ftol = 1e-10
    #      (--ba_tol default) + download of x, wall clock; and config B the same way (time-to-cost beside scipy's R2 record)
    def e2e_converge(args5, x0p, slv):
        t0 = time.perf_counter()
        eng.set_problem(*args5)
        x, r = eng.solve(x0p, ftol=1e-10, solver=slv, max_nfev=200)
        return time.perf_counter() - t0, r
    pg = (n_cam, hi - lo, np.array(ci), np.array(pil), np.array(uv), pb.K)
    x0_pg = np.array(pb.x0)
    e2e_converge(pg, x0_pg, solver)
    barrier()
    dt_pg, r_pg = e2e_converge(pg, x0_pg, solver)
    dtt = torch.tensor([dt_pg], device=f"cuda:{local_rank}", dtype=torch.float64)
    if td:
        td.all_reduce(dtt, op=td.ReduceOp.MAX)
    e2e["solve_to_ftol_pageable"] = {"seconds": float(dtt.item()), "lm_iters": int(r_pg.n_iter), "status": int(r_pg.status),
                                      "cost": r_pg.cost, "value": n_obs * int(r_pg.njev) / float(dtt.item()), "unit": UNIT,
                                      "what": "set_problem + solve(ftol=1e-10) + download with pageable numpy inputs, wall clock"}
    if world == 1:
        # the call of sfm.py:264-268 itself, swapped in: b200ba.least_squares on fresh pageable arrays (cache miss: key,
        # upload, solve to ftol, download), wall clock
        def dropin():
            fresh = (n_cam, n_pts, np.array(ci), np.array(pil), np.array(uv), pb.K)
            mask = b200ba.create_sparsity_matrix(n_cam, n_pts, n_obs, fresh[2], fresh[3])
            t0 = time.perf_counter()
            res = b200ba.least_squares(b200ba.compute_residuals, x0_pg, jac_sparsity=mask, verbose=0, x_scale="jac",
                                       ftol=1e-10, method="trf", args=fresh)
            return time.perf_counter() - t0, res
        dropin()
        dt_di, res_di = dropin()
        e2e["dropin_least_squares"] = {"seconds": dt_di, "cost": float(res_di.cost), "nfev": int(res_di.nfev),
                                       "value": n_obs * int(res_di.njev) / dt_di, "unit": UNIT,
                                       "what": "b200ba.least_squares(compute_residuals, x0, jac_sparsity=..., x_scale='jac', "
                                               "ftol=1e-10, method='trf', args=...) with fresh pageable numpy arrays, wall clock"}
        del res_di
        b200ba.reset()
    del pg, x0_pg
    config_b = None
    if world == 1:
        pbB = b200ba.synthetic.make_config("B")
        argsB = pbB.args()
        e2e_converge(argsB, pbB.x0, "auto")
        dtB, rB = e2e_converge(argsB, pbB.x0, "auto")
        config_b = {"what": "config B (49 / 7776 / 31843) through Engine.set_problem + solve(ftol=1e-10), host arrays, wall clock",
                    "seconds": dtB, "lm_iters": int(rB.n_iter), "cost": rB.cost, "rms_px": rB.rms, "status": int(rB.status)}
        eng.set_problem(n_cam, hi - lo, ci_h, pi_h, uv_h, pb.K)

    # ---- config E inside the default run: k_residual / k_jacobian at 1 M .. 100 M observations
    sweep = None
    if world == 1 and not a.no_sweep:
        pk0 = 6650.0
        try:
            pk0 = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", pk0))
        except Exception:
            pass
        del x_dev, x_work
        torch.cuda.empty_cache()
        sweep = sweep_rows(eng, a.seed, a.sweep_max, pk0)

    # ---- BASELINE configs[0]: the nine global BA calls of the SceauxCastle run (frozen problems, tests/golden) through the
    #      drop-in call of sfm.py:264-268, host arrays in, result.x out; beside it the reference's own times for the same
    #      calls as recorded when the goldens were made (real reference residual + scipy on calls 0-2, scipy with the
    #      vectorised residual on all nine; another machine's CPU, quoted for orientation only)
    config_a = None
    gold = os.path.join(ROOT, "tests", "golden")
    if world == 1 and os.path.exists(os.path.join(gold, "sceaux_ba_0.npz")):
        calls, per_call, ref_vec, ref_real = [], [], [], []
        rr = np.load(os.path.join(gold, "sceaux_scipy_ref.npz"))
        for k in range(9):
            fn = os.path.join(gold, f"sceaux_ba_{k}.npz")
            if not os.path.exists(fn):
                break
            d = np.load(fn)
            calls.append((d["x0"], (int(d["n_cam"]), int(d["n_pts"]), d["ci"], d["pi"], d["uv"], d["K"]), float(d["ftol"])))
            ref_vec.append(float(d["scipy_t"]))
            if f"t_{k}" in rr.files:
                ref_real.append(float(rr[f"t_{k}"]))
        def replay():
            out = []
            for x0, args, ftol in calls:
                t0 = time.perf_counter()
                mask = b200ba.create_sparsity_matrix(args[0], args[1], len(args[2]), args[2], args[3])
                res = b200ba.least_squares(b200ba.compute_residuals, x0, jac_sparsity=mask, verbose=0, x_scale="jac",
                                           ftol=ftol, method="trf", args=args)
                out.append((time.perf_counter() - t0, float(res.cost)))
            return out
        replay()
        per_call = replay()
        config_a = {"what": "SceauxCastle incremental SfM, the 9 frozen global BA calls (3-7 cameras, 563-3034 observations) "
                            "through b200ba.least_squares with host arrays (set_problem + LM solve + download per call)",
                    "calls": len(per_call), "total_ms": 1e3 * sum(t for t, _ in per_call),
                    "per_call_ms": [round(1e3 * t, 3) for t, _ in per_call],
                    "reference_seconds_real_residual_calls_0_2": ref_real,
                    "reference_seconds_vectorised_residual": ref_vec,
                    "reference_note": "scipy times recorded on the CPU of the machine that made the goldens"}
    if rank != 0:
        return
    # ---- roofline of the dominant kernel of the step (SURVEY.md 8d family L)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    n_loc, p_loc = len(ci), hi - lo
    n_ent, n_blk = int(st["schur_entries"]), int(st["schur_blocks"])
    kdeg = deg[lo:hi].astype(np.float64)
    # compulsory bytes of the linearisation contract: observations in, points in, cameras in, per-point factor + gradient
    # out (72 B), reduced system out (block-sparse upper triangle, b, g_c)
    lin_bytes = 24 * n_loc + 24 * p_loc + 48 * n_cam + 72 * p_loc + 288 * n_blk + 96 * n_cam
    # algorithmic flops: 350 per observation + per point [3x3 factor 40 + 120 k + 216 k (k + 1) / 2]
    lin_flops = 350.0 * n_loc + float(np.sum(40.0 + 120.0 * kdeg + 108.0 * kdeg * (kdeg + 1.0)))
    if fused:
        dom = "ms_lin_fused"
        dom_name = "k_lin_flow" if os.environ.get("B200BA_FLOW", "1") != "0" else "k_lin_fused"
        alg = {"ms_lin_fused": lin_bytes}
    else:
        alg = {"ms_linearize_points": 24 * n_loc + 24 * p_loc + 192 * n_cam + (144 + 32) * n_loc + (96 + 24) * p_loc,
               "ms_camera_blocks": (8 + 32 + 24) * n_loc + 192 * n_cam + 33 * 8 * (n_loc // 256 + n_cam),
               "ms_schur": 144 * n_loc + 8 * n_ent + 288 * (n_ent // 512 + n_blk)}
        dom = max(alg, key=lambda k: kern_ms[k])
        dom_name = {"ms_linearize_points": "k_lin_points", "ms_camera_blocks": "k_cam_side", "ms_schur": "k_schur_tiles"}[dom]
    t_dom = max(kern_ms[dom], 1e-9) * 1e-3
    ach = alg[dom] / t_dom / 1e9
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(f"{dom_name}:{a.workload}")
    except Exception:
        pass
    roof = {"bound": "hbm", "kernel": dom_name, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
            "frac_of_nominal_8TBs": ach / NOMINAL_HBM_GBS,
            "traffic": traffic,
            "peak_source": "MEASURED_PEAKS.json hbm_gbs (measured)" if peaks else "fallback 6650 GB/s",
            "algorithmic_bytes_per_launch": alg[dom], "kernel_ms": kern_ms[dom],
            "note": "compulsory bytes of the linearisation contract; the kernel is FP64-pipe-bound, see roofline_fp64"}
    tf = lin_flops / (max(kern_ms.get("ms_lin_fused", ms_step), 1e-9) * 1e-3) / 1e12
    roof64 = {"bound": "fp64", "kernel": dom_name if fused else "linearisation (all kernels)", "achieved": tf,
              "peak": fp64_peak, "unit": "TFLOP/s", "frac": tf / fp64_peak if fp64_peak > 0 else None,
              "peak_source": "b200ba_fp64_peak: unrolled DFMA chains over all SMs, measured in this run",
              "algorithmic_flops_per_launch": lin_flops,
              "flops_per_compulsory_byte": lin_flops / lin_bytes}
    out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
           "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
           "data": "synthetic", "impl": "b200",
           "config": {"workload": WORKLOADS[a.workload], "n_cam": n_cam, "n_pts": n_pts, "n_obs": n_obs,
                      "parallelism": f"points sharded over {world} GPU(s), cameras replicated, 1 all-reduce per linearisation",
                      "cache": "inputs and outputs of a step exceed the 126 MB L2" if n_obs >= 2_000_000 else
                               "working set fits L2 (latency-bound config)", "lambda": lam,
                      "schur_blocks": n_blk, "schur_pair_entries_rank0": n_ent},
           "e2e": e2e, "gpu_launches": launches, "clocks": clocks, "roofline": roof, "roofline_fp64": roof64, "kernel_ms": kern_ms,
           "tiles": {"n_tiles": int(st["n_tiles"]), "pair_slots": int(st["n_pair_slots"]),
                     "self_slots": int(st["n_self_slots"])}}
    if shard_parity is not None:
        out["shard_parity"] = shard_parity
    if hbm_kernels:
        if sweep:
            hbm_kernels["sweep_config_E"] = sweep
        out["roofline_hbm_kernels"] = hbm_kernels
    if config_b:
        out["config_B"] = config_b
    if config_a:
        out["config_A_replay"] = config_a
    out.update(lm)
    if not a.no_cpu_baseline and world == 1:
        out["cpu_baseline"] = cpu_baseline(a.workload, a.seed, a.cpu_sample_obs)
    print(json.dumps(out))


def main():
    a = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if a.impl == "reference":
        run_reference(a, rank)
        return
    if a.sweep:
        if rank == 0:
            run_sweep(a)
        return
    run_b200(a, rank, world, local_rank)
    if world > 1:
        import torch.distributed as td
        td.destroy_process_group()


if __name__ == "__main__":
    main()
