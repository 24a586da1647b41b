/* b200ba.h -- C ABI of libb200ba.so, the B200 (sm_100a) bundle-adjustment engine.
 *
 * The reference (hammershock/SFM-Python) has no FFI; its boundary for this path is two Python
 * imports in sfm_lite/sfm.py (":8" scipy.optimize.least_squares, ":20" compute_residuals /
 * create_sparsity_matrix) used only at sfm_lite/sfm.py:264-268.  The entry points below are what
 * a ctypes binding of that boundary needs; each one names the reference interface it replaces.
 * Plain pointers and sizes only, no C++ or torch types.  All arithmetic is IEEE double.
 *
 * Conventions
 *   x          parameter vector of sfm.py:255-257 / bundle_adjustment.py:38-39:
 *              [ w0(3) T0(3) ... w_{C-1} T_{C-1} | X0(3) ... X_{P-1}(3) ],  6*C + 3*P doubles.
 *   cam_idx    dense camera index per observation (sfm.py:256), 0 <= cam_idx < C.
 *   pt_idx     point index per observation (sfm.py:249); any order is accepted, point-major
 *              (non-decreasing, graph.py:186-191) is the fast path.
 *   uv         observed pixels, N x 2 row-major doubles (the Python wrapper converts the int64
 *              pixels of graph.py:112-113 exactly as numpy's promotion at bundle_adjustment.py:41).
 *   K          3x3 intrinsics, row-major, all nine entries used (bundle_adjustment.py:27).
 *   residuals  (2N) interleaved [du0,dv0,du1,dv1,...], projected - observed, in the CALLER's
 *              observation order (bundle_adjustment.py:41).
 *
 * Return value: 0 = OK, negative = b200ba_status; b200ba_last_error(h) gives the message.
 * No entry point keeps a caller pointer after it returns.  A handle is not re-entrant; different
 * handles may be used from different threads (the reference calls BA from a Tk worker thread,
 * app.py:84,109), every entry sets its CUDA device itself.
 */
#ifndef B200BA_H
#define B200BA_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200ba_handle b200ba_handle;

enum b200ba_status {
  B200BA_OK = 0,
  B200BA_BAD_ARG = -1,
  B200BA_CUDA = -2,
  B200BA_NCCL = -3,
  B200BA_NOT_FINITE = -4,
  B200BA_OOM = -5,
  B200BA_INDEX_RANGE = -6,
  B200BA_NO_PROBLEM = -7
};

enum b200ba_solver { B200BA_SOLVER_AUTO = 0, B200BA_SOLVER_PCG = 1, B200BA_SOLVER_CHOLESKY = 2 };

/* Options of b200ba_solve.  ftol/xtol/gtol/max_nfev have the meaning of the same-named keyword
 * arguments of scipy.optimize.least_squares as called at sfm.py:266-268 (ftol = --ba_tol,
 * main.py:36; xtol = gtol = 1e-8 scipy defaults; max_nfev <= 0 means 100 * n like scipy's TRF). */
typedef struct b200ba_options {
  double ftol;
  double xtol;
  double gtol;
  int64_t max_nfev;
  int32_t verbose;        /* 0 silent, 1 termination line, 2 per-iteration table (scipy's columns) */
  int32_t solver;         /* b200ba_solver */
  double lambda0;         /* initial Marquardt damping, <= 0 -> 1e-4 */
  double pcg_rtol;        /* relative residual of the reduced-camera PCG; <= 0 -> inexact-Newton forcing term chosen by
                             the LM loop (0.3 with the two-level preconditioner, 0.1 with block-Jacobi; tightened to 0.1 /
                             1e-2 once the relative cost reduction drops below 1e-4; 1e-8 for <= 100 cameras) */
  int32_t pcg_max_iter;   /* <= 0 -> 500 */
  int32_t fixed_iters;    /* > 0: run exactly this many LM iterations, ignoring the tolerances (benchmarks) */
} b200ba_options;

/* Result of b200ba_solve: the fields of scipy's OptimizeResult that parity tooling reads. */
typedef struct b200ba_result {
  double cost;            /* 0.5 * sum(r^2) at the returned x (whole job, all ranks) */
  double initial_cost;
  double optimality;      /* ||J^T r||_inf at the returned x */
  double rms;             /* sqrt(mean(r^2)) */
  double lambda_final;
  int64_t nfev;           /* residual evaluations (1 + trial steps) */
  int64_t njev;           /* linearisations */
  int32_t status;         /* scipy codes: 0 max_nfev, 1 gtol, 2 ftol, 3 xtol, 4 ftol and xtol */
  int32_t n_iter;         /* LM iterations (accepted + rejected) */
  int64_t pcg_iters;      /* total PCG iterations */
  double solve_ms;        /* device time of the LM loop (CUDA events) */
} b200ba_result;

/* Device-time counters, milliseconds / launches accumulated since b200ba_reset_stats. */
typedef struct b200ba_stats {
  double ms_cam_prep, ms_linearize_points, ms_camera_blocks, ms_schur, ms_allreduce, ms_camera_solve,
      ms_backsub_trial, ms_residual;
  int64_t launches;       /* kernels launched by this library */
  int64_t n_linearize, n_pcg_iters;
  int64_t schur_entries, schur_blocks;
  int64_t bytes_device;   /* device memory held by the handle */
  double ms_lin_fused, ms_slot_reduce;   /* fused linearisation + Schur kernel and its fixed-order partial reduction */
  int64_t n_tiles, n_pair_slots, n_self_slots;
  double ms_jacobian;     /* k_jacobian alone (config E sweep) */
  int64_t n_tl_builds;    /* rebuilds of the coarse operator of the two-level PCG preconditioner */
  double ms_reproj;       /* k_reproj_err alone */
  double ms_residual_kernel;  /* k_residual alone (ms_residual also covers k_cam_prep and the partial reduction) */
  double ms_chol_mc;      /* k_chol_mc alone: multi-CTA dense Cholesky on the FP64 tensor cores (factor + both substitutions) */
} b200ba_stats;

int b200ba_version(void);

/* One handle = one GPU (one rank).  device = CUDA ordinal. */
int b200ba_create(b200ba_handle** out, int device);
int b200ba_destroy(b200ba_handle* h);
const char* b200ba_last_error(const b200ba_handle* h);

/* Multi-GPU behind the single-process call of sfm.py:264-268: ONE handle that fronts n_devices GPUs of this
 * process (device_ids distinct CUDA ordinals, 1 <= n_devices <= 8).  b200ba_set_problem shards the points over the devices
 * (contiguous point ranges balanced by observation count, cameras replicated), b200ba_solve / b200ba_solve_io /
 * b200ba_residuals / b200ba_normal_equations (S only for n_cam <= 512) take and return the FULL arrays of the reference
 * call; b200ba_set_fixed_cameras / set_loss / set_two_level / get_stats / reset_stats / set_profiling apply to every
 * device.  Every other entry point returns B200BA_BAD_ARG on such a handle.  Each device is driven by its own host thread;
 * the reduced camera system is exchanged through peer memory over NVLink (direct peer pointers), the small sums through
 * NCCL.  The reference has no counterpart (its solve is single-threaded scipy). */
int b200ba_create_multi(b200ba_handle** out, const int* device_ids, int n_devices);

/* Multi-GPU (one process per GPU): rank 0 calls b200ba_nccl_unique_id, the host broadcasts the
 * 128 bytes (torch.distributed), every rank calls b200ba_comm_init.  The reference has no
 * collective; this is the all-reduce of the reduced camera system. */
int b200ba_nccl_unique_id(void* id128);
int b200ba_comm_init(b200ba_handle* h, const void* id128, int rank, int nranks);

/* Replaces the argument packing of sfm.py:248-261 + create_sparsity_matrix (bundle_adjustment.py:6-17):
 * copies the observation arrays to the device, builds the point-major SoA layout, the camera-sorted
 * permutation and the Schur pair lists.  n_pts is the number of points THIS handle owns (its shard);
 * pt_idx is local to the shard.  With a communicator attached this call is collective. */
int b200ba_set_problem(b200ba_handle* h, int64_t n_cam, int64_t n_pts, int64_t n_obs, const int64_t* cam_idx,
                       const int64_t* pt_idx, const double* uv, const double* K);

/* compute_residuals (bundle_adjustment.py:35-42).  x, r_out (2N, may be NULL), cost_out (may be NULL) on the host. */
int b200ba_residuals(b200ba_handle* h, const double* x, double* r_out, double* cost_out);
/* same with DEVICE pointers (torch tensor .data_ptr()); r_out_dev may be NULL */
int b200ba_residuals_dev(b200ba_handle* h, const double* x_dev, double* r_out_dev, double* cost_out);

/* Allocation-free variant of b200ba_residuals.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arithmetic as b200ba_residuals (k_cam_prep + the residual kernel), but it needs no b200ba_set_problem and makes no
 * cudaMalloc / cudaFree / pool allocation: its scratch (per-camera R, J_r, T table + per-CTA cost partials) is a
 * file-scope static __device__ array of 0.77 MB that every GPU loading the library holds until exit.
 * LIMIT: n_cam <= 4096 (B200BA_NOALLOC_MAX_CAM); larger problems are refused with B200BA_BAD_ARG and a message.
 * All array arguments are DEVICE pointers except K (9 host doubles): cam_idx_dev / pt_idx_dev int32 (n_obs), uv_dev
 * n_obs x 2, x_dev the parameter vector (6 n_cam + 3 n_pts), r_out_dev 2 n_obs in the CALLER's observation order (may be
 * NULL), cost_out_dev one double = 0.5 sum r^2 (may be NULL).  Out-of-range indices yield NaN residuals, not an error.
 * The scratch is shared by all calls on a device and keeps the last call's contents (nothing needs pre-clearing: every
 * slot read is rewritten by the same call); calls that overlap on one GPU (other streams / handles / threads) RACE.
 * Runs on the handle's stream and synchronises it before returning. */
#define B200BA_NOALLOC_MAX_CAM 4096
int b200ba_residuals_noalloc(b200ba_handle* h, int64_t n_cam, int64_t n_pts, int64_t n_obs, const int32_t* cam_idx_dev,
                             const int32_t* pt_idx_dev, const double* uv_dev, const double* K, const double* x_dev,
                             double* r_out_dev, double* cost_out_dev);

/* Pose <-> parameter conversion at the boundary.  pack = sfm.py:252-257: for each of n row-major
 * 4x4 poses H, params = [Rot.from_matrix(H[:3,:3]).as_rotvec() (3) | H[:3,3] (3)]; unpack = sfm.py:273-278:
 * H = [[Rot.from_rotvec(w).as_matrix(), T], [0 0 0 1]].  Restates scipy.spatial.transform's arithmetic (orthogonality
 * test + polar factor, Markley's quaternion, canonical sign, series below 1e-3); host buffers; no problem needed. */
int b200ba_pack_cameras(b200ba_handle* h, int64_t n, const double* H /* n x 16 */, double* params /* n x 6 */);
int b200ba_unpack_cameras(b200ba_handle* h, int64_t n, const double* params /* n x 6 */, double* H /* n x 16 */);

/* Perspective-n-point for the cameras of a new edge: cv2.solvePnP(pt3ds, pt2ds, K, zeros) at sfm.py:207-208, the
 * reference's restatement solve_pnp (cv2_lite/solve_pnp.py:17-66): DLT on normalised image points + non-linear
 * refinement of (rvec, tvec) on the reprojection error, R X + t convention (H_from_rtvec, transforms.py:18-23).
 * n_problems independent problems in one launch (one CTA each): counts[k] points of problem k, X (sum counts x 3) and
 * uv (sum counts x 2) concatenated; guess (nullable, n_problems x 6 [rvec | tvec]) skips the DLT (useExtrinsicGuess);
 * out: n_problems x 9 = rvec(3) tvec(3) cost iterations status (1 converged, 0 iteration limit, -1 not finite).
 * Non-planar point sets only (the DLT needs >= 6 points in general position). */
int b200ba_solve_pnp(b200ba_handle* h, int64_t n_problems, const int64_t* counts, const double* X, const double* uv,
                     const double* K, const double* guess, int32_t max_iter, double* out);

/* Two-view triangulation of SFM._apply_increment (sfm.py:214-221): cv2.triangulatePoints(M1, M2, pts1.T, pts2.T)
 * followed by X /= X[3] -- the DLT the reference restates as triangulate_points_linear2
 * (cv2_lite/triangulate_points.py:66-94) -- and, when mask_out is given, the visibility mask
 * (inv(H1) X)[2] > 0 and (inv(H2) X)[2] > 0 of sfm.py:219-221 (H1, H2: row-major 4x4 rigid poses).
 * M1, M2: row-major 3x4; pts1, pts2: n x 2; X_out: n x 4 rows [X Y Z 1]; host buffers; no problem needed. */
int b200ba_triangulate(b200ba_handle* h, int64_t n, const double* M1, const double* M2, const double* pts1,
                       const double* pts2, const double* H1, const double* H2, double* X_out, unsigned char* mask_out);

/* The reference's non-linear triangulation (cv2_lite/triangulate_points.py:122-184, not called by the pipeline itself): the
 * DLT point of b200ba_triangulate refined by minimising the reference's objective (:94-118 -- the squared reprojection
 * errors of the four image coordinates, squared again by its solver: a sum of 4th powers).  The reference hands all points
 * to one scipy call with default tolerances; here every point runs its own damped Gauss-Newton to the minimiser, so the
 * objective is never above the reference's and the points agree to the reference's own stopping accuracy (1e-3).
 * X_out: n x 4 rows [X Y Z 1]; max_iter <= 0 -> 100. */
int b200ba_triangulate_nonlinear(b200ba_handle* h, int64_t n, const double* M1, const double* M2, const double* pts1,
                                 const double* pts2, int32_t max_iter, double* X_out);

/* The per-observation reprojection-error loop of SFM._apply_increment (sfm.py:234-238), which calls
 * calc_reproj_error (sfm.py:38-41) -> reproj_error (cv2_lite/solve_pnp.py:9-14) once per observation in Python:
 * err_i = || pi(K (R_c X_p + t_c)) - uv_i ||  over the observations of the problem on the handle, in the pipeline's
 * R X + t convention (node.H, NOT the R (X - T) of the BA residual).  Rt: n_cam x 12 doubles, per camera R row-major
 * (9) then t (3) = the top three rows of node.H rearranged; X: n_pts x 3.  err_out (N, caller's order, may be NULL);
 * sum_out = sum_i err_i over all ranks (the `errs` of sfm.py:238; the mean is sum / N). */
int b200ba_reproj_errors(b200ba_handle* h, const double* Rt, const double* X, double* err_out, double* sum_out);
/* same with DEVICE pointers */
int b200ba_reproj_errors_dev(b200ba_handle* h, const double* Rt_dev, const double* X_dev, double* err_out_dev,
                             double* sum_out);

/* Analytic replacement of scipy's finite-difference Jacobian (sfm.py:266 jac='2-point' + jac_sparsity):
 * Jc N x 2 x 6 ([d/dw | d/dT]), Jp N x 2 x 3, caller's observation order, host buffers. */
int b200ba_jacobian_blocks(b200ba_handle* h, const double* x, double* Jc, double* Jp);
/* device-pointer variant used by the residual+Jacobian sweep (config E); also writes r_dev (2N) if not NULL */
int b200ba_jacobian_blocks_dev(b200ba_handle* h, const double* x_dev, double* r_dev, double* Jc_dev, double* Jp_dev);

/* One linearisation: residual, Jacobian blocks, J^T J / J^T r blocks, point-block Schur elimination.
 * Outputs (host, each may be NULL): S dense 6C x 6C row-major (only for n_cam <= 512), b (6C),
 * grad (6C + 3P) = J^T r, cost.  With a communicator the reduced system is all-reduced. */
int b200ba_normal_equations(b200ba_handle* h, const double* x, double lambda, double* S_out, double* b_out,
                            double* grad_out, double* cost_out);
/* device-resident linearisation only (benchmark leg): x_dev on the device, nothing copied back. */
int b200ba_linearize_dev(b200ba_handle* h, const double* x_dev, double lambda);
/* One damped step from x: delta (6C + 3P), predicted reduction, cost at x + delta (host buffers; test hook). */
int b200ba_lm_step(b200ba_handle* h, const double* x, double lambda, int solver, double* delta_out,
                   double* predicted_out, double* new_cost_out);

/* The `fixed_camera_indices` hook of create_sparsity_matrix (bundle_adjustment.py:6,13; unused by sfm.py:264, default
 * ()): the listed cameras (dense indices of the current problem) keep their parameters -- in the reference their
 * Jacobian columns are masked out of jac_sparsity, so scipy's finite differences never see them.  Here rows/columns
 * of the reduced camera system are deleted, so their step is exactly zero; the points they observe still move.
 * Call after b200ba_set_problem (which clears the set); n = 0 frees every camera.  On every rank of a multi-GPU job. */
int b200ba_set_fixed_cameras(b200ba_handle* h, const int64_t* cam_idx, int64_t n);

/* Robust loss of the solve: the `loss=` / `f_scale=` keywords of scipy.optimize.least_squares at the call of sfm.py:266
 * (the reference leaves scipy's default 'linear').  scipy's exact scaling is reproduced per residual component:
 * z = (f / f_scale)^2, cost = 0.5 f_scale^2 sum rho(z), J <- sqrt(max(rho' + 2 rho'' z, eps)) J, f <- rho' f / sqrt(...)
 * (scipy _lsq/least_squares.py construct_loss_function, _lsq/common.py scale_for_robust_loss_function).
 * loss: 0 linear, 1 soft_l1, 2 huber, 3 cauchy, 4 arctan.  Call after b200ba_set_problem (which resets it to linear);
 * b200ba_residuals / b200ba_jacobian_blocks stay the plain residual and blocks (scipy's `fun` and `jac`); b200ba_solve,
 * b200ba_normal_equations and b200ba_lm_step use the scaled ones.  On every rank of a multi-GPU job. */
int b200ba_set_loss(b200ba_handle* h, int loss, double f_scale);

/* Preconditioner of the reduced-camera PCG (the reference's scipy step has no counterpart; scipy's LSMR is
 * unpreconditioned): 0 = block-Jacobi only, 1 = automatic (default: block-Jacobi until a solve needs more than 24
 * iterations, then the two-level preconditioner with a lagged coarse operator), 2 = two-level from the first solve,
 * coarse operator rebuilt for every solve.  Only problems with >= 128 cameras carry the two-level structure. */
int b200ba_set_two_level(b200ba_handle* h, int mode);

/* Replaces scipy.optimize.least_squares at sfm.py:266-268: Levenberg-Marquardt on x_inout (host). */
int b200ba_solve(b200ba_handle* h, double* x_inout, const b200ba_options* opt, b200ba_result* out);
/* Page-locked host memory for result / input buffers (cudaHostAlloc): device <-> host copies of the parameter
 * vector (96 MB at config D) run at link speed instead of through the driver's staging copies.  The host mirror keeps
 * a free list of these and wraps them in numpy arrays that own them. */
void* b200ba_host_alloc(int64_t bytes);
void b200ba_host_free(void* p);

/* b200ba_solve with separate input and output buffers (x_in is only read: a pinned caller buffer is uploaded
 * without a staging copy, and the host mirror need not duplicate x0 before the call). */
int b200ba_solve_io(b200ba_handle* h, const double* x_in, double* x_out, const b200ba_options* opt, b200ba_result* out);
/* same with x resident on the device */
int b200ba_solve_dev(b200ba_handle* h, double* x_inout_dev, const b200ba_options* opt, b200ba_result* out);

int b200ba_get_stats(b200ba_handle* h, b200ba_stats* out);
int b200ba_reset_stats(b200ba_handle* h);
/* on != 0: bracket every kernel family with CUDA events and accumulate b200ba_stats.ms_* (adds syncs) */
int b200ba_set_profiling(b200ba_handle* h, int on);
/* Roofline denominator: FP64 FMA throughput of this device in TFLOP/s, measured by an unrolled DFMA-chain kernel
 * (no reference counterpart; MEASURED_PEAKS.json has no FP64 figure). Needs a problem (uses its scratch). */
int b200ba_fp64_peak(b200ba_handle* h, double* tflops_out);
/* stream the library launches on (so callers can record CUDA events around its kernels) */
void* b200ba_stream(b200ba_handle* h);

/* Test hook: copy an internal device array to the host.  name in {"V","gp","U","gc","pp","Sup","blk_keys",
 * "delta_c","rr","Y"}; returns the number of doubles (or int64 for keys) written, or a negative status. */
int64_t b200ba_debug_fetch(b200ba_handle* h, const char* name, void* out, int64_t capacity_bytes);

#ifdef __cplusplus
}
#endif
#endif /* B200BA_H */
