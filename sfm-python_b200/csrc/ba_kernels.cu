// Per-iteration CUDA kernels of the bundle-adjustment hot path (FP64, sm_100a).
//
//   k_cam_prep        w,T -> R, J_r per camera            (replaces Rotation.from_rotvec per OBSERVATION,
//                                                          sfm_lite/bundle_adjustment.py:25)
//   k_residual        reprojection residual + cost        (bundle_adjustment.py:20-42)
//   k_jacobian        materialised 2x6 / 2x3 blocks       (replaces scipy 2-point FD, sfm.py:266)
//   k_lin_points      point-major fused linearisation: V_p, g_p (warp-segmented reduction), damped 3x3
//                     factor, per-observation Y = W L^-T and reduced residual
//   k_cam_side        camera-sorted accumulation of U_c, g_c, b_c without atomics
//   k_schur           S -= sum Y_i Y_j^T over pair entries sorted by block, without atomics
//   k_backsub_trial   delta_p back-substitution fused with the trial residual at x + delta
//   k_lin_fused       (default) k_lin_points + k_cam_side + k_schur in one kernel: per-observation records live only in
//                     shared memory (k_fused.cuh); the three kernels above remain as the fallback / A-B path
#include <type_traits>

#include "ba_internal.h"

namespace b200ba {

namespace {

constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(FULL, v, d);
  return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v = fmax(v, __shfl_xor_sync(FULL, v, d));
  return v;
}

// CTA reduction of NV values per thread; result valid in thread 0. blockDim.x multiple of 32, <= 1024.
template <int NV>
__device__ __forceinline__ void block_sum(double* vals, double* smem /*[32*NV]*/) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    double s = warp_sum(vals[k]);
    if (lane == 0) smem[wid * NV + k] = s;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      double s = 0.0;
      for (int w = 0; w < nw; ++w) s += smem[w * NV + k];
      vals[k] = s;
    }
  }
  __syncthreads();
}

// ------------------------------------------------------------------ camera prep
// With a step dc: prep of the trial cameras x + dc; cstep (nullable, field-major [6][Cs]) receives {J_r(w) dw, dT}
// with J_r taken from camprep_old (the prep of x), the per-camera part of the back-substitution (project_step).
__global__ void k_cam_prep(int C, int64_t Cs, const double* __restrict__ xc, const double* __restrict__ dc,
                           double* __restrict__ x_out, double* __restrict__ camprep,
                           const double* __restrict__ camprep_old, double* __restrict__ cstep) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double cam[6];
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    cam[k] = xc[6 * c + k] + (dc ? dc[6 * c + k] : 0.0);
    if (x_out) x_out[6 * c + k] = cam[k];
  }
  if (cstep) {
    const double d0 = dc[6 * c], d1 = dc[6 * c + 1], d2 = dc[6 * c + 2];
#pragma unroll
    for (int r = 0; r < 3; ++r)
      cstep[(int64_t)r * Cs + c] = camprep_old[(int64_t)(9 + 3 * r) * Cs + c] * d0 +
                                   camprep_old[(int64_t)(10 + 3 * r) * Cs + c] * d1 +
                                   camprep_old[(int64_t)(11 + 3 * r) * Cs + c] * d2;
#pragma unroll
    for (int r = 0; r < 3; ++r) cstep[(int64_t)(3 + r) * Cs + c] = dc[6 * c + 3 + r];
  }
  double out[CAMPREP_STRIDE];
  cam_prep(cam, out);
#pragma unroll
  for (int k = 0; k < CAMPREP_FIELDS; ++k) camprep[(int64_t)k * Cs + c] = out[k];  // field-major: see load_cam
}

// camprep is stored field-major ([21][Cs], Cs = C rounded up to even): the lanes of a warp look at a handful of neighbouring cameras, so one
// load instruction touches 1-2 cache lines instead of one line per distinct camera.
__device__ __forceinline__ void load_cam(const double* __restrict__ camprep, int64_t C, int c, double* __restrict__ cp) {
#pragma unroll
  for (int k = 0; k < CAMPREP_FIELDS; ++k) cp[k] = camprep[(int64_t)k * C + c];
}

// R -> cp[0..8], T -> cp[18..20] (the fields project_residual reads); the J_r fields are skipped
__device__ __forceinline__ void load_rt(const double* __restrict__ camprep, int64_t Cs, int c, double* __restrict__ cp) {
#pragma unroll
  for (int k = 0; k < 9; ++k) cp[k] = camprep[(int64_t)k * Cs + c];
#pragma unroll
  for (int k = 18; k < 21; ++k) cp[k] = camprep[(int64_t)k * Cs + c];
}

// ------------------------------------------------------------------ residual
__global__ void __launch_bounds__(256) k_residual(Problem pb, const double* __restrict__ x,
                                                  const double* __restrict__ camprep, double* __restrict__ r_out,
                                                  double* __restrict__ part) {
  __shared__ double sm[32];
  const double* xp = x + 6 * pb.C;
  double acc = 0.0;
  // The gathers (camera table, point) depend on the index loads: the indices and pixels of the NEXT observation of this
  // thread are requested before the current one's gathers, so a loop iteration waits for one trip to memory, not two
  // (ncu: the kernel is latency-bound, long-scoreboard stalls, DRAM at 41 % of peak without this).
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  int c = 0, p = 0;
  double uo = 0.0, vo = 0.0;
  if (i < pb.N) { c = pb.cam[i]; p = pb.pt[i]; uo = pb.u[i]; vo = pb.v[i]; }
  while (i < pb.N) {
    const int64_t in = i + stride;
    int cn = 0, pn = 0;
    double un = 0.0, vn = 0.0;
    if (in < pb.N) { cn = pb.cam[in]; pn = pb.pt[in]; un = pb.u[in]; vn = pb.v[in]; }
    double r0, r1, cp[CAMPREP_FIELDS];
    load_rt(camprep, pb.Cs, c, cp);   // R and T only: the residual does not read J_r (12 of the 21 gathered fields)
    project_residual(cp, pb.K, xp[3 * p], xp[3 * p + 1], xp[3 * p + 2], uo, vo, r0, r1);
    acc += r0 * r0 + r1 * r1;
    if (r_out) {
      const int64_t d = pb.perm ? pb.perm[i] : i;
      reinterpret_cast<double2*>(r_out)[d] = make_double2(r0, r1);
    }
    i = in; c = cn; p = pn; uo = un; vo = vn;
  }
  block_sum<1>(&acc, sm);
  if (threadIdx.x == 0) part[blockIdx.x * 8] = acc;
}

// ------------------------------------------------------------------ pose <-> parameter conversion at the boundary
// pack   (sfm.py:252-257): per registered camera  [Rot.from_matrix(H[:3,:3]).as_rotvec() | H[:3,3]]
// unpack (sfm.py:273-278): H = [[Rot.from_rotvec(w).as_matrix(), T], [0, 1]]
// The arithmetic restates scipy.spatial.transform (1.18, _rotation_xp.py): from_matrix = orthogonality test
// |R R^T - I| <= 1e-12, else the orthogonal polar factor (scipy: U V^T of the SVD; here Newton's iteration
// X <- (X + X^-T) / 2, which converges to the same matrix), then Markley's choice of the largest of
// {R00, R11, R22, trace}, normalisation; as_rotvec = canonical sign (w >= 0), angle = 2 atan2(|v|, w), series below
// 1e-3; from_rotvec / as_matrix likewise.  One thread per camera: C <= 1e4, the point is that no Python object per
// camera is created (the reference builds C scipy Rotation objects per BA call, each way).
__device__ __forceinline__ void polar3(double* R) {
  for (int it = 0; it < 32; ++it) {
    const double c00 = R[4] * R[8] - R[5] * R[7], c01 = R[5] * R[6] - R[3] * R[8], c02 = R[3] * R[7] - R[4] * R[6];
    const double det = R[0] * c00 + R[1] * c01 + R[2] * c02;
    const double id = 1.0 / det;
    // X^-T = cofactor matrix / det
    double T[9];
    T[0] = c00 * id; T[1] = c01 * id; T[2] = c02 * id;
    T[3] = (R[2] * R[7] - R[1] * R[8]) * id; T[4] = (R[0] * R[8] - R[2] * R[6]) * id; T[5] = (R[1] * R[6] - R[0] * R[7]) * id;
    T[6] = (R[1] * R[5] - R[2] * R[4]) * id; T[7] = (R[2] * R[3] - R[0] * R[5]) * id; T[8] = (R[0] * R[4] - R[1] * R[3]) * id;
    double diff = 0.0;
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      const double n = 0.5 * (R[k] + T[k]);
      diff = fmax(diff, fabs(n - R[k]));
      R[k] = n;
    }
    if (diff < 1e-16) break;
  }
}

__global__ void k_pack_cameras(int64_t n, const double* __restrict__ H, double* __restrict__ params) {
  const int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (c >= n) return;
  const double* h = H + 16 * c;
  double R[9] = {h[0], h[1], h[2], h[4], h[5], h[6], h[8], h[9], h[10]};
  bool orth = true;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const double g = R[3 * i] * R[3 * j] + R[3 * i + 1] * R[3 * j + 1] + R[3 * i + 2] * R[3 * j + 2];
      const double e = (i == j) ? 1.0 : 0.0;
      orth = orth && (fabs(g - e) <= 1e-12 + 1e-5 * fabs(e));   // isclose(gramian, eye, atol = 1e-12), rtol = 1e-5 default
    }
  if (!orth) polar3(R);
  const double tr = R[0] + R[4] + R[8];
  double q[4];
  int choice = 0;
  double best = R[0];
  if (R[4] > best) { best = R[4]; choice = 1; }
  if (R[8] > best) { best = R[8]; choice = 2; }
  if (tr > best) choice = 3;
  if (choice == 0) { q[0] = 1.0 - tr + 2.0 * R[0]; q[1] = R[3] + R[1]; q[2] = R[6] + R[2]; q[3] = R[7] - R[5]; }
  else if (choice == 1) { q[0] = R[3] + R[1]; q[1] = 1.0 - tr + 2.0 * R[4]; q[2] = R[7] + R[5]; q[3] = R[2] - R[6]; }
  else if (choice == 2) { q[0] = R[6] + R[2]; q[1] = R[7] + R[5]; q[2] = 1.0 - tr + 2.0 * R[8]; q[3] = R[3] - R[1]; }
  else { q[0] = R[7] - R[5]; q[1] = R[2] - R[6]; q[2] = R[3] - R[1]; q[3] = 1.0 + tr; }
  const double qn = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
#pragma unroll
  for (int k = 0; k < 4; ++k) q[k] /= qn;
  if (q[3] < 0.0 || (q[3] == 0.0 && (q[0] < 0.0 || (q[0] == 0.0 && (q[1] < 0.0 || (q[1] == 0.0 && q[2] < 0.0))))))
#pragma unroll
    for (int k = 0; k < 4; ++k) q[k] = -q[k];
  const double an = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2]);
  const double angle = 2.0 * atan2(an, q[3]);
  const double a2 = angle * angle;
  const double scale = (angle <= 1e-3) ? 2.0 + a2 / 12.0 + 7.0 * a2 * a2 / 2880.0 : angle / sin(angle / 2.0);
  double* o = params + 6 * c;
  o[0] = scale * q[0]; o[1] = scale * q[1]; o[2] = scale * q[2];
  o[3] = h[3]; o[4] = h[7]; o[5] = h[11];
}

__global__ void k_unpack_cameras(int64_t n, const double* __restrict__ params, double* __restrict__ H) {
  const int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (c >= n) return;
  const double* p = params + 6 * c;
  const double angle = sqrt(p[0] * p[0] + p[1] * p[1] + p[2] * p[2]);
  const double a2 = angle * angle;
  const double scale = (angle <= 1e-3) ? 0.5 - a2 / 48.0 + a2 * a2 / 3840.0 : sin(angle / 2.0) / angle;
  const double x = p[0] * scale, y = p[1] * scale, z = p[2] * scale, w = cos(angle / 2.0);
  const double x2 = x * x, y2 = y * y, z2 = z * z, w2 = w * w;
  const double xy = x * y, zw = z * w, xz = x * z, yw = y * w, yz = y * z, xw = x * w;
  double* h = H + 16 * c;
  h[0] = x2 - y2 - z2 + w2; h[1] = 2.0 * (xy - zw);    h[2] = 2.0 * (xz + yw);     h[3] = p[3];
  h[4] = 2.0 * (xy + zw);   h[5] = -x2 + y2 - z2 + w2; h[6] = 2.0 * (yz - xw);     h[7] = p[4];
  h[8] = 2.0 * (xz - yw);   h[9] = 2.0 * (yz + xw);    h[10] = -x2 - y2 + z2 + w2; h[11] = p[5];
  h[12] = 0.0; h[13] = 0.0; h[14] = 0.0; h[15] = 1.0;
}

#include "k_pnp.cuh"

// ------------------------------------------------------------------ two-view triangulation (sfm.py:214-221)
// cv2.triangulatePoints(M1, M2, pts1.T, pts2.T) as called at sfm.py:216, i.e. the DLT the reference restates as
// triangulate_points_linear2 (cv2_lite/triangulate_points.py:66-94): per point the 4 x 4 system
//   A = [x1 M1[2] - M1[0];  y1 M1[2] - M1[1];  x2 M2[2] - M2[0];  y2 M2[2] - M2[1]],   X = right singular vector of the
// smallest singular value, divided by its last entry (sfm.py:217).  One thread per point; the SVD is a one-sided
// (Hestenes) Jacobi iteration on the columns of A -- it works on A itself (not A^T A), so the null vector keeps full
// relative accuracy.  vis1 / vis2 (nullable): third row of inv(H1) / inv(H2); the mask is sfm.py:219-221.
__global__ void k_triangulate(int64_t n, const double* __restrict__ M1, const double* __restrict__ M2,
                              const double* __restrict__ pts1, const double* __restrict__ pts2,
                              const double* __restrict__ vis1, const double* __restrict__ vis2, double* __restrict__ X_out,
                              unsigned char* __restrict__ mask_out) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double x1 = pts1[2 * i], y1 = pts1[2 * i + 1], x2 = pts2[2 * i], y2 = pts2[2 * i + 1];
  double A[4][4], V[4][4];   // A[row][col]
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    A[0][k] = x1 * M1[8 + k] - M1[k];
    A[1][k] = y1 * M1[8 + k] - M1[4 + k];
    A[2][k] = x2 * M2[8 + k] - M2[k];
    A[3][k] = y2 * M2[8 + k] - M2[4 + k];
#pragma unroll
    for (int j = 0; j < 4; ++j) V[k][j] = (k == j) ? 1.0 : 0.0;
  }
  for (int sweep = 0; sweep < 40; ++sweep) {
    bool rotated = false;
#pragma unroll
    for (int p = 0; p < 3; ++p)
#pragma unroll
      for (int q = p + 1; q < 4; ++q) {
        double al = 0.0, be = 0.0, ga = 0.0;
#pragma unroll
        for (int r = 0; r < 4; ++r) { al += A[r][p] * A[r][p]; be += A[r][q] * A[r][q]; ga += A[r][p] * A[r][q]; }
        if (fabs(ga) > 1e-16 * sqrt(al * be) && ga != 0.0) {
          rotated = true;
          const double zeta = (be - al) / (2.0 * ga);
          const double t = (zeta >= 0.0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
          const double c = 1.0 / sqrt(1.0 + t * t), sn = c * t;
#pragma unroll
          for (int r = 0; r < 4; ++r) {
            const double ap = A[r][p], aq = A[r][q];
            A[r][p] = c * ap - sn * aq; A[r][q] = sn * ap + c * aq;
            const double vp = V[r][p], vq = V[r][q];
            V[r][p] = c * vp - sn * vq; V[r][q] = sn * vp + c * vq;
          }
        }
      }
    if (!rotated) break;
  }
  int best = 0;
  double bn = 1e300;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const double nk = A[0][k] * A[0][k] + A[1][k] * A[1][k] + A[2][k] * A[2][k] + A[3][k] * A[3][k];
    if (nk < bn) { bn = nk; best = k; }
  }
  double X[4];
#pragma unroll
  for (int r = 0; r < 4; ++r) X[r] = best == 0 ? V[r][0] : best == 1 ? V[r][1] : best == 2 ? V[r][2] : V[r][3];
  const double iw = 1.0 / X[3];
  X[0] *= iw; X[1] *= iw; X[2] *= iw; X[3] = 1.0;
#pragma unroll
  for (int r = 0; r < 4; ++r) X_out[4 * i + r] = X[r];
  if (mask_out) {
    const double z1 = vis1[0] * X[0] + vis1[1] * X[1] + vis1[2] * X[2] + vis1[3];
    const double z2 = vis2[0] * X[0] + vis2[1] * X[1] + vis2[2] * X[2] + vis2[3];
    mask_out[i] = (z1 > 0.0 && z2 > 0.0) ? 1 : 0;
  }
}


// ------------------------------------------------------------------ non-linear triangulation refinement
// cv2_lite/triangulate_points.py:122-184: from the DLT point, minimise the reference's objective (:94-118) -- its
// residuals are the SQUARED reprojection errors of the four image coordinates, squared again by the solver, i.e. the sum
// of 4th powers.  The reference passes all points to one scipy call; the points are independent, so one thread runs a
// damped Gauss-Newton on its own 3 unknowns (4 residuals r_k = e_k^2, J_k = 2 e_k de_k/dX) to the minimiser.
__global__ void k_triangulate_refine(int64_t n, const double* __restrict__ M1, const double* __restrict__ M2,
                                     const double* __restrict__ pts1, const double* __restrict__ pts2, int max_iter,
                                     double* __restrict__ X /* n x 4, [X Y Z 1] in and out */) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double obs[4] = {pts1[2 * i], pts1[2 * i + 1], pts2[2 * i], pts2[2 * i + 1]};
  double x[3] = {X[4 * i], X[4 * i + 1], X[4 * i + 2]};
  if (!(x[0] == x[0]) || !(fabs(x[0]) < 1e300)) return;
  auto eval = [&](const double* xx, double* r, double (*J)[3]) -> double {
    double c = 0.0;
#pragma unroll
    for (int v = 0; v < 2; ++v) {
      const double* M = v ? M2 : M1;
      const double a0 = M[0] * xx[0] + M[1] * xx[1] + M[2] * xx[2] + M[3];
      const double a1 = M[4] * xx[0] + M[5] * xx[1] + M[6] * xx[2] + M[7];
      const double a2 = M[8] * xx[0] + M[9] * xx[1] + M[10] * xx[2] + M[11];
      const double ia = 1.0 / a2, pu = a0 * ia, pv = a1 * ia;
      const double e0 = pu - obs[2 * v], e1 = pv - obs[2 * v + 1];
      r[2 * v] = e0 * e0;
      r[2 * v + 1] = e1 * e1;
      c += r[2 * v] * r[2 * v] + r[2 * v + 1] * r[2 * v + 1];
      if (J) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          J[2 * v][k] = 2.0 * e0 * (M[k] - pu * M[8 + k]) * ia;
          J[2 * v + 1][k] = 2.0 * e1 * (M[4 + k] - pv * M[8 + k]) * ia;
        }
      }
    }
    return c;
  };
  double r[4], J[4][3], lam = 1e-3;
  double c = eval(x, r, J);
  for (int it = 0; it < max_iter; ++it) {
    double H[6] = {0, 0, 0, 0, 0, 0}, g[3] = {0, 0, 0};   // H: 00 10 11 20 21 22
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      H[0] += J[k][0] * J[k][0]; H[1] += J[k][1] * J[k][0]; H[2] += J[k][1] * J[k][1];
      H[3] += J[k][2] * J[k][0]; H[4] += J[k][2] * J[k][1]; H[5] += J[k][2] * J[k][2];
      g[0] += J[k][0] * r[k]; g[1] += J[k][1] * r[k]; g[2] += J[k][2] * r[k];
    }
    double Li[6];
    if (!point_factor(H, lam, Li)) { lam *= 4.0; if (lam > 1e12) break; continue; }
    // dx = -(L L^T)^-1 g = -Linv^T (Linv g)
    const double z0 = Li[0] * g[0], z1 = Li[1] * g[0] + Li[2] * g[1], z2 = Li[3] * g[0] + Li[4] * g[1] + Li[5] * g[2];
    const double dx[3] = {-(Li[0] * z0 + Li[1] * z1 + Li[3] * z2), -(Li[2] * z1 + Li[4] * z2), -(Li[5] * z2)};
    const double xn[3] = {x[0] + dx[0], x[1] + dx[1], x[2] + dx[2]};
    double rn[4], Jn[4][3];
    const double cn = eval(xn, rn, Jn);
    if (cn == cn && cn < c) {
      const double dn = sqrt(dx[0] * dx[0] + dx[1] * dx[1] + dx[2] * dx[2]);
      const double xs = sqrt(x[0] * x[0] + x[1] * x[1] + x[2] * x[2]);
      const bool done = (c - cn) < 1e-16 * c || dn < 1e-14 * xs;
      x[0] = xn[0]; x[1] = xn[1]; x[2] = xn[2];
#pragma unroll
      for (int k = 0; k < 4; ++k) { r[k] = rn[k]; J[k][0] = Jn[k][0]; J[k][1] = Jn[k][1]; J[k][2] = Jn[k][2]; }
      c = cn;
      lam = fmax(lam / 3.0, 1e-12);
      if (done) break;
    } else {
      lam *= 4.0;
      if (lam > 1e12) break;
    }
  }
  X[4 * i] = x[0]; X[4 * i + 1] = x[1]; X[4 * i + 2] = x[2];
}

// ------------------------------------------------------------------ per-observation reprojection error (sfm.py:234-238)
// The loop of SFM._apply_increment calls calc_reproj_error (sfm.py:38-41) -> reproj_error (cv2_lite/solve_pnp.py:9-14)
// once per observation: || pi(K (R X + t)) - uv ||, the pipeline's R X + t convention (NOT the R (X - T) of the BA
// residual).  Rt = [R row-major (9) | t (3)] per camera; 24 B/observation in, 8 B out.
__global__ void __launch_bounds__(256) k_reproj_err(Problem pb, const double* __restrict__ Rt, const double* __restrict__ X,
                                                    double* __restrict__ err_out, double* __restrict__ part) {
  __shared__ double sm[32];
  double acc = 0.0;
  // indices and pixels of this thread's next observation are requested ahead of the current gathers (see k_residual)
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  int c = 0, p = 0;
  double uo = 0.0, vo = 0.0;
  if (i < pb.N) { c = pb.cam[i]; p = pb.pt[i]; uo = pb.u[i]; vo = pb.v[i]; }
  while (i < pb.N) {
    const int64_t in = i + stride;
    int cn = 0, pn = 0;
    double un = 0.0, vn = 0.0;
    if (in < pb.N) { cn = pb.cam[in]; pn = pb.pt[in]; un = pb.u[in]; vn = pb.v[in]; }
    const double* q = Rt + 12 * (int64_t)c;
    const double X0 = X[3 * (int64_t)p], X1 = X[3 * (int64_t)p + 1], X2 = X[3 * (int64_t)p + 2];
    const double z0 = q[0] * X0 + q[1] * X1 + q[2] * X2 + q[9];
    const double z1 = q[3] * X0 + q[4] * X1 + q[5] * X2 + q[10];
    const double z2 = q[6] * X0 + q[7] * X1 + q[8] * X2 + q[11];
    const double p0 = pb.K.k[0] * z0 + pb.K.k[1] * z1 + pb.K.k[2] * z2;
    const double p1 = pb.K.k[3] * z0 + pb.K.k[4] * z1 + pb.K.k[5] * z2;
    const double p2 = pb.K.k[6] * z0 + pb.K.k[7] * z1 + pb.K.k[8] * z2;
    const double du = p0 / p2 - uo, dv = p1 / p2 - vo;
    const double e = sqrt(du * du + dv * dv);
    acc += e;
    if (err_out) err_out[pb.perm ? pb.perm[i] : i] = e;
    i = in; c = cn; p = pn; uo = un; vo = vn;
  }
  block_sum<1>(&acc, sm);
  if (threadIdx.x == 0) part[blockIdx.x * 8] = acc;
}

// ------------------------------------------------------------------ materialised Jacobian blocks (config E)
// One observation per thread; outputs staged in shared memory so the 16 + 96 + 48 B per observation leave the
// SM as fully coalesced 8-byte-strided stores (identity order). With a caller permutation the stores scatter.
constexpr int JT = 256;
__global__ void __launch_bounds__(JT) k_jacobian(Problem pb, const double* __restrict__ x,
                                                 const double* __restrict__ camprep, double* __restrict__ r_out,
                                                 double* __restrict__ Jc_out, double* __restrict__ Jp_out) {
  __shared__ double sJc[JT * 13];
  __shared__ double sJp[JT * 7];
  const double* xp = x + 6 * pb.C;
  const int64_t base = blockIdx.x * (int64_t)JT;
  const int64_t i = base + threadIdx.x;
  double r0 = 0, r1 = 0, Jc[12], Jp[6];
  if (i < pb.N) {
    const int c = pb.cam[i], p = pb.pt[i];
    double cp[CAMPREP_FIELDS];
    load_cam(camprep, pb.Cs, c, cp);
    project_jac(cp, pb.K, xp[3 * p], xp[3 * p + 1], xp[3 * p + 2], pb.u[i], pb.v[i], r0, r1, Jc, Jp);
  }
  if (pb.perm) {
    if (i < pb.N) {
      const int64_t d = pb.perm[i];
      if (r_out) reinterpret_cast<double2*>(r_out)[d] = make_double2(r0, r1);
#pragma unroll
      for (int k = 0; k < 12; ++k) Jc_out[12 * d + k] = Jc[k];
#pragma unroll
      for (int k = 0; k < 6; ++k) Jp_out[6 * d + k] = Jp[k];
    }
    return;
  }
  if (i < pb.N && r_out) reinterpret_cast<double2*>(r_out)[i] = make_double2(r0, r1);
#pragma unroll
  for (int k = 0; k < 12; ++k) sJc[threadIdx.x * 13 + k] = Jc[k];
#pragma unroll
  for (int k = 0; k < 6; ++k) sJp[threadIdx.x * 7 + k] = Jp[k];
  __syncthreads();
  const int64_t n_here = min((int64_t)JT, pb.N - base);
  for (int e = threadIdx.x; e < n_here * 12; e += JT) Jc_out[base * 12 + e] = sJc[(e / 12) * 13 + (e % 12)];
  for (int e = threadIdx.x; e < n_here * 6; e += JT) Jp_out[base * 6 + e] = sJp[(e / 6) * 7 + (e % 6)];
}

#include "k_lin.cuh"
#include "k_cam_schur.cuh"
#include "k_backsub.cuh"
#include "k_fused.cuh"
#include "k_flow.cuh"

}  // namespace

#include "k_launch.cuh"
#include "k_noalloc.cuh"

}  // namespace b200ba
