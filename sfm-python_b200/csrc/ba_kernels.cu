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

// ------------------------------------------------------------------ residual
__global__ void __launch_bounds__(256) k_residual(Problem pb, const double* __restrict__ x,
                                                  const double* __restrict__ camprep, double* __restrict__ r_out,
                                                  double* __restrict__ part) {
  __shared__ double sm[32];
  const double* xp = x + 6 * pb.C;
  double acc = 0.0;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < pb.N; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = pb.cam[i], p = pb.pt[i];
    double r0, r1, cp[CAMPREP_FIELDS];
    load_cam(camprep, pb.Cs, c, cp);
    project_residual(cp, pb.K, xp[3 * p], xp[3 * p + 1], xp[3 * p + 2], pb.u[i], pb.v[i],
                     r0, r1);
    acc += r0 * r0 + r1 * r1;
    if (r_out) {
      const int64_t d = pb.perm ? pb.perm[i] : i;
      reinterpret_cast<double2*>(r_out)[d] = make_double2(r0, r1);
    }
  }
  block_sum<1>(&acc, sm);
  if (threadIdx.x == 0) part[blockIdx.x * 8] = acc;
}

// ------------------------------------------------------------------ per-observation reprojection error (sfm.py:234-238)
// The loop of SFM._apply_increment calls calc_reproj_error (sfm.py:38-41) -> reproj_error (cv2_lite/solve_pnp.py:9-14)
// once per observation: || pi(K (R X + t)) - uv ||, the pipeline's R X + t convention (NOT the R (X - T) of the BA
// residual).  Rt = [R row-major (9) | t (3)] per camera; 24 B/observation in, 8 B out.
__global__ void __launch_bounds__(256) k_reproj_err(Problem pb, const double* __restrict__ Rt, const double* __restrict__ X,
                                                    double* __restrict__ err_out, double* __restrict__ part) {
  __shared__ double sm[32];
  double acc = 0.0;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < pb.N; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = pb.cam[i], p = pb.pt[i];
    const double* q = Rt + 12 * (int64_t)c;
    const double X0 = X[3 * (int64_t)p], X1 = X[3 * (int64_t)p + 1], X2 = X[3 * (int64_t)p + 2];
    const double z0 = q[0] * X0 + q[1] * X1 + q[2] * X2 + q[9];
    const double z1 = q[3] * X0 + q[4] * X1 + q[5] * X2 + q[10];
    const double z2 = q[6] * X0 + q[7] * X1 + q[8] * X2 + q[11];
    const double p0 = pb.K.k[0] * z0 + pb.K.k[1] * z1 + pb.K.k[2] * z2;
    const double p1 = pb.K.k[3] * z0 + pb.K.k[4] * z1 + pb.K.k[5] * z2;
    const double p2 = pb.K.k[6] * z0 + pb.K.k[7] * z1 + pb.K.k[8] * z2;
    const double du = p0 / p2 - pb.u[i], dv = p1 / p2 - pb.v[i];
    const double e = sqrt(du * du + dv * dv);
    acc += e;
    if (err_out) err_out[pb.perm ? pb.perm[i] : i] = e;
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

}  // namespace b200ba
