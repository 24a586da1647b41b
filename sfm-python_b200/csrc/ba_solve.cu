// Reduced-camera-system solvers: S dc = b   (replaces scipy's LSMR step, scipy _lsq/trf.py).
//
//   k_pcg       block-Jacobi preconditioned CG on the block-CSR Schur complement, ONE persistent cooperative
//               kernel for the whole solve (grid.sync between phases, fixed-order reductions => the same bits on
//               every rank of a multi-GPU run, which keeps the replicated accept/reject decisions in lock-step).
//   k_cholesky  dense right-looking Cholesky + triangular solves for small camera counts.
#include <cooperative_groups.h>

#include "ba_internal.h"

namespace cg = cooperative_groups;

namespace b200ba {

namespace {

constexpr unsigned FULL = 0xffffffffu;
constexpr int PCG_T = 256;

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(FULL, v, d);
  return v;
}

struct PcgArgs {
  int C;
  const int* rowptr; const int* col; const int* src; const int* diag_blk;
  const double* Sup; const double* b;
  double* x; double* r; double* z; double* p; double* Ap; double* s; double* Minv;
  double* part;  // [2 * gridDim * 4]
  double* scal;
  double rtol; int max_iter;
};

// grid-wide sum of up to 3 values; every thread of every CTA returns the same bits
// `part` is double-buffered ([2][gridDim*4], toggled by `phase`): a fast CTA may start the next reduction while a
// slow one still reads this one's partials; two grid.sync()s always separate reuse of the same half.
__device__ __forceinline__ void grid_sum3(cg::grid_group& grid, double* part_base, int& phase, double* sm, double v0,
                                          double v1, double v2, double& o0, double& o1, double& o2) {
  double* part = part_base + (size_t)(phase & 1) * gridDim.x * 4;
  phase ^= 1;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
  v0 = warp_sum(v0); v1 = warp_sum(v1); v2 = warp_sum(v2);
  if (lane == 0) { sm[wid * 3] = v0; sm[wid * 3 + 1] = v1; sm[wid * 3 + 2] = v2; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double s0 = 0, s1 = 0, s2 = 0;
    for (int w = 0; w < nw; ++w) { s0 += sm[w * 3]; s1 += sm[w * 3 + 1]; s2 += sm[w * 3 + 2]; }
    part[blockIdx.x * 4] = s0; part[blockIdx.x * 4 + 1] = s1; part[blockIdx.x * 4 + 2] = s2;
  }
  grid.sync();
  if (wid == 0) {
    double s0 = 0, s1 = 0, s2 = 0;
    for (int b = lane; b < (int)gridDim.x; b += 32) {
      s0 += part[b * 4]; s1 += part[b * 4 + 1]; s2 += part[b * 4 + 2];
    }
    s0 = warp_sum(s0); s1 = warp_sum(s1); s2 = warp_sum(s2);
    if (lane == 0) { sm[96] = s0; sm[97] = s1; sm[98] = s2; }
  }
  __syncthreads();
  o0 = sm[96]; o1 = sm[97]; o2 = sm[98];
  __syncthreads();
}

// inverse of a symmetric positive definite 6x6 block (row-major in, row-major out); Jacobi fallback
__device__ void inv6(const double* __restrict__ A, double* __restrict__ Mi) {
  double L[6][6], Li[6][6];
  bool ok = true;
#pragma unroll
  for (int i = 0; i < 6; ++i)
#pragma unroll
    for (int j = 0; j < 6; ++j) { L[i][j] = 0.0; Li[i][j] = 0.0; }
#pragma unroll
  for (int j = 0; j < 6; ++j) {
    double d = A[j * 6 + j];
#pragma unroll
    for (int k = 0; k < j; ++k) d -= L[j][k] * L[j][k];
    if (!(d > 0.0)) { ok = false; d = 1.0; }
    const double id = 1.0 / sqrt(d);
    L[j][j] = d * id;
#pragma unroll
    for (int i = j + 1; i < 6; ++i) {
      double s = A[i * 6 + j];
#pragma unroll
      for (int k = 0; k < j; ++k) s -= L[i][k] * L[j][k];
      L[i][j] = s * id;
    }
  }
  // Li = L^-1 (lower)
#pragma unroll
  for (int j = 0; j < 6; ++j) {
    Li[j][j] = 1.0 / L[j][j];
#pragma unroll
    for (int i = j + 1; i < 6; ++i) {
      double s = 0.0;
#pragma unroll
      for (int k = j; k < i; ++k) s -= L[i][k] * Li[k][j];
      Li[i][j] = s / L[i][i];
    }
  }
#pragma unroll
  for (int i = 0; i < 6; ++i)
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      double s = 0.0;
#pragma unroll
      for (int k = 0; k < 6; ++k) s += (k >= i && k >= j) ? Li[k][i] * Li[k][j] : 0.0;
      if (!ok) s = (i == j) ? 1.0 / fmax(fabs(A[i * 6 + i]), 1e-300) : 0.0;
      Mi[i * 6 + j] = s;
    }
}

// y[6c .. 6c+5] = (S v)[camera c] for the block rows of this warp; returns this lane's share of (y, v) (lanes 0..5).
// Lane (kq, r) = (lane / 6, lane % 6), lanes 0..29, multiplies row r of every 5th block of the row; two steps (10
// blocks) are in flight together, B rows and v are read as three 16-byte loads each.
__device__ __forceinline__ double spmv_rows(const PcgArgs& a, const double* __restrict__ v, double* __restrict__ y,
                                            int gwarp, int gwarps, int lane) {
  const int kq = lane / 6, r = lane % 6;
  double dot = 0.0;
  for (int c = gwarp; c < a.C; c += gwarps) {
    const int q0 = a.rowptr[c], q1 = a.rowptr[c + 1];
    double acc = 0.0;
    if (lane < 30) {
      for (int q = q0 + kq; q < q1; q += 10) {
        const int qb = q + 5;
        const bool two = qb < q1;
        const int s0 = a.src[q], c0 = a.col[q];
        const int s1 = two ? a.src[qb] : s0, c1 = two ? a.col[qb] : c0;
        const double* B0 = a.Sup + (int64_t)(s0 & 0x7fffffff) * 36;
        const double* B1 = a.Sup + (int64_t)(s1 & 0x7fffffff) * 36;
        const double2* v0 = reinterpret_cast<const double2*>(v + 6 * c0);
        const double2* v1 = reinterpret_cast<const double2*>(v + 6 * c1);
        double b0[6], b1[6];
        if (s0 < 0) {
#pragma unroll
          for (int j = 0; j < 6; ++j) b0[j] = B0[j * 6 + r];
        } else {
          const double2* p = reinterpret_cast<const double2*>(B0 + 6 * r);
          const double2 t0 = p[0], t1 = p[1], t2 = p[2];
          b0[0] = t0.x; b0[1] = t0.y; b0[2] = t1.x; b0[3] = t1.y; b0[4] = t2.x; b0[5] = t2.y;
        }
        if (s1 < 0) {
#pragma unroll
          for (int j = 0; j < 6; ++j) b1[j] = B1[j * 6 + r];
        } else {
          const double2* p = reinterpret_cast<const double2*>(B1 + 6 * r);
          const double2 t0 = p[0], t1 = p[1], t2 = p[2];
          b1[0] = t0.x; b1[1] = t0.y; b1[2] = t1.x; b1[3] = t1.y; b1[4] = t2.x; b1[5] = t2.y;
        }
        const double2 x0 = v0[0], x1 = v0[1], x2 = v0[2];
        const double2 z0 = v1[0], z1 = v1[1], z2 = v1[2];
        acc += b0[0] * x0.x + b0[1] * x0.y + b0[2] * x1.x + b0[3] * x1.y + b0[4] * x2.x + b0[5] * x2.y;
        if (two) acc += b1[0] * z0.x + b1[1] * z0.y + b1[2] * z1.x + b1[3] * z1.y + b1[4] * z2.x + b1[5] * z2.y;
      }
    }
    double tot = 0.0;
#pragma unroll
    for (int g = 0; g < 5; ++g) tot += __shfl_sync(FULL, acc, (lane % 6) + 6 * g);
    if (lane < 6) { y[6 * c + lane] = tot; dot += tot * v[6 * c + lane]; }
  }
  return dot;
}

// Block-Jacobi preconditioned CG in the Chronopoulos-Gear form: one mat-vec, one fused reduction and TWO grid-wide
// synchronisations per iteration (the textbook form needs four).  With u = M^-1 r, w = S u:
//     p = u + beta p,  s = w + beta s,  x += alpha p,  r -= alpha s,  u = M^-1 r            (local, per camera)
//     w = S u                                                                                (after sync 1)
//     gamma' = (r,u), delta = (w,u), rr = (r,r)                                              (one reduction = sync 2)
//     beta = gamma'/gamma,  alpha = gamma' / (delta - beta gamma' / alpha)
// All reductions are in fixed order, so every rank of a multi-GPU run computes the same bits.
__global__ void __launch_bounds__(PCG_T, 3) k_pcg(PcgArgs a) {
  cg::grid_group grid = cg::this_grid();
  __shared__ double sm[128];
  const int C = a.C, n = 6 * C;
  const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
  const int gthreads = gridDim.x * blockDim.x;
  const int lane = threadIdx.x & 31;
  const int gwarp = gtid >> 5, gwarps = gthreads >> 5;
  double* u = a.z;   // preconditioned residual
  double* w = a.Ap;  // S u
  double* s = a.s;   // S p

  // ---- init: block-Jacobi inverse, x = 0, r = b, u = Minv r, p = s = 0
  double gam = 0.0, bb = 0.0;
  int phase = 0;
  for (int c = gtid; c < C; c += gthreads) {
    inv6(a.Sup + (int64_t)a.diag_blk[c] * 36, a.Minv + (int64_t)c * 36);
    double rv[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      rv[k] = a.b[6 * c + k];
      a.x[6 * c + k] = 0.0; a.r[6 * c + k] = rv[k]; a.p[6 * c + k] = 0.0; s[6 * c + k] = 0.0;
      bb += rv[k] * rv[k];
    }
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      double t = 0.0;
#pragma unroll
      for (int k = 0; k < 6; ++k) t += a.Minv[(int64_t)c * 36 + i * 6 + k] * rv[k];
      u[6 * c + i] = t; gam += t * rv[i];
    }
  }
  grid.sync();
  double del = spmv_rows(a, u, w, gwarp, gwarps, lane);
  double d0, d1, d2;
  grid_sum3(grid, a.part, phase, sm, gam, del, bb, d0, d1, d2);
  gam = d0; del = d1; bb = d2;
  double rr = bb, alpha = 0.0, beta = 0.0;
  int it = 0, fail = 0;
  const double thresh = a.rtol * a.rtol * bb;
  if (bb > 0.0) {
    if (!(del > 0.0)) fail = 1;
    else alpha = gam / del;
  }
  while (!fail && it < a.max_iter && rr > thresh && bb > 0.0) {
    // ---- local updates, one lane per (camera, component) -- 5 cameras per warp step: p, s, x, r elementwise, then
    //      u = Minv r with the camera's six r values exchanged by shuffles; partial (r,u) and (r,r)
    double gam_new = 0.0, rr_new = 0.0;
    {
      const int kq = lane / 6, comp = lane % 6;
      for (int c5 = gwarp * 5; c5 < C; c5 += gwarps * 5) {
        const int c = c5 + kq;
        const bool on = lane < 30 && c < C;
        const int i = 6 * c + comp;
        double rk = 0.0;
        if (on) {
          const double pk = u[i] + beta * a.p[i];
          const double sk = w[i] + beta * s[i];
          a.p[i] = pk; s[i] = sk;
          a.x[i] += alpha * pk;
          rk = a.r[i] - alpha * sk;
          a.r[i] = rk;
          rr_new += rk * rk;
        }
        double t = 0.0;
        const double* mrow = a.Minv + (int64_t)(on ? c : 0) * 36 + comp * 6;
#pragma unroll
        for (int k = 0; k < 6; ++k) {
          const double rj = __shfl_sync(FULL, rk, 6 * kq + k);
          t += mrow[k] * rj;
        }
        if (on) { u[i] = t; gam_new += t * rk; }
      }
    }
    ++it;
    grid.sync();                                   // u complete before the mat-vec reads it
    const double del_part = spmv_rows(a, u, w, gwarp, gwarps, lane);
    grid_sum3(grid, a.part, phase, sm, gam_new, del_part, rr_new, d0, d1, d2);
    gam_new = d0; del = d1; rr = d2;
    if (!(rr > thresh)) break;                     // converged: x already holds the iterate that r belongs to
    beta = gam_new / gam;
    const double denom = del - beta * gam_new / alpha;
    if (!(denom > 0.0)) { fail = 1; break; }       // not positive definite along the search direction
    alpha = gam_new / denom;
    gam = gam_new;
  }
  // ---- dc . (b - S dc) for the predicted reduction
  double dr = 0.0;
  for (int i = gtid; i < n; i += gthreads) dr += a.x[i] * a.r[i];
  grid_sum3(grid, a.part, phase, sm, dr, 0.0, 0.0, d0, d1, d2);
  if (gtid == 0) {
    a.scal[S_DC_RES] = d0;
    a.scal[S_PCG_ITERS] = (double)it;
    a.scal[S_SOLVE_FAIL] = (double)fail;
  }
}

// ------------------------------------------------------------------ dense Cholesky (column-major lower)
constexpr int CH_T = 1024;
__global__ void __launch_bounds__(CH_T) k_cholesky(int n, double* __restrict__ A, const double* __restrict__ b,
                                                  double* __restrict__ x, double* __restrict__ scal) {
  extern __shared__ double y[];  // [n]
  __shared__ int s_fail;
  const int tid = threadIdx.x;
  if (tid == 0) s_fail = 0;
  __syncthreads();
  for (int j = 0; j < n; ++j) {
    if (tid == 0) {
      const double d = A[(int64_t)j * n + j];
      if (!(d > 0.0) || !(d < 1e300)) s_fail = 1;
      A[(int64_t)j * n + j] = sqrt(d > 0.0 ? d : 1.0);
    }
    __syncthreads();
    if (s_fail) break;
    const double inv = 1.0 / A[(int64_t)j * n + j];
    double* colj = A + (int64_t)j * n;
    for (int i = j + 1 + tid; i < n; i += CH_T) colj[i] *= inv;
    __syncthreads();
    const int m = n - j - 1;
    for (int64_t idx = tid; idx < (int64_t)m * m; idx += CH_T) {
      const int i = j + 1 + (int)(idx % m), k = j + 1 + (int)(idx / m);
      if (i >= k) A[(int64_t)k * n + i] -= colj[i] * colj[k];
    }
    __syncthreads();
  }
  if (s_fail) {
    for (int i = tid; i < n; i += CH_T) x[i] = 0.0;
    if (tid == 0) { scal[S_SOLVE_FAIL] = 1.0; scal[S_DC_RES] = 0.0; scal[S_PCG_ITERS] = 0.0; }
    return;
  }
  for (int i = tid; i < n; i += CH_T) y[i] = b[i];
  __syncthreads();
  for (int j = 0; j < n; ++j) {  // L y = b
    if (tid == 0) y[j] /= A[(int64_t)j * n + j];
    __syncthreads();
    const double yj = y[j];
    for (int i = j + 1 + tid; i < n; i += CH_T) y[i] -= A[(int64_t)j * n + i] * yj;
    __syncthreads();
  }
  for (int j = n - 1; j >= 0; --j) {  // L^T x = y : x_j = (y_j - sum_{i>j} L_ij x_i) / L_jj
    double s = 0.0;
    for (int i = j + 1 + tid; i < n; i += CH_T) s += A[(int64_t)j * n + i] * y[i];
    s = warp_sum(s);
    __shared__ double red[32];
    if ((tid & 31) == 0) red[tid >> 5] = s;
    __syncthreads();
    if (tid == 0) {
      double t = 0.0;
      for (int w = 0; w < CH_T / 32; ++w) t += red[w];
      y[j] = (y[j] - t) / A[(int64_t)j * n + j];
    }
    __syncthreads();
  }
  for (int i = tid; i < n; i += CH_T) x[i] = y[i];
  if (tid == 0) { scal[S_SOLVE_FAIL] = 0.0; scal[S_DC_RES] = 0.0; scal[S_PCG_ITERS] = 0.0; }
}

// ------------------------------------------------------------------ blocked dense Cholesky, n <= 768 (C <= 128)
// One CTA; the current 32-column panel lives in shared memory (row stride 33: conflict-free column walks).  Per panel:
// factor in shared memory, forward-substitute the right-hand side with it, write L back, rank-32 update of the
// trailing matrix in global memory.  The backward substitution reloads the panels in reverse order.  Replaces the
// column-at-a-time kernel (every column step paid global-memory latency) for the small real problems (SceauxCastle:
// n <= 54) and config B (n = 294).
constexpr int CHB_NB = 32;
constexpr int CHB_LD = CHB_NB + 1;
constexpr int CHB_MAX_N = 768;
constexpr int CHB_T = 1024;
constexpr int CHB_SMEM = (CHB_MAX_N * CHB_LD + CHB_MAX_N) * 8;

__global__ void __launch_bounds__(CHB_T) k_cholesky_blocked(int n, double* __restrict__ A, const double* __restrict__ b,
                                                           double* __restrict__ x, double* __restrict__ scal) {
  extern __shared__ double chb[];
  double* P = chb;                         // [m][CHB_LD]
  double* y = chb + CHB_MAX_N * CHB_LD;    // [n]
  __shared__ int s_fail;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid == 0) s_fail = 0;
  for (int i = tid; i < n; i += CHB_T) y[i] = b[i];
  __syncthreads();
  for (int k0 = 0; k0 < n; k0 += CHB_NB) {
    const int nb = min(CHB_NB, n - k0), m = n - k0;
    for (int idx = tid; idx < m * nb; idx += CHB_T) {
      const int j = idx / m, i = idx - j * m;
      P[i * CHB_LD + j] = A[(int64_t)(k0 + j) * n + k0 + i];
    }
    __syncthreads();
    // ---- panel factorisation
    for (int j = 0; j < nb; ++j) {
      if (tid == 0) {
        const double d = P[j * CHB_LD + j];
        if (!(d > 0.0) || !(d < 1e300)) s_fail = 1;
        P[j * CHB_LD + j] = sqrt(d > 0.0 ? d : 1.0);
      }
      __syncthreads();
      if (s_fail) break;
      const double inv = 1.0 / P[j * CHB_LD + j];
      for (int i = j + 1 + tid; i < m; i += CHB_T) P[i * CHB_LD + j] *= inv;
      __syncthreads();
      const int mr = m - j - 1, nc = nb - j - 1;
      for (int idx = tid; idx < mr * nc; idx += CHB_T) {
        const int c = j + 1 + idx / mr, i = j + 1 + idx % mr;
        if (i >= c) P[i * CHB_LD + c] -= P[i * CHB_LD + j] * P[c * CHB_LD + j];
      }
      __syncthreads();
    }
    if (s_fail) break;
    // ---- forward substitution with this panel: triangle by warp 0 (lane per row), the rows below by everybody
    if (wid == 0) {
      double yi = lane < nb ? y[k0 + lane] : 0.0;
      const double invd = 1.0 / P[(lane < nb ? lane : 0) * CHB_LD + (lane < nb ? lane : 0)];
      for (int j = 0; j < nb; ++j) {
        double yj = yi * invd;
        yj = __shfl_sync(FULL, yj, j);
        if (lane == j) yi = yj;
        else if (lane > j && lane < nb) yi -= P[lane * CHB_LD + j] * yj;
      }
      if (lane < nb) y[k0 + lane] = yi;
    }
    __syncthreads();
    for (int i = nb + tid; i < m; i += CHB_T) {
      double t = y[k0 + i];
      for (int j = 0; j < nb; ++j) t -= P[i * CHB_LD + j] * y[k0 + j];
      y[k0 + i] = t;
    }
    // ---- write L back, rank-nb update of the trailing lower triangle
    for (int idx = tid; idx < m * nb; idx += CHB_T) {
      const int j = idx / m, i = idx - j * m;
      if (i >= j) A[(int64_t)(k0 + j) * n + k0 + i] = P[i * CHB_LD + j];
    }
    const int mt = m - nb;
    for (int64_t idx = tid; idx < (int64_t)mt * mt; idx += CHB_T) {
      const int c = nb + (int)(idx / mt), i = nb + (int)(idx % mt);
      if (i < c) continue;
      double t = 0.0;
#pragma unroll 8
      for (int j = 0; j < CHB_NB; ++j) t += (j < nb) ? P[i * CHB_LD + j] * P[c * CHB_LD + j] : 0.0;
      A[(int64_t)(k0 + c) * n + k0 + i] -= t;
    }
    __syncthreads();
  }
  if (s_fail) {
    for (int i = tid; i < n; i += CHB_T) x[i] = 0.0;
    if (tid == 0) { scal[S_SOLVE_FAIL] = 1.0; scal[S_DC_RES] = 0.0; scal[S_PCG_ITERS] = 0.0; }
    return;
  }
  // ---- backward substitution L^T x = y, panels in reverse order (y becomes x in place)
  const int last = ((n - 1) / CHB_NB) * CHB_NB;
  for (int k0 = last; k0 >= 0; k0 -= CHB_NB) {
    const int nb = min(CHB_NB, n - k0), m = n - k0;
    __syncthreads();
    for (int idx = tid; idx < m * nb; idx += CHB_T) {
      const int j = idx / m, i = idx - j * m;
      P[i * CHB_LD + j] = (i >= j) ? A[(int64_t)(k0 + j) * n + k0 + i] : 0.0;
    }
    __syncthreads();
    // t_j = y_j - sum_{i >= nb} L[i][j] x_i : warp w handles columns w, w + 32, ... (nb <= 32: one each)
    if (wid < nb) {
      const int j = wid;
      double t = 0.0;
      for (int i = nb + lane; i < m; i += 32) t += P[i * CHB_LD + j] * y[k0 + i];
      t = warp_sum(t);
      if (lane == 0) y[k0 + j] -= t;
    }
    __syncthreads();
    if (wid == 0) {
      double xi = lane < nb ? y[k0 + lane] : 0.0;
      const double invd = 1.0 / P[(lane < nb ? lane : 0) * CHB_LD + (lane < nb ? lane : 0)];
      for (int j = nb - 1; j >= 0; --j) {
        double xj = xi * invd;
        xj = __shfl_sync(FULL, xj, j);
        if (lane == j) xi = xj;
        else if (lane < j) xi -= P[j * CHB_LD + lane] * xj;   // L^T[lane][j] = L[j][lane]
      }
      if (lane < nb) y[k0 + lane] = xi;
    }
  }
  __syncthreads();
  for (int i = tid; i < n; i += CHB_T) x[i] = y[i];
  if (tid == 0) { scal[S_SOLVE_FAIL] = 0.0; scal[S_DC_RES] = 0.0; scal[S_PCG_ITERS] = 0.0; }
}

#include "k_twolevel.cuh"
#include "k_chol_mc.cuh"

}  // namespace

int pcg_grid(Handle* h) {
  // queried once per device (function attributes and occupancy are per device/context: a process may hold handles on
  // several GPUs); the occupancy call sits on the path of every camera solve otherwise
  static int per_sm_dev[MAX_DEVICES] = {};
  int& per_sm = per_sm_dev[h->device % MAX_DEVICES];
  if (per_sm == 0 &&
      (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_pcg, PCG_T, 0) != cudaSuccess || per_sm < 1)) per_sm = 1;
  const int64_t C = h->pb.C;
  int64_t need = std::max<int64_t>((C + (PCG_T / 32) - 1) / (PCG_T / 32), 1);
  return (int)std::min<int64_t>(need, (int64_t)per_sm * h->sm_count);
}

int launch_pcg(Handle* h, double rtol, int max_iter) {
  Problem& pb = h->pb;
  Work& w = h->w;
  const int64_t n = 6 * pb.C;
  PcgArgs a;
  a.C = (int)pb.C;
  a.rowptr = pb.bsr_rowptr; a.col = pb.bsr_col; a.src = pb.bsr_src; a.diag_blk = pb.diag_blk;
  a.Sup = w.Sup; a.b = w.b;
  a.x = w.dc; a.r = w.pcg; a.z = w.pcg + n; a.p = w.pcg + 2 * n; a.Ap = w.pcg + 3 * n; a.s = w.pcg + 4 * n;
  a.Minv = w.pcg + 5 * n;
  a.part = w.pcg_part; a.scal = w.scal; a.rtol = rtol; a.max_iter = max_iter;
  void* args[] = {&a};
  const int grid = pcg_grid(h);
  BA_CUDA(h, cudaLaunchCooperativeKernel((void*)k_pcg, dim3(grid), dim3(PCG_T), args, 0, h->stream));
  h->st.launches += 1;
  return 0;
}

// ------------------------------------------------------------------ two-level preconditioner (k_twolevel.cuh)
static TlArgs tl_args(Handle* h) {
  Problem& pb = h->pb;
  Work& w = h->w;
  TlArgs a;
  a.C = (int)pb.C; a.n_agg = pb.n_agg; a.nc = pb.nc; a.ncp = pb.ncp; a.Cs = pb.Cs;
  a.cam_agg = pb.cam_agg; a.agg_off = pb.agg_off; a.agg_cams = pb.agg_cams;
  a.sp_rowptr = pb.sp_rowptr; a.sp_col = pb.sp_col; a.sp_of = pb.sp_of;
  a.rowptr = pb.bsr_rowptr; a.col = pb.bsr_col; a.src = pb.bsr_src;
  a.Sup = w.Sup; a.camprep = w.camprep; a.fixed = pb.cam_fixed;
  a.G = w.tl_G; a.SP = w.tl_SP; a.A = w.tl_A; a.cb = w.tl_cb; a.rb = w.tl_rb; a.dsc = w.tl_dsc; a.cen = w.tl_cen;
  a.scal = w.scal;
  return a;
}

int launch_tl_build(Handle* h, bool rebuild_coarse) {
  Problem& pb = h->pb;
  if (!pb.tl) return fail(h, B200BA_BAD_ARG, "two-level preconditioner is not set up for this problem");
  TlArgs a = tl_args(h);
  cudaStream_t st = h->stream;
  if (rebuild_coarse) {
    k_tl_gauge<<<pb.n_agg, 64, 0, st>>>(a);
    h->st.launches += 1;
  }
  k_tl_sp<<<(int)((pb.C * 32 + 255) / 256), 256, 0, st>>>(a);
  h->st.launches += 1;
  if (rebuild_coarse) {
    const size_t smem = (size_t)TL_K * pb.ncp * sizeof(double);
    static bool attr_dev[MAX_DEVICES] = {};
    bool& attr = attr_dev[h->device % MAX_DEVICES];
    if (!attr) { cudaFuncSetAttribute(k_tl_assemble, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024); attr = true; }
    const double one = 1.0;
    BA_CUDA(h, cudaMemcpyAsync(h->w.scal + S_TL_OK, &one, sizeof(double), cudaMemcpyHostToDevice, st));
    k_tl_assemble<<<pb.n_agg + 1, 256, smem, st>>>(a);
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_tl_gj, GJ_T, 0) != cudaSuccess || per_sm < 1) per_sm = 1;
    const int grid = std::min(per_sm, 2) * h->sm_count;
    void* args[] = {&a};
    BA_CUDA(h, cudaLaunchCooperativeKernel((void*)k_tl_gj, dim3(grid), dim3(GJ_T), args, 0, st));
    h->st.launches += 2;
  }
  BA_CUDA(h, cudaGetLastError());
  return 0;
}

int launch_pcg2(Handle* h, double rtol, int max_iter) {
  Problem& pb = h->pb;
  Work& w = h->w;
  const int64_t n = 6 * pb.C;
  Pcg2Args g;
  PcgArgs& a = g.p;
  a.C = (int)pb.C;
  a.rowptr = pb.bsr_rowptr; a.col = pb.bsr_col; a.src = pb.bsr_src; a.diag_blk = pb.diag_blk;
  a.Sup = w.Sup; a.b = w.b;
  a.x = w.dc; a.r = w.pcg; a.z = w.pcg + n; a.p = w.pcg + 2 * n; a.Ap = w.pcg + 3 * n; a.s = w.pcg + 4 * n;
  a.Minv = w.pcg + 5 * n;
  a.part = w.pcg_part; a.scal = w.scal; a.rtol = rtol; a.max_iter = max_iter;
  g.n_agg = pb.n_agg; g.nc = pb.nc; g.ncp = pb.ncp;
  g.cam_agg = pb.cam_agg; g.agg_off = pb.agg_off; g.agg_cams = pb.agg_cams; g.sp_rowptr = pb.sp_rowptr; g.sp_col = pb.sp_col;
  g.G = w.tl_G; g.SP = w.tl_SP; g.Ainv = w.tl_A; g.rc = w.tl_rc; g.zc = w.tl_zc; g.uf = w.tl_uf; g.prof = w.tl_prof;
  void* args[] = {&g};
  static int per_sm_dev[MAX_DEVICES] = {};
  int& per_sm = per_sm_dev[h->device % MAX_DEVICES];
  if (per_sm == 0 &&
      (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_pcg2, PCG_T, 0) != cudaSuccess || per_sm < 1)) per_sm = 1;
  const int64_t need = std::max<int64_t>((pb.C + (PCG_T / 32) - 1) / (PCG_T / 32), 1);
  const int grid = (int)std::min<int64_t>(std::min<int64_t>(need, (int64_t)per_sm * h->sm_count), pcg_grid(h));
  BA_CUDA(h, cudaLaunchCooperativeKernel((void*)k_pcg2, dim3(grid), dim3(PCG_T), args, 0, h->stream));
  h->st.launches += 1;
  return 0;
}

// profiling only: CUDA events directly around one launch on the library stream
struct KernelEvents {
  Handle* h; double* acc; bool on;
  KernelEvents(Handle* h_, double* acc_) : h(h_), acc(acc_), on(h_->profile) { if (on) cudaEventRecord(h->ev2, h->stream); }
  void stop() {
    if (!on) return;
    cudaEventRecord(h->ev3, h->stream);
    cudaEventSynchronize(h->ev3);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, h->ev2, h->ev3);
    *acc += ms;
  }
};

int launch_cholesky(Handle* h) {
  Work& w = h->w;
  const int n = (int)(6 * h->pb.C);
  if (h->pb.C > DENSE_MAX_CAM) return fail(h, B200BA_BAD_ARG, "dense Cholesky needs n_cam <= " + std::to_string(DENSE_MAX_CAM));
  if (!w.dense) {   // allocated on first use above the always-resident size (36 C^2 doubles)
    w.dense = (double*)dalloc(h, (size_t)36 * h->pb.C * h->pb.C * sizeof(double));
    if (!w.dense) return B200BA_OOM;
  }
  launch_densify(h, w.dense);
  const char* e_mc = getenv("B200BA_CHOL_MC");   // 1: the multi-CTA DMMA factorisation at every size (tests), 0: never
  const bool want_mc = e_mc ? atoi(e_mc) != 0 : n > 192;   // measured per LM iteration (tools/exp_small_solvers.py): n = 144 single-CTA
                                                         // 0.29 ms vs 0.35; n = 294 (config B) 0.77 vs 0.51; n = 384 1.23 vs 0.57; n = 768 6 vs 1.2
  if (!want_mc && n <= CHB_MAX_N) {
    static bool attr_dev[MAX_DEVICES] = {};
    bool& attr = attr_dev[h->device % MAX_DEVICES];
    if (!attr) { cudaFuncSetAttribute(k_cholesky_blocked, cudaFuncAttributeMaxDynamicSharedMemorySize, CHB_SMEM); attr = true; }
    k_cholesky_blocked<<<1, CHB_T, CHB_SMEM, h->stream>>>(n, w.dense, w.b, w.dc, w.scal);
    h->st.launches += 1;
  } else if (!want_mc) {
    k_cholesky<<<1, CH_T, n * sizeof(double), h->stream>>>(n, w.dense, w.b, w.dc, w.scal);
    h->st.launches += 1;
  } else {
    const int nb = (n + CM_NB - 1) / CM_NB, np = nb * CM_NB;
    if (!w.chol || w.chol_np < np) {
      w.chol = (double*)dalloc(h, ((size_t)np * np + 2 * (size_t)np + (size_t)nb * CM_NB * CM_NB + 8) * sizeof(double));
      if (!w.chol) return B200BA_OOM;
      w.chol_np = np;
    }
    CholMcArgs a;
    a.n = n; a.np = np; a.nb = nb;
    a.A = w.chol;
    a.b = a.A + (size_t)np * np;
    a.x = a.b + np;
    a.Linv = a.x + np;
    a.fail = reinterpret_cast<int*>(a.Linv + (size_t)nb * CM_NB * CM_NB);
    a.scal = w.scal;
    static int per_sm_dev[MAX_DEVICES] = {};
    int& per_sm = per_sm_dev[h->device % MAX_DEVICES];
    if (per_sm == 0) {
      cudaFuncSetAttribute(k_chol_mc, cudaFuncAttributeMaxDynamicSharedMemorySize, CM_SMEM);
      if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_chol_mc, CM_T, CM_SMEM) != cudaSuccess || per_sm < 1) per_sm = 1;
    }
    k_chol_mc_pad<<<(unsigned)(((int64_t)np * np + 255) / 256), 256, 0, h->stream>>>(n, np, w.dense, w.b, a.A, a.b);
    const int grid = std::min(per_sm, 2) * h->sm_count;
    void* args[] = {&a};
    KernelEvents ev(h, &h->st.ms_chol_mc);
    BA_CUDA(h, cudaLaunchCooperativeKernel((void*)k_chol_mc, dim3(grid), dim3(CM_T), args, CM_SMEM, h->stream));
    ev.stop();
    BA_CUDA(h, cudaMemcpyAsync(w.dc, a.x, (size_t)n * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
    h->st.launches += 2;
  }
  BA_CUDA(h, cudaGetLastError());
  return 0;
}

}  // namespace b200ba
