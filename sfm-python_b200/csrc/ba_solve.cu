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
  double* x; double* r; double* z; double* p; double* Ap; double* Minv;
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

__global__ void __launch_bounds__(PCG_T) k_pcg(PcgArgs a) {
  cg::grid_group grid = cg::this_grid();
  __shared__ double sm[128];
  const int C = a.C, n = 6 * C;
  const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
  const int gthreads = gridDim.x * blockDim.x;
  const int lane = threadIdx.x & 31;
  const int gwarp = gtid >> 5, gwarps = gthreads >> 5;

  // ---- init: block-Jacobi inverse, x = 0, r = b, z = Minv r, p = z
  double rz = 0.0, bb = 0.0;
  int phase = 0;
  for (int c = gtid; c < C; c += gthreads) {
    inv6(a.Sup + (int64_t)a.diag_blk[c] * 36, a.Minv + (int64_t)c * 36);
    double rv[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) { rv[k] = a.b[6 * c + k]; a.x[6 * c + k] = 0.0; a.r[6 * c + k] = rv[k]; bb += rv[k] * rv[k]; }
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      double s = 0.0;
#pragma unroll
      for (int k = 0; k < 6; ++k) s += a.Minv[(int64_t)c * 36 + i * 6 + k] * rv[k];
      a.z[6 * c + i] = s; a.p[6 * c + i] = s; rz += s * rv[i];
    }
  }
  double d0, d1, d2;
  grid_sum3(grid, a.part, phase, sm, rz, bb, 0.0, d0, d1, d2);
  rz = d0; bb = d1;
  double rr = bb;
  int it = 0;
  int fail = 0;
  const double thresh = a.rtol * a.rtol * bb;
  while (it < a.max_iter && rr > thresh && bb > 0.0) {
    // ---- Ap = S p  (warp per block row; 5 blocks x 6 rows per step on lanes 0..29)
    double pAp = 0.0;
    for (int c = gwarp; c < C; c += gwarps) {
      const int q0 = a.rowptr[c], q1 = a.rowptr[c + 1];
      const int kq = lane / 6, r = lane % 6;
      double y = 0.0;
      if (lane < 30) {
        for (int q = q0 + kq; q < q1; q += 5) {
          const int s = a.src[q];
          const double* B = a.Sup + (int64_t)(s & 0x7fffffff) * 36;
          const double* pv = a.p + 6 * a.col[q];
          if (s < 0) {
#pragma unroll
            for (int j = 0; j < 6; ++j) y += B[j * 6 + r] * pv[j];
          } else {
#pragma unroll
            for (int j = 0; j < 6; ++j) y += B[r * 6 + j] * pv[j];
          }
        }
      }
      double tot = 0.0;
#pragma unroll
      for (int g = 0; g < 5; ++g) tot += __shfl_sync(FULL, y, (lane % 6) + 6 * g);
      if (lane < 6) { a.Ap[6 * c + lane] = tot; pAp += tot * a.p[6 * c + lane]; }
    }
    grid_sum3(grid, a.part, phase, sm, pAp, 0.0, 0.0, d0, d1, d2);
    pAp = d0;
    if (!(pAp > 0.0)) { fail = 1; break; }  // not positive definite along p (uniform across the grid)
    const double alpha = rz / pAp;
    // ---- x += alpha p, r -= alpha Ap, z = Minv r   (thread per camera)
    double rz_new = 0.0, rr_new = 0.0;
    for (int c = gtid; c < C; c += gthreads) {
      double rv[6];
#pragma unroll
      for (int k = 0; k < 6; ++k) {
        a.x[6 * c + k] += alpha * a.p[6 * c + k];
        rv[k] = a.r[6 * c + k] - alpha * a.Ap[6 * c + k];
        a.r[6 * c + k] = rv[k];
        rr_new += rv[k] * rv[k];
      }
#pragma unroll
      for (int i = 0; i < 6; ++i) {
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < 6; ++k) s += a.Minv[(int64_t)c * 36 + i * 6 + k] * rv[k];
        a.z[6 * c + i] = s; rz_new += s * rv[i];
      }
    }
    grid_sum3(grid, a.part, phase, sm, rz_new, rr_new, 0.0, d0, d1, d2);
    rz_new = d0; rr = d1;
    ++it;
    const double beta = rz_new / rz;
    rz = rz_new;
    for (int i = gtid; i < n; i += gthreads) a.p[i] = a.z[i] + beta * a.p[i];
    grid.sync();
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

}  // namespace

int pcg_grid(Handle* h) {
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_pcg, PCG_T, 0) != cudaSuccess || per_sm < 1) per_sm = 1;
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
  a.x = w.dc; a.r = w.pcg; a.z = w.pcg + n; a.p = w.pcg + 2 * n; a.Ap = w.pcg + 3 * n; a.Minv = w.pcg + 4 * n;
  a.part = w.pcg_part; a.scal = w.scal; a.rtol = rtol; a.max_iter = max_iter;
  void* args[] = {&a};
  const int grid = pcg_grid(h);
  BA_CUDA(h, cudaLaunchCooperativeKernel((void*)k_pcg, dim3(grid), dim3(PCG_T), args, 0, h->stream));
  h->st.launches += 1;
  return 0;
}

int launch_cholesky(Handle* h) {
  Work& w = h->w;
  const int n = (int)(6 * h->pb.C);
  if (!w.dense) return fail(h, B200BA_BAD_ARG, "dense Cholesky needs n_cam <= 512");
  launch_densify(h, w.dense);
  k_cholesky<<<1, CH_T, n * sizeof(double), h->stream>>>(n, w.dense, w.b, w.dc, w.scal);
  h->st.launches += 1;
  BA_CUDA(h, cudaGetLastError());
  return 0;
}

}  // namespace b200ba
