// Multi-CTA blocked dense Cholesky of the reduced camera system on the FP64 tensor cores (included inside ba_solve.cu's
// anonymous namespace).  north_star: "tensor cores (FP64 DMMA) are used only for the dense reduced-camera Cholesky once
// camera count makes it a real dense contraction".  Replaces scipy's LSMR step of sfm.py:266-268 for 128 < C <= 1024
// when solver = cholesky is asked for (the default at that size stays the two-level PCG: S is block-sparse there, see
// DESIGN.md section 5 for the measured comparison).
//
// Right-looking, 64 x 64 blocks, one cooperative kernel, matrix full row-major with leading dimension np (n rounded up
// to 64, the padding carries an identity).  Per block column k:
//   P1  one CTA: L_kk = chol(A_kk) in shared memory, Linv_k = L_kk^-1 (kept for the substitutions), y_k = Linv_k b_k
//   P2  block rows i > k:  A_ik <- A_ik Linv_k^T  (64^3 DMMA tile),  b_i -= A_ik y_k
//   P3  tiles k < j <= i:  A_ij -= A_ik A_jk^T    (64^3 DMMA tile; the dense contraction, ~all of the flops)
// then the backward substitution x_k = Linv_k^T (y_k - sum_{i>k} L_ik^T x_i), right-looking as well.
// DMMA tile: 8 warps, each 16 x 32 of the 64 x 64 output = 2 x 4 m8n8k4 fragments (16 accumulators per lane); operands
// staged in shared memory with leading dimension 68 doubles, which makes the fragment loads (lane -> row lane / 4,
// column k0 + lane % 4) bank-conflict free for both operands of C = A B^T.

constexpr int CM_NB = 64;
constexpr int CM_LD = 68;
constexpr int CM_T = 256;
constexpr int CM_SMEM = 2 * CM_NB * CM_LD * 8;   // two operand tiles

struct CholMcArgs {
  int n, np, nb;
  double* A;       // [np][np]
  double* b;       // [np] right-hand side, overwritten by y
  double* x;       // [np] solution
  double* Linv;    // [nb][64][64] inverses of the diagonal factors
  double* scal;
  int* fail;       // device flag
};

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

// 64 x 64 row-major tile (leading dimension ld) -> shared memory tile with leading dimension CM_LD
__device__ __forceinline__ void cm_load_tile(double* __restrict__ s, const double* __restrict__ g, int64_t ld) {
  for (int i = threadIdx.x; i < CM_NB * CM_NB / 2; i += CM_T) {
    const int r = i / (CM_NB / 2), c2 = i % (CM_NB / 2);
    const double2 v = *reinterpret_cast<const double2*>(g + (int64_t)r * ld + 2 * c2);
    *reinterpret_cast<double2*>(s + r * CM_LD + 2 * c2) = v;
  }
}

// acc (2 x 4 fragments of the warp's 16 x 32 patch) += sA[rows] * sB[rows]^T over K = 64
__device__ __forceinline__ void cm_tile_abt(const double* __restrict__ sA, const double* __restrict__ sB, double (&acc)[2][4][2]) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int r0 = 16 * (w >> 1), c0 = 32 * (w & 1);
  const double* pa = sA + (r0 + (lane >> 2)) * CM_LD + (lane & 3);
  const double* pb = sB + (c0 + (lane >> 2)) * CM_LD + (lane & 3);
#pragma unroll 4
  for (int k0 = 0; k0 < CM_NB; k0 += 4) {
    const double a0 = pa[k0], a1 = pa[8 * CM_LD + k0];
    double bv[4];
#pragma unroll
    for (int t = 0; t < 4; ++t) bv[t] = pb[8 * t * CM_LD + k0];
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      dmma884(acc[0][t][0], acc[0][t][1], a0, bv[t]);
      dmma884(acc[1][t][0], acc[1][t][1], a1, bv[t]);
    }
  }
}

// the warp's accumulators <-> a 64 x 64 global tile: lane holds C[r0 + 8 m + lane / 4][c0 + 8 t + 2 (lane % 4) + {0, 1}]
template <class F>
__device__ __forceinline__ void cm_for_each_acc(double* __restrict__ g, int64_t ld, F f) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int r0 = 16 * (w >> 1) + (lane >> 2), c0 = 32 * (w & 1) + 2 * (lane & 3);
#pragma unroll
  for (int m = 0; m < 2; ++m)
#pragma unroll
    for (int t = 0; t < 4; ++t) f(reinterpret_cast<double2*>(g + (int64_t)(r0 + 8 * m) * ld + c0 + 8 * t), m, t);
}

// P1: factor the 64 x 64 diagonal block in shared memory (T, L: two tiles of leading dimension CM_LD), invert L.
// On return (after a __syncthreads) L holds L_kk (lower, zeros above) and T holds Linv_k (lower, zeros above).
// This is the serial part of every block column (the other CTAs wait at the grid-wide sync), so it is written for
// latency: a 16 x 16 thread grid walks the trailing block without integer divisions, one barrier per column; the
// inverse uses four threads per column (interleaved partial sums, two shuffles per row).
__device__ __forceinline__ void cm_factor_diag(double* __restrict__ T, double* __restrict__ L, int* __restrict__ fail) {
  const int tid = threadIdx.x, ti = tid >> 4, tj = tid & 15;
  for (int i = tid; i < CM_NB * CM_LD; i += CM_T) L[i] = 0.0;
  __syncthreads();
  for (int j = 0; j < CM_NB; ++j) {
    double d = T[j * CM_LD + j];
    if (!(d > 0.0) || !(d < 1e300)) { if (tid == 0) *fail = 1; d = 1.0; }
    const double inv = rsqrt(d), inv2 = inv * inv;
    // column j of L (rows i >= j); the trailing update reads the unscaled column from T, which step j never writes
    if (tid < CM_NB && tid >= j) L[tid * CM_LD + j] = T[tid * CM_LD + j] * inv;
    const int lo = j + 1;
    for (int i = lo + ti; i < CM_NB; i += 16) {
      const double f = -T[i * CM_LD + j] * inv2;
      for (int c = lo + tj; c <= i; c += 16) T[i * CM_LD + c] = fma(f, T[c * CM_LD + j], T[i * CM_LD + c]);
    }
    __syncthreads();
  }
  // Linv into T, row by row: X[j][c] = (delta_jc - sum_{c <= t < j} L[j][t] X[t][c]) / L[j][j]; the four threads
  // 4 c .. 4 c + 3 (one warp) share column c, thread q sums the terms t = c + q, c + q + 4, ...
  for (int i = tid; i < CM_NB * CM_LD; i += CM_T) T[i] = 0.0;
  __syncthreads();
  {
    const int c = tid >> 2, q = tid & 3;
    for (int j = (tid >> 5) * 8; j < CM_NB; ++j) {   // the same trip count in every lane of the warp (full-mask shuffles)
      double s = 0.0;
      for (int t = c + q; t < j; t += 4) s = fma(-L[j * CM_LD + t], T[t * CM_LD + c], s);
      s += __shfl_xor_sync(0xffffffffu, s, 1);
      s += __shfl_xor_sync(0xffffffffu, s, 2);
      if (q == 0 && j >= c) T[j * CM_LD + c] = ((j == c ? 1.0 : 0.0) + s) / L[j * CM_LD + j];
      __syncwarp();
    }
  }
  __syncthreads();
}

__global__ void __launch_bounds__(CM_T, 2) k_chol_mc(CholMcArgs a) {
  cg::grid_group grid = cg::this_grid();
  extern __shared__ __align__(16) double cms[];
  double* sA = cms;
  double* sB = cms + CM_NB * CM_LD;
  __shared__ double sy[CM_NB];
  __shared__ double sred[4][CM_NB];
  const int tid = threadIdx.x, nb = a.nb;
  const int64_t np = a.np;
  if (blockIdx.x == 0 && tid == 0) *a.fail = 0;
  grid.sync();
  for (int k = 0; k < nb; ++k) {
    double* Akk = a.A + ((int64_t)k * CM_NB) * np + (int64_t)k * CM_NB;
    double* Li = a.Linv + (int64_t)k * CM_NB * CM_NB;
    // ---- P1 (one CTA; rotating so the same SM is not always the serial one)
    if ((int)blockIdx.x == k % (int)gridDim.x) {
      cm_load_tile(sA, Akk, np);
      __syncthreads();
      cm_factor_diag(sA, sB, a.fail);
      for (int i = tid; i < CM_NB * CM_NB; i += CM_T) {
        const int r = i / CM_NB, c = i % CM_NB;
        Akk[(int64_t)r * np + c] = sB[r * CM_LD + c];     // L_kk (zeros above the diagonal)
        Li[i] = sA[r * CM_LD + c];                        // Linv_k
      }
      if (tid < CM_NB) {
        double s = 0.0;
        for (int t = 0; t <= tid; ++t) s = fma(sA[tid * CM_LD + t], a.b[(int64_t)k * CM_NB + t], s);
        sy[tid] = s;
      }
      __syncthreads();
      if (tid < CM_NB) a.b[(int64_t)k * CM_NB + tid] = sy[tid];
    }
    grid.sync();
    // ---- P2: panel below the diagonal block
    for (int i = k + 1 + (int)blockIdx.x; i < nb; i += (int)gridDim.x) {
      double* Aik = a.A + ((int64_t)i * CM_NB) * np + (int64_t)k * CM_NB;
      __syncthreads();
      cm_load_tile(sA, Aik, np);
      cm_load_tile(sB, Li, CM_NB);
      if (tid < CM_NB) sy[tid] = a.b[(int64_t)k * CM_NB + tid];
      __syncthreads();
      double acc[2][4][2];
#pragma unroll
      for (int m = 0; m < 2; ++m)
#pragma unroll
        for (int t = 0; t < 4; ++t) acc[m][t][0] = acc[m][t][1] = 0.0;
      cm_tile_abt(sA, sB, acc);
      __syncthreads();
      cm_for_each_acc(Aik, np, [&](double2* p, int m, int t) { *p = make_double2(acc[m][t][0], acc[m][t][1]); });
      cm_for_each_acc(sA, CM_LD, [&](double2* p, int m, int t) { *p = make_double2(acc[m][t][0], acc[m][t][1]); });
      __syncthreads();
      {   // b_i -= A_ik y_k : 4 threads per row, 16 columns each
        const int r = tid & 63, part = tid >> 6;
        double s = 0.0;
#pragma unroll 4
        for (int t = 16 * part; t < 16 * part + 16; ++t) s = fma(sA[r * CM_LD + t], sy[t], s);
        sred[part][r] = s;
      }
      __syncthreads();
      if (tid < CM_NB) a.b[(int64_t)i * CM_NB + tid] -= (sred[0][tid] + sred[1][tid]) + (sred[2][tid] + sred[3][tid]);
    }
    grid.sync();
    // ---- P3: trailing update, lower-triangle tiles (i, j), k < j <= i
    const int m = nb - k - 1;
    const int64_t ntile = (int64_t)m * (m + 1) / 2;
    for (int64_t tix = blockIdx.x; tix < ntile; tix += gridDim.x) {
      int ii = (int)((sqrt(8.0 * (double)tix + 1.0) - 1.0) * 0.5);
      while ((int64_t)(ii + 1) * (ii + 2) / 2 <= tix) ++ii;
      while ((int64_t)ii * (ii + 1) / 2 > tix) --ii;
      const int jj = (int)(tix - (int64_t)ii * (ii + 1) / 2);
      const int i = k + 1 + ii, j = k + 1 + jj;
      __syncthreads();
      cm_load_tile(sA, a.A + ((int64_t)i * CM_NB) * np + (int64_t)k * CM_NB, np);
      cm_load_tile(sB, a.A + ((int64_t)j * CM_NB) * np + (int64_t)k * CM_NB, np);
      __syncthreads();
      double acc[2][4][2];
#pragma unroll
      for (int mm = 0; mm < 2; ++mm)
#pragma unroll
        for (int t = 0; t < 4; ++t) acc[mm][t][0] = acc[mm][t][1] = 0.0;
      cm_tile_abt(sA, sB, acc);
      cm_for_each_acc(a.A + ((int64_t)i * CM_NB) * np + (int64_t)j * CM_NB, np, [&](double2* p, int mm, int t) {
        double2 v = *p;
        v.x -= acc[mm][t][0];
        v.y -= acc[mm][t][1];
        *p = v;
      });
    }
    grid.sync();
  }
  // ---- backward substitution, right-looking: x_k = Linv_k^T y_k, then y_j -= L_kj^T x_k for j < k.  The CTA that
  //      finishes y_{k-1} computes x_{k-1} at once, so one grid-wide sync per block step is enough.
  auto solve_block = [&](int k) {   // whole CTA; x_k from y_k (in a.b)
    const double* Li = a.Linv + (int64_t)k * CM_NB * CM_NB;
    if (tid < CM_NB) sy[tid] = a.b[(int64_t)k * CM_NB + tid];
    __syncthreads();
    {
      const int c = tid & 63, part = tid >> 6;
      double s = 0.0;
      for (int r = 16 * part; r < 16 * part + 16; ++r) s = fma(Li[r * CM_NB + c], sy[r], s);   // (Linv^T y)_c
      sred[part][c] = s;
    }
    __syncthreads();
    if (tid < CM_NB) a.x[(int64_t)k * CM_NB + tid] = (sred[0][tid] + sred[1][tid]) + (sred[2][tid] + sred[3][tid]);
    __syncthreads();
  };
  if (blockIdx.x == 0) solve_block(nb - 1);
  grid.sync();
  for (int k = nb - 1; k > 0; --k) {
    for (int j = (int)blockIdx.x; j < k; j += (int)gridDim.x) {
      const double* Lkj = a.A + ((int64_t)k * CM_NB) * np + (int64_t)j * CM_NB;
      __syncthreads();
      if (tid < CM_NB) sy[tid] = a.x[(int64_t)k * CM_NB + tid];
      __syncthreads();
      {
        const int c = tid & 63, part = tid >> 6;
        double s = 0.0;
        for (int r = 16 * part; r < 16 * part + 16; ++r) s = fma(Lkj[(int64_t)r * np + c], sy[r], s);
        sred[part][c] = s;
      }
      __syncthreads();
      if (tid < CM_NB) a.b[(int64_t)j * CM_NB + tid] -= (sred[0][tid] + sred[1][tid]) + (sred[2][tid] + sred[3][tid]);
      __syncthreads();
      if (j == k - 1) { __threadfence(); solve_block(k - 1); }
    }
    grid.sync();
  }
  if (blockIdx.x == 0 && tid == 0) {
    a.scal[S_SOLVE_FAIL] = *a.fail ? 1.0 : 0.0;
    a.scal[S_DC_RES] = 0.0;
    a.scal[S_PCG_ITERS] = 0.0;
  }
}

// dense[n][n] (both triangles) -> padded [np][np] with an identity in the padding, b -> padded right-hand side
__global__ void k_chol_mc_pad(int n, int np, const double* __restrict__ dense, const double* __restrict__ b,
                              double* __restrict__ A, double* __restrict__ bp) {
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t < (int64_t)np * np) {
    const int r = (int)(t / np), c = (int)(t % np);
    A[t] = (r < n && c < n) ? dense[(int64_t)r * n + c] : (r == c ? 1.0 : 0.0);
  }
  if (t < np) bp[t] = t < n ? b[t] : 0.0;
}
