// Camera-side accumulation and point-block Schur elimination (included inside ba_kernels.cu).
//
// "Atomics-free scatter into 6x6 camera blocks via sorted observation order": observations are visited
// through a camera-sorted permutation, a warp owns a chunk of one camera's run, reduces in registers +
// shuffles and writes one partial; partials are summed per camera in chunk order (bit-reproducible).
// The Schur complement uses the same scheme over pair entries (i,j) sorted by block (cam_i, cam_j).

__device__ __forceinline__ int tri6(int m, int n) { return m * 6 - (m * (m - 1)) / 2 + (n - m); }  // n >= m

__global__ void __launch_bounds__(256)
k_cam_side(Problem pb, const double* __restrict__ x, const double* __restrict__ camprep,
           const double* __restrict__ rr, double* __restrict__ cpart) {
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nwarp = (gridDim.x * blockDim.x) >> 5;
  const double* xp = x + 6 * pb.C;
  for (int ch = warp; ch < pb.n_cch; ch += nwarp) {
    const int c = pb.cch_cam[ch], a = pb.cch_a[ch], b = pb.cch_b[ch];
    double cp[CAMPREP_FIELDS];
    load_cam(camprep, pb.Cs, c, cp);
    double acc[33];
#pragma unroll
    for (int k = 0; k < 33; ++k) acc[k] = 0.0;
    for (int q = a + lane; q < b; q += 32) {
      const int i = pb.cperm[q];
      const int p = pb.pt[i];
      const double2 r = reinterpret_cast<const double2*>(rr)[2 * (int64_t)i];
      const double2 rt = reinterpret_cast<const double2*>(rr)[2 * (int64_t)i + 1];
      double d0, d1, Jc[12], Jp[6];
      project_jac(cp, pb.K, xp[3 * p], xp[3 * p + 1], xp[3 * p + 2], 0.0, 0.0, d0, d1, Jc, Jp);
      int k = 0;
#pragma unroll
      for (int m = 0; m < 6; ++m)
#pragma unroll
        for (int n = m; n < 6; ++n) acc[k++] += Jc[m] * Jc[n] + Jc[6 + m] * Jc[6 + n];
#pragma unroll
      for (int m = 0; m < 6; ++m) {
        acc[21 + m] += Jc[m] * r.x + Jc[6 + m] * r.y;
        acc[27 + m] += Jc[m] * rt.x + Jc[6 + m] * rt.y;
      }
    }
#pragma unroll
    for (int k = 0; k < 33; ++k) acc[k] = warp_sum(acc[k]);
    if (lane == 0) {
#pragma unroll
      for (int k = 0; k < 33; ++k) cpart[(int64_t)ch * 33 + k] = acc[k];
    }
  }
}

// U (21 upper-triangular entries), gc = J_c^T r, b = -J_c^T rt, udiag per camera
__global__ void k_cam_reduce(int C, const int* __restrict__ cam_choff, const double* __restrict__ cpart,
                             double* __restrict__ U, double* __restrict__ gc, double* __restrict__ b,
                             double* __restrict__ udiag) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= C * 33) return;
  const int c = t / 33, k = t % 33;
  double s = 0.0;
  for (int ch = cam_choff[c]; ch < cam_choff[c + 1]; ++ch) s += cpart[(int64_t)ch * 33 + k];
  if (k < 21) {
    U[c * 21 + k] = s;
    const int dm = (k == 0) ? 0 : (k == 6) ? 1 : (k == 11) ? 2 : (k == 15) ? 3 : (k == 18) ? 4 : (k == 20) ? 5 : -1;
    if (dm >= 0) udiag[6 * c + dm] = s;
  } else if (k < 27) {
    gc[6 * c + (k - 21)] = s;
  } else {
    b[6 * c + (k - 27)] = -s;
  }
}

// ---- Schur pair kernel: S_{c_i c_j} -= Y_i Y_j^T summed over the pair entries of a chunk (one block of S).
// A warp streams its chunks through a 3-stage cp.async pipeline: the two 144-byte records of 32 entries are
// gathered global -> shared memory by LDGSTS with 9 consecutive lanes per record (each request touches ~7 lines
// instead of 32), without passing through registers, while the previous group is multiplied out of shared memory.
// Entry slots are padded to 304 B so that the per-lane 16-byte reads are bank-conflict free.
constexpr int SCH_WARPS = 7;
constexpr int SCH_STAGES = 3;
constexpr int SCH_ESTRIDE = 304;
constexpr int SCH_STAGE_BYTES = 32 * SCH_ESTRIDE;
constexpr int SCH_SMEM = SCH_WARPS * SCH_STAGES * SCH_STAGE_BYTES;  // 204288 B

__device__ __forceinline__ void cp_async16(void* dst_smem, const void* src) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(dst_smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

struct ChunkCursor {
  int ch, g, a, b;  // chunk id, 32-entry group inside it, entry range of the chunk
};
__device__ __forceinline__ void cursor_load(const Problem& pb, ChunkCursor& c) {
  c.g = 0;
  if (c.ch < pb.n_sch) { c.a = pb.sch_a[c.ch]; c.b = pb.sch_b[c.ch]; } else { c.a = 0; c.b = 0; }
}
__device__ __forceinline__ bool cursor_advance(const Problem& pb, ChunkCursor& c, int stride) {  // true: chunk finished
  ++c.g;
  if (c.a + 32 * c.g >= c.b) { c.ch += stride; cursor_load(pb, c); return true; }
  return false;
}

template <int MODE>  // 0 = production; 1 = fetch only; 2 = multiply only (diagnostic builds of the same pipeline)
__global__ void __launch_bounds__(SCH_WARPS * 32)
k_schur(Problem pb, const double* __restrict__ Y, double* __restrict__ spart) {
  extern __shared__ __align__(16) unsigned char sch_smem[];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int nwarp = gridDim.x * SCH_WARPS;
  unsigned char* wbase = sch_smem + wid * (SCH_STAGES * SCH_STAGE_BYTES);

  ChunkCursor F, Cc;  // fetch / compute cursors over the same (chunk, group) sequence
  F.ch = blockIdx.x * SCH_WARPS + wid;
  cursor_load(pb, F);
  Cc = F;
  long long ij_pend = 0;
  if (F.ch < pb.n_sch) ij_pend = pb.ent[min(F.a + lane, F.b - 1)];

  auto fetch = [&](int stage) {
    if (F.ch < pb.n_sch) {
      const int i = (int)(ij_pend >> 32), j = (int)(ij_pend & 0xffffffffll);
      unsigned char* st = wbase + stage * SCH_STAGE_BYTES;
#pragma unroll
      for (int t = 0; t < (MODE == 2 ? 0 : 9); ++t) {
        const int id = t * 32 + lane;
        const int rec = id / 9, q = id - rec * 9;
        const int ri = __shfl_sync(FULL, i, rec), rj = __shfl_sync(FULL, j, rec);
        cp_async16(st + rec * SCH_ESTRIDE + q * 16, Y + 18 * (int64_t)ri + 2 * q);
        cp_async16(st + rec * SCH_ESTRIDE + 144 + q * 16, Y + 18 * (int64_t)rj + 2 * q);
      }
      cursor_advance(pb, F, nwarp);
      if (F.ch < pb.n_sch) ij_pend = pb.ent[min(F.a + 32 * F.g + lane, F.b - 1)];  // consumed by the NEXT fetch
    }
    cp_async_commit();
  };

#pragma unroll
  for (int s = 0; s < SCH_STAGES - 1; ++s) fetch(s);
  int stage_c = 0, stage_f = SCH_STAGES - 1;
  double acc[36];
#pragma unroll
  for (int k = 0; k < 36; ++k) acc[k] = 0.0;

  while (Cc.ch < pb.n_sch) {
    fetch(stage_f);
    cp_async_wait<SCH_STAGES - 1>();
    __syncwarp();
    if (MODE != 1 && Cc.a + 32 * Cc.g + lane < Cc.b) {
      const double2* sl = reinterpret_cast<const double2*>(wbase + stage_c * SCH_STAGE_BYTES + lane * SCH_ESTRIDE);
      double yj[18];
#pragma unroll
      for (int k = 0; k < 9; ++k) { const double2 t = sl[9 + k]; yj[2 * k] = t.x; yj[2 * k + 1] = t.y; }
#pragma unroll
      for (int h = 0; h < 3; ++h) {  // rows 2h, 2h+1 of Y_i
        const double2 t0 = sl[3 * h], t1 = sl[3 * h + 1], t2 = sl[3 * h + 2];
        const double yi[6] = {t0.x, t0.y, t1.x, t1.y, t2.x, t2.y};
#pragma unroll
        for (int rr2 = 0; rr2 < 2; ++rr2)
#pragma unroll
          for (int c = 0; c < 6; ++c)
            acc[6 * (2 * h + rr2) + c] += yi[3 * rr2] * yj[3 * c] + yi[3 * rr2 + 1] * yj[3 * c + 1] +
                                          yi[3 * rr2 + 2] * yj[3 * c + 2];
      }
    }
    __syncwarp();
    const int ch_done = Cc.ch;
    if (cursor_advance(pb, Cc, nwarp)) {
#pragma unroll
      for (int k = 0; k < 36; ++k) acc[k] = warp_sum(acc[k]);
      if (lane == 0) {
#pragma unroll
        for (int k = 0; k < 36; ++k) spart[(int64_t)ch_done * 36 + k] = acc[k];
      }
#pragma unroll
      for (int k = 0; k < 36; ++k) acc[k] = 0.0;
    }
    stage_c = (stage_c + 1 == SCH_STAGES) ? 0 : stage_c + 1;
    stage_f = (stage_f + 1 == SCH_STAGES) ? 0 : stage_f + 1;
  }
  cp_async_wait<0>();
}

// ---- Schur tile kernel (default path).  A CTA stages the Y records of one tile of observations (whole points,
// <= 1344 observations = 193.5 KB) into shared memory ONCE with fully coalesced 16-byte LDGSTS copies, then every
// lane owns one "slot" = all pair entries of one (tile, block) cell and accumulates its 6x6 partial in registers,
// reading both records of an entry from shared memory.  Compared with the streaming kernel this cuts the L2->SM
// traffic from 288 B per pair entry to 144 B per observation (6x less at 5 observations per point); no cross-lane
// reduction is needed, slots of a tile are sorted by entry count so the lanes of a warp finish together.
constexpr int SCHT_THREADS = 512;
constexpr int SCHT_G = 4;                                   // lanes cooperating on one slot
constexpr int SCHT_ENT_CAP = 9216;                          // tile entries staged in shared memory (36 KB)
constexpr int SCHT_SMEM_MAX = SCHUR_TILE_MAX * 144 + SCHT_ENT_CAP * 4;   // 230400 B

__device__ __forceinline__ void cp_async4(void* dst_smem, const void* src) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(dst_smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(d), "l"(src) : "memory");
}

__global__ void __launch_bounds__(SCHT_THREADS)
k_schur_tiles(Problem pb, const double* __restrict__ Y, double* __restrict__ spart) {
  extern __shared__ __align__(16) unsigned char sY_raw[];
  const double* sY = reinterpret_cast<const double*>(sY_raw);
  unsigned* sE = reinterpret_cast<unsigned*>(sY_raw + (size_t)pb.tile_obs * 144);
  const int g = threadIdx.x % SCHT_G, sub = threadIdx.x / SCHT_G;
  for (int t = blockIdx.x; t < pb.n_st; t += gridDim.x) {
    const int a = pb.st_off[t], n = pb.st_off[t + 1] - a;
    const int te0 = pb.tile_ent_off[t], nte = pb.tile_ent_off[t + 1] - te0;
    const bool ent_in_smem = nte <= SCHT_ENT_CAP;
    const char* src = reinterpret_cast<const char*>(Y + 18 * (int64_t)a);
    for (int q = threadIdx.x; q < n * 9; q += SCHT_THREADS) cp_async16(sY_raw + 16 * q, src + 16 * (int64_t)q);
    if (ent_in_smem)
      for (int q = threadIdx.x; q < nte; q += SCHT_THREADS) cp_async4(sE + q, pb.tent + te0 + q);
    cp_async_commit();
    cp_async_wait<0>();
    __syncthreads();
    const int s_lo = pb.tile_slot_off[t], s_hi = pb.tile_slot_off[t + 1];
    for (int s0 = s_lo; s0 < s_hi; s0 += SCHT_THREADS / SCHT_G) {
      const int s = s0 + sub;
      const bool active = s < s_hi;
      const int ea = active ? pb.slot_a[s] : 0, en = active ? pb.slot_n[s] : 0;
      const unsigned* ent = ent_in_smem ? (sE + (ea - te0)) : (pb.tent + ea);
      double acc[36];
#pragma unroll
      for (int k = 0; k < 36; ++k) acc[k] = 0.0;
      for (int e = g; e < en; e += SCHT_G) {
        const unsigned ij = ent[e];
        const double2* pi = reinterpret_cast<const double2*>(sY + 18 * (ij >> 16));
        const double2* pj = reinterpret_cast<const double2*>(sY + 18 * (ij & 0xffffu));
        double yj[18];
#pragma unroll
        for (int k = 0; k < 9; ++k) { const double2 v = pj[k]; yj[2 * k] = v.x; yj[2 * k + 1] = v.y; }
#pragma unroll
        for (int h = 0; h < 3; ++h) {
          const double2 t0 = pi[3 * h], t1 = pi[3 * h + 1], t2 = pi[3 * h + 2];
          const double yi[6] = {t0.x, t0.y, t1.x, t1.y, t2.x, t2.y};
#pragma unroll
          for (int r2 = 0; r2 < 2; ++r2)
#pragma unroll
            for (int c = 0; c < 6; ++c)
              acc[6 * (2 * h + r2) + c] += yi[3 * r2] * yj[3 * c] + yi[3 * r2 + 1] * yj[3 * c + 1] +
                                           yi[3 * r2 + 2] * yj[3 * c + 2];
        }
      }
#pragma unroll
      for (int k = 0; k < 36; ++k) {
        acc[k] += __shfl_xor_sync(FULL, acc[k], 1);
        acc[k] += __shfl_xor_sync(FULL, acc[k], 2);
      }
      if (active && g == 0) {
        double2* o = reinterpret_cast<double2*>(spart + 36 * (int64_t)s);
#pragma unroll
        for (int k = 0; k < 18; ++k) o[k] = make_double2(acc[2 * k], acc[2 * k + 1]);
      }
    }
    __syncthreads();  // the next tile overwrites the staged records
  }
}

// tile mode: Sup[g] = [g diagonal ? U*_c : 0] - sum over the slots of block g (fixed order => reproducible)
__global__ void k_schur_reduce_tiles(int n_blk, const int* __restrict__ blk_r, const int* __restrict__ blk_c,
                                     const int* __restrict__ blk_slot_off, const int* __restrict__ blk_slots,
                                     const double* __restrict__ spart, const double* __restrict__ U, double lambda,
                                     double* __restrict__ Sup) {
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t >= (int64_t)n_blk * 36) return;
  const int g = (int)(t / 36), e = (int)(t % 36);
  double s = 0.0;
  for (int q = blk_slot_off[g]; q < blk_slot_off[g + 1]; ++q) s += spart[(int64_t)blk_slots[q] * 36 + e];
  double val = -s;
  if (blk_r[g] == blk_c[g]) {
    const int r = e / 6, c = e % 6;
    double uv = U[blk_r[g] * 21 + tri6(min(r, c), max(r, c))];
    if (r == c) uv *= (1.0 + lambda);
    val += uv;
  }
  Sup[t] = val;
}

// Sup[g] = [g diagonal ? U*_c : 0] - sum of chunk partials      (U* = U + lambda diag U)
__global__ void k_schur_reduce(int n_blk, const int* __restrict__ blk_r, const int* __restrict__ blk_c,
                               const int* __restrict__ blk_choff, const double* __restrict__ spart,
                               const double* __restrict__ U, double lambda, double* __restrict__ Sup) {
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t >= (int64_t)n_blk * 36) return;
  const int g = (int)(t / 36), e = (int)(t % 36);
  double s = 0.0;
  for (int ch = blk_choff[g]; ch < blk_choff[g + 1]; ++ch) s += spart[(int64_t)ch * 36 + e];
  double val = -s;
  if (blk_r[g] == blk_c[g]) {
    const int r = e / 6, c = e % 6;
    double uv = U[blk_r[g] * 21 + tri6(min(r, c), max(r, c))];
    if (r == c) uv *= (1.0 + lambda);
    val += uv;
  }
  Sup[t] = val;
}

// cameras without any observation (globally): unit diagonal so the reduced system stays non-singular
__global__ void k_fix_diag(int C, const int* __restrict__ diag_blk, const double* __restrict__ udiag,
                           double* __restrict__ Sup) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= 6 * C) return;
  if (udiag[t] == 0.0) Sup[(int64_t)diag_blk[t / 6] * 36 + 7 * (t % 6)] = 1.0;
}

// Fixed cameras (the `fixed_camera_indices` hook of bundle_adjustment.py:6,13: their Jacobian columns are masked out,
// so scipy never moves them).  Holding camera c fixed = deleting row/column c of the reduced system, which is
// independent of the point elimination: blocks touching c become 0 (identity on the diagonal), b_c = g_c = 0, hence
// delta_c = 0 exactly and the optimality norm ignores c.
__global__ void k_fix_cameras(int n_blk, int C, const int* __restrict__ blk_r, const int* __restrict__ blk_c,
                              const unsigned char* __restrict__ fixed, double* __restrict__ Sup, double* __restrict__ b,
                              double* __restrict__ gc) {
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t < (int64_t)n_blk * 36) {
    const int q = (int)(t / 36), e = (int)(t % 36);
    const int r = blk_r[q], c = blk_c[q];
    if (fixed[r] || fixed[c]) Sup[t] = (r == c && e % 7 == 0) ? 1.0 : 0.0;
  }
  if (t < 6 * (int64_t)C && fixed[t / 6]) { b[t] = 0.0; gc[t] = 0.0; }
}

// camera-side scalars of the step: g_c.dc, dc^T diag(U) dc, |dc|^2, |x_c|^2, max|g_c|.  Two stages (per-CTA partials over
// contiguous chunks, then one CTA sums them in CTA order): a single CTA over 60 000 entries took 63 us per LM iteration.
__global__ void __launch_bounds__(1024)
k_cam_dots(int n, int chunk, const double* __restrict__ gc, const double* __restrict__ dc, const double* __restrict__ udiag,
           const double* __restrict__ xc, double* __restrict__ part, double* __restrict__ scal) {
  __shared__ double sm[32 * 5];
  double v[5] = {0, 0, 0, 0, 0};
  double mx = 0.0;
  const int lo = blockIdx.x * chunk, hi = min(n, lo + chunk);
#pragma unroll 4
  for (int i = lo + threadIdx.x; i < hi; i += blockDim.x) {
    const double g = gc[i], d = dc[i], xx = xc[i];
    v[0] += g * d;
    v[1] += d * d * udiag[i];
    v[2] += d * d;
    v[3] += xx * xx;
    mx = fmax(mx, fabs(g));
  }
  mx = warp_max(mx);
  block_sum<4>(v, sm);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = mx;
  __syncthreads();
  if (threadIdx.x == 0) {
    double m = 0.0;
    for (int w = 0; w < (blockDim.x >> 5); ++w) m = fmax(m, sm[w]);
    if (gridDim.x == 1) {   // small problems: one CTA, results straight into the scalar block (one launch less)
      scal[S_GC_DC] = v[0]; scal[S_DC_D_DC] = v[1]; scal[S_DC2] = v[2]; scal[S_XC2] = v[3]; scal[S_GCMAX] = m;
    } else {
      double* o = part + 8 * blockIdx.x;
      o[0] = v[0]; o[1] = v[1]; o[2] = v[2]; o[3] = v[3]; o[4] = m;
    }
  }
}

__global__ void k_cam_dots_final(int nparts, const double* __restrict__ part, double* __restrict__ scal) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  double v[4] = {0, 0, 0, 0}, m = 0.0;
  for (int b = 0; b < nparts; ++b) {
#pragma unroll
    for (int k = 0; k < 4; ++k) v[k] += part[8 * b + k];
    m = fmax(m, part[8 * b + 4]);
  }
  scal[S_GC_DC] = v[0];
  scal[S_DC_D_DC] = v[1];
  scal[S_DC2] = v[2];
  scal[S_XC2] = v[3];
  scal[S_GCMAX] = m;
}

// dense symmetric 6C x 6C copy of the block upper triangle (tests, Cholesky)
__global__ void k_densify(int n_blk, int C, const int* __restrict__ blk_r, const int* __restrict__ blk_c,
                          const double* __restrict__ Sup, double* __restrict__ dense) {
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t >= (int64_t)n_blk * 36) return;
  const int g = (int)(t / 36), e = (int)(t % 36);
  const int64_t n = 6 * (int64_t)C;
  const int64_t row = 6 * (int64_t)blk_r[g] + e / 6, col = 6 * (int64_t)blk_c[g] + e % 6;
  const double v = Sup[t];
  dense[row * n + col] = v;
  if (blk_r[g] != blk_c[g]) dense[col * n + row] = v;
}
