// Fused linearisation + point-block Schur elimination (included inside ba_kernels.cu's anonymous namespace).
//
// One CTA owns a tile = a run of whole points whose per-observation records fit shared memory.  The records
//     rec_i = [ Jc_i (2x6) | A_i = Jp_i L^-T (2x3) | r_i (2) | rt_i = r_i - A_i z_p (2) ]        22 doubles
// are produced in phase 1 and consumed in phases 2 and 3 without ever touching HBM:
//   phase 1  one lane per observation (warp tiles of whole points, as k_lin_points): residual, analytic blocks,
//            V_p / g_p by a segmented shuffle scan, damped 3x3 Cholesky; pp / gp go to HBM for the back-substitution.
//   phase 2  "self" slots = the tile's observations of one camera: 4 lanes per slot accumulate in registers
//                N_c  += Jc^T (A A^T - I) Jc   (21 unique)      => S_cc = -N_c - pairs + lambda diag U
//                dU_c += diag(Jc^T Jc),  g_c += Jc^T r,  b_c -= Jc^T rt
//   phase 3  "pair" slots = the ordered pairs (i, j), i != j, cam_i <= cam_j of one point that fall into one block of
//            S: 4 lanes per slot accumulate  Jc_i^T (A_i A_j^T) Jc_j  in registers.
// Every slot partial is written exactly once; k_slot_reduce_* sum them in list order: no atomics, bit-reproducible.
// Replaces the J^T J / J^T r products and the LSMR operator inside scipy's TRF (sfm_lite/sfm.py:266-268).

constexpr int FREC = 22;    // doubles per observation record
#ifndef B200BA_FG
#define B200BA_FG 2
#endif
constexpr int FG = B200BA_FG;
#ifndef FLOW_STORE_INTERLEAVE
#define FLOW_STORE_INTERLEAVE 1
#endif  // lanes cooperating on one slot (2 or 4)
constexpr int DPART = 40;   // doubles per self-slot partial: N (21) | diag U (6) | g_c (6) | b_c (6) | pad

__device__ __forceinline__ void store_record(double* __restrict__ rec, const double* Jc, const double* Jp,
                                             const double* Li, const double* z, double r0, double r1) {
  double A[6];
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    A[3 * r + 0] = Jp[3 * r] * Li[0];
    A[3 * r + 1] = Jp[3 * r] * Li[1] + Jp[3 * r + 1] * Li[2];
    A[3 * r + 2] = Jp[3 * r] * Li[3] + Jp[3 * r + 1] * Li[4] + Jp[3 * r + 2] * Li[5];
  }
  const double rt0 = r0 - (A[0] * z[0] + A[1] * z[1] + A[2] * z[2]);
  const double rt1 = r1 - (A[3] * z[0] + A[4] * z[1] + A[5] * z[2]);
  double2* o = reinterpret_cast<double2*>(rec);
#pragma unroll
  for (int k = 0; k < 6; ++k) o[k] = make_double2(Jc[2 * k], Jc[2 * k + 1]);
  o[6] = make_double2(A[0], A[1]);
  o[7] = make_double2(A[2], A[3]);
  o[8] = make_double2(A[4], A[5]);
  o[9] = make_double2(r0, r1);
  o[10] = make_double2(rt0, rt1);
}

// Sum `acc[2*H]` over the 4 lanes of a slot group so that every lane ends with H/2 of the sums (two exchange steps,
// 3H/2 additions instead of the 4H of an all-reduce butterfly).  Lane g receives entries [(H/2) * g', (H/2) * (g' + 1))
// with g' = 2 * (g & 1) + (g >> 1).
template <int H>
__device__ __forceinline__ void group_reduce_scatter(const double* acc, double* out, int g) {
  const bool odd = g & 1, hi = g & 2;
  double h[H];
#pragma unroll
  for (int k = 0; k < H; ++k) {
    const double mine = odd ? acc[H + k] : acc[k];
    const double give = odd ? acc[k] : acc[H + k];
    h[k] = mine + __shfl_xor_sync(FULL, give, 1);
  }
#pragma unroll
  for (int k = 0; k < H / 2; ++k) {
    const double mine = hi ? h[H / 2 + k] : h[k];
    const double give = hi ? h[k] : h[H / 2 + k];
    out[k] = mine + __shfl_xor_sync(FULL, give, 2);
  }
}

// Sum the NV accumulators of a slot over its FG lanes and store the partial: every lane writes NV / FG consecutive values.
template <int NV>
__device__ __forceinline__ void slot_reduce_store(const double* acc, double* __restrict__ dst, int g, bool active) {
  if constexpr (FG == 4) {
    double q[NV / 4];
    group_reduce_scatter<NV / 2>(acc, q, g);
    if (active) {
      double* o = dst + (NV / 4) * (((g & 1) << 1) | (g >> 1));
#pragma unroll
      for (int k = 0; k < NV / 4; ++k) o[k] = q[k];
    }
  } else {
    static_assert(FG == 2 || FG == 4, "FG must be 2 or 4");
    const bool odd = g & 1;
    double q[NV / 2];
#if FLOW_STORE_INTERLEAVE
    // the two lanes of a slot take alternate 16-byte chunks of the partial (even / odd pairs of values), so together
    // they fill whole 32-byte sectors with every store instruction instead of two half-used ones 144 bytes apart
    static_assert(NV % 4 == 0, "interleaved store needs whole chunk pairs");
#pragma unroll
    for (int c = 0; c < NV / 4; ++c)
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const double mine = odd ? acc[4 * c + 2 + h] : acc[4 * c + h];
        const double give = odd ? acc[4 * c + h] : acc[4 * c + 2 + h];
        q[2 * c + h] = mine + __shfl_xor_sync(FULL, give, 1);
      }
    if (active) {
      double2* o = reinterpret_cast<double2*>(dst) + (g & 1);
#pragma unroll
      for (int c = 0; c < NV / 4; ++c) o[2 * c] = make_double2(q[2 * c], q[2 * c + 1]);
    }
#else
#pragma unroll
    for (int k = 0; k < NV / 2; ++k) {
      const double mine = odd ? acc[NV / 2 + k] : acc[k];
      const double give = odd ? acc[k] : acc[NV / 2 + k];
      q[k] = mine + __shfl_xor_sync(FULL, give, 1);
    }
    if (active) {
      double* o = dst + (NV / 2) * (g & 1);
#pragma unroll
      for (int k = 0; k < NV / 2; ++k) o[k] = q[k];
    }
#endif
  }
}

// ---- slot bodies shared by k_lin_fused and k_lin_flow (pure FMA chains: 2 DFMA per accumulator and entry) ----------
// pair slot: acc(6x6) += Jc_i^T (A_i A_j^T) Jc_j over the entries e = g, g + FG, ... of the slot
__device__ __forceinline__ void pair_slot_body(const double* __restrict__ rec, const unsigned* __restrict__ ent, int en,
                                               int g, double* __restrict__ acc /*[36]*/) {
  for (int e = g; e < en; e += FG) {
    const unsigned cur = ent[e];
    const double2* ri = reinterpret_cast<const double2*>(rec + FREC * (cur >> 16));
    const double2* rj = reinterpret_cast<const double2*>(rec + FREC * (cur & 0xffffu));
    const double2 a01 = ri[6], a23 = ri[7], a45 = ri[8];
    const double2 b01 = rj[6], b23 = rj[7], b45 = rj[8];
    // M = A_i A_j^T (2x2)
    const double m00 = fma(a23.x, b23.x, fma(a01.y, b01.y, a01.x * b01.x));
    const double m01 = fma(a23.x, b45.y, fma(a01.y, b45.x, a01.x * b23.y));
    const double m10 = fma(a45.y, b23.x, fma(a45.x, b01.y, a23.y * b01.x));
    const double m11 = fma(a45.y, b45.y, fma(a45.x, b45.x, a23.y * b23.y));
    double G0[6], G1[6];  // G = M Jc_j (2x6)
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const double2 j0 = rj[k], j1 = rj[3 + k];
      G0[2 * k] = fma(m01, j1.x, m00 * j0.x);     G0[2 * k + 1] = fma(m01, j1.y, m00 * j0.y);
      G1[2 * k] = fma(m11, j1.x, m10 * j0.x);     G1[2 * k + 1] = fma(m11, j1.y, m10 * j0.y);
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const double2 i0 = ri[k], i1 = ri[3 + k];
#pragma unroll
      for (int n = 0; n < 6; ++n) {
        acc[6 * (2 * k) + n] = fma(i1.x, G1[n], fma(i0.x, G0[n], acc[6 * (2 * k) + n]));
        acc[6 * (2 * k + 1) + n] = fma(i1.y, G1[n], fma(i0.y, G0[n], acc[6 * (2 * k + 1) + n]));
      }
    }
  }
}

// self slot: acc[0..20] += Jc^T (A A^T - I) Jc (upper triangle), acc[21..26] += diag(Jc^T Jc), acc[27..32] += Jc^T r,
// acc[33..38] += Jc^T rt over the entries of the slot
__device__ __forceinline__ void self_slot_body(const double* __restrict__ rec, const unsigned short* __restrict__ ent,
                                               int en, int g, double* __restrict__ acc /*[40]*/) {
  for (int e = g; e < en; e += FG) {
    const double2* rc = reinterpret_cast<const double2*>(rec + FREC * (int)ent[e]);
    double J[12];
#pragma unroll
    for (int k = 0; k < 6; ++k) { const double2 v = rc[k]; J[2 * k] = v.x; J[2 * k + 1] = v.y; }
    const double2 A01 = rc[6], A23 = rc[7], A45 = rc[8], rv = rc[9], rtv = rc[10];
    const double m00 = fma(A23.x, A23.x, fma(A01.y, A01.y, fma(A01.x, A01.x, -1.0)));
    const double m01 = fma(A23.x, A45.y, fma(A01.y, A45.x, A01.x * A23.y));
    const double m11 = fma(A45.y, A45.y, fma(A45.x, A45.x, fma(A23.y, A23.y, -1.0)));
    int k = 0;
#pragma unroll
    for (int m = 0; m < 6; ++m) {
      const double h0 = fma(J[6 + m], m01, J[m] * m00);  // row m of Jc^T M
      const double h1 = fma(J[6 + m], m11, J[m] * m01);
#pragma unroll
      for (int n = m; n < 6; ++n, ++k) acc[k] = fma(h1, J[6 + n], fma(h0, J[n], acc[k]));
    }
#pragma unroll
    for (int m = 0; m < 6; ++m) {
      acc[21 + m] = fma(J[6 + m], J[6 + m], fma(J[m], J[m], acc[21 + m]));
      acc[27 + m] = fma(J[6 + m], rv.y, fma(J[m], rv.x, acc[27 + m]));
      acc[33 + m] = fma(J[6 + m], rtv.y, fma(J[m], rtv.x, acc[33 + m]));
    }
  }
}

template <int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB)
k_lin_fused(Problem pb, const double* __restrict__ x, const double* __restrict__ camprep, double lambda,
            double* __restrict__ pp, double* __restrict__ gp, double* __restrict__ spart, double* __restrict__ dpart,
            double* __restrict__ part) {
  extern __shared__ __align__(16) double srec[];
  constexpr int NW = THREADS / 32;
  __shared__ double sred[2 * NW];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int g = threadIdx.x % FG, sub = threadIdx.x / FG;
  const double* xp = x + 6 * pb.C;
  double cost2 = 0.0, gmax = 0.0;

  for (int t = blockIdx.x; t < pb.n_st; t += gridDim.x) {
    const int a0 = pb.st_off[t];
    // ------------------------------------------------ phase 1: records
    for (int wt = pb.st_wt_off[t] + wid; wt < pb.st_wt_off[t + 1]; wt += NW) {
      const int a = pb.wt_off[wt], b = pb.wt_off[wt + 1], n = b - a;
      if (n <= 32) {
        const bool valid = lane < n;
        const int i = a + (valid ? lane : 0);
        const int c = pb.cam[i], p = pb.pt[i];
        double r0, r1, Jc[12], Jp[6], vals[9], cp[CAMPREP_FIELDS];
        load_cam(camprep, pb.Cs, c, cp);
        double c2;
        project_jac_loss(cp, pb.K, pb.loss, xp[3 * p], xp[3 * p + 1], xp[3 * p + 2], pb.u[i], pb.v[i], r0, r1, Jc, Jp, c2);
        point_products(Jp, r0, r1, vals);
        if (!valid) {
#pragma unroll
          for (int k = 0; k < 9; ++k) vals[k] = 0.0;
        }
        const SegInfo s = seg_info(p, valid, lane);
        seg_allsum<9>(vals, s, lane);
        double Li[6], z[3];
        point_factor(vals, lambda, Li);
        z[0] = Li[0] * vals[6];
        z[1] = Li[1] * vals[6] + Li[2] * vals[7];
        z[2] = Li[3] * vals[6] + Li[4] * vals[7] + Li[5] * vals[8];
        if (valid) {
          cost2 += c2;
          store_record(srec + FREC * (i - a0), Jc, Jp, Li, z, r0, r1);
          if (lane == s.first) {
            store_point(pp, gp, p, Li, z, vals);
            gmax = fmax(gmax, fmax(fabs(vals[6]), fmax(fabs(vals[7]), fabs(vals[8]))));
          }
        }
      } else {
        // one point with more than 32 observations: two passes over its run
        const int p = pb.pt[a];
        const double X0 = xp[3 * p], X1 = xp[3 * p + 1], X2 = xp[3 * p + 2];
        double vals[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) vals[k] = 0.0;
        for (int i = a + lane; i < b; i += 32) {
          double r0, r1, Jc[12], Jp[6], v9[9], cp[CAMPREP_FIELDS];
          load_cam(camprep, pb.Cs, pb.cam[i], cp);
          double c2;
        project_jac_loss(cp, pb.K, pb.loss, X0, X1, X2, pb.u[i], pb.v[i], r0, r1, Jc, Jp, c2);
          point_products(Jp, r0, r1, v9);
#pragma unroll
          for (int k = 0; k < 9; ++k) vals[k] += v9[k];
          cost2 += c2;
        }
#pragma unroll
        for (int k = 0; k < 9; ++k) vals[k] = warp_sum(vals[k]);
        double Li[6], z[3];
        point_factor(vals, lambda, Li);
        z[0] = Li[0] * vals[6];
        z[1] = Li[1] * vals[6] + Li[2] * vals[7];
        z[2] = Li[3] * vals[6] + Li[4] * vals[7] + Li[5] * vals[8];
        if (lane == 0) {
          store_point(pp, gp, p, Li, z, vals);
          gmax = fmax(gmax, fmax(fabs(vals[6]), fmax(fabs(vals[7]), fabs(vals[8]))));
        }
        for (int i = a + lane; i < b; i += 32) {
          double r0, r1, Jc[12], Jp[6], cp[CAMPREP_FIELDS];
          load_cam(camprep, pb.Cs, pb.cam[i], cp);
          double c2;
        project_jac_loss(cp, pb.K, pb.loss, X0, X1, X2, pb.u[i], pb.v[i], r0, r1, Jc, Jp, c2);
          store_record(srec + FREC * (i - a0), Jc, Jp, Li, z, r0, r1);
        }
      }
    }
    __syncthreads();

    // ------------------------------------------------ phase 2: self slots (one camera of the tile each)
    {
      const int s_lo = pb.tile_dslot_off[t], s_hi = pb.tile_dslot_off[t + 1];
      for (int s0 = s_lo; s0 < s_hi; s0 += THREADS / FG) {
        const int s = s0 + sub;
        const bool active = s < s_hi;
        const int ea = active ? pb.dslot_a[s] : 0, en = active ? pb.dslot_n[s] : 0;
        double acc[40];
#pragma unroll
        for (int k = 0; k < 40; ++k) acc[k] = 0.0;
        self_slot_body(srec, pb.dent + ea, en, g, acc);
        slot_reduce_store<40>(acc, dpart + DPART * (int64_t)s, g, active);
      }
    }

    // ------------------------------------------------ phase 3: pair slots (one block of S each)
    {
      const int s_lo = pb.tile_slot_off[t], s_hi = pb.tile_slot_off[t + 1];
      for (int s0 = s_lo; s0 < s_hi; s0 += THREADS / FG) {
        const int s = s0 + sub;
        const bool active = s < s_hi;
        const int ea = active ? pb.slot_a[s] : 0, en = active ? pb.slot_n[s] : 0;
        double acc[36];
#pragma unroll
        for (int k = 0; k < 36; ++k) acc[k] = 0.0;
        pair_slot_body(srec, pb.tent + ea, en, g, acc);
        slot_reduce_store<36>(acc, spart + 36 * (int64_t)s, g, active);
      }
    }
    __syncthreads();  // the next tile overwrites the records
  }

  cost2 = warp_sum(cost2);
  gmax = warp_max(gmax);
  if (lane == 0) { sred[wid] = cost2; sred[NW + wid] = gmax; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0, m = 0.0;
    for (int w = 0; w < NW; ++w) { s += sred[w]; m = fmax(m, sred[NW + w]); }
    part[blockIdx.x * 8 + 0] = s;
    part[blockIdx.x * 8 + 1] = m;
  }
}

// Fixed-order sum of a list of strided partials with 8 independent chains (the loads of a round are in flight
// together; the order of the additions is fixed, so the result is bit-reproducible).
#ifndef REDUCE_LDCS
#define REDUCE_LDCS 0
#endif
__device__ __forceinline__ double list_sum8(const int* __restrict__ list, int a, int b, const double* __restrict__ base,
                                            int stride, int e) {
  double s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (int q = a; q < b; q += 8) {
    int id[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) id[u] = (q + u < b) ? list[q + u] : -1;
#pragma unroll
    for (int u = 0; u < 8; ++u)
      if (id[u] >= 0) s[u] += REDUCE_LDCS ? __ldcs(base + (int64_t)id[u] * stride + e) : base[(int64_t)id[u] * stride + e];
  }
  return ((s[0] + s[1]) + (s[2] + s[3])) + ((s[4] + s[5]) + (s[6] + s[7]));
}

// per camera: sums of its self-slot partials -> camsum[c][40] (N 21 | diag U 6 | g_c 6 | b_c 6), and udiag, g_c, b
__global__ void k_cam_sums(int C, const int* __restrict__ cam_dslot_off, const int* __restrict__ cam_dslots,
                           const double* __restrict__ dpart, double* __restrict__ camsum, double* __restrict__ udiag,
                           double* __restrict__ gc, double* __restrict__ b) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= C * DPART) return;
  const int c = t / DPART, k = t % DPART;
  const double s = (k < 39) ? list_sum8(cam_dslots, cam_dslot_off[c], cam_dslot_off[c + 1], dpart, DPART, k) : 0.0;
  camsum[t] = s;
  if (k >= 21 && k < 27) udiag[6 * c + (k - 21)] = s;
  else if (k >= 27 && k < 33) gc[6 * c + (k - 27)] = s;
  else if (k >= 33 && k < 39) b[6 * c + (k - 33)] = -s;
}

// Sup[g] = -(sum of the pair-slot partials of block g) [- N_c + lambda diag U_c on diagonal blocks]
// blk_list == nullptr: all n_blk blocks, Sup indexed by block; else the listed blocks only, Sup indexed by list position
// (sparse exchange: this rank's local blocks, packed)
__global__ void k_slot_reduce_S(int n_blk, const int* __restrict__ blk_list, const int* __restrict__ blk_r,
                                const int* __restrict__ blk_c, const int* __restrict__ blk_slot_off,
                                const int* __restrict__ blk_slots, const double* __restrict__ spart,
                                const double* __restrict__ camsum, double lambda, double* __restrict__ Sup, PeerDst dst) {
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t >= (int64_t)n_blk * 36) return;
  const int e = (int)(t % 36);
  const int g = blk_list ? blk_list[t / 36] : (int)(t / 36);
  double s = list_sum8(blk_slots, blk_slot_off[g], blk_slot_off[g + 1], spart, 36, e);
  if (blk_r[g] == blk_c[g]) {
    const int c = blk_r[g], r = e / 6, cc = e % 6;
    s += camsum[c * DPART + tri6(min(r, cc), max(r, cc))];
    if (r == cc) s -= lambda * camsum[c * DPART + 21 + r];
  }
  if (dst.n > 0) {
    // multi-GPU: the reduced block goes straight into this rank's receive slot on every rank (peer memory over NVLink;
    // own buffer included) -- the transfer overlaps the reduction instead of following it as a separate copy pass
    // (no fence here: the kernel boundary orders these stores before k_p2p_push, whose last CTA publishes the epoch with a
    // system-scope fence + st.release.sys; a per-thread system fence made this kernel 60 % slower)
    for (int k = 0; k < dst.n; ++k) dst.p[k][t] = -s;
  } else {
    Sup[t] = -s;
  }
}

// Sparse exchange, receiving side: every block of S = sum of the ranks' packed contributions in ascending rank order;
// b, g_c, diag U and the cost = sums over ranks of the per-rank tails; max |g_p| = max over ranks.
__global__ void k_xchg_sum(int n_blk, int C, int nranks, long long stride, long long tail_off,
                           const int* __restrict__ xc_off, const long long* __restrict__ xc_src,
                           const double* __restrict__ xrecv, double* __restrict__ comm, double* __restrict__ sums,
                           double* __restrict__ scal) {
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int64_t nS = (int64_t)n_blk * 36, nV = 18 * (int64_t)C;
  if (t < nS) {
    const int g = (int)(t / 36), e = (int)(t % 36);
    double s = 0.0;
    for (int q = xc_off[g]; q < xc_off[g + 1]; ++q) s += xrecv[xc_src[q] + e];
    comm[t] = s;
  } else if (t < nS + nV) {
    const int64_t k = t - nS;
    double s = 0.0;
    for (int r = 0; r < nranks; ++r) s += xrecv[r * stride + tail_off + k];
    comm[t] = s;
  } else if (t == nS + nV) {
    double s = 0.0, m = 0.0;
    for (int r = 0; r < nranks; ++r) {
      s += xrecv[r * stride + tail_off + nV + S_COST2];
      m = fmax(m, xrecv[r * stride + tail_off + nV + S_NSUM]);
    }
    sums[S_COST2] = s;
    scal[S_GPMAX] = m;
  }
}
