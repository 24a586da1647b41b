// Point-major fused linearisation (included inside ba_kernels.cu's anonymous namespace).
//
// A warp owns a "tile" = a run of consecutive observations made of whole points (<= 32 observations,
// or one single point with more).  One observation per lane: residual + analytic 2x6 / 2x3 blocks stay
// in registers; V_p = sum Jp^T Jp and g_p = sum Jp^T r are reduced over each point's contiguous lane run
// with a segmented shuffle scan (no atomics, fixed order => bit-reproducible); every lane of the run then
// factors the damped 3x3 block and emits
//     Y_i  = Jc_i^T Jp_i L^-T        (6x3)  -> Schur:  S_{c_i c_j} -= Y_i Y_j^T
//     rt_i = r_i - Jp_i V*^-1 g_p    (2)    -> rhs:    b_c = -sum_i Jc_i^T rt_i
// which is all the camera side and the Schur kernel need.

struct SegInfo {
  int first, last, maxlen;
};

__device__ __forceinline__ SegInfo seg_info(int pt, bool valid, int lane) {
  const int prev = __shfl_up_sync(FULL, pt, 1);
  const bool head = !valid || lane == 0 || pt != prev;
  const unsigned heads = __ballot_sync(FULL, head);
  const unsigned below = heads & (0xffffffffu >> (31 - lane));
  const unsigned above = heads & ~((2u << lane) - 1u);
  SegInfo s;
  s.first = 31 - __clz(below);
  s.last = above ? (__ffs(above) - 2) : 31;
  s.maxlen = __reduce_max_sync(FULL, valid ? (s.last - s.first + 1) : 1);
  return s;
}

// after the call every lane holds the sum over its segment
template <int NV>
__device__ __forceinline__ void seg_allsum(double* v, const SegInfo& s, int lane) {
  for (int d = 1; d < s.maxlen; d <<= 1) {
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const double t = __shfl_up_sync(FULL, v[k], d);
      if (lane - d >= s.first) v[k] += t;
    }
  }
#pragma unroll
  for (int k = 0; k < NV; ++k) v[k] = __shfl_sync(FULL, v[k], s.last);
}

__device__ __forceinline__ void point_products(const double* Jp, double r0, double r1, double* vals) {
  vals[0] = Jp[0] * Jp[0] + Jp[3] * Jp[3];
  vals[1] = Jp[1] * Jp[0] + Jp[4] * Jp[3];
  vals[2] = Jp[1] * Jp[1] + Jp[4] * Jp[4];
  vals[3] = Jp[2] * Jp[0] + Jp[5] * Jp[3];
  vals[4] = Jp[2] * Jp[1] + Jp[5] * Jp[4];
  vals[5] = Jp[2] * Jp[2] + Jp[5] * Jp[5];
  vals[6] = Jp[0] * r0 + Jp[3] * r1;
  vals[7] = Jp[1] * r0 + Jp[4] * r1;
  vals[8] = Jp[2] * r0 + Jp[5] * r1;
}

// per-observation outputs once the point factor is known
__device__ __forceinline__ void obs_outputs(const double* Jc, const double* Jp, const double* Li, const double* z,
                                            double r0, double r1, double* Yv, double& rt0, double& rt1) {
  double A[6];
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    A[3 * r + 0] = Jp[3 * r] * Li[0];
    A[3 * r + 1] = Jp[3 * r] * Li[1] + Jp[3 * r + 1] * Li[2];
    A[3 * r + 2] = Jp[3 * r] * Li[3] + Jp[3 * r + 1] * Li[4] + Jp[3 * r + 2] * Li[5];
  }
  rt0 = r0 - (A[0] * z[0] + A[1] * z[1] + A[2] * z[2]);
  rt1 = r1 - (A[3] * z[0] + A[4] * z[1] + A[5] * z[2]);
#pragma unroll
  for (int q = 0; q < 6; ++q)
#pragma unroll
    for (int k = 0; k < 3; ++k) Yv[3 * q + k] = Jc[q] * A[k] + Jc[6 + q] * A[3 + k];
}

__device__ __forceinline__ void store_point(double* __restrict__ pp, double* __restrict__ gp, int p, const double* Li,
                                            const double* z, const double* vals) {
  double* o = pp + 12 * (int64_t)p;
#pragma unroll
  for (int k = 0; k < 6; ++k) o[k] = Li[k];
  o[6] = z[0]; o[7] = z[1]; o[8] = z[2];
  o[9] = vals[0]; o[10] = vals[2]; o[11] = vals[5];
  gp[3 * (int64_t)p] = vals[6]; gp[3 * (int64_t)p + 1] = vals[7]; gp[3 * (int64_t)p + 2] = vals[8];
}

// The same outputs written by the first lanes of the point's run together (every lane of the run holds identical Li, z,
// vals after the segmented all-sum): lanes j = 0, 1 write alternate 16-byte chunks of pp[p] (whole 32-byte sectors per
// store instruction, 3 instructions instead of 12 scattered 8-byte ones), lanes j < 3 one value of gp[p] each.
// j = lane - first lane of the run, n = run length; n == 1 falls back to one lane writing everything.
__device__ __forceinline__ void store_point_coop(double* __restrict__ pp, double* __restrict__ gp, int p, const double* Li,
                                                 const double* z, const double* vals, int j, int n) {
  if (n < 2) {
    if (j == 0) store_point(pp, gp, p, Li, z, vals);
    return;
  }
  if (j < 2) {
    const bool odd = j & 1;
    double2* o = reinterpret_cast<double2*>(pp + 12 * (int64_t)p) + j;
    o[0] = odd ? make_double2(Li[2], Li[3]) : make_double2(Li[0], Li[1]);
    o[2] = odd ? make_double2(z[0], z[1]) : make_double2(Li[4], Li[5]);
    o[4] = odd ? make_double2(vals[2], vals[5]) : make_double2(z[2], vals[0]);
  }
  if (j < 3) {
    const double v = j == 0 ? vals[6] : (j == 1 ? vals[7] : vals[8]);
    gp[3 * (int64_t)p + j] = v;
    if (n == 2 && j == 0) gp[3 * (int64_t)p + 2] = vals[8];
  }
}

// stage one warp's Y rows through shared memory and store them as one contiguous run
__device__ __forceinline__ void store_Y(double* sY /*[32*19]*/, const double* Yv, bool valid, int lane, int n,
                                        double* __restrict__ Ydst /* = Y + 18*a */) {
  if (valid) {
#pragma unroll
    for (int k = 0; k < 18; ++k) sY[lane * 19 + k] = Yv[k];
  }
  __syncwarp();
  for (int e = lane; e < n * 18; e += 32) Ydst[e] = sY[(e / 18) * 19 + (e % 18)];
  __syncwarp();
}

__global__ void __launch_bounds__(LIN_WARPS * 32)
k_lin_points(Problem pb, const double* __restrict__ x, const double* __restrict__ camprep, double lambda,
             double* __restrict__ Y, double* __restrict__ rr, double* __restrict__ pp, double* __restrict__ gp,
             double* __restrict__ part) {
  __shared__ double sY[LIN_WARPS][32 * 19];
  __shared__ double sred[32 * 2];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const double* xp = x + 6 * pb.C;
  double cost2 = 0.0, gmax = 0.0;
  for (int t = blockIdx.x * LIN_WARPS + wid; t < pb.n_wt; t += gridDim.x * LIN_WARPS) {
    const int a = pb.wt_off[t], b = pb.wt_off[t + 1], n = b - a;
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
      double Li[6], z[3], Yv[18], rt0, rt1;
      point_factor(vals, lambda, Li);
      z[0] = Li[0] * vals[6];
      z[1] = Li[1] * vals[6] + Li[2] * vals[7];
      z[2] = Li[3] * vals[6] + Li[4] * vals[7] + Li[5] * vals[8];
      obs_outputs(Jc, Jp, Li, z, r0, r1, Yv, rt0, rt1);
      if (valid) {
        cost2 += c2;
        double2* o = reinterpret_cast<double2*>(rr + 4 * (int64_t)i);
        o[0] = make_double2(r0, r1);
        o[1] = make_double2(rt0, rt1);
        if (lane == s.first) {
          store_point(pp, gp, p, Li, z, vals);
          gmax = fmax(gmax, fmax(fabs(vals[6]), fmax(fabs(vals[7]), fabs(vals[8]))));
        }
      }
      store_Y(sY[wid], Yv, valid, lane, n, Y + 18 * (int64_t)a);
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
      for (int base = a; base < b; base += 32) {
        const int nn = min(32, b - base);
        const bool valid = lane < nn;
        const int i = base + (valid ? lane : 0);
        double r0, r1, Jc[12], Jp[6], Yv[18], rt0, rt1, cp[CAMPREP_FIELDS];
        load_cam(camprep, pb.Cs, pb.cam[i], cp);
        double c2;
        project_jac_loss(cp, pb.K, pb.loss, X0, X1, X2, pb.u[i], pb.v[i], r0, r1, Jc, Jp, c2);
        obs_outputs(Jc, Jp, Li, z, r0, r1, Yv, rt0, rt1);
        if (valid) {
          double2* o = reinterpret_cast<double2*>(rr + 4 * (int64_t)i);
          o[0] = make_double2(r0, r1);
          o[1] = make_double2(rt0, rt1);
        }
        store_Y(sY[wid], Yv, valid, lane, nn, Y + 18 * (int64_t)base);
      }
    }
  }
  // per-CTA partials: [0] sum r^2, [1] max |g_p|
  cost2 = warp_sum(cost2);
  gmax = warp_max(gmax);
  if (lane == 0) { sred[wid] = cost2; sred[32 + wid] = gmax; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0, m = 0.0;
    for (int w = 0; w < LIN_WARPS; ++w) { s += sred[w]; m = fmax(m, sred[32 + w]); }
    part[blockIdx.x * 8 + 0] = s;
    part[blockIdx.x * 8 + 1] = m;
  }
}
