// Back-substitution fused with the trial residual (included inside ba_kernels.cu's anonymous namespace).
//
//   delta_p = -V*^-1 (g_p + sum_i Jp_i^T Jc_i delta_c)      
//   cost(x + delta) evaluated in the same pass with the updated cameras (camprep_n) -- this is the
//   `f_new = fun(x_new)` evaluation of scipy's TRF inner loop (scipy _lsq/trf.py), the call that lands in
//   sfm_lite/bundle_adjustment.py:35-42 in the reference.
// Same warp-tile / segmented-scan layout as k_lin_points; Jacobian blocks are recomputed (24 B/observation
// of traffic instead of re-reading 144 B/observation of stored blocks).

__device__ __forceinline__ void point_step(const double* __restrict__ ppp, const double* s3, double* dp, double* sums) {
  const double i00 = ppp[0], i10 = ppp[1], i11 = ppp[2], i20 = ppp[3], i21 = ppp[4], i22 = ppp[5];
  const double z0 = ppp[6], z1 = ppp[7], z2 = ppp[8];
  const double w0 = z0 + i00 * s3[0];
  const double w1 = z1 + i10 * s3[0] + i11 * s3[1];
  const double w2 = z2 + i20 * s3[0] + i21 * s3[1] + i22 * s3[2];
  dp[0] = -(i00 * w0 + i10 * w1 + i20 * w2);
  dp[1] = -(i11 * w1 + i21 * w2);
  dp[2] = -(i22 * w2);
  sums[0] = -(z0 * w0 + z1 * w1 + z2 * w2);                                        // g_p . dp
  sums[1] = dp[0] * dp[0] * ppp[9] + dp[1] * dp[1] * ppp[10] + dp[2] * dp[2] * ppp[11];  // dp^T diag(V) dp
  sums[2] = dp[0] * dp[0] + dp[1] * dp[1] + dp[2] * dp[2];
}

// as load_rt, plus the camera's step fields {J_r dw, dT} -> cp[9..14]
__device__ __forceinline__ void load_step_cam(const double* __restrict__ camprep, const double* __restrict__ cstep, int64_t Cs,
                                              int c, double* __restrict__ cp) {
  load_rt(camprep, Cs, c, cp);
#pragma unroll
  for (int k = 0; k < 6; ++k) cp[9 + k] = cstep[(int64_t)k * Cs + c];
}

__device__ __forceinline__ void jc_dc_jp(const double* Jc, const double* Jp, const double* __restrict__ d, double* s3) {
  const double t0 = Jc[0] * d[0] + Jc[1] * d[1] + Jc[2] * d[2] + Jc[3] * d[3] + Jc[4] * d[4] + Jc[5] * d[5];
  const double t1 = Jc[6] * d[0] + Jc[7] * d[1] + Jc[8] * d[2] + Jc[9] * d[3] + Jc[10] * d[4] + Jc[11] * d[5];
  s3[0] = Jp[0] * t0 + Jp[3] * t1;
  s3[1] = Jp[1] * t0 + Jp[4] * t1;
  s3[2] = Jp[2] * t0 + Jp[5] * t1;
}

__global__ void __launch_bounds__(LIN_WARPS * 32, 4)
k_backsub_trial(Problem pb, const double* __restrict__ x, double* __restrict__ xn, const double* __restrict__ camprep,
                const double* __restrict__ camprep_n, const double* __restrict__ cstep, const double* __restrict__ pp,
                double* __restrict__ part) {
  __shared__ double sred[LIN_WARPS * 5];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const double* xp = x + 6 * pb.C;
  double* xnp = xn + 6 * pb.C;
  double acc[5] = {0, 0, 0, 0, 0};  // newcost2, gp.dp, dp D dp, |dp|^2, |xp|^2
  for (int t = blockIdx.x * LIN_WARPS + wid; t < pb.n_wt; t += gridDim.x * LIN_WARPS) {
    const int a = pb.wt_off[t], b = pb.wt_off[t + 1], n = b - a;
    if (n <= 32) {
      const bool valid = lane < n;
      const int i = a + (valid ? lane : 0);
      const int c = pb.cam[i], p = pb.pt[i];
      const double X0 = xp[3 * p], X1 = xp[3 * p + 1], X2 = xp[3 * p + 2];
      const double uo = pb.u[i], vo = pb.v[i];
      double r0, r1, s3[3], cp[CAMPREP_FIELDS];
      load_step_cam(camprep, cstep, pb.Cs, c, cp);
      project_step(cp, cp + 18, cp + 9, pb.K, pb.loss, X0, X1, X2, uo, vo, s3);
      if (!valid) s3[0] = s3[1] = s3[2] = 0.0;
      const SegInfo s = seg_info(p, valid, lane);
      seg_allsum<3>(s3, s, lane);
      double dp[3], sums[3];
      point_step(pp + 12 * (int64_t)p, s3, dp, sums);
      const double N0 = X0 + dp[0], N1 = X1 + dp[1], N2 = X2 + dp[2];
      load_rt(camprep_n, pb.Cs, c, cp);
      project_residual(cp, pb.K, N0, N1, N2, uo, vo, r0, r1);
      if (valid) {
        acc[0] += robust_cost2(pb.loss, r0, r1);
        if (lane == s.first) {
          xnp[3 * (int64_t)p] = N0; xnp[3 * (int64_t)p + 1] = N1; xnp[3 * (int64_t)p + 2] = N2;
          acc[1] += sums[0]; acc[2] += sums[1]; acc[3] += sums[2];
          acc[4] += X0 * X0 + X1 * X1 + X2 * X2;
        }
      }
    } else {
      const int p = pb.pt[a];
      const double X0 = xp[3 * p], X1 = xp[3 * p + 1], X2 = xp[3 * p + 2];
      double s3[3] = {0, 0, 0};
      for (int i = a + lane; i < b; i += 32) {
        const int c = pb.cam[i];
        double t3[3], cp[CAMPREP_FIELDS];
        load_step_cam(camprep, cstep, pb.Cs, c, cp);
        project_step(cp, cp + 18, cp + 9, pb.K, pb.loss, X0, X1, X2, pb.u[i], pb.v[i], t3);
        s3[0] += t3[0]; s3[1] += t3[1]; s3[2] += t3[2];
      }
      s3[0] = warp_sum(s3[0]); s3[1] = warp_sum(s3[1]); s3[2] = warp_sum(s3[2]);
      double dp[3], sums[3];
      point_step(pp + 12 * (int64_t)p, s3, dp, sums);
      const double N0 = X0 + dp[0], N1 = X1 + dp[1], N2 = X2 + dp[2];
      for (int i = a + lane; i < b; i += 32) {
        double r0, r1, cp[CAMPREP_FIELDS];
        load_rt(camprep_n, pb.Cs, pb.cam[i], cp);
        project_residual(cp, pb.K, N0, N1, N2, pb.u[i], pb.v[i], r0, r1);
        acc[0] += robust_cost2(pb.loss, r0, r1);
      }
      if (lane == 0) {
        xnp[3 * (int64_t)p] = N0; xnp[3 * (int64_t)p + 1] = N1; xnp[3 * (int64_t)p + 2] = N2;
        acc[1] += sums[0]; acc[2] += sums[1]; acc[3] += sums[2];
        acc[4] += X0 * X0 + X1 * X1 + X2 * X2;
      }
    }
  }
#pragma unroll
  for (int k = 0; k < 5; ++k) acc[k] = warp_sum(acc[k]);
  if (lane == 0)
#pragma unroll
    for (int k = 0; k < 5; ++k) sred[wid * 5 + k] = acc[k];
  __syncthreads();
  if (threadIdx.x < 5) {
    double s = 0.0;
    for (int w = 0; w < LIN_WARPS; ++w) s += sred[w * 5 + threadIdx.x];
    part[blockIdx.x * 8 + threadIdx.x] = s;
  }
}
