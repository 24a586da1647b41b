// Perspective-n-point for the two cameras of a new edge (included inside ba_kernels.cu's anonymous namespace).
//
// Reference: SFM._apply_increment calls cv2.solvePnP(pt3ds, pt2ds, K, zeros) twice per edge (sfm.py:207-208) and reads
// the result as H = [Rodrigues(rvec) | tvec] (transforms.py:18-23, the R X + t convention); the reference's own
// restatement is cv2_lite/solve_pnp.py:17-66: a DLT on normalised image points (12-column system, smallest right
// singular vector, rotation part orthogonalised, sign fixed by the determinant) followed by a non-linear least-squares
// refinement of (rvec, tvec) on the reprojection error (reproj_error, solve_pnp.py:9-14).
//
// One CTA per problem.  Stage 1 (DLT): the 12 x 12 normal matrix A^T A is accumulated by 78 threads and diagonalised by
// cyclic Jacobi rotations in shared memory; the eigenvector of the smallest eigenvalue gives [R | t] up to scale, R is
// replaced by its orthogonal polar factor (Newton iteration), t is divided by the scale that was removed.  Stage 2:
// Levenberg-Marquardt on the six parameters with analytic 2 x 6 blocks (d(R X)/dw = -R [X]x J_r(w)), block-wide
// fixed-order reductions of J^T J / J^T r, 6 x 6 Cholesky by thread 0.  The DLT only has to land in the basin of the
// minimum; the minimum is what cv2 / the reference return.

constexpr int PNP_T = 256;

__device__ __forceinline__ double pnp_block_sum(double v, double* sm) {
  // fixed-order sum over the CTA; every thread returns the total
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) sm[wid] = v;
  __syncthreads();
  double s = 0.0;
#pragma unroll
  for (int w = 0; w < PNP_T / 32; ++w) s += sm[w];
  return s;
}

// residual (2) and Jacobian rows (2 x 6, [d/dw | d/dt]) of one point; cp = cam_prep output for (w, *) : R, J_r
__device__ __forceinline__ void pnp_point(const double* __restrict__ cp, const double* __restrict__ t, const double* __restrict__ K,
                                          const double* __restrict__ X, double uo, double vo, double* r, double* J) {
  const double* R = cp;
  const double* Jr = cp + 9;
  const double y0 = R[0] * X[0] + R[1] * X[1] + R[2] * X[2];   // R X
  const double y1 = R[3] * X[0] + R[4] * X[1] + R[5] * X[2];
  const double y2 = R[6] * X[0] + R[7] * X[1] + R[8] * X[2];
  const double z0 = y0 + t[0], z1 = y1 + t[1], z2 = y2 + t[2];
  const double p0 = K[0] * z0 + K[1] * z1 + K[2] * z2;
  const double p1 = K[3] * z0 + K[4] * z1 + K[5] * z2;
  const double p2 = K[6] * z0 + K[7] * z1 + K[8] * z2;
  const double pu = p0 / p2, pv = p1 / p2, ip2 = 1.0 / p2;
  r[0] = pu - uo;
  r[1] = pv - vo;
  if (!J) return;
#pragma unroll
  for (int row = 0; row < 2; ++row) {
    const double q = row == 0 ? pu : pv;
    const double b0 = (K[3 * row + 0] - q * K[6]) * ip2;   // B = d(u,v)/dZ
    const double b1 = (K[3 * row + 1] - q * K[7]) * ip2;
    const double b2 = (K[3 * row + 2] - q * K[8]) * ip2;
    // d(R X)/dw = -R [X]x J_r = -[R X]x R J_r  =>  row: -(b x ... ) ; use  b^T (-[y]x) R J_r = (y x b)^T ... written out:
    // b^T [y]x = (b x y)^T, so  -b^T [y]x (R J_r) = -(b x y)^T R J_r
    const double c0 = b1 * y2 - b2 * y1, c1 = b2 * y0 - b0 * y2, c2 = b0 * y1 - b1 * y0;   // b x y
    // (b x y)^T R
    const double d0 = c0 * R[0] + c1 * R[3] + c2 * R[6];
    const double d1 = c0 * R[1] + c1 * R[4] + c2 * R[7];
    const double d2 = c0 * R[2] + c1 * R[5] + c2 * R[8];
    J[6 * row + 0] = -(d0 * Jr[0] + d1 * Jr[3] + d2 * Jr[6]);
    J[6 * row + 1] = -(d0 * Jr[1] + d1 * Jr[4] + d2 * Jr[7]);
    J[6 * row + 2] = -(d0 * Jr[2] + d1 * Jr[5] + d2 * Jr[8]);
    J[6 * row + 3] = b0; J[6 * row + 4] = b1; J[6 * row + 5] = b2;
  }
}

// rotation matrix -> rotation vector (Markley quaternion as in k_pack_cameras, canonical sign)
__device__ void pnp_rotvec(const double* R, double* w) {
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
  for (int k = 0; k < 4; ++k) q[k] /= qn;
  if (q[3] < 0.0) for (int k = 0; k < 4; ++k) q[k] = -q[k];
  const double an = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2]);
  const double angle = 2.0 * atan2(an, q[3]);
  const double a2 = angle * angle;
  const double scale = (angle <= 1e-3) ? 2.0 + a2 / 12.0 + 7.0 * a2 * a2 / 2880.0 : angle / sin(angle / 2.0);
  w[0] = scale * q[0]; w[1] = scale * q[1]; w[2] = scale * q[2];
}

struct PnpProblem {
  const double* X;    // [n][3]
  const double* uv;   // [n][2]
  int n;
};

// out per problem: rvec(3) tvec(3) cost iterations status   (status 1 = converged, 0 = iteration limit, -1 = failure)
__global__ void __launch_bounds__(PNP_T) k_pnp(const PnpProblem* __restrict__ probs, const double* __restrict__ Kin,
                                               const double* __restrict__ guess /*nullable [6] per problem*/, int max_iter,
                                               double* __restrict__ out) {
  __shared__ double M[12][12], V[12][12];
  __shared__ double sm[PNP_T / 32];
  __shared__ double K[9], Ki[9], par[6], trial[6], cp[CAMPREP_STRIDE], red[28], step_ok;
  const PnpProblem pb = probs[blockIdx.x];
  const int tid = threadIdx.x, n = pb.n;
  if (tid < 9) K[tid] = Kin[tid];
  __syncthreads();
  if (tid == 0) {   // inverse of K (general 3 x 3)
    const double c00 = K[4] * K[8] - K[5] * K[7], c01 = K[5] * K[6] - K[3] * K[8], c02 = K[3] * K[7] - K[4] * K[6];
    const double id = 1.0 / (K[0] * c00 + K[1] * c01 + K[2] * c02);
    Ki[0] = c00 * id; Ki[1] = (K[2] * K[7] - K[1] * K[8]) * id; Ki[2] = (K[1] * K[5] - K[2] * K[4]) * id;
    Ki[3] = c01 * id; Ki[4] = (K[0] * K[8] - K[2] * K[6]) * id; Ki[5] = (K[2] * K[3] - K[0] * K[5]) * id;
    Ki[6] = c02 * id; Ki[7] = (K[1] * K[6] - K[0] * K[7]) * id; Ki[8] = (K[0] * K[4] - K[1] * K[3]) * id;
  }
  __syncthreads();
  if (guess) {
    if (tid < 6) par[tid] = guess[6 * blockIdx.x + tid];
  } else {
    // ---- stage 1: DLT (solve_pnp.py:17-42).  Thread (a, b), a <= b, sums entry (a, b) of A^T A over the points.
    if (tid < 78) {
      int a = 0, b = tid;
      while (b >= 12 - a) { b -= 12 - a; ++a; }
      b += a;
      double s = 0.0;
      for (int i = 0; i < n; ++i) {
        const double X = pb.X[3 * i], Y = pb.X[3 * i + 1], Z = pb.X[3 * i + 2];
        const double x = pb.uv[2 * i], y = pb.uv[2 * i + 1];
        const double w = Ki[6] * x + Ki[7] * y + Ki[8];
        const double u = (Ki[0] * x + Ki[1] * y + Ki[2]) / w, v = (Ki[3] * x + Ki[4] * y + Ki[5]) / w;
        const double r1[12] = {X, Y, Z, 1.0, 0.0, 0.0, 0.0, 0.0, -u * X, -u * Y, -u * Z, -u};
        const double r2[12] = {0.0, 0.0, 0.0, 0.0, X, Y, Z, 1.0, -v * X, -v * Y, -v * Z, -v};
        s += r1[a] * r1[b] + r2[a] * r2[b];
      }
      M[a][b] = s; M[b][a] = s;
    }
    if (tid < 144) V[tid / 12][tid % 12] = (tid / 12 == tid % 12) ? 1.0 : 0.0;
    __syncthreads();
    if (tid == 0) {
      // cyclic Jacobi eigen-decomposition of the symmetric 12 x 12 matrix
      for (int sweep = 0; sweep < 30; ++sweep) {
        double off = 0.0, dia = 0.0;
        for (int i = 0; i < 12; ++i) { dia += M[i][i] * M[i][i]; for (int j = i + 1; j < 12; ++j) off += M[i][j] * M[i][j]; }
        if (off <= 1e-30 * dia) break;
        for (int p = 0; p < 11; ++p)
          for (int q = p + 1; q < 12; ++q) {
            const double apq = M[p][q];
            if (fabs(apq) < 1e-300) continue;
            const double th = (M[q][q] - M[p][p]) / (2.0 * apq);
            const double t = (th >= 0.0 ? 1.0 : -1.0) / (fabs(th) + sqrt(th * th + 1.0));
            const double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
            for (int k = 0; k < 12; ++k) { const double mkp = M[k][p], mkq = M[k][q]; M[k][p] = c * mkp - s * mkq; M[k][q] = s * mkp + c * mkq; }
            for (int k = 0; k < 12; ++k) { const double mpk = M[p][k], mqk = M[q][k]; M[p][k] = c * mpk - s * mqk; M[q][k] = s * mpk + c * mqk; }
            for (int k = 0; k < 12; ++k) { const double vkp = V[k][p], vkq = V[k][q]; V[k][p] = c * vkp - s * vkq; V[k][q] = s * vkp + c * vkq; }
          }
      }
      int best = 0;
      for (int k = 1; k < 12; ++k) if (M[k][k] < M[best][best]) best = k;
      double Rt[9], tt[3];
      for (int r = 0; r < 3; ++r) { for (int c = 0; c < 3; ++c) Rt[3 * r + c] = V[4 * r + c][best]; tt[r] = V[4 * r + 3][best]; }
      double det = Rt[0] * (Rt[4] * Rt[8] - Rt[5] * Rt[7]) - Rt[1] * (Rt[3] * Rt[8] - Rt[5] * Rt[6]) +
                   Rt[2] * (Rt[3] * Rt[7] - Rt[4] * Rt[6]);
      if (det < 0.0) { for (int k = 0; k < 9; ++k) Rt[k] = -Rt[k]; for (int k = 0; k < 3; ++k) tt[k] = -tt[k]; det = -det; }
      const double scale = cbrt(fmax(det, 1e-300));   // [R | t] is known up to this factor
      double R[9];
      for (int k = 0; k < 9; ++k) R[k] = Rt[k] / scale;
      polar3(R);
      pnp_rotvec(R, par);
      for (int k = 0; k < 3; ++k) par[3 + k] = tt[k] / scale;
    }
  }
  __syncthreads();
  // ---- stage 2: Levenberg-Marquardt on (w, t)
  double lambda = 1e-3, cost = 0.0;
  int it = 0, status = 0;
  for (; it < max_iter; ++it) {
    if (tid == 0) cam_prep(par, cp);   // R, J_r of the rotation part (the translation slots are unused)
    __syncthreads();
    double acc[28];
#pragma unroll
    for (int k = 0; k < 28; ++k) acc[k] = 0.0;
    for (int i = tid; i < n; i += PNP_T) {
      double r[2], J[12];
      pnp_point(cp, par + 3, K, pb.X + 3 * i, pb.uv[2 * i], pb.uv[2 * i + 1], r, J);
      int k = 0;
#pragma unroll
      for (int a = 0; a < 6; ++a)
#pragma unroll
        for (int b = a; b < 6; ++b) acc[k++] += J[a] * J[b] + J[6 + a] * J[6 + b];
#pragma unroll
      for (int a = 0; a < 6; ++a) acc[21 + a] += J[a] * r[0] + J[6 + a] * r[1];
      acc[27] += r[0] * r[0] + r[1] * r[1];
    }
#pragma unroll
    for (int k = 0; k < 28; ++k) {
      const double s = pnp_block_sum(acc[k], sm);
      if (tid == 0) red[k] = s;
    }
    __syncthreads();
    cost = 0.5 * red[27];
    // inner loop: raise lambda until a step lowers the cost
    bool done = false;
    for (int tries = 0; tries < 12 && !done; ++tries) {
      if (tid == 0) {
        double A[6][6], L[6][6], g[6], d[6];
        int k = 0;
        for (int a = 0; a < 6; ++a) for (int b = a; b < 6; ++b) { A[a][b] = red[k]; A[b][a] = red[k]; ++k; }
        for (int a = 0; a < 6; ++a) { A[a][a] *= (1.0 + lambda); g[a] = -red[21 + a]; }
        bool ok = true;
        for (int j = 0; j < 6 && ok; ++j) {
          double dd = A[j][j];
          for (int q = 0; q < j; ++q) dd -= L[j][q] * L[j][q];
          if (!(dd > 0.0)) { ok = false; break; }
          L[j][j] = sqrt(dd);
          for (int i2 = j + 1; i2 < 6; ++i2) {
            double s = A[i2][j];
            for (int q = 0; q < j; ++q) s -= L[i2][q] * L[j][q];
            L[i2][j] = s / L[j][j];
          }
        }
        if (ok) {
          for (int i2 = 0; i2 < 6; ++i2) { double s = g[i2]; for (int q = 0; q < i2; ++q) s -= L[i2][q] * d[q]; d[i2] = s / L[i2][i2]; }
          for (int i2 = 5; i2 >= 0; --i2) { double s = d[i2]; for (int q = i2 + 1; q < 6; ++q) s -= L[q][i2] * d[q]; d[i2] = s / L[i2][i2]; }
          for (int a = 0; a < 6; ++a) trial[a] = par[a] + d[a];
        }
        step_ok = ok ? 1.0 : 0.0;
      }
      __syncthreads();
      if (step_ok == 0.0) { lambda *= 10.0; __syncthreads(); continue; }
      if (tid == 0) cam_prep(trial, cp);
      __syncthreads();
      double c2 = 0.0;
      for (int i = tid; i < n; i += PNP_T) {
        double r[2];
        pnp_point(cp, trial + 3, K, pb.X + 3 * i, pb.uv[2 * i], pb.uv[2 * i + 1], r, nullptr);
        c2 += r[0] * r[0] + r[1] * r[1];
      }
      const double cnew = 0.5 * pnp_block_sum(c2, sm);
      double dn = 0.0, pn = 0.0;
      for (int a = 0; a < 6; ++a) { dn += (trial[a] - par[a]) * (trial[a] - par[a]); pn += par[a] * par[a]; }
      __syncthreads();
      if (cnew < cost || (cnew == cost && dn == 0.0)) {
        const bool small = (cost - cnew) <= 1e-14 * cost || dn <= 1e-24 * (pn + 1e-24);
        if (tid < 6) par[tid] = trial[tid];
        lambda = fmax(lambda * 0.2, 1e-12);
        cost = cnew;
        done = true;
        if (small) status = 1;
      } else {
        lambda *= 10.0;
        if (lambda > 1e12) { status = 1; done = true; }   // no descent direction left: at a minimum to working precision
      }
      __syncthreads();
    }
    if (!done) status = 1;
    if (status) { ++it; break; }
  }
  if (tid < 6) out[9 * blockIdx.x + tid] = par[tid];
  if (tid == 0) {
    out[9 * blockIdx.x + 6] = cost;
    out[9 * blockIdx.x + 7] = (double)it;
    out[9 * blockIdx.x + 8] = (cost == cost) ? (double)status : -1.0;
  }
}
