// Per-camera and per-observation FP64 math of the bundle-adjustment hot path.
//
// Camera model of the reference (sfm_lite/bundle_adjustment.py:25-30):
//     (u,v) = pi( K * R(w) * (X - T) ),   residual = (u,v) - observed      (:41)
// with w a rotation vector (scipy Rotation.from_rotvec, :25), T used as a camera
// CENTRE, K the full 3x3 intrinsic matrix.  The analytic blocks replace the
// 2-point finite differences scipy takes at sfm_lite/sfm.py:266-268:
//     Y = X - T, Z = R Y, p = K Z, A = [[1,0,-u],[0,1,-v]]/p2, B = A K
//     J_X = B R (2x3),  J_T = -J_X,  J_w = -J_X [Y]x J_r(w),
//     J_r(w) = I - a [w]x + b [w]x^2,  a = (1-cos t)/t^2,  b = (t - sin t)/t^3.
//
// Everything is __host__ __device__ so tests/hostmath can run the very same code
// on the CPU against the numpy oracle before a GPU is involved.
#pragma once
#include <math.h>

#ifdef __CUDACC__
#define BA_HD __host__ __device__ __forceinline__
#else
#define BA_HD inline
#endif

namespace b200ba {

constexpr int CAMPREP_STRIDE = 24;  // scratch size of cam_prep's output: R[9] Jr[9] T[3] pad[3]
constexpr int CAMPREP_FIELDS = 21;  // stored fields per camera (device layout is field-major [21][C])
constexpr double SMALL_ANGLE = 1e-4;

struct Intrinsics {
  double k[9];  // row-major 3x3
};

// w,T (6 doubles, layout of sfm.py:255-257) -> R, J_r, T
BA_HD void cam_prep(const double* __restrict__ cam6, double* __restrict__ out) {
  const double wx = cam6[0], wy = cam6[1], wz = cam6[2];
  const double t2 = wx * wx + wy * wy + wz * wz;
  const double t = sqrt(t2);
  double s, a, b;
  if (t < SMALL_ANGLE) {
    // series limits: theta = 0 is on the main path (first registered camera, sfm.py:176)
    s = 1.0 - t2 * (1.0 / 6.0) + t2 * t2 * (1.0 / 120.0);
    a = 0.5 - t2 * (1.0 / 24.0) + t2 * t2 * (1.0 / 720.0);
    b = (1.0 / 6.0) - t2 * (1.0 / 120.0) + t2 * t2 * (1.0 / 5040.0);
  } else {
    const double sn = sin(t);
    const double sh = sin(0.5 * t);
    s = sn / t;
    a = 2.0 * sh * sh / t2;  // (1-cos t)/t^2 without cancellation
    b = (t - sn) / (t2 * t);
  }
  // [w]x^2 = w w^T - t2 I
  const double xx = wx * wx - t2, yy = wy * wy - t2, zz = wz * wz - t2;
  const double xy = wx * wy, xz = wx * wz, yz = wy * wz;
  double* R = out;
  double* J = out + 9;
  R[0] = 1.0 + a * xx;      R[1] = -s * wz + a * xy;  R[2] = s * wy + a * xz;
  R[3] = s * wz + a * xy;   R[4] = 1.0 + a * yy;      R[5] = -s * wx + a * yz;
  R[6] = -s * wy + a * xz;  R[7] = s * wx + a * yz;   R[8] = 1.0 + a * zz;
  J[0] = 1.0 + b * xx;      J[1] = a * wz + b * xy;   J[2] = -a * wy + b * xz;
  J[3] = -a * wz + b * xy;  J[4] = 1.0 + b * yy;      J[5] = a * wx + b * yz;
  J[6] = a * wy + b * xz;   J[7] = -a * wx + b * yz;  J[8] = 1.0 + b * zz;
  out[18] = cam6[3]; out[19] = cam6[4]; out[20] = cam6[5];
  out[21] = 0.0; out[22] = 0.0; out[23] = 0.0;
}

// residual only
BA_HD void project_residual(const double* __restrict__ cp, const Intrinsics& K, double X0, double X1, double X2,
                            double uo, double vo, double& r0, double& r1) {
  const double y0 = X0 - cp[18], y1 = X1 - cp[19], y2 = X2 - cp[20];
  const double z0 = cp[0] * y0 + cp[1] * y1 + cp[2] * y2;
  const double z1 = cp[3] * y0 + cp[4] * y1 + cp[5] * y2;
  const double z2 = cp[6] * y0 + cp[7] * y1 + cp[8] * y2;
  const double p0 = K.k[0] * z0 + K.k[1] * z1 + K.k[2] * z2;
  const double p1 = K.k[3] * z0 + K.k[4] * z1 + K.k[5] * z2;
  const double p2 = K.k[6] * z0 + K.k[7] * z1 + K.k[8] * z2;
  r0 = p0 / p2 - uo;
  r1 = p1 / p2 - vo;
}

// residual + Jc (2x6 row-major: [d/dw | d/dT]) + Jp (2x3 row-major: d/dX)
BA_HD void project_jac(const double* __restrict__ cp, const Intrinsics& K, double X0, double X1, double X2,
                       double uo, double vo, double& r0, double& r1, double* __restrict__ Jc,
                       double* __restrict__ Jp) {
  const double* R = cp;
  const double* Jr = cp + 9;
  const double y0 = X0 - cp[18], y1 = X1 - cp[19], y2 = X2 - cp[20];
  const double z0 = R[0] * y0 + R[1] * y1 + R[2] * y2;
  const double z1 = R[3] * y0 + R[4] * y1 + R[5] * y2;
  const double z2 = R[6] * y0 + R[7] * y1 + R[8] * y2;
  const double p0 = K.k[0] * z0 + K.k[1] * z1 + K.k[2] * z2;
  const double p1 = K.k[3] * z0 + K.k[4] * z1 + K.k[5] * z2;
  const double p2 = K.k[6] * z0 + K.k[7] * z1 + K.k[8] * z2;
  const double pu = p0 / p2, pv = p1 / p2;  // same quotient as project_residual -> bit-identical residuals
  const double ip2 = 1.0 / p2;
  r0 = pu - uo;
  r1 = pv - vo;
#pragma unroll
  for (int row = 0; row < 2; ++row) {
    const double q = row == 0 ? pu : pv;
    const double b0 = (K.k[3 * row + 0] - q * K.k[6]) * ip2;
    const double b1 = (K.k[3 * row + 1] - q * K.k[7]) * ip2;
    const double b2 = (K.k[3 * row + 2] - q * K.k[8]) * ip2;
    const double a0 = b0 * R[0] + b1 * R[3] + b2 * R[6];
    const double a1 = b0 * R[1] + b1 * R[4] + b2 * R[7];
    const double a2 = b0 * R[2] + b1 * R[5] + b2 * R[8];
    Jp[3 * row + 0] = a0; Jp[3 * row + 1] = a1; Jp[3 * row + 2] = a2;
    // a [Y]x = a x Y
    const double d0 = a1 * y2 - a2 * y1;
    const double d1 = a2 * y0 - a0 * y2;
    const double d2 = a0 * y1 - a1 * y0;
    Jc[6 * row + 0] = -(d0 * Jr[0] + d1 * Jr[3] + d2 * Jr[6]);
    Jc[6 * row + 1] = -(d0 * Jr[1] + d1 * Jr[4] + d2 * Jr[7]);
    Jc[6 * row + 2] = -(d0 * Jr[2] + d1 * Jr[5] + d2 * Jr[8]);
    Jc[6 * row + 3] = -a0; Jc[6 * row + 4] = -a1; Jc[6 * row + 5] = -a2;
  }
}


// Robust loss of scipy.optimize.least_squares (the `loss=` / `f_scale=` keywords of the call at sfm.py:266; the reference
// leaves the default 'linear').  scipy applies rho to every residual COMPONENT: z = (f / C)^2,
//   cost = 0.5 C^2 sum rho(z),   J_scale = max(rho' + 2 rho'' z, 0.05 rho')  [scipy: EPS; see robust_component],   J <- sqrt(J_scale) J,   f <- rho' f / sqrt(J_scale)
// (scipy _lsq/least_squares.py construct_loss_function, _lsq/common.py scale_for_robust_loss_function).
struct Loss {
  int kind;       // 0 linear, 1 soft_l1, 2 huber, 3 cauchy, 4 arctan
  double c2;      // f_scale^2
  double inv_c2;
};

// one residual component: returns C^2 rho(z), scales r in place, js = sqrt(J_scale) (row factor of the Jacobian)
BA_HD double robust_component(const Loss& L, double& r, double& js) {
  const double z = r * r * L.inv_c2;
  double rho0, rho1, rho2;
  if (L.kind == 1) {            // soft_l1
    const double t = 1.0 + z, st = sqrt(t);
    rho0 = 2.0 * (st - 1.0); rho1 = 1.0 / st; rho2 = -0.5 / (t * st);
  } else if (L.kind == 2) {     // huber
    if (z <= 1.0) { rho0 = z; rho1 = 1.0; rho2 = 0.0; }
    else { const double sz = sqrt(z); rho0 = 2.0 * sz - 1.0; rho1 = 1.0 / sz; rho2 = -0.5 / (z * sz); }
  } else if (L.kind == 3) {     // cauchy
    const double t = 1.0 + z;
    rho0 = log1p(z); rho1 = 1.0 / t; rho2 = -1.0 / (t * t);
  } else {                      // arctan
    const double t = 1.0 + z * z;
    rho0 = atan(z); rho1 = 1.0 / t; rho2 = -2.0 * z / (t * t);
  }
  // scipy clamps the second-order weight at machine epsilon; here it never falls below 5 % of the first-order (IRLS)
  // weight: rows that cancel to ~0 (every Huber outlier) would leave Marquardt-damped point blocks singular and stall
  // the LM loop.  The gradient J_s^T f_s = J^T rho' f stays exact, so the minimiser is the same (oracle: ROBUST_CLAMP).
  double w = fmax(rho1 + 2.0 * rho2 * z, 0.05 * rho1);
  if (w < 2.220446049250313e-16) w = 2.220446049250313e-16;
  js = sqrt(w);
  r *= rho1 / js;
  return L.c2 * rho0;
}

// cost term only (trial point): C^2 (rho(z0) + rho(z1)); linear: r0^2 + r1^2
BA_HD double robust_cost2(const Loss& L, double r0, double r1) {
  if (L.kind == 0) return r0 * r0 + r1 * r1;
  double j;
  return robust_component(L, r0, j) + robust_component(L, r1, j);
}

// scale the residual and the rows of both Jacobian blocks of one observation in place; returns the cost term
BA_HD double apply_loss(const Loss& L, double& r0, double& r1, double* __restrict__ Jc, double* __restrict__ Jp) {
  if (L.kind == 0) return r0 * r0 + r1 * r1;
  double j0, j1;
  const double c = robust_component(L, r0, j0) + robust_component(L, r1, j1);
#pragma unroll
  for (int k = 0; k < 6; ++k) { Jc[k] *= j0; Jc[6 + k] *= j1; }
#pragma unroll
  for (int k = 0; k < 3; ++k) { Jp[k] *= j0; Jp[3 + k] *= j1; }
  return c;
}

// residual + Jacobian blocks of one observation with the loss applied; cost2 = the observation's term of 2 * cost
BA_HD void project_jac_loss(const double* __restrict__ cp, const Intrinsics& K, const Loss& L, double X0, double X1, double X2,
                            double uo, double vo, double& r0, double& r1, double* __restrict__ Jc, double* __restrict__ Jp,
                            double& cost2) {
  project_jac(cp, K, X0, X1, X2, uo, vo, r0, r1, Jc, Jp);
  cost2 = apply_loss(L, r0, r1, Jc, Jp);
}

// sum-term of the back-substitution for one observation: s3 = J_X^T (Jc dc) without forming Jc.
// Jc dc = J_w dw + J_T dT = -J_X ([Y]x J_r dw + dT) = -J_X q,  q = Y x m + dT  with m = J_r(w) dw per CAMERA
// (cs = {m, dT}, k_cam_prep); per observation only J_X = B R is needed (2x3).
// With a robust loss the rows carry the weights J_scale of the residual at x (uo, vo needed); linear: weights 1.
BA_HD void project_step(const double* __restrict__ R, const double* __restrict__ T, const double* __restrict__ cs,
                        const Intrinsics& K, const Loss& L, double X0, double X1, double X2, double uo, double vo,
                        double* __restrict__ s3) {
  const double y0 = X0 - T[0], y1 = X1 - T[1], y2 = X2 - T[2];
  const double z0 = R[0] * y0 + R[1] * y1 + R[2] * y2;
  const double z1 = R[3] * y0 + R[4] * y1 + R[5] * y2;
  const double z2 = R[6] * y0 + R[7] * y1 + R[8] * y2;
  const double p0 = K.k[0] * z0 + K.k[1] * z1 + K.k[2] * z2;
  const double p1 = K.k[3] * z0 + K.k[4] * z1 + K.k[5] * z2;
  const double p2 = K.k[6] * z0 + K.k[7] * z1 + K.k[8] * z2;
  const double ip2 = 1.0 / p2;
  const double pu = p0 * ip2, pv = p1 * ip2;
  const double q0 = y1 * cs[2] - y2 * cs[1] + cs[3];
  const double q1 = y2 * cs[0] - y0 * cs[2] + cs[4];
  const double q2 = y0 * cs[1] - y1 * cs[0] + cs[5];
  s3[0] = 0.0; s3[1] = 0.0; s3[2] = 0.0;
#pragma unroll
  for (int row = 0; row < 2; ++row) {
    const double q = row == 0 ? pu : pv;
    const double b0 = (K.k[3 * row + 0] - q * K.k[6]) * ip2;
    const double b1 = (K.k[3 * row + 1] - q * K.k[7]) * ip2;
    const double b2 = (K.k[3 * row + 2] - q * K.k[8]) * ip2;
    const double a0 = b0 * R[0] + b1 * R[3] + b2 * R[6];
    const double a1 = b0 * R[1] + b1 * R[4] + b2 * R[7];
    const double a2 = b0 * R[2] + b1 * R[5] + b2 * R[8];
    double t = -(a0 * q0 + a1 * q1 + a2 * q2);
    if (L.kind != 0) {
      double rr = q - (row == 0 ? uo : vo), js;
      robust_component(L, rr, js);
      t *= js * js;
    }
    s3[0] = fma(a0, t, s3[0]); s3[1] = fma(a1, t, s3[1]); s3[2] = fma(a2, t, s3[2]);
  }
}

// Damped 3x3 point block: V* = V + lambda*diag(V);  V* = L L^T;  returns Linv (lower, 6 entries:
// i00 i10 i11 i20 i21 i22).  A non-positive pivot (rank-deficient point, e.g. a single observation
// with lambda = 0) freezes the point: Linv = 0 => no Schur contribution, delta_p = 0.
// 1/sqrt(x): the device intrinsic is one MUFU + two Newton steps instead of a full sqrt followed by a full division
// (three of these sit on the dependent chain of every point's 3x3 Cholesky).
BA_HD double inv_sqrt(double x) {
#ifdef __CUDA_ARCH__
  return rsqrt(x);
#else
  return 1.0 / sqrt(x);
#endif
}

BA_HD bool point_factor(const double* __restrict__ V /*v00 v10 v11 v20 v21 v22*/, double lambda,
                        double* __restrict__ Li) {
  const double d = 1.0 + lambda;
  const double v00 = V[0] * d, v10 = V[1], v11 = V[2] * d, v20 = V[3], v21 = V[4], v22 = V[5] * d;
  bool ok = v00 > 0.0;
  const double i00 = ok ? inv_sqrt(v00) : 0.0;
  const double l10 = v10 * i00, l20 = v20 * i00;
  const double s11 = v11 - l10 * l10;
  ok = ok && (s11 > 0.0);
  const double i11 = ok ? inv_sqrt(s11) : 0.0;
  const double l21 = (v21 - l20 * l10) * i11;
  const double s22 = v22 - l20 * l20 - l21 * l21;
  ok = ok && (s22 > 0.0) && (s22 == s22) && (s22 < 1e300);
  if (!ok) {
    Li[0] = Li[1] = Li[2] = Li[3] = Li[4] = Li[5] = 0.0;
    return false;
  }
  const double i22 = inv_sqrt(s22);
  const double i10 = -i11 * l10 * i00;
  const double i21 = -i22 * l21 * i11;
  const double i20 = -(i21 * l10 + i22 * l20) * i00;
  Li[0] = i00; Li[1] = i10; Li[2] = i11; Li[3] = i20; Li[4] = i21; Li[5] = i22;
  return true;
}

}  // namespace b200ba
