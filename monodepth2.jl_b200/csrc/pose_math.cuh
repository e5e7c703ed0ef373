// pose_math.cuh -- so3 exp-map / composeT forward and pullback in double precision.
// Restates src/utils.jl:106-145 (so3_exp_map, hat, hat rrule) and :185-192 (composeT).
// Double is used on purpose: these run on 6 numbers per pose, and `1 - cos(theta)` at
// theta ~ 1e-2 would lose three digits in float.
// Matrices here are math-natural row-major M[3*i + j]; Julia's (3,3) column-major storage
// is the transpose (element (i,j) at j*3+i).
#pragma once
#include "mdb_common.cuh"

#ifdef __CUDACC__
struct So3Aux {
  double theta, theta_inv, f1, f2;
  double s, c, omc;   // sin(theta), cos(theta), 1 - cos(theta)
  double S[9], S2[9];
};

MDB_HD void hat_d(const double r[3], double S[9]) {
  // Julia (1-based): S[2,1]=r3 S[1,2]=-r3 S[3,1]=-r2 S[1,3]=r2 S[3,2]=r1 S[2,3]=-r1
  S[0] = 0;     S[1] = -r[2]; S[2] = r[1];
  S[3] = r[2];  S[4] = 0;     S[5] = -r[0];
  S[6] = -r[1]; S[7] = r[0];  S[8] = 0;
}

MDB_HD void mat3_mul(const double* A, const double* B, double* C) {
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      double s = 0;
      for (int k = 0; k < 3; ++k) s += A[3 * i + k] * B[3 * k + j];
      C[3 * i + j] = s;
    }
}

MDB_HD void mat3_T(const double* A, double* At) {
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) At[3 * j + i] = A[3 * i + j];
}

MDB_HD void so3_exp_d(const double r[3], double R[9], So3Aux& a) {
  hat_d(r, a.S);
  mat3_mul(a.S, a.S, a.S2);
  a.theta = sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
  a.theta_inv = 1.0 / fmax(a.theta, 1e-4);
  // ONE double-precision sincos, at the half angle, serves the forward and the pullback (the pose chains are
  // single-thread latency: five separate sin / cos calls were most of it): sin t = 2 sh ch,
  // 1 - cos t = 2 sh^2 (no cancellation), cos t = 1 - 2 sh^2
  double sh, ch;
  sincos(0.5 * a.theta, &sh, &ch);
  a.s = 2.0 * sh * ch;
  a.omc = 2.0 * sh * sh;
  a.c = 1.0 - a.omc;
  a.f1 = a.theta_inv * a.s;
  a.f2 = a.theta_inv * a.theta_inv * a.omc;
  for (int k = 0; k < 9; ++k) R[k] = a.f1 * a.S[k] + a.f2 * a.S2[k];
  R[0] += 1.0; R[4] += 1.0; R[8] += 1.0;
}

// pullback of so3_exp_map: dR (math row-major) -> dr.  At theta == 0 the reference's
// sqrt pullback yields NaN (README.md:47-51); r/theta reproduces that.
MDB_HD void so3_exp_bwd_d(const double r[3], const So3Aux& a, const double dR[9], double dr[3]) {
  double df1 = 0, df2 = 0;
  for (int k = 0; k < 9; ++k) { df1 += dR[k] * a.S[k]; df2 += dR[k] * a.S2[k]; }
  // dS = f1 dR + f2 (dR S^T + S^T dR)
  double St[9], t1[9], t2[9], dS[9];
  mat3_T(a.S, St);
  mat3_mul(dR, St, t1);
  mat3_mul(St, dR, t2);
  for (int k = 0; k < 9; ++k) dS[k] = a.f1 * dR[k] + a.f2 * (t1[k] + t2[k]);
  // hat pullback (src/utils.jl:139-141), math indices d[i][j]
  dr[0] = dS[3 * 2 + 1] - dS[3 * 1 + 2];
  dr[1] = -dS[3 * 2 + 0] + dS[3 * 0 + 2];
  dr[2] = dS[3 * 1 + 0] - dS[3 * 0 + 1];
  // theta path
  const double dtinv_dtheta = (a.theta > 1e-4) ? -a.theta_inv * a.theta_inv : 0.0;
  const double s = a.s, c = a.c, omc = a.omc;
  const double df1_dtheta = a.theta_inv * c + s * dtinv_dtheta;
  const double df2_dtheta = a.theta_inv * a.theta_inv * s + 2.0 * a.theta_inv * omc * dtinv_dtheta;
  const double dtheta = df1 * df1_dtheta + df2 * df2_dtheta;
  for (int k = 0; k < 3; ++k) dr[k] += dtheta * (r[k] / a.theta);
}

// composeT: R' = R or R^T ; t' = t or R^T (-t)
MDB_HD void compose_T_d(const double r[3], const double t[3], bool invert, double Rp[9],
                        double tp[3], So3Aux& a) {
  double R[9];
  so3_exp_d(r, R, a);
  if (!invert) {
    for (int k = 0; k < 9; ++k) Rp[k] = R[k];
    for (int k = 0; k < 3; ++k) tp[k] = t[k];
  } else {
    mat3_T(R, Rp);
    for (int i = 0; i < 3; ++i) tp[i] = -(Rp[3 * i] * t[0] + Rp[3 * i + 1] * t[1] + Rp[3 * i + 2] * t[2]);
  }
}

// pullback of composeT: (dR', dt') -> (dr, dt)
MDB_HD void compose_T_bwd_d(const double r[3], const double t[3], bool invert, const double dRp[9],
                            const double dtp[3], double dr[3], double dt[3]) {
  So3Aux a;
  double R[9];
  so3_exp_d(r, R, a);
  double dR[9];
  if (!invert) {
    for (int k = 0; k < 9; ++k) dR[k] = dRp[k];
    for (int k = 0; k < 3; ++k) dt[k] = dtp[k];
  } else {
    // R'_{ij} = R_{ji};  t'_i = -sum_j R_{ji} t_j
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) dR[3 * j + i] = dRp[3 * i + j] - dtp[i] * t[j];
    for (int j = 0; j < 3; ++j)
      dt[j] = -(R[3 * j] * dtp[0] + R[3 * j + 1] * dtp[1] + R[3 * j + 2] * dtp[2]);
  }
  so3_exp_bwd_d(r, a, dR, dr);
}
#endif
