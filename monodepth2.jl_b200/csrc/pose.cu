// pose.cu -- so3 / Pose entry points (P2-P4): hat, hat pullback, so3_exp_map, composeT.
// Reference: src/utils.jl:106-145, 185-192.  One thread per pose, double arithmetic.
#include "pose_math.cuh"

namespace {

__global__ void hat_kernel(const float* __restrict__ rvec, int N, float* __restrict__ S) {
  int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  double r[3] = {rvec[3 * n], rvec[3 * n + 1], rvec[3 * n + 2]}, M[9];
  hat_d(r, M);
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) S[9 * n + 3 * j + i] = (float)M[3 * i + j];  // column-major
}

__global__ void hat_pullback_kernel(const float* __restrict__ d, int N, float* __restrict__ g) {
  int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const float* D = d + 9 * n;  // D[j*3+i] = d[i][j]
  auto at = [&](int i, int j) { return D[3 * j + i]; };
  g[3 * n + 0] = at(2, 1) - at(1, 2);
  g[3 * n + 1] = -at(2, 0) + at(0, 2);
  g[3 * n + 2] = at(1, 0) - at(0, 1);
}

__global__ void compose_kernel(const float* __restrict__ rvec, const float* __restrict__ tvec,
                               int invert, int N, float* __restrict__ R, float* __restrict__ t) {
  int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  double r[3] = {rvec[3 * n], rvec[3 * n + 1], rvec[3 * n + 2]};
  double tt[3] = {0, 0, 0};
  if (tvec) { tt[0] = tvec[3 * n]; tt[1] = tvec[3 * n + 1]; tt[2] = tvec[3 * n + 2]; }
  double Rp[9], tp[3];
  So3Aux a;
  compose_T_d(r, tt, invert != 0, Rp, tp, a);
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) R[9 * n + 3 * j + i] = (float)Rp[3 * i + j];
  if (t) for (int k = 0; k < 3; ++k) t[3 * n + k] = (float)tp[k];
}

__global__ void compose_bwd_kernel(const float* __restrict__ rvec, const float* __restrict__ tvec,
                                   int invert, int N, const float* __restrict__ dR,
                                   const float* __restrict__ dt, float* __restrict__ d_rvec,
                                   float* __restrict__ d_tvec) {
  int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  double r[3] = {rvec[3 * n], rvec[3 * n + 1], rvec[3 * n + 2]};
  double tt[3] = {tvec[3 * n], tvec[3 * n + 1], tvec[3 * n + 2]};
  double dRp[9], dtp[3], dr[3], dtt[3];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) dRp[3 * i + j] = dR[9 * n + 3 * j + i];
  for (int k = 0; k < 3; ++k) dtp[k] = dt ? (double)dt[3 * n + k] : 0.0;
  compose_T_bwd_d(r, tt, invert != 0, dRp, dtp, dr, dtt);
  for (int k = 0; k < 3; ++k) {
    d_rvec[3 * n + k] = (float)dr[k];
    if (d_tvec) d_tvec[3 * n + k] = (float)dtt[k];
  }
}

inline int blocks(int n, int b) { return (n + b - 1) / b; }

}  // namespace

extern "C" {

int32_t mdb_hat(mdb_handle* h, const float* rvec, int32_t N, float* S, void* stream) {
  MDB_REQUIRE(h, h && rvec && S && N > 0, MDB_E_INVALID_ARG, "mdb_hat: bad argument");
  hat_kernel<<<blocks(N, 128), 128, 0, (cudaStream_t)stream>>>(rvec, N, S);
  MDB_LAUNCHED(h);
  return MDB_OK;
}

int32_t mdb_hat_pullback(mdb_handle* h, const float* d, int32_t N, float* d_rvec, void* stream) {
  MDB_REQUIRE(h, h && d && d_rvec && N > 0, MDB_E_INVALID_ARG, "mdb_hat_pullback: bad argument");
  hat_pullback_kernel<<<blocks(N, 128), 128, 0, (cudaStream_t)stream>>>(d, N, d_rvec);
  MDB_LAUNCHED(h);
  return MDB_OK;
}

int32_t mdb_so3_exp_map(mdb_handle* h, const float* rvec, int32_t N, float* R, void* stream) {
  MDB_REQUIRE(h, h && rvec && R && N > 0, MDB_E_INVALID_ARG, "mdb_so3_exp_map: bad argument");
  compose_kernel<<<blocks(N, 128), 128, 0, (cudaStream_t)stream>>>(rvec, nullptr, 0, N, R, nullptr);
  MDB_LAUNCHED(h);
  return MDB_OK;
}

int32_t mdb_compose_T(mdb_handle* h, const float* rvec, const float* tvec, int32_t invert,
                      int32_t N, float* R, float* t, void* stream) {
  MDB_REQUIRE(h, h && rvec && tvec && R && t && N > 0, MDB_E_INVALID_ARG, "mdb_compose_T: bad argument");
  compose_kernel<<<blocks(N, 128), 128, 0, (cudaStream_t)stream>>>(rvec, tvec, invert, N, R, t);
  MDB_LAUNCHED(h);
  return MDB_OK;
}

int32_t mdb_compose_T_bwd(mdb_handle* h, const float* rvec, const float* tvec, int32_t invert,
                          int32_t N, const float* dR, const float* dt, float* d_rvec,
                          float* d_tvec, void* stream) {
  MDB_REQUIRE(h, h && rvec && tvec && dR && d_rvec && N > 0, MDB_E_INVALID_ARG,
              "mdb_compose_T_bwd: bad argument");
  compose_bwd_kernel<<<blocks(N, 128), 128, 0, (cudaStream_t)stream>>>(rvec, tvec, invert, N, dR, dt,
                                                                        d_rvec, d_tvec);
  MDB_LAUNCHED(h);
  return MDB_OK;
}

}  // extern "C"
