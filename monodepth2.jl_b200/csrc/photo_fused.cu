// photo_fused.cu -- the fused view-synthesis photometric loss, forward + backward (P5-P12,
// S1-S2 of SURVEY 8a).  Reference: src/training.jl:21-78 (train_loss) and everything it calls
// in src/utils.jl; src/simple_depth.jl:25-41 for the single-scale variant.
//
// Launch sequence of mdb_train_loss_fwdbwd (all on the caller's stream, no host sync):
//   1. pose_prep_kernel        composeT in double -> A = K R' invK, b = K t' per (item, source)
//   2. disp_stats_kernel       per (scale, item): sum of the up-sampled disparity and the two
//                              edge-weighted |gradient| sums of smooth_loss      [HBM-bound]
//   3. disp_stats_finalize     -> mean, smoothness loss and the mean-normalisation constant
//   4. photo_fused_kernel      per tile, all scales: backproject -> project -> bilinear sample
//                              -> SSIM/L1 -> min over sources / automask -> pullback to the
//                              up-sampled disparity and to (A, b); block-level reductions
//   5. upsample_bwd_kernel     transpose of upsample_bilinear for the low-res scales (gather)
//   6. finalize_item_kernel    fixed-order sum of the tile partials, pose chain back to
//                              (rvec, tvec) in double
//   7. finalize_loss_kernel    scalar loss
// No float atomics anywhere: results are bit-reproducible run to run.
#include "photo_fused.cuh"
#include "pose_math.cuh"

namespace {

// ---------------------------------------------------------------------------------------
// 1. pose_prep
// ---------------------------------------------------------------------------------------
struct PoseArgs {
  const float* rvec[MDB_MAX_SOURCES];
  const float* tvec[MDB_MAX_SOURCES];
  int invert[MDB_MAX_SOURCES];
  double K[9], invK[9];  // math row-major
  int N, S;
};

__global__ void pose_prep_kernel(PoseArgs a, PosePrep* __restrict__ out, PosePrepD* __restrict__ outd) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= a.N * a.S) return;
  const int n = idx / a.S, s = idx % a.S;
  double r[3], t[3], Rp[9], tp[3], KR[9], A[9];
  for (int k = 0; k < 3; ++k) { r[k] = a.rvec[s][3 * n + k]; t[k] = a.tvec[s][3 * n + k]; }
  So3Aux aux;
  compose_T_d(r, t, a.invert[s] != 0, Rp, tp, aux);
  mat3_mul(a.K, Rp, KR);
  mat3_mul(KR, a.invK, A);
  PosePrep p;
  PosePrepD pd;
  for (int k = 0; k < 9; ++k) { pd.dA[k] = A[k] - ((k % 4 == 0) ? 1.0 : 0.0); p.dA[k] = (float)pd.dA[k]; }
  for (int i = 0; i < 3; ++i) {
    pd.b[i] = a.K[3 * i] * tp[0] + a.K[3 * i + 1] * tp[1] + a.K[3 * i + 2] * tp[2];
    p.b[i] = (float)pd.b[i];
  }
  out[idx] = p;
  outd[idx] = pd;
}

// ---------------------------------------------------------------------------------------
// 2./3. disparity statistics + smoothness loss value  (src/training.jl:64-67, utils.jl:163-177)
// ---------------------------------------------------------------------------------------
constexpr int STAT_ROWS = 8;       // image rows per CTA
constexpr int STAT_THREADS = 256;

template <int C>
__global__ void __launch_bounds__(STAT_THREADS)
disp_stats_kernel(FusedArgs a, float* __restrict__ part /* [N][rowblocks][Ls][3] */) {
  const int n = blockIdx.y, rb = blockIdx.x;
  const int W = a.W, H = a.H;
  const size_t plane = (size_t)W * H;
  const float* tgt = a.x + ((size_t)n * a.L + a.tgt) * C * plane;
  float acc[MDB_MAX_SCALES][3];
#pragma unroll
  for (int s = 0; s < MDB_MAX_SCALES; ++s) acc[s][0] = acc[s][1] = acc[s][2] = 0.0f;

  const int y0 = rb * STAT_ROWS, y1 = min(y0 + STAT_ROWS, H);
  for (int i = threadIdx.x; i < (y1 - y0) * W; i += STAT_THREADS) {
    const int gy = y0 + i / W, gx = i % W;
    const bool hx = gx + 1 < W, hy = gy + 1 < H;
    float ex = 0.0f, ey = 0.0f;
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const float* p = tgt + c * plane + (size_t)gy * W + gx;
      const float v = ldg(p);
      if (hx) ex += fabsf(v - ldg(p + 1));
      if (hy) ey += fabsf(v - ldg(p + W));
    }
    const float wx = hx ? expf(-ex * (1.0f / C)) : 0.0f;
    const float wy = hy ? expf(-ey * (1.0f / C)) : 0.0f;
    for (int s = 0; s < a.Ls; ++s) {
      const float* d = a.disp[s] + (size_t)n * a.dw[s] * a.dh[s];
      float v, vx = 0.0f, vy = 0.0f;
      if (a.native[s]) {
        v = ldg(d + (size_t)gy * W + gx);
        if (hx) vx = ldg(d + (size_t)gy * W + gx + 1);
        if (hy) vy = ldg(d + (size_t)(gy + 1) * W + gx);
      } else {
        v = upsample_at(d, a.dw[s], a.dh[s], a.rx[s], a.ry[s], gx, gy);
        if (hx) vx = upsample_at(d, a.dw[s], a.dh[s], a.rx[s], a.ry[s], gx + 1, gy);
        if (hy) vy = upsample_at(d, a.dw[s], a.dh[s], a.rx[s], a.ry[s], gx, gy + 1);
      }
      acc[s][0] += v;
      if (hx) acc[s][1] += fabsf(v - vx) * wx;
      if (hy) acc[s][2] += fabsf(v - vy) * wy;
    }
  }
  __shared__ float red[STAT_THREADS / 32][MDB_MAX_SCALES * 3];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int s = 0; s < a.Ls; ++s)
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const float v = warp_sum(acc[s][k]);
      if (lane == 0) red[warp][s * 3 + k] = v;
    }
  __syncthreads();
  if (threadIdx.x < a.Ls * 3) {
    float v = 0.0f;
    for (int w = 0; w < STAT_THREADS / 32; ++w) v += red[w][threadIdx.x];
    part[((size_t)n * gridDim.x + rb) * (a.Ls * 3) + threadIdx.x] = v;
  }
}

__global__ void disp_stats_finalize_kernel(FusedArgs a, const float* __restrict__ part, int rowblocks,
                                           int normalize, ScaleStat* __restrict__ stats) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= a.Ls * a.N) return;
  const int s = idx / a.N, n = idx % a.N;
  double sum = 0, tx = 0, ty = 0;
  for (int rb = 0; rb < rowblocks; ++rb) {
    const float* p = part + ((size_t)n * rowblocks + rb) * (a.Ls * 3) + s * 3;
    sum += p[0]; tx += p[1]; ty += p[2];
  }
  const double wh = (double)a.W * a.H;
  const double inv_mp = normalize ? 1.0 / (sum / wh + 1e-7) : 1.0;
  const double l = ((double)a.kx[s] * tx + (double)a.ky[s] * ty) * inv_mp;
  ScaleStat st;
  st.inv_mp = (float)inv_mp;
  st.cconst = normalize ? (float)(-l * inv_mp / wh) : 0.0f;
  st.loss = (float)l;
  st.pad = 0.0f;
  stats[idx] = st;
}

}  // namespace

#include "photo_fused_main.inl"
#include "photo_strip.inl"
#include "photo_fused_host.inl"
