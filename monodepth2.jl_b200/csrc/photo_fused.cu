// photo_fused.cu -- the fused view-synthesis photometric loss, forward + backward (P5-P12,
// S1-S2 of SURVEY 8a).  Reference: src/training.jl:21-78 (train_loss) and everything it calls
// in src/utils.jl; src/simple_depth.jl:25-41 for the single-scale variant.
//
// Launch sequence of mdb_train_loss_fwdbwd (all on the caller's stream, no host sync):
//   1. disp_mean_kernel        per (scale, item): sum of the up-sampled disparity (native maps only,
//                              no image traffic)                                  [L2-bound, tiny]
//   2. prep_kernel             composeT in double -> A = K R' invK, b = K t' per (item, source);
//                              mean -> 1 / (mean + 1e-7) per (scale, item)
//   3. photo_strip_kernel      one warp per 28-column strip: backproject -> project -> bilinear
//                              sample -> SSIM/L1 -> min over sources / automask -> pullback to the
//                              up-sampled disparity and to (A, b); smoothness value and pullback;
//                              one partial record per warp                       [issue-bound]
//   4. finalize_item_kernel    fixed-order sum of the records (double): photometric sum, pose chain
//                              back to (rvec, tvec), smoothness loss and the constant the
//                              mean-normalisation adds to every disparity gradient
//   5. upsample_bwd_band_kernel transpose of upsample_bilinear for the low-res scales (one streaming
//                              gather pass per band of native rows) and the constant for the native ones
//   6. (inside 4)              scalar loss, by the finalize CTA that finishes last
// No float atomics anywhere: results are bit-reproducible run to run.
#include "photo_fused.cuh"
#include "pose_math.cuh"

namespace {

struct PoseArgs {
  const float* rvec[MDB_MAX_SOURCES];
  const float* tvec[MDB_MAX_SOURCES];
  int invert[MDB_MAX_SOURCES];
  double K[9], invK[9];  // math row-major
  int N, S;
};

// ---------------------------------------------------------------------------------------
// 1. sum of the up-sampled disparity (src/training.jl:44-46, 64) from the native maps alone
// ---------------------------------------------------------------------------------------
constexpr int MEAN_ROWS = 16;
constexpr int MEAN_THREADS = 256;

__global__ void __launch_bounds__(MEAN_THREADS)
disp_mean_kernel(const __grid_constant__ FusedArgs a, double* __restrict__ part /* [Ls][N][rowblocks] */) {
  const int rb = blockIdx.x, sc = blockIdx.y, n = blockIdx.z;
  const int W = a.W, H = a.H, dw = a.dw[sc], dh = a.dh[sc];
  const float* d = a.disp[sc] + (size_t)n * dw * dh;
  const int y0 = rb * MEAN_ROWS, y1 = min(y0 + MEAN_ROWS, H);
  double acc = 0.0;
  if (a.native[sc]) {
    const int i0 = y0 * W, i1 = y1 * W;
    if ((((uintptr_t)(d + i0)) & 15) == 0) {   // 16-byte aligned slab: 128-bit loads, 4 in flight per thread
      const float4* d4 = reinterpret_cast<const float4*>(d + i0);
      const int n4 = (i1 - i0) / 4;
      float s0 = 0.0f, s1 = 0.0f, s2 = 0.0f, s3 = 0.0f;   // <= 16 rows x W / 256 / 4 terms each: float is ample here
      // batches of 8 independent 128-bit loads per thread (all 16 of a 1024-wide slab in two batches)
      for (int b = threadIdx.x; b < n4; b += 8 * MEAN_THREADS) {
        float4 v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int i = b + k * MEAN_THREADS;
          v[k] = i < n4 ? __ldg(d4 + i) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) { s0 += v[k].x; s1 += v[k].y; s2 += v[k].z; s3 += v[k].w; }
      }
      acc = ((double)s0 + (double)s1) + ((double)s2 + (double)s3);
      for (int i = i0 + 4 * n4 + threadIdx.x; i < i1; i += MEAN_THREADS) acc += (double)ldg(d + i);
    } else {
      for (int i = i0 + threadIdx.x; i < i1; i += MEAN_THREADS) acc += (double)ldg(d + i);
    }
  } else {
    // sum over the full-res grid of the bilinear interpolant = sum_j d_j Wx(jx) Wy(jy), with
    // W(j) = sum_p max(0, 1 - |p r - j|) on the exact integer numerator: native pixels only
    const UpAxis uy = up_axis(dh, H, a.hp), ux = up_axis(dw, W, a.hp);
    const int rows_per = (dh + gridDim.x - 1) / gridDim.x;
    const int j0 = rb * rows_per, j1 = min(j0 + rows_per, dh);
    // the row weights are the same for every column: one thread per row forms them once
    __shared__ double s_wy[MEAN_THREADS];
    for (int jy = j0 + threadIdx.x; jy < j1; jy += MEAN_THREADS) {
      int py_lo, py_hi;
      up_range(uy, jy, py_lo, py_hi);
      long long numy = 0;
      for (int py = py_lo; py <= py_hi; ++py) numy += max(0ll, uy.D - llabs(up_num(uy, py) - jy * uy.D));
      s_wy[jy - j0] = (double)numy / (double)uy.D;
    }
    __syncthreads();
    for (int jx = threadIdx.x; jx < dw; jx += MEAN_THREADS) {
      int px_lo, px_hi;
      up_range(ux, jx, px_lo, px_hi);
      long long numx = 0;   // sum of (D - |num(p) - jx D|) over the support, exact
      for (int px = px_lo; px <= px_hi; ++px) numx += max(0ll, ux.D - llabs(up_num(ux, px) - jx * ux.D));
      const double wx = (double)numx / (double)ux.D;
      double col = 0.0;
      for (int jy = j0; jy < j1; ++jy) col = fma(s_wy[jy - j0], (double)ldg(d + jy * dw + jx), col);
      acc = fma(wx, col, acc);
    }
  }
  __shared__ double red[MEAN_THREADS / 32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(MDB_FULL, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double v = 0.0;
    for (int w = 0; w < MEAN_THREADS / 32; ++w) v += red[w];
    part[((size_t)sc * a.N + n) * gridDim.x + rb] = v;
  }
}

// ---------------------------------------------------------------------------------------
// 2. pose prep (threads [0, N S)) and mean -> normalisation factor (threads [N S, N S + Ls N))
// ---------------------------------------------------------------------------------------
__global__ void prep_kernel(PoseArgs a, PoseStrip* __restrict__ outd, const double* __restrict__ mean_part,
                            int rowblocks, int Ls, int W, int H, int normalize, ScaleStat* __restrict__ stats,
                            unsigned* __restrict__ done_counter) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx == 0) *done_counter = 0u;   // finalize_item_kernel's last-CTA election
  if (idx < a.N * a.S) {
    const int n = idx / a.S, s = idx % a.S;
    double r[3], t[3], Rp[9], tp[3], KR[9], A[9];
    for (int k = 0; k < 3; ++k) { r[k] = a.rvec[s][3 * n + k]; t[k] = a.tvec[s][3 * n + k]; }
    So3Aux aux;
    compose_T_d(r, t, a.invert[s] != 0, Rp, tp, aux);
    mat3_mul(a.K, Rp, KR);
    mat3_mul(KR, a.invK, A);
    PoseStrip pd;
    for (int k = 0; k < 9; ++k) pd.dA[k] = A[k] - ((k % 4 == 0) ? 1.0 : 0.0);
    for (int i = 0; i < 3; ++i) pd.b[i] = a.K[3 * i] * tp[0] + a.K[3 * i + 1] * tp[1] + a.K[3 * i + 2] * tp[2];
    pd.i2 = 1.0 + pd.dA[4];
    pd.b3e = pd.b[2] + 1e-7;
    pd.b0f = (float)pd.b[0]; pd.b1f = (float)pd.b[1]; pd.b3f = (float)pd.b3e; pd.pad = 0.0f;
    outd[idx] = pd;
    return;
  }
  const int j = idx - a.N * a.S;
  if (j >= Ls * a.N) return;
  double sum = 0.0;
  if (normalize)
    for (int rb = 0; rb < rowblocks; ++rb) sum += mean_part[(size_t)j * rowblocks + rb];
  ScaleStat st;
  st.inv_mp = normalize ? (float)(1.0 / (sum / ((double)W * H) + 1e-7)) : 1.0f;   // src/training.jl:64-65
  st.cconst = 0.0f;
  st.loss = 0.0f;
  st.pad = 0.0f;
  stats[j] = st;
}

}  // namespace

#include "photo_strip.inl"
#include "photo_fused_host.inl"
