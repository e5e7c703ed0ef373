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
//   5. upsample_bwd_kernel     transpose of upsample_bilinear for the low-res scales (gather) and
//                              the constant for the native ones
//   6. finalize_loss_kernel    scalar loss
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
// 1. sum of the up-sampled disparity (src/training.jl:44-46, 64): thread = full-res column,
//    walking down the rows of its block with the column-interpolated native rows cached
// ---------------------------------------------------------------------------------------
constexpr int MEAN_ROWS = 32;
constexpr int MEAN_THREADS = 256;

__global__ void __launch_bounds__(MEAN_THREADS)
disp_mean_kernel(const __grid_constant__ FusedArgs a, double* __restrict__ part /* [Ls][N][rowblocks] */) {
  const int rb = blockIdx.x, sc = blockIdx.y, n = blockIdx.z;
  const int W = a.W, H = a.H, dw = a.dw[sc], dh = a.dh[sc];
  const float* d = a.disp[sc] + (size_t)n * dw * dh;
  const int y0 = rb * MEAN_ROWS, y1 = min(y0 + MEAN_ROWS, H);
  double acc = 0.0;
  if (a.native[sc]) {
    for (int i = y0 * W + threadIdx.x; i < y1 * W; i += MEAN_THREADS) acc += (double)ldg(d + i);
  } else {
    const double rxd = a.rxd[sc], ryd = a.ryd[sc];
    for (int gx = threadIdx.x; gx < W; gx += MEAN_THREADS) {
      const double f = (double)gx * rxd;
      const int x0 = min((int)f, dw - 1), x1 = min(x0 + 1, dw - 1);
      const double lx = f - (double)x0;
      int cy = -1;
      double top = 0.0, dif = 0.0;
      for (int gy = y0; gy < y1; ++gy) {
        const double fy = (double)gy * ryd;
        const int r0 = min((int)fy, dh - 1);
        if (r0 != cy) {
          cy = r0;
          const int r1 = min(r0 + 1, dh - 1);
          const double v00 = ldg(d + r0 * dw + x0), v01 = ldg(d + r0 * dw + x1);
          const double v10 = ldg(d + r1 * dw + x0), v11 = ldg(d + r1 * dw + x1);
          top = fma(lx, v01 - v00, v00);
          dif = fma(lx, v11 - v10, v10) - top;
        }
        // the strip kernel rounds the up-sampled value to float before using it; sum the same numbers
        acc += (double)(float)fma(fy - (double)r0, dif, top);
      }
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
__global__ void prep_kernel(PoseArgs a, PosePrepD* __restrict__ outd, const double* __restrict__ mean_part,
                            int rowblocks, int Ls, int W, int H, int normalize, ScaleStat* __restrict__ stats) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < a.N * a.S) {
    const int n = idx / a.S, s = idx % a.S;
    double r[3], t[3], Rp[9], tp[3], KR[9], A[9];
    for (int k = 0; k < 3; ++k) { r[k] = a.rvec[s][3 * n + k]; t[k] = a.tvec[s][3 * n + k]; }
    So3Aux aux;
    compose_T_d(r, t, a.invert[s] != 0, Rp, tp, aux);
    mat3_mul(a.K, Rp, KR);
    mat3_mul(KR, a.invK, A);
    PosePrepD pd;
    for (int k = 0; k < 9; ++k) pd.dA[k] = A[k] - ((k % 4 == 0) ? 1.0 : 0.0);
    for (int i = 0; i < 3; ++i) pd.b[i] = a.K[3 * i] * tp[0] + a.K[3 * i + 1] * tp[1] + a.K[3 * i + 2] * tp[2];
    outd[idx] = pd;
    return;
  }
  const int j = idx - a.N * a.S;
  if (j >= Ls * a.N) return;
  double sum = 0.0;
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
