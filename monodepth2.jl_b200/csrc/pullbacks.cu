// pullbacks.cu -- pullbacks of the fine-grained operators that `slow_depth` differentiates as written
// (src/simple_depth.jl:25-41: warp -> prediction_loss -> mean, smooth_loss) and of SSIM / _apply_mask, so that a
// ChainRulesCore.rrule can sit on each reference function separately (SURVEY 8b export list).  The training hot path
// does not come through here: train_loss / slow_depth use the fused kernel, which produces value and gradient in
// one pass.  These operators are two short streaming passes each:
//   pass 1  per window q: SSIM statistics (sums shifted by the window centre, as ssim_window() wants them, so that
//           sigma = E[x'^2] - E[x']^2 does not cancel in float32) -> pullback coefficients (A, B, C) of the window,
//           scaled by the incoming cotangent; for prediction_loss also the first-arg-min over the sources
//   pass 2  per pixel p: gather over the <= 9 windows that contain p, with the multiplicities pad_reflect(1) gives
//           border pixels:  d x_p = sum_q m_pq / 9 (A_q + 2 (x_p - x_q) B_q + (y_p - y_q) C_q)
#include "photo_fused.cuh"

namespace {

constexpr int PB_THREADS = 256;

// multiplicity of pixel p among the three taps q-1, q, q+1 of a window centred on q, under pad_reflect(1)
// (src/utils.jl:30-31: index -1 -> 1, n -> n-2)
MDB_DEV int tap_mult(int q, int p, int n) {
  int m = 0;
#pragma unroll
  for (int k = -1; k <= 1; ++k) {
    int t = q + k;
    t = t < 0 ? -t : (t >= n ? 2 * n - 2 - t : t);
    m += (t == p);
  }
  return m;
}

// SSIM of the window centred on (gx, gy) of one channel plane, and its pullback coefficients in the basis shifted
// by the window centre (ssim_window, photo_fused.cuh)
MDB_DEV float window_coeffs(const float* __restrict__ x, const float* __restrict__ y, int gx, int gy, int W, int H,
                            float c1, float c2, float& A, float& B, float& Cc) {
  int cx[3], ry[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) { cx[k] = reflect1(gx + k, W) - 1; ry[k] = (reflect1(gy + k, H) - 1) * W; }
  const float xc = ldg(x + gy * W + gx), yc = ldg(y + gy * W + gx);
  float Sx = 0, Sy = 0, Sxx = 0, Syy = 0, Sxy = 0;
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    const float xv = ldg(x + ry[k / 3] + cx[k % 3]) - xc, yv = ldg(y + ry[k / 3] + cx[k % 3]) - yc;
    Sx += xv; Sy += yv; Sxx = fmaf(xv, xv, Sxx); Syy = fmaf(yv, yv, Syy); Sxy = fmaf(xv, yv, Sxy);
  }
  return ssim_window(Sx, Sy, Sxx, Syy, Sxy, xc, yc, c1, c2, A, B, Cc);
}

// ---- pass 1 ---------------------------------------------------------------------------------------------------
// SSIM: coef[n][c][3][H][W] = g[n][c][q] (A, B, C)
__global__ void __launch_bounds__(PB_THREADS)
ssim_coeff_kernel(const float* __restrict__ x, const float* __restrict__ y, const float* __restrict__ g, int W, int H,
                  int C, int N, float c1, float c2, float* __restrict__ coef) {
  const size_t wh = (size_t)W * H, total = wh * C * N;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const size_t nc = i / wh;
    const int pix = (int)(i % wh);
    float A, B, Cc;
    window_coeffs(x + nc * wh, y + nc * wh, pix % W, pix / W, W, H, c1, c2, A, B, Cc);
    const float gq = ldg(g + i);
    float* o = coef + nc * 3 * wh + pix;
    o[0] = gq * A; o[wh] = gq * B; o[2 * wh] = gq * Cc;
  }
}

struct PredArgs {
  const float* pred[MDB_MAX_SOURCES];
  float* d_pred[MDB_MAX_SOURCES];
  const float* target; const float* g;
  int S, W, H, C, N;
  float alpha, c1, c2, tie;
  float* coef;          // [N][C][3][H][W]: the winner's coefficients times g alpha / C
  signed char* sel;     // [N][H][W]: winning source of every window
};

// prediction_loss = minimum over the sources of alpha mean_c SSIM + (1 - alpha) mean_c |target - pred|
// (src/training.jl:1-15); the pullback of `minimum(cat(...; dims=3); dims=3)` goes to the FIRST arg-min
__global__ void __launch_bounds__(PB_THREADS) pred_coeff_kernel(const __grid_constant__ PredArgs a) {
  const size_t wh = (size_t)a.W * a.H, total = wh * a.N;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const size_t n = i / wh;
    const int pix = (int)(i % wh), gx = pix % a.W, gy = pix / a.W;
    float best = 0.0f, bA[3], bB[3], bC[3];
    int bs = 0;
    for (int s = 0; s < a.S; ++s) {
      float ls = 0.0f, l1 = 0.0f, A[3], B[3], Cc[3];
      for (int c = 0; c < a.C; ++c) {
        const float* xp = a.pred[s] + (n * a.C + c) * wh;
        const float* yp = a.target + (n * a.C + c) * wh;
        ls += window_coeffs(xp, yp, gx, gy, a.W, a.H, a.c1, a.c2, A[c], B[c], Cc[c]);
        l1 += fabsf(ldg(yp + pix) - ldg(xp + pix));
      }
      const float v = a.alpha * (ls / a.C) + (1.0f - a.alpha) * (l1 / a.C);
      if (s == 0 || v < best) {
        best = v; bs = s;
        for (int c = 0; c < a.C; ++c) { bA[c] = A[c]; bB[c] = B[c]; bC[c] = Cc[c]; }
      }
    }
    a.sel[i] = (signed char)bs;
    const float k = ldg(a.g + i) * a.alpha / (float)a.C;
    for (int c = 0; c < a.C; ++c) {
      float* o = a.coef + (n * a.C + c) * 3 * wh + pix;
      o[0] = k * bA[c]; o[wh] = k * bB[c]; o[2 * wh] = k * bC[c];
    }
  }
}

// ---- pass 2 ---------------------------------------------------------------------------------------------------
// d x[n][c][p] for source `s` (sel == nullptr: SSIM alone, every window counts)
__global__ void __launch_bounds__(PB_THREADS)
ssim_gather_kernel(const float* __restrict__ x, const float* __restrict__ y, const float* __restrict__ coef,
                   const signed char* __restrict__ sel, int s, const float* __restrict__ g_l1, float kl1, float tie,
                   int W, int H, int C, int N, float* __restrict__ dx) {
  const size_t wh = (size_t)W * H, total = wh * C * N;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const size_t nc = i / wh, n = nc / C;
    const int pix = (int)(i % wh), gx = pix % W, gy = pix / W;
    const float* xp = x + nc * wh;
    const float* yp = y + nc * wh;
    const float* cf = coef + nc * 3 * wh;
    const float xv = ldg(xp + pix), yv = ldg(yp + pix);
    float acc = 0.0f;
    for (int qy = max(gy - 1, 0); qy <= min(gy + 1, H - 1); ++qy) {
      const int my = tap_mult(qy, gy, H);
      for (int qx = max(gx - 1, 0); qx <= min(gx + 1, W - 1); ++qx) {
        const int q = qy * W + qx;
        if (sel && sel[n * wh + q] != s) continue;
        const float m = (float)(my * tap_mult(qx, gx, W));
        const float t = fmaf(2.0f * (xv - ldg(xp + q)), ldg(cf + wh + q), fmaf(yv - ldg(yp + q), ldg(cf + 2 * wh + q), ldg(cf + q)));
        acc = fmaf(m, t, acc);
      }
    }
    acc *= (1.0f / 9.0f);
    // L1 of the winner at this pixel: d |y - x| / dx = -abs'(y - x)
    if (g_l1 && sel[n * wh + pix] == s) acc = fmaf(ldg(g_l1 + n * wh + pix) * kl1, sgnt(-tie, xv - yv), acc);
    dx[i] = acc;
  }
}

// _apply_mask(mask, warp_loss) = minimum(cat(mask, warp_loss; dims=3); dims=3) (src/training.jl:17-19): the mask
// comes first, so it keeps ties and warp_loss gets the cotangent only where it is strictly smaller
__global__ void apply_mask_bwd_kernel(const float* __restrict__ mask, const float* __restrict__ wl, const float* __restrict__ g,
                                      int64_t n, float* __restrict__ d_mask, float* __restrict__ d_wl) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const bool w = wl[i] < mask[i];
    if (d_wl) d_wl[i] = w ? g[i] : 0.0f;
    if (d_mask) d_mask[i] = w ? 0.0f : g[i];
  }
}

inline int pb_grid(mdb_handle* h, size_t n) {
  const size_t want = (n + PB_THREADS - 1) / PB_THREADS;
  return (int)std::min<size_t>(want, (size_t)h->num_sms * 32);
}

}  // namespace

extern "C" {

size_t mdb_ssim_bwd_workspace_bytes(int32_t W, int32_t H, int32_t C, int32_t N) {
  if (W < 2 || H < 2 || C < 1 || N < 1) return 0;
  return mdb_align_up(sizeof(float) * 3 * (size_t)W * H * C * N, 256) + mdb_align_up((size_t)W * H * N, 256);
}

int32_t mdb_ssim_bwd(mdb_handle* h, const float* x, const float* y, const float* g, int32_t W, int32_t H, int32_t C,
                     int32_t N, float c1, float c2, float* dx, void* workspace, size_t workspace_bytes, void* stream) {
  MDB_REQUIRE(h, h && x && y && g && dx, MDB_E_INVALID_ARG, "mdb_ssim_bwd: null argument");
  MDB_REQUIRE(h, W >= 2 && H >= 2 && C >= 1 && N >= 1, MDB_E_SHAPE, "mdb_ssim_bwd: need W,H >= 2");
  MDB_REQUIRE(h, workspace && workspace_bytes >= mdb_ssim_bwd_workspace_bytes(W, H, C, N) && ((uintptr_t)workspace & 255) == 0,
              MDB_E_WORKSPACE, "mdb_ssim_bwd: workspace too small or misaligned");
  cudaStream_t st = (cudaStream_t)stream;
  const size_t total = (size_t)W * H * C * N;
  float* coef = (float*)workspace;
  ssim_coeff_kernel<<<pb_grid(h, total), PB_THREADS, 0, st>>>(x, y, g, W, H, C, N, c1, c2, coef);
  MDB_LAUNCHED(h);
  ssim_gather_kernel<<<pb_grid(h, total), PB_THREADS, 0, st>>>(x, y, coef, nullptr, 0, nullptr, 0.0f, 0.0f, W, H, C, N, dx);
  MDB_LAUNCHED(h);
  return MDB_OK;
}

int32_t mdb_prediction_loss_bwd(mdb_handle* h, const float* const* predictions, int32_t S, const float* target,
                                const float* g, int32_t W, int32_t H, int32_t C, int32_t N, float alpha, float c1,
                                float c2, int32_t abs_subgradient, float* const* d_predictions, void* workspace,
                                size_t workspace_bytes, void* stream) {
  MDB_REQUIRE(h, h && predictions && target && g && d_predictions, MDB_E_INVALID_ARG, "mdb_prediction_loss_bwd: null argument");
  MDB_REQUIRE(h, S >= 1 && S <= MDB_MAX_SOURCES && W >= 2 && H >= 2 && C >= 1 && C <= 3 && N >= 1, MDB_E_SHAPE,
              "mdb_prediction_loss_bwd: need S <= %d, W,H >= 2, C <= 3", MDB_MAX_SOURCES);
  MDB_REQUIRE(h, workspace && workspace_bytes >= mdb_ssim_bwd_workspace_bytes(W, H, C, N) && ((uintptr_t)workspace & 255) == 0,
              MDB_E_WORKSPACE, "mdb_prediction_loss_bwd: workspace too small or misaligned (mdb_ssim_bwd_workspace_bytes)");
  cudaStream_t st = (cudaStream_t)stream;
  PredArgs a{};
  for (int s = 0; s < S; ++s) {
    MDB_REQUIRE(h, predictions[s] && d_predictions[s], MDB_E_INVALID_ARG, "mdb_prediction_loss_bwd: null prediction %d", s);
    a.pred[s] = predictions[s]; a.d_pred[s] = d_predictions[s];
  }
  a.target = target; a.g = g; a.S = S; a.W = W; a.H = H; a.C = C; a.N = N;
  a.alpha = alpha; a.c1 = c1; a.c2 = c2; a.tie = abs_subgradient ? 1.0f : 0.0f;
  a.coef = (float*)workspace;
  a.sel = (signed char*)workspace + mdb_align_up(sizeof(float) * 3 * (size_t)W * H * C * N, 256);
  pred_coeff_kernel<<<pb_grid(h, (size_t)W * H * N), PB_THREADS, 0, st>>>(a);
  MDB_LAUNCHED(h);
  const size_t total = (size_t)W * H * C * N;
  for (int s = 0; s < S; ++s) {
    ssim_gather_kernel<<<pb_grid(h, total), PB_THREADS, 0, st>>>(a.pred[s], target, a.coef, a.sel, s, g, (1.0f - alpha) / (float)C,
                                                                 a.tie, W, H, C, N, a.d_pred[s]);
    MDB_LAUNCHED(h);
  }
  return MDB_OK;
}

int32_t mdb_apply_mask_bwd(mdb_handle* h, const float* mask, const float* warp_loss, const float* g, int64_t n,
                           float* d_mask, float* d_warp_loss, void* stream) {
  MDB_REQUIRE(h, h && mask && warp_loss && g && n > 0 && (d_mask || d_warp_loss), MDB_E_INVALID_ARG, "mdb_apply_mask_bwd: bad argument");
  apply_mask_bwd_kernel<<<pb_grid(h, (size_t)n), PB_THREADS, 0, (cudaStream_t)stream>>>(mask, warp_loss, g, n, d_mask, d_warp_loss);
  MDB_LAUNCHED(h);
  return MDB_OK;
}

}  // extern "C"
