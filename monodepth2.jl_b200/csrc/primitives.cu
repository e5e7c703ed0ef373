// primitives.cu -- un-fused entry points that mirror the reference's fine-grained functions
// (P5-P11, P13, S2 of SURVEY 8a).  They exist so that every function the reference's tests
// and drivers call by name has a device implementation behind the C ABI; the training hot
// path uses the fused kernel in photo_fused.cu instead.
#include "photo_fused.cuh"
#include <mutex>
#include "pose_math.cuh"

namespace {

inline int nblk(int64_t n, int b) { return (int)((n + b - 1) / b); }

// disparity_to_depth -- src/utils.jl:179-183
__global__ void d2d_kernel(const float* __restrict__ d, int64_t n, float mind, float dd, float* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = 1.0f / fmaf(d[i], dd, mind);
}

struct Mat3 { float m[9]; };  // math row-major

// Backproject -- src/utils.jl:45-69: points(:, w + (h-1) W, n) = depth * invK [w, h, 1]
__global__ void backproject_kernel(const float* __restrict__ depth, Mat3 iK, int W, int H, int N,
                                   float* __restrict__ pts) {
  const int64_t total = (int64_t)W * H * N;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int pix = (int)(i % ((int64_t)W * H));
    const float w = (float)(pix % W + 1), hh = (float)(pix / W + 1);
    const float d = depth[i];
#pragma unroll
    for (int r = 0; r < 3; ++r) pts[3 * i + r] = d * (iK.m[3 * r] * w + iK.m[3 * r + 1] * hh + iK.m[3 * r + 2]);
  }
}

// Project + normalize -- src/utils.jl:71-103
__global__ void project_kernel(const float* __restrict__ pts, Mat3 K, const float* __restrict__ R,
                               const float* __restrict__ t, int W, int H, int N, float* __restrict__ uv) {
  const int64_t wh = (int64_t)W * H, total = wh * N;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int n = (int)(i / wh);
    const float* Rn = R + 9 * n;  // column-major: R[i][j] at j*3+i
    const float X0 = pts[3 * i], X1 = pts[3 * i + 1], X2 = pts[3 * i + 2];
    float Y[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) Y[r] = Rn[r] * X0 + Rn[3 + r] * X1 + Rn[6 + r] * X2 + t[3 * n + r];
    float P[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) P[r] = K.m[3 * r] * Y[0] + K.m[3 * r + 1] * Y[1] + K.m[3 * r + 2] * Y[2];
    const float den = 1.0f / (P[2] + 1e-7f);
    uv[2 * i] = (((P[0] * den) - 1.0f) / (float)(W - 1) - 0.5f) * 2.0f;
    uv[2 * i + 1] = (((P[1] * den) - 1.0f) / (float)(H - 1) - 0.5f) * 2.0f;
  }
}

// NNlib grid_sample (bilinear, :border, align-corners=true): input (W,H,C,N), grid (2,Wo,Ho,N)
__global__ void grid_sample_kernel(const float* __restrict__ in, const float* __restrict__ grid, int W, int H,
                                   int C, int N, int Wo, int Ho, float* __restrict__ out) {
  const int64_t who = (int64_t)Wo * Ho, total = who * N;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int n = (int)(i / who);
    const int64_t pix = i % who;
    float ix = (grid[2 * i] + 1.0f) * 0.5f * (float)(W - 1);   // 0-based
    float iy = (grid[2 * i + 1] + 1.0f) * 0.5f * (float)(H - 1);
    ix = fminf(fmaxf(ix, 0.0f), (float)(W - 1));
    iy = fminf(fmaxf(iy, 0.0f), (float)(H - 1));
    const float fx0 = floorf(ix), fy0 = floorf(iy);
    const float fx = ix - fx0, fy = iy - fy0;
    const int x0 = (int)fx0, y0 = (int)fy0, x1 = min(x0 + 1, W - 1), y1 = min(y0 + 1, H - 1);
    for (int c = 0; c < C; ++c) {
      const float* p = in + ((size_t)n * C + c) * W * H;
      const float v00 = p[y0 * W + x0], v01 = p[y0 * W + x1], v10 = p[y1 * W + x0], v11 = p[y1 * W + x1];
      const float top = fmaf(fx, v01 - v00, v00), bot = fmaf(fx, v11 - v10, v10);
      out[((size_t)n * C + c) * who + pix] = fmaf(fy, bot - top, top);
    }
  }
}

// warp (P13): disparity -> depth -> backproject -> project -> sample, per source.
struct WarpArgs {
  const float* disp; const float* x; float* out[MDB_MAX_SOURCES];
  const PosePrep* pp; const PosePrepD* ppd; int src[MDB_MAX_SOURCES];
  int S, W, H, C, L, N; float mind, dd; double mind_d, dd_d;
};

__global__ void warp_kernel(const __grid_constant__ WarpArgs a) {
  const int64_t wh = (int64_t)a.W * a.H, total = wh * a.N;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int n = (int)(i / wh);
    const int pix = (int)(i % wh), gx = pix % a.W, gy = pix / a.W;
    const double invd = fma((double)a.disp[i], a.dd_d, a.mind_d);
    for (int s = 0; s < a.S; ++s) {
      const Samp sp = sample_coords(a.ppd[n * a.S + s], invd, gx, gy, a.W, a.H);
      const Taps& t = sp.t;
      for (int c = 0; c < a.C; ++c) {
        const float* p = a.x + (((size_t)n * a.L + a.src[s]) * a.C + c) * wh;
        const float v00 = p[t.o00], v01 = p[t.o01], v10 = p[t.o10], v11 = p[t.o11];
        const float top = fmaf(t.fx, v01 - v00, v00), bot = fmaf(t.fx, v11 - v10, v10);
        a.out[s][((size_t)n * a.C + c) * wh + pix] = fmaf(t.fy, bot - top, top);
      }
    }
  }
}

// pose constants from already-composed (R, t) (Julia layout), for mdb_warp
struct WarpPoseArgs { const float* R[MDB_MAX_SOURCES]; const float* t[MDB_MAX_SOURCES]; double K[9], invK[9]; int N, S; };
__global__ void warp_pose_kernel(WarpPoseArgs a, PosePrep* __restrict__ out, PosePrepD* __restrict__ outd) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= a.N * a.S) return;
  const int n = idx / a.S, s = idx % a.S;
  double R[9], KR[9], A[9];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) R[3 * i + j] = a.R[s][9 * n + 3 * j + i];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) { double v = 0; for (int k = 0; k < 3; ++k) v += a.K[3 * i + k] * R[3 * k + j]; KR[3 * i + j] = v; }
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) { double v = 0; for (int k = 0; k < 3; ++k) v += KR[3 * i + k] * a.invK[3 * k + j]; A[3 * i + j] = v; }
  PosePrep p;
  PosePrepD pd;
  for (int k = 0; k < 9; ++k) { pd.dA[k] = A[k] - ((k % 4 == 0) ? 1.0 : 0.0); p.dA[k] = (float)pd.dA[k]; }
  for (int i = 0; i < 3; ++i) {
    double v = 0; for (int k = 0; k < 3; ++k) v += a.K[3 * i + k] * (double)a.t[s][3 * n + k];
    pd.b[i] = v; p.b[i] = (float)v;
  }
  out[idx] = p;
  outd[idx] = pd;
}

// pullback of warp (P13): d warped[s] (W,H,C,N) -> d disp (1,W,H,N), and per (item, source) the 12 sums dL/dA (9),
// dL/db (3) of P' = A q + invd b, reduced warp -> block -> per-block record (no float atomics: deterministic)
struct WarpBwdArgs {
  const float* disp; const float* x; const float* g[MDB_MAX_SOURCES];
  const PosePrepD* ppd; int src[MDB_MAX_SOURCES];
  int S, W, H, C, L, N; double mind_d, dd_d;
  float* d_disp; float* partials;   // [N][blocks][S][12]
};
constexpr int WB_THREADS = 256;

__global__ void __launch_bounds__(WB_THREADS) warp_bwd_kernel(const __grid_constant__ WarpBwdArgs a) {
  const int n = blockIdx.y, wh = a.W * a.H;
  const int pix = blockIdx.x * WB_THREADS + threadIdx.x;
  const bool on = pix < wh;
  const int gx = on ? pix % a.W : 0, gy = on ? pix / a.W : 0;
  __shared__ float red[WB_THREADS / 32][12];
  double invd = 0.0;
  if (on) invd = fma((double)a.disp[(size_t)n * wh + pix], a.dd_d, a.mind_d);
  float dinvd = 0.0f;
  for (int s = 0; s < a.S; ++s) {
    float acc[12];
#pragma unroll
    for (int k = 0; k < 12; ++k) acc[k] = 0.0f;
    if (on) {
      const PosePrepD& pp = a.ppd[n * a.S + s];
      const Samp sp = sample_coords(pp, invd, gx, gy, a.W, a.H);
      const Taps& t = sp.t;
      float gxs = 0.0f, gys = 0.0f;   // d L / d (sample position)
      for (int c = 0; c < a.C; ++c) {
        const float* p = a.x + (((size_t)n * a.L + a.src[s]) * a.C + c) * wh;
        const float v00 = p[t.o00], v01 = p[t.o01], v10 = p[t.o10], v11 = p[t.o11];
        const float h0 = v01 - v00, h1 = v11 - v10;
        const float dxv = fmaf(t.fy, h1 - h0, h0);
        const float dyv = fmaf(t.fx, h1, v10) - fmaf(t.fx, h0, v00);
        const float g = a.g[s][((size_t)n * a.C + c) * wh + pix];
        gxs = fmaf(g, dxv, gxs); gys = fmaf(g, dyv, gys);
      }
      // u = P'1 / z'', v = P'2 / z'' (1-based); no gradient where the sample was clamped (NNlib :border)
      const float dP1 = gxs * t.mx * sp.zinv, dP2 = gys * t.my * sp.zinv;
      const float u = (float)(gx + 1) + sp.fu, v = (float)(gy + 1) + sp.fv;
      const float dP3 = -(u * dP1 + v * dP2);
      const float q0 = (float)(gx + 1), q1 = (float)(gy + 1), fi = (float)invd;
      acc[0] = dP1 * q0; acc[1] = dP1 * q1; acc[2] = dP1;
      acc[3] = dP2 * q0; acc[4] = dP2 * q1; acc[5] = dP2;
      acc[6] = dP3 * q0; acc[7] = dP3 * q1; acc[8] = dP3;
      acc[9] = dP1 * fi; acc[10] = dP2 * fi; acc[11] = dP3 * fi;
      dinvd += dP1 * (float)pp.b[0] + dP2 * (float)pp.b[1] + dP3 * (float)(pp.b[2] + 1e-7);
    }
#pragma unroll
    for (int k = 0; k < 12; ++k) {
      const float v = warp_sum(acc[k]);
      if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5][k] = v;
    }
    __syncthreads();
    if (threadIdx.x < 12) {
      float v = 0.0f;
      for (int w = 0; w < WB_THREADS / 32; ++w) v += red[w][threadIdx.x];
      a.partials[(((size_t)n * gridDim.x + blockIdx.x) * a.S + s) * 12 + threadIdx.x] = v;
    }
    __syncthreads();
  }
  if (on) a.d_disp[(size_t)n * wh + pix] = dinvd * (float)a.dd_d;   // d invd / d disparity = 1/min - 1/max
}

// (dA, db) -> (dR, dt):  A = K R invK  =>  dR = K^T dA invK^T;  b = K t  =>  dt = K^T db.  Julia layouts out.
struct WarpBwdFin { float* dR[MDB_MAX_SOURCES]; float* dt[MDB_MAX_SOURCES]; double K[9], invK[9]; int N, S, blocks; };
__global__ void warp_bwd_finalize_kernel(const float* __restrict__ partials, WarpBwdFin f) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= f.N * f.S) return;
  const int n = idx / f.S, s = idx % f.S;
  double t[12];
  for (int k = 0; k < 12; ++k) t[k] = 0.0;
  for (int b = 0; b < f.blocks; ++b)
    for (int k = 0; k < 12; ++k) t[k] += (double)partials[(((size_t)n * f.blocks + b) * f.S + s) * 12 + k];
  double Kt[9], iKt[9], t1[9], dR[9];
  mat3_T(f.K, Kt);
  mat3_T(f.invK, iKt);
  mat3_mul(Kt, t, t1);
  mat3_mul(t1, iKt, dR);
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j)
      if (f.dR[s]) f.dR[s][9 * n + 3 * j + i] = (float)dR[3 * i + j];
  for (int i = 0; i < 3; ++i)
    if (f.dt[s]) f.dt[s][3 * n + i] = (float)(Kt[3 * i] * t[9] + Kt[3 * i + 1] * t[10] + Kt[3 * i + 2] * t[11]);
}

// ---- SSIM / photometric, one thread per output pixel ----------------------------------
// mode 0: write SSIM per channel (W,H,C,N); mode 1: photometric loss (W,H,1,N);
// mode 2: min over S predictions of the photometric loss (W,H,1,N)
struct PhotoArgs {
  const float* pred[4]; int64_t pred_nstride[4];  // per-item stride of each prediction
  const float* target; int64_t tgt_nstride;
  int S, W, H, C, N, mode; float alpha, c1, c2; float* out;
};

__device__ float photo_pixel(const float* __restrict__ x, const float* __restrict__ y, int gx, int gy, int W, int H,
                             int C, float alpha, float c1, float c2, float* ssim_out, int64_t ssim_cstride) {
  const size_t plane = (size_t)W * H;
  int cx[3], ry[3];
  for (int k = 0; k < 3; ++k) { cx[k] = reflect1(gx + k, W) - 1; ry[k] = (reflect1(gy + k, H) - 1) * W; }
  float ls = 0.0f, l1 = 0.0f;
  for (int c = 0; c < C; ++c) {
    float Sx = 0, Sy = 0, Sxx = 0, Syy = 0, Sxy = 0;
    const float xc = x[c * plane + gy * W + gx], yc = y[c * plane + gy * W + gx];
    for (int k = 0; k < 9; ++k) {
      const float xv = x[c * plane + ry[k / 3] + cx[k % 3]] - xc, yv = y[c * plane + ry[k / 3] + cx[k % 3]] - yc;
      Sx += xv; Sy += yv; Sxx = fmaf(xv, xv, Sxx); Syy = fmaf(yv, yv, Syy); Sxy = fmaf(xv, yv, Sxy);
    }
    float A, B, Cc;
    const float s = ssim_window(Sx, Sy, Sxx, Syy, Sxy, xc, yc, c1, c2, A, B, Cc);
    if (ssim_out) ssim_out[c * ssim_cstride] = s;
    ls += s;
    l1 += fabsf(y[c * plane + gy * W + gx] - x[c * plane + gy * W + gx]);
  }
  return alpha * (ls / C) + (1.0f - alpha) * (l1 / C);
}

__global__ void photo_kernel(PhotoArgs a) {
  const int64_t wh = (int64_t)a.W * a.H, total = wh * a.N;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int n = (int)(i / wh);
    const int pix = (int)(i % wh), gx = pix % a.W, gy = pix / a.W;
    const float* y = a.target + n * a.tgt_nstride;
    if (a.mode == 0) {
      photo_pixel(a.pred[0] + n * a.pred_nstride[0], y, gx, gy, a.W, a.H, a.C, a.alpha, a.c1, a.c2,
                  a.out + (size_t)n * a.C * wh + pix, wh);
    } else {
      float best = 0.0f;
      for (int s = 0; s < a.S; ++s) {
        const float v = photo_pixel(a.pred[s] + n * a.pred_nstride[s], y, gx, gy, a.W, a.H, a.C, a.alpha, a.c1,
                                    a.c2, nullptr, 0);
        best = (s == 0 || v < best) ? v : best;
      }
      a.out[i] = best;
    }
  }
}

}  // namespace
#include "photo_stream.inl"
namespace {

__global__ void min2_kernel(const float* __restrict__ a, const float* __restrict__ b, int64_t n, float* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = fminf(a[i], b[i]);
}

// ---- smooth_loss (S2) -- src/utils.jl:163-177 -----------------------------------------
constexpr int SM_THREADS = 256, SM_ROWS = 16, SM_ROWS_MIN = 8;   // the workspace holds partials for 8-row blocks

// Thread = column, walking down SM_ROWS rows: the row below is loaded one step ahead, the vertical edge
// weight is carried to the next row, the horizontal one comes from the left lane by shuffle -- each edge's
// exp(-mean_c |dI|) is evaluated once (twice only at warp / block seams), each element loaded once.
template <int C>
__global__ void __launch_bounds__(SM_THREADS)
smooth_kernel(const float* __restrict__ disp, const float* __restrict__ img, int W, int H, int N,
              float kx, float ky, float tie, float* __restrict__ part, float* __restrict__ grad) {
  const int n = blockIdx.y, rb = blockIdx.x;
  const size_t plane = (size_t)W * H;
  const float* d = disp + n * plane;
  const float* im = img + (size_t)n * C * plane;
  const int y0 = rb * SM_ROWS, y1 = min(y0 + SM_ROWS, H);
  const int lane = threadIdx.x & 31;
  const float invC = 1.0f / (float)C;
  float ax = 0.0f, ay = 0.0f;
  auto edge = [&](const float* Ia, const float* Ib) {     // exp(-mean_c |Ia - Ib|)   src/utils.jl:168-172
    float e = 0.0f;
#pragma unroll
    for (int c = 0; c < C; ++c) e += fabsf(Ia[c] - Ib[c]);
    return __expf(-e * invC) ;
  };
  for (int gx0 = 0; gx0 < W; gx0 += SM_THREADS) {
    const int gx = gx0 + threadIdx.x;
    const bool cv = gx < W;
    const int gxc = cv ? gx : W - 1;
    float v, I[C], vd = 0.0f, Id[C], tup = 0.0f;          // tup: ky sign(df) w of the edge to the row above
    v = ldg(d + (size_t)y0 * W + gxc);
#pragma unroll
    for (int c = 0; c < C; ++c) { I[c] = ldg(im + c * plane + (size_t)y0 * W + gxc); Id[c] = 0.0f; }
    if (y0 > 0) {                                          // seam: the edge to the previous block's last row
      float Iu[C];
#pragma unroll
      for (int c = 0; c < C; ++c) Iu[c] = ldg(im + c * plane + (size_t)(y0 - 1) * W + gxc);
      const float df = ldg(d + (size_t)(y0 - 1) * W + gxc) - v;
      tup = ky * sgnt(tie, df) * edge(Iu, I);
    }
    for (int gy = y0; gy < y1; ++gy) {
      const bool has_dn = gy + 1 < H;
      if (has_dn) {
        vd = ldg(d + (size_t)(gy + 1) * W + gxc);
#pragma unroll
        for (int c = 0; c < C; ++c) Id[c] = ldg(im + c * plane + (size_t)(gy + 1) * W + gxc);
      }
      // right neighbour: lane + 1, or a load at the warp seam
      float vr = __shfl_down_sync(MDB_FULL, v, 1), Ir[C];
#pragma unroll
      for (int c = 0; c < C; ++c) Ir[c] = __shfl_down_sync(MDB_FULL, I[c], 1);
      const bool has_r = gx + 1 < W;
      if (lane == 31 && has_r) {
        vr = ldg(d + (size_t)gy * W + gx + 1);
#pragma unroll
        for (int c = 0; c < C; ++c) Ir[c] = ldg(im + c * plane + (size_t)gy * W + gx + 1);
      }
      float g = 0.0f, tr = 0.0f;
      if (cv && has_r) { const float df = v - vr, w = edge(I, Ir); ax += fabsf(df) * w; tr = kx * sgnt(tie, df) * w; g += tr; }
      // left edge = the left lane's right edge
      float tl = __shfl_up_sync(MDB_FULL, tr, 1);
      if (lane == 0 && cv && gx > 0) {
        float Il[C];
#pragma unroll
        for (int c = 0; c < C; ++c) Il[c] = ldg(im + c * plane + (size_t)gy * W + gx - 1);
        const float df = ldg(d + (size_t)gy * W + gx - 1) - v;
        tl = kx * sgnt(tie, df) * edge(Il, I);
      }
      if (cv && gx > 0) g -= tl;
      g -= tup;
      tup = 0.0f;
      if (cv && has_dn) { const float df = v - vd, w = edge(I, Id); ay += fabsf(df) * w; tup = ky * sgnt(tie, df) * w; g += tup; }
      if (grad && cv) grad[n * plane + (size_t)gy * W + gx] = g;
      v = vd;
#pragma unroll
      for (int c = 0; c < C; ++c) I[c] = Id[c];
    }
  }
  __shared__ float red[SM_THREADS / 32][2];
  const int warp = threadIdx.x >> 5;
  ax = warp_sum(ax); ay = warp_sum(ay);
  if (lane == 0) { red[warp][0] = ax; red[warp][1] = ay; }
  __syncthreads();
  if (threadIdx.x < 2) {
    float v = 0.0f;
    for (int w = 0; w < SM_THREADS / 32; ++w) v += red[w][threadIdx.x];
    part[((size_t)n * gridDim.x + rb) * 2 + threadIdx.x] = v;
  }
}

// The same walk with four columns per thread (W % 4 == 0, 16-byte aligned rows): 128-bit loads and stores --
// four times the bytes in flight per load instruction -- three of the four horizontal edges inside the thread,
// one shuffle exchange per four columns.
template <int C>
__global__ void __launch_bounds__(SM_THREADS)
smooth4_kernel(const float* __restrict__ disp, const float* __restrict__ img, int W, int H, int N, int rows,
               float kx, float ky, float tie, float* __restrict__ part, float* __restrict__ grad) {
  const int n = blockIdx.y, rb = blockIdx.x;
  const size_t plane = (size_t)W * H;
  const float* d = disp + n * plane;
  const float* im = img + (size_t)n * C * plane;
  const int y0 = rb * rows, y1 = min(y0 + rows, H);
  const int lane = threadIdx.x & 31;
  const float invC = 1.0f / (float)C;
  float ax = 0.0f, ay = 0.0f;
  auto ld4 = [](const float* p, float* o) { const float4 q = __ldg(reinterpret_cast<const float4*>(p)); o[0] = q.x; o[1] = q.y; o[2] = q.z; o[3] = q.w; };
  for (int gx0 = 0; gx0 < W; gx0 += 4 * SM_THREADS) {
    const int gx = gx0 + 4 * threadIdx.x;                  // first of this thread's four columns
    const bool cv = gx < W;
    const int gxc = cv ? gx : W - 4;
    float v[5], I[C][5], vd[4], Id[C][4], tup[4] = {0.0f, 0.0f, 0.0f, 0.0f};
    ld4(d + (size_t)y0 * W + gxc, v);
#pragma unroll
    for (int c = 0; c < C; ++c) ld4(im + c * plane + (size_t)y0 * W + gxc, I[c]);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      vd[j] = 0.0f;
#pragma unroll
      for (int c = 0; c < C; ++c) Id[c][j] = 0.0f;
    }
    if (y0 > 0) {                                          // seam: the edges to the previous block's last row
      float vu[4], Iu[C][4];
      ld4(d + (size_t)(y0 - 1) * W + gxc, vu);
#pragma unroll
      for (int c = 0; c < C; ++c) ld4(im + c * plane + (size_t)(y0 - 1) * W + gxc, Iu[c]);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float e = 0.0f;
#pragma unroll
        for (int c = 0; c < C; ++c) e += fabsf(Iu[c][j] - I[c][j]);
        tup[j] = ky * sgnt(tie, vu[j] - v[j]) * __expf(-e * invC);
      }
    }
    for (int gy = y0; gy < y1; ++gy) {
      const bool has_dn = gy + 1 < H;
      if (has_dn) {
        ld4(d + (size_t)(gy + 1) * W + gxc, vd);
#pragma unroll
        for (int c = 0; c < C; ++c) ld4(im + c * plane + (size_t)(gy + 1) * W + gxc, Id[c]);
      }
      // fifth column = the right neighbour's first: lane + 1, or a load at the warp seam
      v[4] = __shfl_down_sync(MDB_FULL, v[0], 1);
#pragma unroll
      for (int c = 0; c < C; ++c) I[c][4] = __shfl_down_sync(MDB_FULL, I[c][0], 1);
      const bool has_r = gx + 4 < W;
      if (lane == 31 && has_r) {
        v[4] = ldg(d + (size_t)gy * W + gx + 4);
#pragma unroll
        for (int c = 0; c < C; ++c) I[c][4] = ldg(im + c * plane + (size_t)gy * W + gx + 4);
      }
      float tr[4], g[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float e = 0.0f;
#pragma unroll
        for (int c = 0; c < C; ++c) e += fabsf(I[c][j] - I[c][j + 1]);
        const float w = __expf(-e * invC), df = v[j] - v[j + 1];
        const bool on = cv && (j < 3 || has_r);
        ax += on ? fabsf(df) * w : 0.0f;
        tr[j] = on ? kx * sgnt(tie, df) * w : 0.0f;
      }
      // left edge of column 0 = the left lane's right edge of its column 3
      float tl = __shfl_up_sync(MDB_FULL, tr[3], 1);
      if (lane == 0) {
        tl = 0.0f;
        if (cv && gx > 0) {
          float e = 0.0f;
#pragma unroll
          for (int c = 0; c < C; ++c) e += fabsf(ldg(im + c * plane + (size_t)gy * W + gx - 1) - I[c][0]);
          tl = kx * sgnt(tie, ldg(d + (size_t)gy * W + gx - 1) - v[0]) * __expf(-e * invC);
        }
      }
      g[0] = tr[0] - tl; g[1] = tr[1] - tr[0]; g[2] = tr[2] - tr[1]; g[3] = tr[3] - tr[2];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        g[j] -= tup[j];
        tup[j] = 0.0f;
        if (cv && has_dn) {
          float e = 0.0f;
#pragma unroll
          for (int c = 0; c < C; ++c) e += fabsf(I[c][j] - Id[c][j]);
          const float w = __expf(-e * invC), df = v[j] - vd[j];
          ay += fabsf(df) * w;
          tup[j] = ky * sgnt(tie, df) * w;
          g[j] += tup[j];
        }
      }
      if (grad && cv) *reinterpret_cast<float4*>(grad + n * plane + (size_t)gy * W + gx) = make_float4(g[0], g[1], g[2], g[3]);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        v[j] = vd[j];
#pragma unroll
        for (int c = 0; c < C; ++c) I[c][j] = Id[c][j];
      }
    }
  }
  __shared__ float red[SM_THREADS / 32][2];
  const int warp = threadIdx.x >> 5;
  ax = warp_sum(ax); ay = warp_sum(ay);
  if (lane == 0) { red[warp][0] = ax; red[warp][1] = ay; }
  __syncthreads();
  if (threadIdx.x < 2) {
    float s = 0.0f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) s += red[w][threadIdx.x];
    part[((size_t)n * gridDim.x + rb) * 2 + threadIdx.x] = s;
  }
}

__global__ void smooth_finalize_kernel(const float* __restrict__ part, int count, double kx, double ky,
                                       float* __restrict__ loss) {
  __shared__ double red[8];
  double v = 0.0;
  for (int i = threadIdx.x; i < count; i += blockDim.x) v += kx * (double)part[2 * i] + ky * (double)part[2 * i + 1];
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(MDB_FULL, v, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x == 0) { double s = 0; for (int w = 0; w < (int)(blockDim.x >> 5); ++w) s += red[w]; *loss = (float)s; }
}

int grid_for(mdb_handle* h, int64_t n, int b) { return (int)min((int64_t)h->num_sms * 16, (n + b - 1) / b); }
Mat3 to_math(const float* cm) { Mat3 m; for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) m.m[3 * i + j] = cm[3 * j + i]; return m; }

}  // namespace

extern "C" {

int32_t mdb_disparity_to_depth(mdb_handle* h, const float* disparity, int64_t n, double min_depth,
                               double max_depth, float* depth, void* stream) {
  MDB_REQUIRE(h, h && disparity && depth && n > 0, MDB_E_INVALID_ARG, "mdb_disparity_to_depth: bad argument");
  const double mind = 1.0 / max_depth, maxd = 1.0 / min_depth;
  d2d_kernel<<<grid_for(h, n, 256), 256, 0, (cudaStream_t)stream>>>(disparity, n, (float)mind, (float)(maxd - mind), depth);
  MDB_LAUNCHED(h);
  return MDB_OK;
}

int32_t mdb_backproject(mdb_handle* h, const float* depth, const float* invK, int32_t W, int32_t H,
                        int32_t N, float* points, void* stream) {
  MDB_REQUIRE(h, h && depth && invK && points && W > 0 && H > 0 && N > 0, MDB_E_INVALID_ARG, "mdb_backproject: bad argument");
  backproject_kernel<<<grid_for(h, (int64_t)W * H * N, 256), 256, 0, (cudaStream_t)stream>>>(depth, to_math(invK), W, H, N, points);
  MDB_LAUNCHED(h);
  return MDB_OK;
}

int32_t mdb_project(mdb_handle* h, const float* points, const float* K, const float* R, const float* t,
                    int32_t W, int32_t H, int32_t N, float* uv, void* stream) {
  MDB_REQUIRE(h, h && points && K && R && t && uv && W > 1 && H > 1 && N > 0, MDB_E_INVALID_ARG, "mdb_project: bad argument");
  project_kernel<<<grid_for(h, (int64_t)W * H * N, 256), 256, 0, (cudaStream_t)stream>>>(points, to_math(K), R, t, W, H, N, uv);
  MDB_LAUNCHED(h);
  return MDB_OK;
}

int32_t mdb_grid_sample(mdb_handle* h, const float* input, const float* grid, int32_t W, int32_t H,
                        int32_t C, int32_t N, int32_t Wo, int32_t Ho, float* out, void* stream) {
  MDB_REQUIRE(h, h && input && grid && out && W > 0 && H > 0 && C > 0 && N > 0 && Wo > 0 && Ho > 0,
              MDB_E_INVALID_ARG, "mdb_grid_sample: bad argument");
  grid_sample_kernel<<<grid_for(h, (int64_t)Wo * Ho * N, 256), 256, 0, (cudaStream_t)stream>>>(input, grid, W, H, C, N, Wo, Ho, out);
  MDB_LAUNCHED(h);
  return MDB_OK;
}

int32_t mdb_warp(mdb_handle* h, const float* disp, const float* x, const float* const* R,
                 const float* const* t, const int32_t* source_ids, int32_t S, const float* K,
                 const float* invK, int32_t W, int32_t H, int32_t C, int32_t L, int32_t N,
                 double min_depth, double max_depth, float* const* warped, void* stream) {
  MDB_REQUIRE(h, h && disp && x && R && t && source_ids && K && invK && warped, MDB_E_INVALID_ARG, "mdb_warp: null argument");
  MDB_REQUIRE(h, S >= 1 && S <= MDB_MAX_SOURCES && W > 1 && H > 1 && C > 0 && L > 0 && N > 0, MDB_E_SHAPE, "mdb_warp: bad shape");
  cudaStream_t st = (cudaStream_t)stream;
  WarpPoseArgs pa{};
  WarpArgs wa{};
  pa.N = N; pa.S = S;
  for (int s = 0; s < S; ++s) {
    MDB_REQUIRE(h, R[s] && t[s] && warped[s] && source_ids[s] >= 1 && source_ids[s] <= L, MDB_E_INVALID_ARG, "mdb_warp: bad source %d", s);
    pa.R[s] = R[s]; pa.t[s] = t[s]; wa.out[s] = warped[s]; wa.src[s] = source_ids[s] - 1;
  }
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) { pa.K[3 * i + j] = K[3 * j + i]; pa.invK[3 * i + j] = invK[3 * j + i]; }
  char* buf = nullptr;
  const size_t ppd_off = mdb_align_up(sizeof(PosePrep) * N * S, 256);
  MDB_CUDA(h, cudaMallocAsync((void**)&buf, ppd_off + sizeof(PosePrepD) * N * S, st));
  PosePrep* pp = (PosePrep*)buf;
  PosePrepD* ppd = (PosePrepD*)(buf + ppd_off);
  warp_pose_kernel<<<nblk(N * S, 128), 128, 0, st>>>(pa, pp, ppd);
  MDB_LAUNCHED(h);
  wa.disp = disp; wa.x = x; wa.pp = pp; wa.ppd = ppd; wa.S = S; wa.W = W; wa.H = H; wa.C = C; wa.L = L; wa.N = N;
  wa.mind_d = 1.0 / max_depth; wa.dd_d = 1.0 / min_depth - 1.0 / max_depth;
  wa.mind = (float)wa.mind_d; wa.dd = (float)wa.dd_d;
  warp_kernel<<<grid_for(h, (int64_t)W * H * N, 256), 256, 0, st>>>(wa);
  MDB_LAUNCHED(h);
  MDB_CUDA(h, cudaFreeAsync(buf, st));
  return MDB_OK;
}

int32_t mdb_warp_bwd(mdb_handle* h, const float* disp, const float* x, const float* const* R, const float* const* t,
                     const int32_t* source_ids, int32_t S, const float* K, const float* invK, int32_t W, int32_t H,
                     int32_t C, int32_t L, int32_t N, double min_depth, double max_depth, const float* const* g_warped,
                     float* d_disp, float* const* dR, float* const* dt, void* stream) {
  MDB_REQUIRE(h, h && disp && x && R && t && source_ids && K && invK && g_warped && d_disp, MDB_E_INVALID_ARG, "mdb_warp_bwd: null argument");
  MDB_REQUIRE(h, S >= 1 && S <= MDB_MAX_SOURCES && W > 1 && H > 1 && C > 0 && L > 0 && N > 0 && N <= 65535, MDB_E_SHAPE, "mdb_warp_bwd: bad shape");
  cudaStream_t st = (cudaStream_t)stream;
  WarpPoseArgs pa{};
  WarpBwdArgs wa{};
  WarpBwdFin fin{};
  pa.N = N; pa.S = S;
  for (int s = 0; s < S; ++s) {
    MDB_REQUIRE(h, R[s] && t[s] && g_warped[s] && source_ids[s] >= 1 && source_ids[s] <= L, MDB_E_INVALID_ARG, "mdb_warp_bwd: bad source %d", s);
    pa.R[s] = R[s]; pa.t[s] = t[s]; wa.g[s] = g_warped[s]; wa.src[s] = source_ids[s] - 1;
    fin.dR[s] = dR ? dR[s] : nullptr; fin.dt[s] = dt ? dt[s] : nullptr;
  }
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) { pa.K[3 * i + j] = K[3 * j + i]; pa.invK[3 * i + j] = invK[3 * j + i]; fin.K[3 * i + j] = pa.K[3 * i + j]; fin.invK[3 * i + j] = pa.invK[3 * i + j]; }
  const int blocks = (W * H + WB_THREADS - 1) / WB_THREADS;
  char* buf = nullptr;
  const size_t o_ppd = mdb_align_up(sizeof(PosePrep) * N * S, 256);
  const size_t o_part = o_ppd + mdb_align_up(sizeof(PosePrepD) * N * S, 256);
  MDB_CUDA(h, cudaMallocAsync((void**)&buf, o_part + sizeof(float) * 12 * (size_t)N * blocks * S, st));
  warp_pose_kernel<<<nblk(N * S, 128), 128, 0, st>>>(pa, (PosePrep*)buf, (PosePrepD*)(buf + o_ppd));
  MDB_LAUNCHED(h);
  wa.disp = disp; wa.x = x; wa.ppd = (const PosePrepD*)(buf + o_ppd); wa.S = S; wa.W = W; wa.H = H; wa.C = C; wa.L = L; wa.N = N;
  wa.mind_d = 1.0 / max_depth; wa.dd_d = 1.0 / min_depth - 1.0 / max_depth;
  wa.d_disp = d_disp; wa.partials = (float*)(buf + o_part);
  warp_bwd_kernel<<<dim3(blocks, N), WB_THREADS, 0, st>>>(wa);
  MDB_LAUNCHED(h);
  fin.N = N; fin.S = S; fin.blocks = blocks;
  warp_bwd_finalize_kernel<<<nblk(N * S, 64), 64, 0, st>>>((const float*)(buf + o_part), fin);
  MDB_LAUNCHED(h);
  MDB_CUDA(h, cudaFreeAsync(buf, st));
  return MDB_OK;
}

static int32_t photo_launch(mdb_handle* h, PhotoArgs& a, void* stream) {
  MDB_REQUIRE(h, a.W >= 2 && a.H >= 2 && a.C >= 1 && a.N >= 1, MDB_E_SHAPE, "SSIM needs W,H >= 2 (reflect pad)");
  // the streaming kernel covers the shapes the path uses (C in {1,3}, one or two sources); anything else
  // (C = 2, 4, ...; 3-4 predictions) runs the per-pixel kernel
  cudaStream_t st = (cudaStream_t)stream;
  if (a.C == 3 && a.S == 2) return photo_stream_launch<3, 2>(h, a, st);
  if (a.C == 3 && a.S == 1) return photo_stream_launch<3, 1>(h, a, st);
  if (a.C == 1 && a.S == 2) return photo_stream_launch<1, 2>(h, a, st);
  if (a.C == 1 && a.S == 1) return photo_stream_launch<1, 1>(h, a, st);
  photo_kernel<<<grid_for(h, (int64_t)a.W * a.H * a.N, 128), 128, 0, (cudaStream_t)stream>>>(a);
  MDB_LAUNCHED(h);
  return MDB_OK;
}

int32_t mdb_ssim(mdb_handle* h, const float* x, const float* y, int32_t W, int32_t H, int32_t C,
                 int32_t N, float c1, float c2, float* out, void* stream) {
  MDB_REQUIRE(h, h && x && y && out, MDB_E_INVALID_ARG, "mdb_ssim: null argument");
  PhotoArgs a{};
  a.pred[0] = x; a.pred_nstride[0] = (int64_t)W * H * C; a.target = y; a.tgt_nstride = (int64_t)W * H * C;
  a.S = 1; a.W = W; a.H = H; a.C = C; a.N = N; a.mode = 0; a.alpha = 1.0f; a.c1 = c1; a.c2 = c2; a.out = out;
  return photo_launch(h, a, stream);
}

int32_t mdb_photometric_loss(mdb_handle* h, const float* predicted, const float* target, int32_t W,
                             int32_t H, int32_t C, int32_t N, float alpha, float c1, float c2,
                             float* out, void* stream) {
  MDB_REQUIRE(h, h && predicted && target && out, MDB_E_INVALID_ARG, "mdb_photometric_loss: null argument");
  PhotoArgs a{};
  a.pred[0] = predicted; a.pred_nstride[0] = (int64_t)W * H * C; a.target = target; a.tgt_nstride = (int64_t)W * H * C;
  a.S = 1; a.W = W; a.H = H; a.C = C; a.N = N; a.mode = 1; a.alpha = alpha; a.c1 = c1; a.c2 = c2; a.out = out;
  return photo_launch(h, a, stream);
}

int32_t mdb_prediction_loss(mdb_handle* h, const float* const* predictions, int32_t S,
                            const float* target, int32_t W, int32_t H, int32_t C, int32_t N,
                            float alpha, float c1, float c2, float* out, void* stream) {
  MDB_REQUIRE(h, h && predictions && target && out && S >= 1 && S <= 4, MDB_E_INVALID_ARG, "mdb_prediction_loss: bad argument");
  PhotoArgs a{};
  for (int s = 0; s < S; ++s) { MDB_REQUIRE(h, predictions[s], MDB_E_INVALID_ARG, "null prediction"); a.pred[s] = predictions[s]; a.pred_nstride[s] = (int64_t)W * H * C; }
  a.target = target; a.tgt_nstride = (int64_t)W * H * C;
  a.S = S; a.W = W; a.H = H; a.C = C; a.N = N; a.mode = 2; a.alpha = alpha; a.c1 = c1; a.c2 = c2; a.out = out;
  return photo_launch(h, a, stream);
}

int32_t mdb_automasking_loss(mdb_handle* h, const float* inputs, const float* target, int64_t target_item_stride,
                             const int32_t* source_ids, int32_t S, int32_t W, int32_t H, int32_t C, int32_t L,
                             int32_t N, float alpha, float c1, float c2, float* out, void* stream) {
  MDB_REQUIRE(h, h && inputs && target && source_ids && out && S >= 1 && S <= 4, MDB_E_INVALID_ARG,
              "mdb_automasking_loss: bad argument");
  PhotoArgs a{};
  const int64_t frame = (int64_t)W * H * C;
  MDB_REQUIRE(h, target_item_stride == 0 || target_item_stride >= frame, MDB_E_INVALID_ARG,
              "mdb_automasking_loss: target_item_stride must be 0 (dense) or >= W H C");
  for (int s = 0; s < S; ++s) {
    MDB_REQUIRE(h, source_ids[s] >= 1 && source_ids[s] <= L, MDB_E_INVALID_ARG, "source id out of range");
    a.pred[s] = inputs + (source_ids[s] - 1) * frame; a.pred_nstride[s] = frame * L;
  }
  a.target = target; a.tgt_nstride = target_item_stride > 0 ? target_item_stride : frame;
  a.S = S; a.W = W; a.H = H; a.C = C; a.N = N; a.mode = 2; a.alpha = alpha; a.c1 = c1; a.c2 = c2; a.out = out;
  return photo_launch(h, a, stream);
}

int32_t mdb_apply_mask(mdb_handle* h, const float* mask, const float* warp_loss, int64_t n, float* out, void* stream) {
  MDB_REQUIRE(h, h && mask && warp_loss && out && n > 0, MDB_E_INVALID_ARG, "mdb_apply_mask: bad argument");
  min2_kernel<<<grid_for(h, n, 256), 256, 0, (cudaStream_t)stream>>>(mask, warp_loss, n, out);
  MDB_LAUNCHED(h);
  return MDB_OK;
}

size_t mdb_smooth_loss_workspace_bytes(int32_t W, int32_t H, int32_t N) {
  if (W < 1 || H < 1 || N < 1) return 0;
  return mdb_align_up(sizeof(float) * 2 * (size_t)N * ((H + SM_ROWS_MIN - 1) / SM_ROWS_MIN), 256);
}

int32_t mdb_smooth_loss(mdb_handle* h, const float* disparity, const float* image, int32_t W, int32_t H,
                        int32_t C, int32_t N, float* loss, float* d_disparity, void* workspace,
                        size_t workspace_bytes, void* stream) {
  return mdb_smooth_loss_ex(h, disparity, image, W, H, C, N, 0, loss, d_disparity, workspace, workspace_bytes, stream);
}

int32_t mdb_smooth_loss_ex(mdb_handle* h, const float* disparity, const float* image, int32_t W, int32_t H,
                           int32_t C, int32_t N, int32_t abs_subgradient, float* loss, float* d_disparity,
                           void* workspace, size_t workspace_bytes, void* stream) {
  const float tie = abs_subgradient ? 1.0f : 0.0f;   // d|x|/dx at +0.0 (every difference here is first - second of a pair)
  MDB_REQUIRE(h, h && disparity && image && loss, MDB_E_INVALID_ARG, "mdb_smooth_loss: null argument");
  MDB_REQUIRE(h, W >= 2 && H >= 2 && C >= 1 && C <= 4 && N >= 1 && N <= 65535, MDB_E_SHAPE,
              "mdb_smooth_loss: need W,H >= 2, 1 <= C <= 4, N <= 65535");
  MDB_REQUIRE(h, workspace && workspace_bytes >= mdb_smooth_loss_workspace_bytes(W, H, N) && ((uintptr_t)workspace & 255) == 0,
              MDB_E_WORKSPACE, "mdb_smooth_loss: workspace too small or misaligned");
  int rb = (H + SM_ROWS - 1) / SM_ROWS;
  const double kx = 1.0 / ((double)(W - 1) * H * N), ky = 1.0 / ((double)W * (H - 1) * N);
  cudaStream_t st = (cudaStream_t)stream;
  float* part = (float*)workspace;
  const bool vec4 = W % 4 == 0 && W >= 8 && (((uintptr_t)disparity | (uintptr_t)image | (uintptr_t)d_disparity) & 15) == 0;
  if (vec4) {
    // four columns per thread: a CTA covers 4 x threads columns; rows per CTA 16, or 8 when that fills the SMs better
    const int threads = std::min(SM_THREADS, ((W / 4 + 31) / 32) * 32);
    int rows = SM_ROWS;
    if ((long long)rb * N < (long long)h->num_sms * 8) rows = SM_ROWS_MIN;
    rb = (H + rows - 1) / rows;
    dim3 grid4(rb, N);
    switch (C) {
      case 1: smooth4_kernel<1><<<grid4, threads, 0, st>>>(disparity, image, W, H, N, rows, (float)kx, (float)ky, tie, part, d_disparity); break;
      case 2: smooth4_kernel<2><<<grid4, threads, 0, st>>>(disparity, image, W, H, N, rows, (float)kx, (float)ky, tie, part, d_disparity); break;
      case 3: smooth4_kernel<3><<<grid4, threads, 0, st>>>(disparity, image, W, H, N, rows, (float)kx, (float)ky, tie, part, d_disparity); break;
      default: smooth4_kernel<4><<<grid4, threads, 0, st>>>(disparity, image, W, H, N, rows, (float)kx, (float)ky, tie, part, d_disparity); break;
    }
    MDB_LAUNCHED(h);
    smooth_finalize_kernel<<<1, 256, 0, st>>>((const float*)workspace, rb * N, kx, ky, loss);
    MDB_LAUNCHED(h);
    return MDB_OK;
  }
  dim3 grid(rb, N);
  switch (C) {
    case 1: smooth_kernel<1><<<grid, SM_THREADS, 0, st>>>(disparity, image, W, H, N, (float)kx, (float)ky, tie, part, d_disparity); break;
    case 2: smooth_kernel<2><<<grid, SM_THREADS, 0, st>>>(disparity, image, W, H, N, (float)kx, (float)ky, tie, part, d_disparity); break;
    case 3: smooth_kernel<3><<<grid, SM_THREADS, 0, st>>>(disparity, image, W, H, N, (float)kx, (float)ky, tie, part, d_disparity); break;
    default: smooth_kernel<4><<<grid, SM_THREADS, 0, st>>>(disparity, image, W, H, N, (float)kx, (float)ky, tie, part, d_disparity); break;
  }
  MDB_LAUNCHED(h);
  smooth_finalize_kernel<<<1, 256, 0, (cudaStream_t)stream>>>((const float*)workspace, rb * N, kx, ky, loss);
  MDB_LAUNCHED(h);
  return MDB_OK;
}

}  // extern "C"

// ---- smooth_loss without any allocation -----------------------------------------------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same kernels and same result as
// mdb_smooth_loss_ex, but the per-block partial sums live in the file-scope array below instead of a caller-provided
// workspace.  Every GPU that loads the library holds these 256 KiB until the process exits.
//  * The array's NAME is a host-side stand-in in host code, not a device address: the address the kernels get is
//    looked up once per device, with that device current, by cudaGetSymbolAddress (smooth_scratch_ptr below).
//  * All calls on a device share the array and it keeps the last call's contents.  Nothing in it has to start at zero:
//    a call's kernels write every partial its finalize kernel then reads (2 floats per (row block, item)), so stale
//    contents are never read.  Calls that OVERLAP on the GPU (two streams, two host threads) race on it: serialise them.
constexpr size_t SMOOTH_NOALLOC_FLOATS = MDB_SMOOTH_NOALLOC_MAX_BYTES / sizeof(float);
static __device__ __align__(256) float g_smooth_scratch[SMOOTH_NOALLOC_FLOATS];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_smooth_scratch_mu;
float* g_smooth_scratch_ptr[kMaxDevices] = {nullptr};   // device address of g_smooth_scratch per device, nullptr = not looked up

// once per device: the symbol's address ON THAT DEVICE (cudaGetSymbolAddress answers for the current device)
cudaError_t smooth_scratch_ptr(int device, float** out) {
  std::lock_guard<std::mutex> lock(g_smooth_scratch_mu);
  if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
  if (!g_smooth_scratch_ptr[device]) {
    int prev = -1;
    cudaError_t e = cudaGetDevice(&prev);
    if (e != cudaSuccess) return e;
    if (prev != device && (e = cudaSetDevice(device)) != cudaSuccess) return e;
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_smooth_scratch);
    if (prev != device) cudaSetDevice(prev);
    if (e != cudaSuccess) return e;
    g_smooth_scratch_ptr[device] = (float*)p;
  }
  *out = g_smooth_scratch_ptr[device];
  return cudaSuccess;
}
}  // namespace

extern "C" int32_t mdb_smooth_loss_noalloc(mdb_handle* h, const float* disparity, const float* image, int32_t W, int32_t H,
                                           int32_t C, int32_t N, int32_t abs_subgradient, float* loss, float* d_disparity,
                                           void* stream) {
  MDB_REQUIRE(h, h != nullptr, MDB_E_INVALID_ARG, "null handle");
  MDB_REQUIRE(h, W >= 2 && H >= 2 && N >= 1, MDB_E_SHAPE, "mdb_smooth_loss_noalloc: need W,H >= 2, N >= 1");
  const size_t need = mdb_smooth_loss_workspace_bytes(W, H, N);
  MDB_REQUIRE(h, need != 0 && need <= MDB_SMOOTH_NOALLOC_MAX_BYTES, MDB_E_WORKSPACE,
              "mdb_smooth_loss_noalloc: N * ceil(H / 8) = %lld exceeds the built-in scratch (%d): this shape needs %zu bytes, the "
              "static array holds %d; use mdb_smooth_loss_ex with a workspace of your own",
              (long long)N * ((H + 7) / 8), (int)(MDB_SMOOTH_NOALLOC_MAX_BYTES / 8), need, (int)MDB_SMOOTH_NOALLOC_MAX_BYTES);
  float* scratch = nullptr;
  MDB_CUDA(h, smooth_scratch_ptr(h->device, &scratch));
  // the existing entry with the static array as its workspace: kernel launches only, no cudaMalloc / cudaFree anywhere
  return mdb_smooth_loss_ex(h, disparity, image, W, H, C, N, abs_subgradient, loss, d_disparity, scratch,
                            MDB_SMOOTH_NOALLOC_MAX_BYTES, stream);
}
