// photo_fused_host.inl -- finalize kernels, up-sample transpose and the C-ABI entry of the fused loss
#include <cmath>
#include <mutex>
namespace {

// test hooks (mdb_debug_option): process-global, read on the host only, default 0 = off
int g_dbg_strip_segs = 0;       // force this many row segments per strip
int g_dbg_upsample_split = 0;   // force the two-kernel V / H transpose (the W > 4096 path)
int g_dbg_strip_taper = 0;      // experiment 20: this many row segments of linearly decreasing length, long ones first
int32_t* g_dbg_strip_seg_dump = nullptr;   // tests: mdb_train_loss_strip_geometry also writes { taper, seg_y[0 .. 8] } here (host memory)
int g_dbg_strip_taper_pow10 = 20;   // lengths ~ (k - i)^(pow10 / 10): 20 = quadratic (shipped), measured 5 .. 40
signed char* g_dbg_decisions = nullptr;   // device buffer [Ls][N][2C+2][H][W]: run the DBG instantiation and dump into it

// ---------------------------------------------------------------------------------------
// 4. per-item finalize: fixed-order sum of the warp records (double); pose chain to rvec/tvec;
//    smoothness loss per scale and the mean-normalisation constant of its gradient
// ---------------------------------------------------------------------------------------
#ifndef MDB_FIN_THREADS
#define MDB_FIN_THREADS 256
#endif
constexpr int FIN_THREADS = MDB_FIN_THREADS;   // thread = record: all of a thread's loads are independent

struct FinArgs {
  float* d_rvec[MDB_MAX_SOURCES];
  float* d_tvec[MDB_MAX_SOURCES];
  float kx[MDB_MAX_SCALES], ky[MDB_MAX_SCALES];
  int Ls, N, W, H, normalize;
  double up;          // 1 / (W H Ng Ls): weight of the photometric sum in the loss
  float* loss_out;    // non-null only when N == 1: the scalar loss is written here by the item's own CTA
  // N > 1: the CTA that finishes last (done_counter, zeroed by prep_kernel) sums the items in fixed order
  unsigned* done_counter;
  float* loss_final;
  int adam_on;          // slow_depth: ADAM step of the poses + loss history here (mdb_pose_adam)
  mdb_pose_adam adam;
};

MDB_DEV float adam_bias_corr(double beta, int t) { return (float)(1.0 / (1.0 - pow(beta, (double)t))); }

template <int S>
__global__ void __launch_bounds__(FIN_THREADS)
finalize_item_kernel(const float* __restrict__ partials, int records, PoseArgs pa, FinArgs fa,
                     ScaleStat* __restrict__ stats, double* __restrict__ item_loss) {
  constexpr int NPART = STRIP_NHEAD + NPOSE * S;
  constexpr int NRED = 1 + NPOSE * S;   // photometric sum + pose sums, reduced over all records of the item
  const int n = blockIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __shared__ double red[FIN_THREADS / 32][NRED];
  __shared__ double tot[NRED];
  __shared__ double sred[FIN_THREADS / 32][MDB_MAX_SCALES][2];

  double acc[NRED];
#pragma unroll
  for (int k = 0; k < NRED; ++k) acc[k] = 0.0;
  for (int t = threadIdx.x; t < records; t += FIN_THREADS) {
    const float* p = partials + ((size_t)n * records + t) * NPART;
    acc[0] += (double)p[0];
#pragma unroll
    for (int k = 1; k < NRED; ++k) acc[k] += (double)p[STRIP_NHEAD - 1 + k];
  }
#pragma unroll
  for (int k = 0; k < NRED; ++k) {
    double v = acc[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(MDB_FULL, v, o);
    if (lane == 0) red[warp][k] = v;
  }
  // smoothness sums per scale (src/utils.jl:163-177, src/training.jl:64-67): records are ordered [..][scale]
  const int per_scale = records / fa.Ls;
#pragma unroll
  for (int sc = 0; sc < MDB_MAX_SCALES; ++sc) {
    if (sc < fa.Ls) {
      double sx = 0.0, sy = 0.0;
      for (int t = threadIdx.x; t < per_scale; t += FIN_THREADS) {
        const float* p = partials + ((size_t)n * records + (size_t)t * fa.Ls + sc) * NPART;
        sx += (double)p[1]; sy += (double)p[2];
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) { sx += __shfl_xor_sync(MDB_FULL, sx, o); sy += __shfl_xor_sync(MDB_FULL, sy, o); }
      if (lane == 0) { sred[warp][sc][0] = sx; sred[warp][sc][1] = sy; }
    }
  }
  __syncthreads();
  if (threadIdx.x >= 64 && threadIdx.x < 64 + NRED) {
    double v = 0.0;
    for (int w = 0; w < FIN_THREADS / 32; ++w) v += red[w][threadIdx.x - 64];
    tot[threadIdx.x - 64] = v;
  }
  __syncthreads();
  // from here two chains run side by side on different warps: thread 0 closes the smoothness terms and the
  // item's loss, threads 32 .. 32+S-1 pull the pose sums back to (rvec, tvec) -- ~550 dependent FP64 operations
  if (threadIdx.x == 0) {
    double sloss = 0.0;
    for (int sc = 0; sc < fa.Ls; ++sc) {
      double tx = 0.0, ty = 0.0;
      for (int w = 0; w < FIN_THREADS / 32; ++w) { tx += sred[w][sc][0]; ty += sred[w][sc][1]; }
      ScaleStat st = stats[(size_t)sc * fa.N + n];
      const double inv_mp = (double)st.inv_mp;
      const double l = ((double)fa.kx[sc] * tx + (double)fa.ky[sc] * ty) * inv_mp;
      st.loss = (float)l;
      sloss += (double)st.loss;
      // d/dd of S(d / (m + eps)): the part through the mean is the constant -S / ((m + eps) W H)
      st.cconst = fa.normalize ? (float)(-l * inv_mp / ((double)fa.W * fa.H)) : 0.0f;
      stats[(size_t)sc * fa.N + n] = st;
    }
    item_loss[n] = tot[0];
    if (fa.loss_out) {   // N == 1: same sum as below
      const float l = (float)(tot[0] * fa.up + sloss);
      *fa.loss_out = l;
      if (fa.adam_on && fa.adam.hist && *fa.adam.step < fa.adam.hist_len) fa.adam.hist[*fa.adam.step] = l;
    }
  }
  // slow_depth: the bias corrections of the NEXT step, on a warp of their own (two pow() beside the pose chain)
  if (fa.adam_on && n == 0 && warp == 2 && lane < 2) {
    const int t1 = *fa.adam.step + 2;
    fa.adam.cbuf[(t1 & 1) * 2 + lane] = adam_bias_corr(lane ? fa.adam.b2 : fa.adam.b1, t1);
  }
  if (threadIdx.x >= 32 && threadIdx.x < 32 + S) {
    const int s = threadIdx.x - 32;
    const double* dA = tot + 1 + s * NPOSE;  // row-major dL/dA
    const double* db = dA + 9;
    // A = K R' invK  ->  dR' = K^T dA invK^T ;  b = K t' -> dt' = K^T db
    double Kt[9], iKt[9], t1[9], dRp[9], dtp[3];
    mat3_T(pa.K, Kt);
    mat3_T(pa.invK, iKt);
    mat3_mul(Kt, dA, t1);
    mat3_mul(t1, iKt, dRp);
    for (int i = 0; i < 3; ++i) dtp[i] = Kt[3 * i] * db[0] + Kt[3 * i + 1] * db[1] + Kt[3 * i + 2] * db[2];
    double r[3], t[3], dr[3], dt[3];
    for (int k = 0; k < 3; ++k) { r[k] = pa.rvec[s][3 * n + k]; t[k] = pa.tvec[s][3 * n + k]; }
    compose_T_bwd_d(r, t, pa.invert[s] != 0, dRp, dtp, dr, dt);
    for (int k = 0; k < 3; ++k) {
      if (fa.d_rvec[s]) fa.d_rvec[s][3 * n + k] = (float)dr[k];
      if (fa.d_tvec[s]) fa.d_tvec[s][3 * n + k] = (float)dt[k];
    }
    if (fa.adam_on) {   // Flux ADAM (apply! + update!), the expressions of adam_kernel (optim.cu)
      const int t = *fa.adam.step + 1;
      float c1 = fa.adam.cbuf[(t & 1) * 2], c2 = fa.adam.cbuf[(t & 1) * 2 + 1];
      if (c1 == 0.0f || c2 == 0.0f) {   // first step after create / load_state: nothing handed on yet
        c1 = adam_bias_corr(fa.adam.b1, t); c2 = adam_bias_corr(fa.adam.b2, t);
      }
      const float b1 = (float)fa.adam.b1, b2 = (float)fa.adam.b2, eta = (float)fa.adam.eta, eps = (float)fa.adam.eps;
      auto step1 = [&](float* p, float* mm, float* vv, float g) {
        const float m = b1 * *mm + (1.0f - b1) * g;
        const float v = b2 * *vv + (1.0f - b2) * g * g;
        *mm = m; *vv = v;
        *p -= m * c1 / (sqrtf(v * c2) + eps) * eta;
      };
      for (int k = 0; k < 3; ++k) {
        const int j = 3 * n + k;
        step1(fa.adam.p_r[s] + j, fa.adam.m_r[s] + j, fa.adam.v_r[s] + j, (float)dr[k]);
        step1(fa.adam.p_t[s] + j, fa.adam.m_t[s] + j, fa.adam.v_t[s] + j, (float)dt[k]);
      }
    }
  }
  // 6. scalar loss = up * sum(item photometric sums) + sum(smoothness terms), by the last CTA to get here
  //    (fixed summation order whichever CTA that is: bit-reproducible)
  if (fa.done_counter) {
    __shared__ bool last;
    __syncthreads();   // thread 0's item_loss / stats stores precede its fence below
    if (threadIdx.x == 0) {
      __threadfence();
      last = atomicAdd(fa.done_counter, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (last) {
      __threadfence();
      double v = 0.0;
      for (int i = threadIdx.x; i < fa.N; i += FIN_THREADS) v += __ldcg(item_loss + i) * fa.up;
      for (int i = threadIdx.x; i < fa.Ls * fa.N; i += FIN_THREADS) v += (double)__ldcg(&stats[i].loss);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(MDB_FULL, v, o);
      if (lane == 0) red[warp][0] = v;
      __syncthreads();
      if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < FIN_THREADS / 32; ++w) t += red[w][0];
        *fa.loss_final = (float)t;
        if (fa.adam_on) {   // every CTA has read the step counter before it took its ticket
          if (fa.adam.hist && *fa.adam.step < fa.adam.hist_len) fa.adam.hist[*fa.adam.step] = (float)t;
          *fa.adam.step += 1;
        }
      }
    }
  } else if (fa.adam_on) {   // N == 1: the one CTA advances its step counter once all its warps have read it
    __syncthreads();
    if (threadIdx.x == 0) *fa.adam.step += 1;
  }
}

// ---------------------------------------------------------------------------------------
// 5. transpose of upsample_bilinear (align-corners), separable and atomic-free (deterministic):
//    V: tmp[jy][px] = sum_py wy(py, jy) (g[py][px] + c)   thread = full-res column (coalesced, the
//       <= 2/r + 1 row loads of a thread are independent), and the constant alone for native scales
//    H: out[jy][jx] = sum_px wx(px, jx) tmp[jy][px]       thread = native pixel, on the small tmp
//    The weight of full-res index p for native index j is the hat function max(0, 1 - |p r - j|),
//    r = (dn-1)/(Dn-1), evaluated on the exact integer numerator p (dn-1) - j (Dn-1).
// ---------------------------------------------------------------------------------------
struct UpArgs {
  const float* gfull[MDB_MAX_SCALES];   // null for native scales
  float* gnat[MDB_MAX_SCALES];
  float* tmp[MDB_MAX_SCALES];           // [N][dh][W] per low-res scale
  int dw[MDB_MAX_SCALES], dh[MDB_MAX_SCALES];
  int row_start[MDB_MAX_SCALES + 1];    // prefix sums of dh: blockIdx.y -> (scale, native row)
  int Ls, W, H, N;
  int hp;                               // half-pixel up-sampling (upsample_align_corners == 0)
};

constexpr int UP_ROWS = 4;   // (scale, native row) pairs per CTA: keeps the CTA count in the thousands

__global__ void __launch_bounds__(256)
upsample_bwd_v_kernel(const __grid_constant__ UpArgs a, const ScaleStat* __restrict__ stats) {
  const int n = blockIdx.y, W = a.W, H = a.H;
  const int rows_total = a.row_start[a.Ls];
  for (int rr = blockIdx.x * UP_ROWS; rr < min(rows_total, (int)(blockIdx.x + 1) * UP_ROWS); ++rr) {
    int sc = 0;
    while (sc + 1 < a.Ls && rr >= a.row_start[sc + 1]) ++sc;
    const int jy = rr - a.row_start[sc], dh = a.dh[sc];
    const float cc = stats[(size_t)sc * a.N + n].cconst;
    const float* g = a.gfull[sc];
    if (g == nullptr) {   // native: the strip kernel wrote the gradient in place; add the constant
      if (cc != 0.0f) {
        float* o = a.gnat[sc] + ((size_t)n * H + jy) * W;
        for (int px = threadIdx.x; px < W; px += 256) o[px] += cc;
      }
      continue;
    }
    g += (size_t)n * W * H;
    const UpAxis uy = up_axis(dh, H, a.hp);
    const float ih = 1.0f / (float)uy.D;
    int py_lo, py_hi;
    up_range(uy, jy, py_lo, py_hi);
    float* t = a.tmp[sc] + ((size_t)n * dh + jy) * W;
    // four columns per thread with independent accumulators: 4 x 4 loads in flight per thread
    for (int px0 = threadIdx.x; px0 < W; px0 += 4 * 256) {
      float acc[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll 4
      for (int py = py_lo; py <= py_hi; ++py) {
        const float wy = fmaxf(1.0f - (float)llabs(up_num(uy, py) - jy * uy.D) * ih, 0.0f);
        const float* row = g + (size_t)py * W;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int px = px0 + k * 256;
          if (px < W) acc[k] = fmaf(wy, ldg(row + px) + cc, acc[k]);
        }
      }
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (px0 + k * 256 < W) t[px0 + k * 256] = acc[k];
    }
  }
}

__global__ void __launch_bounds__(256)
upsample_bwd_h_kernel(const __grid_constant__ UpArgs a) {
  const int n = blockIdx.y, W = a.W;
  const int rows_total = a.row_start[a.Ls];
  for (int rr = blockIdx.x * UP_ROWS; rr < min(rows_total, (int)(blockIdx.x + 1) * UP_ROWS); ++rr) {
    int sc = 0;
    while (sc + 1 < a.Ls && rr >= a.row_start[sc + 1]) ++sc;
    if (a.gfull[sc] == nullptr) continue;
    const int jy = rr - a.row_start[sc], dw = a.dw[sc], dh = a.dh[sc];
    const float* t = a.tmp[sc] + ((size_t)n * dh + jy) * W;
    const UpAxis ux = up_axis(dw, W, a.hp);
    const float iw = 1.0f / (float)ux.D;
    for (int jx = threadIdx.x; jx < dw; jx += 256) {
      int px_lo, px_hi;
      up_range(ux, jx, px_lo, px_hi);
      float acc = 0.0f;
#pragma unroll 4
      for (int px = px_lo; px <= px_hi; ++px) {
        const float wx = fmaxf(1.0f - (float)llabs(up_num(ux, px) - jx * ux.D) * iw, 0.0f);
        acc = fmaf(wx, ldg(t + px), acc);
      }
      a.gnat[sc][((size_t)n * dh + jy) * dw + jx] = acc;
    }
  }
}

// 5b. The same transpose as ONE streaming kernel (W <= 256 KC, KC <= 16): a CTA owns a band of native rows
//     of one (item, scale), walks down the full-res rows that touch them (each element is fetched once by
//     the CTA; neighbouring bands re-read 2/r rows of each other, from L2), keeps the running vertical sums
//     of the two native rows under the current full-res row in registers (thread = KC full-res columns),
//     and every time a native row is complete hands it through shared memory to the horizontal pass.
//     No tmp[dh][W] round trip and no second launch; summation order per output element is the one of
//     the V and H kernels above (ascending py, then ascending px), which stay as the path for W > 4096.
struct UpBandArgs {
  const float* gfull[MDB_MAX_SCALES];
  float* gnat[MDB_MAX_SCALES];
  int dw[MDB_MAX_SCALES], dh[MDB_MAX_SCALES];
  int band[MDB_MAX_SCALES];             // native rows per CTA (0: native scale)
  int blk_start[MDB_MAX_SCALES + 1];    // blockIdx.x -> (scale, band); native scales own H / UPB_NATIVE_ROWS blocks
  int Ls, W, H, N;
};
constexpr int UPB_NATIVE_ROWS = 8;

#ifndef UPB_ROWS
#define UPB_ROWS 8      // full-res rows per batch of loads (independent 128-bit loads in flight per thread)
#endif
#ifndef UPB_MINB
#define UPB_MINB 4
#endif
#ifndef UPB_NG
#define UPB_NG 1        // 128-bit column groups per thread at W <= 1024 (the CTA has 256 / UPB_NG threads); measured
                        // 1 / 2 / 4 -> segment 0.108 / 0.124 / 0.196 ms: parallelism matters more than the per-row bookkeeping
#endif
#ifndef UPB_BAND_ROWS
#define UPB_BAND_ROWS 4   // measured 4 / 6 / 8 -> segment 0.102 / 0.108 / 0.109 ms at config 5 (more CTAs beat less overlap)
#endif
constexpr int UPB_BAND = UPB_BAND_ROWS;     // native rows per CTA at most (accumulators of the horizontal pass; <= 8)
constexpr int UPB_SMEM_BYTES = 32 * 1024;   // the band's vertically summed rows: band x W floats

// G = floats per load (4: 128-bit, needs W % 4 == 0), NG = column groups per thread, TPB threads: W <= TPB G NG.
template <int G, int NG, int TPB>
__global__ void __launch_bounds__(TPB, (G * NG * TPB <= 1024) ? (UPB_MINB * 256 / TPB) : 1)
upsample_bwd_band_kernel(const __grid_constant__ UpBandArgs a, const ScaleStat* __restrict__ stats) {
  extern __shared__ float4 s_rows4[];   // [band][W]
  float* s_rows = reinterpret_cast<float*>(s_rows4);
  const int n = blockIdx.y, W = a.W, H = a.H, tid = threadIdx.x;
  int sc = 0;
  while (sc + 1 < a.Ls && (int)blockIdx.x >= a.blk_start[sc + 1]) ++sc;
  const int blk = blockIdx.x - a.blk_start[sc];
  const float cc = stats[(size_t)sc * a.N + n].cconst;
  if (a.band[sc] == 0) {   // native: the strip kernel wrote the gradient in place; add the constant
    if (cc == 0.0f) return;
    const int r0 = blk * UPB_NATIVE_ROWS, r1 = min(H, r0 + UPB_NATIVE_ROWS);
    float* o = a.gnat[sc] + ((size_t)n * H + r0) * W;
    const int cnt = (r1 - r0) * W;
    if (G == 4) {
      float4* o4 = reinterpret_cast<float4*>(o);
      for (int i = tid; i < cnt / 4; i += TPB) {
        float4 q = o4[i];
        q.x += cc; q.y += cc; q.z += cc; q.w += cc;
        o4[i] = q;
      }
    } else {
      for (int i = tid; i < cnt; i += TPB) o[i] += cc;
    }
    return;
  }
  const int dw = a.dw[sc], dh = a.dh[sc], dy1 = dh - 1, H1 = H - 1, dx1 = dw - 1, W1 = W - 1;
  const float ih = 1.0f / (float)H1, iw = 1.0f / (float)W1;
  const int j0 = blk * a.band[sc], j1 = min(dh, j0 + a.band[sc]);   // native rows [j0, j1) of this CTA
  // full-res rows whose lower native neighbour floor(py dy1 / H1) lies in [j0 - 1, j1 - 1]
  const int py_s = j0 > 0 ? ((j0 - 1) * H1 + dy1 - 1) / dy1 : 0;
  const int py_e = min(H, (j1 * H1 + dy1 - 1) / dy1);
  const float* g = a.gfull[sc] + (size_t)n * W * H;

  // ---- vertical pass: thread = G NG full-res columns, no barrier; lo / hi are the running sums of the
  // native rows jcur and jcur + 1 = the two neighbours of the current full-res row.  (jcur, rem) = floor
  // and remainder of py dy1 / H1, carried down the rows without divisions.
  float lo[NG][G], hi[NG][G];
#pragma unroll
  for (int q = 0; q < NG; ++q)
#pragma unroll
    for (int e = 0; e < G; ++e) lo[q][e] = hi[q][e] = 0.0f;
  auto park = [&](int j) {   // row j is complete: into the band's buffer
    if (j < j0 || j >= j1) return;
#pragma unroll
    for (int q = 0; q < NG; ++q) {
      const int px = G * (tid + TPB * q);
      if (px < W) {
        if (G == 4) s_rows4[((j - j0) * W + px) >> 2] = make_float4(lo[q][0], lo[q][1], lo[q][2], lo[q][3]);
        else s_rows[(j - j0) * W + px] = lo[q][0];
      }
    }
  };
  constexpr int R = ((G == 4 ? UPB_ROWS : 2 * UPB_ROWS) / NG) > 0 ? ((G == 4 ? UPB_ROWS : 2 * UPB_ROWS) / NG) : 1;
  int jcur = (py_s * dy1) / H1, rem = py_s * dy1 - jcur * H1;
  for (int pb = py_s; pb < py_e; pb += R) {
    float v[R][NG][G];
#pragma unroll
    for (int i = 0; i < R; ++i) {
      const float* row = g + (size_t)(pb + i) * W;
#pragma unroll
      for (int q = 0; q < NG; ++q) {
        const int px = G * (tid + TPB * q);
        const bool on = pb + i < py_e && px < W;
        if (G == 4) {
          const float4 t = on ? __ldg(reinterpret_cast<const float4*>(row + px)) : make_float4(0.f, 0.f, 0.f, 0.f);
          v[i][q][0] = t.x; v[i][q][1] = t.y; v[i][q][2] = t.z; v[i][q][3] = t.w;
        } else {
          v[i][q][0] = on ? ldg(row + px) : 0.0f;
        }
      }
    }
#pragma unroll
    for (int i = 0; i < R; ++i) {
      if (pb + i < py_e) {   // CTA-uniform
        if (rem >= H1) {     // moved on by one native row (dh <= H): row jcur is complete
          park(jcur);
#pragma unroll
          for (int q = 0; q < NG; ++q)
#pragma unroll
            for (int e = 0; e < G; ++e) { lo[q][e] = hi[q][e]; hi[q][e] = 0.0f; }
          ++jcur;
          rem -= H1;
        }
        // the hat weights of upsample_bwd_v_kernel, for j = jcur and j = jcur + 1
        const float wlo = fmaxf(1.0f - (float)rem * ih, 0.0f), whi = fmaxf(1.0f - (float)(H1 - rem) * ih, 0.0f);
#pragma unroll
        for (int q = 0; q < NG; ++q)
#pragma unroll
          for (int e = 0; e < G; ++e) {
            const float x = v[i][q][e] + cc;
            lo[q][e] = fmaf(wlo, x, lo[q][e]); hi[q][e] = fmaf(whi, x, hi[q][e]);
          }
        rem += dy1;
      }
    }
  }
  park(jcur);   // the last row the band's rows end in; its upper neighbour jcur + 1 belongs to the next band
  __syncthreads();

  // ---- horizontal pass: thread = native column jx of ALL the band's rows, so the tap range and the hat
  // weights (which do not depend on the row) are formed once per column and tap
  const int nrows = j1 - j0;
  float* out = a.gnat[sc] + ((size_t)n * dh + j0) * dw;
  for (int jx = tid; jx < dw; jx += TPB) {
    const int p0 = max(0, ((jx - 1) * W1) / dx1), p1 = min(W1, ((jx + 1) * W1 + dx1 - 1) / dx1);
    float acc[UPB_BAND];
#pragma unroll
    for (int r = 0; r < UPB_BAND; ++r) acc[r] = 0.0f;
    int num = p0 * dx1 - jx * W1;
    for (int px = p0; px <= p1; ++px, num += dx1) {
      const float wx = fmaxf(1.0f - (float)abs(num) * iw, 0.0f);
#pragma unroll
      for (int r = 0; r < UPB_BAND; ++r)
        if (r < nrows) acc[r] = fmaf(wx, s_rows[r * W + px], acc[r]);
    }
#pragma unroll
    for (int r = 0; r < UPB_BAND; ++r)
      if (r < nrows) out[r * dw + jx] = acc[r];
  }
}

struct FusedLayout {
  size_t o_ppd, o_meanpart, o_stats, o_partials, o_itemloss, o_counter, o_gfull[MDB_MAX_SCALES], o_tmp[MDB_MAX_SCALES], total;
  int rowblocks, npart;
  StripGeom strip;   // decomposition of the streaming kernel
  int records;       // partial-sum records per item = warps per item
};

// Rows per segment.  Every warp costs (rows + 4 halo rows) row-times, and the launch runs in w = warps / resident
// waves of 13 warps per SM.  Measured at 1024x320 for 8 / 16 / 32 / 64 items and 1 .. 20 segments (tools/sweep_segs.py,
// profiles/r2_strip_summary.md): below ~3 waves the waves are what the clock sees, time = ceil(w) x (rows + 4) -- a
// 1.2-wave launch is as slow as a 2-wave one; above it the block scheduler smooths them out and what is left is the
// tail, about half a segment: time = (w + 0.5) x (rows + 4).  The segment count minimising that model is within 0.3 %
// of the measured best at all four batch sizes (3 / 5 / 4 / 3 segments), where the old "16 resident sets" rule lost
// 1.5 - 2.5 % below batch 64.  Segments never go below 8 rows unless one wave cannot be filled otherwise (config 1).
inline StripGeom strip_geom(const mdb_loss_io* io, int num_sms) {
  StripGeom g{};
  g.nstrips = (io->W + STRIP_PX - 1) / STRIP_PX;
  const int sms = num_sms > 0 ? num_sms : 148;
  const long long base = (long long)io->N * io->Ls * g.nstrips;
  const double resident = (double)(65536 / (STRIP_MAXREG * 32)) * sms;   // warps the register file holds per SM
  int best_n = 1;
  if (g_dbg_strip_segs > 0) {
    best_n = std::min(g_dbg_strip_segs, std::max(1, io->H / 2));
  } else {
    double best_cost = 1e300;
    const int n_max = std::max(1, io->H / 4);
    for (int n = 1; n <= n_max; ++n) {
      const int rows = (io->H + n - 1) / n;
      const int ne = (io->H + rows - 1) / rows;          // segments this row count really gives
      if (ne != n) continue;
      const double w = (double)(base * ne) / resident;
      if (rows < 8 && w > 1.0) break;                      // 4-row segments only to fill the first wave
      const double cost = (w >= 3.0 ? w + 0.5 : std::ceil(w)) * (rows + 4);
      if (cost < best_cost * 0.999) { best_cost = cost; best_n = n; }   // ties: fewer, longer segments
    }
  }
  g.seg_rows = (io->H + best_n - 1) / best_n;
  g.nsegs = (io->H + g.seg_rows - 1) / g.seg_rows;
  // Tapered segments (round 2, experiment 20): lengths proportional to k^2, (k-1)^2, ..., 1, launched long ones first, so that the
  // launch ends evenly instead of with a tail of full-length warps.  Measured at 1024x320 with L2 flushed (tools/sweep_taper.py,
  // fused kernel ms, equal -> tapered): 8 items 0.585 -> 0.569 (k = 6), 16: 1.106 -> 1.082 (k = 4), 32: 2.138 -> 2.099 (k = 3),
  // 64: 4.207 -> 4.144 (k = 3); 12 items (exactly one resident wave of whole-height warps) and <= 4 items are better off
  // with equal segments.  Those figures are for LINEAR lengths; quadratic ones (tools/sweep_taper_shape.py) take another 2.3 % at
  // 8 items (0.567 -> 0.554), 3 % at 16, 0.2 % at 64.  k by the number of resident waves ONE segment index fills.
  // "strip_taper" = -1 switches it off.
  int k_taper = 0;
  if (g_dbg_strip_taper >= 2) k_taper = g_dbg_strip_taper;
  else if (g_dbg_strip_taper == 0 && g_dbg_strip_segs == 0) {
    const double wb = (double)base / resident;
    // 1.5 and not 2: 20 .. 28 items at 1024x320 and config 3 (32 x 640x192, wb = 1.66: 0.798 against 0.809 ms) prefer three segments
    k_taper = wb >= 1.5 ? 3 : (wb >= 1.2 ? 4 : ((wb >= 0.55 && wb < 0.9) ? 6 : 0));
  }
#if STRIP_TAPER
  if (k_taper >= 2 && k_taper <= STRIP_MAX_TAPER && io->H >= 16 * k_taper) {
    const int k = k_taper;
    const double pw = g_dbg_strip_taper_pow10 / 10.0;
    double tot = 0.0, acc = 0.0;
    for (int i = 0; i < k; ++i) tot += std::pow((double)(k - i), pw);
    g.taper = 1; g.nsegs = k; g.seg_y[0] = 0;
    for (int i = 0; i < k; ++i) {
      acc += std::pow((double)(k - i), pw);
      g.seg_y[i + 1] = (i == k - 1) ? io->H : std::min(io->H - 4 * (k - 1 - i), std::max(g.seg_y[i] + 4, (int)(io->H * acc / tot)));
    }
    g.seg_rows = g.seg_y[1];
  }
#else
  (void)k_taper;
#endif
  g.total = (int)(base * g.nsegs);
  return g;
}

FusedLayout fused_layout(const mdb_loss_io* io, int num_sms = 148) {
  FusedLayout L{};
  L.strip = strip_geom(io, num_sms);
  L.rowblocks = (io->H + MEAN_ROWS - 1) / MEAN_ROWS;
  L.npart = STRIP_NHEAD + NPOSE * io->S;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = mdb_align_up(off + bytes, 256); return o; };
  L.o_ppd = take(sizeof(PoseStrip) * io->N * io->S);
  L.o_meanpart = take(sizeof(double) * (size_t)io->N * L.rowblocks * io->Ls);
  L.o_stats = take(sizeof(ScaleStat) * (size_t)io->Ls * io->N);
  L.records = L.strip.nsegs * L.strip.nstrips * io->Ls;
  L.o_partials = take(sizeof(float) * (size_t)io->N * L.records * L.npart);
  L.o_itemloss = take(sizeof(double) * io->N);
  L.o_counter = take(sizeof(unsigned));
  for (int s = 0; s < io->Ls; ++s) {
    const bool native = io->dw[s] == io->W && io->dh[s] == io->H;
    L.o_gfull[s] = native ? 0 : take(sizeof(float) * (size_t)io->N * io->W * io->H);
    L.o_tmp[s] = native ? 0 : take(sizeof(float) * (size_t)io->N * io->W * io->dh[s]);
  }
  L.total = off;
  return L;
}

bool io_shape_ok(const mdb_loss_io* io) {
  if (!io || io->W < 2 || io->H < 2 || io->N < 1 || io->N > 65535 || io->L < 1) return false;   // N rides in gridDim.y/z
  if (io->C != 1 && io->C != 3) return false;
  if ((long long)io->W * io->H * io->C * io->L >= (1ll << 31)) return false;   // 32-bit offsets inside an item
  if (io->S < 1 || io->S > MDB_MAX_SOURCES || io->Ls < 1 || io->Ls > MDB_MAX_SCALES) return false;
  for (int s = 0; s < io->Ls; ++s)
    if (io->dw[s] < 2 || io->dh[s] < 2 || io->dw[s] > io->W || io->dh[s] > io->H) return false;
  return true;
}

#if STRIP_CONSTPOSE
// Regions of c_pose (per device): handed out round-robin; an event per region orders "the last kernel that read it" before
// "the next copy that overwrites it", also across streams.  Calls made under stream capture do not use the bank at all
// (launch_fused picks the global-memory instantiation), so a graph never holds a region.
struct PosePool {
  std::mutex mu;
  cudaEvent_t ev[STRIP_CONST_POOL] = {nullptr};
  bool used[STRIP_CONST_POOL] = {false};
  int next = 0;
};
PosePool* pose_pool(int device) {
  static PosePool pools[64];
  return &pools[device & 63];
}
#endif

template <int C, int S>
int32_t launch_fused(mdb_handle* h, const FusedArgs& fa, const FusedLayout& L, cudaStream_t st) {
  constexpr size_t smem = ((size_t)3 * (S * (2 * C + 2) / 4) + (STRIP_SMEM_RING ? 2 * ((C * S + 3) / 4) + 2 * ((3 * C + 3) / 4) : 0)) *
                          STRIP_THREADS * sizeof(float4);
  StripGeom g = L.strip;
  const int per_item = g.nsegs * g.nstrips * fa.Ls;
#if STRIP_CONSTPOSE
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  MDB_CUDA(h, cudaStreamIsCapturing(st, &cap));
  // under capture (slow_depth's per-iteration graph) the constants are read from global memory: the graph then holds
  // the kernel alone, no copy node and no event nodes (config 1: 30.1 -> 22.7 us per replayed iteration)
  const bool use_bank = cap == cudaStreamCaptureStatusNone && !fa.dbg && !fa.hp;
  const int chunk = use_bank ? STRIP_CONST_CHUNK / S : fa.N;
  PosePool* pool = pose_pool(h->device);
#else
  const bool use_bank = false;
  const int chunk = fa.N;
#endif
  MDB_REQUIRE(h, !fa.x_u8 || use_bank, MDB_E_INVALID_ARG,
              "x_format = 1 (uint8 frames) is served by the throughput instantiation only: not under stream capture, not with "
              "upsample_align_corners = 0, not with the decision dump");
  for (int n0 = 0; n0 < fa.N; n0 += chunk) {
    const int items = std::min(chunk, fa.N - n0);
    g.n0 = n0;
#if STRIP_TAPER || STRIP_TAPER_FIELDS
    g.items = items;
#endif
    g.total = items * per_item;
    g.cbase = 0;
    const int blocks = (g.total + STRIP_WARPS - 1) / STRIP_WARPS;
#if STRIP_CONSTPOSE
    if (use_bank) {
      int region;
      cudaEvent_t ev;
      {
        std::lock_guard<std::mutex> lock(pool->mu);
        region = pool->next;
        pool->next = (pool->next + 1) % STRIP_CONST_POOL;
        if (!pool->ev[region]) MDB_CUDA(h, cudaEventCreateWithFlags(&pool->ev[region], cudaEventDisableTiming));
        ev = pool->ev[region];
        if (pool->used[region]) MDB_CUDA(h, cudaStreamWaitEvent(st, ev, 0));
        pool->used[region] = true;
      }
      g.cbase = region * STRIP_CONST_CHUNK;
      MDB_CUDA(h, cudaMemcpyToSymbolAsync(c_pose, fa.ppd + (size_t)n0 * S, sizeof(PoseStrip) * items * S,
                                          sizeof(PoseStrip) * g.cbase, cudaMemcpyDeviceToDevice, st));
      if (fa.x_u8) photo_strip_kernel<C, S, false, true, true><<<blocks, STRIP_THREADS, smem, st>>>(fa, g);
      else photo_strip_kernel<C, S, false, true><<<blocks, STRIP_THREADS, smem, st>>>(fa, g);
      MDB_LAUNCHED(h);
      MDB_CUDA(h, cudaEventRecord(ev, st));
      continue;
    }
#endif
    if (fa.dbg) photo_strip_kernel<C, S, true, false><<<blocks, STRIP_THREADS, smem, st>>>(fa, g);
    else photo_strip_kernel<C, S, false, false><<<blocks, STRIP_THREADS, smem, st>>>(fa, g);
    MDB_LAUNCHED(h);
  }
  return MDB_OK;
}

}  // namespace

extern "C" {

int32_t mdb_debug_option(const char* name, int64_t value) {
  if (!name) return MDB_E_INVALID_ARG;
  if (!strcmp(name, "strip_segs")) g_dbg_strip_segs = (int)value;
  else if (!strcmp(name, "upsample_split")) g_dbg_upsample_split = (int)value;
  else if (!strcmp(name, "decision_dump")) g_dbg_decisions = (signed char*)(intptr_t)value;
  else if (!strcmp(name, "render_precise")) mdb_dbg_render_precise = (int)value;
  else if (!strcmp(name, "strip_taper")) g_dbg_strip_taper = (int)value;
  else if (!strcmp(name, "strip_seg_dump")) g_dbg_strip_seg_dump = reinterpret_cast<int32_t*>((uintptr_t)value);
  else if (!strcmp(name, "strip_taper_pow10")) g_dbg_strip_taper_pow10 = value > 0 ? (int)value : 20;
  else return MDB_E_INVALID_ARG;
  return MDB_OK;
}

int32_t mdb_train_loss_strip_geometry(const mdb_loss_io* io, int32_t* out4) {
  if (!io_shape_ok(io) || !out4) return MDB_E_INVALID_ARG;
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const StripGeom g = strip_geom(io, sms);
  out4[0] = g.nstrips; out4[1] = g.nsegs; out4[2] = g.seg_rows; out4[3] = g.total;
  if (g_dbg_strip_seg_dump) {
    g_dbg_strip_seg_dump[0] = g.taper;
    for (int i = 0; i <= STRIP_MAX_TAPER; ++i) g_dbg_strip_seg_dump[1 + i] = g.taper ? g.seg_y[i] : 0;
  }
  return MDB_OK;
}

size_t mdb_train_loss_workspace_bytes(const mdb_loss_io* io) {
  if (!io_shape_ok(io)) return 0;
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  return fused_layout(io, sms).total;
}

int32_t mdb_train_loss_fwdbwd(mdb_handle* h, const mdb_loss_io* io, const mdb_loss_params* p,
                              void* workspace, size_t workspace_bytes, void* stream) {
  MDB_REQUIRE(h, h != nullptr, MDB_E_INVALID_ARG, "null handle");
  MDB_REQUIRE(h, io && p, MDB_E_INVALID_ARG, "mdb_train_loss_fwdbwd: null io/params");
  MDB_REQUIRE(h, io_shape_ok(io), MDB_E_SHAPE,
              "mdb_train_loss_fwdbwd: unsupported shape (need W,H>=2, C in {1,3}, S<=%d, Ls<=%d, 2<=dw<=W)",
              MDB_MAX_SOURCES, MDB_MAX_SCALES);
  MDB_REQUIRE(h, io->x && io->loss, MDB_E_INVALID_ARG, "mdb_train_loss_fwdbwd: null x/loss");
  MDB_REQUIRE(h, io->target_id >= 1 && io->target_id <= io->L, MDB_E_INVALID_ARG, "target_id out of range");
  MDB_REQUIRE(h, p->upsample_align_corners == 0 || p->upsample_align_corners == 1, MDB_E_INVALID_ARG,
              "upsample_align_corners must be 1 (NNlib's default) or 0 (half-pixel centres)");
  MDB_REQUIRE(h, !p->automasking || io->auto_loss, MDB_E_INVALID_ARG, "automasking needs auto_loss");
  for (int s = 0; s < io->S; ++s) {
    MDB_REQUIRE(h, io->source_ids[s] >= 1 && io->source_ids[s] <= io->L, MDB_E_INVALID_ARG, "source id out of range");
    MDB_REQUIRE(h, io->rvec[s] && io->tvec[s], MDB_E_INVALID_ARG, "null pose pointer");
  }
  for (int s = 0; s < io->Ls; ++s)
    MDB_REQUIRE(h, io->disparity[s] && io->d_disparity[s], MDB_E_INVALID_ARG, "null disparity pointer (scale %d)", s);
  const FusedLayout L = fused_layout(io, h->num_sms);
  MDB_REQUIRE(h, workspace && workspace_bytes >= L.total && ((uintptr_t)workspace & 255) == 0, MDB_E_WORKSPACE,
              "workspace must be >= %zu bytes and 256-byte aligned", L.total);
  cudaStream_t st = (cudaStream_t)stream;
  char* ws = (char*)workspace;
  const int N = io->N, S = io->S, Ls = io->Ls, W = io->W, H = io->H;
  const int Ng = p->n_global > 0 ? p->n_global : N;

  PoseArgs pa{};
  pa.N = N; pa.S = S;
  for (int s = 0; s < S; ++s) { pa.rvec[s] = io->rvec[s]; pa.tvec[s] = io->tvec[s]; pa.invert[s] = io->invert[s]; }
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) { pa.K[3 * i + j] = io->K[3 * j + i]; pa.invK[3 * i + j] = io->invK[3 * j + i]; }

  FusedArgs fa{};
  FinArgs fin{};
  UpArgs ua{};
  fa.x = io->x; fa.W = W; fa.H = H; fa.L = io->L; fa.N = N;
  fa.tgt = io->target_id - 1;
  fa.hp = p->upsample_align_corners == 0 ? 1 : 0;
  fa.x_u8 = io->x_format == 1 ? 1 : 0;
  ua.hp = fa.hp;
  for (int s = 0; s < S; ++s) fa.src[s] = io->source_ids[s] - 1;
  fa.Ls = Ls;
  for (int s = 0; s < Ls; ++s) {
    fa.disp[s] = io->disparity[s]; fa.dw[s] = io->dw[s]; fa.dh[s] = io->dh[s];
    fa.native[s] = (io->dw[s] == W && io->dh[s] == H) ? 1 : 0;
    if (fa.hp) {   // half-pixel centres: coordinate = max(0, (p + 0.5) dn / Dn - 0.5)
      fa.rxd[s] = (double)io->dw[s] / (double)W; fa.oxd[s] = 0.5 * fa.rxd[s] - 0.5;
      fa.ryd[s] = (double)io->dh[s] / (double)H; fa.oyd[s] = 0.5 * fa.ryd[s] - 0.5;
    } else {
      fa.rxd[s] = (double)(io->dw[s] - 1) / (double)(W - 1);
      fa.ryd[s] = (double)(io->dh[s] - 1) / (double)(H - 1);
    }
    fa.rx[s] = (float)fa.rxd[s];
    fa.ry[s] = (float)fa.ryd[s];
    fa.gout[s] = fa.native[s] ? io->d_disparity[s] : (float*)(ws + L.o_gfull[s]);
    const double wsm = p->smooth_weight_mode == 0 ? (double)p->disparity_smoothness * (double)io->scale[s] : 1.0;
    fa.kx[s] = (float)(wsm / ((double)Ls * (double)(W - 1) * H * Ng));
    fa.ky[s] = (float)(wsm / ((double)Ls * (double)W * (H - 1) * Ng));
    fin.kx[s] = fa.kx[s]; fin.ky[s] = fa.ky[s];
    ua.gfull[s] = fa.native[s] ? nullptr : fa.gout[s];
    ua.gnat[s] = io->d_disparity[s];
    ua.tmp[s] = fa.native[s] ? nullptr : (float*)(ws + L.o_tmp[s]);
    ua.dw[s] = io->dw[s]; ua.dh[s] = io->dh[s];
    ua.row_start[s + 1] = ua.row_start[s] + io->dh[s];
  }
  fa.auto_loss = p->automasking ? io->auto_loss : nullptr;
  fa.automask = p->automasking ? 1 : 0;
  fa.mind_d = 1.0 / p->max_depth;
  fa.dd_d = 1.0 / p->min_depth - 1.0 / p->max_depth;
  fa.mind = (float)fa.mind_d;
  fa.dd = (float)fa.dd_d;
  fa.ppd = (const PoseStrip*)(ws + L.o_ppd);
  fa.alpha = p->alpha; fa.c1 = p->ssim_c1; fa.c2 = p->ssim_c2;
  fa.abs_tie = p->abs_subgradient ? 1.0f : 0.0f;
  const double up = 1.0 / ((double)W * H * Ng * Ls);
  fa.up = (float)up;
  fa.ksu = p->alpha / (9.0f * io->C) * fa.up;
  fa.klu = (1.0f - p->alpha) / io->C * fa.up;
  fa.kl = (1.0f - p->alpha) / io->C;
  fa.ka = p->alpha / io->C;
  fa.stats = (const ScaleStat*)(ws + L.o_stats);
  fa.partials = (float*)(ws + L.o_partials);
  fa.vis = io->vis_warp_loss;
  fa.dbg = g_dbg_decisions;
  for (int s = 0; s < S; ++s) { fin.d_rvec[s] = io->d_rvec[s]; fin.d_tvec[s] = io->d_tvec[s]; }
  fin.Ls = Ls; fin.N = N; fin.W = W; fin.H = H; fin.normalize = p->normalize_disparity;
  fin.up = up;
  fin.loss_out = (N == 1) ? io->loss : nullptr;
  fin.done_counter = (N == 1) ? nullptr : (unsigned*)(ws + L.o_counter);
  fin.loss_final = io->loss;
  fin.adam_on = h->pose_adam ? 1 : 0;
  if (h->pose_adam) fin.adam = *h->pose_adam;
  ua.Ls = Ls; ua.W = W; ua.H = H; ua.N = N;
  ScaleStat* stats = (ScaleStat*)(ws + L.o_stats);
  double* mean_part = (double*)(ws + L.o_meanpart);

  const int pslot = (h->prof_ev && h->prof_next < h->prof_slots) ? h->prof_next++ : -1;
  mdb_prof_mark(h, pslot, 0, st);
  {
    if (p->normalize_disparity) {   // the mean only feeds the normalisation (src/training.jl:64-65)
      dim3 grid(L.rowblocks, Ls, N);
      disp_mean_kernel<<<grid, MEAN_THREADS, 0, st>>>(fa, mean_part);
      MDB_LAUNCHED(h);
    }
    const int threads = N * S + Ls * N;
    prep_kernel<<<(threads + 127) / 128, 128, 0, st>>>(pa, (PoseStrip*)(ws + L.o_ppd), mean_part, L.rowblocks, Ls, W, H,
                                                       p->normalize_disparity, stats, (unsigned*)(ws + L.o_counter));
    MDB_LAUNCHED(h);
  }
  mdb_prof_mark(h, pslot, 1, st);
  int32_t rc;
  if (io->C == 3 && S == 2) rc = launch_fused<3, 2>(h, fa, L, st);
  else if (io->C == 3 && S == 1) rc = launch_fused<3, 1>(h, fa, L, st);
  else if (io->C == 1 && S == 2) rc = launch_fused<1, 2>(h, fa, L, st);
  else rc = launch_fused<1, 1>(h, fa, L, st);
  if (rc != MDB_OK) return rc;
  if (h->fork_event) MDB_CUDA(h, cudaEventRecord(h->fork_event, st));   // slow_depth's graph forks here
  mdb_prof_mark(h, pslot, 2, st);

  double* item_loss = (double*)(ws + L.o_itemloss);
  if (S == 2) finalize_item_kernel<2><<<N, FIN_THREADS, 0, st>>>(fa.partials, L.records, pa, fin, stats, item_loss);
  else finalize_item_kernel<1><<<N, FIN_THREADS, 0, st>>>(fa.partials, L.records, pa, fin, stats, item_loss);
  MDB_LAUNCHED(h);
  mdb_prof_mark(h, pslot, 3, st);
  {
    bool any_low = false;
    for (int s = 0; s < Ls; ++s) any_low = any_low || !fa.native[s];
    const bool force_split = g_dbg_upsample_split != 0;   // tests: the two-kernel path (W > 4096)
    // native scales only receive the normalisation constant here: nothing to do without it (slow_depth)
    if (!(any_low || p->normalize_disparity)) {
    } else if (W <= 4096 && !force_split && !fa.hp) {   // the band kernel carries align-corners remainders down the rows
      UpBandArgs ub{};
      ub.Ls = Ls; ub.W = W; ub.H = H; ub.N = N;
      // native rows per band: what the shared-memory row buffer holds (8 at W = 1024), fewer when the launch
      // could not give every SM a few CTAs otherwise
      int band = std::min(UPB_BAND, std::max(1, UPB_SMEM_BYTES / (4 * W)));
      auto blocks_for = [&](int band_) {
        int b = 0;
        for (int s = 0; s < Ls; ++s)
          if (!fa.native[s]) b += (io->dh[s] + band_ - 1) / band_;
        return b;
      };
      while (band > 2 && (long long)blocks_for(band) * N < 4ll * h->num_sms) band /= 2;   // stays <= UPB_BAND
      for (int s = 0; s < Ls; ++s) {
        ub.gfull[s] = ua.gfull[s]; ub.gnat[s] = ua.gnat[s]; ub.dw[s] = io->dw[s]; ub.dh[s] = io->dh[s];
        ub.band[s] = fa.native[s] ? 0 : band;
        const int nb = fa.native[s] ? (p->normalize_disparity ? (H + UPB_NATIVE_ROWS - 1) / UPB_NATIVE_ROWS : 0)
                                    : (io->dh[s] + band - 1) / band;
        ub.blk_start[s + 1] = ub.blk_start[s] + nb;
      }
      dim3 gb(ub.blk_start[Ls], N);
      const size_t smem = (size_t)band * W * sizeof(float);
      // 128-bit accesses need W % 4 == 0 and 16-byte aligned caller buffers (the native maps are updated in place)
      bool vec = W % 4 == 0;
      for (int s = 0; s < Ls; ++s) vec = vec && (!fa.native[s] || ((uintptr_t)io->d_disparity[s] & 15) == 0);
      if (vec && W <= 1024) upsample_bwd_band_kernel<4, UPB_NG, 256 / UPB_NG><<<gb, 256 / UPB_NG, smem, st>>>(ub, stats);
      else if (vec) upsample_bwd_band_kernel<4, 4, 256><<<gb, 256, smem, st>>>(ub, stats);
      else if (W <= 1024) upsample_bwd_band_kernel<1, 4, 256><<<gb, 256, smem, st>>>(ub, stats);
      else upsample_bwd_band_kernel<1, 16, 256><<<gb, 256, smem, st>>>(ub, stats);
      MDB_LAUNCHED(h);
    } else {
      dim3 gv((ua.row_start[Ls] + UP_ROWS - 1) / UP_ROWS, N), gh = gv;
      upsample_bwd_v_kernel<<<gv, 256, 0, st>>>(ua, stats);
      MDB_LAUNCHED(h);
      if (any_low) {
        upsample_bwd_h_kernel<<<gh, 256, 0, st>>>(ua);
        MDB_LAUNCHED(h);
      }
    }
  }
  // the scalar loss was written by finalize_item_kernel (its last CTA): no launch of its own
  mdb_prof_mark(h, pslot, 4, st);
  return MDB_OK;
}

}  // extern "C"
