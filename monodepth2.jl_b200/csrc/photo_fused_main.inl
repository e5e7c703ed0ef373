// photo_fused_main.inl -- the per-tile fused forward+backward kernel (included by photo_fused.cu)
namespace {

template <int C, int S>
struct FusedSmem {
  static constexpr int NPART = 1 + NPOSE * S;
  static constexpr int WXN = (TW + 1) * TH, WYN = TW * (TH + 1);
  static constexpr int o_tgt = 0;
  static constexpr int o_wrp = o_tgt + C * R2N;
  static constexpr int o_dsp = o_wrp + S * C * R2N;
  static constexpr int o_coef = o_dsp + R2N;
  static constexpr int o_wx = o_coef + 3 * C * R1N;
  static constexpr int o_wy = o_wx + WXN;
  static constexpr int o_red = o_wy + WYN;
  static constexpr int o_pp = o_red + (FUSED_THREADS / 32) * NPART;
  static constexpr int o_ppd = (o_pp + S * 12 + 1) / 2 * 2;   // 8-byte aligned doubles
  static constexpr int o_invd = o_ppd + S * 24;
  static constexpr int o_sel = o_invd + 2 * R2N;  // bytes from here on
  static constexpr size_t bytes = (size_t)o_sel * 4 + ((R1N + 15) / 16) * 16;
};

template <int C, int S>
__global__ void __launch_bounds__(FUSED_THREADS, 2)
photo_fused_kernel(const __grid_constant__ FusedArgs a) {
  using SM = FusedSmem<C, S>;
  extern __shared__ __align__(128) float smem[];
  float* s_tgt = smem + SM::o_tgt;
  float* s_wrp = smem + SM::o_wrp;
  float* s_dsp = smem + SM::o_dsp;
  float* s_coef = smem + SM::o_coef;
  float* s_wx = smem + SM::o_wx;
  float* s_wy = smem + SM::o_wy;
  float* s_red = smem + SM::o_red;
  PosePrep* s_pp = reinterpret_cast<PosePrep*>(smem + SM::o_pp);
  PosePrepD* s_ppd = reinterpret_cast<PosePrepD*>(smem + SM::o_ppd);
  double* s_invd = reinterpret_cast<double*>(smem + SM::o_invd);
  unsigned char* s_sel = reinterpret_cast<unsigned char*>(smem + SM::o_sel);

  const int tid = threadIdx.x;
  const int n = blockIdx.z;
  const int X0 = blockIdx.x * TW, Y0 = blockIdx.y * TH;
  const int W = a.W, H = a.H;
  const size_t plane = (size_t)W * H;
  const float* tgt = a.x + ((size_t)n * a.L + a.tgt) * C * plane;
  const float* srcp[S];
#pragma unroll
  for (int s = 0; s < S; ++s) srcp[s] = a.x + ((size_t)n * a.L + a.src[s]) * C * plane;

  if (tid < S * 12) {
    reinterpret_cast<float*>(s_pp)[tid] = reinterpret_cast<const float*>(a.pp + (size_t)n * S)[tid];
    reinterpret_cast<double*>(s_ppd)[tid] = reinterpret_cast<const double*>(a.ppd + (size_t)n * S)[tid];
  }

  // ---- phase 0: target tile + 2-pixel halo, smoothness edge weights ------------------
  for (int i = tid; i < R2N; i += FUSED_THREADS) {
    const int gx = X0 - 2 + i % R2W, gy = Y0 - 2 + i / R2W;
    const bool in = gx >= 0 && gx < W && gy >= 0 && gy < H;
#pragma unroll
    for (int c = 0; c < C; ++c) s_tgt[c * R2N + i] = in ? ldg(tgt + c * plane + (size_t)gy * W + gx) : 0.0f;
  }
  __syncthreads();
  for (int i = tid; i < SM::WXN; i += FUSED_THREADS) {  // weight between (gx,gy) and (gx+1,gy)
    const int lx = i % (TW + 1), ly = i / (TW + 1);
    const int gx = X0 - 1 + lx, gy = Y0 + ly;
    float w = 0.0f;
    if (gx >= 0 && gx + 1 < W && gy < H) {
      const int r = (ly + 2) * R2W + lx + 1;
      float e = 0.0f;
#pragma unroll
      for (int c = 0; c < C; ++c) e += fabsf(s_tgt[c * R2N + r] - s_tgt[c * R2N + r + 1]);
      w = expf(-e * (1.0f / C));
    }
    s_wx[i] = w;
  }
  for (int i = tid; i < SM::WYN; i += FUSED_THREADS) {  // weight between (gx,gy) and (gx,gy+1)
    const int lx = i % TW, ly = i / TW;
    const int gx = X0 + lx, gy = Y0 - 1 + ly;
    float w = 0.0f;
    if (gy >= 0 && gy + 1 < H && gx < W) {
      const int r = (ly + 1) * R2W + lx + 2;
      float e = 0.0f;
#pragma unroll
      for (int c = 0; c < C; ++c) e += fabsf(s_tgt[c * R2N + r] - s_tgt[c * R2N + r + R2W]);
      w = expf(-e * (1.0f / C));
    }
    s_wy[i] = w;
  }

  float loss_acc = 0.0f;
  float pacc[S][NPOSE];
#pragma unroll
  for (int s = 0; s < S; ++s)
#pragma unroll
    for (int k = 0; k < NPOSE; ++k) pacc[s][k] = 0.0f;

  const float ks = a.alpha / (9.0f * C);   // SSIM pullback factor  alpha / C / 9
  const float kl = (1.0f - a.alpha) / C;   // L1 factor

  for (int sc = 0; sc < a.Ls; ++sc) {
    const float* dsp = a.disp[sc] + (size_t)n * a.dw[sc] * a.dh[sc];
    const bool native = a.native[sc] != 0;
    __syncthreads();  // previous scale's phase 3 is done with s_wrp / s_dsp / s_coef

    // ---- phase 1: disparity -> depth -> project -> bilinear sample, on R2 -------------
    for (int i = tid; i < R2N; i += FUSED_THREADS) {
      const int gx = X0 - 2 + i % R2W, gy = Y0 - 2 + i / R2W;
      if (gx < 0 || gx >= W || gy < 0 || gy >= H) continue;
      const double dd_ = native ? (double)ldg(dsp + (size_t)gy * W + gx)
                                : upsample_at_d(dsp, a.dw[sc], a.dh[sc], a.rxd[sc], a.ryd[sc], gx, gy);
      s_dsp[i] = (float)dd_;
      s_invd[i] = fma(dd_, a.dd_d, a.mind_d);   // 1 / depth  (src/utils.jl:179-183)
      const double invd = s_invd[i];
#pragma unroll
      for (int s = 0; s < S; ++s) {
        const Samp sp = sample_coords(s_ppd[s], invd, gx, gy, W, H);
        const Taps& t = sp.t;
#pragma unroll
        for (int c = 0; c < C; ++c) {
          const float* p = srcp[s] + c * plane;
          const float v00 = ldg(p + t.o00), v01 = ldg(p + t.o01), v10 = ldg(p + t.o10), v11 = ldg(p + t.o11);
          const float top = fmaf(t.fx, v01 - v00, v00), bot = fmaf(t.fx, v11 - v10, v10);
          s_wrp[(s * C + c) * R2N + i] = fmaf(t.fy, bot - top, top);
        }
      }
    }
    __syncthreads();

    // ---- phase 2: window statistics, per-source loss, arg-min, pullback coefficients, on R1
    for (int i = tid; i < R1N; i += FUSED_THREADS) {
      const int l1x = i % R1W, l1y = i / R1W;
      const int gx = X0 - 1 + l1x, gy = Y0 - 1 + l1y;
      if (gx < 0 || gx >= W || gy < 0 || gy >= H) { s_sel[i] = 255; continue; }
      int cx[3], ry[3];
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        cx[k] = reflect1(gx + k, W) - 1 - (X0 - 2);          // reflect1 works on 1-based coords
        ry[k] = (reflect1(gy + k, H) - 1 - (Y0 - 2)) * R2W;
      }
      float lossv[S];
      float cf[S][3 * C];
#pragma unroll
      for (int s = 0; s < S; ++s) lossv[s] = 0.0f;
#pragma unroll
      for (int c = 0; c < C; ++c) {
        float yv[9], Sy = 0.0f, Syy = 0.0f;
        const float yc = s_tgt[c * R2N + ry[1] + cx[1]];
#pragma unroll
        for (int k = 0; k < 9; ++k) {
          yv[k] = s_tgt[c * R2N + ry[k / 3] + cx[k % 3]] - yc;   // shifted by the window centre
          Sy += yv[k];
          Syy = fmaf(yv[k], yv[k], Syy);
        }
#pragma unroll
        for (int s = 0; s < S; ++s) {
          const float* w = s_wrp + (s * C + c) * R2N;
          const float xc = w[ry[1] + cx[1]];
          float Sx = 0.0f, Sxx = 0.0f, Sxy = 0.0f;
#pragma unroll
          for (int k = 0; k < 9; ++k) {
            const float xv = w[ry[k / 3] + cx[k % 3]] - xc;
            Sx += xv;
            Sxx = fmaf(xv, xv, Sxx);
            Sxy = fmaf(xv, yv[k], Sxy);
          }
          float A, B, Cc;
          const float sv = ssim_window(Sx, Sy, Sxx, Syy, Sxy, xc, yc, a.c1, a.c2, A, B, Cc);
          cf[s][3 * c + 0] = A; cf[s][3 * c + 1] = B; cf[s][3 * c + 2] = Cc;
          lossv[s] += a.alpha * (1.0f / C) * sv + kl * fabsf(yc - xc);
        }
      }
      int best = 0;
      float val = lossv[0];
#pragma unroll
      for (int s = 1; s < S; ++s)
        if (lossv[s] < val) { val = lossv[s]; best = s; }   // first arg-min wins ties
      if (a.automask) {
        const float am = ldg(a.auto_loss + (size_t)n * plane + (size_t)gy * W + gx);
        if (am <= val) { val = am; best = 255; }             // mask precedes warp_loss
      }
      s_sel[i] = (unsigned char)best;
#pragma unroll
      for (int k = 0; k < 3 * C; ++k) {
        float v = cf[0][k];
#pragma unroll
        for (int s = 1; s < S; ++s) v = (best == s) ? cf[s][k] : v;
        s_coef[k * R1N + i] = v * ks;
      }
      if (l1x >= 1 && l1x <= TW && l1y >= 1 && l1y <= TH) {
        loss_acc += val;
        if (a.vis && sc == a.Ls - 1) a.vis[(size_t)n * plane + (size_t)gy * W + gx] = val;
      }
    }
    __syncthreads();

    // ---- phase 3: pullback to the warped samples, the sample position, depth and pose --
    const ScaleStat st = a.stats[(size_t)sc * a.N + n];
    float* gout = a.gout[sc] + (size_t)n * plane;
    for (int i = tid; i < TW * TH; i += FUSED_THREADS) {
      const int ix = i % TW, iy = i / TW;
      const int gx = X0 + ix, gy = Y0 + iy;
      if (gx >= W || gy >= H) continue;
      const int r2 = (iy + 2) * R2W + ix + 2;
      float xp[S][C], yp[C], G[S][C];
#pragma unroll
      for (int c = 0; c < C; ++c) {
        yp[c] = s_tgt[c * R2N + r2];
#pragma unroll
        for (int s = 0; s < S; ++s) { xp[s][c] = s_wrp[(s * C + c) * R2N + r2]; G[s][c] = 0.0f; }
      }
#pragma unroll
      for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
        for (int dx = -1; dx <= 1; ++dx) {
          const int r1 = (iy + 1 + dy) * R1W + ix + 1 + dx;
          const int sel = s_sel[r1];
          if (sel == 255) continue;
          float m = 1.0f;
          if ((dx == -1 && gx == 1) || (dx == 1 && gx == W - 2)) m = 2.0f;
          if ((dy == -1 && gy == 1) || (dy == 1 && gy == H - 2)) m *= 2.0f;
#pragma unroll
          for (int c = 0; c < C; ++c) {
            float xs = xp[0][c];
#pragma unroll
            for (int s = 1; s < S; ++s) xs = (sel == s) ? xp[s][c] : xs;
            const float A = s_coef[(3 * c + 0) * R1N + r1], B = s_coef[(3 * c + 1) * R1N + r1],
                        Cc = s_coef[(3 * c + 2) * R1N + r1];
            const int q2 = r2 + dy * R2W + dx;                       // the window's centre pixel
            const float xq = s_wrp[(sel * C + c) * R2N + q2], yq = s_tgt[c * R2N + q2];
            float v = m * fmaf(2.0f * (xs - xq), B, fmaf(yp[c] - yq, Cc, A));
            if (dx == 0 && dy == 0) v += kl * sgnf(xs - yp[c]);
#pragma unroll
            for (int s = 0; s < S; ++s) G[s][c] += (sel == s) ? v : 0.0f;
          }
        }
      const float d = s_dsp[r2];
      const double invd_d = s_invd[r2];
      const float invd = (float)invd_d;
      const float px = (float)(gx + 1), py = (float)(gy + 1);
      float dinvd = 0.0f;
#pragma unroll
      for (int s = 0; s < S; ++s) {
        bool any = false;
#pragma unroll
        for (int c = 0; c < C; ++c) any |= (G[s][c] != 0.0f);
        if (!any) continue;
        const PosePrep& pp = s_pp[s];
        const Samp sp = sample_coords(s_ppd[s], invd_d, gx, gy, W, H);
        const Taps& t = sp.t;
        float gu = 0.0f, gv = 0.0f;
#pragma unroll
        for (int c = 0; c < C; ++c) {
          const float* p = srcp[s] + c * plane;
          const float v00 = ldg(p + t.o00), v01 = ldg(p + t.o01), v10 = ldg(p + t.o10), v11 = ldg(p + t.o11);
          const float dxv = fmaf(t.fy, (v11 - v10) - (v01 - v00), v01 - v00);
          const float dyv = fmaf(t.fx, (v11 - v01) - (v10 - v00), v10 - v00);
          gu = fmaf(G[s][c], dxv, gu);
          gv = fmaf(G[s][c], dyv, gv);
        }
        gu *= t.mx * a.up;
        gv *= t.my * a.up;
        // u = P'1 / z'', v = P'2 / z''  with  P' = A q + invd b,  z'' = P'3 + invd 1e-7
        const float u = px + sp.fu, v = py + sp.fv;
        const float dP0 = gu * sp.zinv, dP1 = gv * sp.zinv, dP2 = -(gu * u + gv * v) * sp.zinv;
        dinvd += dP0 * pp.b[0] + dP1 * pp.b[1] + dP2 * (pp.b[2] + 1e-7f);
        pacc[s][0] = fmaf(dP0, px, pacc[s][0]); pacc[s][1] = fmaf(dP0, py, pacc[s][1]); pacc[s][2] += dP0;
        pacc[s][3] = fmaf(dP1, px, pacc[s][3]); pacc[s][4] = fmaf(dP1, py, pacc[s][4]); pacc[s][5] += dP1;
        pacc[s][6] = fmaf(dP2, px, pacc[s][6]); pacc[s][7] = fmaf(dP2, py, pacc[s][7]); pacc[s][8] += dP2;
        pacc[s][9] = fmaf(dP0, invd, pacc[s][9]); pacc[s][10] = fmaf(dP1, invd, pacc[s][10]);
        pacc[s][11] = fmaf(dP2, invd, pacc[s][11]);
      }
      // d invd / d disparity = 1/min_depth - 1/max_depth
      float g = dinvd * a.dd;
      // smoothness pullback (src/utils.jl:163-177), mean-normalisation constant from disp_stats
      float gsx = 0.0f, gsy = 0.0f;
      if (gx + 1 < W) gsx += sgnf(d - s_dsp[r2 + 1]) * s_wx[iy * (TW + 1) + ix + 1];
      if (gx > 0) gsx -= sgnf(s_dsp[r2 - 1] - d) * s_wx[iy * (TW + 1) + ix];
      if (gy + 1 < H) gsy += sgnf(d - s_dsp[r2 + R2W]) * s_wy[(iy + 1) * TW + ix];
      if (gy > 0) gsy -= sgnf(s_dsp[r2 - R2W] - d) * s_wy[iy * TW + ix];
      g += st.inv_mp * (a.kx[sc] * gsx + a.ky[sc] * gsy) + st.cconst;
      gout[(size_t)gy * W + gx] = g;
    }
  }

  // ---- phase 4: block reduction of the loss and pose partials (no atomics) -------------
  const int lane = tid & 31, warp = tid >> 5;
  {
    const float v = warp_sum(loss_acc);
    if (lane == 0) s_red[warp * SM::NPART] = v;
  }
#pragma unroll
  for (int s = 0; s < S; ++s)
#pragma unroll
    for (int k = 0; k < NPOSE; ++k) {
      const float v = warp_sum(pacc[s][k]);
      if (lane == 0) s_red[warp * SM::NPART + 1 + s * NPOSE + k] = v;
    }
  __syncthreads();
  if (tid < SM::NPART) {
    float v = 0.0f;
#pragma unroll
    for (int w = 0; w < FUSED_THREADS / 32; ++w) v += s_red[w * SM::NPART + tid];
    const size_t tile = (size_t)blockIdx.y * a.tiles_x + blockIdx.x;
    a.partials[((size_t)n * a.tiles_x * a.tiles_y + tile) * SM::NPART + tid] = v;
  }
}

// ---------------------------------------------------------------------------------------
// 5. transpose of upsample_bilinear: native-resolution gradient by gather (deterministic)
// ---------------------------------------------------------------------------------------
__global__ void upsample_bwd_kernel(const float* __restrict__ gfull, int W, int H, int N, int dw, int dh,
                                    float rx, float ry, float* __restrict__ gnat) {
  const int jx = blockIdx.x * blockDim.x + threadIdx.x;
  const int jy = blockIdx.y * blockDim.y + threadIdx.y;
  const int n = blockIdx.z;
  if (jx >= dw || jy >= dh) return;
  // full-res pixels whose source coordinate g*r lies in (j-1, j+1)
  const int px_lo = max(0, (int)floorf((float)(jx - 1) / rx) - 1), px_hi = min(W - 1, (int)ceilf((float)(jx + 1) / rx) + 1);
  const int py_lo = max(0, (int)floorf((float)(jy - 1) / ry) - 1), py_hi = min(H - 1, (int)ceilf((float)(jy + 1) / ry) + 1);
  const float* g = gfull + (size_t)n * W * H;
  float acc = 0.0f;
  for (int py = py_lo; py <= py_hi; ++py) {
    int y0, y1; float ly;
    upsample_coords(py, ry, dh, y0, y1, ly);
    const float wy = (y0 == jy ? 1.0f - ly : 0.0f) + (y1 == jy ? ly : 0.0f);
    if (wy == 0.0f) continue;
    float row = 0.0f;
    for (int px = px_lo; px <= px_hi; ++px) {
      int x0, x1; float lx;
      upsample_coords(px, rx, dw, x0, x1, lx);
      const float wx = (x0 == jx ? 1.0f - lx : 0.0f) + (x1 == jx ? lx : 0.0f);
      if (wx != 0.0f) row = fmaf(wx, ldg(g + (size_t)py * W + px), row);
    }
    acc = fmaf(wy, row, acc);
  }
  gnat[((size_t)n * dh + jy) * dw + jx] = acc;
}

}  // namespace
