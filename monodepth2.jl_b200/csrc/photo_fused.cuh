// photo_fused.cuh -- argument blocks and shared device math of the fused photometric path.
#pragma once
#include "mdb_common.cuh"

// ---- tile geometry of the fused kernel ------------------------------------------------
// One CTA owns a TW x TH block of target pixels.  SSIM windows need warped samples one
// pixel around every window centre, and the backward pass needs windows one pixel around
// every interior pixel, so warped samples are produced on the tile + 2 pixel halo (R2) and
// window statistics on the tile + 1 pixel halo (R1).
constexpr int TW = 32, TH = 16;
constexpr int R2W = TW + 4, R2H = TH + 4, R2N = R2W * R2H;  // 36 x 20
constexpr int R1W = TW + 2, R1H = TH + 2, R1N = R1W * R1H;  // 34 x 18
constexpr int FUSED_THREADS = 256;
constexpr int NPOSE = 12;  // dL/dA (9) + dL/db (3) per (item, source)

// Per (item, source) projection constants produced by pose_prep_kernel:
//   P = depth * (A q) + b,  q = [w, h, 1],  A = K R' invK,  b = K t'   (src/utils.jl:67-69,99-103
//   with K (R X + t) pre-composed).  dA = A - I is stored so the pixel *displacement* can be
//   computed without cancellation (see project_flow).
struct PosePrep {
  float dA[9];  // row-major
  float b[3];
};

struct PosePrepD {  // the same constants in double: the geometric chain runs in double
  double dA[9];
  double b[3];
};

// What the strip kernel reads per (item, source): PosePrepD plus the derived constants its row loop uses, so that --
// served from the constant bank (STRIP_CONSTPOSE) -- they are operands, not registers.
struct PoseStrip {
  double dA[9];   // row-major
  double b[3];
  double i2;      // 1 + dA[4]: the row step of P'2
  double b3e;     // b[2] + 1e-7 (src/utils.jl:101, in units of depth)
  float b0f, b1f, b3f, pad;
};

struct ScaleStat {
  float inv_mp;  // 1 / (mean(disparity) + 1e-7), or 1 when the disparity is not normalised
  float cconst;  // constant added to every full-res disparity gradient (mean-normalisation term)
  float loss;    // this (scale, item)'s smoothness loss, all weights applied
  float pad;
};

struct FusedArgs {
  const float* x;
  int W, H, L, N;
  int tgt;                      // 0-based frame index of the target
  int src[MDB_MAX_SOURCES];     // 0-based frame indices of the sources
  int Ls;
  const float* disp[MDB_MAX_SCALES];
  int dw[MDB_MAX_SCALES], dh[MDB_MAX_SCALES];
  float rx[MDB_MAX_SCALES], ry[MDB_MAX_SCALES];  // (w_k-1)/(W-1), (h_k-1)/(H-1)
  int native[MDB_MAX_SCALES];
  float* gout[MDB_MAX_SCALES];  // full-res disparity gradient target (output or workspace)
  float kx[MDB_MAX_SCALES], ky[MDB_MAX_SCALES];  // smoothness weights incl. 1/Ls and the means
  const float* auto_loss;
  float mind, dd;               // 1/max_depth, 1/min_depth - 1/max_depth
  double mind_d, dd_d;          // the same in double
  double rxd[MDB_MAX_SCALES], ryd[MDB_MAX_SCALES];
  double oxd[MDB_MAX_SCALES], oyd[MDB_MAX_SCALES];   // half-pixel up-sampling (hp): source coordinate = max(0, p r + o)
  int hp;                       // upsample_align_corners == 0; served by the CP = false instantiations only
  int x_u8;                     // x holds N0f8 bytes instead of Float32 (mdb_loss_io.x_format == 1); constant-bank instantiation only
  const PoseStrip* ppd;         // [N][S]
  float alpha, c1, c2;
  float ksu, klu, kl, ka;       // loss / pullback factors (photo_strip.inl)
  float abs_tie;                // d|x|/dx at +0.0: 0 (sign(0) = 0) or 1 (ForwardDiff's Dual rule)
  int automask;
  float up;                     // 1 / (W H N_global Ls)
  const ScaleStat* stats;       // [Ls][N]
  float* partials;              // [N][tiles][1 + NPOSE*S]
  float* vis;
  signed char* dbg;             // decision dump (photo_strip.inl; DBG instantiation only)
  int tiles_x, tiles_y;
};

#ifdef __CUDACC__
// bilinear upsample (align-corners) of a native disparity map at the 0-based full-res pixel
// (gx, gy); mirrors NNlib upsample_bilinear / torch interpolate(align_corners=True).
MDB_DEV float upsample_at(const float* __restrict__ d, int dw, int dh, float rx, float ry, int gx,
                          int gy) {
  const float fx = (float)gx * rx, fy = (float)gy * ry;
  int x0 = (int)fx, y0 = (int)fy;
  x0 = min(x0, dw - 1);
  y0 = min(y0, dh - 1);
  const float lx = fx - (float)x0, ly = fy - (float)y0;
  const int x1 = min(x0 + 1, dw - 1), y1 = min(y0 + 1, dh - 1);
  const float v00 = ldg(d + (size_t)y0 * dw + x0), v01 = ldg(d + (size_t)y0 * dw + x1);
  const float v10 = ldg(d + (size_t)y1 * dw + x0), v11 = ldg(d + (size_t)y1 * dw + x1);
  const float top = v00 + lx * (v01 - v00), bot = v10 + lx * (v11 - v10);
  return top + ly * (bot - top);
}

// Up-sampling coordinate of full-res index p on a native axis of dn cells (Dn full-res): num(p) / D with the exact integer
// numerator num(p) = clamp(A p + B, 0, (dn - 1) D).   align-corners (NNlib's default): A = dn - 1, B = 0, D = Dn - 1;
// half-pixel centres (align_corners = false: max(0, (p + 0.5) dn / Dn - 0.5)): A = 2 dn, B = dn - Dn, D = 2 Dn.
// The weight of p for native cell j is the hat function max(0, 1 - |num(p) - j D| / D) in both modes.
struct UpAxis { long long A, B, D, nmax; int Dn; };
MDB_DEV UpAxis up_axis(int dn, int Dn, int hp) {
  UpAxis u;
  if (hp) { u.A = 2ll * dn; u.B = (long long)dn - Dn; u.D = 2ll * Dn; }
  else { u.A = dn - 1; u.B = 0; u.D = Dn - 1; }
  u.nmax = (long long)(dn - 1) * u.D;
  u.Dn = Dn;
  return u;
}
MDB_DEV long long up_num(const UpAxis& u, int p) { return min(max(u.A * p + u.B, 0ll), u.nmax); }
MDB_DEV long long up_fdiv(long long a, long long b) { return a >= 0 ? a / b : -((-a + b - 1) / b); }   // floor, b > 0
// full-res indices that can carry weight for native cell j (a superset by at most one on either side)
MDB_DEV void up_range(const UpAxis& u, int j, int& lo, int& hi) {
  lo = (int)max(0ll, up_fdiv((long long)(j - 1) * u.D - u.B, u.A));
  hi = (int)min((long long)u.Dn - 1, up_fdiv((long long)(j + 1) * u.D - u.B, u.A) + 1);
  if (j == 0) lo = 0;                                    // clamped head: every p with A p + B <= 0 sits on cell 0
  if ((long long)j * u.D >= u.nmax) hi = u.Dn - 1;       // clamped tail
}

// the same interpolation weights, for the transposed (backward) gather
MDB_DEV void upsample_coords(int g, float r, int dn, int& i0, int& i1, float& l) {
  const float f = (float)g * r;
  i0 = min((int)f, dn - 1);
  l = f - (float)i0;
  i1 = min(i0 + 1, dn - 1);
}

// MUFU.RCP: 1 ulp, no slow path (callers guarantee a normal, positive argument)
MDB_DEV float rcp_approx(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

struct Taps {
  int o00, o01, o10, o11;  // offsets inside one channel plane
  float fx, fy;            // fractional weights
  float mx, my;            // 1 if the coordinate was not clamped (gradient flows), else 0
};

struct Samp {
  Taps t;
  float fu, fv;  // displacement (u - w, v - h) of the projected point, unclamped
  float zinv;    // 1 / z'',  z'' = A3.q + invd (b3 + 1e-7)
};

// Projection of the 0-based pixel (gx, gy) with inverse depth `invd` into a source frame, and
// the bilinear taps there.  Reference chain: disparity_to_depth (src/utils.jl:179-183) ->
// Backproject (:67-69) -> Project (:99-103) -> normalize -> grid_sample's unnormalize, which
// cancel to the pixel coordinate u = P1 / (P3 + 1e-7), P = depth (A q) + b.  Dividing through by
// depth (invd = 1/depth = d (1/min - 1/max) + 1/max needs no division) and writing A = I + dA:
//   u - w = [ (dA1.q - w dA3.q) + invd (b1 - w (b3 + 1e-7)) ] / [ 1 + dA3.q + invd (b3 + 1e-7) ]
// This short chain runs in DOUBLE.  Displacements reach hundreds of pixels (f t / depth), so a
// float32 chain leaves ~1e-5 px of error, part of it coherent over the image (rounded
// constants); the SSIM pullback amplifies that to ~1e-4 relative in the pose gradients, and it
// flips which bilinear cell a sample falls in (the only discontinuity of the pullback) a few
// times per image.  In double both effects vanish, for ~30 FP64 ops per sample on a part whose
// FP64 rate is half its FP32 rate.  Then NNlib grid_sample(padding_mode=:border),
// align-corners=true: sample position = clamp(u, 1, W); gradient is zero where clamped.
MDB_DEV Samp sample_coords(const PosePrepD& pp, double invd, int gx, int gy, int W, int H) {
  Samp r;
  const double px = (double)(gx + 1), py = (double)(gy + 1);
  const double d0 = fma(pp.dA[0], px, fma(pp.dA[1], py, pp.dA[2]));
  const double d1 = fma(pp.dA[3], px, fma(pp.dA[4], py, pp.dA[5]));
  const double d2 = fma(pp.dA[6], px, fma(pp.dA[7], py, pp.dA[8]));
  const double b3e = pp.b[2] + 1e-7;
  const double z = fma(invd, b3e, 1.0 + d2);
  double zi = (double)__frcp_rn((float)z);     // float seed + one Newton step = ~1e-14 relative
  zi = fma(zi, fma(-z, zi, 1.0), zi);
  const double fu = fma(invd, fma(-px, b3e, pp.b[0]), fma(-px, d2, d0)) * zi;
  const double fv = fma(invd, fma(-py, b3e, pp.b[1]), fma(-py, d2, d1)) * zi;
  const double lox = 1.0 - px, hix = (double)W - px, loy = 1.0 - py, hiy = (double)H - py;
  Taps& t = r.t;
  t.mx = (fu > lox && fu < hix) ? 1.0f : 0.0f;
  t.my = (fv > loy && fv < hiy) ? 1.0f : 0.0f;
  const double cu = fmin(fmax(fu, lox), hix), cv = fmin(fmax(fv, loy), hiy);
  const double flx = floor(cu), fly = floor(cv);
  t.fx = (float)(cu - flx);
  t.fy = (float)(cv - fly);
  const int x0 = gx + (int)flx, y0 = gy + (int)fly;
  const int x1 = min(x0 + 1, W - 1), y1 = min(y0 + 1, H - 1);
  t.o00 = y0 * W + x0; t.o01 = y0 * W + x1;
  t.o10 = y1 * W + x0; t.o11 = y1 * W + x1;
  r.fu = (float)fu;
  r.fv = (float)fv;
  r.zinv = (float)zi;
  return r;
}

// double-precision bilinear upsample (align-corners) of a native disparity map
MDB_DEV double upsample_at_d(const float* __restrict__ d, int dw, int dh, double rx, double ry, int gx, int gy) {
  const double fx = gx * rx, fy = gy * ry;
  const int x0 = min((int)fx, dw - 1), y0 = min((int)fy, dh - 1);
  const double lx = fx - x0, ly = fy - y0;
  const int x1 = min(x0 + 1, dw - 1), y1 = min(y0 + 1, dh - 1);
  const double v00 = ldg(d + (size_t)y0 * dw + x0), v01 = ldg(d + (size_t)y0 * dw + x1);
  const double v10 = ldg(d + (size_t)y1 * dw + x0), v11 = ldg(d + (size_t)y1 * dw + x1);
  const double top = fma(lx, v01 - v00, v00), bot = fma(lx, v11 - v10, v10);
  return fma(ly, bot - top, top);
}

// SSIM of one 3x3 window and the coefficients of its pullback.  Reference: src/utils.jl:29-43.
// The five sums are taken over values SHIFTED by the window's centre pixel (x' = x - xc,
// y' = y - yc): the variances sigma = E[x'^2] - E[x']^2 are then free of the catastrophic
// cancellation E[x^2] - mu^2 has in float32 on flat image regions, while
// being the same numbers mathematically.  The pullback is returned in the shifted basis too:
//   d s / d x_p = (m_pq / 9) (A + 2 (x_p - xc) B + (y_p - yc) C)
// with A = ds/dmu_x - 2 mu_x' B - mu_y' C, B = ds/dsigma_x, C = ds/dsigma_xy.
MDB_DEV float ssim_window(float Sx, float Sy, float Sxx, float Syy, float Sxy, float xc, float yc,
                          float c1, float c2, float& A, float& B, float& C) {
  const float k = 1.0f / 9.0f;
  const float mxs = Sx * k, mys = Sy * k;            // shifted means
  const float sx = fmaf(Sxx, k, -mxs * mxs), sy = fmaf(Syy, k, -mys * mys), sxy = fmaf(Sxy, k, -mxs * mys);
  const float mx = xc + mxs, my = yc + mys;
  const float n1 = fmaf(2.0f * mx, my, c1), n2 = fmaf(2.0f, sxy, c2);
  const float d1 = fmaf(mx, mx, fmaf(my, my, c1)), d2 = sx + sy + c2;
  // d1 >= c1 > 0 and d2 >= c2 - rounding > 0: the 1-ulp hardware reciprocal is enough here
  const float i1 = rcp_approx(d1), i2 = rcp_approx(d2);
  const float r1 = n1 * i1, r2 = n2 * i2;            // n/d = r1 r2
  const float raw = (1.0f - r1 * r2) * 0.5f;
  const bool inside = (raw >= 0.0f) && (raw <= 1.0f);
  // ds/dmu_x = n2 (mu_x - mu_y) (mu_y (mu_x + mu_y) + c1) / (d1^2 d2), with mu_x - mu_y formed
  // from the raw centre difference so no large terms cancel
  const float dmu = (xc - yc) + (mxs - mys);
  const float Abar = r2 * i1 * i1 * dmu * fmaf(my, mx + my, c1);
  B = inside ? 0.5f * r1 * r2 * i2 : 0.0f;           // ds/dsigma_x  = n1 n2 / (2 d1 d2^2)
  C = inside ? -r1 * i2 : 0.0f;                       // ds/dsigma_xy = -n1 / (d1 d2)
  A = inside ? Abar - 2.0f * mxs * B - mys * C : 0.0f;
  return fminf(fmaxf(raw, 0.0f), 1.0f);
}

MDB_DEV float sgnf(float v) { return (float)((v > 0.0f) - (v < 0.0f)); }
// pullback of abs with a configurable value at 0 (mdb_loss_params.abs_subgradient)
MDB_DEV float sgnt(float tie, float v) { return v == 0.0f ? tie : copysignf(1.0f, v); }
#endif
