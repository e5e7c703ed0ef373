// render.cu -- plane-disparity (MPI) volume rendering, R1-R7 of SURVEY 8a.
// Reference: src/render.jl:3-17 (inv), :21-23 (create_meshgrid), :25-30, :51-62 (plane xyz),
// :64-90 (sample), :32-49 (plane_volume_rendering), :92-114 (render_tgt_rgb_depth).
// The reference's quirks are reproduced as written: 0-based target meshgrid, grid
// normalisation (x+0.5)/(W/2) without "-1" fed to an align-corners=true sampler, valid mask
// x > 0 & y > 0, "+ 1e-6" inside the cumulative product, "depth" output = transparency_acc.
//
// Layout (Julia column-major == these C orders): rgb [B][Np][3][H][W], sigma [B][Np][1][H][W],
// xyz [B][Np][H][W][3], disparity_src [B][Np], rvec [B][3], tvec [B][3];
// rgb_out [B][3][H][W], transparency_acc [B][Np][1][H][W], mask [B][1][1][H][W] (int32).
#include "mdb_common.cuh"
#include "pose_math.cuh"

int mdb_dbg_render_precise = 0;

namespace {

// per (plane, batch item) constants (16-byte aligned: Ax/tx and hf are read with 128-bit loads)
struct alignas(16) PlanePrep {
  double Hst[9];   // H_src_tgt = inv(K (R + t n^T / depth) K_inv), row-major
  float pad0_[2];  // G1 | disp | pad: 48 bytes at offset 80, read by the pullback with three 128-bit loads
  float G1[9];     // K^T H_src_tgt^T : pulls d/d(homogeneous source coords) back to d/dM
  float disp;      // 1 / depth_src
  float pad1_[2];
  // xyz_tgt of this plane as a function of the 1-based source pixel m = (w, h, 1):  Ax m + tx  with
  // Ax = R K_inv / disparity  (src/render.jl:25-30, 51-62) -- what the producers folded into the kernels use
  float Ax[9];
  float tx[3];
  // H_src_tgt / H22 - I for the forward kernel:  x - u = (a0 u + h01 v + h02 - u e) / (1 + e),
  // y - v = (h10 u + a1 v + h12 - v e) / (1 + e),  e = h20 u + h21 v.   hd = {a0, h01, h02, h10, a1, h12, h20, h21}.
  // The two numerators and e are formed in double (8 operations, no cancellation error), everything after them in float.
  double hd[8];   // byte offset 176
  int fast;       // |e| < 1/4 over the whole image: the short tap path applies (render_taps_fast); otherwise -- strong perspective,
                  // a plane close to the camera's horizon -- the plane goes through the general double-precision homography
  int pad_[3];
};
static_assert(offsetof(PlanePrep, G1) % 16 == 0 && offsetof(PlanePrep, Ax) % 16 == 0 && offsetof(PlanePrep, hd) % 16 == 0 &&
              sizeof(PlanePrep) % 16 == 0, "PlanePrep: the blocks read with 128-bit loads must be 16-byte aligned");

// variants of the sampler / outputs (mdb_render_opts in include/mdb200.h)
constexpr int RV_MINE_GRID = 1, RV_MINE_MASK = 2;

struct RenderConst {
  double K[9], Kinv[9];  // math row-major
  float Kinv_f[9];
  int W, H;              // image size (render_prep_kernel: range of the float sampling path); 0 = unknown
};

MDB_HD bool inv3_d(const double* A, double* B) {
  const double c00 = A[4] * A[8] - A[5] * A[7], c01 = A[5] * A[6] - A[3] * A[8], c02 = A[3] * A[7] - A[4] * A[6];
  const double det = A[0] * c00 + A[1] * c01 + A[2] * c02;
  const double id = 1.0 / det;
  B[0] = c00 * id; B[1] = (A[2] * A[7] - A[1] * A[8]) * id; B[2] = (A[1] * A[5] - A[2] * A[4]) * id;
  B[3] = c01 * id; B[4] = (A[0] * A[8] - A[2] * A[6]) * id; B[5] = (A[2] * A[3] - A[0] * A[5]) * id;
  B[6] = c02 * id; B[7] = (A[1] * A[6] - A[0] * A[7]) * id; B[8] = (A[0] * A[4] - A[1] * A[3]) * id;
  return det != 0.0;
}

// R1: batched 3x3 inverse, column-major in and out (src/render.jl:3-17 uses getrf/getri)
__global__ void inv3_kernel(const float* __restrict__ A, int M, float* __restrict__ out) {
  const int m = blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= M) return;
  double a[9], b[9];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) a[3 * i + j] = A[9 * m + 3 * j + i];
  inv3_d(a, b);
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) out[9 * m + 3 * j + i] = (float)b[3 * i + j];
}

// create_meshgrid(H, W): (3,W,H) = [w,h,1], 1-based (src/render.jl:21-23)
__global__ void meshgrid_kernel(int W, int H, float* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= W * H) return;
  out[3 * i] = (float)(i % W + 1);
  out[3 * i + 1] = (float)(i / W + 1);
  out[3 * i + 2] = 1.0f;
}

// R3: xyz_src = (K_inv meshgrid) / disparity   (src/render.jl:25-30)
__global__ void plane_xyz_src_kernel(const float* __restrict__ mesh, const float* __restrict__ disp, RenderConst rc,
                                     int WH, int Np, int B, float* __restrict__ out) {
  const int64_t total = (int64_t)WH * Np * B;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int pix = (int)(i % WH);
    const int kb = (int)(i / WH);
    const float depth = 1.0f / disp[kb];
    const float m0 = mesh[3 * pix], m1 = mesh[3 * pix + 1], m2 = mesh[3 * pix + 2];
#pragma unroll
    for (int r = 0; r < 3; ++r)
      out[3 * i + r] = (rc.Kinv_f[3 * r] * m0 + rc.Kinv_f[3 * r + 1] * m1 + rc.Kinv_f[3 * r + 2] * m2) * depth;
  }
}

// R4: xyz_tgt = R(rvec) xyz_src + tvec   (src/render.jl:51-62)
__global__ void plane_xyz_tgt_kernel(const float* __restrict__ xyz, const float* __restrict__ rvec,
                                     const float* __restrict__ tvec, int64_t per_b, int B, float* __restrict__ out) {
  __shared__ float Rs[9], ts[3];
  const int b = blockIdx.y;
  if (threadIdx.x == 0) {
    double r[3] = {rvec[3 * b], rvec[3 * b + 1], rvec[3 * b + 2]}, R[9];
    So3Aux aux;
    so3_exp_d(r, R, aux);
    for (int k = 0; k < 9; ++k) Rs[k] = (float)R[k];
    for (int k = 0; k < 3; ++k) ts[k] = tvec[3 * b + k];
  }
  __syncthreads();
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < per_b; i += (int64_t)gridDim.x * blockDim.x) {
    const float* p = xyz + ((int64_t)b * per_b + i) * 3;
    const float x = p[0], y = p[1], z = p[2];
    float* o = out + ((int64_t)b * per_b + i) * 3;
#pragma unroll
    for (int r = 0; r < 3; ++r) o[r] = Rs[3 * r] * x + Rs[3 * r + 1] * y + Rs[3 * r + 2] * z + ts[r];
  }
}

// homographies per (plane, item)   (src/render.jl:66-74)
__global__ void render_prep_kernel(const float* __restrict__ disp, int is_depth, const float* __restrict__ rvec,
                                   const float* __restrict__ tvec, RenderConst rc, int Np, int B,
                                   PlanePrep* __restrict__ out) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= Np * B) return;
  const int b = idx / Np;
  double r[3] = {rvec[3 * b], rvec[3 * b + 1], rvec[3 * b + 2]}, R[9], M[9], KM[9], Hts[9];
  So3Aux aux;
  so3_exp_d(r, R, aux);
  const double dsp = is_depth ? 1.0 / (double)disp[idx] : (double)disp[idx];  // 1 / depth_src
  for (int k = 0; k < 9; ++k) M[k] = R[k];
  // R - (t n^T) / (-depth) = R + t n^T * disparity,  n = [0,0,1]
  for (int i = 0; i < 3; ++i) M[3 * i + 2] += (double)tvec[3 * b + i] * dsp;
  mat3_mul(rc.K, M, KM);
  mat3_mul(KM, rc.Kinv, Hts);
  PlanePrep p;
  inv3_d(Hts, p.Hst);
  // G1 = K^T Hst^T
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      double v = 0;
      for (int k = 0; k < 3; ++k) v += rc.K[3 * k + i] * p.Hst[3 * j + k];
      p.G1[3 * i + j] = (float)v;
    }
  p.disp = (float)dsp;
  {
    double RK[9];
    mat3_mul(R, rc.Kinv, RK);
    for (int k = 0; k < 9; ++k) p.Ax[k] = (float)(RK[k] / dsp);
    for (int k = 0; k < 3; ++k) p.tx[k] = tvec[3 * b + k];
  }
  {
    const double i22 = 1.0 / p.Hst[8];
    double hn[8];
    for (int k = 0; k < 8; ++k) hn[k] = p.Hst[k] * i22;
    hn[0] -= 1.0; hn[4] -= 1.0;
    for (int k = 0; k < 8; ++k) p.hd[k] = hn[k];
    // the short path needs 1 + e = 1 + h20 u + h21 v away from zero over the whole image (its reciprocal starts from a float
    // seed and takes one Newton step): |e| < 1/4.  The size of the displacement does not matter: floor and fraction of the
    // coordinate are taken in double.  NaN / inf compare false.
    const double em = fabs(hn[6]) * (rc.W - 1) + fabs(hn[7]) * (rc.H - 1);
    double big = 0.0;
    for (int k = 0; k < 8; ++k) big = fmax(big, fabs(hn[k]));
    p.fast = (rc.W > 0 && em < 0.25 && big < 1e6) ? 1 : 0;
    p.pad_[0] = p.pad_[1] = p.pad_[2] = 0;
    p.pad0_[0] = p.pad0_[1] = p.pad1_[0] = p.pad1_[1] = 0.0f;
  }
  out[idx] = p;
}

struct RTaps {
  int o00, o01, o10, o11;
  float fx, fy, mx, my;
  double X0, X1, X2;  // homogeneous source coordinates
  float x, y;         // source pixel (0-based, as the reference computes it)
  float q0, q1;       // clamped sample point, 1-based
  int x0, y0;         // top-left cell
  bool valid;
};

// sample coordinates of target pixel (u, v) (0-based) under H_src_tgt; src/render.jl:76-88.
// Done in double: with random-texture inputs a 1e-5 px coordinate error is a 1e-5 output error.
template <int FLAGS = 0>
MDB_DEV RTaps render_taps(const double* __restrict__ H, int u, int v, int W, int Hh) {
  RTaps t;
  const double du = u, dv = v;
  t.X0 = fma(H[0], du, fma(H[1], dv, H[2]));
  t.X1 = fma(H[3], du, fma(H[4], dv, H[5]));
  t.X2 = fma(H[6], du, fma(H[7], dv, H[8]));
  // perspective divide: MUFU.RCP64H seed (2^-20) + two Newton steps = full double precision, ~6 FP64 ops
  // instead of two IEEE divisions (~60); the result differs from x / z by an ulp at most
  double zi;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(zi) : "d"(t.X2));
  zi = fma(zi, fma(-t.X2, zi, 1.0), zi);
  zi = fma(zi, fma(-t.X2, zi, 1.0), zi);
  const double x = t.X0 * zi, y = t.X1 * zi;
  if (FLAGS & RV_MINE_MASK)   // test/test_sample.jl:123-125
    t.valid = (x < (double)W) && (x > -1.0) && (y < (double)Hh) && (y > -1.0);
  else   // Julia: (x .< W .* x .>= 0) == (x < W x) & (W x >= 0)  ==  x > 0 for W > 1   
    t.valid = (x < W * x) && (W * x >= 0.0) && (y < Hh * y) && (Hh * y >= 0.0);
  double ix, iy;
  if (FLAGS & RV_MINE_GRID) {
    // grid = (x + 0.5) / (W/2) - 1 sampled with align_corners=false, border padding (test/test_sample.jl:131-137):
    // ((g + 1) W - 1) / 2 = x, the source pixel itself
    ix = x; iy = y;
  } else {
    // grid = (x + 0.5) / (W/2) [no "-1"], then align-corners=true un-normalisation
    ix = ((x + 0.5) * (2.0 / W) + 1.0) * 0.5 * (W - 1); iy = ((y + 0.5) * (2.0 / Hh) + 1.0) * 0.5 * (Hh - 1);
  }
  t.mx = (ix > 0.0 && ix < (double)(W - 1)) ? 1.0f : 0.0f;
  t.my = (iy > 0.0 && iy < (double)(Hh - 1)) ? 1.0f : 0.0f;
  const double cx = fmin(fmax(ix, 0.0), (double)(W - 1)), cy = fmin(fmax(iy, 0.0), (double)(Hh - 1));
  const double fx0 = floor(cx), fy0 = floor(cy);
  t.fx = (float)(cx - fx0);
  t.fy = (float)(cy - fy0);
  const int x0 = (int)fx0, y0 = (int)fy0, x1 = min(x0 + 1, W - 1), y1 = min(y0 + 1, Hh - 1);
  t.o00 = y0 * W + x0; t.o01 = y0 * W + x1; t.o10 = y1 * W + x0; t.o11 = y1 * W + x1;
  t.x0 = x0; t.y0 = y0;
  t.x = (float)x;
  t.y = (float)y;
  t.q0 = (float)(cx + 1.0);   // 1-based (create_meshgrid) coordinates of the clamped sample point
  t.q1 = (float)(cy + 1.0);
  return t;
}

MDB_DEV float bil(const RTaps& t, float v00, float v01, float v10, float v11) {
  const float top = fmaf(t.fx, v01 - v00, v00), bot = fmaf(t.fx, v11 - v10, v10);
  return fmaf(t.fy, bot - top, top);
}

struct RenderArgs {
  const float* rgb; const float* sigma; const float* xyz;   // xyz unused by the FUSED instantiations
  const PlanePrep* prep;
  int W, H, Np, B;
  float* rgb_out; float* tacc; int* mask;
  float* edepth;      // [B][1][H][W] expected depth sum w z / (sum w + depth_eps), or null
  float* wsum;        // [B][1][H][W] sum w, or null
  float depth_eps;
  double sx, sy, cx0, cy0;   // grid as written: coordinate = x sx + cx0, sx = (W-1)/W, cx0 = 0.5 sx + 0.5 (W-1) (host-computed)
};

// R7 forward.  A pixel's 32 planes are a serial chain (front-to-back compositing, and the distance to the
// NEXT plane's sampled xyz is needed before a plane can be composited), and W H B threads alone do not
// fill the machine (106 K at config 4), so the chain is cut into RF_G plane groups that run as separate
// threads: each composites its own planes from transparency 1, then the groups are stitched through
// shared memory -- transparency_acc_k = (product of the earlier groups' (T + 1e-6)) x local value,
// rgb = sum_g prefix_g rgb_g (and likewise sum w, sum w z).  The first plane of the following group is
// sampled for its xyz only.
#ifndef RF_GROUPS
#define RF_GROUPS 4      // measured at config 4 (as written / MINE, us): 4 groups 59.6 / 69.6, 8: 64.5 / 74.7, 16: 76.9 / 88.5
#endif
#ifndef RF_MINB
#define RF_MINB 8        // 128-thread CTAs per SM the register allocation aims for (64 registers; 6 -> 85 registers: 70.6 us)
#endif
constexpr int RF_PIX = 32, RF_G = RF_GROUPS, RF_THREADS = RF_PIX * RF_G;
constexpr int RF_MAXKPG = (64 + RF_G - 1) / RF_G;   // planes per group at Np = 64

struct RSample { float rgb[3], sig, xyz[3]; };

// ---- forward taps -------------------------------------------------------------------------------------------------
// Top-left cell (x0 <= W-2, y0 <= H-2: the other three taps are +1, +W, +W+1, immediates and one add) and the weights.
struct FTap {
  int o;            // y0 W + x0
  float fx, fy;     // in [0, 1]; 1 only on the last column / row
  float q0, q1;     // clamped sample point, 1-based (create_meshgrid coordinates; FUSED xyz)
  bool valid;
};

// What a thread knows about its own pixel: the sample coordinate of plane k is  bi + (bf + t s)  with t = x - u (or y - v)
// the plane's displacement.   MINE grid: the coordinate is the source pixel itself, bi = u, bf = 0, s = 1.
// Grid as written (no "-1", align-corners un-normalisation):  (x + 0.5)(W-1)/W + (W-1)/2, whose
// plane-independent part u (W-1)/W + 0.5 (W-1)/W + (W-1)/2 is split ONCE per thread, in double, into integer + fraction.
struct PixBase { int u, v, bxi, byi; float uf, vf; double ud, vd, bxf, byf; };

template <int FLAGS>
MDB_DEV PixBase pix_base(int u, int v, double sx, double sy, double cx0, double cy0) {
  PixBase p;
  p.u = u; p.v = v; p.uf = (float)u; p.vf = (float)v; p.ud = (double)u; p.vd = (double)v;
  if (FLAGS & RV_MINE_GRID) {
    p.bxi = u; p.byi = v; p.bxf = p.byf = 0.0;
  } else {
    const double bx = fma(p.ud, sx, cx0), by = fma(p.vd, sy, cy0);
    const double fx = floor(bx), fy = floor(by);
    p.bxi = (int)fx; p.byi = (int)fy; p.bxf = bx - fx; p.byf = by - fy;
  }
  return p;
}

// One axis, coordinate = bi + r with r (double) small: cell index (<= n-2), weight in [0, 1], the clamp mask of the pullback
// (1 where 0 < coordinate < n-1) and the 1-based clamped point.  Floor and fraction in double -- the LAST plane's
// transmittance exp(-1000 sigma) (src/render.jl:36) turns 1e-7 of sampled sigma into 4e-5 of output, so the interpolation
// weight has to be good to a few 1e-8 -- the clamp on the integer part.
struct GridC { double sx, sy, cx0, cy0; };
MDB_DEV void axis_cell(double r, int bi, int n, int& i0, float& f, float& m, float& q) {
  const double fl = floor(r);
  i0 = bi + (int)fl;
  f = (float)(r - fl);
  m = (i0 >= 0 && i0 <= n - 2 && (i0 > 0 || f > 0.0f)) ? 1.0f : 0.0f;
  if (i0 < 0) { i0 = 0; f = 0.0f; }                  // coordinate < 0: clamped onto the first sample
  if (i0 > n - 2) { i0 = n - 2; f = 1.0f; }          // coordinate >= n - 1: the last sample = cell n-2 with weight 1
  q = (float)(i0 + 1) + f;
}

// Short path (PlanePrep::fast): displacement form, all in double but cheap: numerators 8 operations, reciprocal from a
// float seed + one Newton step (|e| < 1/4), 2 products, 2 FMAs; no comparisons, clamps or un-normalisation in double.
// Fills everything the forward AND the pullback want (the forward's unused fields cost nothing: dead code): the four tap
// offsets in the cell convention (top-left cell <= (W-2, H-2), the others +1, +W, +W+1), the clamp masks, homogeneous
// coordinates up to the common factor 1 / H22 (the pullback is invariant to it).
template <int FLAGS>
MDB_DEV RTaps render_taps_fast(const double* __restrict__ hd, const PixBase& p, const GridC& gc, int W, int H) {
  const double2 c0 = __ldg(reinterpret_cast<const double2*>(hd)), c1 = __ldg(reinterpret_cast<const double2*>(hd) + 1);
  const double2 c2 = __ldg(reinterpret_cast<const double2*>(hd) + 2), c3 = __ldg(reinterpret_cast<const double2*>(hd) + 3);
  const double e = fma(c3.x, p.ud, c3.y * p.vd), zd = 1.0 + e;      // zd in (3/4, 5/4)
  const double N0 = fma(-p.ud, e, fma(c0.x, p.ud, fma(c0.y, p.vd, c1.x)));
  const double N1 = fma(-p.vd, e, fma(c1.y, p.ud, fma(c2.x, p.vd, c2.y)));
  float zf;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(zf) : "f"((float)zd));
  double zi = (double)zf;
  zi = fma(zi, fma(-zd, zi, 1.0), zi);                              // relative error ~1e-14
  const double txd = N0 * zi, tyd = N1 * zi;                        // x - u, y - v
  const float tx = (float)txd, ty = (float)tyd;
  RTaps t;
  t.X0 = fma(p.ud, zd, N0); t.X1 = fma(p.vd, zd, N1); t.X2 = zd;
  t.x = p.uf + tx; t.y = p.vf + ty;
  if (FLAGS & RV_MINE_MASK) t.valid = (tx < (float)(W - p.u)) && (tx > (float)(-1 - p.u)) && (ty < (float)(H - p.v)) && (ty > (float)(-1 - p.v));
  else t.valid = (tx > -p.uf) && (ty > -p.vf);    // x > 0 & y > 0
  axis_cell((FLAGS & RV_MINE_GRID) ? txd : fma(txd, gc.sx, p.bxf), p.bxi, W, t.x0, t.fx, t.mx, t.q0);
  axis_cell((FLAGS & RV_MINE_GRID) ? tyd : fma(tyd, gc.sy, p.byf), p.byi, H, t.y0, t.fy, t.my, t.q1);
  t.o00 = t.y0 * W + t.x0; t.o01 = t.o00 + 1; t.o10 = t.o00 + W; t.o11 = t.o10 + 1;
  return t;
}

template <int FLAGS>
MDB_DEV FTap ftap_fast(const double* __restrict__ hd, const PixBase& p, const RenderArgs& a) {
  const GridC gc{a.sx, a.sy, a.cx0, a.cy0};
  const RTaps r = render_taps_fast<FLAGS>(hd, p, gc, a.W, a.H);
  FTap t;
  t.o = r.o00; t.fx = r.fx; t.fy = r.fy; t.q0 = r.q0; t.q1 = r.q1; t.valid = r.valid;
  return t;
}

// double path: any homography (the pullback and mdb_sample use render_taps, the same arithmetic)
template <int FLAGS>
__device__ __noinline__ FTap ftap_precise(const double* __restrict__ Hm, const PixBase& p, int W, int H) {
  const RTaps r = render_taps<FLAGS>(Hm, p.u, p.v, W, H);
  FTap t;
  t.valid = r.valid;
  int x0 = r.x0, y0 = r.y0;
  t.fx = r.fx; t.fy = r.fy;
  if (x0 > W - 2) { x0 = W - 2; t.fx = 1.0f; }
  if (y0 > H - 2) { y0 = H - 2; t.fy = 1.0f; }
  t.q0 = r.q0; t.q1 = r.q1;
  t.o = y0 * W + x0;
  return t;
}

template <int FLAGS>
MDB_DEV FTap ftap(const PlanePrep* __restrict__ pp, const PixBase& p, const RenderArgs& a) {
  if (__ldg(&pp->fast)) return ftap_fast<FLAGS>(pp->hd, p, a);   // warp-uniform
  return ftap_precise<FLAGS>(pp->Hst, p, a.W, a.H);
}

struct FW { float w00, w01, w10, w11; };
MDB_DEV FW ftap_weights(const FTap& t) {
  FW w;
  w.w11 = t.fx * t.fy; w.w01 = t.fx - w.w11; w.w10 = t.fy - w.w11; w.w00 = (1.0f - t.fx) - w.w10;
  return w;
}

// FUSED: xyz_tgt is affine in the source pixel, and the bilinear interpolant of an affine map at the
// (border-clamped) sample point is the map itself there: no xyz loads (12 of the 28 per sample)
// the arrays of the current plane: advanced by one plane per iteration instead of re-deriving 64-bit products
struct PlanePtr { const float* rgb; const float* sig; const float* xyz; const PlanePrep* prep; };

template <bool FUSED>
MDB_DEV void render_sample_xyz(const PlanePtr& pl, const FTap& t, const FW& w, int W, float* xyz) {
  if (FUSED) {
    const float4* ax4 = reinterpret_cast<const float4*>(pl.prep->Ax);   // Ax[9], tx[3]: 48 bytes at offset 128
    const float4 c0 = __ldg(ax4), c1 = __ldg(ax4 + 1), c2 = __ldg(ax4 + 2);
    xyz[0] = fmaf(c0.x, t.q0, fmaf(c0.y, t.q1, c0.z + c2.y));
    xyz[1] = fmaf(c0.w, t.q0, fmaf(c1.x, t.q1, c1.y + c2.z));
    xyz[2] = fmaf(c1.z, t.q0, fmaf(c1.w, t.q1, c2.x + c2.w));
    return;
  }
  const float* p0 = pl.xyz + 3 * t.o;
  const float* p1 = p0 + 3 * W;
#pragma unroll
  for (int c = 0; c < 3; ++c)
    xyz[c] = fmaf(w.w11, ldg(p1 + 3 + c), fmaf(w.w10, ldg(p1 + c), fmaf(w.w01, ldg(p0 + 3 + c), w.w00 * ldg(p0 + c))));
}

template <bool FUSED>
MDB_DEV void render_sample7(const PlanePtr& pl, const FTap& t, int W, int WH, RSample& s) {
  const FW w = ftap_weights(t);
  const float* p0 = pl.rgb + t.o;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const float* q0 = p0 + c * WH;
    const float* q1 = q0 + W;
    s.rgb[c] = fmaf(w.w11, ldg(q1 + 1), fmaf(w.w10, ldg(q1), fmaf(w.w01, ldg(q0 + 1), w.w00 * ldg(q0))));
  }
  const float* q0 = pl.sig + t.o;
  const float* q1 = q0 + W;
  const float sg = fmaf(w.w11, ldg(q1 + 1), fmaf(w.w10, ldg(q1), fmaf(w.w01, ldg(q0 + 1), w.w00 * ldg(q0))));
  s.sig = sg >= 0.0f ? sg : 0.0f;                                  // sigma .* (sigma .>= 0)
  render_sample_xyz<FUSED>(pl, t, w, W, s.xyz);
}

#ifndef RF_PREFETCH
#define RF_PREFETCH 1    // 1: prefetch.global.L2, 2: prefetch.global.L1 (same result) of the next plane's tap lines, 0: none
#endif
MDB_DEV void prefetch_line(const void* p) {
#if RF_PREFETCH == 2
  asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
#else
  asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
#endif
}
template <bool FUSED>
MDB_DEV void render_prefetch(const PlanePtr& pl, int o, int W, int WH) {
  const float* r = pl.rgb + 3 * WH + o;      // the next plane's arrays at this plane's cell
#pragma unroll
  for (int c = 0; c < 3; ++c) { prefetch_line(r + c * WH); prefetch_line(r + c * WH + W); }
  const float* sg = pl.sig + WH + o;
  prefetch_line(sg); prefetch_line(sg + W);
  if (!FUSED) {
    const float* x = pl.xyz + 3 * WH + 3 * o;
    prefetch_line(x); prefetch_line(x + 3 * W);
  }
}

// MUFU-based square root and exponential: ~2 ulp each.  The transmittance exp(-dist sigma) is in (0, 1], so its ABSOLUTE
// error stays below 2e-7 whatever the argument (the relative error of ex2 grows with |argument|, the value shrinks faster).
MDB_DEV float sqrt_approx(float x) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
MDB_DEV float exp_neg_approx(float x) { float r; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x * -1.4426950408889634f)); return r; }

template <int FLAGS, bool FUSED>
__global__ void __launch_bounds__(RF_THREADS, RF_MINB) render_fwd_kernel(const __grid_constant__ RenderArgs a) {
  __shared__ float s_T[RF_G][RF_PIX];            // product of (T + 1e-6) over the group's planes
  __shared__ float s_out[RF_G][5][RF_PIX];       // rgb, sum w, sum w z of the group
  __shared__ int s_nv[RF_G][RF_PIX];
  __shared__ float s_tl[RF_MAXKPG][RF_THREADS];  // local transparency_acc of the group's planes
  const int lp = threadIdx.x, g = threadIdx.y, tid = g * RF_PIX + lp;   // a warp = 32 consecutive pixels of a row, one group
  const int u = blockIdx.x * RF_PIX + lp, v = blockIdx.y, b = blockIdx.z;
  const int WH = a.W * a.H, pix = v * a.W + u;
  const bool live = u < a.W;
  const int kpg = (a.Np + RF_G - 1) / RF_G, k0 = min(g * kpg, a.Np), k1 = min(k0 + kpg, a.Np);
  float out[5] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f}, tacc = 1.0f;
  int nvalid = 0;
  if (live && k0 < k1) {
    const PixBase pb = pix_base<FLAGS>(u, v, a.sx, a.sy, a.cx0, a.cy0);
    const size_t kb0 = (size_t)b * a.Np + k0;
    PlanePtr pl{a.rgb + kb0 * 3 * WH, a.sigma + kb0 * WH, FUSED ? nullptr : a.xyz + kb0 * WH * 3, a.prep + kb0};
    RSample p;
    {
      const FTap t = ftap<FLAGS>(pl.prep, pb, a);
      nvalid += t.valid ? 1 : 0;
      render_sample7<FUSED>(pl, t, a.W, WH, p);
    }
    for (int k = k0 + 1; k <= k1; ++k) {
      RSample c;
      float dist = 1e3f;                                             // last plane (src/render.jl:36)
      if (k < a.Np) {
        pl.rgb += 3 * WH; pl.sig += WH; pl.prep += 1;
        if (!FUSED) pl.xyz += 3 * WH;
        const FTap t = ftap<FLAGS>(pl.prep, pb, a);
#if RF_PREFETCH
        // the NEXT plane's taps sit within a few pixels of this plane's (adjacent planes differ by a small parallax): ask for
        // their lines now, one plane (~250 instructions x the resident warps) ahead of the loads -- no registers held
        // (pays only where the taps really go to DRAM: MINE grid with xyz_tgt given 72.7 -> 61.6 us; the other variants lose 5 %)
        if ((FLAGS & RV_MINE_GRID) && !FUSED && k + 1 < k1) render_prefetch<FUSED>(pl, t.o, a.W, WH);
#endif
        if (k < k1) { nvalid += t.valid ? 1 : 0; render_sample7<FUSED>(pl, t, a.W, WH, c); }
        else render_sample_xyz<FUSED>(pl, t, ftap_weights(t), a.W, c.xyz);  // first plane of the next group
        const float dx = c.xyz[0] - p.xyz[0], dy = c.xyz[1] - p.xyz[1], dz = c.xyz[2] - p.xyz[2];
        dist = sqrt_approx(fmaf(dx, dx, fmaf(dy, dy, dz * dz)));
      }
      const float T = exp_neg_approx(dist * p.sig);
      s_tl[k - 1 - k0][tid] = tacc;
      const float w = tacc * (1.0f - T);
#pragma unroll
      for (int ch = 0; ch < 3; ++ch) out[ch] = fmaf(w, p.rgb[ch], out[ch]);
      out[3] += w;
      out[4] = fmaf(w, p.xyz[2], out[4]);
      tacc *= (T + 1e-6f);
      p = c;
    }
  }
  s_T[g][lp] = tacc;
#pragma unroll
  for (int ch = 0; ch < 5; ++ch) s_out[g][ch][lp] = out[ch];
  s_nv[g][lp] = nvalid;
  __syncthreads();
  if (!live) return;
  float prefix = 1.0f;
  for (int j = 0; j < g; ++j) prefix *= s_T[j][lp];
  for (int k = k0; k < k1; ++k) st_cs(a.tacc + (size_t)(b * a.Np + k) * WH + pix, prefix * s_tl[k - k0][tid]);
  if (g == 0) {
    float pf = 1.0f, o[5] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
    int nv = 0;
    for (int j = 0; j < RF_G; ++j) {
#pragma unroll
      for (int ch = 0; ch < 5; ++ch) o[ch] = fmaf(pf, s_out[j][ch][lp], o[ch]);
      pf *= s_T[j][lp];
      nv += s_nv[j][lp];
    }
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) a.rgb_out[((size_t)b * 3 + ch) * WH + pix] = o[ch];
    a.mask[(size_t)b * WH + pix] = nv;
    if (a.edepth) a.edepth[(size_t)b * WH + pix] = o[4] / (o[3] + a.depth_eps);
    if (a.wsum) a.wsum[(size_t)b * WH + pix] = o[3];
  }
}

}  // namespace

#include "render_bwd.inl"
