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

namespace {

// per (plane, batch item) constants
struct PlanePrep {
  double Hst[9];   // H_src_tgt = inv(K (R + t n^T / depth) K_inv), row-major
  float G1[9];     // K^T H_src_tgt^T : pulls d/d(homogeneous source coords) back to d/dM
  float disp;      // 1 / depth_src
  // xyz_tgt of this plane as a function of the 1-based source pixel m = (w, h, 1):  Ax m + tx  with
  // Ax = R K_inv / disparity  (src/render.jl:25-30, 51-62) -- what the producers folded into the kernels use
  float Ax[9];
  float tx[3];
};

// variants of the sampler / outputs (mdb_render_opts in include/mdb200.h)
constexpr int RV_MINE_GRID = 1, RV_MINE_MASK = 2;

struct RenderConst {
  double K[9], Kinv[9];  // math row-major
  float Kinv_f[9];
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
  out[idx] = p;
}

struct RTaps {
  int o00, o01, o10, o11;
  float fx, fy, mx, my;
  double X0, X1, X2;  // homogeneous source coordinates
  float x, y;         // source pixel (0-based, as the reference computes it)
  float q0, q1;       // clamped sample point, 1-based
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
};

// R7 forward.  A pixel's 32 planes are a serial chain (front-to-back compositing, and the distance to the
// NEXT plane's sampled xyz is needed before a plane can be composited), and W H B threads alone do not
// fill the machine (106 K at config 4), so the chain is cut into RF_G plane groups that run as separate
// threads: each composites its own planes from transparency 1, then the groups are stitched through
// shared memory -- transparency_acc_k = (product of the earlier groups' (T + 1e-6)) x local value,
// rgb = sum_g prefix_g rgb_g (and likewise sum w, sum w z).  The first plane of the following group is
// sampled for its xyz only.
constexpr int RF_PIX = 64, RF_G = 4, RF_THREADS = RF_PIX * RF_G;
constexpr int RF_MAXKPG = 16;          // planes per group: ceil(64 / 4)

struct RSample { float rgb[3], sig, xyz[3]; };

// FUSED: xyz_tgt is affine in the source pixel, and the bilinear interpolant of an affine map at the
// (border-clamped) sample point is the map itself there: no xyz loads (12 of the 28 per sample)
template <bool FUSED>
MDB_DEV void render_sample_xyz(const RenderArgs& a, const RTaps& t, int kb, int WH, float* xyz) {
  if (FUSED) {
    const PlanePrep& pp = a.prep[kb];
#pragma unroll
    for (int c = 0; c < 3; ++c)
      xyz[c] = fmaf(ldg(pp.Ax + 3 * c), t.q0, fmaf(ldg(pp.Ax + 3 * c + 1), t.q1, ldg(pp.Ax + 3 * c + 2) + ldg(pp.tx + c)));
    return;
  }
  const float* px = a.xyz + (size_t)kb * WH * 3;
#pragma unroll
  for (int c = 0; c < 3; ++c)
    xyz[c] = bil(t, ldg(px + 3 * t.o00 + c), ldg(px + 3 * t.o01 + c), ldg(px + 3 * t.o10 + c), ldg(px + 3 * t.o11 + c));
}

template <bool FUSED>
MDB_DEV void render_sample7(const RenderArgs& a, const RTaps& t, int kb, int WH, RSample& s) {
  const float* pr = a.rgb + (size_t)kb * 3 * WH;
#pragma unroll
  for (int c = 0; c < 3; ++c)
    s.rgb[c] = bil(t, ldg(pr + c * WH + t.o00), ldg(pr + c * WH + t.o01), ldg(pr + c * WH + t.o10), ldg(pr + c * WH + t.o11));
  const float* ps = a.sigma + (size_t)kb * WH;
  const float sg = bil(t, ldg(ps + t.o00), ldg(ps + t.o01), ldg(ps + t.o10), ldg(ps + t.o11));
  s.sig = sg >= 0.0f ? sg : 0.0f;                                  // sigma .* (sigma .>= 0)
  render_sample_xyz<FUSED>(a, t, kb, WH, s.xyz);
}

template <int FLAGS, bool FUSED>
__global__ void __launch_bounds__(RF_THREADS) render_fwd_kernel(const __grid_constant__ RenderArgs a) {
  __shared__ float s_T[RF_G][RF_PIX];            // product of (T + 1e-6) over the group's planes
  __shared__ float s_out[RF_G][5][RF_PIX];       // rgb, sum w, sum w z of the group
  __shared__ int s_nv[RF_G][RF_PIX];
  __shared__ float s_tl[RF_MAXKPG][RF_THREADS];  // local transparency_acc of the group's planes
  const int lp = threadIdx.x, g = threadIdx.y, tid = g * RF_PIX + lp;   // a warp = 32 consecutive pixels, one group
  const int pix = blockIdx.x * RF_PIX + lp, b = blockIdx.y;
  const int WH = a.W * a.H;
  const bool live = pix < WH;
  const int kpg = (a.Np + RF_G - 1) / RF_G, k0 = min(g * kpg, a.Np), k1 = min(k0 + kpg, a.Np);
  float out[5] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f}, tacc = 1.0f;
  int nvalid = 0;
  if (live && k0 < k1) {
    const int u = pix % a.W, v = pix / a.W;
    RSample p;
    {
      const RTaps t = render_taps<FLAGS>(a.prep[b * a.Np + k0].Hst, u, v, a.W, a.H);
      nvalid += t.valid ? 1 : 0;
      render_sample7<FUSED>(a, t, b * a.Np + k0, WH, p);
    }
    for (int k = k0 + 1; k <= k1; ++k) {
      RSample c;
      float dist = 1e3f;                                             // last plane (src/render.jl:36)
      if (k < a.Np) {
        const RTaps t = render_taps<FLAGS>(a.prep[b * a.Np + k].Hst, u, v, a.W, a.H);
        if (k < k1) { nvalid += t.valid ? 1 : 0; render_sample7<FUSED>(a, t, b * a.Np + k, WH, c); }
        else render_sample_xyz<FUSED>(a, t, b * a.Np + k, WH, c.xyz);  // first plane of the next group
        const float dx = c.xyz[0] - p.xyz[0], dy = c.xyz[1] - p.xyz[1], dz = c.xyz[2] - p.xyz[2];
        dist = sqrtf(dx * dx + dy * dy + dz * dz);
      }
      const float T = expf(-dist * p.sig);
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
