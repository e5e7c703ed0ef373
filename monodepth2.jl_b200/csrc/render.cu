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
  float pad[5];
};

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
  for (int k = 0; k < 5; ++k) p.pad[k] = 0.0f;
  out[idx] = p;
}

struct RTaps {
  int o00, o01, o10, o11;
  float fx, fy, mx, my;
  double X0, X1, X2;  // homogeneous source coordinates
  float x, y;         // source pixel (0-based, as the reference computes it)
  bool valid;
};

// sample coordinates of target pixel (u, v) (0-based) under H_src_tgt; src/render.jl:76-88.
// Done in double: with random-texture inputs a 1e-5 px coordinate error is a 1e-5 output error.
MDB_DEV RTaps render_taps(const double* __restrict__ H, int u, int v, int W, int Hh) {
  RTaps t;
  const double du = u, dv = v;
  t.X0 = fma(H[0], du, fma(H[1], dv, H[2]));
  t.X1 = fma(H[3], du, fma(H[4], dv, H[5]));
  t.X2 = fma(H[6], du, fma(H[7], dv, H[8]));
  const double x = t.X0 / t.X2, y = t.X1 / t.X2;
  // Julia: (x .< W .* x .>= 0) == (x < W x) & (W x >= 0)  ==  x > 0 for W > 1   
  t.valid = (x < W * x) && (W * x >= 0.0) && (y < Hh * y) && (Hh * y >= 0.0);
  // grid = (x + 0.5) / (W/2) [no "-1"], then align-corners=true un-normalisation
  const double ix = ((x + 0.5) / (0.5 * W) + 1.0) * 0.5 * (W - 1), iy = ((y + 0.5) / (0.5 * Hh) + 1.0) * 0.5 * (Hh - 1);
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
  return t;
}

MDB_DEV float bil(const RTaps& t, float v00, float v01, float v10, float v11) {
  const float top = fmaf(t.fx, v01 - v00, v00), bot = fmaf(t.fx, v11 - v10, v10);
  return fmaf(t.fy, bot - top, top);
}

struct RenderArgs {
  const float* rgb; const float* sigma; const float* xyz;
  const PlanePrep* prep;
  int W, H, Np, B;
  float* rgb_out; float* tacc; int* mask;
};

constexpr int RENDER_THREADS = 128;

// R7 forward: sample 7 channels per plane, composite front to back with a one-plane lag (the
// distance to the NEXT plane's sampled xyz is needed before a plane can be composited).
__global__ void __launch_bounds__(RENDER_THREADS) render_fwd_kernel(const __grid_constant__ RenderArgs a) {
  const int pix = blockIdx.x * RENDER_THREADS + threadIdx.x;
  const int b = blockIdx.y;
  const int WH = a.W * a.H;
  if (pix >= WH) return;
  const int u = pix % a.W, v = pix / a.W;
  float out[3] = {0.0f, 0.0f, 0.0f};
  float tacc = 1.0f;
  float p_rgb[3] = {0, 0, 0}, p_sig = 0.0f, p_xyz[3] = {0, 0, 0};
  int nvalid = 0;
  for (int k = 0; k <= a.Np; ++k) {
    float c_rgb[3] = {0, 0, 0}, c_sig = 0.0f, c_xyz[3] = {0, 0, 0};
    if (k < a.Np) {
      const int kb = b * a.Np + k;
      const RTaps t = render_taps(a.prep[kb].Hst, u, v, a.W, a.H);
      nvalid += t.valid ? 1 : 0;
      const float* pr = a.rgb + (size_t)kb * 3 * WH;
#pragma unroll
      for (int c = 0; c < 3; ++c)
        c_rgb[c] = bil(t, ldg(pr + c * WH + t.o00), ldg(pr + c * WH + t.o01), ldg(pr + c * WH + t.o10), ldg(pr + c * WH + t.o11));
      const float* ps = a.sigma + (size_t)kb * WH;
      c_sig = bil(t, ldg(ps + t.o00), ldg(ps + t.o01), ldg(ps + t.o10), ldg(ps + t.o11));
      c_sig = c_sig >= 0.0f ? c_sig : 0.0f;                       // sigma .* (sigma .>= 0)
      const float* px = a.xyz + (size_t)kb * WH * 3;
#pragma unroll
      for (int c = 0; c < 3; ++c)
        c_xyz[c] = bil(t, ldg(px + 3 * t.o00 + c), ldg(px + 3 * t.o01 + c), ldg(px + 3 * t.o10 + c), ldg(px + 3 * t.o11 + c));
    }
    if (k > 0) {  // composite plane k-1
      float dist = 1e3f;                                           // last plane (src/render.jl:36)
      if (k < a.Np) {
        const float dx = c_xyz[0] - p_xyz[0], dy = c_xyz[1] - p_xyz[1], dz = c_xyz[2] - p_xyz[2];
        dist = sqrtf(dx * dx + dy * dy + dz * dz);
      }
      const float T = expf(-dist * p_sig);
      st_cs(a.tacc + (size_t)(b * a.Np + k - 1) * WH + pix, tacc);
      const float w = tacc * (1.0f - T);
#pragma unroll
      for (int c = 0; c < 3; ++c) out[c] = fmaf(w, p_rgb[c], out[c]);
      tacc *= (T + 1e-6f);
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) { p_rgb[c] = c_rgb[c]; p_xyz[c] = c_xyz[c]; }
    p_sig = c_sig;
  }
#pragma unroll
  for (int c = 0; c < 3; ++c) a.rgb_out[((size_t)b * 3 + c) * WH + pix] = out[c];
  a.mask[(size_t)b * WH + pix] = nvalid;
}

}  // namespace

#include "render_bwd.inl"
