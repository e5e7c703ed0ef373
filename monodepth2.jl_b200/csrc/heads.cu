// heads.cu -- SURVEY 8(f) rank 4: the small operators on the network side of the path.
//  * PoseDecoder tail (src/pose_decoder.jl:27-32): spatial mean of the (w,h,6,N) conv output, x 1e-2,
//    split into Pose(rvec (3,N), tvec (3,1,N)); and its pullback.
//  * per-item mean of a (W,H,1,N) map: the reduction of `find_static` (src/dtk.jl:51-69),
//    mean(automasking_loss(ssim, x, target; source_ids)) > alpha.
// Both are deterministic (fixed-order sums in double).
#include "mdb_common.cuh"
#include <cooperative_groups.h>

namespace {

constexpr int HEAD_THREADS = 256;

__device__ double block_sum_d(double v, double* red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(MDB_FULL, v, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  double s = 0.0;
  if (threadIdx.x == 0)
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) s += red[w];
  __syncthreads();
  return s;   // valid in thread 0
}

// one block per (channel c of 6, item n): conv_out is [N][6][h][w]
__global__ void __launch_bounds__(HEAD_THREADS)
pose_head_kernel(const float* __restrict__ y, int wh, float scale, float* __restrict__ rvec, float* __restrict__ tvec) {
  __shared__ double red[HEAD_THREADS / 32];
  const int c = blockIdx.x, n = blockIdx.y;
  const float* p = y + ((size_t)n * 6 + c) * wh;
  double acc = 0.0;
  for (int i = threadIdx.x; i < wh; i += HEAD_THREADS) acc += (double)ldg(p + i);
  const double s = block_sum_d(acc, red);
  if (threadIdx.x == 0) {
    // eltype(T)(1e-2) .* mean(...): the mean is formed first, then scaled, both in Float32 in the reference
    const float m = (float)(s / (double)wh);
    const float v = scale * m;
    if (c < 3) rvec[3 * n + c] = v; else tvec[3 * n + (c - 3)] = v;
  }
}

__global__ void pose_head_bwd_kernel(const float* __restrict__ d_rvec, const float* __restrict__ d_tvec, int wh, int N,
                                     float scale, float* __restrict__ d_y) {
  const size_t total = (size_t)N * 6 * wh;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int nc = (int)(i / wh), n = nc / 6, c = nc - 6 * n;
    const float g = c < 3 ? d_rvec[3 * n + c] : d_tvec[3 * n + (c - 3)];
    d_y[i] = g * scale / (float)wh;
  }
}

// One thread-block CLUSTER of 8 CTAs per item: each CTA sums a slab (128-bit loads where aligned), the
// partial sums meet in CTA 0 through distributed shared memory in rank order -- one launch, no workspace,
// no atomics, bit-reproducible.
constexpr int IM_CLUSTER = 8;

__global__ void __cluster_dims__(IM_CLUSTER, 1, 1) __launch_bounds__(HEAD_THREADS)
item_mean_kernel(const float* __restrict__ map, long long per_item, float* __restrict__ out) {
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  __shared__ double red[HEAD_THREADS / 32];
  __shared__ double slab_sum;
  const unsigned rank = cluster.block_rank();
  const long long item = blockIdx.x / IM_CLUSTER;
  const float* p = map + (size_t)item * per_item;
  const long long chunk = (per_item + IM_CLUSTER - 1) / IM_CLUSTER;
  const long long lo = rank * chunk, hi = lo + chunk < per_item ? lo + chunk : per_item;
  double acc = 0.0;
  if (lo < hi) {
    const float* q = p + lo;
    const long long n = hi - lo;
    long long head = (4 - (((uintptr_t)q >> 2) & 3)) & 3;          // elements before the first 16-byte boundary
    if (head > n) head = n;
    for (long long i = threadIdx.x; i < head; i += HEAD_THREADS) acc += (double)ldg(q + i);
    const float4* q4 = reinterpret_cast<const float4*>(q + head);
    const long long n4 = (n - head) / 4;
    float s0 = 0.0f, s1 = 0.0f, s2 = 0.0f, s3 = 0.0f;
    long long i = threadIdx.x;
    for (; i < n4; i += HEAD_THREADS) {
      const float4 v = __ldg(q4 + i);
      s0 += v.x; s1 += v.y; s2 += v.z; s3 += v.w;
      if ((i / HEAD_THREADS & 63) == 63) { acc += ((double)s0 + s1) + ((double)s2 + s3); s0 = s1 = s2 = s3 = 0.0f; }   // bound the float run
    }
    acc += ((double)s0 + s1) + ((double)s2 + s3);
    for (long long j = head + 4 * n4 + threadIdx.x; j < n; j += HEAD_THREADS) acc += (double)ldg(q + j);
  }
  const double s = block_sum_d(acc, red);
  if (threadIdx.x == 0) slab_sum = s;
  cluster.sync();
  if (rank == 0 && threadIdx.x == 0) {
    double tot = 0.0;
    for (unsigned r = 0; r < IM_CLUSTER; ++r) tot += *cluster.map_shared_rank(&slab_sum, r);
    out[item] = (float)(tot / (double)per_item);
  }
  cluster.sync();   // keep every CTA's shared memory alive until rank 0 has read it
}

}  // namespace

extern "C" {

int32_t mdb_pose_head(mdb_handle* h, const float* conv_out, int32_t w, int32_t hh, int32_t N, float scale, float* rvec,
                      float* tvec, void* stream) {
  MDB_REQUIRE(h, h != nullptr, MDB_E_INVALID_ARG, "null handle");
  MDB_REQUIRE(h, conv_out && rvec && tvec, MDB_E_INVALID_ARG, "mdb_pose_head: null pointer");
  MDB_REQUIRE(h, w >= 1 && hh >= 1 && N >= 1 && N <= 65535, MDB_E_SHAPE, "mdb_pose_head: need w,h,N >= 1, N <= 65535");
  pose_head_kernel<<<dim3(6, N), HEAD_THREADS, 0, (cudaStream_t)stream>>>(conv_out, w * hh, scale, rvec, tvec);
  MDB_LAUNCHED(h);
  return MDB_OK;
}

int32_t mdb_pose_head_bwd(mdb_handle* h, const float* d_rvec, const float* d_tvec, int32_t w, int32_t hh, int32_t N,
                          float scale, float* d_conv_out, void* stream) {
  MDB_REQUIRE(h, h != nullptr, MDB_E_INVALID_ARG, "null handle");
  MDB_REQUIRE(h, d_rvec && d_tvec && d_conv_out, MDB_E_INVALID_ARG, "mdb_pose_head_bwd: null pointer");
  MDB_REQUIRE(h, w >= 1 && hh >= 1 && N >= 1, MDB_E_SHAPE, "mdb_pose_head_bwd: need w,h,N >= 1");
  const size_t total = (size_t)N * 6 * w * hh;
  const int blocks = (int)std::min<size_t>((total + 255) / 256, (size_t)h->num_sms * 8);
  pose_head_bwd_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(d_rvec, d_tvec, w * hh, N, scale, d_conv_out);
  MDB_LAUNCHED(h);
  return MDB_OK;
}

int32_t mdb_item_mean(mdb_handle* h, const float* map, int64_t per_item, int32_t N, float* out, void* stream) {
  MDB_REQUIRE(h, h != nullptr, MDB_E_INVALID_ARG, "null handle");
  MDB_REQUIRE(h, map && out, MDB_E_INVALID_ARG, "mdb_item_mean: null pointer");
  MDB_REQUIRE(h, per_item >= 1 && N >= 1, MDB_E_SHAPE, "mdb_item_mean: need per_item, N >= 1");
  item_mean_kernel<<<N * IM_CLUSTER, HEAD_THREADS, 0, (cudaStream_t)stream>>>(map, (long long)per_item, out);
  MDB_LAUNCHED(h);
  return MDB_OK;
}

}  // extern "C"

// ---- DepthDecoder output heads (src/depth_decoder.jl:46, 52-68) ----------------------------------------------------
// Every scale level ends in `DecoderBlock(Conv((3,3), ch => 1, σ))`: the disparity of a level is σ of a one-channel
// conv output.  The convolution is the network's (cuDNN, out of scope); the activation is the op on the boundary of
// the loss path: all levels in ONE launch, logits (w_k,h_k,1,N) -> disparity (w_k,h_k,1,N), and its pullback
// d logit = g σ (1 - σ) from the saved output, so the loss's d_disparity flows back into the network without an extra
// elementwise pass per level.  σ is NNlib's overflow-safe form: t = exp(-|x|); x >= 0 ? 1 / (1 + t) : t / (1 + t).
namespace {
struct HeadsArgs {
  const float* in[MDB_MAX_SCALES];      // fwd: logits;  bwd: saved disparities
  const float* g[MDB_MAX_SCALES];       // bwd: d loss / d disparity
  float* out[MDB_MAX_SCALES];
  long long start[MDB_MAX_SCALES + 1];  // prefix sums of the element counts (in units of 4 floats where vectorised)
  int Ls, vec;
};

MDB_DEV float sigmoid_nnlib(float x) {
  const float t = expf(-fabsf(x));
  return x >= 0.0f ? 1.0f / (1.0f + t) : t / (1.0f + t);
}

template <bool BWD>
__global__ void __launch_bounds__(256) disparity_heads_kernel(const __grid_constant__ HeadsArgs a) {
  const long long total = a.start[a.Ls];
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    int k = 0;
    while (k + 1 < a.Ls && i >= a.start[k + 1]) ++k;
    const long long j = i - a.start[k];
    if (a.vec) {
      const float4 v = __ldg(reinterpret_cast<const float4*>(a.in[k]) + j);
      float4 o;
      if (BWD) {
        const float4 g = __ldg(reinterpret_cast<const float4*>(a.g[k]) + j);
        o = make_float4(g.x * v.x * (1.0f - v.x), g.y * v.y * (1.0f - v.y), g.z * v.z * (1.0f - v.z), g.w * v.w * (1.0f - v.w));
      } else {
        o = make_float4(sigmoid_nnlib(v.x), sigmoid_nnlib(v.y), sigmoid_nnlib(v.z), sigmoid_nnlib(v.w));
      }
      reinterpret_cast<float4*>(a.out[k])[j] = o;
    } else {
      const float v = ldg(a.in[k] + j);
      a.out[k][j] = BWD ? ldg(a.g[k] + j) * v * (1.0f - v) : sigmoid_nnlib(v);
    }
  }
}

int32_t heads_launch(mdb_handle* h, bool bwd, const float* const* in, const float* const* g, float* const* out,
                     const int32_t* dw, const int32_t* dh, int32_t Ls, int32_t N, void* stream, const char* who) {
  MDB_REQUIRE(h, h != nullptr, MDB_E_INVALID_ARG, "null handle");
  MDB_REQUIRE(h, in && out && dw && dh && (!bwd || g) && Ls >= 1 && Ls <= MDB_MAX_SCALES && N >= 1, MDB_E_INVALID_ARG, "%s: bad argument", who);
  HeadsArgs a{};
  a.Ls = Ls;
  a.vec = 1;
  for (int k = 0; k < Ls; ++k) {
    MDB_REQUIRE(h, in[k] && out[k] && (!bwd || g[k]) && dw[k] >= 1 && dh[k] >= 1, MDB_E_INVALID_ARG, "%s: bad level %d", who, k);
    const long long n = (long long)dw[k] * dh[k] * N;
    if (n % 4 || (((uintptr_t)in[k] | (uintptr_t)out[k] | (uintptr_t)(bwd ? g[k] : in[k])) & 15)) a.vec = 0;
  }
  for (int k = 0; k < Ls; ++k) {
    const long long n = (long long)dw[k] * dh[k] * N;
    a.in[k] = in[k]; a.g[k] = bwd ? g[k] : nullptr; a.out[k] = out[k];
    a.start[k + 1] = a.start[k] + (a.vec ? n / 4 : n);
  }
  const long long blocks = std::max<long long>(1, std::min<long long>((a.start[Ls] + 255) / 256, (long long)h->num_sms * 16));
  if (bwd) disparity_heads_kernel<true><<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(a);
  else disparity_heads_kernel<false><<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(a);
  MDB_LAUNCHED(h);
  return MDB_OK;
}
}  // namespace

extern "C" {

int32_t mdb_disparity_heads(mdb_handle* h, const float* const* logits, const int32_t* dw, const int32_t* dh, int32_t Ls,
                            int32_t N, float* const* disparity, void* stream) {
  return heads_launch(h, false, logits, nullptr, disparity, dw, dh, Ls, N, stream, "mdb_disparity_heads");
}

int32_t mdb_disparity_heads_bwd(mdb_handle* h, const float* const* disparity, const float* const* d_disparity, const int32_t* dw,
                                const int32_t* dh, int32_t Ls, int32_t N, float* const* d_logits, void* stream) {
  return heads_launch(h, true, disparity, d_disparity, d_logits, dw, dh, Ls, N, stream, "mdb_disparity_heads_bwd");
}

}  // extern "C"
