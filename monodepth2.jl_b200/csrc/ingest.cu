// ingest.cu -- SURVEY 8(f) rank 3: the data format on the input side of the path.
// Depth10k stores a sample as ONE decoded image holding the three frames side by side
// (src/dtk.jl:29-46): load -> [optional Gray{Float32}] -> split into 3 column blocks of width W ->
// [optional FlipX] -> channelview -> cat(dims=4) -> permutedims((3,2,1,4)) -> Float32.  Doing that on the
// host and shipping Float32 costs 12 bytes per pixel over PCIe; this kernel takes the decoder's raster
// (uint8, row-major [H][3W][RGB]) and writes the (W,H,C,3,N) Float32 stack the loss reads: 3 bytes per
// pixel over PCIe, one HBM-bound pass on the device (1 B read + 4 B written per output float).
#include "mdb_common.cuh"

namespace {

constexpr int ING_THREADS = 256;   // == the size of the N0f8 table

// Float32(::N0f8) = u / 255 in float32, correctly rounded (FixedPointNumbers, third-party)
MDB_DEV float n0f8(unsigned char u) { return __fdiv_rn((float)u, 255.0f); }

// One block per (strip row h, item n): stage the 9W source bytes in shared memory with 128-bit streaming
// loads, then write the 3 (frames) x C plane rows coalesced along w.  Float32(::N0f8) is an IEEE division;
// it is done 256 times per block into a table instead of once per output, and the output loop has no
// integer division (plane index outside, w inside).
__global__ void __launch_bounds__(ING_THREADS)
ingest_triplet_u8_kernel(const unsigned char* __restrict__ strips, const unsigned char* __restrict__ flip,
                         int W, int H, int gray, float* __restrict__ x) {
  extern __shared__ __align__(16) unsigned char row[];
  __shared__ float lut[256];
  const int h = blockIdx.x, n = blockIdx.y;
  const int nbytes = 9 * W;
  const unsigned char* src = strips + ((size_t)n * H + h) * nbytes;
  lut[threadIdx.x] = n0f8((unsigned char)threadIdx.x);        // ING_THREADS == 256
  if ((((uintptr_t)src) & 15) == 0 && (nbytes & 15) == 0) {
    const uint4* s4 = reinterpret_cast<const uint4*>(src);
    uint4* r4 = reinterpret_cast<uint4*>(row);
    for (int i = threadIdx.x; i < nbytes / 16; i += ING_THREADS) r4[i] = __ldcs(s4 + i);
  } else {
    for (int i = threadIdx.x; i < nbytes; i += ING_THREADS) row[i] = src[i];
  }
  __syncthreads();
  const bool f = flip && flip[n];                 // FlipX: reverse the width axis of every frame alike
  const int C = gray ? 1 : 3;
  const size_t plane = (size_t)H * W;
  float* out = x + (size_t)n * 3 * C * plane + (size_t)h * W;
  for (int l = 0; l < 3; ++l) {
    const unsigned char* fr = row + (size_t)l * W * 3;
    if (gray) {
      // Gray{Float32}(::RGB{N0f8}) = 0.299f0 r + 0.587f0 g + 0.114f0 b, products and sums rounded one by one
      // (Colors.jl, third-party; Julia does not contract to FMA)                    src/dtk.jl:33-35
      float* o = out + (size_t)l * plane;
      for (int w = threadIdx.x; w < W; w += ING_THREADS) {
        const unsigned char* p = fr + (f ? W - 1 - w : w) * 3;
        const float v = __fadd_rn(__fadd_rn(__fmul_rn(0.299f, lut[p[0]]), __fmul_rn(0.587f, lut[p[1]])),
                                  __fmul_rn(0.114f, lut[p[2]]));
        __stcs(o + w, v);
      }
    } else {
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        float* o = out + (size_t)(l * 3 + c) * plane;
        for (int w = threadIdx.x; w < W; w += ING_THREADS) __stcs(o + w, lut[fr[(f ? W - 1 - w : w) * 3 + c]]);
      }
    }
  }
}


// ---- KittyDataset.getindex, src/kitty.jl:45-61 ---------------------------------------------------------------------
// Three consecutive `image_0` PNGs (Gray{N0f8}, e.g. 1241 x 376) -> imresize to (height, width) -> Float32 -> (W,H,1,3).
// `imresize` is ImageTransformations.jl (third-party): per axis p = s i + (0.5 - 0.5 s) clamped to [1, m], s = m / n,
// linear B-spline, rows outermost, Float64 arithmetic, result rounded back to N0f8 (ties to even) -- restated in
// oracle/ingest_ref.py, which this kernel matches bit for bit: every Float64 operation below is an explicit
// round-to-nearest intrinsic so that nvcc contracts nothing into an FMA numpy would not do.
// One thread per output pixel: 4 byte taps + ~20 FP64 operations; (m / n)^2 bytes read + 4 written per output, HBM-bound.
struct AxisTap { int i0, i1; double w0, w1; };

MDB_DEV AxisTap axis_tap(int i /*0-based output index*/, int n_new, int m_old) {
  const double s = __ddiv_rn((double)m_old, (double)n_new);
  const double off = __dsub_rn(0.5, __dmul_rn(0.5, s));
  double p = __dadd_rn(__dmul_rn(s, (double)(i + 1)), off);
  p = fmin(fmax(p, 1.0), (double)m_old);
  double f = floor(p);
  if (f > (double)(m_old - 1)) f = __dsub_rn(f, 1.0);
  const double d = __dsub_rn(p, f);
  AxisTap t;
  t.i0 = (int)f - 1;
  t.i1 = min(t.i0 + 1, m_old - 1);
  t.w0 = __dsub_rn(1.0, d);
  t.w1 = d;
  return t;
}

__global__ void __launch_bounds__(ING_THREADS)
ingest_kitti_u8_kernel(const unsigned char* __restrict__ frames, const unsigned char* __restrict__ flip, int Wo, int Ho,
                       int W, int H, float* __restrict__ x) {
  __shared__ double lut64[256];   // Float64(::N0f8) = u / 255
  __shared__ float lut32[256];    // Float32(::N0f8)
  lut64[threadIdx.x] = __ddiv_rn((double)threadIdx.x, 255.0);
  lut32[threadIdx.x] = n0f8((unsigned char)threadIdx.x);
  __syncthreads();
  const int w = blockIdx.x * ING_THREADS + threadIdx.x, h = blockIdx.y, fn = blockIdx.z;   // fn = 3 n + frame
  if (w >= W) return;
  const unsigned char* src = frames + (size_t)fn * Ho * Wo;
  unsigned char q;
  if (Wo == W && Ho == H) {
    q = src[(size_t)h * Wo + w];                                    // imresize copies when the size already matches
  } else {
    const AxisTap ty = axis_tap(h, H, Ho), tx = axis_tap(w, W, Wo);
    const unsigned char* r0 = src + (size_t)ty.i0 * Wo;
    const unsigned char* r1 = src + (size_t)ty.i1 * Wo;
    const double a00 = lut64[r0[tx.i0]], a01 = lut64[r0[tx.i1]], a10 = lut64[r1[tx.i0]], a11 = lut64[r1[tx.i1]];
    const double top = __dadd_rn(__dmul_rn(tx.w0, a00), __dmul_rn(tx.w1, a01));
    const double bot = __dadd_rn(__dmul_rn(tx.w0, a10), __dmul_rn(tx.w1, a11));
    const double v = __dadd_rn(__dmul_rn(ty.w0, top), __dmul_rn(ty.w1, bot));
    q = (unsigned char)(int)rint(__dmul_rn(255.0, v));             // N0f8(v): round(255 v), ties to even
  }
  const bool f = flip && flip[fn / 3];
  __stcs(x + ((size_t)fn * H + h) * W + (f ? W - 1 - w : w), lut32[q]);
}


// The transpose alone (mdb_loss_io.x_format = 1: the fused kernel converts on the fly): same staging of the strip row,
// bytes out.  One block per (strip row, item); 4 output bytes per thread and store where W % 4 == 0.
__global__ void __launch_bounds__(ING_THREADS)
ingest_triplet_u8_planar_kernel(const unsigned char* __restrict__ strips, const unsigned char* __restrict__ flip, int W, int H,
                                unsigned char* __restrict__ x) {
  extern __shared__ __align__(16) unsigned char row[];
  const int h = blockIdx.x, n = blockIdx.y;
  const int nbytes = 9 * W;
  const unsigned char* src = strips + ((size_t)n * H + h) * nbytes;
  if ((((uintptr_t)src) & 15) == 0 && (nbytes & 15) == 0) {
    const uint4* s4 = reinterpret_cast<const uint4*>(src);
    uint4* r4 = reinterpret_cast<uint4*>(row);
    for (int i = threadIdx.x; i < nbytes / 16; i += ING_THREADS) r4[i] = __ldcs(s4 + i);
  } else {
    for (int i = threadIdx.x; i < nbytes; i += ING_THREADS) row[i] = src[i];
  }
  __syncthreads();
  const bool f = flip && flip[n];
  const size_t plane = (size_t)H * W;
  unsigned char* out = x + (size_t)n * 9 * plane + (size_t)h * W;
  const bool vec = (W & 3) == 0 && (((uintptr_t)x) & 3) == 0;
  for (int lc = 0; lc < 9; ++lc) {
    const int l = lc / 3, c = lc % 3;
    const unsigned char* fr = row + (size_t)l * W * 3 + c;
    unsigned char* o = out + (size_t)lc * plane;
    if (vec) {
      for (int w4 = threadIdx.x; w4 < W / 4; w4 += ING_THREADS) {
        unsigned v = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int w = 4 * w4 + k;
          v |= (unsigned)fr[(f ? W - 1 - w : w) * 3] << (8 * k);
        }
        reinterpret_cast<unsigned*>(o)[w4] = v;
      }
    } else {
      for (int w = threadIdx.x; w < W; w += ING_THREADS) o[w] = fr[(f ? W - 1 - w : w) * 3];
    }
  }
}

}  // namespace

extern "C" int32_t mdb_ingest_triplet_u8(mdb_handle* h, const uint8_t* strips, const uint8_t* flip, int32_t W,
                                         int32_t H, int32_t N, int32_t grayscale, float* x, void* stream) {
  MDB_REQUIRE(h, h != nullptr, MDB_E_INVALID_ARG, "null handle");
  MDB_REQUIRE(h, strips && x, MDB_E_INVALID_ARG, "mdb_ingest_triplet_u8: null pointer");
  MDB_REQUIRE(h, W >= 1 && H >= 1 && N >= 1 && H <= 0x7fffffff && N <= 65535, MDB_E_SHAPE,
              "mdb_ingest_triplet_u8: need W,H,N >= 1 and N <= 65535");
  const size_t smem = mdb_align_up((size_t)9 * W, 16);
  MDB_REQUIRE(h, smem <= 200 * 1024, MDB_E_SHAPE, "mdb_ingest_triplet_u8: strip row of %d px does not fit shared memory", 3 * W);
  if (smem > 48 * 1024)
    MDB_CUDA(h, cudaFuncSetAttribute(ingest_triplet_u8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid(H, N);
  ingest_triplet_u8_kernel<<<grid, ING_THREADS, smem, (cudaStream_t)stream>>>(strips, flip, W, H, grayscale ? 1 : 0, x);
  MDB_LAUNCHED(h);
  return MDB_OK;
}

extern "C" int32_t mdb_ingest_kitti_u8(mdb_handle* h, const uint8_t* frames, const uint8_t* flip, int32_t W_orig,
                                       int32_t H_orig, int32_t W, int32_t H, int32_t N, float* x, void* stream) {
  MDB_REQUIRE(h, h != nullptr, MDB_E_INVALID_ARG, "null handle");
  MDB_REQUIRE(h, frames && x, MDB_E_INVALID_ARG, "mdb_ingest_kitti_u8: null pointer");
  MDB_REQUIRE(h, W >= 1 && H >= 1 && N >= 1 && W_orig >= 1 && H_orig >= 1 && H <= 65535 && N <= 21845, MDB_E_SHAPE,
              "mdb_ingest_kitti_u8: need W,H,N,W_orig,H_orig >= 1, H <= 65535 and N <= 21845");
  dim3 grid((W + ING_THREADS - 1) / ING_THREADS, H, 3 * N);
  ingest_kitti_u8_kernel<<<grid, ING_THREADS, 0, (cudaStream_t)stream>>>(frames, flip, W_orig, H_orig, W, H, x);
  MDB_LAUNCHED(h);
  return MDB_OK;
}

extern "C" int32_t mdb_ingest_triplet_u8_planar(mdb_handle* h, const uint8_t* strips, const uint8_t* flip, int32_t W,
                                                int32_t H, int32_t N, uint8_t* x_u8, void* stream) {
  MDB_REQUIRE(h, h != nullptr, MDB_E_INVALID_ARG, "null handle");
  MDB_REQUIRE(h, strips && x_u8, MDB_E_INVALID_ARG, "mdb_ingest_triplet_u8_planar: null pointer");
  MDB_REQUIRE(h, W >= 1 && H >= 1 && N >= 1 && N <= 65535, MDB_E_SHAPE, "mdb_ingest_triplet_u8_planar: need W,H,N >= 1 and N <= 65535");
  const size_t smem = mdb_align_up((size_t)9 * W, 16);
  MDB_REQUIRE(h, smem <= 200 * 1024, MDB_E_SHAPE, "mdb_ingest_triplet_u8_planar: strip row of %d px does not fit shared memory", 3 * W);
  if (smem > 48 * 1024)
    MDB_CUDA(h, cudaFuncSetAttribute(ingest_triplet_u8_planar_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid(H, N);
  ingest_triplet_u8_planar_kernel<<<grid, ING_THREADS, smem, (cudaStream_t)stream>>>(strips, flip, W, H, x_u8);
  MDB_LAUNCHED(h);
  return MDB_OK;
}
