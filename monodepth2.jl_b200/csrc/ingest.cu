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
