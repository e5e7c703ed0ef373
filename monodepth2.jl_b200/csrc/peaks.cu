// peaks.cu -- the FP32 roofline denominator, measured on the device the bench runs on.
// MEASURED_PEAKS.json (driver-written) holds HBM and bf16 tensor peaks only; the fused loss is bound by FP32
// issue (SURVEY F10, 8d "Honest roofline"), so bench.py needs P_fp32: a kernel of independent FFMA chains
// (8 per thread, 8 CTAs of 256 threads per SM), timed with CUDA events on the caller's stream.
// No reference counterpart (the reference's only timing is `CUDA.@time` in scratch files, src/test.jl:30).
#include "mdb_common.cuh"

namespace {
constexpr int PK_ITER = 4096, PK_CH = 8, PK_THREADS = 256, PK_CTAS_PER_SM = 8;

__global__ void __launch_bounds__(PK_THREADS) ffma_peak_kernel(float* out, float seed) {
  float a[PK_CH];
  const float b = seed, c = 1.0f - seed;
#pragma unroll
  for (int i = 0; i < PK_CH; ++i) a[i] = seed * i + threadIdx.x;
  for (int it = 0; it < PK_ITER; ++it) {
#pragma unroll
    for (int i = 0; i < PK_CH; ++i) a[i] = fmaf(a[i], b, c);
  }
  float r = 0.0f;
#pragma unroll
  for (int i = 0; i < PK_CH; ++i) r += a[i];
  out[blockIdx.x * PK_THREADS + threadIdx.x] = r;
}
}  // namespace

extern "C" int32_t mdb_measure_fp32_peak(mdb_handle* h, void* scratch, size_t scratch_bytes, int32_t repeats,
                                         double* tflops_out, double* ms_out, void* stream) {
  MDB_REQUIRE(h, h != nullptr, MDB_E_INVALID_ARG, "null handle");
  MDB_REQUIRE(h, tflops_out && repeats >= 1, MDB_E_INVALID_ARG, "mdb_measure_fp32_peak: bad argument");
  const int blocks = h->num_sms * PK_CTAS_PER_SM;
  const size_t need = (size_t)blocks * PK_THREADS * sizeof(float);
  MDB_REQUIRE(h, scratch && scratch_bytes >= need, MDB_E_WORKSPACE, "mdb_measure_fp32_peak: scratch must hold %zu bytes", need);
  cudaStream_t st = (cudaStream_t)stream;
  cudaEvent_t e0, e1;
  MDB_CUDA(h, cudaEventCreate(&e0));
  MDB_CUDA(h, cudaEventCreate(&e1));
  float best = 1e30f;
  for (int r = 0; r <= repeats; ++r) {   // run 0 warms up
    cudaEventRecord(e0, st);
    ffma_peak_kernel<<<blocks, PK_THREADS, 0, st>>>((float*)scratch, 0.5f);
    cudaEventRecord(e1, st);
    h->launches++;
    cudaError_t e = cudaEventSynchronize(e1);
    if (e != cudaSuccess) {
      cudaEventDestroy(e0); cudaEventDestroy(e1);
      return mdb_fail(h, (int32_t)e, "mdb_measure_fp32_peak: %s", cudaGetErrorString(e));
    }
    float ms = 0.0f;
    cudaEventElapsedTime(&ms, e0, e1);
    if (r > 0 && ms < best) best = ms;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  const double flops = 2.0 * (double)blocks * PK_THREADS * PK_ITER * PK_CH;
  *tflops_out = flops / (best * 1e-3) / 1e12;
  if (ms_out) *ms_out = best;
  return MDB_OK;
}
