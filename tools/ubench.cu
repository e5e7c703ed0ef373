// ubench.cu -- instruction-throughput microbenchmarks for sm_100a (B200): FFMA vs FFMA2 (packed
// f32x2), FADD2, DFMA, MUFU.RCP, SHFL, LDS.  Prints warp-instructions per clock per SM.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench ubench.cu
#include <cuda_runtime.h>
#include <cstdio>
constexpr int ITER = 4096, CH = 8;
template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, float seed) {
  __shared__ float sm[1024];
  const int t = threadIdx.x;
  sm[t] = seed * t; sm[t + 256] = t; sm[t + 512] = -t; sm[t + 768] = 1.0f;
  __syncthreads();
  float a[CH], b = seed, c = 1.0f - seed;
  float2 a2[CH]; float2 b2 = make_float2(seed, seed * 0.5f), c2 = make_float2(c, c * 0.5f);
  double d[CH]; double bd = seed, cd = 1.0 - seed;
  for (int i = 0; i < CH; ++i) { a[i] = seed * i + t; a2[i] = make_float2(a[i], -a[i]); d[i] = a[i]; }
  int idx = t;
  for (int it = 0; it < ITER; ++it) {
#pragma unroll
    for (int i = 0; i < CH; ++i) {
      if (MODE == 0) a[i] = fmaf(a[i], b, c);
      if (MODE == 1) a2[i] = __ffma2_rn(a2[i], b2, c2);
      if (MODE == 2) a2[i] = __fadd2_rn(a2[i], b2);
      if (MODE == 3) d[i] = fma(d[i], bd, cd);
      if (MODE == 4) asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
      if (MODE == 5) a[i] = __shfl_down_sync(0xffffffffu, a[i], 1);
      if (MODE == 6) { a[i] += sm[(idx + i * 32) & 1023]; }
      if (MODE == 7) a[i] = a[i] + b;
      if (MODE == 8) { a[i] = fmaf(a[i], b, c); a2[i] = __ffma2_rn(a2[i], b2, c2); }   // mix
      if (MODE == 9) { a[i] = fmaf(a[i], b, c); d[i] = fma(d[i], bd, cd); }          // mix
    }
    if (MODE == 6) idx = (idx + 1) & 1023;
  }
  float r = 0;
  for (int i = 0; i < CH; ++i) r += a[i] + a2[i].x + a2[i].y + (float)d[i];
  out[blockIdx.x * blockDim.x + t] = r;
}
template <int MODE>
void run(const char* name, float per_iter) {
  int dev; cudaGetDevice(&dev); cudaDeviceProp p; cudaGetDeviceProperties(&p, dev);
  const int sms = p.multiProcessorCount, blocks = sms * 8;
  float* out; cudaMalloc(&out, blocks * 256 * 4);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MODE><<<blocks, 256>>>(out, 0.5f); cudaDeviceSynchronize();
  cudaEventRecord(e0);
  k<MODE><<<blocks, 256>>>(out, 0.5f);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  int khz; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
  const double winstr = (double)blocks * 8 * ITER * CH * per_iter;
  const double clk = ms * 1e-3 * khz * 1e3;
  printf("%-28s %8.3f ms  %6.3f warp-instr/clk/SM (at %d MHz nominal)\n", name, ms, winstr / clk / sms, khz / 1000);
  cudaFree(out);
}
int main() {
  run<0>("FFMA", 1); run<1>("FFMA2 (f32x2)", 1); run<2>("FADD2", 1); run<7>("FADD", 1); run<3>("DFMA", 1);
  run<4>("MUFU.RCP", 1); run<5>("SHFL", 1); run<6>("LDS+FADD", 2); run<8>("FFMA+FFMA2", 2); run<9>("FFMA+DFMA", 2);
  return 0;
}
