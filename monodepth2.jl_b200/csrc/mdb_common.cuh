// mdb_common.cuh -- shared host/device helpers for the mdb200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <algorithm>

#include "../../include/mdb200.h"

// slow_depth only (optim.cu).  When the handle carries one, finalize_item_kernel applies Flux's ADAM step to the
// poses right where it forms their gradients and records the loss history, so the graph's ADAM kernel is left
// with the disparity map and can run BESIDE the finalize kernel (forked after the strip kernel).
struct mdb_pose_adam {
  float* p_r[MDB_MAX_SOURCES]; float* p_t[MDB_MAX_SOURCES];   // parameters (3,N) each
  float* m_r[MDB_MAX_SOURCES]; float* v_r[MDB_MAX_SOURCES];
  float* m_t[MDB_MAX_SOURCES]; float* v_t[MDB_MAX_SOURCES];
  int* step;        // this kernel's own step counter (the ADAM kernel keeps the one the API documents)
  float* cbuf;      // [2][2] bias corrections 1 / (1 - beta^t), slot t & 1; 0 = not formed yet
  float* hist; int hist_len;
  double eta, b1, b2, eps;
};

struct mdb_handle {
  int device = 0;
  int num_sms = 0;
  int cc_major = 0, cc_minor = 0;
  size_t smem_optin = 0;
  int64_t launches = 0;
  struct mdb_comm* comm = nullptr;   // comm.cu: NCCL communicator and / or the peer-mapped buffers of the one-shot reduce
  // optional per-call segment timing (mdb_profile_*): events recorded on the caller's stream
  static constexpr int kProfSegments = 4;  // 0 prep+stats, 1 fused kernel, 2 finalize, 3 upsample_bwd
  int prof_slots = 0, prof_next = 0;
  cudaEvent_t* prof_ev = nullptr;  // [slots][kProfSegments + 1]
  const mdb_pose_adam* pose_adam = nullptr;   // set by mdb_slow_depth_create around its capture of the loss call
  cudaEvent_t fork_event = nullptr;           // ditto: recorded right after the strip kernel (graph fork point)
  char err[512] = {0};
};

inline int32_t mdb_fail(mdb_handle* h, int32_t code, const char* fmt, ...) {
  if (h) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(h->err, sizeof(h->err), fmt, ap);
    va_end(ap);
  }
  return code;
}

#define MDB_REQUIRE(h, cond, code, ...)                   \
  do {                                                    \
    if (!(cond)) return mdb_fail((h), (code), __VA_ARGS__); \
  } while (0)

#define MDB_CUDA(h, expr)                                                            \
  do {                                                                               \
    cudaError_t _e = (expr);                                                         \
    if (_e != cudaSuccess)                                                           \
      return mdb_fail((h), (int32_t)_e, "%s failed: %s (%s:%d)", #expr,              \
                      cudaGetErrorString(_e), __FILE__, __LINE__);                   \
  } while (0)

// launch check: peek (do not clear sticky errors of other work) + count
#define MDB_LAUNCHED(h)                      \
  do {                                       \
    (h)->launches++;                         \
    MDB_CUDA((h), cudaPeekAtLastError());    \
  } while (0)

// record the boundary event `seg` of the current profiling slot (no-op when profiling is off)
inline void mdb_prof_mark(mdb_handle* h, int slot, int seg, cudaStream_t st) {
  if (h->prof_ev && slot >= 0) cudaEventRecord(h->prof_ev[slot * (mdb_handle::kProfSegments + 1) + seg], st);
}

static inline size_t mdb_align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

#ifdef __CUDACC__
#define MDB_DEV __device__ __forceinline__
#define MDB_HD __host__ __device__ __forceinline__

constexpr unsigned MDB_FULL = 0xffffffffu;

MDB_DEV float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(MDB_FULL, v, o);
  return v;
}

// reflect-pad index (no edge repeat) for a 1-based coordinate r in [0, n+1], n >= 2:
// NNlib pad_reflect(x, 1): xp[0] = x[2], xp[n+1] = x[n-1]  (src/utils.jl:30-31)
MDB_DEV int reflect1(int r, int n) { return r < 1 ? 2 - r : (r > n ? 2 * n - r : r); }

MDB_DEV float ldg(const float* p) { return __ldg(p); }

// streaming store (write-once outputs that are not re-read by this kernel)
MDB_DEV void st_cs(float* p, float v) { __stcs(p, v); }
#endif

// test hook (mdb_debug_option "render_precise"): the render forward samples every plane through the double-precision
// homography instead of the float displacement form (defined in render.cu)
extern int mdb_dbg_render_precise;
