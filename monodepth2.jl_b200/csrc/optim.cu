// optim.cu -- SURVEY 8(f) rank 1: the caller of the hot path in config 1.  `slow_depth`
// (src/simple_depth.jl:1-62) runs 500 iterations of { Zygote.gradient of the single-scale loss;
// Flux.Optimise.update!(ADAM(3e-4), theta, grads) } with a host round trip per op.  Here one iteration
// is a CUDA graph, replayed back to back: pose prep -> strip kernel -> { finalize kernel, which also applies the
// ADAM step of the poses where it forms their gradients (mdb_pose_adam)  ||  the ADAM kernel of the disparity
// map, whose gradient is complete after the strip kernel }.  Step counters live on the device.
#include "mdb_common.cuh"

namespace {

struct AdamArgs {
  float* p[MDB_ADAM_MAX_TENSORS];
  const float* g[MDB_ADAM_MAX_TENSORS];
  float* m[MDB_ADAM_MAX_TENSORS];
  float* v[MDB_ADAM_MAX_TENSORS];
  long long end[MDB_ADAM_MAX_TENSORS];   // exclusive prefix ends of the flattened index space
  int n;
  double eta, b1, b2, eps;
};

constexpr int ADAM_THREADS = 256;

// Flux ADAM (apply! + update!): the state (m, v) and the parameters are float32 like the reference's
// arrays; the bias corrections 1 - beta^t are formed in double (Flux keeps beta^t as Float64).
__global__ void __launch_bounds__(ADAM_THREADS)
adam_kernel(const __grid_constant__ AdamArgs a, int* __restrict__ step, int* __restrict__ ticket,
            const float* __restrict__ loss, float* __restrict__ hist, int hist_len) {
  const int t = *step + 1;
  __shared__ float sc[2];
  if (threadIdx.x < 2) sc[threadIdx.x] = (float)(1.0 / (1.0 - pow(threadIdx.x ? a.b2 : a.b1, (double)t)));
  __syncthreads();
  const float c1 = sc[0], c2 = sc[1];
  const float b1 = (float)a.b1, b2 = (float)a.b2, eta = (float)a.eta, eps = (float)a.eps;
  const long long total = a.end[a.n - 1];
  for (long long i = (long long)blockIdx.x * ADAM_THREADS + threadIdx.x; i < total;
       i += (long long)gridDim.x * ADAM_THREADS) {
    int k = 0;
    while (i >= a.end[k]) ++k;
    const long long j = i - (k ? a.end[k - 1] : 0);
    const float g = a.g[k][j];
    const float m = b1 * a.m[k][j] + (1.0f - b1) * g;
    const float v = b2 * a.v[k][j] + (1.0f - b2) * g * g;
    a.m[k][j] = m;
    a.v[k][j] = v;
    a.p[k][j] -= m * c1 / (sqrtf(v * c2) + eps) * eta;
  }
  // last block out advances the counter: every block has read *step by then
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(ticket, 1) == (int)gridDim.x - 1) {
      if (hist && t - 1 < hist_len) hist[t - 1] = *loss;
      *ticket = 0;
      *step = t;
      __threadfence();
    }
  }
}

int32_t launch_adam(mdb_handle* h, const AdamArgs& a, int* step, int* ticket, const float* loss, float* hist,
                    int hist_len, cudaStream_t st) {
  const long long total = a.end[a.n - 1];
  const int blocks = (int)std::min<long long>((total + ADAM_THREADS - 1) / ADAM_THREADS, 4 * 148);
  adam_kernel<<<std::max(blocks, 1), ADAM_THREADS, 0, st>>>(a, step, ticket, loss, hist, hist_len);
  MDB_LAUNCHED(h);
  return MDB_OK;
}

}  // namespace

struct mdb_slow_depth_plan {
  mdb_handle* h = nullptr;
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t exec = nullptr;
  cudaStream_t capture_stream = nullptr, side_stream = nullptr;
  cudaEvent_t fork_ev = nullptr, join_ev = nullptr;
  int kernels_per_iter = 0;
};

extern "C" {

int32_t mdb_adam_update(mdb_handle* h, int32_t n_tensors, float* const* params, const float* const* grads,
                        float* const* m, float* const* v, const int64_t* sizes, const mdb_adam_params* ap,
                        int32_t* step, int32_t* ticket, const float* loss, float* loss_history, int32_t history_len,
                        void* stream) {
  MDB_REQUIRE(h, h != nullptr, MDB_E_INVALID_ARG, "null handle");
  MDB_REQUIRE(h, n_tensors >= 1 && n_tensors <= MDB_ADAM_MAX_TENSORS, MDB_E_INVALID_ARG,
              "mdb_adam_update: 1..%d tensors", MDB_ADAM_MAX_TENSORS);
  MDB_REQUIRE(h, params && grads && m && v && sizes && ap && step && ticket, MDB_E_INVALID_ARG,
              "mdb_adam_update: null argument");
  MDB_REQUIRE(h, !loss_history || loss, MDB_E_INVALID_ARG, "mdb_adam_update: loss_history needs loss");
  AdamArgs a{};
  a.n = n_tensors;
  long long end = 0;
  for (int k = 0; k < n_tensors; ++k) {
    MDB_REQUIRE(h, params[k] && grads[k] && m[k] && v[k] && sizes[k] > 0, MDB_E_INVALID_ARG,
                "mdb_adam_update: null pointer / empty tensor %d", k);
    a.p[k] = params[k]; a.g[k] = grads[k]; a.m[k] = m[k]; a.v[k] = v[k];
    end += sizes[k];
    a.end[k] = end;
  }
  a.eta = ap->eta; a.b1 = ap->beta1; a.b2 = ap->beta2; a.eps = ap->eps;
  return launch_adam(h, a, step, ticket, loss, loss_history, history_len, (cudaStream_t)stream);
}

size_t mdb_slow_depth_state_bytes(const mdb_loss_io* io) {
  if (!io || io->W < 2 || io->H < 2 || io->N < 1 || io->S < 1 || io->S > MDB_MAX_SOURCES) return 0;
  const size_t n = (size_t)io->W * io->H * io->N + (size_t)6 * io->S * io->N;   // disp + rvec/tvec per source
  return mdb_align_up(2 * n * sizeof(float), 256) + 256;                          // (m, v) + counters
}

int32_t mdb_slow_depth_create(mdb_handle* h, const mdb_loss_io* io, const mdb_loss_params* params,
                              const mdb_adam_params* adam, void* state, size_t state_bytes, void* workspace,
                              size_t workspace_bytes, float* loss_history, int32_t history_len,
                              mdb_slow_depth_plan** out) {
  MDB_REQUIRE(h, h != nullptr, MDB_E_INVALID_ARG, "null handle");
  MDB_REQUIRE(h, io && params && adam && out, MDB_E_INVALID_ARG, "mdb_slow_depth_create: null argument");
  MDB_REQUIRE(h, io->Ls == 1 && io->dw[0] == io->W && io->dh[0] == io->H, MDB_E_SHAPE,
              "mdb_slow_depth_create: needs one full-resolution scale (src/simple_depth.jl:8)");
  const size_t need = mdb_slow_depth_state_bytes(io);
  MDB_REQUIRE(h, need && state && state_bytes >= need && ((uintptr_t)state & 255) == 0, MDB_E_WORKSPACE,
              "mdb_slow_depth_create: state must be >= %zu bytes and 256-byte aligned", need);
  // the graph forks the disparity map's ADAM kernel off right after the fused kernel: that is only safe when nothing
  // later in the loss call still writes d_disparity -- i.e. no mean-normalisation constant (added after finalize)
  MDB_REQUIRE(h, params->normalize_disparity == 0 && params->smooth_weight_mode == 1 && !params->automasking, MDB_E_INVALID_ARG,
              "mdb_slow_depth_create: needs normalize_disparity = 0, smooth_weight_mode = 1, automasking = 0 "
              "(the closure of src/simple_depth.jl:25-41)");
  MDB_REQUIRE(h, !h->prof_ev, MDB_E_INVALID_ARG, "mdb_slow_depth_create: switch segment profiling off first");
  const int S = io->S, N = io->N;
  const size_t nd = (size_t)io->W * io->H * N;
  // tensors: disp, then (rvec, tvec) per source
  float* mv = (float*)state;
  AdamArgs a{};
  long long end = 0;
  size_t off = 0;
  auto add = [&](const float* p, float* g, size_t n) {
    a.p[a.n] = const_cast<float*>(p); a.g[a.n] = g;
    a.m[a.n] = mv + off; a.v[a.n] = mv + off + n;
    off += 2 * n;
    end += (long long)n; a.end[a.n] = end;
    ++a.n;
  };
  MDB_REQUIRE(h, io->disparity[0] && io->d_disparity[0], MDB_E_INVALID_ARG, "null disparity pointer");
  add(io->disparity[0], io->d_disparity[0], nd);
  // the poses keep their (m, v) slots in the state, in this order, but their ADAM step runs inside the loss
  // call's finalize kernel: the graph's ADAM kernel is left with the disparity map alone
  mdb_pose_adam pad{};
  for (int s = 0; s < S; ++s) {
    MDB_REQUIRE(h, io->rvec[s] && io->tvec[s] && io->d_rvec[s] && io->d_tvec[s], MDB_E_INVALID_ARG, "null pose pointer");
    const size_t n3 = (size_t)3 * N;
    pad.p_r[s] = const_cast<float*>(io->rvec[s]); pad.m_r[s] = mv + off; pad.v_r[s] = mv + off + n3; off += 2 * n3;
    pad.p_t[s] = const_cast<float*>(io->tvec[s]); pad.m_t[s] = mv + off; pad.v_t[s] = mv + off + n3; off += 2 * n3;
  }
  a.eta = adam->eta; a.b1 = adam->beta1; a.b2 = adam->beta2; a.eps = adam->eps;
  // counters (int32): [0] step of the ADAM kernel (the documented one), [1] its ticket, [2] step of the finalize
  // kernel's pose ADAM (kept equal to [0]); floats [4..7]: bias corrections handed from one iteration to the next
  int* counters = (int*)((char*)state + mdb_align_up(off * sizeof(float), 256));
  pad.step = counters + 2;
  pad.cbuf = (float*)(counters + 4);
  pad.hist = loss_history; pad.hist_len = history_len;
  pad.eta = adam->eta; pad.b1 = adam->beta1; pad.b2 = adam->beta2; pad.eps = adam->eps;

  mdb_slow_depth_plan* pl = new mdb_slow_depth_plan();
  pl->h = h;
  auto fail = [&](int32_t rc) { mdb_slow_depth_destroy(pl); return rc; };
  cudaError_t e = cudaStreamCreateWithFlags(&pl->capture_stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&pl->side_stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&pl->fork_ev, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&pl->join_ev, cudaEventDisableTiming);
  if (e != cudaSuccess) return fail(mdb_fail(h, (int32_t)e, "stream / event creation: %s", cudaGetErrorString(e)));
  e = cudaStreamBeginCapture(pl->capture_stream, cudaStreamCaptureModeThreadLocal);
  if (e != cudaSuccess) return fail(mdb_fail(h, (int32_t)e, "cudaStreamBeginCapture: %s", cudaGetErrorString(e)));
  const int64_t before = h->launches;
  // the loss call records fork_ev right after the strip kernel: the disparity's ADAM kernel branches off there
  // and joins after the finalize kernel (which steps the poses and writes the loss history itself)
  h->pose_adam = &pad;
  h->fork_event = pl->fork_ev;
  int32_t rc = mdb_train_loss_fwdbwd(h, io, params, workspace, workspace_bytes, pl->capture_stream);
  h->pose_adam = nullptr;
  h->fork_event = nullptr;
  if (rc == MDB_OK) {
    cudaError_t e2 = cudaStreamWaitEvent(pl->side_stream, pl->fork_ev, 0);
    if (e2 == cudaSuccess) rc = launch_adam(h, a, counters, counters + 1, nullptr, nullptr, 0, pl->side_stream);
    if (e2 == cudaSuccess) e2 = cudaEventRecord(pl->join_ev, pl->side_stream);
    if (e2 == cudaSuccess) e2 = cudaStreamWaitEvent(pl->capture_stream, pl->join_ev, 0);
    if (e2 != cudaSuccess && rc == MDB_OK) rc = mdb_fail(h, (int32_t)e2, "graph fork / join: %s", cudaGetErrorString(e2));
  }
  e = cudaStreamEndCapture(pl->capture_stream, &pl->graph);
  pl->kernels_per_iter = (int)(h->launches - before);
  h->launches = before;                       // nothing ran yet: launches are counted per replay
  if (rc != MDB_OK) return fail(rc);
  if (e != cudaSuccess) return fail(mdb_fail(h, (int32_t)e, "cudaStreamEndCapture: %s", cudaGetErrorString(e)));
  e = cudaGraphInstantiate(&pl->exec, pl->graph, 0);
  if (e != cudaSuccess) return fail(mdb_fail(h, (int32_t)e, "cudaGraphInstantiate: %s", cudaGetErrorString(e)));
  *out = pl;
  return MDB_OK;
}

int32_t mdb_slow_depth_run(mdb_slow_depth_plan* pl, int32_t iters, void* stream) {
  if (!pl || !pl->exec) return MDB_E_INVALID_ARG;
  mdb_handle* h = pl->h;
  MDB_REQUIRE(h, iters >= 0, MDB_E_INVALID_ARG, "mdb_slow_depth_run: negative iteration count");
  for (int i = 0; i < iters; ++i) {
    MDB_CUDA(h, cudaGraphLaunch(pl->exec, (cudaStream_t)stream));
    h->launches += pl->kernels_per_iter;
  }
  return MDB_OK;
}

int32_t mdb_slow_depth_destroy(mdb_slow_depth_plan* pl) {
  if (!pl) return MDB_OK;
  if (pl->exec) cudaGraphExecDestroy(pl->exec);
  if (pl->graph) cudaGraphDestroy(pl->graph);
  if (pl->capture_stream) cudaStreamDestroy(pl->capture_stream);
  if (pl->side_stream) cudaStreamDestroy(pl->side_stream);
  if (pl->fork_ev) cudaEventDestroy(pl->fork_ev);
  if (pl->join_ev) cudaEventDestroy(pl->join_ev);
  delete pl;
  return MDB_OK;
}

}  // extern "C"
