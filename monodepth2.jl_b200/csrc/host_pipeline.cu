// host_pipeline.cu -- train_loss value + gradients for inputs that live in HOST memory (bench "e2e").
// The reference ships a batch to the device with `transfer = gpu . f32` (src/utils.jl:1-3; callers
// src/simple_depth.jl:81-92, scripts/script.jl:65-90) and then calls train_loss; this entry is that pair behind the
// C ABI for a host that is not Python: the batch is cut into chunks of items, chunk k+1 is copied host->device on a
// copy stream while chunk k runs mdb_train_loss_fwdbwd on the caller's stream and chunk k-1's disparity gradients
// drain device->host on a third stream, so a step costs about max(PCIe, compute) instead of their sum.  Every chunk
// is evaluated with params.n_global = the whole batch (the multi-GPU contract), so chunk losses add.
// Optional last step (first_item >= 0): the path's one all-reduce (mdb_allreduce_loss_posegrad) of the packed
// [loss, pose gradients], after which every rank holds the GLOBAL loss and pose gradients.
//
// The plan owns its device buffers, streams and events (cudaMalloc at create, freed at destroy): one of the two places
// where the library allocates (the other is the peer-mapped buffer of comm.cu).
#include "mdb_common.cuh"

#include <vector>

namespace {

__global__ void chunk_loss_sum_kernel(const float* __restrict__ part, int n, float* __restrict__ out) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    double v = 0.0;   // fixed order: deterministic
    for (int i = 0; i < n; ++i) v += (double)part[i];
    *out = (float)v;
  }
}

struct Chunk { int a, b; cudaEvent_t ready, done; };

}  // namespace

struct mdb_host_plan {
  mdb_handle* h = nullptr;
  mdb_loss_io shape{};          // sizes / ids / K of the whole local batch (pointers unused)
  mdb_loss_params params{};
  int x_format = 0;             // 0: float32 (W,H,C,L,N); 1: uint8 triplet strips [N][H][3W][3] (mdb_ingest_triplet_u8)
  int first_item = -1, n_global = 0;
  std::vector<Chunk> chunks;
  cudaStream_t copy_in = nullptr, copy_out = nullptr;
  cudaEvent_t ev_start = nullptr, ev_small = nullptr;
  // device buffers, whole local batch
  float* d_x = nullptr; unsigned char* d_u8 = nullptr; unsigned char* d_flip = nullptr;
  unsigned char* d_xu8 = nullptr;   // x_format 1 with C = 3: the planar uint8 stack the fused kernel reads directly (no Float32 stack)
  float* d_disp[MDB_MAX_SCALES] = {nullptr}; float* g_disp[MDB_MAX_SCALES] = {nullptr};
  float* d_rv[MDB_MAX_SOURCES] = {nullptr}; float* d_tv[MDB_MAX_SOURCES] = {nullptr};
  float* g_rv[MDB_MAX_SOURCES] = {nullptr}; float* g_tv[MDB_MAX_SOURCES] = {nullptr};
  float* d_auto = nullptr;
  float* g_loss = nullptr;      // one partial loss per chunk, then [chunks] = their sum
  float* packed_send = nullptr; float* packed_recv = nullptr;   // first_item >= 0
  void* ws = nullptr; size_t ws_bytes = 0;
  float* h_small = nullptr;     // pinned staging of the small results (loss / packed vector)
  size_t h2d_bytes = 0, d2h_bytes = 0;
  std::vector<void*> owned;     // everything cudaMalloc'ed here
};

namespace {

template <typename T>
cudaError_t plan_alloc(mdb_host_plan* p, T** out, size_t count) {
  void* q = nullptr;
  cudaError_t e = cudaMalloc(&q, std::max<size_t>(count * sizeof(T), 256));
  if (e == cudaSuccess) { p->owned.push_back(q); *out = (T*)q; }
  return e;
}

void plan_free(mdb_host_plan* p) {
  if (!p) return;
  for (void* q : p->owned) cudaFree(q);
  for (Chunk& c : p->chunks) { if (c.ready) cudaEventDestroy(c.ready); if (c.done) cudaEventDestroy(c.done); }
  if (p->ev_start) cudaEventDestroy(p->ev_start);
  if (p->ev_small) cudaEventDestroy(p->ev_small);
  if (p->copy_in) cudaStreamDestroy(p->copy_in);
  if (p->copy_out) cudaStreamDestroy(p->copy_out);
  if (p->h_small) cudaFreeHost(p->h_small);
  delete p;
}

}  // namespace

extern "C" {

int32_t mdb_host_plan_create(mdb_handle* h, const mdb_loss_io* shape, const mdb_loss_params* params, int32_t chunk_items,
                             int32_t x_format, int32_t first_item, mdb_host_plan** out) {
  MDB_REQUIRE(h, h != nullptr, MDB_E_INVALID_ARG, "null handle");
  MDB_REQUIRE(h, shape && params && out, MDB_E_INVALID_ARG, "mdb_host_plan_create: null argument");
  MDB_REQUIRE(h, x_format == 0 || (x_format == 1 && shape->L == 3), MDB_E_INVALID_ARG,
              "mdb_host_plan_create: x_format 0 (float32 stack) or 1 (uint8 triplet strips, L = 3)");
  *out = nullptr;
  {
    mdb_loss_io probe = *shape;
    MDB_REQUIRE(h, mdb_train_loss_workspace_bytes(&probe) != 0, MDB_E_SHAPE, "mdb_host_plan_create: unsupported shape");
  }
  const int N = shape->N, S = shape->S, Ls = shape->Ls, W = shape->W, H = shape->H, C = shape->C, L = shape->L;
  const int ng = params->n_global > 0 ? params->n_global : N;
  MDB_REQUIRE(h, first_item < 0 || first_item + N <= ng, MDB_E_INVALID_ARG, "mdb_host_plan_create: first_item + N exceeds n_global");
  MDB_CUDA(h, cudaSetDevice(h->device));
  mdb_host_plan* p = new mdb_host_plan();
  p->h = h; p->shape = *shape; p->params = *params; p->params.n_global = ng;
  p->x_format = x_format; p->first_item = first_item; p->n_global = ng;
  // ---- pipeline shape: chunks of `chunk_items`, then a geometric tail (what is not hidden is the work on the LAST
  // chunk after its copy lands); compute-bound pipelines (uint8 input) also ramp up from one item
  size_t per_item = (x_format == 0 ? (size_t)4 * L * C * H * W : (size_t)9 * H * W) + (params->automasking ? (size_t)4 * H * W : 0);
  for (int k = 0; k < Ls; ++k) per_item += (size_t)4 * shape->dw[k] * shape->dh[k];
  int chunk = chunk_items;
  if (chunk <= 0) chunk = std::max(std::max(1, N / 8), (int)std::min<size_t>((size_t)N, ((size_t)16 << 20) / per_item + 1));
  std::vector<int> sizes;
  int left = N;
  if (x_format == 1)
    for (int z = 1; z < chunk && left - z > chunk; z *= 2) { sizes.push_back(z); left -= z; }
  while (left > chunk) { sizes.push_back(chunk); left -= chunk; }
  while (left > 1) { const int half = (left + 1) / 2; sizes.push_back(half); left -= half; }
  if (left == 1) sizes.push_back(1);
  int maxc = 1, s0 = 0;
  for (int z : sizes) { p->chunks.push_back(Chunk{s0, s0 + z, nullptr, nullptr}); s0 += z; maxc = std::max(maxc, z); }
  cudaError_t e = cudaSuccess;
  auto ok = [&](cudaError_t r) { if (e == cudaSuccess) e = r; return r == cudaSuccess; };
  ok(cudaStreamCreateWithFlags(&p->copy_in, cudaStreamNonBlocking));
  ok(cudaStreamCreateWithFlags(&p->copy_out, cudaStreamNonBlocking));
  ok(cudaEventCreateWithFlags(&p->ev_start, cudaEventDisableTiming));
  ok(cudaEventCreateWithFlags(&p->ev_small, cudaEventDisableTiming));
  for (Chunk& c : p->chunks) { ok(cudaEventCreateWithFlags(&c.ready, cudaEventDisableTiming)); ok(cudaEventCreateWithFlags(&c.done, cudaEventDisableTiming)); }
  const size_t frame = (size_t)W * H * C * L;
  const bool u8_direct = x_format == 1 && C == 3;   // Gray luma (C = 1) is not an 8-bit value: that case keeps the Float32 stack
  if (u8_direct) ok(plan_alloc(p, &p->d_xu8, frame * N));
  else ok(plan_alloc(p, &p->d_x, frame * N));
  if (x_format == 1) { ok(plan_alloc(p, &p->d_u8, (size_t)9 * W * H * N)); ok(plan_alloc(p, &p->d_flip, (size_t)N)); }
  for (int k = 0; k < Ls; ++k) {
    ok(plan_alloc(p, &p->d_disp[k], (size_t)shape->dw[k] * shape->dh[k] * N));
    ok(plan_alloc(p, &p->g_disp[k], (size_t)shape->dw[k] * shape->dh[k] * N));
  }
  for (int s = 0; s < S; ++s) { ok(plan_alloc(p, &p->d_rv[s], (size_t)3 * N)); ok(plan_alloc(p, &p->d_tv[s], (size_t)3 * N)); }
  if (params->automasking) ok(plan_alloc(p, &p->d_auto, (size_t)W * H * N));
  ok(plan_alloc(p, &p->g_loss, p->chunks.size() + 1));
  size_t small = p->chunks.size() + 1;
  if (first_item >= 0) {
    const size_t cnt = (size_t)mdb_packed_size(ng, S);
    ok(plan_alloc(p, &p->packed_send, cnt));
    ok(plan_alloc(p, &p->packed_recv, cnt));
    if (e == cudaSuccess) ok(cudaMemset(p->packed_send, 0, cnt * sizeof(float)));   // foreign slots stay zero for good
    small = std::max(small, cnt);
    for (int s = 0; s < S; ++s) {
      p->g_rv[s] = p->packed_send + 1 + 3 * ((size_t)s * ng + first_item);
      p->g_tv[s] = p->packed_send + 1 + 3 * (size_t)S * ng + 3 * ((size_t)s * ng + first_item);
    }
  } else {
    for (int s = 0; s < S; ++s) { ok(plan_alloc(p, &p->g_rv[s], (size_t)3 * N)); ok(plan_alloc(p, &p->g_tv[s], (size_t)3 * N)); }
    small += (size_t)6 * S * N;
  }
  {
    mdb_loss_io big = *shape;
    big.N = maxc;
    p->ws_bytes = mdb_train_loss_workspace_bytes(&big);
    char* w = nullptr;
    ok(plan_alloc(p, &w, p->ws_bytes));
    p->ws = w;
  }
  ok(cudaMallocHost((void**)&p->h_small, small * sizeof(float)));
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    plan_free(p);
    return mdb_fail(h, (int32_t)e, "mdb_host_plan_create: %s", cudaGetErrorString(e));
  }
  size_t grads = 0;
  for (int k = 0; k < Ls; ++k) grads += (size_t)shape->dw[k] * shape->dh[k] * N;
  p->h2d_bytes = per_item * N + (size_t)4 * 6 * S * N + (x_format == 1 ? (size_t)N : 0);
  p->d2h_bytes = 4 * (grads + (first_item >= 0 ? (size_t)mdb_packed_size(ng, S) : p->chunks.size() + (size_t)6 * S * N));
  *out = p;
  return MDB_OK;
}

int32_t mdb_train_loss_host(mdb_host_plan* p, const mdb_host_io* io, void* stream) {
  if (!p) return MDB_E_INVALID_ARG;
  mdb_handle* h = p->h;
  MDB_REQUIRE(h, io && io->x && io->loss, MDB_E_INVALID_ARG, "mdb_train_loss_host: null io / x / loss");
  const mdb_loss_io& sh = p->shape;
  const int N = sh.N, S = sh.S, Ls = sh.Ls, W = sh.W, H = sh.H, C = sh.C, L = sh.L;
  for (int k = 0; k < Ls; ++k) MDB_REQUIRE(h, io->disparity[k] && io->d_disparity[k], MDB_E_INVALID_ARG, "mdb_train_loss_host: null disparity %d", k);
  for (int s = 0; s < S; ++s) MDB_REQUIRE(h, io->rvec[s] && io->tvec[s] && io->d_rvec[s] && io->d_tvec[s], MDB_E_INVALID_ARG, "mdb_train_loss_host: null pose %d", s);
  MDB_REQUIRE(h, !p->params.automasking || io->auto_loss, MDB_E_INVALID_ARG, "mdb_train_loss_host: automasking needs auto_loss");
  cudaStream_t st = (cudaStream_t)stream;
  // the copy streams start after whatever the caller has queued on its stream
  MDB_CUDA(h, cudaEventRecord(p->ev_start, st));
  MDB_CUDA(h, cudaStreamWaitEvent(p->copy_in, p->ev_start, 0));
  MDB_CUDA(h, cudaStreamWaitEvent(p->copy_out, p->ev_start, 0));
  for (int s = 0; s < S; ++s) {   // 6 floats per item and source: the whole batch at once
    MDB_CUDA(h, cudaMemcpyAsync(p->d_rv[s], io->rvec[s], sizeof(float) * 3 * N, cudaMemcpyHostToDevice, p->copy_in));
    MDB_CUDA(h, cudaMemcpyAsync(p->d_tv[s], io->tvec[s], sizeof(float) * 3 * N, cudaMemcpyHostToDevice, p->copy_in));
  }
  if (p->x_format == 1 && io->flip) MDB_CUDA(h, cudaMemcpyAsync(p->d_flip, io->flip, (size_t)N, cudaMemcpyHostToDevice, p->copy_in));
  const size_t frame = (size_t)W * H * C * L, strip = (size_t)9 * W * H, wh = (size_t)W * H;
  int k = 0;
  for (Chunk& c : p->chunks) {
    const int a = c.a, n = c.b - c.a;
    if (p->x_format == 0)
      MDB_CUDA(h, cudaMemcpyAsync(p->d_x + frame * a, (const float*)io->x + frame * a, sizeof(float) * frame * n, cudaMemcpyHostToDevice, p->copy_in));
    else
      MDB_CUDA(h, cudaMemcpyAsync(p->d_u8 + strip * a, (const unsigned char*)io->x + strip * a, strip * n, cudaMemcpyHostToDevice, p->copy_in));
    for (int j = 0; j < Ls; ++j) {
      const size_t per = (size_t)sh.dw[j] * sh.dh[j];
      MDB_CUDA(h, cudaMemcpyAsync(p->d_disp[j] + per * a, io->disparity[j] + per * a, sizeof(float) * per * n, cudaMemcpyHostToDevice, p->copy_in));
    }
    if (p->params.automasking)
      MDB_CUDA(h, cudaMemcpyAsync(p->d_auto + wh * a, io->auto_loss + wh * a, sizeof(float) * wh * n, cudaMemcpyHostToDevice, p->copy_in));
    MDB_CUDA(h, cudaEventRecord(c.ready, p->copy_in));
    MDB_CUDA(h, cudaStreamWaitEvent(st, c.ready, 0));
    if (p->x_format == 1) {
      const int32_t rc = p->d_xu8
          ? mdb_ingest_triplet_u8_planar(h, p->d_u8 + strip * a, io->flip ? p->d_flip + a : nullptr, W, H, n, p->d_xu8 + frame * a, st)
          : mdb_ingest_triplet_u8(h, p->d_u8 + strip * a, io->flip ? p->d_flip + a : nullptr, W, H, n, C == 1, p->d_x + frame * a, st);
      if (rc != MDB_OK) return rc;
    }
    mdb_loss_io ci = sh;
    ci.N = n;
    ci.x = p->d_xu8 ? reinterpret_cast<const float*>(p->d_xu8 + frame * a) : p->d_x + frame * a;
    ci.x_format = p->d_xu8 ? 1 : 0;
    for (int j = 0; j < Ls; ++j) {
      const size_t per = (size_t)sh.dw[j] * sh.dh[j];
      ci.disparity[j] = p->d_disp[j] + per * a;
      ci.d_disparity[j] = p->g_disp[j] + per * a;
    }
    for (int s = 0; s < S; ++s) {
      ci.rvec[s] = p->d_rv[s] + 3 * a; ci.tvec[s] = p->d_tv[s] + 3 * a;
      ci.d_rvec[s] = p->g_rv[s] + 3 * a; ci.d_tvec[s] = p->g_tv[s] + 3 * a;
    }
    ci.auto_loss = p->params.automasking ? p->d_auto + wh * a : nullptr;
    ci.loss = p->g_loss + k;
    ci.vis_warp_loss = nullptr;
    const int32_t rc = mdb_train_loss_fwdbwd(h, &ci, &p->params, p->ws, p->ws_bytes, st);
    if (rc != MDB_OK) return rc;
    MDB_CUDA(h, cudaEventRecord(c.done, st));
    MDB_CUDA(h, cudaStreamWaitEvent(p->copy_out, c.done, 0));
    for (int j = 0; j < Ls; ++j) {
      const size_t per = (size_t)sh.dw[j] * sh.dh[j];
      MDB_CUDA(h, cudaMemcpyAsync(io->d_disparity[j] + per * a, p->g_disp[j] + per * a, sizeof(float) * per * n, cudaMemcpyDeviceToHost, p->copy_out));
    }
    ++k;
  }
  // ---- the small results: chunk losses summed on the device (fixed order), then either the all-reduce of the packed
  // vector or the local pose gradients, through one pinned staging buffer
  const int nk = (int)p->chunks.size();
  float* total = p->first_item >= 0 ? p->packed_send : p->g_loss + nk;
  chunk_loss_sum_kernel<<<1, 32, 0, st>>>(p->g_loss, nk, total);
  MDB_LAUNCHED(h);
  size_t cnt;
  if (p->first_item >= 0) {
    cnt = (size_t)mdb_packed_size(p->n_global, S);
    const int32_t rc = mdb_allreduce_loss_posegrad(h, p->packed_send, p->packed_recv, (int64_t)cnt, st);
    if (rc != MDB_OK) return rc;
    MDB_CUDA(h, cudaEventRecord(p->ev_small, st));
    MDB_CUDA(h, cudaStreamWaitEvent(p->copy_out, p->ev_small, 0));
    MDB_CUDA(h, cudaMemcpyAsync(p->h_small, p->packed_recv, sizeof(float) * cnt, cudaMemcpyDeviceToHost, p->copy_out));
  } else {
    MDB_CUDA(h, cudaEventRecord(p->ev_small, st));
    MDB_CUDA(h, cudaStreamWaitEvent(p->copy_out, p->ev_small, 0));
    MDB_CUDA(h, cudaMemcpyAsync(p->h_small, total, sizeof(float), cudaMemcpyDeviceToHost, p->copy_out));
    for (int s = 0; s < S; ++s) {
      MDB_CUDA(h, cudaMemcpyAsync(p->h_small + 1 + 3 * (size_t)N * (2 * s), p->g_rv[s], sizeof(float) * 3 * N, cudaMemcpyDeviceToHost, p->copy_out));
      MDB_CUDA(h, cudaMemcpyAsync(p->h_small + 1 + 3 * (size_t)N * (2 * s + 1), p->g_tv[s], sizeof(float) * 3 * N, cudaMemcpyDeviceToHost, p->copy_out));
    }
  }
  MDB_CUDA(h, cudaStreamSynchronize(p->copy_out));   // the call returns once the results are in host memory
  // the caller's stream continues after the drain as well (its next call may overwrite the device buffers)
  MDB_CUDA(h, cudaEventRecord(p->ev_small, p->copy_out));
  MDB_CUDA(h, cudaStreamWaitEvent(st, p->ev_small, 0));
  *io->loss = p->h_small[0];
  if (p->first_item >= 0) {
    const size_t ng = (size_t)p->n_global;
    for (int s = 0; s < S; ++s) {
      memcpy(io->d_rvec[s], p->h_small + 1 + 3 * s * ng, sizeof(float) * 3 * ng);
      memcpy(io->d_tvec[s], p->h_small + 1 + 3 * S * ng + 3 * s * ng, sizeof(float) * 3 * ng);
    }
  } else {
    for (int s = 0; s < S; ++s) {
      memcpy(io->d_rvec[s], p->h_small + 1 + 3 * (size_t)N * (2 * s), sizeof(float) * 3 * N);
      memcpy(io->d_tvec[s], p->h_small + 1 + 3 * (size_t)N * (2 * s + 1), sizeof(float) * 3 * N);
    }
  }
  return MDB_OK;
}

int32_t mdb_host_plan_info(const mdb_host_plan* p, int64_t* out4) {
  if (!p || !out4) return MDB_E_INVALID_ARG;
  out4[0] = (int64_t)p->chunks.size(); out4[1] = (int64_t)p->h2d_bytes; out4[2] = (int64_t)p->d2h_bytes;
  out4[3] = p->first_item >= 0 ? p->n_global : p->shape.N;   // items the returned d_rvec / d_tvec cover
  return MDB_OK;
}

int32_t mdb_host_plan_destroy(mdb_host_plan* p) {
  plan_free(p);
  return MDB_OK;
}

}  // extern "C"
