// comm.cu -- the path's only exchange step: one sum over the ranks of the packed vector
//   [ loss | d_rvec[s][n] (3 floats) for s, n | d_tvec[s][n] (3 floats) for s, n ]     (1 + 6 S N_global floats)
// in which a rank fills only the slots of its own batch shard.  The reference has no multi-GPU code at all; the
// two lines that couple the items are the means over N (src/training.jl:69, src/utils.jl:176), which the kernels
// divide by n_global so that shard losses add.
//
// Two transports behind the same call:
//   * NCCL (mdb_comm_init): libnccl.so.2 is resolved at run time with dlopen -- the library has no link-time
//     dependency on it, and inside a PyTorch process the call binds to the copy torch already loaded.
//   * one-shot peer writes (mdb_comm_peers_export / _attach): every rank owns a small symmetric buffer that its peers
//     map through CUDA IPC; ONE kernel writes the rank's vector into its slot on every peer over NVLink, raises a
//     flag there, waits for the peers' flags and adds the slots in rank order (deterministic, bit-identical on
//     all ranks).  No library call, ~one NVLink round trip of latency for the 3 KB the step exchanges.
#include "mdb_common.cuh"

#include <dlfcn.h>

namespace {

// ---- the handful of NCCL declarations we bind (ABI of nccl.h 2.x) ---------------------------------------
struct NcclUniqueId { char internal[128]; };
typedef void* NcclComm;
typedef int (*fn_GetUniqueId)(NcclUniqueId*);
typedef int (*fn_CommInitRank)(NcclComm*, int, NcclUniqueId, int);
typedef int (*fn_CommDestroy)(NcclComm);
typedef int (*fn_AllReduce)(const void*, void*, size_t, int /*dtype*/, int /*op*/, NcclComm, cudaStream_t);
typedef const char* (*fn_GetErrorString)(int);
constexpr int kNcclFloat32 = 7, kNcclSum = 0;

struct NcclApi {
  void* so = nullptr;
  fn_GetUniqueId get_id = nullptr;
  fn_CommInitRank init_rank = nullptr;
  fn_CommDestroy destroy = nullptr;
  fn_AllReduce all_reduce = nullptr;
  fn_GetErrorString err_str = nullptr;
};

NcclApi* nccl_api() {
  static NcclApi api;
  static bool tried = false;
  if (tried) return api.so ? &api : nullptr;
  tried = true;
  // the copy that is already in the process (torch's) wins; else the system one
  void* so = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
  if (!so) so = dlopen("libnccl.so.2", RTLD_NOW);
  if (!so) so = dlopen("libnccl.so", RTLD_NOW);
  if (!so) return nullptr;
  api.get_id = (fn_GetUniqueId)dlsym(so, "ncclGetUniqueId");
  api.init_rank = (fn_CommInitRank)dlsym(so, "ncclCommInitRank");
  api.destroy = (fn_CommDestroy)dlsym(so, "ncclCommDestroy");
  api.all_reduce = (fn_AllReduce)dlsym(so, "ncclAllReduce");
  api.err_str = (fn_GetErrorString)dlsym(so, "ncclGetErrorString");
  if (!api.get_id || !api.init_rank || !api.destroy || !api.all_reduce) return nullptr;
  api.so = so;
  return &api;
}

// ---- one-shot peer-write all-reduce ------------------------------------------------------------------
// Symmetric buffer of every rank (allocated by mdb_comm_peers_export):
//   int   flag[2][MDB_COMM_MAX_RANKS]      flag[p][r] = sequence number of rank r's last write into parity p
//   float slot[2][world][cap]              slot[p][r] = rank r's vector of the call with parity p
// Call number q (1, 2, ...) uses parity q & 1.  A rank can only start call q + 2 after it has finished call
// q + 1, i.e. after every peer has WRITTEN its call q + 1 vector, which a peer does after it has finished
// reading call q (stream order) -- so a slot is never overwritten while a peer still reads it.
constexpr int kPeerThreads = 256;

struct PeerArgs {
  float* peer[MDB_COMM_MAX_RANKS];   // base of every rank's symmetric buffer as mapped HERE (own buffer at [rank])
  int rank, world, cap;
  long long timeout_cycles;
};

MDB_DEV int* peer_flags(float* base, int parity) { return reinterpret_cast<int*>(base) + parity * MDB_COMM_MAX_RANKS; }
MDB_DEV float* peer_slot(float* base, int parity, int r, int world, int cap) {
  return base + 2 * MDB_COMM_MAX_RANKS + ((size_t)parity * world + r) * cap;
}

__global__ void __launch_bounds__(kPeerThreads)
peer_allreduce_kernel(const __grid_constant__ PeerArgs a, const float* __restrict__ send, float* __restrict__ recv,
                      int count, int seq, int* __restrict__ status) {
  const int parity = seq & 1;
  // 1. this rank's vector into its slot on every rank (its own included): warp w serves peers w, w + 8, ...
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int p = warp; p < a.world; p += kPeerThreads / 32) {
    float* dst = peer_slot(a.peer[p], parity, a.rank, a.world, a.cap);
    for (int i = lane; i < count; i += 32) dst[i] = send[i];
  }
  __threadfence_system();
  __syncthreads();
  // 2. raise the flag on every rank
  if (threadIdx.x < a.world) {
    volatile int* f = peer_flags(a.peer[threadIdx.x], parity) + a.rank;
    *f = seq;
  }
  // 3. wait for every rank's flag in OUR buffer (bounded: a missing peer sets *status instead of hanging the GPU)
  __shared__ int ok;
  if (threadIdx.x == 0) ok = 1;
  __syncthreads();
  if (threadIdx.x < a.world) {
    volatile int* f = peer_flags(a.peer[a.rank], parity) + threadIdx.x;
    const long long t0 = clock64();
    while (*f != seq) {
      if (clock64() - t0 > a.timeout_cycles) { ok = 0; break; }
      __nanosleep(20);
    }
  }
  __syncthreads();
  __threadfence_system();
  if (!ok) {
    if (threadIdx.x == 0) *status = 1;
    return;
  }
  // 4. sum of the slots in rank order: the same order on every rank, so the results are bit-identical
  float* mine = a.peer[a.rank];
  for (int i = threadIdx.x; i < count; i += kPeerThreads) {
    float v = 0.0f;
    for (int r = 0; r < a.world; ++r) v += __ldcv(peer_slot(mine, parity, r, a.world, a.cap) + i);
    recv[i] = v;
  }
}

}  // namespace

struct mdb_comm {
  int rank = 0, world = 1;
  NcclComm nccl = nullptr;
  // one-shot transport
  float* own = nullptr;            // this rank's symmetric buffer (cudaMalloc'ed here: IPC exports whole allocations)
  size_t own_bytes = 0;
  void* mapped[MDB_COMM_MAX_RANKS] = {nullptr};   // peers' buffers opened through IPC (null at [rank])
  PeerArgs peer{};
  int seq = 0;
  int* status = nullptr;           // device int inside `own`'s tail: set by a timed-out wait
  bool peers_on = false;
};

extern "C" {

int64_t mdb_packed_size(int32_t n_global, int32_t S) { return 1 + 6ll * S * n_global; }

int32_t mdb_packed_bind(mdb_loss_io* io, float* packed, int32_t n_global, int32_t first_item) {
  if (!io || !packed || io->S < 1 || io->S > MDB_MAX_SOURCES || first_item < 0 || first_item + io->N > n_global)
    return MDB_E_INVALID_ARG;
  io->loss = packed;
  for (int s = 0; s < io->S; ++s) {
    io->d_rvec[s] = packed + 1 + 3ll * ((long long)s * n_global + first_item);
    io->d_tvec[s] = packed + 1 + 3ll * io->S * n_global + 3ll * ((long long)s * n_global + first_item);
  }
  return MDB_OK;
}

int32_t mdb_comm_get_unique_id(void* out, size_t bytes) {
  NcclApi* api = nccl_api();
  if (!api || !out || bytes < sizeof(NcclUniqueId)) return MDB_E_INVALID_ARG;
  NcclUniqueId id;
  const int rc = api->get_id(&id);
  if (rc != 0) return MDB_E_INTERNAL;
  memcpy(out, &id, sizeof(id));
  return MDB_OK;
}

static mdb_comm* comm_of(mdb_handle* h) {
  if (!h->comm) h->comm = new mdb_comm();
  return h->comm;
}

int32_t mdb_comm_init(mdb_handle* h, const void* unique_id, int32_t rank, int32_t world) {
  MDB_REQUIRE(h, h != nullptr, MDB_E_INVALID_ARG, "null handle");
  MDB_REQUIRE(h, unique_id && world >= 1 && rank >= 0 && rank < world, MDB_E_INVALID_ARG, "mdb_comm_init: bad rank / world / id");
  NcclApi* api = nccl_api();
  MDB_REQUIRE(h, api != nullptr, MDB_E_INTERNAL, "mdb_comm_init: libnccl.so.2 could not be loaded (%s)", dlerror());
  mdb_comm* c = comm_of(h);
  MDB_REQUIRE(h, c->nccl == nullptr, MDB_E_INVALID_ARG, "mdb_comm_init: communicator already initialised");
  MDB_CUDA(h, cudaSetDevice(h->device));
  NcclUniqueId id;
  memcpy(&id, unique_id, sizeof(id));
  const int rc = api->init_rank(&c->nccl, world, id, rank);
  if (rc != 0) {
    c->nccl = nullptr;
    return mdb_fail(h, MDB_E_INTERNAL, "ncclCommInitRank failed: %s", api->err_str ? api->err_str(rc) : "?");
  }
  c->rank = rank; c->world = world;
  return MDB_OK;
}

static size_t peer_buffer_bytes(int world, int64_t max_count) {
  return mdb_align_up(sizeof(float) * (2 * MDB_COMM_MAX_RANKS + 2ull * world * (size_t)max_count) + 256, 256);
}

// The symmetric buffer is the one allocation the library makes itself: CUDA IPC exports whole cudaMalloc
// allocations, so a sub-allocation of the caller's pool (torch, CUDA.jl) could not be mapped by the peers.
int32_t mdb_comm_peers_export(mdb_handle* h, int32_t world, int64_t max_count, void* ipc_handle_out, size_t bytes) {
  MDB_REQUIRE(h, h != nullptr, MDB_E_INVALID_ARG, "null handle");
  MDB_REQUIRE(h, ipc_handle_out && bytes >= sizeof(cudaIpcMemHandle_t) && world >= 1 && world <= MDB_COMM_MAX_RANKS && max_count >= 1,
              MDB_E_INVALID_ARG, "mdb_comm_peers_export: bad argument (world <= %d, handle needs %zu bytes)", MDB_COMM_MAX_RANKS,
              sizeof(cudaIpcMemHandle_t));
  mdb_comm* c = comm_of(h);
  MDB_REQUIRE(h, c->own == nullptr, MDB_E_INVALID_ARG, "mdb_comm_peers_export: already exported");
  MDB_CUDA(h, cudaSetDevice(h->device));
  const size_t need = peer_buffer_bytes(world, max_count);
  void* buf = nullptr;
  MDB_CUDA(h, cudaMalloc(&buf, need));
  cudaError_t e = cudaMemset(buf, 0, need);   // flags and status start at zero, before any peer can map the buffer
  cudaIpcMemHandle_t hd;
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&hd, buf);
  if (e != cudaSuccess) {
    cudaFree(buf);
    return mdb_fail(h, (int32_t)e, "mdb_comm_peers_export: %s", cudaGetErrorString(e));
  }
  memcpy(ipc_handle_out, &hd, sizeof(hd));
  c->own = (float*)buf;
  c->own_bytes = need;
  c->status = (int*)((char*)buf + need - 256);
  c->peer.world = world; c->peer.cap = (int)max_count;
  return MDB_OK;
}

int32_t mdb_comm_peers_attach(mdb_handle* h, const void* ipc_handles, int32_t rank) {
  MDB_REQUIRE(h, h != nullptr, MDB_E_INVALID_ARG, "null handle");
  mdb_comm* c = h->comm;
  MDB_REQUIRE(h, c && c->own && !c->peers_on, MDB_E_INVALID_ARG, "mdb_comm_peers_attach: call mdb_comm_peers_export first (once)");
  const int world = c->peer.world;
  MDB_REQUIRE(h, ipc_handles && rank >= 0 && rank < world, MDB_E_INVALID_ARG, "mdb_comm_peers_attach: bad rank / handles");
  MDB_CUDA(h, cudaSetDevice(h->device));
  const cudaIpcMemHandle_t* hs = (const cudaIpcMemHandle_t*)ipc_handles;
  for (int r = 0; r < world; ++r) {
    if (r == rank) { c->peer.peer[r] = c->own; continue; }
    void* p = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&p, hs[r], cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      (void)cudaGetLastError();
      for (int q = 0; q < r; ++q)
        if (c->mapped[q]) { cudaIpcCloseMemHandle(c->mapped[q]); c->mapped[q] = nullptr; }
      return mdb_fail(h, (int32_t)e, "cudaIpcOpenMemHandle(rank %d) failed: %s", r, cudaGetErrorString(e));
    }
    c->mapped[r] = p;
    c->peer.peer[r] = (float*)p;
  }
  c->peer.rank = rank;
  c->peer.timeout_cycles = 20ll * 1000 * 1000 * 1000;   // ~10 s at 1.9 GHz: a peer that never shows up
  c->rank = rank; c->world = world;
  c->seq = 0;
  c->peers_on = true;
  return MDB_OK;
}

int32_t mdb_allreduce_loss_posegrad(mdb_handle* h, const float* send, float* recv, int64_t count, void* stream) {
  MDB_REQUIRE(h, h != nullptr, MDB_E_INVALID_ARG, "null handle");
  MDB_REQUIRE(h, send && recv && count >= 1, MDB_E_INVALID_ARG, "mdb_allreduce_loss_posegrad: bad argument");
  mdb_comm* c = h->comm;
  MDB_REQUIRE(h, c && (c->peers_on || c->nccl), MDB_E_INVALID_ARG,
              "mdb_allreduce_loss_posegrad: call mdb_comm_init or mdb_comm_peers_attach first");
  cudaStream_t st = (cudaStream_t)stream;
  if (c->peers_on) {
    MDB_REQUIRE(h, count <= c->peer.cap, MDB_E_INVALID_ARG, "mdb_allreduce_loss_posegrad: %lld floats exceed the peer buffer's %d",
                (long long)count, c->peer.cap);
    c->seq += 1;
    peer_allreduce_kernel<<<1, kPeerThreads, 0, st>>>(c->peer, send, recv, (int)count, c->seq, c->status);
    MDB_LAUNCHED(h);
    return MDB_OK;
  }
  NcclApi* api = nccl_api();
  const int rc = api->all_reduce(send, recv, (size_t)count, kNcclFloat32, kNcclSum, c->nccl, st);
  if (rc != 0) return mdb_fail(h, MDB_E_INTERNAL, "ncclAllReduce failed: %s", api->err_str ? api->err_str(rc) : "?");
  return MDB_OK;
}

int32_t mdb_comm_status(mdb_handle* h, int32_t* timed_out) {
  MDB_REQUIRE(h, h && timed_out, MDB_E_INVALID_ARG, "mdb_comm_status: bad argument");
  *timed_out = 0;
  mdb_comm* c = h->comm;
  if (c && c->peers_on) MDB_CUDA(h, cudaMemcpy(timed_out, c->status, sizeof(int), cudaMemcpyDeviceToHost));
  return MDB_OK;
}

int32_t mdb_comm_destroy(mdb_handle* h) {
  if (!h || !h->comm) return MDB_OK;
  mdb_comm* c = h->comm;
  if (c->nccl) {
    if (NcclApi* api = nccl_api()) api->destroy(c->nccl);
  }
  for (int r = 0; r < MDB_COMM_MAX_RANKS; ++r)
    if (c->mapped[r]) cudaIpcCloseMemHandle(c->mapped[r]);
  if (c->own) cudaFree(c->own);
  delete c;
  h->comm = nullptr;
  return MDB_OK;
}

}  // extern "C"
