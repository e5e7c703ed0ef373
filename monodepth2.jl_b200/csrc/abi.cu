// abi.cu -- handle life-cycle and error reporting of the C ABI (include/mdb200.h).
#include "mdb_common.cuh"

extern "C" {

int32_t mdb_version(void) { return MDB_VERSION; }

int32_t mdb_create(mdb_handle** out, int32_t device) {
  if (!out) return MDB_E_INVALID_ARG;
  *out = nullptr;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count <= 0 || device < 0 || device >= count) {
    (void)cudaGetLastError();
    return MDB_E_NO_DEVICE;  // no CPU fallback exists
  }
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return MDB_E_NO_DEVICE;
  if (prop.major != 10) return MDB_E_NO_DEVICE;  // the library only contains sm_100a code
  mdb_handle* h = new mdb_handle();
  h->device = device;
  h->num_sms = prop.multiProcessorCount;
  h->cc_major = prop.major;
  h->cc_minor = prop.minor;
  h->smem_optin = prop.sharedMemPerBlockOptin;
  *out = h;
  return MDB_OK;
}

static void prof_free(mdb_handle* h) {
  if (!h->prof_ev) return;
  for (int i = 0; i < h->prof_slots * (mdb_handle::kProfSegments + 1); ++i) cudaEventDestroy(h->prof_ev[i]);
  delete[] h->prof_ev;
  h->prof_ev = nullptr;
  h->prof_slots = h->prof_next = 0;
}

int32_t mdb_destroy(mdb_handle* h) {
  if (h) {
    prof_free(h);
    mdb_comm_destroy(h);
  }
  delete h;
  return MDB_OK;
}

int32_t mdb_profile_begin(mdb_handle* h, int32_t slots) {
  MDB_REQUIRE(h, h != nullptr, MDB_E_INVALID_ARG, "null handle");
  prof_free(h);
  if (slots <= 0) return MDB_OK;  // profiling off
  const int n = slots * (mdb_handle::kProfSegments + 1);
  h->prof_ev = new cudaEvent_t[n];
  for (int i = 0; i < n; ++i) MDB_CUDA(h, cudaEventCreate(&h->prof_ev[i]));
  h->prof_slots = slots;
  h->prof_next = 0;
  return MDB_OK;
}

int32_t mdb_profile_read(mdb_handle* h, int32_t slot, float* ms_out) {
  MDB_REQUIRE(h, h && ms_out, MDB_E_INVALID_ARG, "mdb_profile_read: bad argument");
  MDB_REQUIRE(h, h->prof_ev && slot >= 0 && slot < h->prof_slots && slot < h->prof_next, MDB_E_INVALID_ARG,
              "mdb_profile_read: slot %d was not recorded", slot);
  cudaEvent_t* e = h->prof_ev + slot * (mdb_handle::kProfSegments + 1);
  MDB_CUDA(h, cudaEventSynchronize(e[mdb_handle::kProfSegments]));
  for (int s = 0; s < mdb_handle::kProfSegments; ++s) MDB_CUDA(h, cudaEventElapsedTime(&ms_out[s], e[s], e[s + 1]));
  return MDB_OK;
}

const char* mdb_last_error(const mdb_handle* h) { return h ? h->err : "null handle"; }

int64_t mdb_launch_count(const mdb_handle* h) { return h ? h->launches : 0; }

void mdb_loss_params_default(mdb_loss_params* p) {
  if (!p) return;
  p->min_depth = 0.1;
  p->max_depth = 100.0;
  p->disparity_smoothness = 1e-3;
  p->alpha = 0.85f;
  p->ssim_c1 = 0.01f * 0.01f;
  p->ssim_c2 = 0.03f * 0.03f;
  p->automasking = 1;
  p->normalize_disparity = 1;
  p->smooth_weight_mode = 0;
  p->upsample_align_corners = 1;
  p->n_global = 0;
  p->abs_subgradient = 0;
}

}  // extern "C"
