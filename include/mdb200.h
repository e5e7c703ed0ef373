/* mdb200.h -- C ABI of the B200-native view-synthesis loss path for Monodepth2.jl.
 *
 * This is the only public native surface.  The reference (jumerckx/Monodepth2.jl) has no
 * FFI: the path is ordinary Julia functions differentiated by Zygote, and the extension
 * mechanism it already uses is `ChainRulesCore.rrule` (src/Monodepth.jl:28-29,
 * src/utils.jl:134-145).  Each entry point below replaces the body of one reference
 * function (cited per entry as file:line under /root/reference); the Julia glue that binds
 * them with `ccall` + `rrule` is shown in INTEGRATION.md and shipped (unexecuted: no Julia
 * on the build box) under monodepth2.jl_b200/julia/.
 *
 * Conventions
 *  - All array arguments are DEVICE pointers to dense float32 data in Julia column-major
 *    order (first Julia dimension fastest): images (W,H,C,N), frame stacks (W,H,C,L,N),
 *    disparities (w,h,1,N), rvec (3,N), tvec (3,1,N), rotation stacks (3,3,N).
 *    `K` / `invK` are HOST pointers to 9 floats, column-major 3x3 (as Julia stores them).
 *  - Pixel coordinates are 1-based inside the math, as in the reference.
 *  - Every call only ENQUEUES work on `stream` (a cudaStream_t passed as void*); nothing
 *    synchronises the device.  The caller owns every buffer, including the workspace.
 *  - Return value: 0 = OK; negative = MDB_E_* (invalid argument / shape / state);
 *    positive = a cudaError_t passed through.  `mdb_last_error` describes the last
 *    failure on a handle.  Nothing throws across the ABI.  There is NO CPU fallback: on a
 *    machine without a usable CUDA device `mdb_create` fails.
 */
#ifndef MDB200_H
#define MDB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* every entry point below is exported with default visibility; the library is built with
 * -fvisibility=hidden so nothing else leaks */
#if defined(__GNUC__)
#pragma GCC visibility push(default)
#endif

#define MDB_VERSION 100 /* 0.1.0 */

#define MDB_OK 0
#define MDB_E_INVALID_ARG (-1)   /* null pointer, non-positive size, unsupported C / S / Ls */
#define MDB_E_SHAPE (-2)         /* W,H < 2 (reflect pad needs 2), scale larger than target */
#define MDB_E_WORKSPACE (-3)     /* workspace pointer null / too small / misaligned */
#define MDB_E_NO_DEVICE (-4)     /* no CUDA device, or device is not sm_100 */
#define MDB_E_INTERNAL (-5)

#define MDB_MAX_SCALES 8
#define MDB_MAX_SOURCES 2
#define MDB_COMM_MAX_RANKS 16 /* ranks of the one-shot peer-write reduce (one box) */

typedef struct mdb_handle mdb_handle; /* opaque; caches device properties, profiling events, the communicator */

/* ---- handle ------------------------------------------------------------------------ */
int32_t mdb_version(void);
int32_t mdb_create(mdb_handle** out, int32_t device);
int32_t mdb_destroy(mdb_handle* h);
const char* mdb_last_error(const mdb_handle* h); /* valid until the next failing call */
/* number of kernels launched through this handle since creation (bench "gpu_launches") */
int64_t mdb_launch_count(const mdb_handle* h);

/* Optional segment timing of mdb_train_loss_fwdbwd for the bench roofline: after
 * mdb_profile_begin(h, slots) every call records CUDA events on the caller's stream around its
 * 4 segments (0 pose/stats pre-pass, 1 fused kernel, 2 per-item finalize + scalar loss, 3 upsample transpose) into the
 * next slot; mdb_profile_read waits for that slot and returns the 4 durations in ms.
 * mdb_profile_begin(h, 0) switches it off.  No reference counterpart (the reference's only
 * timing is `CUDA.@time` in scratch files, src/test.jl:30).                                 */
int32_t mdb_profile_begin(mdb_handle* h, int32_t slots);
int32_t mdb_profile_read(mdb_handle* h, int32_t slot, float* ms_out /* 4 floats */);

/* FP32 roofline denominator for the bench (MEASURED_PEAKS.json has none): best of `repeats` runs of a kernel of
 * independent FFMA chains filling every SM, timed with CUDA events; returns 2 x FMAs / time in TFLOP/s.  `scratch`:
 * device memory of at least num_sms x 8 x 256 floats.  Synchronises `stream` (bench-only entry).            */
int32_t mdb_measure_fp32_peak(mdb_handle* h, void* scratch, size_t scratch_bytes, int32_t repeats,
                              double* tflops_out, double* ms_out, void* stream);

/* Test hooks, process-global, never read from the environment: "strip_segs" = n forces n row segments per strip
 * in the fused kernel (0 = the library's choice), "upsample_split" = 1 forces the two-kernel transpose that
 * W > 4096 takes; "decision_dump" = device pointer to Ls N (2C+2) H W int8 makes the fused kernel (a separate
 * instantiation) record every discrete choice its pullback depends on (layout in csrc/photo_strip.inl), 0 = off;
 * "render_precise" = 1 makes the render forward sample every plane through the double-precision homography.
 * "strip_taper" = k (2..8) forces k row segments of decreasing length (long ones launched first), -1 forces equal segments,
 * 0 = the library's choice; "strip_taper_pow10" = 10 p makes those lengths proportional to (k - i)^p (0 = shipped: p = 2).
 * "strip_seg_dump" = HOST pointer to 10 int32 (0 = off): mdb_train_loss_strip_geometry also writes { tapered?, the 9 row
 * boundaries } there (tests of the host-side segment rule; needs no GPU).
 * Set BEFORE mdb_train_loss_workspace_bytes (the record count depends on the segments).                       */
int32_t mdb_debug_option(const char* name, int64_t value);

/* ---- parameters of the loss ---------------------------------------------------------- */
typedef struct mdb_loss_params {
  double min_depth;           /* Params.min_depth (Float64)  src/Monodepth.jl:38 (0.1)   */
  double max_depth;           /* Params.max_depth (Float64)  src/Monodepth.jl:39 (100)   */
  double disparity_smoothness;/* Params.disparity_smoothness src/Monodepth.jl:40 (1e-3)  */
  float alpha;                /* photometric_loss alpha      src/training.jl:2   (0.85)  */
  float ssim_c1;              /* SSIM c1 = 0.01^2            src/utils.jl:23            */
  float ssim_c2;              /* SSIM c2 = 0.03^2            src/utils.jl:23            */
  int32_t automasking;        /* Params.automasking          src/training.jl:60-62      */
  int32_t normalize_disparity;/* 1: train_loss (src/training.jl:64-65); 0: slow_depth   */
  int32_t smooth_weight_mode; /* 0: disparity_smoothness*scale[k] (src/training.jl:66-67)
                                 1: weight 1 (src/simple_depth.jl:39)                    */
  int32_t upsample_align_corners; /* src/training.jl:45 upsample_bilinear: 1 = align-corners (NNlib's default, the
                                   * throughput path); 0 = half-pixel centres max(0, (p + 0.5) w_k / W - 0.5) -- nothing in
                                   * the reference pins which (unpinned NNlib); 0 runs the global-memory kernel instantiation */
  int32_t n_global;           /* batch size the means divide by; 0 = N (single GPU)      */
  int32_t abs_subgradient;    /* pullback of abs.() at exactly 0 (Zygote, unpinned): 0 = sign(0) = 0
                                 (ChainRules' scalar rule); 1 = ForwardDiff's Dual rule, which Zygote's
                                 broadcast uses on CuArrays: +1 at +0.0.  Decides the first iterations
                                 of slow_depth, whose disparity starts constant (src/simple_depth.jl:8) */
} mdb_loss_params;

void mdb_loss_params_default(mdb_loss_params* p);

/* ---- P2-P4: so3 / pose  (src/utils.jl:106-145, 185-192) ------------------------------ */
/* hat(rvec): (3,N) -> (3,3,N)                                   src/utils.jl:123-132 */
int32_t mdb_hat(mdb_handle* h, const float* rvec, int32_t N, float* S, void* stream);
/* rrule(hat) pullback: d (3,3,N) -> d_rvec (3,N)                 src/utils.jl:134-145 */
int32_t mdb_hat_pullback(mdb_handle* h, const float* d, int32_t N, float* d_rvec, void* stream);
/* so3_exp_map(rvec): (3,N) -> (3,3,N)                            src/utils.jl:106-121 */
int32_t mdb_so3_exp_map(mdb_handle* h, const float* rvec, int32_t N, float* R, void* stream);
/* composeT(rvec, t, invert) -> R (3,3,N), t (3,1,N)              src/utils.jl:185-192 */
int32_t mdb_compose_T(mdb_handle* h, const float* rvec, const float* tvec, int32_t invert,
                      int32_t N, float* R, float* t, void* stream);
/* pullback of composeT: (dR, dt) -> (d_rvec, d_tvec)                                    */
int32_t mdb_compose_T_bwd(mdb_handle* h, const float* rvec, const float* tvec, int32_t invert,
                          int32_t N, const float* dR, const float* dt, float* d_rvec,
                          float* d_tvec, void* stream);

/* ---- P5-P8: depth / backproject / project / sample ---------------------------------- */
/* disparity_to_depth, elementwise over n values                  src/utils.jl:179-183 */
int32_t mdb_disparity_to_depth(mdb_handle* h, const float* disparity, int64_t n, double min_depth,
                               double max_depth, float* depth, void* stream);
/* (b::Backproject)(depth (1,WH,N), invK) -> points (3,WH,N)      src/utils.jl:45-69   */
int32_t mdb_backproject(mdb_handle* h, const float* depth, const float* invK, int32_t W,
                        int32_t H, int32_t N, float* points, void* stream);
/* (p::Project)(points, K, R, t) -> normalised uv (2,WH,N)        src/utils.jl:71-103  */
int32_t mdb_project(mdb_handle* h, const float* points, const float* K, const float* R,
                    const float* t, int32_t W, int32_t H, int32_t N, float* uv, void* stream);
/* NNlib grid_sample(input (W,H,C,N), grid (2,Wo,Ho,N); padding_mode=:border),
 * align-corners=true                          call sites src/training.jl:56, src/render.jl:88 */
int32_t mdb_grid_sample(mdb_handle* h, const float* input, const float* grid, int32_t W,
                        int32_t H, int32_t C, int32_t N, int32_t Wo, int32_t Ho, float* out,
                        void* stream);
/* warp(disp, x, Ps, ...; source_ids): called at src/simple_depth.jl:30-32, undefined in the
 * fork; semantics = the inlined body at src/training.jl:48-57.  disp (1,W,H,N) full res,
 * x (W,H,C,L,N), R[s] (3,3,N), t[s] (3,1,N) -> warped[s] (W,H,C,N)                       */
int32_t mdb_warp(mdb_handle* h, const float* disp, const float* x, const float* const* R,
                 const float* const* t, const int32_t* source_ids, int32_t S, const float* K,
                 const float* invK, int32_t W, int32_t H, int32_t C, int32_t L, int32_t N,
                 double min_depth, double max_depth, float* const* warped, void* stream);
/* pullback of warp: g_warped[s] (W,H,C,N) -> d_disp (1,W,H,N), dR[s] (3,3,N), dt[s] (3,1,N) (dR / dt or single
 * entries may be NULL).  With mdb_prediction_loss_bwd and mdb_smooth_loss this differentiates the closure of
 * src/simple_depth.jl:25-41 function by function, as Zygote does in the reference (the fused
 * mdb_train_loss_fwdbwd computes the same gradient in one pass).  Deterministic: no float atomics.            */
int32_t mdb_warp_bwd(mdb_handle* h, const float* disp, const float* x, const float* const* R, const float* const* t,
                     const int32_t* source_ids, int32_t S, const float* K, const float* invK, int32_t W, int32_t H,
                     int32_t C, int32_t L, int32_t N, double min_depth, double max_depth, const float* const* g_warped,
                     float* d_disp, float* const* dR, float* const* dt, void* stream);

/* ---- P9-P11: SSIM / photometric / min-reprojection / automask ------------------------ */
/* (ssim::SSIM)(x, y) -> (W,H,C,N)                                src/utils.jl:17-43   */
int32_t mdb_ssim(mdb_handle* h, const float* x, const float* y, int32_t W, int32_t H, int32_t C,
                 int32_t N, float c1, float c2, float* out, void* stream);
/* photometric_loss(ssim, predicted, target; alpha) -> (W,H,1,N)  src/training.jl:1-7  */
int32_t mdb_photometric_loss(mdb_handle* h, const float* predicted, const float* target,
                             int32_t W, int32_t H, int32_t C, int32_t N, float alpha, float c1,
                             float c2, float* out, void* stream);
/* prediction_loss(ssim, predictions, target) -> (W,H,1,N)        src/training.jl:13-15 */
int32_t mdb_prediction_loss(mdb_handle* h, const float* const* predictions, int32_t S,
                            const float* target, int32_t W, int32_t H, int32_t C, int32_t N,
                            float alpha, float c1, float c2, float* out, void* stream);
/* automasking_loss(ssim, inputs (W,H,C,L,N), target (W,H,C,N); source_ids)  src/training.jl:9-11;
 * callers src/dtk.jl:59-61 pass target = x[:, :, :, target_id, :].  `target` is the target ARRAY as in the
 * reference; target_item_stride = elements between two items of it: 0 or W H C for a dense (W,H,C,N) array
 * (what the reference's slicing copy produces), W H C L for a view into `inputs` (no copy).              */
int32_t mdb_automasking_loss(mdb_handle* h, const float* inputs, const float* target, int64_t target_item_stride,
                             const int32_t* source_ids, int32_t S, int32_t W, int32_t H, int32_t C, int32_t L,
                             int32_t N, float alpha, float c1, float c2, float* out, void* stream);
/* _apply_mask(mask, warp_loss): elementwise min over n values    src/training.jl:17-19 */
int32_t mdb_apply_mask(mdb_handle* h, const float* mask, const float* warp_loss, int64_t n,
                       float* out, void* stream);
/* pullbacks of SSIM / prediction_loss / _apply_mask w.r.t. the predictions (the target gets no gradient on the
 * path).  g is the cotangent of the output: (W,H,C,N) for SSIM, (W,H,1,N) for prediction_loss.  The arg-min of
 * `minimum(cat(...; dims=3); dims=3)` takes the whole cotangent (first arg-min on ties); abs_subgradient as in
 * mdb_loss_params.  workspace: mdb_ssim_bwd_workspace_bytes(W,H,C,N) bytes, 256-byte aligned, for both.      */
size_t mdb_ssim_bwd_workspace_bytes(int32_t W, int32_t H, int32_t C, int32_t N);
int32_t mdb_ssim_bwd(mdb_handle* h, const float* x, const float* y, const float* g, int32_t W, int32_t H, int32_t C,
                     int32_t N, float c1, float c2, float* dx, void* workspace, size_t workspace_bytes, void* stream);
int32_t mdb_prediction_loss_bwd(mdb_handle* h, const float* const* predictions, int32_t S, const float* target,
                                const float* g, int32_t W, int32_t H, int32_t C, int32_t N, float alpha, float c1,
                                float c2, int32_t abs_subgradient, float* const* d_predictions, void* workspace,
                                size_t workspace_bytes, void* stream);
/* d_mask / d_warp_loss may be NULL; the mask keeps ties (it comes first in the cat, src/training.jl:18)        */
int32_t mdb_apply_mask_bwd(mdb_handle* h, const float* mask, const float* warp_loss, const float* g, int64_t n,
                           float* d_mask, float* d_warp_loss, void* stream);

/* ---- S1-S2: smoothness  (src/utils.jl:163-177) ---------------------------------------- */
/* smooth_loss(disparity (W,H,N), image (W,H,C,N)) -> scalar *loss (device); if d_disparity
 * is non-null also writes d loss / d disparity (W,H,N).  Deterministic (no float atomics).
 * workspace: mdb_smooth_loss_workspace_bytes(W,H,N) bytes, 256-byte aligned.              */
size_t mdb_smooth_loss_workspace_bytes(int32_t W, int32_t H, int32_t N);
int32_t mdb_smooth_loss(mdb_handle* h, const float* disparity, const float* image, int32_t W,
                        int32_t H, int32_t C, int32_t N, float* loss, float* d_disparity,
                        void* workspace, size_t workspace_bytes, void* stream);
/* the same with the pullback of abs at 0 chosen (see mdb_loss_params.abs_subgradient)   */
int32_t mdb_smooth_loss_ex(mdb_handle* h, const float* disparity, const float* image, int32_t W,
                           int32_t H, int32_t C, int32_t N, int32_t abs_subgradient, float* loss,
                           float* d_disparity, void* workspace, size_t workspace_bytes, void* stream);
/* smooth_loss with NO allocation and no workspace argument.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same kernels and result as mdb_smooth_loss_ex; the scratch (per-block partial sums) is a
 * static __device__ array inside the library of MDB_SMOOTH_NOALLOC_MAX_BYTES = 256 KiB, held by every GPU that loads the
 * library until exit.  Limit: N * ceil(H / 8) <= 32768 (e.g. 64 items of 1024x320 need 2560); larger shapes are
 * rejected with MDB_E_WORKSPACE.  All calls on a device share the array: calls that overlap on the GPU (different
 * streams / threads) race on it -- serialise them.  Nothing in it needs clearing between calls.            */
#define MDB_SMOOTH_NOALLOC_MAX_BYTES (256 * 1024)
int32_t mdb_smooth_loss_noalloc(mdb_handle* h, const float* disparity, const float* image, int32_t W,
                                int32_t H, int32_t C, int32_t N, int32_t abs_subgradient, float* loss,
                                float* d_disparity, void* stream);

/* ---- P12: the fused hot path  (src/training.jl:21-78; src/simple_depth.jl:25-41) ------ */
typedef struct mdb_loss_io {
  /* inputs */
  const float* x;                 /* frame stack (W,H,C,L,N)                             */
  int32_t W, H, C, L, N;
  int32_t target_id;              /* 1-based, cache.target_id                            */
  int32_t S;                      /* number of sources (<= MDB_MAX_SOURCES)              */
  int32_t source_ids[MDB_MAX_SOURCES]; /* 1-based, cache.source_ids                     */
  int32_t invert[MDB_MAX_SOURCES];     /* source_id < target_id, src/training.jl:28     */
  const float* rvec[MDB_MAX_SOURCES];  /* Pose.rvec (3,N)                               */
  const float* tvec[MDB_MAX_SOURCES];  /* Pose.tvec (3,1,N)                             */
  int32_t Ls;                     /* number of scales (<= MDB_MAX_SCALES)                */
  const float* disparity[MDB_MAX_SCALES]; /* (w_k,h_k,1,N), native resolution           */
  int32_t dw[MDB_MAX_SCALES], dh[MDB_MAX_SCALES];
  float scale[MDB_MAX_SCALES];    /* cache.scales                                        */
  const float* auto_loss;         /* (W,H,1,N) identity-reprojection loss or NULL        */
  float K[9], invK[9];            /* column-major 3x3                                     */
  /* outputs */
  float* loss;                    /* 1 float: loss / Ls summed over the LOCAL items      */
  float* d_disparity[MDB_MAX_SCALES]; /* d loss / d disparity[k], native resolution     */
  float* d_rvec[MDB_MAX_SOURCES]; /* (3,N)                                               */
  float* d_tvec[MDB_MAX_SOURCES]; /* (3,1,N)                                             */
  float* vis_warp_loss;           /* optional (W,H,1,N): warp_loss of the last scale     */
  int32_t x_format;               /* 0: x is Float32; 1: x points to uint8 (N0f8) samples in the same (W,H,C,L,N) layout,
                                   * read in the kernel as Float32(::N0f8) = u / 255 correctly rounded (src/dtk.jl:45) --
                                   * bit-identical results, a quarter of the image bytes.  Not under stream capture, not
                                   * with upsample_align_corners = 0 (MDB_E_INVALID_ARG).  mdb_ingest_triplet_u8_planar
                                   * produces this layout from the dataset's strips.                                  */
} mdb_loss_io;

size_t mdb_train_loss_workspace_bytes(const mdb_loss_io* io);
/* launch decomposition of the fused kernel for this shape on the current device: out4 = { 28-column strips per
 * row, row segments per strip, rows per segment, warps in the launch } (bench roofline: halo overhead)       */
int32_t mdb_train_loss_strip_geometry(const mdb_loss_io* io, int32_t* out4);
/* train_loss value AND gradient in one pass (forward of the rrule; the pullback scales the
 * cached gradients by the incoming scalar cotangent).                                    */
int32_t mdb_train_loss_fwdbwd(mdb_handle* h, const mdb_loss_io* io, const mdb_loss_params* p,
                              void* workspace, size_t workspace_bytes, void* stream);

/* ---- train_loss for inputs in HOST memory (the reference: `transfer(x)` = gpu . f32, src/utils.jl:1-3, then
 * train_loss).  A plan cuts the batch into chunks of items and pipelines host->device copies, the fused loss and the
 * device->host drain of the disparity gradients over three streams, one cudaMemcpyAsync per array and chunk; every chunk
 * is one mdb_train_loss_fwdbwd with n_global = the whole batch, so chunk losses add.  All pointers of mdb_host_io are
 * HOST pointers (pinned memory gives the full PCIe rate); layouts as in mdb_loss_io.  x_format 0: x is the float32
 * (W,H,C,L,N) stack; 1: x is uint8 triplet strips [N][H][3W][3] converted on the device (mdb_ingest_triplet_u8, L = 3),
 * `flip` NULL or N bytes.  chunk_items <= 0 picks N/8 but at least ~16 MB of input per chunk.  first_item >= 0: the
 * batch is a shard of params.n_global items starting there; the call then ends with the path's all-reduce
 * (mdb_comm_* must be set up on the handle) and d_rvec / d_tvec receive the GLOBAL (3,n_global) gradients, loss the
 * global loss; the disparity gradients stay the shard's.  The plan owns its device buffers (cudaMalloc at create).
 * mdb_train_loss_host returns when every result is in host memory.                                               */
typedef struct mdb_host_plan mdb_host_plan;
typedef struct mdb_host_io {
  const void* x;
  const uint8_t* flip;
  const float* disparity[MDB_MAX_SCALES];
  const float* rvec[MDB_MAX_SOURCES];
  const float* tvec[MDB_MAX_SOURCES];
  const float* auto_loss;
  float* loss;                          /* 1 float */
  float* d_disparity[MDB_MAX_SCALES];
  float* d_rvec[MDB_MAX_SOURCES];
  float* d_tvec[MDB_MAX_SOURCES];
} mdb_host_io;
int32_t mdb_host_plan_create(mdb_handle* h, const mdb_loss_io* shape /* sizes, ids, K; pointers ignored */,
                             const mdb_loss_params* params, int32_t chunk_items, int32_t x_format, int32_t first_item,
                             mdb_host_plan** out);
int32_t mdb_train_loss_host(mdb_host_plan* plan, const mdb_host_io* io, void* stream);
/* out4 = { chunks, host->device bytes per call, device->host bytes per call, items d_rvec / d_tvec cover }      */
int32_t mdb_host_plan_info(const mdb_host_plan* plan, int64_t* out4);
int32_t mdb_host_plan_destroy(mdb_host_plan* plan);

/* ---- the batch sharded over the GPUs of one box ------------------------------------------
 * The reference is single-GPU; the only coupling between items is the mean over N (src/training.jl:69,
 * src/utils.jl:176).  Rank r evaluates its contiguous shard with params.n_global = N_global, so shard losses
 * add, and ONE sum over the ranks of the packed vector
 *     [ loss | d_rvec[s][n] (3 floats) for s, n | d_tvec[s][n] (3 floats) for s, n ]    1 + 6 S N_global floats
 * gives every rank the batch loss and the full (3,N) pose gradients; disparity gradients stay sharded.
 * mdb_packed_bind points io->loss / d_rvec / d_tvec INTO the rank's slots of a packed SEND buffer (all other
 * slots zero, written once by the caller), so mdb_train_loss_fwdbwd fills it in place and the step is the five
 * loss kernels + one all-reduce: no packing kernels.                                                     */
int64_t mdb_packed_size(int32_t n_global, int32_t S);
int32_t mdb_packed_bind(mdb_loss_io* io, float* packed_send, int32_t n_global, int32_t first_item);
/* transport 1: NCCL (libnccl.so.2 resolved at run time with dlopen; in a PyTorch process the already loaded
 * copy).  Rank 0 calls mdb_comm_get_unique_id (128 bytes) and the launcher hands the bytes to every rank.   */
int32_t mdb_comm_get_unique_id(void* out, size_t bytes);
int32_t mdb_comm_init(mdb_handle* h, const void* unique_id, int32_t rank, int32_t world);
/* transport 2: one-shot peer writes over NVLink, no library.  _export allocates this rank's symmetric buffer
 * (the one allocation the library makes itself: CUDA IPC exports whole cudaMalloc allocations) and returns its
 * 64-byte cudaIpcMemHandle; the launcher gathers the handles of all ranks (rank order) and calls _attach, then
 * synchronises the ranks once.  Every all-reduce is then ONE kernel: write own vector into its slot on every
 * peer, raise a flag there, wait (bounded) for the peers' flags, add the slots in rank order -- deterministic
 * and bit-identical on all ranks.  When both transports are set up this one is used.                         */
int32_t mdb_comm_peers_export(mdb_handle* h, int32_t world, int64_t max_count, void* ipc_handle_out, size_t bytes);
int32_t mdb_comm_peers_attach(mdb_handle* h, const void* ipc_handles /* world x 64 bytes */, int32_t rank);
/* recv[i] = sum over ranks of send[i] (send != recv allowed and intended: the send buffer keeps its zeros).
 * Enqueued on `stream`.                                                                                     */
int32_t mdb_allreduce_loss_posegrad(mdb_handle* h, const float* send, float* recv, int64_t count, void* stream);
/* *timed_out = 1 if a one-shot reduce gave up waiting for a peer (~10 s); synchronises the device           */
int32_t mdb_comm_status(mdb_handle* h, int32_t* timed_out);
int32_t mdb_comm_destroy(mdb_handle* h); /* also done by mdb_destroy */

/* ---- R1-R7: plane-disparity volume rendering  (src/render.jl) ------------------------------
 * Layouts: rgb (W,H,3,Np,B), sigma (W,H,1,Np,B), xyz (3,W,H,Np,B), disparity_src (Np,B),
 * pose rvec (3,B), tvec (3,B) (2-D, as test/test_sample.jl:46 builds it).  The reference's
 * quirks are reproduced as written; 1 <= Np <= 64.                              */
/* Monodepth.inv: batched 3x3 inverse (3,3,M)                       src/render.jl:3-17    */
int32_t mdb_inv3x3_batched(mdb_handle* h, const float* A, int32_t M, float* out, void* stream);
/* create_meshgrid(H, W) -> (3,W,H) = [w,h,1], 1-based              src/render.jl:21-23   */
int32_t mdb_create_meshgrid(mdb_handle* h, int32_t H, int32_t W, float* out, void* stream);
/* get_src_xyz_from_plane_disparity(meshgrid, disparity, K_inv)     src/render.jl:25-30   */
int32_t mdb_plane_xyz_src(mdb_handle* h, const float* meshgrid, const float* disparity, const float* K_inv,
                          int32_t W, int32_t H, int32_t Np, int32_t B, float* out, void* stream);
/* get_tgt_xyz_from_plane_disparity(xyz_src, pose)                  src/render.jl:51-62   */
int32_t mdb_plane_xyz_tgt(mdb_handle* h, const float* xyz_src, const float* rvec, const float* tvec, int32_t W,
                          int32_t H, int32_t Np, int32_t B, float* out, void* stream);
/* workspace for mdb_sample / mdb_render_* (0 if the shape is unsupported)                   */
size_t mdb_render_workspace_bytes(int32_t W, int32_t H, int32_t Np, int32_t B);
/* sample(src (W,H,Cs,Np*B), depth_src (Np,B), pose, K, K_inv) -> tgt, valid_mask (W*H, Np*B)
 * as bytes                                                          src/render.jl:64-90   */
int32_t mdb_sample(mdb_handle* h, const float* src, const float* depth_src, const float* rvec, const float* tvec,
                   const float* K, const float* K_inv, int32_t W, int32_t H, int32_t Cs, int32_t Np, int32_t B,
                   float* tgt, uint8_t* valid_mask, void* workspace, size_t workspace_bytes, void* stream);
/* plane_volume_rendering(rgb, sigma, xyz) -> rgb_out (W,H,3,B), transparency_acc, weights
 * (W,H,1,Np,B); weights may be NULL                                 src/render.jl:32-49   */
int32_t mdb_plane_volume_rendering(mdb_handle* h, const float* rgb, const float* sigma, const float* xyz, int32_t W,
                                   int32_t H, int32_t Np, int32_t B, float* rgb_out, float* transparency_acc,
                                   float* weights, void* stream);
/* pullbacks of the two (the reference has none: `sample` mutates in place).  mdb_sample_bwd: g_tgt
 * (W,H,Cs,Np*B) -> d_src, the bilinear gather transposed with the homography held fixed (pose / depth gradients
 * flow through mdb_render_tgt_rgb_depth_bwd); d_src is zeroed here and accumulated with red.global.add.f32.
 * mdb_plane_volume_rendering_bwd: cotangents of (rgb_out, transparency_acc, weights), any of them NULL ->
 * d_rgb, d_sigma, d_xyz (d_xyz may be NULL); NaN where two planes coincide, as the reference's sqrt would give. */
int32_t mdb_sample_bwd(mdb_handle* h, const float* g_tgt, const float* depth_src, const float* rvec, const float* tvec,
                       const float* K, const float* K_inv, int32_t W, int32_t H, int32_t Cs, int32_t Np, int32_t B,
                       float* d_src, void* workspace, size_t workspace_bytes, void* stream);
int32_t mdb_plane_volume_rendering_bwd(mdb_handle* h, const float* rgb, const float* sigma, const float* xyz, int32_t W,
                                       int32_t H, int32_t Np, int32_t B, const float* g_rgb_out,
                                       const float* g_transparency_acc, const float* g_weights, float* d_rgb,
                                       float* d_sigma, float* d_xyz, void* stream);
/* render_tgt_rgb_depth(rgb, sigma, disparity_src, xyz_tgt, pose, K_inv, K) -> rgb (W,H,3,B),
 * "depth" = transparency_acc (W,H,1,Np,B), mask (W,H,1,1,B) int32   src/render.jl:92-114  */
int32_t mdb_render_tgt_rgb_depth_fwd(mdb_handle* h, const float* rgb, const float* sigma, const float* disparity_src,
                                     const float* xyz_tgt, const float* rvec, const float* tvec, const float* K_inv,
                                     const float* K, int32_t W, int32_t H, int32_t Np, int32_t B, float* rgb_out,
                                     float* depth_out, int32_t* mask_out, void* workspace, size_t workspace_bytes,
                                     void* stream);
/* pullback of the above (the reference has none: its `sample` mutates in place).  depth_saved is
 * the forward's depth output; g_depth may be NULL; d_disparity_src / d_rvec / d_tvec may be NULL.
 * d_rgb / d_sigma / d_xyz_tgt are zeroed here and accumulated with red.global.add.f32.        */
int32_t mdb_render_tgt_rgb_depth_bwd(mdb_handle* h, const float* rgb, const float* sigma, const float* disparity_src,
                                     const float* xyz_tgt, const float* rvec, const float* tvec, const float* K_inv,
                                     const float* K, int32_t W, int32_t H, int32_t Np, int32_t B, const float* depth_saved,
                                     const float* g_rgb_out, const float* g_depth, float* d_rgb, float* d_sigma,
                                     float* d_xyz_tgt, float* d_disparity_src, float* d_rvec, float* d_tvec,
                                     void* workspace, size_t workspace_bytes, void* stream);

/* ---- render variants ------------------------------------------------------
 * What the reference's own comparison script quotes from MINE as the intended sampler
 * (test/test_sample.jl:123-137) and the TODOs of src/render.jl:22,36,40,47,52, behind options; all zero
 * (or opts == NULL) is the code as written.
 *   mine_grid  grid = (x + 0.5) / (W/2) - 1 sampled with align_corners=false, border padding: the sample
 *              lands on source pixel x itself (as written: no "- 1", align-corners=true; SURVEY F8a,b)
 *   mine_mask  valid = -1 < x < W and -1 < y < H   (as written: x > 0 and y > 0)
 *   depth_eps  expected depth = sum_k w_k z_k / (sum_k w_k + depth_eps), written to `expected_depth`
 *              (W,H,1,B) when that pointer is given ("TODO: return depth?", src/render.jl:47);
 *              `weights_sum` (W,H,1,B) = sum_k w_k is what the pullback needs back
 * xyz_tgt == NULL folds the producers in: xyz_tgt is what get_tgt_xyz_from_plane_disparity(
 * get_src_xyz_from_plane_disparity(create_meshgrid(H, W), disparity_src, K_inv), pose) would hold
 * (src/render.jl:21-30, 51-62), evaluated at the sample point -- the 3 W H Np B array is never
 * materialised, and in the pullback its gradient flows on into d_disparity_src / d_rvec / d_tvec
 * (d_xyz_tgt must be NULL too).                                                                  */
typedef struct mdb_render_opts {
  int32_t mine_grid;
  int32_t mine_mask;
  float depth_eps;   /* 1e-5 when opts == NULL */
  int32_t reserved;
} mdb_render_opts;
int32_t mdb_render_tgt_rgb_depth_fwd_ex(mdb_handle* h, const float* rgb, const float* sigma, const float* disparity_src,
                                        const float* xyz_tgt, const float* rvec, const float* tvec, const float* K_inv,
                                        const float* K, int32_t W, int32_t H, int32_t Np, int32_t B,
                                        const mdb_render_opts* opts, float* rgb_out, float* depth_out, int32_t* mask_out,
                                        float* expected_depth, float* weights_sum, void* workspace, size_t workspace_bytes,
                                        void* stream);
int32_t mdb_render_tgt_rgb_depth_bwd_ex(mdb_handle* h, const float* rgb, const float* sigma, const float* disparity_src,
                                        const float* xyz_tgt, const float* rvec, const float* tvec, const float* K_inv,
                                        const float* K, int32_t W, int32_t H, int32_t Np, int32_t B,
                                        const mdb_render_opts* opts, const float* depth_saved,
                                        const float* expected_depth_saved, const float* weights_sum_saved,
                                        const float* g_rgb_out, const float* g_depth, const float* g_expected_depth,
                                        float* d_rgb, float* d_sigma, float* d_xyz_tgt, float* d_disparity_src,
                                        float* d_rvec, float* d_tvec, void* workspace, size_t workspace_bytes, void* stream);

/* ---- the optimiser loop of slow_depth  (src/simple_depth.jl:14-16, 22-42) ----
 * Flux.Optimise.update!(ADAM(eta, (beta1, beta2)), theta, grads) over a list of tensors (third-party
 * Flux.jl, unpinned; semantics restated in oracle/torch_ref.py::flux_adam_step):
 *   m = b1 m + (1-b1) g;  v = b2 v + (1-b2) g^2;  x -= eta * m / (1 - b1^t) / (sqrt(v / (1 - b2^t)) + eps)
 * with t = *step + 1.  `step` is a DEVICE int32 that the kernel itself advances (last block out), so the
 * call can be replayed from a CUDA graph.  If loss_history != NULL, loss_history[*step] = *loss (both
 * device) is recorded before the counter advances (history_len bounds the index).                 */
typedef struct mdb_adam_params {
  double eta;    /* 3e-4, src/simple_depth.jl:16 */
  double beta1;  /* 0.9   (Flux default)         */
  double beta2;  /* 0.999 (Flux default)         */
  double eps;    /* 1e-8  (Flux <= 0.13 default) */
} mdb_adam_params;
#define MDB_ADAM_MAX_TENSORS 8
/* state: n_tensors pairs (m, v) with the tensors' sizes, all zero before the first step; `ticket` is a
 * DEVICE int32 initialised to 0 (scratch of the last-block-out counter).                          */
int32_t mdb_adam_update(mdb_handle* h, int32_t n_tensors, float* const* params, const float* const* grads,
                        float* const* m, float* const* v, const int64_t* sizes, const mdb_adam_params* a,
                        int32_t* step, int32_t* ticket, const float* loss, float* loss_history,
                        int32_t history_len, void* stream);

/* `slow_depth` (src/simple_depth.jl:1-62) without its logging: `iters` iterations of
 *   grads = d/d(disp, poses) [ mean(prediction_loss(warp(disp, x, Ps))) + smooth_loss(disp, target) ];
 *   update!(ADAM(eta), (disp, poses), grads)
 * One iteration is captured ONCE into a CUDA graph at plan creation -- the loss kernels of
 * mdb_train_loss_fwdbwd, whose finalize kernel also applies the ADAM step of the poses, and beside it
 * (forked after the fused kernel) the ADAM kernel of the disparity map; mdb_slow_depth_run replays it
 * `iters` times on the caller's stream with no host round trip.  `io` must describe one full-resolution scale (Ls = 1, dw = W, dh = H); its
 * disparity / rvec / tvec buffers are the parameters and are UPDATED IN PLACE; its d_* buffers and
 * `loss` hold the last iteration's gradients and loss.  `params` MUST have normalize_disparity = 0,
 * smooth_weight_mode = 1 and automasking = 0 (src/simple_depth.jl:39; MDB_E_INVALID_ARG otherwise).  `state` is caller-owned device memory of
 * mdb_slow_depth_state_bytes(io) bytes, zeroed by the caller before the first run (ADAM moments +
 * counters); `workspace` as for mdb_train_loss_fwdbwd.  All pointers must stay valid and unchanged
 * for the life of the plan.  Layout of `state`: per parameter tensor (disparity, then rvec, tvec per
 * source) its m then its v, then at the next 256-byte boundary the int32 counter block
 * [0] step of the disparity's ADAM kernel, [1] its ticket, [2] step of the pose ADAM inside the finalize
 * kernel, floats [4..7] bias corrections handed from one iteration to the next.  To resume from a
 * checkpoint set [0] = [2] = step and zero [1] and [4..7].                                           */
typedef struct mdb_slow_depth_plan mdb_slow_depth_plan;
size_t mdb_slow_depth_state_bytes(const mdb_loss_io* io);
int32_t mdb_slow_depth_create(mdb_handle* h, const mdb_loss_io* io, const mdb_loss_params* params,
                              const mdb_adam_params* adam, void* state, size_t state_bytes, void* workspace,
                              size_t workspace_bytes, float* loss_history, int32_t history_len,
                              mdb_slow_depth_plan** out);
int32_t mdb_slow_depth_run(mdb_slow_depth_plan* plan, int32_t iters, void* stream);
int32_t mdb_slow_depth_destroy(mdb_slow_depth_plan* plan);

/* ---- input ingestion  (src/dtk.jl:29-46) ---------------------------------------
 * Depth10k's `getindex`: one decoded image holds the three frames side by side.  `strips` is the decoder's
 * raster, uint8 row-major [N][H][3W][3] (R,G,B interleaved), DEVICE memory; `flip` is NULL or N bytes
 * (non-zero = FlipX applied to that sample: the width axis of all three frames reversed, Augmentations.jl,
 * third-party).  Writes x (W,H,C,3,N) Float32 = value/255 (C = 3), or Gray{Float32} luma
 * 0.299 r + 0.587 g + 0.114 b (C = 1, `grayscale` != 0; Colors.jl, third-party).  Bit-exact against the
 * numpy restatement oracle/ingest_ref.py.                                                          */
int32_t mdb_ingest_triplet_u8(mdb_handle* h, const uint8_t* strips, const uint8_t* flip, int32_t W, int32_t H,
                              int32_t N, int32_t grayscale, float* x, void* stream);

/* KittyDataset's `getindex` (src/kitty.jl:45-61): `frames` = the three consecutive `image_0` rasters of every sample as
 * decoded, uint8 gray row-major [N][3][H_orig][W_orig], DEVICE memory.  Each is resized to (H, W) as
 * `imresize(x, (height, width))` does for Matrix{Gray{N0f8}} (ImageTransformations.jl, third-party: half-pixel-centre
 * bilinear in Float64, result rounded back to N0f8; restated in oracle/ingest_ref.py), then `Float32.`, transposed and
 * stacked: x (W,H,1,3,N).  `flip` as for mdb_ingest_triplet_u8 (`d.augmentations`, :53-55).  Bit-exact against the numpy
 * restatement.                                                                                                     */
int32_t mdb_ingest_kitti_u8(mdb_handle* h, const uint8_t* frames, const uint8_t* flip, int32_t W_orig, int32_t H_orig,
                            int32_t W, int32_t H, int32_t N, float* x, void* stream);

/* The same split / FlipX / channel-planar transpose as mdb_ingest_triplet_u8 but WITHOUT the conversion: x_u8 is the uint8
 * (W,H,3,3,N) stack that mdb_loss_io.x_format = 1 reads (C = 3 only: Gray luma is not an 8-bit value).  9 W H N bytes read
 * and written instead of 9 read + 36 written.                                                                        */
int32_t mdb_ingest_triplet_u8_planar(mdb_handle* h, const uint8_t* strips, const uint8_t* flip, int32_t W, int32_t H,
                                     int32_t N, uint8_t* x_u8, void* stream);

/* ---- the operators on the network side of the path ---------------------------
 * PoseDecoder tail (src/pose_decoder.jl:27-32): pose = scale * mean(conv_out (w,h,6,N); dims=(1,2)),
 * Pose(pose[1:3,1,:], pose[4:6,:,:]) -> rvec (3,N), tvec (3,1,N); scale = 1e-2 in the reference.     */
int32_t mdb_pose_head(mdb_handle* h, const float* conv_out, int32_t w, int32_t hh, int32_t N, float scale,
                      float* rvec, float* tvec, void* stream);
/* its pullback: d_conv_out[w,h,c,n] = scale / (w h) * (c <= 3 ? d_rvec[c,n] : d_tvec[c-3,1,n])        */
int32_t mdb_pose_head_bwd(mdb_handle* h, const float* d_rvec, const float* d_tvec, int32_t w, int32_t hh,
                          int32_t N, float scale, float* d_conv_out, void* stream);
/* DepthDecoder output heads (src/depth_decoder.jl:46, 52-68): every scale level ends in Conv((3,3), ch => 1, σ); this is
 * the σ of all Ls levels in one launch, logits[k] (w_k,h_k,1,N) -> disparity[k] (the loss path's input), with NNlib's
 * overflow-safe σ; and its pullback d_logits[k] = d_disparity[k] .* disparity[k] .* (1 - disparity[k]) from the saved
 * output, so the gradients of mdb_train_loss_fwdbwd reach the network's conv outputs in one more launch.         */
int32_t mdb_disparity_heads(mdb_handle* h, const float* const* logits, const int32_t* dw, const int32_t* dh, int32_t Ls,
                            int32_t N, float* const* disparity, void* stream);
int32_t mdb_disparity_heads_bwd(mdb_handle* h, const float* const* disparity, const float* const* d_disparity, const int32_t* dw,
                                const int32_t* dh, int32_t Ls, int32_t N, float* const* d_logits, void* stream);
/* per-item mean of a map with `per_item` floats per item (deterministic): the reduction of find_static,
 * mean(automasking_loss(ssim, x, target; source_ids)) > alpha                    src/dtk.jl:51-69     */
int32_t mdb_item_mean(mdb_handle* h, const float* map, int64_t per_item, int32_t N, float* out, void* stream);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* MDB200_H */
