// photo_fused_host.inl -- finalize kernels and the C-ABI entry of the fused loss
namespace {

// ---------------------------------------------------------------------------------------
// 6. per-item finalize: fixed-order sum of tile partials (double), pose chain to rvec/tvec
// ---------------------------------------------------------------------------------------
constexpr int FIN_THREADS = 128;

template <int S>
__global__ void __launch_bounds__(FIN_THREADS)
finalize_item_kernel(const float* __restrict__ partials, int tiles, PoseArgs pa,
                     float* const* d_rvec_unused, float* dr0, float* dr1, float* dt0, float* dt1,
                     double* __restrict__ item_loss) {
  constexpr int NPART = 1 + NPOSE * S;
  const int n = blockIdx.x;
  double acc[NPART];
#pragma unroll
  for (int k = 0; k < NPART; ++k) acc[k] = 0.0;
  for (int t = threadIdx.x; t < tiles; t += FIN_THREADS) {
    const float* p = partials + ((size_t)n * tiles + t) * NPART;
#pragma unroll
    for (int k = 0; k < NPART; ++k) acc[k] += (double)p[k];
  }
  __shared__ double red[FIN_THREADS / 32][NPART];
  __shared__ double tot[NPART];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < NPART; ++k) {
    double v = acc[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(MDB_FULL, v, o);
    if (lane == 0) red[warp][k] = v;
  }
  __syncthreads();
  if (threadIdx.x < NPART) {
    double v = 0.0;
    for (int w = 0; w < FIN_THREADS / 32; ++w) v += red[w][threadIdx.x];
    tot[threadIdx.x] = v;
  }
  __syncthreads();
  if (threadIdx.x == 0) item_loss[n] = tot[0];
  if (threadIdx.x < S) {
    const int s = threadIdx.x;
    const double* dA = tot + 1 + s * NPOSE;  // row-major dL/dA
    const double* db = dA + 9;
    // A = K R' invK  ->  dR' = K^T dA invK^T ;  b = K t' -> dt' = K^T db
    double Kt[9], iKt[9], t1[9], dRp[9], dtp[3];
    mat3_T(pa.K, Kt);
    mat3_T(pa.invK, iKt);
    mat3_mul(Kt, dA, t1);
    mat3_mul(t1, iKt, dRp);
    for (int i = 0; i < 3; ++i) dtp[i] = Kt[3 * i] * db[0] + Kt[3 * i + 1] * db[1] + Kt[3 * i + 2] * db[2];
    double r[3], t[3], dr[3], dt[3];
    for (int k = 0; k < 3; ++k) { r[k] = pa.rvec[s][3 * n + k]; t[k] = pa.tvec[s][3 * n + k]; }
    compose_T_bwd_d(r, t, pa.invert[s] != 0, dRp, dtp, dr, dt);
    float* orv = s == 0 ? dr0 : dr1;
    float* otv = s == 0 ? dt0 : dt1;
    for (int k = 0; k < 3; ++k) {
      if (orv) orv[3 * n + k] = (float)dr[k];
      if (otv) otv[3 * n + k] = (float)dt[k];
    }
  }
}

// 7. scalar loss = up * sum(item photometric sums) + sum(smoothness terms)
__global__ void finalize_loss_kernel(const double* __restrict__ item_loss, int N, const ScaleStat* __restrict__ stats,
                                     int nstats, double up, float* __restrict__ loss) {
  __shared__ double red[32];
  double v = 0.0;
  for (int i = threadIdx.x; i < N; i += blockDim.x) v += item_loss[i] * up;
  for (int i = threadIdx.x; i < nstats; i += blockDim.x) v += (double)stats[i].loss;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(MDB_FULL, v, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) s += red[w];
    *loss = (float)s;
  }
}

struct FusedLayout {
  size_t o_pp, o_ppd, o_statpart, o_stats, o_partials, o_itemloss, o_gfull[MDB_MAX_SCALES], total;
  int tiles_x, tiles_y, rowblocks, npart;
  StripGeom strip;   // decomposition of the streaming kernel
  int records;       // partial-sum records per item (tile kernel: tiles; strip kernel: warps per item)
  bool use_strip;
};

// MDB_FUSED_IMPL=tile selects the first-generation tiled kernel (kept for A/B measurements)
inline bool mdb_use_strip() {
  static const int v = [] { const char* e = getenv("MDB_FUSED_IMPL"); return (e && !strcmp(e, "tile")) ? 0 : 1; }();
  return v != 0;
}

// Rows per segment: long segments amortise the 4 halo rows, short ones give the machine enough
// warps.  Aim for ~16 resident sets of 12 warps per SM, never below 32 rows.
inline StripGeom strip_geom(const mdb_loss_io* io, int num_sms) {
  StripGeom g{};
  g.nstrips = (io->W + STRIP_PX - 1) / STRIP_PX;
  const long long base = (long long)io->N * io->Ls * g.nstrips;
  long long want = ((long long)16 * 12 * (num_sms > 0 ? num_sms : 148) + base - 1) / base;
  if (const char* e = getenv("MDB_STRIP_SEGS")) want = atoll(e);
  const int max_segs = std::max(1, io->H / 32);
  g.nsegs = (int)std::min<long long>(std::max<long long>(want, 1), max_segs);
  g.seg_rows = (io->H + g.nsegs - 1) / g.nsegs;
  g.nsegs = (io->H + g.seg_rows - 1) / g.seg_rows;
  g.total = (int)(base * g.nsegs);
  return g;
}

FusedLayout fused_layout(const mdb_loss_io* io, int num_sms = 148) {
  FusedLayout L{};
  L.use_strip = mdb_use_strip();
  L.strip = strip_geom(io, num_sms);
  L.tiles_x = (io->W + TW - 1) / TW;
  L.tiles_y = (io->H + TH - 1) / TH;
  L.rowblocks = (io->H + STAT_ROWS - 1) / STAT_ROWS;
  L.npart = 1 + NPOSE * io->S;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = mdb_align_up(off + bytes, 256); return o; };
  L.o_pp = take(sizeof(PosePrep) * io->N * io->S);
  L.o_ppd = take(sizeof(PosePrepD) * io->N * io->S);
  L.o_statpart = take(sizeof(float) * (size_t)io->N * L.rowblocks * io->Ls * 3);
  L.o_stats = take(sizeof(ScaleStat) * (size_t)io->Ls * io->N);
  L.records = L.use_strip ? L.strip.nsegs * L.strip.nstrips * io->Ls : L.tiles_x * L.tiles_y;
  L.o_partials = take(sizeof(float) * (size_t)io->N * L.records * L.npart);
  L.o_itemloss = take(sizeof(double) * io->N);
  for (int s = 0; s < io->Ls; ++s) {
    const bool native = io->dw[s] == io->W && io->dh[s] == io->H;
    L.o_gfull[s] = native ? 0 : take(sizeof(float) * (size_t)io->N * io->W * io->H);
  }
  L.total = off;
  return L;
}

bool io_shape_ok(const mdb_loss_io* io) {
  if (!io || io->W < 2 || io->H < 2 || io->N < 1 || io->L < 1) return false;
  if (io->C != 1 && io->C != 3) return false;
  if (io->S < 1 || io->S > MDB_MAX_SOURCES || io->Ls < 1 || io->Ls > MDB_MAX_SCALES) return false;
  for (int s = 0; s < io->Ls; ++s)
    if (io->dw[s] < 2 || io->dh[s] < 2 || io->dw[s] > io->W || io->dh[s] > io->H) return false;
  return true;
}

template <int C, int S>
int32_t launch_fused(mdb_handle* h, const FusedArgs& fa, const FusedLayout& L, cudaStream_t st) {
  using SM = FusedSmem<C, S>;
  static bool configured = false;  // per-process; the attribute is per-function, per-device context
  (void)configured;
  MDB_CUDA(h, cudaFuncSetAttribute(photo_fused_kernel<C, S>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)SM::bytes));
  if (L.use_strip) {
    constexpr size_t smem = (size_t)3 * (S * (2 * C + 2) / 4) * STRIP_THREADS * sizeof(float4);
    const int blocks = (L.strip.total + STRIP_WARPS - 1) / STRIP_WARPS;
    photo_strip_kernel<C, S><<<blocks, STRIP_THREADS, smem, st>>>(fa, L.strip);
    MDB_LAUNCHED(h);
    return MDB_OK;
  }
  dim3 grid(L.tiles_x, L.tiles_y, fa.N);
  photo_fused_kernel<C, S><<<grid, FUSED_THREADS, SM::bytes, st>>>(fa);
  MDB_LAUNCHED(h);
  return MDB_OK;
}

}  // namespace

extern "C" {

size_t mdb_train_loss_workspace_bytes(const mdb_loss_io* io) {
  if (!io_shape_ok(io)) return 0;
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  return fused_layout(io, sms).total;
}

int32_t mdb_train_loss_fwdbwd(mdb_handle* h, const mdb_loss_io* io, const mdb_loss_params* p,
                              void* workspace, size_t workspace_bytes, void* stream) {
  MDB_REQUIRE(h, h != nullptr, MDB_E_INVALID_ARG, "null handle");
  MDB_REQUIRE(h, io && p, MDB_E_INVALID_ARG, "mdb_train_loss_fwdbwd: null io/params");
  MDB_REQUIRE(h, io_shape_ok(io), MDB_E_SHAPE,
              "mdb_train_loss_fwdbwd: unsupported shape (need W,H>=2, C in {1,3}, S<=%d, Ls<=%d, 2<=dw<=W)",
              MDB_MAX_SOURCES, MDB_MAX_SCALES);
  MDB_REQUIRE(h, io->x && io->loss, MDB_E_INVALID_ARG, "mdb_train_loss_fwdbwd: null x/loss");
  MDB_REQUIRE(h, io->target_id >= 1 && io->target_id <= io->L, MDB_E_INVALID_ARG, "target_id out of range");
  MDB_REQUIRE(h, p->upsample_align_corners == 1, MDB_E_INVALID_ARG,
              "only upsample_align_corners=1 (NNlib default) is implemented");
  MDB_REQUIRE(h, !p->automasking || io->auto_loss, MDB_E_INVALID_ARG, "automasking needs auto_loss");
  for (int s = 0; s < io->S; ++s) {
    MDB_REQUIRE(h, io->source_ids[s] >= 1 && io->source_ids[s] <= io->L, MDB_E_INVALID_ARG, "source id out of range");
    MDB_REQUIRE(h, io->rvec[s] && io->tvec[s], MDB_E_INVALID_ARG, "null pose pointer");
  }
  for (int s = 0; s < io->Ls; ++s)
    MDB_REQUIRE(h, io->disparity[s] && io->d_disparity[s], MDB_E_INVALID_ARG, "null disparity pointer (scale %d)", s);
  const FusedLayout L = fused_layout(io, h->num_sms);
  MDB_REQUIRE(h, workspace && workspace_bytes >= L.total && ((uintptr_t)workspace & 255) == 0, MDB_E_WORKSPACE,
              "workspace must be >= %zu bytes and 256-byte aligned", L.total);
  cudaStream_t st = (cudaStream_t)stream;
  char* ws = (char*)workspace;
  const int N = io->N, S = io->S, Ls = io->Ls, W = io->W, H = io->H;
  const int Ng = p->n_global > 0 ? p->n_global : N;

  PoseArgs pa{};
  pa.N = N; pa.S = S;
  for (int s = 0; s < S; ++s) { pa.rvec[s] = io->rvec[s]; pa.tvec[s] = io->tvec[s]; pa.invert[s] = io->invert[s]; }
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) { pa.K[3 * i + j] = io->K[3 * j + i]; pa.invK[3 * i + j] = io->invK[3 * j + i]; }

  FusedArgs fa{};
  fa.x = io->x; fa.W = W; fa.H = H; fa.L = io->L; fa.N = N;
  fa.tgt = io->target_id - 1;
  for (int s = 0; s < S; ++s) fa.src[s] = io->source_ids[s] - 1;
  fa.Ls = Ls;
  for (int s = 0; s < Ls; ++s) {
    fa.disp[s] = io->disparity[s]; fa.dw[s] = io->dw[s]; fa.dh[s] = io->dh[s];
    fa.native[s] = (io->dw[s] == W && io->dh[s] == H) ? 1 : 0;
    fa.rxd[s] = (double)(io->dw[s] - 1) / (double)(W - 1);
    fa.ryd[s] = (double)(io->dh[s] - 1) / (double)(H - 1);
    fa.rx[s] = (float)fa.rxd[s];
    fa.ry[s] = (float)fa.ryd[s];
    fa.gout[s] = fa.native[s] ? io->d_disparity[s] : (float*)(ws + L.o_gfull[s]);
    const double wsm = p->smooth_weight_mode == 0 ? (double)p->disparity_smoothness * (double)io->scale[s] : 1.0;
    fa.kx[s] = (float)(wsm / ((double)Ls * (double)(W - 1) * H * Ng));
    fa.ky[s] = (float)(wsm / ((double)Ls * (double)W * (H - 1) * Ng));
  }
  fa.auto_loss = p->automasking ? io->auto_loss : nullptr;
  fa.automask = p->automasking ? 1 : 0;
  fa.mind_d = 1.0 / p->max_depth;
  fa.dd_d = 1.0 / p->min_depth - 1.0 / p->max_depth;
  fa.mind = (float)fa.mind_d;
  fa.dd = (float)fa.dd_d;
  fa.ppd = (const PosePrepD*)(ws + L.o_ppd);
  fa.alpha = p->alpha; fa.c1 = p->ssim_c1; fa.c2 = p->ssim_c2;
  const double up = 1.0 / ((double)W * H * Ng * Ls);
  fa.up = (float)up;
  fa.pp = (const PosePrep*)(ws + L.o_pp);
  fa.stats = (const ScaleStat*)(ws + L.o_stats);
  fa.partials = (float*)(ws + L.o_partials);
  fa.vis = io->vis_warp_loss;
  fa.tiles_x = L.tiles_x; fa.tiles_y = L.tiles_y;

  const int pslot = (h->prof_ev && h->prof_next < h->prof_slots) ? h->prof_next++ : -1;
  mdb_prof_mark(h, pslot, 0, st);
  pose_prep_kernel<<<(N * S + 127) / 128, 128, 0, st>>>(pa, (PosePrep*)(ws + L.o_pp), (PosePrepD*)(ws + L.o_ppd));
  MDB_LAUNCHED(h);
  {
    dim3 grid(L.rowblocks, N);
    if (io->C == 3) disp_stats_kernel<3><<<grid, STAT_THREADS, 0, st>>>(fa, (float*)(ws + L.o_statpart));
    else disp_stats_kernel<1><<<grid, STAT_THREADS, 0, st>>>(fa, (float*)(ws + L.o_statpart));
    MDB_LAUNCHED(h);
    disp_stats_finalize_kernel<<<(Ls * N + 127) / 128, 128, 0, st>>>(fa, (const float*)(ws + L.o_statpart), L.rowblocks,
                                                                     p->normalize_disparity, (ScaleStat*)(ws + L.o_stats));
    MDB_LAUNCHED(h);
  }
  mdb_prof_mark(h, pslot, 1, st);
  int32_t rc;
  if (io->C == 3 && S == 2) rc = launch_fused<3, 2>(h, fa, L, st);
  else if (io->C == 3 && S == 1) rc = launch_fused<3, 1>(h, fa, L, st);
  else if (io->C == 1 && S == 2) rc = launch_fused<1, 2>(h, fa, L, st);
  else rc = launch_fused<1, 1>(h, fa, L, st);
  if (rc != MDB_OK) return rc;
  mdb_prof_mark(h, pslot, 2, st);

  for (int s = 0; s < Ls; ++s) {
    if (fa.native[s]) continue;
    dim3 blk(32, 8), grid((io->dw[s] + 31) / 32, (io->dh[s] + 7) / 8, N);
    upsample_bwd_kernel<<<grid, blk, 0, st>>>(fa.gout[s], W, H, N, io->dw[s], io->dh[s], fa.rx[s], fa.ry[s],
                                              io->d_disparity[s]);
    MDB_LAUNCHED(h);
  }
  mdb_prof_mark(h, pslot, 3, st);
  const int tiles = L.records;
  double* item_loss = (double*)(ws + L.o_itemloss);
  if (S == 2)
    finalize_item_kernel<2><<<N, FIN_THREADS, 0, st>>>(fa.partials, tiles, pa, nullptr, io->d_rvec[0], io->d_rvec[1],
                                                       io->d_tvec[0], io->d_tvec[1], item_loss);
  else
    finalize_item_kernel<1><<<N, FIN_THREADS, 0, st>>>(fa.partials, tiles, pa, nullptr, io->d_rvec[0], nullptr,
                                                       io->d_tvec[0], nullptr, item_loss);
  MDB_LAUNCHED(h);
  finalize_loss_kernel<<<1, 256, 0, st>>>(item_loss, N, fa.stats, Ls * N, up, io->loss);
  MDB_LAUNCHED(h);
  mdb_prof_mark(h, pslot, 4, st);
  return MDB_OK;
}

}  // extern "C"
