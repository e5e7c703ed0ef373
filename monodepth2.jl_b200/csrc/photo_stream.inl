// photo_stream.inl -- SSIM / photometric loss / min over sources, forward only (P9-P11 as stand-alone
// operators: mdb_ssim, mdb_photometric_loss, mdb_prediction_loss, mdb_automasking_loss).
// Reference: src/utils.jl:29-43, src/training.jl:1-15.
//
// Same streaming scheme as the fused kernel's stage C: one warp owns a 30-pixel-wide column strip and walks
// down the rows of its segment; the last three rows live in registers, neighbours' column statistics arrive
// by shuffle (reflect padding = a lane index choice), and the 3x3 moments are built with Chan's pairwise
// update from differences, so nothing of the size of x^2 is ever subtracted (no float32 cancellation on
// flat regions).  Every input element is loaded once per strip (+ 2 halo columns in 32, 2 halo rows per
// segment) instead of 18 times per pixel, source and channel as in the first version of this operator.
namespace {

constexpr int PS_PX = 30;          // output columns per warp
constexpr int PS_WARPS = 4;

template <int C, int S>
__global__ void __launch_bounds__(32 * PS_WARPS)
photo_stream_kernel(const PhotoArgs a, int nstrips, int nsegs, int seg_rows) {
  const int lane = threadIdx.x & 31;
  long long w = (long long)blockIdx.x * PS_WARPS + (threadIdx.x >> 5);
  if (w >= (long long)a.N * nsegs * nstrips) return;
  const int strip = (int)(w % nstrips); w /= nstrips;
  const int seg = (int)(w % nsegs);
  const int n = (int)(w / nsegs);
  const int W = a.W, H = a.H;
  const size_t plane = (size_t)W * H;
  const int gx = strip * PS_PX - 1 + lane;
  const bool colv = gx >= 0 && gx < W;
  const bool ocol = colv && lane >= 1 && lane <= PS_PX;          // a column this warp writes
  const int gxc = min(max(gx, 0), W - 1);
  // pad_reflect(1) (src/utils.jl:30-31): column -1 is column 1, column W is column W-2
  const int nl = (gx == 0) ? lane + 1 : max(lane - 1, 0);
  const int nr = (gx == W - 1) ? lane - 1 : min(lane + 1, 31);
  const int Y0 = seg * seg_rows, Y1 = min(Y0 + seg_rows, H);
  const float* yb = a.target + (size_t)n * a.tgt_nstride;
  const float* xb[S];
#pragma unroll
  for (int s = 0; s < S; ++s) xb[s] = a.pred[s] + (size_t)n * a.pred_nstride[s];

  float y1[C], y2[C], x1[S][C], x2[S][C];     // rows r-1 and r-2
#pragma unroll
  for (int c = 0; c < C; ++c) {
    y1[c] = y2[c] = 0.0f;
#pragma unroll
    for (int s = 0; s < S; ++s) x1[s][c] = x2[s][c] = 0.0f;
  }
  const float third = 1.0f / 3.0f, ninth = 1.0f / 9.0f;
  // the loads of row r+1 are issued before row r is consumed (9 registers in flight hide the DRAM latency)
  float yn[C], xn[S][C];
  auto load_row = [&](int r) {
    const int rr = r < 0 ? 1 : (r >= H ? H - 2 : r);            // reflected row
#pragma unroll
    for (int c = 0; c < C; ++c) {
      yn[c] = ldg(yb + c * plane + (size_t)rr * W + gxc);
#pragma unroll
      for (int s = 0; s < S; ++s) xn[s][c] = ldg(xb[s] + c * plane + (size_t)rr * W + gxc);
    }
  };
  load_row(Y0 - 1);
  for (int r = Y0 - 1; r <= Y1; ++r) {
    float y0[C], x0[S][C];
#pragma unroll
    for (int c = 0; c < C; ++c) {
      y0[c] = yn[c];
#pragma unroll
      for (int s = 0; s < S; ++s) x0[s][c] = xn[s][c];
    }
    if (r < Y1) load_row(r + 1);
    if (r >= Y0 + 1) {   // rows r-2, r-1, r are in registers: emit output row r-1
      float loss[S];
#pragma unroll
      for (int s = 0; s < S; ++s) loss[s] = 0.0f;
#pragma unroll
      for (int c = 0; c < C; ++c) {
        // vertical triple of this column around its middle element
        const float dy1 = y2[c] - y1[c], dy2 = y0[c] - y1[c], sy = dy1 + dy2;
        const float my = fmaf(sy, third, y1[c]);
        const float qy = fmaf(dy1, dy1, dy2 * dy2) - sy * sy * third;
        const float my_l = __shfl_sync(MDB_FULL, my, nl), my_r = __shfl_sync(MDB_FULL, my, nr);
        const float qy_l = __shfl_sync(MDB_FULL, qy, nl), qy_r = __shfl_sync(MDB_FULL, qy, nr);
        const float ely = my_l - my, ery = my_r - my, dly = (ely + ery) * third;
        const float M2y = (qy_l + qy + qy_r) + 3.0f * fmaf(ely, ely, ery * ery) - 9.0f * dly * dly;
        const float muy = my + dly, vy = M2y * ninth;
#pragma unroll
        for (int s = 0; s < S; ++s) {
          const float dx1 = x2[s][c] - x1[s][c], dx2 = x0[s][c] - x1[s][c], sx = dx1 + dx2;
          const float mx = fmaf(sx, third, x1[s][c]);
          const float qx = fmaf(dx1, dx1, dx2 * dx2) - sx * sx * third;
          const float cxy = fmaf(dx1, dy1, dx2 * dy2) - sx * sy * third;
          const float mx_l = __shfl_sync(MDB_FULL, mx, nl), mx_r = __shfl_sync(MDB_FULL, mx, nr);
          const float qx_l = __shfl_sync(MDB_FULL, qx, nl), qx_r = __shfl_sync(MDB_FULL, qx, nr);
          const float c_l = __shfl_sync(MDB_FULL, cxy, nl), c_r = __shfl_sync(MDB_FULL, cxy, nr);
          const float elx = mx_l - mx, erx = mx_r - mx, dlx = (elx + erx) * third;
          const float M2x = (qx_l + qx + qx_r) + 3.0f * fmaf(elx, elx, erx * erx) - 9.0f * dlx * dlx;
          const float Cxy = (c_l + cxy + c_r) + 3.0f * fmaf(elx, ely, erx * ery) - 9.0f * dlx * dly;
          const float mux = mx + dlx, vx = M2x * ninth, vxy = Cxy * ninth;
          // src/utils.jl:39-42
          const float n1 = fmaf(2.0f * mux, muy, a.c1), n2 = fmaf(2.0f, vxy, a.c2);
          const float d1 = fmaf(mux, mux, fmaf(muy, muy, a.c1)), d2 = vx + vy + a.c2;
          const float raw = (1.0f - (n1 * rcp_approx(d1)) * (n2 * rcp_approx(d2))) * 0.5f;
          const float sv = fminf(fmaxf(raw, 0.0f), 1.0f);
          if (a.mode == 0) {
            if (ocol) a.out[((size_t)n * C + c) * plane + (size_t)(r - 1) * W + gx] = sv;
          } else {
            loss[s] += a.alpha * sv + (1.0f - a.alpha) * fabsf(y1[c] - x1[s][c]);   // src/training.jl:3-6 (sum over c; / C below)
          }
        }
      }
      if (a.mode != 0 && ocol) {
        float best = loss[0];
#pragma unroll
        for (int s = 1; s < S; ++s) best = loss[s] < best ? loss[s] : best;        // minimum over sources, :13-15
        a.out[(size_t)n * plane + (size_t)(r - 1) * W + gx] = best * (1.0f / C);
      }
    }
#pragma unroll
    for (int c = 0; c < C; ++c) {
      y2[c] = y1[c]; y1[c] = y0[c];
#pragma unroll
      for (int s = 0; s < S; ++s) { x2[s][c] = x1[s][c]; x1[s][c] = x0[s][c]; }
    }
  }
}

template <int C, int S>
int32_t photo_stream_launch(mdb_handle* h, const PhotoArgs& a, cudaStream_t st) {
  const int nstrips = (a.W + PS_PX - 1) / PS_PX;
  // enough warps for ~16 per SM, segments of at least 8 rows (2 halo rows each)
  const long long base = (long long)a.N * nstrips;
  long long want = ((long long)16 * h->num_sms + base - 1) / base;
  const int nsegs0 = (int)std::min<long long>(std::max<long long>(want, 1), std::max(1, a.H / 8));
  const int seg_rows = (a.H + nsegs0 - 1) / nsegs0;
  const int nsegs = (a.H + seg_rows - 1) / seg_rows;
  const long long warps = base * nsegs;
  photo_stream_kernel<C, S><<<(unsigned)((warps + PS_WARPS - 1) / PS_WARPS), 32 * PS_WARPS, 0, st>>>(a, nstrips, nsegs, seg_rows);
  MDB_LAUNCHED(h);
  return MDB_OK;
}

}  // namespace
