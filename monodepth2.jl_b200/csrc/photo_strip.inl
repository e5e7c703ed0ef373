// photo_strip.inl -- the streaming "warp-strip" fused forward+backward kernel (included by
// photo_fused.cu).  Reference path: src/training.jl:42-69 with everything it calls in
// src/utils.jl (P5-P12, S2 of SURVEY 8a).
//
// One WARP owns a 28-pixel-wide column strip of one (item, scale) and streams down the rows of
// its segment.  Lane l handles sample column X0 - 2 + l: 32 sample columns feed 30 window
// columns feed 28 pixel columns, so the only redundancy is the 2+2 column halo.  Everything a
// row needs from the rows above it lives in REGISTERS (three-row rings), everything it needs
// from the columns beside it arrives by warp shuffle.  There is no block barrier and no
// shared-memory tile; the only shared memory is a thread-private three-row ring that parks the
// bilinear-sample derivatives between the forward row and its pullback two rows later.
//
//   row r arrives:  A  disparity -> depth -> project (double) -> 4 taps x C x S  -> x(r), dx/du, dx/dv
//                   C  window row r-1: 3x3 moments as a vertical triple (registers) combined
//                      across the three columns (shuffle) with Chan's update, which keeps the
//                      variances free of the float32 cancellation of E[x^2]-mu^2; SSIM + L1 per
//                      source, first-arg-min, automask, pullback coefficients of the winner
//                   D  pixel row r-2: 3x3 box sums of the coefficient fields (vertical in
//                      registers, horizontal by shuffle) -> d/d(sample) -> d/d(position) ->
//                      dP' -> d disparity (written once) and the 12 pose sums per source
//
// The two sources ride in the two halves of packed f32x2 registers (FFMA2/FADD2/FMUL2, new in
// sm_100): one issue slot does the arithmetic of both sources.
#include <type_traits>
namespace {

constexpr int STRIP_PX = 28;
#ifndef STRIP_NWARPS
#define STRIP_NWARPS 1   // one warp per CTA: occupancy is then set by the register cap alone
#endif
constexpr int STRIP_WARPS = STRIP_NWARPS;   // warps per CTA (independent of each other)
constexpr int STRIP_THREADS = 32 * STRIP_WARPS;
constexpr int SEL_NONE = 255;
constexpr int STRIP_NHEAD = 3;   // record head: photometric sum, smoothness x-sum, smoothness y-sum
#ifndef STRIP_CBASE
#define STRIP_CBASE 1   // one 64-bit base per channel plane: two more register pairs, one IMAD.WIDE per tap address (4.51 -> 4.47 ms)
#endif
#ifndef STRIP_AM_EARLY
#define STRIP_AM_EARLY 1   // fetch the automask value with the row's tap loads (stage A): ptxas sinks a load placed in stage C to 28-80
                           // instructions before its use and the warp eats the latency every row (config 3: 0.939 -> 0.879 ms)
#endif
#ifndef STRIP_DPF
#define STRIP_DPF 1     // low-res disparity: the next native row is fetched one step ahead (its use no longer waits)
#endif
#ifndef STRIP_CONSTPOSE
#define STRIP_CONSTPOSE 1   // the per-(item, source) projection constants are served from the constant bank: they reach the
                            // FP64 pipe as uniform-register operands instead of occupying ~36 vector registers per thread.
                            // With STRIP_PEEL this is what makes the 160-register build spill-free (4.33 -> 4.18 ms at config 5;
                            // every spilling build of the peeled loop is SLOWER than the unpeeled one, profiles/r2_strip_summary.md)
#endif
#ifndef STRIP_HOSTK
#define STRIP_HOSTK 0
#endif
#ifndef STRIP_FENCE
#define STRIP_FENCE 0   // experiment: scheduling fences between the stages of the steady-state body (1: after A, 2: after A and C)
#endif
#ifndef STRIP_PEEL
#define STRIP_PEEL 1   // the row loop body exists twice: a generic copy for the <= 6 rows at a segment's / the image's edges, and
                       // a steady-state copy in which every row-position test is a compile-time constant (no reflection, no
                       // clamps, no stage gating): fewer instructions and one long basic block for ptxas to schedule
#endif
// ---- switches of recorded experiments (profiles/r2_strip_summary.md); the defaults are the shipped kernel -------------
#ifndef STRIP_TAPER_FIELDS
#define STRIP_TAPER_FIELDS 0
#endif
#ifndef STRIP_TAPER
#define STRIP_TAPER 1      // experiment 20, shipped: tapered row segments (host side: strip_geom)
#endif
#ifndef STRIP_UNROLL
#define STRIP_UNROLL 1     // experiment 19: 2 or 3 steady-state rows per loop trip
#endif
#ifndef STRIP_FAKE_RGBX
#define STRIP_FAKE_RGBX 0  // experiment 18 (timing only, wrong results)
#endif
#ifndef STRIP_ROWBAR
#define STRIP_ROWBAR 0     // experiment 17: named barrier per row across the CTA's warps
#endif
#ifndef STRIP_TMA_TGT
#define STRIP_TMA_TGT 0    // experiment 15: target rows by cp.async.bulk into a shared-memory ring
#endif
#ifndef STRIP_PEEL_CAPTURED
#define STRIP_PEEL_CAPTURED 0
#endif
#ifndef STRIP_PF
#define STRIP_PF 0         // experiment 22: prefetch of the source row the NEXT target row's bottom taps will read (1: into L1, 2: into
                           // L2; +4: the next target row as well), steady-state rows only
#endif
#ifndef STRIP_ROTATE
#define STRIP_ROTATE 0     // experiment 23: the row loop rotated by half a row -- the NEXT row's loads are issued between stages C and D, so
                           // that stage D and the next row's smoothness block run in their shadow (today: the smoothness block alone)
#endif
#ifndef STRIP_SMEM_RING
#define STRIP_SMEM_RING 0  // experiment 21: the two-row rings of the warped samples and of the pullback coefficients (30 registers that
                           // only cross rows) parked in thread-private shared memory, to run more warps per SM at a lower register cap
#endif
#ifndef STRIP_MAXREG_CAPTURED
#define STRIP_MAXREG_CAPTURED 216   // spill-free; config 1 per iteration: 152 (24 B spilled) 23.95 us, 184 (16 B) 24.18, 216 (0 B) 22.58;
                                    // peeled copy at 160 / 200: 23.94 / 24.62.  Occupancy is no concern at 480 warps.
#endif
#if !defined(STRIP_MAXREG) && !defined(STRIP_MINB)
#define STRIP_MAXREG 160   // 12 warps per SM, no spills.  Round 1 (unpeeled loop, constants in registers): best of 128 .. 184 was
                           // 152.  Round 2 (peeled loop, constants in the bank): 152 spills 148 B (4.83 ms), 160 and 168 are
                           // spill-free (4.18 / 4.20 ms), 176 .. 224 spill again (ptxas trades) -- spills decide, not the 13th warp
#endif
#ifndef STRIP_MINB
#define STRIP_MINB 4   // CTAs per SM the register allocation aims for (128 registers; measured best of 2/3/4)
#endif

#if STRIP_CONSTPOSE
// A launch covers at most STRIP_CONST_CHUNK (item, source) pairs and reads them from one of STRIP_CONST_POOL regions of
// the bank; the host side (pose_pool, photo_fused_host.inl) hands the regions out round-robin and orders their reuse
// with events, so calls on different streams of one device never see each other's constants.
constexpr int STRIP_CONST_CHUNK = 128, STRIP_CONST_POOL = 3;
__constant__ PoseStrip c_pose[STRIP_CONST_POOL * STRIP_CONST_CHUNK];
#endif

// ---- S floats packed: S = 2 uses the sm_100 f32x2 pipe ----------------------------------
template <int S> struct Pk;
template <> struct Pk<1> {
  float v;
  MDB_DEV Pk() {}
  MDB_DEV explicit Pk(float f) : v(f) {}
  MDB_DEV float get(int) const { return v; }
  MDB_DEV void set(int, float f) { v = f; }
  friend MDB_DEV Pk operator+(Pk a, Pk b) { return Pk(a.v + b.v); }
  friend MDB_DEV Pk operator-(Pk a, Pk b) { return Pk(a.v - b.v); }
  friend MDB_DEV Pk operator*(Pk a, Pk b) { return Pk(a.v * b.v); }
  friend MDB_DEV Pk pfma(Pk a, Pk b, Pk c) { return Pk(fmaf(a.v, b.v, c.v)); }
  MDB_DEV Pk shfl(int src) const { return Pk(__shfl_sync(MDB_FULL, v, src)); }
};
template <> struct Pk<2> {
  float2 v;
  MDB_DEV Pk() {}
  MDB_DEV explicit Pk(float f) : v(make_float2(f, f)) {}
  MDB_DEV Pk(float a, float b) : v(make_float2(a, b)) {}
  MDB_DEV explicit Pk(float2 f) : v(f) {}
  MDB_DEV float get(int s) const { return s == 0 ? v.x : v.y; }
  MDB_DEV void set(int s, float f) { if (s == 0) v.x = f; else v.y = f; }
  friend MDB_DEV Pk operator+(Pk a, Pk b) { return Pk(__fadd2_rn(a.v, b.v)); }
  friend MDB_DEV Pk operator-(Pk a, Pk b) { return Pk(__fadd2_rn(a.v, make_float2(-b.v.x, -b.v.y))); }
  friend MDB_DEV Pk operator*(Pk a, Pk b) { return Pk(__fmul2_rn(a.v, b.v)); }
  friend MDB_DEV Pk pfma(Pk a, Pk b, Pk c) { return Pk(__ffma2_rn(a.v, b.v, c.v)); }
  MDB_DEV Pk shfl(int src) const {
    return Pk(__shfl_sync(MDB_FULL, v.x, src), __shfl_sync(MDB_FULL, v.y, src));
  }
};
// sign(v) with sign(0) = 0, without an int->float conversion
MDB_DEV float sgn1(float v) { return v == 0.0f ? 0.0f : copysignf(1.0f, v); }

// frame samples: Float32 as stored, or N0f8 bytes carried as 2^23 + u until first use (see the kernel's U8 parameter)
MDB_DEV float ldx(const float* p) { return ldg(p); }
MDB_DEV float ldx(const unsigned char* p) { return __uint_as_float(0x4B000000u | (unsigned)__ldg(p)); }
MDB_DEV void pf_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }
MDB_DEV void pf_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
constexpr float N0F8_R1 = 0.003921568859368562698f;      // fl(1 / 255)
constexpr float N0F8_R2 = -2.3191758e-10f;               // fl(1 / 255 - r1)
template <bool U8> MDB_DEV float cvx(float g) {
  if (!U8) return g;
  const float f = g - 8388608.0f;
  return fmaf(f, N0F8_R1, f * N0F8_R2);
}
template <bool U8, int S> MDB_DEV Pk<S> cvx(Pk<S> g) {
  if (!U8) return g;
  const Pk<S> f = g + Pk<S>(-8388608.0f);
  return pfma(f, Pk<S>(N0F8_R1), f * Pk<S>(N0F8_R2));
}

template <int S> MDB_DEV Pk<S> pfma(Pk<S> a, float b, Pk<S> c) { return pfma(a, Pk<S>(b), c); }
template <int S> MDB_DEV Pk<S> operator*(Pk<S> a, float b) { return a * Pk<S>(b); }
template <int S> MDB_DEV Pk<S> operator+(Pk<S> a, float b) { return a + Pk<S>(b); }

// The image row that ring row r stands for: the rows just outside the image are the reflected rows
// 1 and H-2 (pad_reflect(1) of the WARPED image, src/utils.jl:30-31 -- the warp of target row 1 is computed
// a second time instead of being copied), everything further out is never read.
MDB_DEV int geo_row(int r, int H) { return r == -1 ? 1 : (r == H ? H - 2 : r); }

constexpr int STRIP_MAX_TAPER = 8;
struct StripGeom {
  int nstrips, nsegs, seg_rows, total;  // total = items * nsegs * nstrips * Ls warps of THIS launch
  int n0;                               // first item of this launch (a launch covers a chunk of the batch's items)
  int cbase;                            // first entry of this launch's region of c_pose
#if STRIP_TAPER || STRIP_TAPER_FIELDS
  int items;                            // items of this launch
  // experiment 20 (test hook "strip_taper"): row segments of decreasing length, the long ones first in launch order, so that a
  // launch of only ~2 resident waves ends evenly; seg_y[] holds the boundaries.  0 = equal segments, item-major order.
  int taper;
  int seg_y[STRIP_MAX_TAPER + 1];
#endif
};

// Decision dump (DBG instantiation only; tests/test_gpu_forced.py): every discrete choice the pullback
// depends on, one int8 plane each, layout [scale][item][plane][H][W] with 2 C + 2 planes:
//   0            winner of the window centred on the pixel: source index, -1 = the automask took it
//   1 .. C       the winner's SSIM clamp gate per channel (1: inside [0,1], gradient passes)
//   C+1 .. 2C    sign(x - y) of the winner's sample per channel (the L1 pullback)
//   2C+1         smoothness signs: (sign(d - d_right) + 1) + 3 (sign(d - d_down) + 1)
// The float64 oracle can be run with these forced (oracle/c_port mdbo_set_decisions), which separates
// arithmetic error from sub-ulp ties that legitimately fall the other way in float32.
MDB_DEV signed char sgn_code(float v) { return v == 0.0f ? 0 : (v > 0.0f ? 1 : -1); }

// CP: the projection constants come from the constant bank (c_pose, filled by the host before the launch) instead of
// global memory.  The throughput path uses CP = true; a call made under stream capture (slow_depth's graph: one item,
// latency-bound) uses CP = false, which needs no copy node and no event nodes in the graph.
// The captured instantiation (CP = false, DBG = false) keeps round 1's loop (no steady-state copy), its per-row constants in
// registers, and a register cap under which nothing spills (STRIP_MAXREG_CAPTURED): 22.6 us per iteration at config 1, where
// the peeled loop with constants re-loaded through a.ppd took 24.2 - 26.0 us.  It also serves upsample_align_corners = 0.
// U8: the frame stack x holds 8-bit samples (N0f8, same (W,H,C,L,N) layout, 1 byte each) instead of Float32: a quarter of the
// image bytes in HBM and in the gathers.  A loaded byte u travels as the float 2^23 + u (one LOP3 on the load's result), and
// where the value is first used it becomes Float32(::N0f8) = u / 255 correctly rounded: f = (2^23 + u) - 2^23 (exact),
// then fma(f, r1, f r2) with r1 = fl(1/255), r2 = fl(1/255 - r1) -- equal to the IEEE quotient for all 256 values (checked
// exhaustively in tests/test_ingest.py), so the u8 path is bit-identical to the float path fed the converted stack.
// (u * (1/255f) alone is NOT: it differs from u / 255f for 126 of the 256 values.)  Throughput instantiation only.
template <int C, int S, bool DBG = false, bool CP = (STRIP_CONSTPOSE != 0), bool U8 = false>
#ifdef STRIP_MAXREG
__global__ void __maxnreg__((CP || DBG) ? STRIP_MAXREG : STRIP_MAXREG_CAPTURED)
#else
__global__ void __launch_bounds__(STRIP_THREADS, STRIP_MINB)
#endif
photo_strip_kernel(const __grid_constant__ FusedArgs a, const StripGeom g) {
  using P = Pk<S>;
  constexpr int NPART = STRIP_NHEAD + NPOSE * S;
  constexpr int DRV = S * (2 * C + 2);  // floats parked per (thread, row): dx/du, dx/dv per channel, u, v
  constexpr int DRV4 = DRV / 4;
  extern __shared__ float4 s_drv[];     // [3 rows][DRV4][STRIP_THREADS]
  constexpr bool SR = (STRIP_SMEM_RING != 0) && CP && !DBG;
  constexpr int NX4 = (C * S + 3) / 4, NF4 = (3 * C + 3) / 4;
  float4* const s_rx = s_drv + 3 * DRV4 * STRIP_THREADS;   // SR: [2 rows][NX4][STRIP_THREADS] warped samples
  float4* const s_rf = s_rx + 2 * NX4 * STRIP_THREADS;     // SR: [2 rows][NF4][STRIP_THREADS] pullback coefficients

  // the shuffle makes the warp index provably warp-uniform, so every branch on the strip geometry
  // below is uniform for the compiler and the row shuffles need no WARPSYNC
  const int lane = threadIdx.x & 31, warp = __shfl_sync(MDB_FULL, (int)(threadIdx.x >> 5), 0);
  int w = blockIdx.x * STRIP_WARPS + warp;
  if (w >= g.total) return;
  const int w_lin = w;
  const int sc = w % a.Ls; w /= a.Ls;
  const int strip = w % g.nstrips; w /= g.nstrips;
  // launch order: scales, strips, then segments inside an item -- or, tapered, items inside a segment (long segments first)
#if STRIP_TAPER
  // ONE runtime division either way (a second one costs the 480-warp launch of config 1 two microseconds of prologue latency)
  const int dv = g.taper ? g.items : g.nsegs, qd = w / dv, rd = w - qd * dv;
  const int seg = g.taper ? qd : rd;
  const int nl_item = g.taper ? rd : qd;        // item inside this launch's chunk
#else
  const int seg = w % g.nsegs, nl_item = w / g.nsegs;
#endif
  const int n = g.n0 + nl_item;
  // record index in the whole batch, layout [item][seg][strip][scale]: the launch order itself unless the segments are tapered
#if STRIP_TAPER
  const int rec = g.taper ? ((n * g.nsegs + seg) * g.nstrips + strip) * a.Ls + sc : g.n0 * (g.nsegs * g.nstrips * a.Ls) + w_lin;
#else
  const int rec = g.n0 * (g.nsegs * g.nstrips * a.Ls) + w_lin;
#endif

  const int W = a.W, H = a.H;
  const size_t plane = (size_t)W * H;
  const int planei = W * H;   // io_shape_ok guarantees C * W * H < 2^31: plane offsets are 32-bit
  const int X0 = strip * STRIP_PX;
#if STRIP_TAPER
  const int Y0 = g.taper ? g.seg_y[seg] : seg * g.seg_rows, Y1 = g.taper ? g.seg_y[seg + 1] : min(Y0 + g.seg_rows, H);
#else
  const int Y0 = seg * g.seg_rows, Y1 = min(Y0 + g.seg_rows, H);
#endif
  const int gx = X0 - 2 + lane;
  const bool colv = gx >= 0 && gx < W;                          // a real sample column
  const bool wcol = colv && lane >= 1 && lane <= 30;            // a window column of this strip
  const bool pcol = colv && lane >= 2 && lane <= 29;            // a pixel column this warp owns
  // neighbour lanes of a window column, with pad_reflect(1) folded in (src/utils.jl:30-31)
  const int nl = (gx == 0) ? lane + 1 : max(lane - 1, 0);
  const int nr = (gx == W - 1) ? lane - 1 : min(lane + 1, 31);
  // multiplicity of this pixel inside its left / right neighbour windows (reflected twice at the border)
  const float wl = (gx == 1) ? 2.0f : 1.0f, wr_ = (gx == W - 2) ? 2.0f : 1.0f;

  // one 64-bit base per item; everything inside the item is addressed with 32-bit element offsets
  // (io_shape_ok guarantees L * C * W * H < 2^31)
  using XT = typename std::conditional<U8, unsigned char, float>::type;
  const XT* xb;
  {
    // opaque to the optimiser: otherwise it re-derives this base from n in 64-bit arithmetic at every load
    unsigned long long xbi = (unsigned long long)(reinterpret_cast<const XT*>(a.x) + (size_t)n * a.L * C * plane);
    asm volatile("" : "+l"(xbi));
    xb = reinterpret_cast<const XT*>(xbi);
  }
#if STRIP_CBASE
  const XT* xbc[C];   // per-channel bases (opaque, like xb)
#pragma unroll
  for (int c = 0; c < C; ++c) {
    unsigned long long t = (unsigned long long)(xb + (size_t)c * plane);
    asm volatile("" : "+l"(t));
    xbc[c] = reinterpret_cast<const XT*>(t);
  }
#endif
  const unsigned tgt_off = a.tgt * C * planei;
#if STRIP_TMA_TGT
  // ---- experiment: the target rows (the one REGULAR image operand: 3 x 128 B per row, the same for every lane pattern) arrive
  // by TMA bulk copies (cp.async.bulk, 144 B = 36 floats per channel from a 16-byte aligned column 2 left of lane 0) into a
  // 4-row shared-memory ring, two rows ahead of their use, completion on one mbarrier per ring slot.  Interior strips only
  // (a bulk copy cannot clamp at the image border); edge strips keep the register path.
  static_assert(STRIP_WARPS == 1, "the TMA experiment assumes one-warp CTAs");
  constexpr int TMA_SLOTS = 4, TMA_ROWF = 36;
  __shared__ __align__(16) float s_tgt[TMA_SLOTS][C][TMA_ROWF];
  __shared__ __align__(8) unsigned long long s_bar[TMA_SLOTS];
  const bool tma_ok = !U8 && strip >= 1 && X0 + 32 <= W && (W & 3) == 0 && ((reinterpret_cast<uintptr_t>(a.x) & 15) == 0);
  const unsigned bar0 = (unsigned)__cvta_generic_to_shared(&s_bar[0]);
  const unsigned tg0 = (unsigned)__cvta_generic_to_shared(&s_tgt[0][0][0]);
  if (tma_ok) {
    if (lane == 0) {
#pragma unroll
      for (int i = 0; i < TMA_SLOTS; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0 + 8 * i));
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
  }
  auto tma_issue = [&](int rr) {   // target row of ring row rr into slot (rr - (Y0 - 2)) & 3
    __syncwarp();                  // every lane has read the slot's previous contents (row rr - 4)
    if (lane == 0) {
      const int rc2 = min(max(geo_row(rr, H), 0), H - 1);
      const unsigned sl = (unsigned)(rr - (Y0 - 2)) & 3u, bar = bar0 + 8 * sl;
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((unsigned)(C * TMA_ROWF * 4)) : "memory");
#pragma unroll
      for (int c = 0; c < C; ++c) {
        const XT* src = xb + (unsigned)(tgt_off + c * planei + rc2 * W + (X0 - 4));
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(tg0 + (unsigned)((sl * C + c) * TMA_ROWF * 4)), "l"(src), "r"((unsigned)(TMA_ROWF * 4)), "r"(bar) : "memory");
      }
    }
  };
#endif
  unsigned src_off[S];
#pragma unroll
  for (int s = 0; s < S; ++s) src_off[s] = a.src[s] * C * planei;
  const float* dsp = a.disp[sc] + (size_t)n * a.dw[sc] * a.dh[sc];
  const float* amb = a.automask ? a.auto_loss + (size_t)n * plane : nullptr;
  float* gout = a.gout[sc] + (size_t)n * plane;
  const bool native = a.native[sc] != 0;
  const int dw = a.dw[sc], dh = a.dh[sc];
  const double ryd = a.ryd[sc];
  signed char* dbg = nullptr;
  if (DBG) dbg = a.dbg + ((size_t)sc * a.N + n) * (2 * C + 2) * plane;

  // ---- per-lane projection state: P' = A q + invd b, q = [px, py, 1], A = I + dA -------------
  // p_i = A_i . q is carried down the rows by adding column 1 of A (double: no drift that matters)
  double p1[S], p2[S], p3[S];
#if STRIP_CONSTPOSE
  // constant bank, warp-uniform index: operands, not registers
  const PoseStrip* const ppn = CP ? c_pose + g.cbase + nl_item * S : a.ppd + (size_t)n * S;
#else
  const PoseStrip* const ppn = a.ppd + (size_t)n * S;
#endif
  // CP = false: the per-row constants are copied into registers once (through a.ppd the compiler would have to re-load
  // them every row: the gradient stores may alias); CP = true: they stay constant-bank operands and `pregs` vanishes
  struct PoseRegs { double b0, b1, b3e, a1, i2, a7; float b0f, b1f, b3f; };
  PoseRegs pregs[S];
  if (!CP) {
#pragma unroll
    for (int s = 0; s < S; ++s) {
      const PoseStrip& pp = ppn[s];
      pregs[s].b0 = pp.b[0]; pregs[s].b1 = pp.b[1]; pregs[s].b3e = pp.b3e; pregs[s].a1 = pp.dA[1]; pregs[s].i2 = pp.i2; pregs[s].a7 = pp.dA[7];
      pregs[s].b0f = pp.b0f; pregs[s].b1f = pp.b1f; pregs[s].b3f = pp.b3f;
    }
  }
#define PPV(s, bank, reg) (CP ? ppn[s].bank : pregs[s].reg)
  {
    const double px = (double)(gx + 1), py = (double)(geo_row(Y0 - 2, H) + 1);
#pragma unroll
    for (int s = 0; s < S; ++s) {
      const PoseStrip& pp = ppn[s];
      p1[s] = fma(1.0 + pp.dA[0], px, fma(pp.dA[1], py, pp.dA[2]));
      p2[s] = fma(pp.dA[3], px, fma(pp.i2, py, pp.dA[5]));
      p3[s] = fma(pp.dA[6], px, fma(pp.dA[7], py, 1.0 + pp.dA[8]));
    }
  }
  // bilinear up-sampling of the disparity (align-corners): the column part is per lane
#if STRIP_DPF
  int ux0 = 0, ux1 = 0, uy0 = -2;          // uy0: native row whose column-interpolated values are cached (-2: none)
  double ulx = 0.0, utop = 0.0, ubot = 0.0;
  float nv0 = 0.0f, nv1 = 0.0f;            // native row uy0 + 2 at columns ux0, ux1, fetched one step ahead
#else
  int ux0 = 0, ux1 = 0, uy0 = -1;          // uy0: native row whose column-interpolated values are cached
  double ulx = 0.0, utop = 0.0, udif = 0.0;
#endif
  if (!native) {
    double f = (double)min(max(gx, 0), W - 1) * a.rxd[sc];
    if (!CP && a.hp) f = fmax(f + a.oxd[sc], 0.0);   // half-pixel centres: max(0, (p + 0.5) dw / W - 0.5)
    ux0 = min((int)f, dw - 1);
    ulx = f - (double)ux0;
    ux1 = min(ux0 + 1, dw - 1);
  }

#if STRIP_HOSTK
  // experiment: the four factors formed on the host and read as constant-bank operands
  const float ksu = a.ksu, klu = a.klu, kl = a.kl, ka = a.ka;
#else
  const float ksu = a.alpha / (9.0f * C) * a.up;   // SSIM pullback factor alpha / C / 9, times 1/(W H N Ls)
  const float klu = (1.0f - a.alpha) / C * a.up;   // L1 factor
  const float kl = (1.0f - a.alpha) / C, ka = a.alpha / C;
#endif
  const ScaleStat st = a.stats[(size_t)sc * a.N + n];
  const float skx = st.inv_mp * a.kx[sc], sky = st.inv_mp * a.ky[sc];

  // ---- rings (registers) ---------------------------------------------------------------------
  P xr1[C], xr2[C];            // warped samples of rows r-1, r-2 (both sources packed)
  float yr1[C], yr2[C];        // target rows r-1, r-2
  float F1[3 * C], F2[3 * C];  // winner's pullback coefficients of window rows r-2 (F1), r-3 (F2)
  int sel1 = SEL_NONE, sel2 = SEL_NONE;
  float d1 = 0.0f, d2 = 0.0f, d3 = 0.0f;  // disparity rows r-1, r-2, r-3
  float wy_dn = 0.0f;                      // smoothness weight between rows r-2 and r-1
#pragma unroll
  for (int c = 0; c < C; ++c) { xr1[c] = P(0.0f); xr2[c] = P(0.0f); yr1[c] = yr2[c] = 0.0f; }
#pragma unroll
  for (int k = 0; k < 3 * C; ++k) F1[k] = F2[k] = 0.0f;

  float loss_acc = 0.0f, smx_acc = 0.0f, smy_acc = 0.0f;
  P S0[3], SP[3], SI[3];  // sum dP_i, sum dP_i py, sum dP_i invd  (times px at the end for column 0)
#pragma unroll
  for (int i = 0; i < 3; ++i) { S0[i] = P(0.0f); SP[i] = P(0.0f); SI[i] = P(0.0f); }

  int slot = 0;  // ring slot of row r in s_drv
  int par = 0;   // SR: slot of row r (and, until it is overwritten, of row r-2) in s_rx / s_rf
  if (SR) {
#pragma unroll
    for (int k = 0; k < 2 * NX4; ++k) s_rx[k * STRIP_THREADS + threadIdx.x] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
#pragma unroll
    for (int k = 0; k < 2 * NF4; ++k) s_rf[k * STRIP_THREADS + threadIdx.x] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
  }

  // ---- stage A in two halves: issue() starts the row's loads, consume() runs after the smoothness block ----
  // in flight between issue() and consume(): 4 taps x C channels (sources packed), the target row,
  // the bilinear weights and masks
  P tv00[C], tv01[C], tv10[C], tv11[C];
  float ty_[C], td_ = 0.0f;
#if STRIP_AM_EARLY
  float tam = 0.0f;
#endif
  P tfx, tfy, tmzx, tmzy, tuu, tvv;
  float dq = 0.0f;   // native disparity of the NEXT row to be issued, loaded one call ahead
  const int gxc = min(max(gx, 0), W - 1);   // loads of dead lanes / rows go to a clamped (valid) address
  if (native) dq = ldg(dsp + (min(max(geo_row(Y0 - 2, H), 0), H - 1) * W + gxc));

  auto issue = [&](int r, auto steady) {
    constexpr bool ST = decltype(steady)::value;   // steady-state row: 0 <= r, r + 1 <= H - 1, nothing reflected or clamped
    const int rg = ST ? r : geo_row(r, H), rgn = ST ? r + 1 : geo_row(r + 1, H);   // warp-uniform
    const int rc = ST ? r : min(max(rg, 0), H - 1);
    {   // unconditional (clamped addresses): nothing in flight is carried from one row to the next
      double dd_;
      if (native) {
        dd_ = (double)dq;
      } else {
        // rows of the low-res map change every 1/ry full-res rows: keep the column-interpolated pair
        double fy = (double)rc * ryd;
        if (!CP && a.hp) fy = fmax(fy + a.oyd[sc], 0.0);
        const int y0 = min((int)fy, dh - 1);
#if STRIP_DPF
        if (y0 != uy0) {   // warp-uniform
          if (y0 == uy0 + 1) {
            // one native row further down (the usual step): the new bottom row was fetched when the
            // previous step happened, >= 1 full-res row ago -- nothing to wait for
            utop = ubot;
            ubot = fma(ulx, (double)nv1 - (double)nv0, (double)nv0);
          } else {   // first row of the segment, or a reflected row stepping back
            const int y1 = min(y0 + 1, dh - 1);
            const double v00 = ldg(dsp + (y0 * dw + ux0)), v01 = ldg(dsp + (y0 * dw + ux1));
            const double v10 = ldg(dsp + (y1 * dw + ux0)), v11 = ldg(dsp + (y1 * dw + ux1));
            utop = fma(ulx, v01 - v00, v00);
            ubot = fma(ulx, v11 - v10, v10);
          }
          uy0 = y0;
          const int y2 = min(y0 + 2, dh - 1);   // bottom row of the next step
          nv0 = ldg(dsp + (y2 * dw + ux0)); nv1 = ldg(dsp + (y2 * dw + ux1));
        }
        dd_ = fma(fy - (double)y0, ubot - utop, utop);
#else
        if (y0 != uy0) {   // warp-uniform
          uy0 = y0;
          const int y1 = min(y0 + 1, dh - 1);
          const double v00 = ldg(dsp + (y0 * dw + ux0)), v01 = ldg(dsp + (y0 * dw + ux1));
          const double v10 = ldg(dsp + (y1 * dw + ux0)), v11 = ldg(dsp + (y1 * dw + ux1));
          utop = fma(ulx, v01 - v00, v00);
          udif = fma(ulx, v11 - v10, v10) - utop;
        }
        dd_ = fma(fy - (double)y0, udif, utop);
#endif
      }
      td_ = (float)dd_;
      const double invd = fma(dd_, a.dd_d, a.mind_d);   // 1 / depth (src/utils.jl:179-183)
#if STRIP_TMA_TGT
      if (tma_ok) {
        if (r + 2 <= Y1 + 1) tma_issue(r + 2);
      } else
#endif
#pragma unroll
      for (int c = 0; c < C; ++c) ty_[c] = ldx(xb + (unsigned)(tgt_off + c * planei + rc * W + gxc));
#if STRIP_PF & 4
      if (ST) {
#pragma unroll
        for (int c = 0; c < C; ++c) {
          if ((STRIP_PF & 3) == 1) pf_l1(xb + (unsigned)(tgt_off + c * planei + (rc + 1) * W + gxc));
          else pf_l2(xb + (unsigned)(tgt_off + c * planei + (rc + 1) * W + gxc));
        }
      }
#endif
#if STRIP_AM_EARLY
      if (a.automask) tam = ldg(amb + (unsigned)((ST ? r - 1 : min(max(r - 1, 0), H - 1)) * W + gxc));   // window row r-1, used in stage C
#endif

      unsigned ot[S];   // top-left tap; the other three are +1, +W, +W+1 (x0 <= W-2, y0 <= H-2 always)
#if STRIP_PF & 3
      unsigned opf[S];
#endif
#pragma unroll
      for (int s = 0; s < S; ++s) {
        // Backproject -> Project -> normalize -> grid_sample's unnormalize (src/utils.jl:67-69,
        // 99-103; src/training.jl:56) cancel to u = P'1 / z'', z'' = P'3 + invd 1e-7 (in units of depth)
        const double P1 = fma(invd, PPV(s, b[0], b0), p1[s]), P2 = fma(invd, PPV(s, b[1], b1), p2[s]), z = fma(invd, PPV(s, b3e, b3e), p3[s]);
        double zi;
        asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(zi) : "d"(z));   // MUFU.RCP64H: ~2^-20, squared by the Newton step
        zi = fma(zi, fma(-z, zi, 1.0), zi);
        const double X = fma(P1, zi, -1.0), Y = fma(P2, zi, -1.0);   // 0-based source coordinates
        // floor with a saturating convert (far-outside coordinates saturate and clamp to the border below)
        int ix = __double2int_rd(X), iy = __double2int_rd(Y);
        float fxs = (float)(X - (double)ix), fys = (float)(Y - (double)iy);
        // NNlib grid_sample(padding_mode=:border), align-corners: clamp to [0, W-1]; no gradient where
        // clamped (position <= 0 or >= W-1).  The right border is the cell W-2 with weight 1.
        const bool rx = (unsigned)ix <= (unsigned)(W - 2), ry = (unsigned)iy <= (unsigned)(H - 2);   // inside the cell range
        const bool inx = rx && ((ix | __float_as_int(fxs)) != 0), iny = ry && ((iy | __float_as_int(fys)) != 0);
        if (!rx) fxs = ix < 0 ? 0.0f : 1.0f;
        if (!ry) fys = iy < 0 ? 0.0f : 1.0f;
        ix = min(max(ix, 0), W - 2);
        iy = min(max(iy, 0), H - 2);
#if STRIP_PF & 3
        if (ST) opf[s] = (unsigned)(min(iy + 2, H - 1) * W + ix) + src_off[s];   // iy already clamped to [0, H-2]
#endif
        ot[s] = (unsigned)(iy * W + ix) + src_off[s];
        const float zf = (float)zi;
        tfx.set(s, fxs); tfy.set(s, fys);
        tmzx.set(s, inx ? zf : 0.0f); tmzy.set(s, iny ? zf : 0.0f);
        // u, v only multiply dP'1, dP'2, which vanish where clamped: the clamped position is good enough
        tuu.set(s, (float)ix + (fxs + 1.0f)); tvv.set(s, (float)iy + (fys + 1.0f));
      }
#if STRIP_CBASE
      unsigned ob[S];
#pragma unroll
      for (int s = 0; s < S; ++s) ob[s] = ot[s] + (unsigned)W;
#pragma unroll
      for (int c = 0; c < C; ++c) {
#pragma unroll
        for (int s = 0; s < S; ++s) {
#if STRIP_FAKE_RGBX
          // TIMING ONLY (wrong results): channels 1, 2 reuse channel 0's taps -- the loads and address arithmetic an
          // interleaved RGBX source layout would save, without its wider registers: an upper bound of that layout's gain
          if (c > 0) { tv00[c] = tv00[0]; tv01[c] = tv01[0]; tv10[c] = tv10[0]; tv11[c] = tv11[0]; continue; }
#endif
          const XT* p = xbc[c] + ot[s];    // one IMAD.WIDE per address: the channel plane is in the base
          const XT* pb = xbc[c] + ob[s];
          tv00[c].set(s, ldx(p)); tv01[c].set(s, ldx(p + 1));
          tv10[c].set(s, ldx(pb)); tv11[c].set(s, ldx(pb + 1));
#if STRIP_PF & 3
          if (ST) {
            if ((STRIP_PF & 3) == 1) pf_l1(xbc[c] + opf[s]); else pf_l2(xbc[c] + opf[s]);
          }
#endif
        }
      }
#else
#pragma unroll
      for (int c = 0; c < C; ++c) {
#pragma unroll
        for (int s = 0; s < S; ++s) {
          const XT* p = xb + (ot[s] + (unsigned)(c * planei));
          const XT* pb = xb + (ot[s] + (unsigned)(c * planei + W));
          tv00[c].set(s, ldx(p)); tv01[c].set(s, ldx(p + 1));
          tv10[c].set(s, ldx(pb)); tv11[c].set(s, ldx(pb + 1));
        }
      }
#endif
    }
    if (native) dq = ldg(dsp + ((ST ? rgn : min(max(rgn, 0), H - 1)) * W + gxc));
    if (ST) {
#pragma unroll
      for (int s = 0; s < S; ++s) { p1[s] += PPV(s, dA[1], a1); p2[s] += PPV(s, i2, i2); p3[s] += PPV(s, dA[7], a7); }
    } else {
      const double stp = (double)(rgn - rg);   // 1 except around the two reflected rows
#pragma unroll
      for (int s = 0; s < S; ++s) { p1[s] = fma(stp, PPV(s, dA[1], a1), p1[s]); p2[s] = fma(stp, PPV(s, i2, i2), p2[s]); p3[s] = fma(stp, PPV(s, dA[7], a7), p3[s]); }
    }
  };

  // bilinear value and its derivatives of the issued row -> xn, yn, dn; park the derivatives in `wslot`
  auto consume = [&](int r, int wslot, P (&xn)[C], float (&yn)[C], float& dn) {
    {
      // rows outside the image and columns outside it carry the (finite) values of the clamped addresses:
      // nothing reads them unmasked -- window rows reflect (stage C), window columns reflect (nl, nr), the
      // smoothness weights vanish across the border, and columns that are not window columns select no source
      dn = td_;
#if STRIP_TMA_TGT
      if (tma_ok) {
        const unsigned k = (unsigned)(r - (Y0 - 2)), sl = k & 3u, par = (k >> 2) & 1u, bar = bar0 + 8 * sl;
        unsigned done = 0;
        int spins = 0;
        do {
          asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                       : "=r"(done) : "r"(bar), "r"(par) : "memory");
        } while (!done && ++spins < (1 << 22));
        if (!done) __trap();   // a copy that never lands must not hang the GPU
#pragma unroll
        for (int c = 0; c < C; ++c) ty_[c] = s_tgt[sl][c][lane + 2];
      }
#else
      (void)r;
#endif
      float park[DRV];
#pragma unroll
      for (int c = 0; c < C; ++c) {
        yn[c] = cvx<U8>(ty_[c]);
        const P v00 = cvx<U8, S>(tv00[c]), v01 = cvx<U8, S>(tv01[c]), v10 = cvx<U8, S>(tv10[c]), v11 = cvx<U8, S>(tv11[c]);
        const P h0 = v01 - v00, h1 = v11 - v10;
        const P top = pfma(tfx, h0, v00), bot = pfma(tfx, h1, v10);
        const P dyv = bot - top;
        xn[c] = pfma(tfy, dyv, top);
        const P dxv = pfma(tfy, h1 - h0, h0);
        const P dxz = dxv * tmzx, dyz = dyv * tmzy;   // d sample / d P'1, d P'2 up to the common 1/z''
#pragma unroll
        for (int s = 0; s < S; ++s) { park[(2 * c) * S + s] = dxz.get(s); park[(2 * c + 1) * S + s] = dyz.get(s); }
      }
#pragma unroll
      for (int s = 0; s < S; ++s) { park[2 * C * S + s] = tuu.get(s); park[(2 * C + 1) * S + s] = tvv.get(s); }
#pragma unroll
      for (int k = 0; k < DRV4; ++k)
        s_drv[(wslot * DRV4 + k) * STRIP_THREADS + threadIdx.x] =
            make_float4(park[4 * k], park[4 * k + 1], park[4 * k + 2], park[4 * k + 3]);
    }
  };

  P xn[C];
  float yn[C], dn = 0.0f;
  float Fn[3 * C];
  int seln = SEL_NONE;
#pragma unroll
  for (int c = 0; c < C; ++c) { xn[c] = P(0.0f); yn[c] = 0.0f; }
#pragma unroll
  for (int k = 0; k < 3 * C; ++k) Fn[k] = 0.0f;
  // ---- rotate the rings: row r becomes row r-1 ... ---------------------------------------------
  auto rotate = [&]() {
#pragma unroll
    for (int c = 0; c < C; ++c) {
      if (!SR) { xr2[c] = xr1[c]; xr1[c] = xn[c]; }
      yr2[c] = yr1[c]; yr1[c] = yn[c];
    }
    if (!SR) {
#pragma unroll
      for (int k = 0; k < 3 * C; ++k) { F2[k] = F1[k]; F1[k] = Fn[k]; }
    }
    par ^= 1;
    sel2 = sel1; sel1 = seln;
    d3 = d2; d2 = d1; d1 = dn;
    slot = (slot + 1) % 3;
  };
  auto row = [&](int r, auto steady) {
    constexpr bool ST = decltype(steady)::value;
#if STRIP_ROWBAR
    // experiment: what a per-row rendezvous of the four scale-warps of a strip would cost BEFORE anything is shared
    // (STRIP_NWARPS = 4, Ls = 4: the warps of a CTA are the four scales of one strip segment and walk the same rows)
    asm volatile("bar.sync 1, %0;" ::"n"(STRIP_THREADS) : "memory");
#endif
#if !STRIP_ROTATE
    issue(r, steady);
#endif
    // ---- smoothness of pixel row r-2 (src/utils.jl:163-177): needs nothing of row r, so it runs in the
    // shadow of the loads just issued; one register (its pullback) crosses stages C and D
    const int pr = r - 2;
    float gsm = 0.0f;
    {
      const float wy_up = wy_dn;
      if (ST || r >= Y0 + 1) {  // weight between rows r-2 and r-1 (and wy_up of the next iteration)
        float e = 0.0f;
#pragma unroll
        for (int c = 0; c < C; ++c) e += fabsf(yr2[c] - yr1[c]);
        wy_dn = (ST || (pr >= 0 && pr + 1 < H)) ? __expf(-e * (1.0f / C)) : 0.0f;
      }
      if (ST || r >= Y0 + 2) {  // warp-uniform; pr in [Y0, Y1)
        const float dl = __shfl_sync(MDB_FULL, d2, max(lane - 1, 0)), dr = __shfl_sync(MDB_FULL, d2, min(lane + 1, 31));
        float ex = 0.0f;
#pragma unroll
        for (int c = 0; c < C; ++c) ex += fabsf(yr2[c] - __shfl_sync(MDB_FULL, yr2[c], min(lane + 1, 31)));
        const float wxr = (gx >= 0 && gx + 1 < W) ? __expf(-ex * (1.0f / C)) : 0.0f;
        const float wxl = __shfl_sync(MDB_FULL, wxr, max(lane - 1, 0));
        if (pcol) {
          // the weights are already zero across the image border
          // every difference is (first - second) of a pair, so an exact tie is +0.0 and takes +abs_tie
          const float gsx = sgnt(a.abs_tie, d2 - dr) * wxr - sgnt(a.abs_tie, dl - d2) * wxl;
          const float gsy = sgnt(a.abs_tie, d2 - d1) * wy_dn - sgnt(a.abs_tie, d3 - d2) * wy_up;
          gsm = skx * gsx + sky * gsy;
          if (DBG) dbg[(size_t)(2 * C + 1) * plane + (size_t)pr * W + gx] =
              (signed char)((sgn_code(d2 - dr) + 1) + 3 * (sgn_code(d2 - d1) + 1));
          smx_acc = fmaf(fabsf(d2 - dr), wxr, smx_acc);
          smy_acc = fmaf(fabsf(d2 - d1), wy_dn, smy_acc);
        }
      }
    }
    consume(r, slot, xn, yn, dn);
    if (SR) {   // rows r-1 and r-2 come back from shared memory, row r takes row r-2's place
      float4* const o1 = s_rx + ((par ^ 1) * NX4) * STRIP_THREADS + threadIdx.x;
      float4* const o2 = s_rx + (par * NX4) * STRIP_THREADS + threadIdx.x;
      float t1[4 * NX4], t2[4 * NX4], tn[4 * NX4];
#pragma unroll
      for (int k = 0; k < NX4; ++k) {
        const float4 q1 = o1[k * STRIP_THREADS], q2 = o2[k * STRIP_THREADS];
        t1[4 * k] = q1.x; t1[4 * k + 1] = q1.y; t1[4 * k + 2] = q1.z; t1[4 * k + 3] = q1.w;
        t2[4 * k] = q2.x; t2[4 * k + 1] = q2.y; t2[4 * k + 2] = q2.z; t2[4 * k + 3] = q2.w;
      }
#pragma unroll
      for (int k = 0; k < 4 * NX4; ++k) tn[k] = 0.0f;
#pragma unroll
      for (int c = 0; c < C; ++c)
#pragma unroll
        for (int s = 0; s < S; ++s) { xr1[c].set(s, t1[c * S + s]); xr2[c].set(s, t2[c * S + s]); tn[c * S + s] = xn[c].get(s); }
#pragma unroll
      for (int k = 0; k < NX4; ++k) o2[k * STRIP_THREADS] = make_float4(tn[4 * k], tn[4 * k + 1], tn[4 * k + 2], tn[4 * k + 3]);
    }
    if (ST && STRIP_FENCE >= 1) asm volatile("bar.warp.sync 0xffffffff;" ::: "memory");

    // =========================== C: window row r-1 =============================================
    const int wr = r - 1;
    seln = SEL_NONE;
#pragma unroll
    for (int k = 0; k < 3 * C; ++k) Fn[k] = 0.0f;
    if (ST || (r >= Y0 && wr >= 0 && wr < H)) {             // warp-uniform
      // pad_reflect(1) along H (src/utils.jl:30-31) costs nothing here: the rows -1 and H that the ring
      // delivers ARE rows 1 and H-2 (geo_row below), so the first and last window read top / bottom as usual
#if STRIP_AM_EARLY
      const float am = tam;
#else
      // the automask value of this window row, fetched before the ~300 instructions that precede its use
      float am = 0.0f;
      if (a.automask) am = ldg(amb + (unsigned)(wr * W + gxc));
#endif
      P Mxa[C], Qxa[C], Cxya[C];
      float Mya[C], Qya[C];
      P vm[C], vq[C], vc[C];
      float vmy[C], vqy[C];
#pragma unroll
      for (int c = 0; c < C; ++c) {
        const P top = xr2[c], bot = xn[c], mid = xr1[c];
        const float ty = yr2[c], by = yn[c], my_ = yr1[c];
        const float e1 = ty - my_, e2 = by - my_, se = e1 + e2, se3 = se * (-1.0f / 3.0f);
        vmy[c] = fmaf(se, 1.0f / 3.0f, my_);
        vqy[c] = fmaf(e1, e1, fmaf(e2, e2, se * se3));
        const P g1 = top - mid, g2 = bot - mid, sg = g1 + g2, sg3 = sg * (-1.0f / 3.0f);
        vm[c] = pfma(sg, 1.0f / 3.0f, mid);
        vq[c] = pfma(g1, g1, pfma(g2, g2, sg * sg3));
        vc[c] = pfma(g1, e1, pfma(g2, e2, sg3 * se));
      }
#pragma unroll
      for (int c = 0; c < C; ++c) {
        // combine the column triples l-1, l, l+1 (Chan): M2 = sum q_j + 3 sum (m_j - M)^2
        const float myL = __shfl_sync(MDB_FULL, vmy[c], nl), myR = __shfl_sync(MDB_FULL, vmy[c], nr);
        const float qyL = __shfl_sync(MDB_FULL, vqy[c], nl), qyR = __shfl_sync(MDB_FULL, vqy[c], nr);
        const P mL = vm[c].shfl(nl), mR = vm[c].shfl(nr);
        const P qL = vq[c].shfl(nl), qR = vq[c].shfl(nr);
        const P cL = vc[c].shfl(nl), cR = vc[c].shfl(nr);
        const float e1 = myL - vmy[c], e2 = myR - vmy[c], se = e1 + e2, se3 = se * (-1.0f / 3.0f);
        Mya[c] = fmaf(se, 1.0f / 3.0f, vmy[c]);
        Qya[c] = fmaf(3.0f, fmaf(e1, e1, fmaf(e2, e2, se * se3)), qyL + vqy[c] + qyR);   // 9 var(y)
        const P g1 = mL - vm[c], g2 = mR - vm[c], sg = g1 + g2, sg3 = sg * (-1.0f / 3.0f);
        Mxa[c] = pfma(sg, 1.0f / 3.0f, vm[c]);
        Qxa[c] = pfma(pfma(g1, g1, pfma(g2, g2, sg * sg3)), 3.0f, qL + vq[c] + qR);          // 9 var(x)
        Cxya[c] = pfma(pfma(g1, e1, pfma(g2, e2, sg3 * se)), 3.0f, cL + vc[c] + cR);         // 9 cov(x,y)
      }
      P lossv(0.0f);
      P cA[C], cB[C], cC[C];
      P dbg_ins[DBG ? C : 1];
#pragma unroll
      for (int c = 0; c < C; ++c) {
        const P Mx = Mxa[c], Qx = Qxa[c], Cxy = Cxya[c];
        const float My = Mya[c], Qy = Qya[c];
        // SSIM (src/utils.jl:29-43) and its pullback wrt mu_x (A), sigma_x (B), sigma_xy (Cc)
        const P n1 = pfma(Mx, 2.0f * My, P(a.c1));
        const P n2 = pfma(Cxy, 2.0f / 9.0f, P(a.c2));
        const P dd1 = pfma(Mx, Mx, P(fmaf(My, My, a.c1)));
        const P dd2 = pfma(Qx + Qy, 1.0f / 9.0f, P(a.c2));
        P r1, r2, q1, q2;
#pragma unroll
        for (int s = 0; s < S; ++s) { q1.set(s, rcp_approx(dd1.get(s))); q2.set(s, rcp_approx(dd2.get(s))); }
        r1 = n1 * q1; r2 = n2 * q2;
        const P r12 = r1 * r2;
        const P raw = pfma(r12, -0.5f, P(0.5f));
        const P dmu = Mx + (-My);
        const P Abar = (r2 * q1) * (q1 * dmu) * pfma(Mx + My, My, P(a.c1));
        const P Bb = (r12 * q2) * 0.5f;
        const P Cb = (r1 * q2) * (-1.0f);
        const P A0 = pfma(Mx * Bb, -2.0f, pfma(Cb, -My, Abar));   // A - 2 mu_x B - mu_y C
        const P l1 = xr1[c] + (-yr1[c]);
        // clamp(raw, 0, 1) passes gradient on the closed interval only: raw in [0,1] <=> |r1 r2| <= 1
        P sv, ab, ins;
#pragma unroll
        for (int s = 0; s < S; ++s) {
          sv.set(s, __saturatef(raw.get(s)));
          ab.set(s, fabsf(l1.get(s)));
          ins.set(s, fabsf(r12.get(s)) <= 1.0f ? 1.0f : 0.0f);
        }
        cA[c] = A0 * ins; cB[c] = Bb * ins; cC[c] = Cb * ins;
        if (DBG) dbg_ins[c] = ins;
        lossv = pfma(sv, ka, pfma(ab, kl, lossv));
      }
      int best = 0;
      float val = lossv.get(0);
#pragma unroll
      for (int s = 1; s < S; ++s)
        if (lossv.get(s) < val) { val = lossv.get(s); best = s; }   // first arg-min wins ties
      if (a.automask && wcol && am <= val) { val = am; best = SEL_NONE; }   // mask precedes warp_loss
      if (!wcol) best = SEL_NONE;
      seln = best;
#pragma unroll
      for (int c = 0; c < C; ++c) {
        float fa = cA[c].get(0), fb = cB[c].get(0), fc = cC[c].get(0);
#pragma unroll
        for (int s = 1; s < S; ++s) {
          fa = (best == s) ? cA[c].get(s) : fa; fb = (best == s) ? cB[c].get(s) : fb; fc = (best == s) ? cC[c].get(s) : fc;
        }
        Fn[3 * c] = fa; Fn[3 * c + 1] = fb; Fn[3 * c + 2] = fc;   // SEL_NONE windows are gated out by the row masks in D
      }
      if (DBG && pcol && (ST || (wr >= Y0 && wr < Y1))) {
        const size_t o = (size_t)wr * W + gx;
        dbg[o] = best == SEL_NONE ? (signed char)-1 : (signed char)best;
#pragma unroll
        for (int c = 0; c < C; ++c)
          dbg[(size_t)(1 + c) * plane + o] = best == SEL_NONE ? (signed char)1 : (signed char)(dbg_ins[c].get(best == SEL_NONE ? 0 : best) != 0.0f);
      }
      if (pcol && (ST || (wr >= Y0 && wr < Y1))) {
        loss_acc += val;
        if (a.vis && sc == a.Ls - 1) a.vis[(size_t)n * plane + (size_t)wr * W + gx] = val;
      }
    }

    if (ST && STRIP_FENCE >= 2) asm volatile("bar.warp.sync 0xffffffff;" ::: "memory");
#if STRIP_ROTATE
    issue(r + 1, steady);   // steady rows end one row earlier under this switch, so that row r + 1 is a steady row for issue() too
#endif
    // =========================== D: pixel row r-2 ==============================================
    if (SR) {
      float4* const o1 = s_rf + ((par ^ 1) * NF4) * STRIP_THREADS + threadIdx.x;
      float4* const o2 = s_rf + (par * NF4) * STRIP_THREADS + threadIdx.x;
      float t1[4 * NF4], t2[4 * NF4], tn[4 * NF4];
#pragma unroll
      for (int k = 0; k < NF4; ++k) {
        const float4 q1 = o1[k * STRIP_THREADS], q2 = o2[k * STRIP_THREADS];
        t1[4 * k] = q1.x; t1[4 * k + 1] = q1.y; t1[4 * k + 2] = q1.z; t1[4 * k + 3] = q1.w;
        t2[4 * k] = q2.x; t2[4 * k + 1] = q2.y; t2[4 * k + 2] = q2.z; t2[4 * k + 3] = q2.w;
      }
#pragma unroll
      for (int k = 0; k < 4 * NF4; ++k) tn[k] = 0.0f;
#pragma unroll
      for (int k = 0; k < 3 * C; ++k) { F1[k] = t1[k]; F2[k] = t2[k]; tn[k] = Fn[k]; }
#pragma unroll
      for (int k = 0; k < NF4; ++k) o2[k * STRIP_THREADS] = make_float4(tn[4 * k], tn[4 * k + 1], tn[4 * k + 2], tn[4 * k + 3]);
    }
    if (ST || r >= Y0 + 2) {  // warp-uniform; pr in [Y0, Y1)
      // box sums of the coefficient fields over the 3x3 windows containing this pixel, per source
      const float wu = (!ST && pr == 1) ? 2.0f : 1.0f, wd = (!ST && pr == H - 2) ? 2.0f : 1.0f;
      P ma, mb, mc;   // row weights gated by which source won the window
#pragma unroll
      for (int s = 0; s < S; ++s) {
        ma.set(s, sel2 == s ? wu : 0.0f); mb.set(s, sel1 == s ? 1.0f : 0.0f); mc.set(s, seln == s ? wd : 0.0f);
      }
      P T[3 * C];
#pragma unroll
      for (int k = 0; k < 3 * C; ++k) {
        const P V = pfma(ma, F2[k], pfma(mb, F1[k], mc * Fn[k]));
        const P VL = V.shfl(lane - 1 < 0 ? 0 : lane - 1), VR = V.shfl(lane + 1 > 31 ? 31 : lane + 1);
        T[k] = pfma(VL, wl, pfma(VR, wr_, V));
      }
      if (pcol) {
        const float4* dv = s_drv + (size_t)((slot + 1) % 3 * DRV4) * STRIP_THREADS + threadIdx.x;
        float park[DRV];
#pragma unroll
        for (int k = 0; k < DRV4; ++k) {
          const float4 q = dv[k * STRIP_THREADS];
          park[4 * k] = q.x; park[4 * k + 1] = q.y; park[4 * k + 2] = q.z; park[4 * k + 3] = q.w;
        }
        P dP0(0.0f), dP1(0.0f);
#pragma unroll
        for (int c = 0; c < C; ++c) {
          const P x = xr2[c];
          const float y = yr2[c];
          const P df = x + (-y);
          P sg;
#pragma unroll
          // |target - pred| (src/training.jl:5): target - pred = +0.0 at a tie, so d/d pred = -abs_tie there
          for (int s = 0; s < S; ++s) sg.set(s, sgnt(-a.abs_tie, df.get(s)));
          if (DBG) dbg[(size_t)(C + 1 + c) * plane + (size_t)pr * W + gx] = sel1 == SEL_NONE ? (signed char)0 : sgn_code(df.get(sel1 == SEL_NONE ? 0 : sel1));
          sg = sg * mb;
          // d loss / d sample = ks [T_A + 2 x T_B + y T_C] + kl sign(x - y) [this window chose s]
          const P G = pfma(pfma(x * 2.0f, T[3 * c + 1], pfma(T[3 * c + 2], y, T[3 * c])), ksu, sg * klu);
          P dxz, dyz;
#pragma unroll
          for (int s = 0; s < S; ++s) { dxz.set(s, park[(2 * c) * S + s]); dyz.set(s, park[(2 * c + 1) * S + s]); }
          dP0 = pfma(G, dxz, dP0);
          dP1 = pfma(G, dyz, dP1);
        }
        P uu, vv;
#pragma unroll
        for (int s = 0; s < S; ++s) { uu.set(s, park[2 * C * S + s]); vv.set(s, park[(2 * C + 1) * S + s]); }
        // u = P'1 / z'', v = P'2 / z''  =>  dP'3 = -(u dP'1 + v dP'2)   (the 1/z'' is already inside dxz, dyz)
        const P dP2 = pfma(dP0, uu, dP1 * vv) * (-1.0f);
        const float invd = fmaf(d2, a.dd, a.mind);
        const float py = (float)(pr + 1);
        S0[0] = S0[0] + dP0; S0[1] = S0[1] + dP1; S0[2] = S0[2] + dP2;
        SP[0] = pfma(dP0, py, SP[0]); SP[1] = pfma(dP1, py, SP[1]); SP[2] = pfma(dP2, py, SP[2]);
        SI[0] = pfma(dP0, invd, SI[0]); SI[1] = pfma(dP1, invd, SI[1]); SI[2] = pfma(dP2, invd, SI[2]);
        float dinvd = 0.0f;
#pragma unroll
        for (int s = 0; s < S; ++s)
          dinvd += dP0.get(s) * PPV(s, b0f, b0f) + dP1.get(s) * PPV(s, b1f, b1f) + dP2.get(s) * PPV(s, b3f, b3f);
        // d invd / d disparity = 1/min_depth - 1/max_depth; plus the smoothness pullback formed above
        // (the mean-normalisation constant is added after the item's sum is known)
        gout[(size_t)pr * W + gx] = fmaf(dinvd, a.dd, gsm);
      }
    }

    rotate();
  };
  {
    // steady-state rows: all three stages active (r >= Y0 + 2), window row r-1 and pixel row r-2 inside this segment
    // (r <= Y1), no reflected row in sight (pixel row >= 2 and <= H-3, r + 1 <= H - 1)
#if STRIP_TMA_TGT
    if (tma_ok) { tma_issue(Y0 - 2); tma_issue(Y0 - 1); }
#endif
    constexpr bool PEEL = STRIP_PEEL && (CP || DBG || STRIP_PEEL_CAPTURED);
    const int s_lo = max(Y0 + 2, 4), s_hi = PEEL ? min(Y1, H - 2 - STRIP_ROTATE) : -1000000;
    int r = Y0 - 2;
#if STRIP_ROTATE
    issue(r, std::false_type{});
#endif
    while (r <= Y1 + 1) {
      if (PEEL && r >= s_lo && r <= s_hi) {
#if STRIP_UNROLL == 3
        // experiment 19: three rows per trip, so that the compiler can rename the three-deep rings instead of moving them
        for (; r + 2 <= s_hi; r += 3) { row(r, std::true_type{}); row(r + 1, std::true_type{}); row(r + 2, std::true_type{}); }
#elif STRIP_UNROLL == 2
        for (; r + 1 <= s_hi; r += 2) { row(r, std::true_type{}); row(r + 1, std::true_type{}); }
#endif
        for (; r <= s_hi; ++r) row(r, std::true_type{});
      } else {
        row(r, std::false_type{});
        ++r;
      }
    }
  }

  // ---- one record per warp: loss + 12 pose sums per source (no atomics) -------------------------
  float* out = a.partials + (size_t)rec * NPART;
  {
    const float v = warp_sum(loss_acc), vx = warp_sum(smx_acc), vy = warp_sum(smy_acc);
    if (lane == 0) { out[0] = v; out[1] = vx; out[2] = vy; }
  }
  const float px = (float)(gx + 1);
#pragma unroll
  for (int s = 0; s < S; ++s)
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const float s0 = S0[i].get(s);
      const float v0 = warp_sum(s0 * px), v1 = warp_sum(SP[i].get(s)), v2 = warp_sum(s0), v3 = warp_sum(SI[i].get(s));
      if (lane == 0) {
        out[STRIP_NHEAD + s * NPOSE + 3 * i] = v0; out[STRIP_NHEAD + s * NPOSE + 3 * i + 1] = v1; out[STRIP_NHEAD + s * NPOSE + 3 * i + 2] = v2;
        out[STRIP_NHEAD + s * NPOSE + 9 + i] = v3;
      }
    }
}

}  // namespace
