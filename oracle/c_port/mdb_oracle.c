/* mdb_oracle.c -- plain-C restatement of the Monodepth2.jl view-synthesis loss path, value AND
 * hand-derived gradient.  TEST INFRASTRUCTURE ONLY: the checker for tests/, for
 * __graft_entry__.smoke() and the timed CPU baseline of bench.py.  Nothing under
 * monodepth2.jl_b200/ links, loads or calls it, and the product path has no CPU fallback.
 *
 * It is a second, independent restatement next to oracle/torch_ref.py (which lets torch
 * autograd differentiate the forward).  Here every pullback is written out by hand, op by op in
 * the reference's own (unfused) order, so that the two oracles only agree if both the forward
 * restatement and the derivative of every op agree (tests/test_oracle_c.py, 1e-9).
 *
 * Pinning status: same as oracle/torch_ref.py -- the reference is Julia and cannot run here;
 * the pieces test/runtests.jl pins (so3, composeT, SSIM identities, smooth loss 0.2542, depth
 * range, identity warp) are checked against this file too; train_loss as a whole and the
 * gradients are "parity unpinned" against the reference itself.
 *
 * Layout: Julia column-major arrays are passed as the same bytes: x (W,H,C,L,N) is
 * x[n][l][c][h][w]; disparity (w,h,1,N) is d[n][h][w]; rvec (3,N) is r[n][3]; K, invK are the
 * 9 numbers Julia stores (column-major, i.e. the transpose of the math-natural row-major matrix).
 * Pixel coordinates are 1-based inside the math, as in the reference.
 *
 * Build:  make -C oracle/c_port          (gcc -O2 -pthread; -DMDBO_FLOAT for the fp32 variant the
 *                                         CPU baseline times -- the reference computes in Float32)
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include <stdatomic.h>

#ifdef MDBO_FLOAT
typedef float real;
#define R_ABS fabsf
#define R_EXP expf
#define R_FLOOR floorf
#else
typedef double real;
#define R_ABS fabs
#define R_EXP exp
#define R_FLOOR floor
#endif

#define MDBO_MAX_S 4
#define MDBO_MAX_SCALES 8

/* ------------------------------------------------------------------------------------------
 * so3 / pose, always in double (6 numbers per pose)           src/utils.jl:106-145, 185-192
 * matrices are math-natural row-major M[3 i + j]
 * ---------------------------------------------------------------------------------------- */
static void mat3_mul(const double* A, const double* B, double* C) {
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      double s = 0;
      for (int k = 0; k < 3; ++k) s += A[3 * i + k] * B[3 * k + j];
      C[3 * i + j] = s;
    }
}

/* hat(rvec)                                                        src/utils.jl:123-132 */
static void hat_rm(const double r[3], double S[9]) {
  S[0] = 0;     S[1] = -r[2]; S[2] = r[1];
  S[3] = r[2];  S[4] = 0;     S[5] = -r[0];
  S[6] = -r[1]; S[7] = r[0];  S[8] = 0;
}

typedef struct { double theta, tinv, f1, f2, S[9], S2[9]; } so3_aux;

/* so3_exp_map: R = f1 S + f2 S^2 + I, f1 = sin(t)/max(t,1e-4), f2 = (1-cos t)/max(t,1e-4)^2
 *                                                                   src/utils.jl:106-121 */
static void so3_exp_rm(const double r[3], double R[9], so3_aux* a) {
  hat_rm(r, a->S);
  mat3_mul(a->S, a->S, a->S2);
  a->theta = sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
  a->tinv = 1.0 / fmax(a->theta, 1e-4);
  a->f1 = a->tinv * sin(a->theta);
  a->f2 = a->tinv * a->tinv * (1.0 - cos(a->theta));
  for (int k = 0; k < 9; ++k) R[k] = a->f1 * a->S[k] + a->f2 * a->S2[k];
  R[0] += 1.0; R[4] += 1.0; R[8] += 1.0;
}

/* pullback of so3_exp_map (broadcast products, batched_mul, hat rrule, sqrt(sum(abs2)))
 * dR -> dr.  theta == 0 gives NaN exactly as the reference's sqrt pullback (README.md:47-51). */
static void so3_exp_bwd_rm(const double r[3], const so3_aux* a, const double dR[9], double dr[3]) {
  double df1 = 0, df2 = 0, dS[9];
  for (int k = 0; k < 9; ++k) { df1 += dR[k] * a->S[k]; df2 += dR[k] * a->S2[k]; }
  /* S2 = S S  =>  dS += f2 (dR S^T + S^T dR) */
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      double s = 0;
      for (int k = 0; k < 3; ++k) s += dR[3 * i + k] * a->S[3 * j + k] + a->S[3 * k + i] * dR[3 * k + j];
      dS[3 * i + j] = a->f1 * dR[3 * i + j] + a->f2 * s;
    }
  /* rrule(hat)                                                    src/utils.jl:134-145 */
  dr[0] = dS[7] - dS[5];
  dr[1] = -dS[6] + dS[2];
  dr[2] = dS[3] - dS[1];
  /* theta_inv = 1/max(theta,1e-4); f1 = theta_inv sin; f2 = theta_inv^2 (1-cos) */
  const double dtinv = (a->theta > 1e-4) ? -a->tinv * a->tinv : 0.0;
  const double s = sin(a->theta), c = cos(a->theta);
  const double df1dt = a->tinv * c + s * dtinv;
  const double df2dt = a->tinv * a->tinv * s + 2.0 * a->tinv * (1.0 - c) * dtinv;
  const double dtheta = df1 * df1dt + df2 * df2dt;
  for (int k = 0; k < 3; ++k) dr[k] += dtheta * (r[k] / a->theta);
}

/* composeT(rvec, t, invert): R' = R or R^T, t' = t or R^T (-t)       src/utils.jl:185-192 */
static void compose_T_rm(const double r[3], const double t[3], int invert, double Rp[9], double tp[3],
                         so3_aux* a) {
  double R[9];
  so3_exp_rm(r, R, a);
  if (!invert) {
    memcpy(Rp, R, sizeof(R));
    memcpy(tp, t, 3 * sizeof(double));
  } else {
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) Rp[3 * i + j] = R[3 * j + i];
    for (int i = 0; i < 3; ++i) tp[i] = -(Rp[3 * i] * t[0] + Rp[3 * i + 1] * t[1] + Rp[3 * i + 2] * t[2]);
  }
}

static void compose_T_bwd_rm(const double r[3], const double t[3], int invert, const double dRp[9],
                             const double dtp[3], double dr[3], double dt[3]) {
  so3_aux a;
  double R[9], dR[9];
  so3_exp_rm(r, R, &a);
  if (!invert) {
    memcpy(dR, dRp, sizeof(dR));
    memcpy(dt, dtp, 3 * sizeof(double));
  } else {
    /* R'_ij = R_ji ; t'_i = -sum_j R_ji t_j */
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) dR[3 * j + i] = dRp[3 * i + j] - dtp[i] * t[j];
    for (int j = 0; j < 3; ++j) dt[j] = -(R[3 * j] * dtp[0] + R[3 * j + 1] * dtp[1] + R[3 * j + 2] * dtp[2]);
  }
  so3_exp_bwd_rm(r, &a, dR, dr);
}

/* exported for the known-answer tests (Julia layout in and out: (3,3,N) column-major) */
void mdbo_so3_exp_map(const double* rvec, int N, double* R_jl) {
  for (int n = 0; n < N; ++n) {
    double R[9];
    so3_aux a;
    so3_exp_rm(rvec + 3 * n, R, &a);
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) R_jl[9 * n + 3 * j + i] = R[3 * i + j];
  }
}

void mdbo_compose_T(const double* rvec, const double* tvec, int invert, int N, double* R_jl, double* t_out) {
  for (int n = 0; n < N; ++n) {
    double R[9];
    so3_aux a;
    compose_T_rm(rvec + 3 * n, tvec + 3 * n, invert, R, t_out + 3 * n, &a);
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) R_jl[9 * n + 3 * j + i] = R[3 * i + j];
  }
}

void mdbo_compose_T_bwd(const double* rvec, const double* tvec, int invert, int N, const double* dR_jl,
                        const double* dt_in, double* d_rvec, double* d_tvec) {
  for (int n = 0; n < N; ++n) {
    double dR[9];
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) dR[3 * i + j] = dR_jl[9 * n + 3 * j + i];
    compose_T_bwd_rm(rvec + 3 * n, tvec + 3 * n, invert, dR, dt_in + 3 * n, d_rvec + 3 * n, d_tvec + 3 * n);
  }
}

/* ------------------------------------------------------------------------------------------
 * small helpers
 * ---------------------------------------------------------------------------------------- */
static inline int reflect1(int i, int n) {          /* pad_reflect by 1: -1 -> 1, n -> n-2 */
  return i < 0 ? -i : (i >= n ? 2 * n - 2 - i : i);
}
static inline real sgn(real v) { return (real)((v > 0) - (v < 0)); }   /* abs pullback = sign, 0 at 0 */

/* source coordinate of upsample_bilinear: full-res index p -> (j0, j1, frac).  align-corners (NNlib's default):
 * p (nat-1)/(full-1); half_pixel (align_corners = false, NNlib >= 0.8 / torch): max(0, (p + 0.5) nat/full - 0.5) */
static inline void up_coord(int p, int full, int nat, int half_pixel, int* j0, int* j1, real* f) {
  if (full <= 1 || nat <= 1) { *j0 = 0; *j1 = 0; *f = 0; return; }
  double s;
  if (half_pixel) {
    s = ((double)p + 0.5) * (double)nat / (double)full - 0.5;
    if (s < 0) s = 0;
  } else {
    s = (double)p * (double)(nat - 1) / (double)(full - 1);
  }
  int a = (int)floor(s);
  if (a > nat - 1) a = nat - 1;
  *j0 = a;
  *j1 = a + 1 < nat ? a + 1 : nat - 1;
  *f = (real)(s - (double)a);
}

typedef struct {
  /* per-item scratch, all H*W unless noted */
  real *up, *gup, *wl, *gdn;
  real *warped[MDBO_MAX_S];      /* [C][H][W] */
  real *gwarp[MDBO_MAX_S];       /* [C][H][W] */
  real *ix[MDBO_MAX_S], *iy[MDBO_MAX_S];   /* clamped 0-based sample coordinates */
  unsigned char *mx[MDBO_MAX_S], *my[MDBO_MAX_S];  /* clip multiplier (1 = inside) */
  real *pl[MDBO_MAX_S];
  int *sel;
} scratch_t;

static int scratch_alloc(scratch_t* s, int W, int H, int C, int S) {
  const size_t hw = (size_t)W * H;
  memset(s, 0, sizeof(*s));
  s->up = malloc(hw * sizeof(real)); s->gup = malloc(hw * sizeof(real));
  s->wl = malloc(hw * sizeof(real)); s->gdn = malloc(hw * sizeof(real));
  s->sel = malloc(hw * sizeof(int));
  int ok = s->up && s->gup && s->wl && s->gdn && s->sel;
  for (int k = 0; k < S; ++k) {
    s->warped[k] = malloc(hw * C * sizeof(real)); s->gwarp[k] = malloc(hw * C * sizeof(real));
    s->ix[k] = malloc(hw * sizeof(real)); s->iy[k] = malloc(hw * sizeof(real));
    s->mx[k] = malloc(hw); s->my[k] = malloc(hw);
    s->pl[k] = malloc(hw * sizeof(real));
    ok = ok && s->warped[k] && s->gwarp[k] && s->ix[k] && s->iy[k] && s->mx[k] && s->my[k] && s->pl[k];
  }
  return ok;
}
static void scratch_free(scratch_t* s, int S) {
  free(s->up); free(s->gup); free(s->wl); free(s->gdn); free(s->sel);
  for (int k = 0; k < S; ++k) {
    free(s->warped[k]); free(s->gwarp[k]); free(s->ix[k]); free(s->iy[k]);
    free(s->mx[k]); free(s->my[k]); free(s->pl[k]);
  }
}

/* (ssim::SSIM)(x = pred, y = target) for one channel plane, reflect pad 1 + 3x3 mean pool.
 * Returns the clamped value at (px,py) and, if g != NULL, the 6 window statistics needed by the
 * pullback: g = {mx, my, dval/dmx, dval/dsx, dval/dsxy, gate}.          src/utils.jl:29-43 */
static inline real ssim_at(const real* X, const real* Y, int W, int H, int px, int py, real c1, real c2,
                           real* g) {
  real sx = 0, sy = 0, sxx = 0, syy = 0, sxy = 0;
  for (int dy = -1; dy <= 1; ++dy) {
    const int yy = reflect1(py + dy, H);
    for (int dx = -1; dx <= 1; ++dx) {
      const int xx = reflect1(px + dx, W);
      const real a = X[(size_t)yy * W + xx], b = Y[(size_t)yy * W + xx];
      sx += a; sy += b; sxx += a * a; syy += b * b; sxy += a * b;
    }
  }
  const real ninth = (real)1 / (real)9;
  const real mx = sx * ninth, my = sy * ninth;
  const real vx = sxx * ninth - mx * mx, vy = syy * ninth - my * my, cxy = sxy * ninth - mx * my;
  const real n1 = 2 * mx * my + c1, n2 = 2 * cxy + c2;
  const real d1 = mx * mx + my * my + c1, d2 = vx + vy + c2;
  const real n = n1 * n2, d = d1 * d2;
  const real raw = (1 - n / d) * (real)0.5;
  if (g) {
    /* val = (1 - n/d)/2 ; partials with (mx, vx, cxy) as the independent statistics of X */
    const real inv_d = 1 / d;
    const real dn_dmx = 2 * my * n2, dn_dc = 2 * n1;
    const real dd_dmx = 2 * mx * d2, dd_dv = d1;
    g[0] = mx; g[1] = my;
    g[2] = (real)-0.5 * (dn_dmx * inv_d - n * dd_dmx * inv_d * inv_d);
    g[3] = (real)-0.5 * (-n * dd_dv * inv_d * inv_d);
    g[4] = (real)-0.5 * (dn_dc * inv_d);
    g[5] = (raw >= 0 && raw <= 1) ? (real)1 : (real)0;     /* clamp passes the gradient inside [0,1] */
  }
  return raw < 0 ? 0 : (raw > 1 ? 1 : raw);
}

/* ------------------------------------------------------------------------------------------
 * train_loss (src/training.jl:21-78) / the slow_depth closure (src/simple_depth.jl:25-41)
 * ---------------------------------------------------------------------------------------- */
typedef struct mdbo_desc {
  int32_t W, H, C, L, N, target_id, S;
  int32_t source_ids[MDBO_MAX_S];         /* 1-based */
  int32_t Ls;
  int32_t dw[MDBO_MAX_SCALES], dh[MDBO_MAX_SCALES];
  double scales[MDBO_MAX_SCALES];
  double min_depth, max_depth, smoothness, alpha, c1, c2;
  int32_t automask;        /* Params.automasking                          src/training.jl:60-62 */
  int32_t normalize;       /* 1: disparity / (mean + 1e-7)                src/training.jl:64-65 */
  int32_t smooth_mode;     /* 0: smoothness * scale[k]; 1: weight 1       src/simple_depth.jl:39 */
  int32_t n_global;        /* batch size the means divide by; 0 = N */
  int32_t abs_subgradient; /* pullback of abs at +-0 (Zygote, unpinned): 0 = sign(0) = 0 (ChainRules' rule);
                              1 = ForwardDiff's Dual rule, which Zygote's broadcast uses on CuArrays:
                              abs(d) = signbit(value) ? -d : d, i.e. +1 at +0.0 and -1 at -0.0 */
  int32_t upsample_half_pixel; /* 0: upsample_bilinear with align-corners (NNlib's default, src/training.jl:45);
                                  1: half-pixel centres (align_corners = false) -- nothing in the reference pins it */
} mdbo_desc;

/* Decision forcing / dumping (tests/test_gpu_forced.py).  Every discrete choice the pullback depends on, one
 * int8 plane each, layout [scale][item][2C+2][H][W] (the CUDA kernel's DBG dump has the same layout):
 *   0          winner of the window: source index, -1 = the automask took it
 *   1..C       the winner's SSIM clamp gate per channel
 *   C+1..2C    sign(pred - target) of the winner per channel (L1 pullback)
 *   2C+1       (sign(d - d_right) + 1) + 3 (sign(d - d_down) + 1)
 * `force`: use these instead of the oracle's own; `dump`: record the oracle's own; `margin` [scale][item][H][W]:
 * gap between the winner and the runner-up candidate (how close the arg-min was to a tie).  Static pointers,
 * read by the next mdbo_train_loss call: test infrastructure, not re-entrant. */
static const int8_t* g_force = NULL;
static int8_t* g_dump = NULL;
static float* g_margin = NULL;
void mdbo_set_decisions(const int8_t* force, int8_t* dump, float* margin) { g_force = force; g_dump = dump; g_margin = margin; }
static inline int sgn_i(real v) { return (v > 0) - (v < 0); }
/* abs pullback with the configured convention at +-0 (v is the argument of abs) */
static inline real abs_grad(real v, int mode) {
  if (v > 0) return 1;
  if (v < 0) return -1;
  return mode ? (signbit((double)v) ? (real)-1 : (real)1) : (real)0;
}

typedef struct { double R[9], t[3]; } pose_t;

/* one item, one scale; returns the item's contribution to loss_k (already divided by the global
 * element counts) and accumulates the gradients when gdisp != NULL */
static double item_scale(const mdbo_desc* d, int n, int k, const real* x, const pose_t* P, const double* Km,
                         const double* iKm, const real* disp_k, const real* auto_n, scratch_t* s,
                         real* gdisp_k, double (*dRacc)[9], double (*dtacc)[3]) {
  const int W = d->W, H = d->H, C = d->C, S = d->S, dw = d->dw[k], dh = d->dh[k];
  const size_t hw = (size_t)W * H;
  const int Ng = d->n_global > 0 ? d->n_global : d->N;
  const real* xn = x + (size_t)n * d->L * C * hw;
  const real* tgt = xn + (size_t)(d->target_id - 1) * C * hw;
  const real c1 = (real)d->c1, c2 = (real)d->c2, alpha = (real)d->alpha;
  const double min_disp = 1.0 / d->max_depth, max_disp = 1.0 / d->min_depth;   /* src/utils.jl:179-183 */
  const real dd = (real)(max_disp - min_disp), mind = (real)min_disp;
  const int native = (dw == W && dh == H);

  /* upsample_bilinear(disparity; size=(W,H)) when the sizes differ     src/training.jl:44-46 */
  double msum = 0;
  for (int py = 0; py < H; ++py) {
    int y0, y1; real fy;
    up_coord(py, H, dh, d->upsample_half_pixel, &y0, &y1, &fy);
    for (int px = 0; px < W; ++px) {
      real v;
      if (native) v = disp_k[(size_t)py * W + px];
      else {
        int x0, x1; real fx;
        up_coord(px, W, dw, d->upsample_half_pixel, &x0, &x1, &fx);
        const real a = disp_k[(size_t)y0 * dw + x0], b = disp_k[(size_t)y0 * dw + x1];
        const real c = disp_k[(size_t)y1 * dw + x0], e = disp_k[(size_t)y1 * dw + x1];
        v = (a * (1 - fx) + b * fx) * (1 - fy) + (c * (1 - fx) + e * fx) * fy;
      }
      s->up[(size_t)py * W + px] = v;
      msum += v;
    }
  }

  /* disparity_to_depth -> Backproject -> Project -> grid_sample       src/training.jl:48-57 */
  for (int py = 0; py < H; ++py)
    for (int px = 0; px < W; ++px) {
      const size_t p = (size_t)py * W + px;
      const real depth = 1 / (s->up[p] * dd + mind);
      const real q0 = (real)(px + 1), q1 = (real)(py + 1);
      real X[3];
      for (int i = 0; i < 3; ++i) X[i] = depth * ((real)iKm[3 * i] * q0 + (real)iKm[3 * i + 1] * q1 + (real)iKm[3 * i + 2]);
      for (int si = 0; si < S; ++si) {
        real Y[3], cam[3];                                               /* K (R X + t), src/utils.jl:83-85 */
        for (int i = 0; i < 3; ++i)
          Y[i] = (real)P[si].R[3 * i] * X[0] + (real)P[si].R[3 * i + 1] * X[1] + (real)P[si].R[3 * i + 2] * X[2] + (real)P[si].t[i];
        for (int i = 0; i < 3; ++i)
          cam[i] = (real)Km[3 * i] * Y[0] + (real)Km[3 * i + 1] * Y[1] + (real)Km[3 * i + 2] * Y[2];
        const real den = 1 / (cam[2] + (real)1e-7);
        const real u = cam[0] * den, v = cam[1] * den;                   /* 1-based pixel coords */
        /* normalize (src/utils.jl:99-103), then NNlib's un-normalisation (align corners) */
        const real gx = (((u - 1) / (real)(W - 1)) - (real)0.5) * 2, gy = (((v - 1) / (real)(H - 1)) - (real)0.5) * 2;
        real ix = (gx + 1) * (real)0.5 * (real)(W - 1), iy = (gy + 1) * (real)0.5 * (real)(H - 1);
        unsigned char mx = 1, my = 1;
        if (!(ix >= 0)) { ix = 0; mx = 0; } else if (ix > (real)(W - 1)) { ix = (real)(W - 1); mx = 0; }
        if (!(iy >= 0)) { iy = 0; my = 0; } else if (iy > (real)(H - 1)) { iy = (real)(H - 1); my = 0; }
        s->ix[si][p] = ix; s->iy[si][p] = iy; s->mx[si][p] = mx; s->my[si][p] = my;
        const int x0 = (int)R_FLOOR(ix), y0 = (int)R_FLOOR(iy);
        const real fx = ix - (real)x0, fy = iy - (real)y0;
        const int xin = x0 + 1 < W, yin = y0 + 1 < H;
        const real* src = xn + (size_t)(d->source_ids[si] - 1) * C * hw;
        for (int c = 0; c < C; ++c) {
          const real* pl = src + (size_t)c * hw;
          real v = pl[(size_t)y0 * W + x0] * (1 - fx) * (1 - fy);
          if (xin) v += pl[(size_t)y0 * W + x0 + 1] * fx * (1 - fy);
          if (yin) v += pl[(size_t)(y0 + 1) * W + x0] * (1 - fx) * fy;
          if (xin && yin) v += pl[(size_t)(y0 + 1) * W + x0 + 1] * fx * fy;
          s->warped[si][(size_t)c * hw + p] = v;
        }
      }
    }

  /* photometric_loss per source (src/training.jl:1-7), minimum over sources (:13-15, first
   * arg-min takes the gradient), _apply_mask (:17-19, the mask comes first so it wins ties) */
  double lsum = 0;
  for (int py = 0; py < H; ++py)
    for (int px = 0; px < W; ++px) {
      const size_t p = (size_t)py * W + px;
      real best = 0; int bs = 0;
      for (int si = 0; si < S; ++si) {
        real ss = 0, l1 = 0;
        for (int c = 0; c < C; ++c) {
          ss += ssim_at(s->warped[si] + (size_t)c * hw, tgt + (size_t)c * hw, W, H, px, py, c1, c2, NULL);
          l1 += R_ABS(tgt[(size_t)c * hw + p] - s->warped[si][(size_t)c * hw + p]);
        }
        const real v = alpha * (ss / (real)C) + (1 - alpha) * (l1 / (real)C);
        s->pl[si][p] = v;
        if (si == 0 || v < best) { best = v; bs = si; }
      }
      if (d->automask && auto_n && auto_n[p] <= best) { best = auto_n[p]; bs = -1; }
      if (g_margin) {          /* distance of the winner to the closest other candidate */
        real gap = (real)1e30;
        for (int si = 0; si < S; ++si)
          if (si != bs && s->pl[si][p] - best < gap) gap = s->pl[si][p] - best;
        if (d->automask && auto_n && bs >= 0 && auto_n[p] - best < gap) gap = auto_n[p] - best;
        g_margin[((size_t)k * d->N + n) * hw + p] = (float)gap;
      }
      if (g_dump) g_dump[((size_t)k * d->N + n) * (2 * C + 2) * hw + p] = (int8_t)bs;
      if (g_force) {
        bs = g_force[((size_t)k * d->N + n) * (2 * C + 2) * hw + p];
        best = bs < 0 ? auto_n[p] : s->pl[bs][p];
      }
      s->wl[p] = best; s->sel[p] = bs;
      lsum += best;
    }
  double loss = lsum / ((double)hw * Ng);

  /* smooth_loss(disparity ./ (mean + 1e-7), target) * weight    src/training.jl:64-67, utils.jl:163-177 */
  const double m = msum / (double)hw;
  const real inv = d->normalize ? (real)(1.0 / (m + 1e-7)) : (real)1;
  const double wk = d->smooth_mode == 0 ? d->smoothness * d->scales[k] : 1.0;
  const double cx = wk / ((double)(W - 1) * H * Ng), cy = wk / ((double)W * (H - 1) * Ng);
  double sxs = 0, sys = 0;
  if (gdisp_k) memset(s->gdn, 0, hw * sizeof(real));
  const real gscale = (real)(1.0 / d->Ls);
  const size_t dec0 = ((size_t)k * d->N + n) * (2 * C + 2) * hw;     /* this (scale, item)'s decision planes */
  for (int py = 0; py < H; ++py)
    for (int px = 0; px < W; ++px) {
      const size_t p = (size_t)py * W + px;
      int cx_code = 1, cy_code = 1;
      const int fcode = g_force ? g_force[dec0 + (size_t)(2 * C + 1) * hw + p] : 0;
      if (px + 1 < W) {
        real e = 0;
        for (int c = 0; c < C; ++c) e += R_ABS(tgt[(size_t)c * hw + p] - tgt[(size_t)c * hw + p + 1]);
        const real w = R_EXP(-(e / (real)C));
        const real df = s->up[p] * inv - s->up[p + 1] * inv;
        sxs += R_ABS(df) * w;
        real sg = abs_grad(df, d->abs_subgradient);
        cx_code = sgn_i(df) + 1;
        if (g_force) sg = (fcode % 3 - 1) != 0 ? (real)(fcode % 3 - 1) : abs_grad((real)0, d->abs_subgradient);
        if (gdisp_k) { const real g = sg * w * (real)cx * gscale; s->gdn[p] += g; s->gdn[p + 1] -= g; }
      }
      if (py + 1 < H) {
        real e = 0;
        for (int c = 0; c < C; ++c) e += R_ABS(tgt[(size_t)c * hw + p] - tgt[(size_t)c * hw + p + W]);
        const real w = R_EXP(-(e / (real)C));
        const real df = s->up[p] * inv - s->up[p + W] * inv;
        sys += R_ABS(df) * w;
        real sg = abs_grad(df, d->abs_subgradient);
        cy_code = sgn_i(df) + 1;
        if (g_force) sg = (fcode / 3 - 1) != 0 ? (real)(fcode / 3 - 1) : abs_grad((real)0, d->abs_subgradient);
        if (gdisp_k) { const real g = sg * w * (real)cy * gscale; s->gdn[p] += g; s->gdn[p + W] -= g; }
      }
      if (g_dump) g_dump[dec0 + (size_t)(2 * C + 1) * hw + p] = (int8_t)(cx_code + 3 * cy_code);
    }
  loss += sxs * cx + sys * cy;
  if (!gdisp_k) return loss;

  /* ---------------- pullback, op by op in reverse ---------------- */
  const real gmean = gscale / (real)((double)hw * Ng);     /* d loss / d warp_loss[p] */
  for (int si = 0; si < S; ++si) memset(s->gwarp[si], 0, hw * C * sizeof(real));
  const real ninth = (real)1 / (real)9;
  for (int py = 0; py < H; ++py)
    for (int px = 0; px < W; ++px) {
      const size_t p = (size_t)py * W + px;
      const int si = s->sel[p];
      if (si < 0) {                                         /* the automask took this pixel */
        if (g_dump) for (int c = 0; c < C; ++c) { g_dump[dec0 + (size_t)(1 + c) * hw + p] = 1; g_dump[dec0 + (size_t)(C + 1 + c) * hw + p] = 0; }
        continue;
      }
      for (int c = 0; c < C; ++c) {
        const real* Xc = s->warped[si] + (size_t)c * hw;
        const real* Yc = tgt + (size_t)c * hw;
        real* G = s->gwarp[si] + (size_t)c * hw;
        /* L1: d |t - w| / dw = -abs'(t - w) */
        real l1g = -abs_grad(Yc[p] - Xc[p], d->abs_subgradient);
        if (g_dump) g_dump[dec0 + (size_t)(C + 1 + c) * hw + p] = (int8_t)sgn_i(Xc[p] - Yc[p]);
        if (g_force) {
          const int fs = g_force[dec0 + (size_t)(C + 1 + c) * hw + p];
          l1g = fs != 0 ? (real)fs : -abs_grad((real)0, d->abs_subgradient);
        }
        G[p] += gmean * (1 - alpha) / (real)C * l1g;
        /* SSIM through the 9 (reflected) taps of the window */
        real st[6];
        ssim_at(Xc, Yc, W, H, px, py, c1, c2, st);
        if (g_dump) g_dump[dec0 + (size_t)(1 + c) * hw + p] = (int8_t)(st[5] != 0);
        if (g_force) st[5] = (real)g_force[dec0 + (size_t)(1 + c) * hw + p];
        if (st[5] == 0) continue;
        const real q = gmean * alpha / (real)C;
        for (int dy = -1; dy <= 1; ++dy) {
          const int yy = reflect1(py + dy, H);
          for (int dx = -1; dx <= 1; ++dx) {
            const int xx = reflect1(px + dx, W);
            const size_t j = (size_t)yy * W + xx;
            G[j] += q * ninth * (st[2] + st[3] * 2 * (Xc[j] - st[0]) + st[4] * (Yc[j] - st[1]));
          }
        }
      }
    }

  /* grid_sample pullback to the grid, Project, Backproject, disparity_to_depth */
  memset(s->gup, 0, hw * sizeof(real));
  for (int py = 0; py < H; ++py)
    for (int px = 0; px < W; ++px) {
      const size_t p = (size_t)py * W + px;
      const real depth = 1 / (s->up[p] * dd + mind);
      const real q0 = (real)(px + 1), q1 = (real)(py + 1);
      real ray[3], X[3], gX[3] = {0, 0, 0};
      for (int i = 0; i < 3; ++i) {
        ray[i] = (real)iKm[3 * i] * q0 + (real)iKm[3 * i + 1] * q1 + (real)iKm[3 * i + 2];
        X[i] = depth * ray[i];
      }
      for (int si = 0; si < S; ++si) {
        const real ix = s->ix[si][p], iy = s->iy[si][p];
        const int x0 = (int)R_FLOOR(ix), y0 = (int)R_FLOOR(iy);
        const real fx = ix - (real)x0, fy = iy - (real)y0;
        const int xin = x0 + 1 < W, yin = y0 + 1 < H;
        const real* src = xn + (size_t)(d->source_ids[si] - 1) * C * hw;
        real gix = 0, giy = 0;
        for (int c = 0; c < C; ++c) {
          const real g = s->gwarp[si][(size_t)c * hw + p];
          if (g == 0) continue;
          const real* pl = src + (size_t)c * hw;
          const real nw = pl[(size_t)y0 * W + x0];
          gix -= nw * (1 - fy) * g; giy -= nw * (1 - fx) * g;
          if (xin) { const real ne = pl[(size_t)y0 * W + x0 + 1]; gix += ne * (1 - fy) * g; giy -= ne * fx * g; }
          if (yin) { const real sw = pl[(size_t)(y0 + 1) * W + x0]; gix -= sw * fy * g; giy += sw * (1 - fx) * g; }
          if (xin && yin) { const real se = pl[(size_t)(y0 + 1) * W + x0 + 1]; gix += se * fy * g; giy += se * fx * g; }
        }
        /* clip multiplier, un-normalisation (W-1)/2 and normalize 2/(W-1) cancel to 1 */
        const real gu = s->mx[si][p] ? gix * ((real)(W - 1) * (real)0.5) * (2 / (real)(W - 1)) : 0;
        const real gv = s->my[si][p] ? giy * ((real)(H - 1) * (real)0.5) * (2 / (real)(H - 1)) : 0;
        if (gu == 0 && gv == 0) continue;
        real Y[3], cam[3];
        for (int i = 0; i < 3; ++i)
          Y[i] = (real)P[si].R[3 * i] * X[0] + (real)P[si].R[3 * i + 1] * X[1] + (real)P[si].R[3 * i + 2] * X[2] + (real)P[si].t[i];
        for (int i = 0; i < 3; ++i)
          cam[i] = (real)Km[3 * i] * Y[0] + (real)Km[3 * i + 1] * Y[1] + (real)Km[3 * i + 2] * Y[2];
        const real den = 1 / (cam[2] + (real)1e-7);
        const real gcam[3] = {gu * den, gv * den, -(gu * cam[0] + gv * cam[1]) * den * den};
        real gY[3];
        for (int j = 0; j < 3; ++j) gY[j] = (real)Km[j] * gcam[0] + (real)Km[3 + j] * gcam[1] + (real)Km[6 + j] * gcam[2];
        for (int i = 0; i < 3; ++i) {
          dtacc[si][i] += (double)gY[i];
          for (int j = 0; j < 3; ++j) dRacc[si][3 * i + j] += (double)gY[i] * (double)X[j];
        }
        for (int j = 0; j < 3; ++j)
          gX[j] += (real)P[si].R[j] * gY[0] + (real)P[si].R[3 + j] * gY[1] + (real)P[si].R[6 + j] * gY[2];
      }
      const real gdepth = gX[0] * ray[0] + gX[1] * ray[1] + gX[2] * ray[2];
      s->gup[p] += -depth * depth * gdepth * dd;
    }

  /* normalisation pullback: dn = up / (mean + eps) */
  if (d->normalize) {
    double dot = 0;
    for (size_t p = 0; p < hw; ++p) dot += (double)s->gdn[p] * (double)s->up[p];
    const real cc = (real)(-dot * (double)inv * (double)inv / (double)hw);
    for (size_t p = 0; p < hw; ++p) s->gup[p] += s->gdn[p] * inv + cc;
  } else {
    for (size_t p = 0; p < hw; ++p) s->gup[p] += s->gdn[p];
  }

  /* transpose of upsample_bilinear (or identity) into the native-resolution gradient */
  if (native) {
    for (size_t p = 0; p < hw; ++p) gdisp_k[p] = s->gup[p];
  } else {
    memset(gdisp_k, 0, (size_t)dw * dh * sizeof(real));
    for (int py = 0; py < H; ++py) {
      int y0, y1; real fy;
      up_coord(py, H, dh, d->upsample_half_pixel, &y0, &y1, &fy);
      for (int px = 0; px < W; ++px) {
        int x0, x1; real fx;
        up_coord(px, W, dw, d->upsample_half_pixel, &x0, &x1, &fx);
        const real g = s->gup[(size_t)py * W + px];
        gdisp_k[(size_t)y0 * dw + x0] += g * (1 - fx) * (1 - fy);
        gdisp_k[(size_t)y0 * dw + x1] += g * fx * (1 - fy);
        gdisp_k[(size_t)y1 * dw + x0] += g * (1 - fx) * fy;
        gdisp_k[(size_t)y1 * dw + x1] += g * fx * fy;
      }
    }
  }
  return loss;
}

/* the items are independent: a pool of pthreads pulls item indices from an atomic counter */
typedef struct {
  const mdbo_desc* d;
  const real *x, *rvec, *tvec, *auto_loss;
  const real* const* disp;
  real* const* d_disp;
  real *d_rvec, *d_tvec;
  double Km[9], iKm[9];
  double* item_loss;
  atomic_int next, fail;
} job_t;

static void* worker(void* arg) {
  job_t* J = (job_t*)arg;
  const mdbo_desc* d = J->d;
  const int W = d->W, H = d->H, N = d->N, S = d->S;
  const size_t hw = (size_t)W * H;
  scratch_t sc;
  if (!scratch_alloc(&sc, W, H, d->C, S)) { atomic_store(&J->fail, 1); scratch_free(&sc, S); return NULL; }
  for (;;) {
    const int n = atomic_fetch_add(&J->next, 1);
    if (n >= N) break;
    pose_t P[MDBO_MAX_S];
    double r[MDBO_MAX_S][3], t[MDBO_MAX_S][3];
    double dRacc[MDBO_MAX_S][9], dtacc[MDBO_MAX_S][3];
    memset(dRacc, 0, sizeof(dRacc));
    memset(dtacc, 0, sizeof(dtacc));
    for (int si = 0; si < S; ++si) {
      for (int k = 0; k < 3; ++k) {
        r[si][k] = (double)J->rvec[((size_t)si * N + n) * 3 + k];
        t[si][k] = (double)J->tvec[((size_t)si * N + n) * 3 + k];
      }
      so3_aux a;
      compose_T_rm(r[si], t[si], d->source_ids[si] < d->target_id, P[si].R, P[si].t, &a);   /* src/training.jl:28 */
    }
    double acc = 0;
    for (int k = 0; k < d->Ls; ++k) {
      const real* dk = J->disp[k] + (size_t)n * d->dw[k] * d->dh[k];
      real* gk = J->d_disp ? J->d_disp[k] + (size_t)n * d->dw[k] * d->dh[k] : NULL;
      acc += item_scale(d, n, k, J->x, P, J->Km, J->iKm, dk, J->auto_loss ? J->auto_loss + (size_t)n * hw : NULL,
                        &sc, gk, dRacc, dtacc);
    }
    J->item_loss[n] = acc / d->Ls;                                                         /* src/training.jl:76 */
    if (J->d_disp && J->d_rvec && J->d_tvec)
      for (int si = 0; si < S; ++si) {
        double dr[3], dt[3];
        compose_T_bwd_rm(r[si], t[si], d->source_ids[si] < d->target_id, dRacc[si], dtacc[si], dr, dt);
        for (int k = 0; k < 3; ++k) {
          J->d_rvec[((size_t)si * N + n) * 3 + k] = (real)dr[k];
          J->d_tvec[((size_t)si * N + n) * 3 + k] = (real)dt[k];
        }
      }
  }
  scratch_free(&sc, S);
  return NULL;
}

/* value (and, when d_disp != NULL, gradient) of the multi-scale loss.
 *   x            (W,H,C,L,N) frame stacks
 *   rvec, tvec   [S][N][3]   one Pose per source (src/pose_decoder.jl:1-5)
 *   disp[k]      (w_k,h_k,1,N);  auto_loss (W,H,1,N) or NULL
 *   K_jl,invK_jl 9 numbers each as Julia stores a 3x3
 *   loss_out     1 double:  sum_k (mean(warp_loss_k) + smooth_k) / Ls           src/training.jl:69-76
 *   d_disp[k]    same shape as disp[k];  d_rvec, d_tvec [S][N][3]
 * Returns 0, or -1 on bad arguments / out of memory. */
int mdbo_train_loss(const mdbo_desc* d, const real* x, const real* rvec, const real* tvec,
                    const real* const* disp, const real* auto_loss, const real* K_jl, const real* invK_jl,
                    double* loss_out, real* const* d_disp, real* d_rvec, real* d_tvec, int nthreads) {
  if (!d || !x || !rvec || !tvec || !disp || !K_jl || !invK_jl || !loss_out) return -1;
  if (d->S < 1 || d->S > MDBO_MAX_S || d->Ls < 1 || d->Ls > MDBO_MAX_SCALES || d->W < 2 || d->H < 2 || d->N < 1)
    return -1;
  job_t J;
  memset(&J, 0, sizeof(J));
  J.d = d; J.x = x; J.rvec = rvec; J.tvec = tvec; J.auto_loss = auto_loss; J.disp = disp;
  J.d_disp = d_disp; J.d_rvec = d_rvec; J.d_tvec = d_tvec;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) { J.Km[3 * i + j] = (double)K_jl[3 * j + i]; J.iKm[3 * i + j] = (double)invK_jl[3 * j + i]; }
  J.item_loss = calloc((size_t)d->N, sizeof(double));
  if (!J.item_loss) return -1;
  atomic_init(&J.next, 0);
  atomic_init(&J.fail, 0);
  int nt = nthreads < 1 ? 1 : (nthreads > d->N ? d->N : nthreads);
  if (nt > 256) nt = 256;
  pthread_t th[256];
  int started = 0;
  for (int i = 1; i < nt; ++i)
    if (pthread_create(&th[started], NULL, worker, &J) == 0) ++started;
  worker(&J);
  for (int i = 0; i < started; ++i) pthread_join(th[i], NULL);
  double total = 0;
  for (int n = 0; n < d->N; ++n) total += J.item_loss[n];      /* fixed order: deterministic */
  free(J.item_loss);
  *loss_out = total;
  return atomic_load(&J.fail) ? -1 : 0;
}

int mdbo_sizeof_real(void) { return (int)sizeof(real); }

/* ==========================================================================================
 * plane-disparity volume rendering -- src/render.jl, quirks reproduced as written.
 * Second, independent restatement next to oracle/torch_ref.py::render_tgt_rgb_depth (autograd):
 * the pullback below is written by hand.
 * Layouts (Julia column-major == these C orders): rgb [B][Np][3][H][W], sigma [B][Np][1][H][W],
 * xyz_tgt [B][Np][H][W][3], disparity_src [B][Np], rvec [B][3], tvec [B][3];
 * rgb_out [B][3][H][W], depth_out = transparency_acc [B][Np][1][H][W], mask [B][1][1][H][W] int32.
 * ======================================================================================== */
static int inv3(const double* A, double* B) {          /* Monodepth.inv, src/render.jl:3-17 */
  const double c00 = A[4] * A[8] - A[5] * A[7], c01 = A[5] * A[6] - A[3] * A[8], c02 = A[3] * A[7] - A[4] * A[6];
  const double det = A[0] * c00 + A[1] * c01 + A[2] * c02, id = 1.0 / det;
  B[0] = c00 * id; B[1] = (A[2] * A[7] - A[1] * A[8]) * id; B[2] = (A[1] * A[5] - A[2] * A[4]) * id;
  B[3] = c01 * id; B[4] = (A[0] * A[8] - A[2] * A[6]) * id; B[5] = (A[2] * A[3] - A[0] * A[5]) * id;
  B[6] = c02 * id; B[7] = (A[1] * A[6] - A[0] * A[7]) * id; B[8] = (A[0] * A[4] - A[1] * A[3]) * id;
  return det != 0.0;
}

typedef struct {
  int W, H, Np, B;
  const real *rgb, *sigma, *xyz, *disp;
  const double *Km, *iKm;               /* math row-major */
  const real *rvec, *tvec;
  real *rgb_out, *tacc;
  int32_t* mask;
  /* backward (NULL for forward only) */
  const real *g_out, *g_tacc;
  real *d_rgb, *d_sigma, *d_xyz;        /* zero-initialised by the driver */
  double* Gacc;                         /* [B][Np][9]: d loss / d H_src_tgt, math row-major */
  real* gs;                             /* [B][Np][H W][9]: per-sample grads (7 channels) + d/d(cx), d/d(cy) */
  /* variants */
  int flags;                            /* MDBO_RENDER_* */
  double depth_eps;
  real* dsyn;                           /* [B][1][H][W] expected depth, or NULL */
  const real* g_dsyn;                   /* its cotangent, or NULL */
  atomic_int next;
} rjob_t;

/* The reference as written is flags = 0.  The variants restate what the reference's own comparison
 * script holds up as the intended behaviour (MINE's HomographySample.sample, quoted in
 * test/test_sample.jl:123-137) and the TODOs of src/render.jl:22,36,40,47,52:
 *   MINE_GRID  grid = (x + 0.5) / (W/2) - 1, grid_sample(..., padding_mode="border", align_corners=false)
 *              (test/test_sample.jl:131-137): the sample lands at source pixel x itself
 *   MINE_MASK  valid = (x < W) & (x > -1) & (y < H) & (y > -1)              (test/test_sample.jl:123-125)
 *   FUSED_XYZ  xyz_tgt is not an input: it is what get_src/get_tgt_xyz_from_plane_disparity
 *              (src/render.jl:25-30, 51-62) make of create_meshgrid (:21-23), disparity_src, K_inv, pose
 * and, independent of the flags, the expected depth  sum_k w_k z_k / (sum_k w_k + depth_eps)  that
 * src/render.jl:47 leaves as "TODO: return depth?" (MINE's weighted_sum_mpi).                       */
#define MDBO_RENDER_MINE_GRID 1
#define MDBO_RENDER_MINE_MASK 2
#define MDBO_RENDER_FUSED_XYZ 4

typedef struct { int o[4]; real fx, fy, mx, my; double X[3], x, y, sx, sy; int valid; } rtap_t;

/* sample coordinates of target pixel (u,v) (0-based) under H_src_tgt    src/render.jl:76-88 */
static rtap_t render_tap(const double* Hst, int u, int v, int W, int H, int flags) {
  rtap_t t;
  for (int i = 0; i < 3; ++i) t.X[i] = Hst[3 * i] * u + Hst[3 * i + 1] * v + Hst[3 * i + 2];
  t.x = t.X[0] / t.X[2]; t.y = t.X[1] / t.X[2];
  if (flags & MDBO_RENDER_MINE_MASK)
    t.valid = (t.x < W) && (t.x > -1.0) && (t.y < H) && (t.y > -1.0);
  else
    /* Julia parses `a .< W .* a .>= 0` as (a < W a) & (W a >= 0)          src/render.jl:81, SURVEY F8c */
    t.valid = (t.x < W * t.x) && (W * t.x >= 0.0) && (t.y < H * t.y) && (H * t.y >= 0.0);
  double ix, iy;
  if (flags & MDBO_RENDER_MINE_GRID) {
    /* g = (x + 0.5) / (W/2) - 1;  align_corners=false un-normalisation ((g + 1) W - 1) / 2 */
    const double gx = (t.x + 0.5) / (W / 2.0) - 1.0, gy = (t.y + 0.5) / (H / 2.0) - 1.0;
    ix = ((gx + 1.0) * W - 1.0) / 2.0; iy = ((gy + 1.0) * H - 1.0) / 2.0;
    t.sx = 1.0; t.sy = 1.0;
  } else {
    /* grid = (xy + 0.5) / (W/2) without "-1" (:83-84), then grid_sample's align-corners un-normalisation */
    ix = ((t.x + 0.5) / (W / 2.0) + 1.0) * 0.5 * (W - 1); iy = ((t.y + 0.5) / (H / 2.0) + 1.0) * 0.5 * (H - 1);
    t.sx = (double)(W - 1) / (double)W; t.sy = (double)(H - 1) / (double)H;
  }
  t.mx = (ix > 0.0 && ix < (double)(W - 1)) ? 1 : 0;       /* clamp multiplier of the sampling gradient */
  t.my = (iy > 0.0 && iy < (double)(H - 1)) ? 1 : 0;
  const double cx = fmin(fmax(ix, 0.0), (double)(W - 1)), cy = fmin(fmax(iy, 0.0), (double)(H - 1));
  const int x0 = (int)floor(cx), y0 = (int)floor(cy);
  const int x1 = x0 + 1 < W ? x0 + 1 : W - 1, y1 = y0 + 1 < H ? y0 + 1 : H - 1;
  t.fx = (real)(cx - x0); t.fy = (real)(cy - y0);
  t.o[0] = y0 * W + x0; t.o[1] = y0 * W + x1; t.o[2] = y1 * W + x0; t.o[3] = y1 * W + x1;
  return t;
}

static inline real bil4(const rtap_t* t, real a, real b, real c, real d) {
  const real top = a + t->fx * (b - a), bot = c + t->fx * (d - c);
  return top + t->fy * (bot - top);
}

static void plane_homography(const rjob_t* J, int b, int k, double* Hst) {   /* src/render.jl:66-74 */
  double r[3] = {(double)J->rvec[3 * b], (double)J->rvec[3 * b + 1], (double)J->rvec[3 * b + 2]}, R[9], M[9], KM[9], Hts[9];
  so3_aux a;
  so3_exp_rm(r, R, &a);
  const double dsp = (double)J->disp[b * J->Np + k];      /* 1 / depth_src, depth_src = 1 ./ disparity_src (:93) */
  memcpy(M, R, sizeof(M));
  for (int i = 0; i < 3; ++i) M[3 * i + 2] += (double)J->tvec[3 * b + i] * dsp;   /* R - t n^T / (-depth) */
  mat3_mul(J->Km, M, KM);
  mat3_mul(KM, J->iKm, Hts);
  inv3(Hts, Hst);
}

/* one (b, row) unit of work: forward compositing and, if asked, the per-sample pullback */
static void render_row(rjob_t* J, int b, int v, const double* Hst_all) {
  const int W = J->W, H = J->H, Np = J->Np, WH = W * H;
  real* s7 = malloc((size_t)Np * 7 * sizeof(real));
  real* tacc = malloc((size_t)(Np + 1) * sizeof(real));
  real* Tk = malloc((size_t)Np * sizeof(real));
  real* dist = malloc((size_t)Np * sizeof(real));
  rtap_t* taps = malloc((size_t)Np * sizeof(rtap_t));
  for (int u = 0; u < W; ++u) {
    const int pix = v * W + u;
    int nvalid = 0;
    for (int k = 0; k < Np; ++k) {
      const int kb = b * Np + k;
      taps[k] = render_tap(Hst_all + 9 * kb, u, v, W, H, J->flags);
      const rtap_t* t = &taps[k];
      nvalid += t->valid;
      const real* pr = J->rgb + (size_t)kb * 3 * WH;
      for (int c = 0; c < 3; ++c) s7[7 * k + c] = bil4(t, pr[c * WH + t->o[0]], pr[c * WH + t->o[1]], pr[c * WH + t->o[2]], pr[c * WH + t->o[3]]);
      const real* ps = J->sigma + (size_t)kb * WH;
      s7[7 * k + 3] = bil4(t, ps[t->o[0]], ps[t->o[1]], ps[t->o[2]], ps[t->o[3]]);
      const real* px = J->xyz + (size_t)kb * WH * 3;
      for (int c = 0; c < 3; ++c) s7[7 * k + 4 + c] = bil4(t, px[3 * t->o[0] + c], px[3 * t->o[1] + c], px[3 * t->o[2] + c], px[3 * t->o[3] + c]);
    }
    /* plane_volume_rendering                                              src/render.jl:32-49 */
    real out[3] = {0, 0, 0}, wsum = 0, wz = 0;
    tacc[0] = 1;
    for (int k = 0; k < Np; ++k) {
      real d = (real)1e3;
      if (k + 1 < Np) {
        const real dx = s7[7 * (k + 1) + 4] - s7[7 * k + 4], dy = s7[7 * (k + 1) + 5] - s7[7 * k + 5], dz = s7[7 * (k + 1) + 6] - s7[7 * k + 6];
        d = (real)sqrt((double)(dx * dx + dy * dy + dz * dz));
      }
      dist[k] = d;
      const real sg = s7[7 * k + 3] >= 0 ? s7[7 * k + 3] : 0;              /* sigma .* (sigma .>= 0), :101 */
      Tk[k] = R_EXP(-d * sg);
      const real w = tacc[k] * (1 - Tk[k]);
      for (int c = 0; c < 3; ++c) out[c] += w * s7[7 * k + c];
      wsum += w; wz += w * s7[7 * k + 6];
      J->tacc[(size_t)(b * Np + k) * WH + pix] = tacc[k];
      tacc[k + 1] = tacc[k] * (Tk[k] + (real)1e-6);
    }
    for (int c = 0; c < 3; ++c) J->rgb_out[((size_t)b * 3 + c) * WH + pix] = out[c];
    J->mask[(size_t)b * WH + pix] = nvalid;
    const real Dn = wsum + (real)J->depth_eps, dsyn = wz / Dn;
    if (J->dsyn) J->dsyn[(size_t)b * WH + pix] = dsyn;
    if (!J->g_out) continue;
    const real gds = J->g_dsyn ? J->g_dsyn[(size_t)b * WH + pix] : 0;

    /* ---- pullback of the compositing, planes in reverse ---- */
    real go[3];
    for (int c = 0; c < 3; ++c) go[c] = J->g_out[((size_t)b * 3 + c) * WH + pix];
    real gT_next = 0;                   /* d loss / d tacc[k+1] */
    real gxyz_next[3] = {0, 0, 0};      /* d loss / d xyz_{k+1} from dist_k */
    for (int k = Np - 1; k >= 0; --k) {
      const real sgraw = s7[7 * k + 3], sg = sgraw >= 0 ? sgraw : 0;
      const real alpha = 1 - Tk[k];
      real gw = 0;
      for (int c = 0; c < 3; ++c) gw += go[c] * s7[7 * k + c];
      gw += gds * (s7[7 * k + 6] - dsyn) / Dn;                             /* depth = sum w z / (sum w + eps) */
      real* g = J->gs + ((size_t)(b * Np + k) * WH + pix) * 9;
      for (int c = 0; c < 3; ++c) g[c] = go[c] * tacc[k] * alpha;          /* d/d rgb_k = g_out w_k */
      const real gtk = (J->g_tacc ? J->g_tacc[(size_t)(b * Np + k) * WH + pix] : 0) + gw * alpha + gT_next * (Tk[k] + (real)1e-6);
      const real gTk = -gw * tacc[k] + gT_next * tacc[k];                   /* d/d T_k */
      const real ge = gTk * Tk[k];                                          /* d/d(-dist sigma) */
      g[3] = sgraw >= 0 ? -ge * dist[k] : 0;
      real gxyz[3] = {gxyz_next[0], gxyz_next[1], gxyz_next[2]};
      gxyz_next[0] = gxyz_next[1] = gxyz_next[2] = 0;
      if (k + 1 < Np) {
        const real gd = -ge * sg;
        for (int c = 0; c < 3; ++c) {
          const real dv = (s7[7 * (k + 1) + 4 + c] - s7[7 * k + 4 + c]) / dist[k];
          /* xyz_{k+1} was handled one iteration ago: add to its stored slot */
          J->gs[((size_t)(b * Np + k + 1) * WH + pix) * 9 + 4 + c] += gd * dv;
          gxyz[c] -= gd * dv;
        }
      }
      gxyz[2] += gds * tacc[k] * alpha / Dn;                                /* d depth / d z_k = w_k / D */
      for (int c = 0; c < 3; ++c) g[4 + c] = gxyz[c];
      gT_next = gtk;
    }
    /* ---- per-sample: d/d(cx), d/d(cy) from the 7 channel grads (needs the FINAL g[4..6]) ---- */
    for (int k = 0; k < Np; ++k) {
      const int kb = b * Np + k;
      const rtap_t* t = &taps[k];
      real* g = J->gs + ((size_t)kb * WH + pix) * 9;
      real gx = 0, gy = 0;
      const real* pr = J->rgb + (size_t)kb * 3 * WH;
      const real* ps = J->sigma + (size_t)kb * WH;
      const real* px = J->xyz + (size_t)kb * WH * 3;
      for (int ch = 0; ch < 7; ++ch) {
        real a, bq, c, dq;
        if (ch < 3) { a = pr[ch * WH + t->o[0]]; bq = pr[ch * WH + t->o[1]]; c = pr[ch * WH + t->o[2]]; dq = pr[ch * WH + t->o[3]]; }
        else if (ch == 3) { a = ps[t->o[0]]; bq = ps[t->o[1]]; c = ps[t->o[2]]; dq = ps[t->o[3]]; }
        else { const int c3 = ch - 4; a = px[3 * t->o[0] + c3]; bq = px[3 * t->o[1] + c3]; c = px[3 * t->o[2] + c3]; dq = px[3 * t->o[3] + c3]; }
        gx += g[ch] * ((bq - a) * (1 - t->fy) + (dq - c) * t->fy);
        gy += g[ch] * ((c - a) * (1 - t->fx) + (dq - bq) * t->fx);
      }
      g[7] = gx * t->mx;
      g[8] = gy * t->my;
    }
  }
  free(s7); free(tacc); free(Tk); free(dist); free(taps);
}

/* phase 2, one (b, k) plane: transpose of the bilinear gather into d_rgb / d_sigma / d_xyz (this plane's
 * arrays only: race-free across planes) and the sum G = d loss / d H_src_tgt over the plane's pixels */
static void render_scatter_plane(rjob_t* J, int kb, const double* Hst) {
  const int W = J->W, H = J->H, WH = W * H;
  real* dr = J->d_rgb + (size_t)kb * 3 * WH;
  real* ds = J->d_sigma + (size_t)kb * WH;
  real* dx = J->d_xyz + (size_t)kb * WH * 3;
  double G[9] = {0};
  for (int v = 0; v < H; ++v)
    for (int u = 0; u < W; ++u) {
      const rtap_t t = render_tap(Hst, u, v, W, H, J->flags);
      const real* g = J->gs + ((size_t)kb * WH + v * W + u) * 9;
      const real w[4] = {(1 - t.fx) * (1 - t.fy), t.fx * (1 - t.fy), (1 - t.fx) * t.fy, t.fx * t.fy};
      for (int q = 0; q < 4; ++q) {
        for (int c = 0; c < 3; ++c) dr[c * WH + t.o[q]] += g[c] * w[q];
        ds[t.o[q]] += g[3] * w[q];
        for (int c = 0; c < 3; ++c) dx[3 * t.o[q] + c] += g[4 + c] * w[q];
      }
      /* (cx, cy) <- (x, y): d ix / d x = (W-1)/W (no "-1" grid, align-corners un-normalisation), 1 for the MINE grid */
      const double gx = (double)g[7] * t.sx, gy = (double)g[8] * t.sy;
      if (gx == 0.0 && gy == 0.0) continue;
      const double gX[3] = {gx / t.X[2], gy / t.X[2], -(gx * t.x + gy * t.y) / t.X[2]};
      const double q3[3] = {(double)u, (double)v, 1.0};
      for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) G[3 * i + j] += gX[i] * q3[j];
    }
  memcpy(J->Gacc + 9 * kb, G, sizeof(G));
}

typedef struct { rjob_t* J; const double* Hst; int phase; } rwork_t;

static void* render_worker(void* arg) {
  rwork_t* Wk = (rwork_t*)arg;
  rjob_t* J = Wk->J;
  const int units = Wk->phase == 1 ? J->B * J->H : J->B * J->Np;
  for (;;) {
    const int i = atomic_fetch_add(&J->next, 1);
    if (i >= units) break;
    if (Wk->phase == 1) render_row(J, i / J->H, i % J->H, Wk->Hst);
    else render_scatter_plane(J, i, Wk->Hst + 9 * i);
  }
  return NULL;
}

static void render_run(rjob_t* J, const double* Hst, int phase, int nthreads) {
  rwork_t Wk = {J, Hst, phase};
  atomic_store(&J->next, 0);
  int nt = nthreads < 1 ? 1 : (nthreads > 256 ? 256 : nthreads);
  pthread_t th[256];
  int started = 0;
  for (int i = 1; i < nt; ++i)
    if (pthread_create(&th[started], NULL, render_worker, &Wk) == 0) ++started;
  render_worker(&Wk);
  for (int i = 0; i < started; ++i) pthread_join(th[i], NULL);
}

/* create_meshgrid -> get_src_xyz_from_plane_disparity -> get_tgt_xyz_from_plane_disparity
 * (src/render.jl:21-23, 25-30, 51-62): xyz_tgt [B][Np][H][W][3].  xyz_src (same layout) is kept when asked. */
static void plane_xyz(int W, int H, int Np, int B, const real* disp, const real* rvec, const real* tvec, const double* iKm,
                      real* xyz_src, real* xyz_tgt) {
  for (int b = 0; b < B; ++b) {
    double r[3] = {(double)rvec[3 * b], (double)rvec[3 * b + 1], (double)rvec[3 * b + 2]}, R[9];
    so3_aux a;
    so3_exp_rm(r, R, &a);
    for (int k = 0; k < Np; ++k) {
      const real depth = (real)1 / disp[b * Np + k];                          /* :28 */
      for (int v = 0; v < H; ++v)
        for (int u = 0; u < W; ++u) {
          const real m[3] = {(real)(u + 1), (real)(v + 1), (real)1};         /* 1-based meshgrid, :22 */
          real xs[3];
          for (int i = 0; i < 3; ++i) xs[i] = ((real)iKm[3 * i] * m[0] + (real)iKm[3 * i + 1] * m[1] + (real)iKm[3 * i + 2] * m[2]) * depth;
          const size_t o = (((size_t)(b * Np + k) * H + v) * W + u) * 3;
          if (xyz_src) for (int i = 0; i < 3; ++i) xyz_src[o + i] = xs[i];
          for (int i = 0; i < 3; ++i)                                         /* R xyz_src + t, :60 */
            xyz_tgt[o + i] = (real)R[3 * i] * xs[0] + (real)R[3 * i + 1] * xs[1] + (real)R[3 * i + 2] * xs[2] + tvec[3 * b + i];
        }
    }
  }
}

/* render_tgt_rgb_depth (src/render.jl:92-114) forward and, when g_rgb_out != NULL, its pullback.
 * K_jl / Kinv_jl: the 9 numbers Julia stores.  g_depth may be NULL.  All d_* must be given for the
 * pullback (d_disp [B][Np], d_rvec [B][3], d_tvec [B][3]).  Returns 0, -1 on bad arguments / OOM.
 * `flags`, `depth_eps`, `dsyn_out` / `g_dsyn`: the variants described at MDBO_RENDER_* above.  With
 * FUSED_XYZ, xyz_tgt and d_xyz are NULL and the pose / disparity gradients include the path through xyz. */
int mdbo_render_tgt_rgb_depth_ex(int W, int H, int Np, int B, const real* rgb, const real* sigma, const real* disparity_src,
                                 const real* xyz_tgt, const real* rvec, const real* tvec, const real* Kinv_jl, const real* K_jl,
                                 int flags, double depth_eps, real* rgb_out, real* depth_out, int32_t* mask_out, real* dsyn_out,
                                 const real* g_rgb_out, const real* g_depth, const real* g_dsyn,
                                 real* d_rgb, real* d_sigma, real* d_disp, real* d_xyz, real* d_rvec, real* d_tvec, int nthreads) {
  const int fused = (flags & MDBO_RENDER_FUSED_XYZ) != 0;
  if (W < 2 || H < 2 || Np < 1 || B < 1 || !rgb || !sigma || !disparity_src || !rvec || !tvec || !rgb_out ||
      !depth_out || !mask_out || (!fused && !xyz_tgt))
    return -1;
  if (g_rgb_out && (!d_rgb || !d_sigma || !d_disp || !d_rvec || !d_tvec || (!fused && !d_xyz))) return -1;
  double Km[9], iKm[9];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) { Km[3 * i + j] = (double)K_jl[3 * j + i]; iKm[3 * i + j] = (double)Kinv_jl[3 * j + i]; }
  const size_t WH = (size_t)W * H, planes = (size_t)B * Np;
  real *xs_own = NULL, *xt_own = NULL, *dx_own = NULL;
  if (fused) {
    xs_own = malloc(planes * WH * 3 * sizeof(real));
    xt_own = malloc(planes * WH * 3 * sizeof(real));
    if (g_rgb_out) dx_own = malloc(planes * WH * 3 * sizeof(real));
    if (!xs_own || !xt_own || (g_rgb_out && !dx_own)) { free(xs_own); free(xt_own); free(dx_own); return -1; }
    plane_xyz(W, H, Np, B, disparity_src, rvec, tvec, iKm, xs_own, xt_own);
    xyz_tgt = xt_own;
    d_xyz = dx_own;
  }
  rjob_t J;
  memset(&J, 0, sizeof(J));
  J.W = W; J.H = H; J.Np = Np; J.B = B;
  J.rgb = rgb; J.sigma = sigma; J.xyz = xyz_tgt; J.disp = disparity_src; J.Km = Km; J.iKm = iKm; J.rvec = rvec; J.tvec = tvec;
  J.rgb_out = rgb_out; J.tacc = depth_out; J.mask = mask_out;
  J.flags = flags; J.depth_eps = depth_eps; J.dsyn = dsyn_out; J.g_dsyn = g_dsyn;
  double* Hst = malloc(planes * 9 * sizeof(double));
  if (!Hst) { free(xs_own); free(xt_own); free(dx_own); return -1; }
  for (int b = 0; b < B; ++b)
    for (int k = 0; k < Np; ++k) plane_homography(&J, b, k, Hst + 9 * (b * Np + k));
  int rc = 0;
  if (g_rgb_out) {
    J.g_out = g_rgb_out; J.g_tacc = g_depth; J.d_rgb = d_rgb; J.d_sigma = d_sigma; J.d_xyz = d_xyz;
    J.gs = malloc(planes * WH * 9 * sizeof(real));
    J.Gacc = calloc(planes * 9, sizeof(double));
    if (!J.gs || !J.Gacc) { free(J.gs); free(J.Gacc); free(Hst); free(xs_own); free(xt_own); free(dx_own); return -1; }
    memset(d_rgb, 0, planes * 3 * WH * sizeof(real));
    memset(d_sigma, 0, planes * WH * sizeof(real));
    memset(d_xyz, 0, planes * 3 * WH * sizeof(real));
  }
  atomic_init(&J.next, 0);
  render_run(&J, Hst, 1, nthreads);
  if (g_rgb_out) {
    render_run(&J, Hst, 2, nthreads);
    /* H_src_tgt = inv(H_tgt_src), H_tgt_src = K M K^-1, M = R + t n^T disparity    src/render.jl:66-74 */
    for (int b = 0; b < B; ++b) {
      double gR[9] = {0}, gt[3] = {0};
      double r[3] = {(double)rvec[3 * b], (double)rvec[3 * b + 1], (double)rvec[3 * b + 2]}, R[9], dr[3];
      so3_aux a;
      so3_exp_rm(r, R, &a);
      for (int k = 0; k < Np; ++k) {
        const double* Hs = Hst + 9 * (b * Np + k);
        const double* G = J.Gacc + 9 * (b * Np + k);
        double HT[9], t1[9], gHts[9], KT[9], iKT[9], t2[9], gM[9];
        for (int i = 0; i < 3; ++i)
          for (int j = 0; j < 3; ++j) { HT[3 * i + j] = Hs[3 * j + i]; KT[3 * i + j] = Km[3 * j + i]; iKT[3 * i + j] = iKm[3 * j + i]; }
        mat3_mul(HT, G, t1);
        mat3_mul(t1, HT, gHts);
        for (int q = 0; q < 9; ++q) gHts[q] = -gHts[q];              /* d inv(A) = -A^-T G A^-T */
        mat3_mul(KT, gHts, t2);
        mat3_mul(t2, iKT, gM);
        const double dsp = (double)disparity_src[b * Np + k];
        double gd = 0;
        for (int i = 0; i < 3; ++i) {
          for (int j = 0; j < 3; ++j) gR[3 * i + j] += gM[3 * i + j];
          gt[i] += gM[3 * i + 2] * dsp;
          gd += gM[3 * i + 2] * (double)tvec[3 * b + i];
        }
        if (fused) {
          /* xyz_tgt = R xyz_src + t (:60), xyz_src = K_inv m / disparity (:28-29) */
          const real* gx = d_xyz + (size_t)(b * Np + k) * WH * 3;
          const real* xs = xs_own + (size_t)(b * Np + k) * WH * 3;
          double sd = 0;
          for (size_t p = 0; p < WH; ++p) {
            const double g3[3] = {(double)gx[3 * p], (double)gx[3 * p + 1], (double)gx[3 * p + 2]};
            const double x3[3] = {(double)xs[3 * p], (double)xs[3 * p + 1], (double)xs[3 * p + 2]};
            for (int i = 0; i < 3; ++i) {
              gt[i] += g3[i];
              for (int j = 0; j < 3; ++j) gR[3 * i + j] += g3[i] * x3[j];
            }
            for (int j = 0; j < 3; ++j) sd += (R[j] * g3[0] + R[3 + j] * g3[1] + R[6 + j] * g3[2]) * x3[j];   /* (R^T g) . xyz_src */
          }
          gd -= sd / dsp;
        }
        d_disp[b * Np + k] = (real)gd;
      }
      so3_exp_bwd_rm(r, &a, gR, dr);
      for (int i = 0; i < 3; ++i) { d_rvec[3 * b + i] = (real)dr[i]; d_tvec[3 * b + i] = (real)gt[i]; }
    }
    free(J.gs); free(J.Gacc);
  }
  free(Hst); free(xs_own); free(xt_own); free(dx_own);
  return rc;
}

int mdbo_render_tgt_rgb_depth(int W, int H, int Np, int B, const real* rgb, const real* sigma, const real* disparity_src,
                              const real* xyz_tgt, const real* rvec, const real* tvec, const real* Kinv_jl, const real* K_jl,
                              real* rgb_out, real* depth_out, int32_t* mask_out, const real* g_rgb_out, const real* g_depth,
                              real* d_rgb, real* d_sigma, real* d_disp, real* d_xyz, real* d_rvec, real* d_tvec, int nthreads) {
  return mdbo_render_tgt_rgb_depth_ex(W, H, Np, B, rgb, sigma, disparity_src, xyz_tgt, rvec, tvec, Kinv_jl, K_jl, 0, 0.0, rgb_out,
                                      depth_out, mask_out, NULL, g_rgb_out, g_depth, NULL, d_rgb, d_sigma, d_disp, d_xyz, d_rvec,
                                      d_tvec, nthreads);
}

/* the producers alone, for the tests of mdb_plane_xyz_src / _tgt */
int mdbo_plane_xyz(int W, int H, int Np, int B, const real* disparity_src, const real* rvec, const real* tvec,
                   const real* Kinv_jl, real* xyz_src, real* xyz_tgt) {
  if (W < 1 || H < 1 || Np < 1 || B < 1 || !disparity_src || !rvec || !tvec || !Kinv_jl || !xyz_tgt) return -1;
  double iKm[9];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) iKm[3 * i + j] = (double)Kinv_jl[3 * j + i];
  plane_xyz(W, H, Np, B, disparity_src, rvec, tvec, iKm, xyz_src, xyz_tgt);
  return 0;
}
