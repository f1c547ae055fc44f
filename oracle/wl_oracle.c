/*
 * oracle/wl_oracle.c -- TEST INFRASTRUCTURE ONLY (see oracle/oracle.h).
 *
 * Restatement of the wrench-linearisation stage:
 *   quadratic wrench model   uprightmpc2/funapprox.c:35-96 (== wlcontroller/c/funapprox.c:15-56)
 *   WLCon projected gradient uprightmpc2/funapprox.c:98-165
 *   true WLQP                wlcontroller/c/wlqp.c:17-123, wlcontroller.c:17-45
 *                            (C++ twin wlcontroller/cpp/wlqp.cpp:24-78: check_termination=0)
 */
#include <math.h>
#include <string.h>
#include "oracle.h"

#define NDU 4
#define NW 6

typedef struct { real a0, a1[NDU], A2[NDU * NDU]; } fa_t;

typedef struct {
  fa_t fa[NW];
  real u0[NDU], umin[NDU], umax[NDU], dumax[NDU], Qw[NW];
  /* true-WLQP solver state */
  oq_set_t set;
  oq_t qp;
} wl_t;

static void fa_init(fa_t *f, const real *popts) {
  int i, j, kk = 0;
  f->a0 = popts[0];
  for (i = 0; i < NDU; ++i) f->a1[i] = popts[1 + i];
  for (i = 0; i < NDU; ++i)
    for (j = i; j < NDU; ++j) { f->A2[i + j * NDU] = f->A2[j + i * NDU] = popts[NDU + 1 + kk]; ++kk; }
}

static real fa_f(const fa_t *f, const real *x) {
  int i, j; real res = f->a0, d = 0, v[NDU], h = 0;
  for (i = 0; i < NDU; ++i) d += x[i] * f->a1[i];
  res += d;
  for (i = 0; i < NDU; ++i) { real a = 0; for (j = 0; j < NDU; ++j) a += f->A2[i + j * NDU] * x[j]; v[i] = a; }
  for (i = 0; i < NDU; ++i) h += x[i] * v[i];
  return res + (real)0.5 * h;
}

static void fa_df(const fa_t *f, const real *x, real *df) {
  int i, j;
  for (i = 0; i < NDU; ++i) { real a = 0; for (j = 0; j < NDU; ++j) a += f->A2[i + j * NDU] * x[j]; df[i] = f->a1[i] + a; }
}

int orc_wl_sizeof(void) { return (int)sizeof(wl_t); }

/* wlConInit argument order (funapprox.h:45); dumax is per second, divided by controlRate */
void orc_wl_init(wl_t *wl, const double *u0, const double *umin, const double *umax, const double *dumax,
                 const double *Qw, double controlRate, const double *popts90) {
  int i; real po[90];
  real P0[16], A0[16];
  memset(wl, 0, sizeof(*wl));
  for (i = 0; i < 90; ++i) po[i] = (real)popts90[i];
  for (i = 0; i < NW; ++i) { fa_init(&wl->fa[i], &po[15 * i]); wl->Qw[i] = (real)Qw[i]; }
  for (i = 0; i < NDU; ++i) {
    wl->u0[i] = (real)u0[i]; wl->umin[i] = (real)umin[i]; wl->umax[i] = (real)umax[i];
    wl->dumax[i] = (real)dumax[i] / (real)controlRate;
  }
  /* wlqp/workspace.c: n=m=4, P dense (placeholder ones, genqp.py:8-17 style), A = I */
  wl->set.rho0 = (real)0.1; wl->set.sigma = (real)1e-6; wl->set.alpha = (real)1.6;
  wl->set.scaling_iters = 10; wl->set.adaptive_rho_interval = 0; wl->set.adaptive_rho_tol = 5;
  /* wlqp.c:43-45: osqp_set_default_settings (eps 1e-3, check 25) then max_iter 40 */
  wl->set.eps_abs = wl->set.eps_rel = (real)1e-3;
  wl->set.max_iter = 40; wl->set.check_termination = 25;
  for (i = 0; i < 16; ++i) { P0[i] = 0; A0[i] = 0; }
  for (i = 0; i < 4; ++i) { P0[i * 4 + i] = 1; A0[i * 4 + i] = 1; }
  oq_init(&wl->qp, &wl->set, 4, 4, P0, A0);
}
void orc_wl_set_check_termination(wl_t *wl, int k) { wl->set.check_termination = k; }
void orc_wl_set_u0(wl_t *wl, const double *u0) { int i; for (i = 0; i < NDU; ++i) wl->u0[i] = (real)u0[i]; }

static void wl_common(const wl_t *wl, const real *u0, const double *h0, const double *pdotdes,
                      real *w0, real *A1 /* 6x4 col-major */, real *a0, real *L, real *U) {
  int i, j; real df[NDU];
  for (i = 0; i < NW; ++i) {
    w0[i] = fa_f(&wl->fa[i], u0);
    fa_df(&wl->fa[i], u0, df);
    for (j = 0; j < NDU; ++j) A1[i + j * NW] = df[j];
    a0[i] = w0[i] - (real)h0[i] - (real)pdotdes[i];
  }
  for (i = 0; i < NDU; ++i) {
    L[i] = -wl->dumax[i]; U[i] = wl->dumax[i];
    if (u0[i] < wl->umin[i]) L[i] = 0;
    else if (u0[i] > wl->umax[i]) U[i] = 0;
  }
}

/* wlConUpdate: one clipped gradient step, stateful in u0 */
void orc_wlcon_update(wl_t *wl, double *u1, double *w0o, const double *h0, const double *pdotdes) {
  real w0[NW], A1[NW * NDU], a0[NW], L[NDU], U[NDU];
  int i, j;
  wl_common(wl, wl->u0, h0, pdotdes, w0, A1, a0, L, U);
  for (j = 0; j < NDU; ++j) {
    real acc = 0, d;
    for (i = 0; i < NW; ++i) acc += A1[i + j * NW] * (wl->Qw[i] * a0[i]);
    d = (real)-1e3 * acc;
    if (d < L[j]) d = L[j]; else if (d > U[j]) d = U[j];
    wl->u0[j] += d;
    u1[j] = (double)wl->u0[j];
  }
  for (i = 0; i < NW; ++i) w0o[i] = (double)w0[i];
}

/* wlqpUpdate + wlControllerUpdate: u = u0 + argmin 0.5 du'P du + q'du, L<=du<=U */
void orc_wlqp_update(wl_t *wl, double *u1, double *w0o, const double *h0, const double *pdotdes) {
  real w0[NW], A1[NW * NDU], a0[NW], L[NDU], U[NDU], P[16], q[NDU];
  int i, j, k;
  oq_t *qp = &wl->qp;
  wl_common(wl, wl->u0, h0, pdotdes, w0, A1, a0, L, U);
  for (i = 0; i < NDU; ++i) {
    for (j = 0; j < NDU; ++j) {
      real acc = 0;
      for (k = 0; k < NW; ++k) acc += A1[k + i * NW] * (wl->Qw[k] * A1[k + j * NW]);
      P[i * 4 + j] = acc;
    }
    { real acc = 0; for (k = 0; k < NW; ++k) acc += A1[k + i * NW] * (wl->Qw[k] * a0[k]); q[i] = acc; }
  }
  /* wlqp.c:64-68 order: update_P (re-equilibrates with the PREVIOUS q), lin_cost, bounds */
  oq_unscale_data(qp);
  for (i = 0; i < NDU; ++i) for (j = 0; j < NDU; ++j) qp->P[i * OQ_NMAX + j] = i <= j ? P[i * 4 + j] : P[j * 4 + i];
  oq_scale_and_factor(qp, &wl->set);
  oq_update_lin_cost(qp, q);
  oq_update_bounds(qp, &wl->set, L, U);
  oq_solve(qp, &wl->set);
  for (i = 0; i < NDU; ++i) { wl->u0[i] = qp->sol_x[i] + wl->u0[i]; u1[i] = (double)wl->u0[i]; }
  for (i = 0; i < NW; ++i) w0o[i] = (double)w0[i];
}

void orc_wl_get_info(const wl_t *wl, double *info4) {
  info4[0] = wl->qp.iter; info4[1] = wl->qp.status; info4[2] = wl->qp.pri_res; info4[3] = wl->qp.dua_res;
}
