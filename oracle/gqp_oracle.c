/*
 * oracle/gqp_oracle.c -- TEST INFRASTRUCTURE ONLY (see oracle/oracle.h).
 *
 * A generic "OSQP object" on top of the restated solver of osqp_oracle.c (pinned to the reference's vendored OSQP by
 * tests/test_oracle_vs_ref.py): setup once, then update(P, A, q, l, u) + solve, the call sequence of the reference's
 * Python controllers (template_controllers.py:190-191,216-217; uprightmpc2_withWLQP.py:243-244,283-284;
 * planar/mpc_osqp.py:104-140).  Checker of robobee3d_b200/csrc/gqp.cu.  Dense row-major P (n x n) and A (m x n).
 */
#include <stdlib.h>
#include <string.h>
#include "oracle.h"

typedef struct { oq_t w; oq_set_t set; int n, m; double infty; } gqo_t;

/* st = eps_abs, eps_rel, rho, sigma, alpha, max_iter, check_termination, adaptive_rho_interval, adaptive_rho_tol, scaling_iters, infty */
void *gqo_create(int n, int m, const double *st) {
  gqo_t *g;
  real *Pz, *Az;
  int i;
  if (n > OQ_NMAX || m > OQ_MMAX) return 0;
  g = (gqo_t *)calloc(1, sizeof(gqo_t));
  g->n = n; g->m = m;
  g->set.eps_abs = (real)st[0]; g->set.eps_rel = (real)st[1]; g->set.rho0 = (real)st[2]; g->set.sigma = (real)st[3];
  g->set.alpha = (real)st[4]; g->set.max_iter = (int)st[5]; g->set.check_termination = (int)st[6];
  g->set.adaptive_rho_interval = (int)st[7]; g->set.adaptive_rho_tol = (real)st[8]; g->set.scaling_iters = (int)st[9];
  g->infty = st[10];
  Pz = (real *)calloc((size_t)n * n, sizeof(real)); Az = (real *)calloc((size_t)m * n, sizeof(real));
  oq_init(&g->w, &g->set, n, m, Pz, Az);
  for (i = 0; i < m; ++i) { g->w.l[i] = 0; g->w.u[i] = 0; }
  free(Pz); free(Az);
  return g;
}
void gqo_destroy(void *h) { free(h); }

/* osqp.update(...) + solve(): bounds, linear cost (old scaling), then -- when P / A are given -- unscale, new data,
 * re-equilibrate, refactor (the order of uprightmpc2.c:247-250 == osqp_update_* in the Python wrapper). */
int gqo_update_solve(void *h, const double *P, const double *A, const double *q, const double *l, const double *u,
                     double *x, double *y, double *info) {
  gqo_t *g = (gqo_t *)h;
  int n = g->n, m = g->m, i, j;
  real *lr = (real *)malloc(sizeof(real) * m), *ur = (real *)malloc(sizeof(real) * m), *qr = (real *)malloc(sizeof(real) * n);
  for (i = 0; i < m; ++i) {
    double lo = l[i] < -g->infty ? -g->infty : l[i], hi = u[i] > g->infty ? g->infty : u[i];
    lr[i] = (real)lo; ur[i] = (real)hi;
  }
  for (j = 0; j < n; ++j) qr[j] = (real)q[j];
  oq_update_bounds(&g->w, &g->set, lr, ur);
  oq_update_lin_cost(&g->w, qr);
  if (P && A) {
    oq_unscale_data(&g->w);
    for (i = 0; i < n; ++i) for (j = 0; j < n; ++j) g->w.P[i * OQ_NMAX + j] = (real)P[i * n + j];
    for (i = 0; i < m; ++i) for (j = 0; j < n; ++j) g->w.A[i * OQ_NMAX + j] = (real)A[i * n + j];
    oq_scale_and_factor(&g->w, &g->set);
  }
  oq_solve(&g->w, &g->set);
  for (j = 0; j < n; ++j) x[j] = (double)g->w.sol_x[j];
  for (i = 0; i < m; ++i) y[i] = (double)g->w.sol_y[i];
  info[0] = g->w.iter; info[1] = g->w.status; info[2] = (double)g->w.pri_res; info[3] = (double)g->w.dua_res;
  info[4] = (double)g->w.rho; info[5] = g->w.rho_updates;
  free(lr); free(ur); free(qr);
  return g->w.status;
}
