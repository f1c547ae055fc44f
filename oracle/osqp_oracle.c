/*
 * oracle/osqp_oracle.c -- TEST INFRASTRUCTURE ONLY (see oracle/oracle.h).
 *
 * Restatement of the ADMM QP algorithm the reference runs (vendored OSQP 0.6.0,
 * EMBEDDED=2) on DENSE storage with a plain no-pivot LDL' of the quasi-definite
 * KKT matrix (elimination order selectable, see get_perm).  It follows the reference's order of operations where that order
 * changes results (warm start kept in stale scaling, Ruiz + cost scaling inside
 * one loop, bounds scaled with the OLD E before the re-equilibration ...):
 *
 *   oq_update_bounds     osqp.c:784-832  + update_rho_vec auxil.c:103-146
 *   oq_update_lin_cost   osqp.c:752-782
 *   oq_unscale_data      scaling.c:160-175
 *   oq_scale_and_factor  scaling.c:44-156 (scale_data), kkt.c:184-222 (KKT values),
 *                        qdldl.c:86-246 (numeric LDL'; here dense)
 *   oq_solve             osqp.c:288-641; steps auxil.c:164-228; residuals
 *                        auxil.c:243-361; termination auxil.c:684-790; rho
 *                        adaptation auxil.c:13-75 + osqp.c:1270-1330; solution
 *                        auxil.c:527-562 + scaling.c:177-192
 *
 * Not restated: the primal/dual infeasibility certificates (auxil.c:365-505).
 * On this path P > 0 and the feasible set is never empty, so those statuses
 * cannot occur; oq_solve reports ST_MAX_ITER instead if nothing converges.
 */
#include <math.h>
#include <string.h>
#include "oracle.h"

#define OQ_INFTY ((real)1e30)
#define MIN_SCALING ((real)1e-4)
#define MAX_SCALING ((real)1e4)
#define RHO_MIN ((real)1e-6)
#define RHO_MAX ((real)1e6)
#define RHO_TOL ((real)1e-4)
#define RHO_EQ_OVER_RHO_INEQ ((real)1e3)

#define P_(w, i, j) ((w)->P[(i) * OQ_NMAX + (j)])
#define A_(w, i, j) ((w)->A[(i) * OQ_NMAX + (j)])
#define L_(w, i, j) ((w)->L[(i) * OQ_KMAX + (j)])

static real r_abs(real v) { return v < 0 ? -v : v; }
static real r_max(real a, real b) { return a > b ? a : b; }
static real r_min(real a, real b) { return a < b ? a : b; }

static real limit_scaling1(real d) {
  d = d < MIN_SCALING ? (real)1.0 : d;
  d = d > MAX_SCALING ? MAX_SCALING : d;
  return d;
}

void oq_init(oq_t *w, const oq_set_t *s, int n, int m, const real *P0, const real *A0) {
  int i, j;
  memset(w, 0, sizeof(*w));
  w->n = n; w->m = m;
  for (i = 0; i < n; ++i) for (j = 0; j < n; ++j) P_(w, i, j) = P0[i * n + j];
  for (i = 0; i < m; ++i) for (j = 0; j < n; ++j) A_(w, i, j) = A0[i * n + j];
  for (j = 0; j < n; ++j) { w->D[j] = w->Dinv[j] = 1; }
  for (i = 0; i < m; ++i) {
    w->E[i] = w->Einv[i] = 1;
    /* generated workspaces start all-inequality (workspace.c:2479-2601):
     * codegen was run with l=0, u=+inf */
    w->ctype[i] = 0; w->rho_vec[i] = s->rho0; w->rho_inv_vec[i] = (real)1.0 / s->rho0;
    w->l[i] = 0; w->u[i] = OQ_INFTY;
  }
  w->c = w->cinv = 1; w->rho = s->rho0; w->status = ST_UNSOLVED;
}

/* optional symmetric permutation of the KKT system (elimination order), per size.
 * perm[k] = index of the KKT row/col eliminated k-th (same convention as QDLDL's
 * permute_x, qdldl_interface.c:325-331).  Identity if never set. */
static int g_perm[OQ_KMAX + 1][OQ_KMAX];
static int g_perm_set[OQ_KMAX + 1];
void orc_set_kkt_perm(int N, const int *perm) {
  int i;
  for (i = 0; i < N; ++i) g_perm[N][i] = perm[i];
  g_perm_set[N] = 1;
}
/* default order: the m constraint rows first (exact pivots -1/rho), then x.
 * This is the reduced form (P + sigma I + A' diag(rho) A) and is measurably the
 * most accurate no-pivot order in float32 (tests/test_oracle_vs_ref.py). */
static const int *get_perm(int n, int m) {
  int i, N = n + m;
  if (!g_perm_set[N]) {
    for (i = 0; i < m; ++i) g_perm[N][i] = n + i;
    for (i = 0; i < n; ++i) g_perm[N][m + i] = i;
    g_perm_set[N] = 1;
  }
  return g_perm[N];
}

/* numeric LDL' of K = [[P + sigma I, A'], [A, -diag(1/rho)]] */
static void oq_factor(oq_t *w, const oq_set_t *s) {
  int n = w->n, m = w->m, N = n + m, i, j, k;
  static real K[OQ_KMAX * OQ_KMAX], K0[OQ_KMAX * OQ_KMAX];
  const int *perm = get_perm(n, m);
  for (i = 0; i < N; ++i) for (j = 0; j < N; ++j) K[i * OQ_KMAX + j] = 0;
  for (i = 0; i < n; ++i) {
    for (j = 0; j < n; ++j) K[i * OQ_KMAX + j] = P_(w, i, j);
    K[i * OQ_KMAX + i] += s->sigma;
  }
  for (i = 0; i < m; ++i) {
    for (j = 0; j < n; ++j) K[(n + i) * OQ_KMAX + j] = K[j * OQ_KMAX + n + i] = A_(w, i, j);
    K[(n + i) * OQ_KMAX + n + i] = -w->rho_inv_vec[i];
  }
  memcpy(K0, K, sizeof(K));
  for (i = 0; i < N; ++i) for (j = 0; j < N; ++j) K[i * OQ_KMAX + j] = K0[perm[i] * OQ_KMAX + perm[j]];
  for (j = 0; j < N; ++j) {
    real d = K[j * OQ_KMAX + j];
    for (k = 0; k < j; ++k) d -= L_(w, j, k) * L_(w, j, k) * w->Dg[k];
    w->Dg[j] = d;
    L_(w, j, j) = 1;
    for (i = j + 1; i < N; ++i) {
      real v = K[i * OQ_KMAX + j];
      if (v != 0 || j > 0) {
        for (k = 0; k < j; ++k) {
          real lik = L_(w, i, k);
          if (lik != 0) v -= lik * L_(w, j, k) * w->Dg[k];
        }
      }
      L_(w, i, j) = v / d;
    }
  }
}

static void oq_kkt_solve(const oq_t *w, real *b0) {
  int N = w->n + w->m, i, j;
  const int *perm = get_perm(w->n, w->m);
  real b[OQ_KMAX];
  for (i = 0; i < N; ++i) b[i] = b0[perm[i]];
  for (i = 0; i < N; ++i) {
    real v = b[i];
    for (j = 0; j < i; ++j) { real l = L_(w, i, j); if (l != 0) v -= l * b[j]; }
    b[i] = v;
  }
  for (i = 0; i < N; ++i) b[i] /= w->Dg[i];
  for (i = N - 1; i >= 0; --i) {
    real v = b[i];
    for (j = i + 1; j < N; ++j) { real l = L_(w, j, i); if (l != 0) v -= l * b[j]; }
    b[i] = v;
  }
  for (i = 0; i < N; ++i) b0[perm[i]] = b[i];
}

static void set_rho_entry(oq_t *w, int i, int type) {
  w->ctype[i] = type;
  w->rho_vec[i] = type == -1 ? RHO_MIN : (type == 1 ? RHO_EQ_OVER_RHO_INEQ * w->rho : w->rho);
  w->rho_inv_vec[i] = (real)1.0 / w->rho_vec[i];
}

int oq_update_bounds(oq_t *w, const oq_set_t *s, const real *l, const real *u) {
  int i, m = w->m, changed = 0;
  for (i = 0; i < m; ++i) if (l[i] > u[i]) return 1; /* osqp.c:801-808: refuses, keeps old */
  for (i = 0; i < m; ++i) { w->l[i] = w->E[i] * l[i]; w->u[i] = w->E[i] * u[i]; }
  w->status = ST_UNSOLVED;
  for (i = 0; i < m; ++i) {
    int type;
    if (w->l[i] < -OQ_INFTY * MIN_SCALING && w->u[i] > OQ_INFTY * MIN_SCALING) type = -1;
    else if (w->u[i] - w->l[i] < RHO_TOL) type = 1;
    else type = 0;
    if (type != w->ctype[i]) { set_rho_entry(w, i, type); changed = 1; }
  }
  if (changed) oq_factor(w, s);
  return 0;
}

void oq_update_lin_cost(oq_t *w, const real *q) {
  int j;
  for (j = 0; j < w->n; ++j) { w->q[j] = w->D[j] * q[j]; w->q[j] *= w->c; }
  w->status = ST_UNSOLVED;
}

void oq_unscale_data(oq_t *w) {
  int n = w->n, m = w->m, i, j;
  for (i = 0; i < n; ++i) for (j = 0; j < n; ++j) P_(w, i, j) *= w->cinv;
  for (i = 0; i < n; ++i) for (j = 0; j < n; ++j) P_(w, i, j) *= w->Dinv[i];
  for (i = 0; i < n; ++i) for (j = 0; j < n; ++j) P_(w, i, j) *= w->Dinv[j];
  for (j = 0; j < n; ++j) { w->q[j] *= w->cinv; w->q[j] = w->Dinv[j] * w->q[j]; }
  for (i = 0; i < m; ++i) for (j = 0; j < n; ++j) A_(w, i, j) *= w->Einv[i];
  for (i = 0; i < m; ++i) for (j = 0; j < n; ++j) A_(w, i, j) *= w->Dinv[j];
  for (i = 0; i < m; ++i) { w->l[i] = w->Einv[i] * w->l[i]; w->u[i] = w->Einv[i] * w->u[i]; }
}

static void pcol_norms(const oq_t *w, real *d) {
  int n = w->n, i, j;
  for (j = 0; j < n; ++j) {
    real v = 0;
    for (i = 0; i < n; ++i) v = r_max(v, r_abs(P_(w, i, j)));
    d[j] = v;
  }
}

void oq_scale_and_factor(oq_t *w, const oq_set_t *s) {
  int n = w->n, m = w->m, i, j, it;
  real Dt[OQ_NMAX], Et[OQ_MMAX];
  w->c = 1;
  for (j = 0; j < n; ++j) w->D[j] = w->Dinv[j] = 1;
  for (i = 0; i < m; ++i) w->E[i] = w->Einv[i] = 1;
  for (it = 0; it < s->scaling_iters; ++it) {
    real c_temp, nq;
    pcol_norms(w, Dt);
    for (j = 0; j < n; ++j) for (i = 0; i < m; ++i) Dt[j] = r_max(Dt[j], r_abs(A_(w, i, j)));
    for (i = 0; i < m; ++i) {
      real v = 0;
      for (j = 0; j < n; ++j) v = r_max(v, r_abs(A_(w, i, j)));
      Et[i] = v;
    }
    for (j = 0; j < n; ++j) Dt[j] = (real)1.0 / R_SQRT(limit_scaling1(Dt[j]));
    for (i = 0; i < m; ++i) Et[i] = (real)1.0 / R_SQRT(limit_scaling1(Et[i]));
    for (i = 0; i < n; ++i) for (j = 0; j < n; ++j) P_(w, i, j) *= Dt[i];
    for (i = 0; i < n; ++i) for (j = 0; j < n; ++j) P_(w, i, j) *= Dt[j];
    for (i = 0; i < m; ++i) for (j = 0; j < n; ++j) A_(w, i, j) *= Et[i];
    for (i = 0; i < m; ++i) for (j = 0; j < n; ++j) A_(w, i, j) *= Dt[j];
    for (j = 0; j < n; ++j) { w->q[j] = Dt[j] * w->q[j]; w->D[j] = w->D[j] * Dt[j]; }
    for (i = 0; i < m; ++i) w->E[i] = w->E[i] * Et[i];
    /* cost normalisation */
    pcol_norms(w, Dt);
    c_temp = 0;
    for (j = 0; j < n; ++j) c_temp += Dt[j];
    c_temp /= (real)n;
    nq = 0;
    for (j = 0; j < n; ++j) nq = r_max(nq, r_abs(w->q[j]));
    nq = limit_scaling1(nq);
    c_temp = r_max(c_temp, nq);
    c_temp = limit_scaling1(c_temp);
    c_temp = (real)1.0 / c_temp;
    for (i = 0; i < n; ++i) for (j = 0; j < n; ++j) P_(w, i, j) *= c_temp;
    for (j = 0; j < n; ++j) w->q[j] *= c_temp;
    w->c *= c_temp;
  }
  w->cinv = (real)1.0 / w->c;
  for (j = 0; j < n; ++j) w->Dinv[j] = (real)1.0 / w->D[j];
  for (i = 0; i < m; ++i) w->Einv[i] = (real)1.0 / w->E[i];
  for (i = 0; i < m; ++i) { w->l[i] = w->E[i] * w->l[i]; w->u[i] = w->E[i] * w->u[i]; }
  oq_factor(w, s);
  w->status = ST_UNSOLVED;
}

/* ---- residuals / termination (auxil.c:243-361, 684-790) -------------------- */
typedef struct { real Ax[OQ_MMAX], Px[OQ_NMAX], Aty[OQ_NMAX], rp[OQ_MMAX], rd[OQ_NMAX]; } oq_res_t;

static void update_info(oq_t *w, oq_res_t *r, int iter) {
  int n = w->n, m = w->m, i, j;
  real v;
  w->iter = iter;
  for (i = 0; i < m; ++i) {
    real a = 0;
    for (j = 0; j < n; ++j) { real aij = A_(w, i, j); if (aij != 0) a += aij * w->x[j]; }
    r->Ax[i] = a; r->rp[i] = a - w->z[i];
  }
  v = 0;
  for (i = 0; i < m; ++i) v = r_max(v, r_abs(w->Einv[i] * r->rp[i]));
  w->pri_res = v;
  for (j = 0; j < n; ++j) {
    real px = 0, aty = 0;
    for (i = 0; i < n; ++i) { real pij = P_(w, j, i); if (pij != 0) px += pij * w->x[i]; }
    for (i = 0; i < m; ++i) { real aij = A_(w, i, j); if (aij != 0) aty += aij * w->y[i]; }
    r->Px[j] = px; r->Aty[j] = aty; r->rd[j] = w->q[j] + px + aty;
  }
  v = 0;
  for (j = 0; j < n; ++j) v = r_max(v, r_abs(w->Dinv[j] * r->rd[j]));
  w->dua_res = w->cinv * v;
}

static int check_termination(oq_t *w, const oq_set_t *s, const oq_res_t *r, int approximate) {
  int n = w->n, m = w->m, i, j;
  real eps_abs = s->eps_abs, eps_rel = s->eps_rel, eps_prim, eps_dual, a, b;
  if (w->pri_res > OQ_INFTY || w->dua_res > OQ_INFTY) { w->status = ST_NON_CVX; return 1; }
  if (approximate) { eps_abs *= 10; eps_rel *= 10; }
  a = 0; b = 0;
  for (i = 0; i < m; ++i) { a = r_max(a, r_abs(w->Einv[i] * w->z[i])); b = r_max(b, r_abs(w->Einv[i] * r->Ax[i])); }
  eps_prim = eps_abs + eps_rel * r_max(a, b);
  a = 0;
  for (j = 0; j < n; ++j) a = r_max(a, r_abs(w->Dinv[j] * w->q[j]));
  b = 0;
  for (j = 0; j < n; ++j) b = r_max(b, r_abs(w->Dinv[j] * r->Aty[j]));
  a = r_max(a, b);
  b = 0;
  for (j = 0; j < n; ++j) b = r_max(b, r_abs(w->Dinv[j] * r->Px[j]));
  a = r_max(a, b);
  a *= w->cinv;
  eps_dual = eps_abs + eps_rel * a;
  if ((m == 0 || w->pri_res < eps_prim) && w->dua_res < eps_dual) {
    w->status = approximate ? ST_SOLVED_INACCURATE : ST_SOLVED;
    return 1;
  }
  return 0;
}

/* auxil.c:13-75 and osqp_update_rho (osqp.c:1270-1330) */
static void adapt_rho(oq_t *w, const oq_set_t *s, const oq_res_t *r) {
  int n = w->n, m = w->m, i, j;
  real pri = 0, dua = 0, pn = 0, dn = 0, est;
  for (i = 0; i < m; ++i) { pri = r_max(pri, r_abs(r->rp[i])); pn = r_max(pn, r_max(r_abs(w->z[i]), r_abs(r->Ax[i]))); }
  for (j = 0; j < n; ++j) {
    dua = r_max(dua, r_abs(r->rd[j]));
    dn = r_max(dn, r_max(r_abs(w->q[j]), r_max(r_abs(r->Aty[j]), r_abs(r->Px[j]))));
  }
  pri /= (pn + (real)1e-10);
  dua /= (dn + (real)1e-10);
  est = w->rho * R_SQRT(pri / (dua + (real)1e-10));
  est = r_min(r_max(est, RHO_MIN), RHO_MAX);
  if (est > w->rho * s->adaptive_rho_tol || est < w->rho / s->adaptive_rho_tol) {
    w->rho = est;
    for (i = 0; i < m; ++i) if (w->ctype[i] != -1) set_rho_entry(w, i, w->ctype[i]);
    oq_factor(w, s);
    w->rho_updates += 1;
  }
}

void oq_solve(oq_t *w, const oq_set_t *s) {
  int n = w->n, m = w->m, i, j, iter, checked = 0;
  real xp[OQ_NMAX], zp[OQ_MMAX], t[OQ_KMAX];
  oq_res_t res;
  w->rho_updates = 0;
  for (iter = 1; iter <= s->max_iter; ++iter) {
    memcpy(xp, w->x, sizeof(xp)); memcpy(zp, w->z, sizeof(zp));
    /* update_xz_tilde */
    for (j = 0; j < n; ++j) t[j] = s->sigma * xp[j] - w->q[j];
    for (i = 0; i < m; ++i) t[n + i] = zp[i] - w->rho_inv_vec[i] * w->y[i];
    oq_kkt_solve(w, t);
    for (i = 0; i < m; ++i) t[n + i] = zp[i] + w->rho_inv_vec[i] * (t[n + i] - w->y[i]);
    /* update_x, update_z (+ projection), update_y */
    for (j = 0; j < n; ++j) w->x[j] = s->alpha * t[j] + ((real)1.0 - s->alpha) * xp[j];
    for (i = 0; i < m; ++i) {
      real v = s->alpha * t[n + i] + ((real)1.0 - s->alpha) * zp[i] + w->rho_inv_vec[i] * w->y[i];
      w->z[i] = r_min(r_max(v, w->l[i]), w->u[i]);
    }
    for (i = 0; i < m; ++i)
      w->y[i] += w->rho_vec[i] * (s->alpha * t[n + i] + ((real)1.0 - s->alpha) * zp[i] - w->z[i]);

    checked = s->check_termination && (iter % s->check_termination == 0);
    if (checked) {
      update_info(w, &res, iter);
      if (check_termination(w, s, &res, 0)) break;
    }
    if (s->adaptive_rho_interval && (iter % s->adaptive_rho_interval == 0)) {
      if (!checked) update_info(w, &res, iter);
      adapt_rho(w, s, &res);
    }
  }
  if (!checked) {
    update_info(w, &res, iter - 1);
    check_termination(w, s, &res, 0);
  }
  if (w->status == ST_UNSOLVED) {
    if (!check_termination(w, s, &res, 1)) w->status = ST_MAX_ITER;
  }
  if (w->status == ST_NON_CVX) {
    /* auxil.c:538-561: NaN solution, cold start */
    for (j = 0; j < n; ++j) { w->sol_x[j] = (real)NAN; w->x[j] = 0; }
    for (i = 0; i < m; ++i) { w->sol_y[i] = (real)NAN; w->z[i] = 0; w->y[i] = 0; }
  } else {
    for (j = 0; j < n; ++j) w->sol_x[j] = w->D[j] * w->x[j];
    for (i = 0; i < m; ++i) { w->sol_y[i] = w->E[i] * w->y[i]; w->sol_y[i] *= w->cinv; }
  }
}
