/*
 * oracle/dirtran_oracle.c -- TEST INFRASTRUCTURE ONLY (see oracle/oracle.h).
 *
 * CPU restatement of the planar direct-transcription MPC (BASELINE config 5):
 *
 *   model            planar/FlappingModels.py:27-129   PlanarThrustStrokeDev(rescale=True): dydt, getLinearDynamics, limits
 *   QP template      controlutils/py/csc.py:6-54 (A), ltvsystem.py:29-77 (P, q), :107-165 (l, u), :181-232 (updateTrajectory,
 *                    trajMode GIVEN_POINT_OR_TRAJ with a single point: ONE linearisation at (x0, previous input) for all N)
 *   controller       controlutils/py/mpc.py:62-91 LTVMPC.update (costMode FINAL), returns u_0 and keeps it as the next
 *                    linearisation input
 *   loop             planar/mpc_thrust_strokedev.py:97-199 runSim(CTRL_MPC): tgoal=(i+N)dt, update, int(dt/SIM_DT) Euler
 *                    substeps of dydt at SIM_DT=0.0005
 *   solver           the generic dense OSQP restatement of osqp_oracle.c (compiled into this library with larger size
 *                    caps), settings of mpc_thrust_strokedev.py:124: scaling=0, eps_abs=eps_rel=eps, polish off, warm start,
 *                    rho adapted (and kept across solves, as one OSQP workspace does).
 *
 * x = (y_0..y_N, u_0..u_{N-1});  rows: (N+1)nx dynamics equalities, (N+1)nx state bounds, N nu input bounds.
 *
 * Pinning: the solver part is checked against the reference's own generated instance of this QP (planar/code, OSQP 0.5.0
 * EMBEDDED=1, float32, n=46 m=82) compiled in place as oracle/_ref/libref_planar_f32.so -- tests/test_dirtran_oracle.py.
 * The QP data layout is checked there against the same instance's CSC pattern.  The Python pieces above (model, loop)
 * need autograd/osqp/matplotlib, absent here: restated by reading, parity of those unpinned.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "oracle.h"

#define DT_NX 6
#define DT_NU 2
#define DT_INF ((real)1e30)

typedef struct {
  int N, n, m;
  real dt, wx[DT_NX], wu[DT_NU];
  real g, ibmb, umin[DT_NU], umax[DT_NU], xmin[DT_NX], xmax[DT_NX];
  oq_set_t set;
  oq_t *qp;
  real ctrl[DT_NU];
  real *l, *u, *q;
  int fresh;
} dto_t;

static void dt_dydt(const dto_t *d, const real *y, const real *u, real *f) {
  /* FlappingModels.py:114-126 (rescale) */
  const real phi = y[2], thrust = d->g + u[0];
  f[0] = y[3]; f[1] = y[4]; f[2] = y[5];
  f[3] = -thrust * R_SIN(phi);
  f[4] = -d->g + thrust * R_COS(phi);
  f[5] = thrust * u[1] / d->ibmb;
}

/* FlappingModels.py:51-102: Ad = I + dt dfdy, Bd = dt dfdu, fd = dt (f - dfdy y - dfdu u); row-major */
static void dt_linearise(const dto_t *d, const real *y, const real *u, real *Ad, real *Bd, real *fd) {
  real fy[DT_NX], dfdy[DT_NX * DT_NX], dfdu[DT_NX * DT_NU];
  const real phi = y[2], c = R_COS(phi), s = R_SIN(phi), thrust = d->g + u[0];
  int i, j;
  dt_dydt(d, y, u, fy);
  memset(dfdy, 0, sizeof(dfdy)); memset(dfdu, 0, sizeof(dfdu));
  for (i = 0; i < 3; ++i) dfdy[i * DT_NX + 3 + i] = 1;
  dfdy[3 * DT_NX + 2] = -thrust * c;
  dfdy[4 * DT_NX + 2] = -thrust * s;
  dfdu[3 * DT_NU + 0] = -s;
  dfdu[4 * DT_NU + 0] = c;
  dfdu[5 * DT_NU + 0] = u[1] / d->ibmb;
  dfdu[5 * DT_NU + 1] = thrust / d->ibmb;
  for (i = 0; i < DT_NX; ++i) {
    real a = fy[i];
    for (j = 0; j < DT_NX; ++j) a -= dfdy[i * DT_NX + j] * y[j];
    for (j = 0; j < DT_NU; ++j) a -= dfdu[i * DT_NU + j] * u[j];
    fd[i] = d->dt * a;
    for (j = 0; j < DT_NX; ++j) Ad[i * DT_NX + j] = (i == j ? 1 : 0) + d->dt * dfdy[i * DT_NX + j];
    for (j = 0; j < DT_NU; ++j) Bd[i * DT_NU + j] = d->dt * dfdu[i * DT_NU + j];
  }
}

/* settings: eps_abs, eps_rel, rho0, sigma, alpha, max_iter, check_termination, adaptive_rho_interval, adaptive_rho_tol */
void *dto_create(int N, double dt, const double *wx, const double *wu, const double *st) {
  dto_t *d = (dto_t *)calloc(1, sizeof(dto_t));
  int i, j, k, n, m;
  const real PI = (real)3.14159265358979323846;
  real *P0, *A0;
  d->N = N; d->dt = (real)dt;
  n = d->n = (N + 1) * DT_NX + N * DT_NU;
  m = d->m = n + (N + 1) * DT_NX;
  if (n > OQ_NMAX || m > OQ_MMAX) { free(d); return 0; }
  for (i = 0; i < DT_NX; ++i) d->wx[i] = (real)wx[i];
  for (i = 0; i < DT_NU; ++i) d->wu[i] = (real)wu[i];
  /* FlappingModels.py:29-50,104-112 */
  d->g = (real)9.81; d->ibmb = (real)(1.0 / 12.0 * 0.1 * 0.1);
  d->umin[0] = -d->g; d->umin[1] = (real)(-0.05 * 0.1);
  d->umax[0] = 3 * d->g; d->umax[1] = (real)(0.05 * 0.1);
  for (i = 0; i < DT_NX; ++i) { d->xmin[i] = -DT_INF; d->xmax[i] = DT_INF; }
  d->xmin[2] = -10 * PI; d->xmax[2] = 10 * PI;
  d->set.eps_abs = (real)st[0]; d->set.eps_rel = (real)st[1]; d->set.rho0 = (real)st[2]; d->set.sigma = (real)st[3];
  d->set.alpha = (real)st[4]; d->set.max_iter = (int)st[5]; d->set.check_termination = (int)st[6];
  d->set.adaptive_rho_interval = (int)st[7]; d->set.adaptive_rho_tol = (real)st[8]; d->set.scaling_iters = 0;
  d->qp = (oq_t *)malloc(sizeof(oq_t));
  d->l = (real *)calloc(m, sizeof(real)); d->u = (real *)calloc(m, sizeof(real)); d->q = (real *)calloc(n, sizeof(real));
  P0 = (real *)calloc((size_t)n * n, sizeof(real)); A0 = (real *)calloc((size_t)m * n, sizeof(real));
  /* P = blkdiag(I_{N+1} (x) Q, I_N (x) R), ltvsystem.py:52 (kdamping 0) */
  for (k = 0; k <= N; ++k) for (i = 0; i < DT_NX; ++i) P0[(size_t)(k * DT_NX + i) * n + k * DT_NX + i] = d->wx[i];
  for (k = 0; k < N; ++k) for (i = 0; i < DT_NU; ++i) {
    j = (N + 1) * DT_NX + k * DT_NU + i;
    P0[(size_t)j * n + j] = d->wu[i];
  }
  /* A pattern with csc.init's placeholder ones in the Ad/Bd blocks (csc.py:16-22); identity below */
  for (k = 0; k <= N; ++k) for (i = 0; i < DT_NX; ++i) A0[(size_t)(k * DT_NX + i) * n + k * DT_NX + i] = -1;
  for (k = 0; k < N; ++k) for (i = 0; i < DT_NX; ++i) {
    for (j = 0; j < DT_NX; ++j) A0[(size_t)((k + 1) * DT_NX + i) * n + k * DT_NX + j] = 1;
    for (j = 0; j < DT_NU; ++j) A0[(size_t)((k + 1) * DT_NX + i) * n + (N + 1) * DT_NX + k * DT_NU + j] = 1;
  }
  for (i = 0; i < n; ++i) A0[(size_t)((N + 1) * DT_NX + i) * n + i] = 1;
  oq_init(d->qp, &d->set, n, m, P0, A0);
  /* bounds of ltvsystem.py:133-150 */
  for (k = 0; k <= N; ++k) for (i = 0; i < DT_NX; ++i) {
    d->l[(N + 1) * DT_NX + k * DT_NX + i] = d->xmin[i]; d->u[(N + 1) * DT_NX + k * DT_NX + i] = d->xmax[i];
  }
  for (k = 0; k < N; ++k) for (i = 0; i < DT_NU; ++i) {
    d->l[2 * (N + 1) * DT_NX + k * DT_NU + i] = d->umin[i]; d->u[2 * (N + 1) * DT_NX + k * DT_NU + i] = d->umax[i];
  }
  oq_update_bounds(d->qp, &d->set, d->l, d->u);
  d->fresh = 1;
  free(P0); free(A0);
  return d;
}

void dto_destroy(void *h) {
  dto_t *d = (dto_t *)h;
  if (!d) return;
  free(d->qp); free(d->l); free(d->u); free(d->q); free(d);
}

void dto_set_ctrl(void *h, const double *ctrl) {
  dto_t *d = (dto_t *)h;
  d->ctrl[0] = (real)ctrl[0]; d->ctrl[1] = (real)ctrl[1];
}

/* one LTVMPC.update(x0, xr, costMode=FINAL): returns OSQP's status value */
int dto_update(void *h, const double *x0d, const double *xrd, double *u_out) {
  dto_t *d = (dto_t *)h;
  const int N = d->N, n = d->n;
  real x0[DT_NX], xr[DT_NX], Ad[DT_NX * DT_NX], Bd[DT_NX * DT_NU], fd[DT_NX];
  int i, j, k;
  oq_t *w = d->qp;
  for (i = 0; i < DT_NX; ++i) { x0[i] = (real)x0d[i]; xr[i] = (real)xrd[i]; }
  dt_linearise(d, x0, d->ctrl, Ad, Bd, fd);
  for (i = 0; i < DT_NX; ++i) d->l[i] = d->u[i] = -x0[i];
  for (k = 0; k < N; ++k) for (i = 0; i < DT_NX; ++i) d->l[(k + 1) * DT_NX + i] = d->u[(k + 1) * DT_NX + i] = -fd[i];
  /* q: ltvsystem.py:68 -- -Q xr at all N+1 knots (the costMode/trajMode test at :76 is false for FINAL + a single point) */
  for (k = 0; k <= N; ++k) for (i = 0; i < DT_NX; ++i) d->q[k * DT_NX + i] = -d->wx[i] * xr[i];
  for (i = (N + 1) * DT_NX; i < n; ++i) d->q[i] = 0;
  oq_update_bounds(w, &d->set, d->l, d->u);
  oq_update_lin_cost(w, d->q);
  oq_unscale_data(w);
  for (k = 0; k < N; ++k) for (i = 0; i < DT_NX; ++i) {
    for (j = 0; j < DT_NX; ++j) w->A[(size_t)((k + 1) * DT_NX + i) * OQ_NMAX + k * DT_NX + j] = Ad[i * DT_NX + j];
    for (j = 0; j < DT_NU; ++j) w->A[(size_t)((k + 1) * DT_NX + i) * OQ_NMAX + (N + 1) * DT_NX + k * DT_NU + j] = Bd[i * DT_NU + j];
  }
  oq_scale_and_factor(w, &d->set);
  oq_solve(w, &d->set);
  /* mpc.py:89 */
  for (i = 0; i < DT_NU; ++i) { d->ctrl[i] = w->sol_x[(N + 1) * DT_NX + i]; u_out[i] = (double)d->ctrl[i]; }
  return w->status;
}

/* info: iter, status, pri_res, dua_res, rho, rho_updates */
void dto_get(void *h, double *solx, double *soly, double *info) {
  dto_t *d = (dto_t *)h;
  oq_t *w = d->qp;
  int i;
  if (solx) for (i = 0; i < d->n; ++i) solx[i] = (double)w->sol_x[i];
  if (soly) for (i = 0; i < d->m; ++i) soly[i] = (double)w->sol_y[i];
  if (info) { info[0] = w->iter; info[1] = w->status; info[2] = (double)w->pri_res; info[3] = (double)w->dua_res;
              info[4] = (double)w->rho; info[5] = w->rho_updates; }
}

/* the QP data as the solver holds it (dense, row-major) */
void dto_qpdata(void *h, double *Pdiag, double *A, double *q, double *l, double *u) {
  dto_t *d = (dto_t *)h;
  oq_t *w = d->qp;
  int i, j;
  for (i = 0; i < d->n; ++i) { Pdiag[i] = (double)w->P[(size_t)i * OQ_NMAX + i]; q[i] = (double)w->q[i]; }
  for (i = 0; i < d->m; ++i) {
    l[i] = (double)w->l[i]; u[i] = (double)w->u[i];
    for (j = 0; j < d->n; ++j) A[(size_t)i * d->n + j] = (double)w->A[(size_t)i * OQ_NMAX + j];
  }
}

/* mpc_thrust_strokedev.py:189-192 */
void dto_plant(void *h, double *x, const double *u, double sim_dt, int nsub) {
  dto_t *d = (dto_t *)h;
  real y[DT_NX], uu[DT_NU], f[DT_NX];
  int i, s;
  for (i = 0; i < DT_NX; ++i) y[i] = (real)x[i];
  uu[0] = (real)u[0]; uu[1] = (real)u[1];
  for (s = 0; s < nsub; ++s) {
    dt_dydt(d, y, uu, f);
    for (i = 0; i < DT_NX; ++i) y[i] += f[i] * (real)sim_dt;
  }
  for (i = 0; i < DT_NX; ++i) x[i] = (double)y[i];
}

/* nsteps of runSim's loop from x (in/out); X[nsteps][6], U[nsteps][2], iters[nsteps] logs may be NULL */
int dto_rollout(void *h, double *x, const double *xr, int nsteps, double sim_dt, int nsub, double *X, double *U, double *iters) {
  dto_t *d = (dto_t *)h;
  int s, i, bad = 0;
  double u[DT_NU];
  for (s = 0; s < nsteps; ++s) {
    int st = dto_update(h, x, xr, u);
    if (st != ST_SOLVED) ++bad;
    dto_plant(h, x, u, sim_dt, nsub);
    if (X) for (i = 0; i < DT_NX; ++i) X[s * DT_NX + i] = x[i];
    if (U) for (i = 0; i < DT_NU; ++i) U[s * DT_NU + i] = u[i];
    if (iters) iters[s] = d->qp->iter;
  }
  return bad;
}

/* ---- raw QP entry: the solver restatement alone on caller-supplied dense data (pinning against planar/code) ---- */
/* st as dto_create; rho_vec0 (may be NULL): per-row rho to start from, as a generated workspace carries it */
int dto_raw_solve(int n, int m, const double *P, const double *A, const double *q, const double *l, const double *u,
                  const double *st, double *x_out, double *y_out, double *info) {
  oq_set_t set;
  oq_t *w;
  real *Pr, *Ar, *qr, *lr, *ur;
  int i, status;
  if (n > OQ_NMAX || m > OQ_MMAX) return -100;
  set.eps_abs = (real)st[0]; set.eps_rel = (real)st[1]; set.rho0 = (real)st[2]; set.sigma = (real)st[3]; set.alpha = (real)st[4];
  set.max_iter = (int)st[5]; set.check_termination = (int)st[6]; set.adaptive_rho_interval = (int)st[7];
  set.adaptive_rho_tol = (real)st[8]; set.scaling_iters = 0;
  w = (oq_t *)malloc(sizeof(oq_t));
  Pr = (real *)malloc(sizeof(real) * n * n); Ar = (real *)malloc(sizeof(real) * m * n);
  qr = (real *)malloc(sizeof(real) * n); lr = (real *)malloc(sizeof(real) * m); ur = (real *)malloc(sizeof(real) * m);
  for (i = 0; i < n * n; ++i) Pr[i] = (real)P[i];
  for (i = 0; i < m * n; ++i) Ar[i] = (real)A[i];
  for (i = 0; i < n; ++i) qr[i] = (real)q[i];
  for (i = 0; i < m; ++i) { lr[i] = (real)l[i]; ur[i] = (real)u[i]; }
  oq_init(w, &set, n, m, Pr, Ar);
  oq_update_bounds(w, &set, lr, ur);
  oq_update_lin_cost(w, qr);
  oq_scale_and_factor(w, &set);
  oq_solve(w, &set);
  for (i = 0; i < n; ++i) x_out[i] = (double)w->sol_x[i];
  for (i = 0; i < m; ++i) y_out[i] = (double)w->sol_y[i];
  info[0] = w->iter; info[1] = w->status; info[2] = (double)w->pri_res; info[3] = (double)w->dua_res;
  info[4] = (double)w->rho; info[5] = w->rho_updates;
  status = w->status;
  free(w); free(Pr); free(Ar); free(qr); free(lr); free(ur);
  return status;
}
