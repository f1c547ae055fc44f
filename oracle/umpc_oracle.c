/*
 * oracle/umpc_oracle.c -- TEST INFRASTRUCTURE ONLY (see oracle/oracle.h).
 *
 * Restatement of the UprightMPC2 controller update:
 *   pattern / Ax_idx   template_controllers.py:28-63,86-121  (== uprightmpc2.c:65-113)
 *   init               uprightmpc2.c:19-63,116-117; createMPC template_controllers.py:260-280
 *   template projection uprightmpc2.c:214-241   (update2, template_controllers.py:226-237)
 *   constraint data    uprightmpc2.c:126-180    (updateConstraint, :65-125)
 *   objective data     uprightmpc2.c:182-207    (updateObjective, :127-143)
 *   OSQP call order    uprightmpc2.c:247-250
 *   extraction         uprightmpc2.c:252-269    (getAccDes, template_controllers.py:245-251)
 */
#include <math.h>
#include <string.h>
#include "oracle.h"

static umpc_pattern_t g_pat;

/* Build the dense placeholder A exactly as initConstraint does, then read the
 * CSC order (column-major scan) off it. */
static void build_dense_A(real *A /* NC x NX row-major */, real dt, real T0, const real *s0,
                          const real *Btau /* 3x2 col-major */) {
  int k, i, N = UM_N, ny = UM_NY, nu = UM_NU;
  int n1 = N * ny, n2 = 2 * N * ny, nc1 = N * ny, nc2 = 2 * N * ny;
  memset(A, 0, sizeof(real) * UM_NC * UM_NX);
#define AD(r, c) A[(r) * UM_NX + (c)]
  for (k = 0; k < N; ++k) {
    for (i = 0; i < ny; ++i) {
      AD(k * ny + i, k * ny + i) = -1;
      AD(k * ny + i, n1 + k * ny + i) = dt;
      if (k > 0) AD(k * ny + i, (k - 1) * ny + i) = 1;
      AD(nc1 + k * ny + i, n1 + k * ny + i) = -1;
      if (k > 0) AD(nc1 + k * ny + i, n1 + (k - 1) * ny + i) = 1;
    }
    /* B0 = [[s0, 0],[0, Btau]] */
    for (i = 0; i < 3; ++i) {
      AD(nc1 + k * ny + i, n2 + k * nu) = dt * s0[i];
      AD(nc1 + k * ny + 3 + i, n2 + k * nu + 1) = dt * Btau[i];
      AD(nc1 + k * ny + 3 + i, n2 + k * nu + 2) = dt * Btau[3 + i];
    }
    /* A0(T0 dt): upper-right 3x3 block T0*dt*I acting on y_{k-2}'s s part */
    if (k > 1)
      for (i = 0; i < 3; ++i) AD(nc1 + k * ny + i, (k - 2) * ny + 3 + i) = T0 * dt;
    AD(nc2 + k, n2 + 3 * k) = 1;
  }
#undef AD
}

const umpc_pattern_t *orc_umpc_pattern(void) {
  if (!g_pat.built) {
    static real A[UM_NC * UM_NX];
    real ones3[3] = {1, 1, 1}, ones6[6] = {1, 1, 1, 1, 1, 1};
    int r, c, k = 0, i, offs, n1, n2;
    build_dense_A(A, 1, 1, ones3, ones6);
    for (c = 0; c < UM_NX; ++c)
      for (r = 0; r < UM_NC; ++r)
        if (A[r * UM_NX + c] != 0) { g_pat.row[k] = r; g_pat.col[k] = c; g_pat.init_val[k] = A[r * UM_NX + c]; ++k; }
    /* k must be 111 (assert A.nnz == n1 + n2*N, template_controllers.py:113) */
    /* Ax_idx, restated from the index arithmetic of template_controllers.py:86-121 */
    offs = 0;
    n2 = 2 * UM_NY + 3;
    for (i = 0; i < UM_N - 2; ++i) { g_pat.ax_idx[offs++] = n2 * i + 8; g_pat.ax_idx[offs++] = n2 * i + 11; g_pat.ax_idx[offs++] = n2 * i + 14; }
    n1 = (2 * UM_N - 1) * UM_NY + (UM_N - 2) * 3;
    n2 = 3 * UM_NY;
    for (i = 0; i < UM_N; ++i) {
      int j, step = i < UM_N - 1 ? 3 : 2;
      for (j = 0; j < 6; ++j) g_pat.ax_idx[offs++] = n1 + n2 * i + step * j;
    }
    n1 += 3 * UM_NY * (UM_N - 1) + 2 * UM_NY;
    n2 = 10;
    for (i = 0; i < UM_N; ++i) { int j; for (j = 0; j < 3; ++j) g_pat.ax_idx[offs++] = n1 + n2 * i + j; }
    for (i = 0; i < UM_N; ++i) { int j; for (j = 0; j < 6; ++j) g_pat.ax_idx[offs++] = n1 + n2 * i + 4 + j; }
    g_pat.built = (k == UM_NNZA && offs == UM_NADATA) ? 1 : -1;
  }
  return &g_pat;
}

int orc_umpc_sizeof_par(void) { return (int)sizeof(umpc_par_t); }
int orc_umpc_sizeof_state(void) { return (int)sizeof(umpc_state_t); }
int orc_sizeof_real(void) { return (int)sizeof(real); }

/* pat_out: row[111], col[111], ax_idx[48]; returns built flag (1 ok) */
int orc_umpc_get_pattern(int *row, int *col, int *ax_idx) {
  const umpc_pattern_t *p = orc_umpc_pattern();
  memcpy(row, p->row, sizeof(p->row)); memcpy(col, p->col, sizeof(p->col));
  memcpy(ax_idx, p->ax_idx, sizeof(p->ax_idx));
  return p->built;
}

/* args13 = dt,g,TtoWmax,ws,wds,wpr,wpf,wvr,wvf,wthrust,wmom,... (umpcInit order) */
void orc_umpc_init(umpc_par_t *par, umpc_state_t *st, const double *a11, const double *Ib, int maxIter) {
  int i;
  real dt = (real)a11[0], g = (real)a11[1], TtoW = (real)a11[2], ws = (real)a11[3], wds = (real)a11[4];
  real wpr = (real)a11[5], wpf = (real)a11[6], wvr = (real)a11[7], wvf = (real)a11[8];
  real wthrust = (real)a11[9], wmom = (real)a11[10];
  static real P0[UM_NX * UM_NX], A0[UM_NC * UM_NX];
  real ones3[3] = {1, 1, 1}, ones6[6] = {1, 1, 1, 1, 1, 1};
  memset(par, 0, sizeof(*par));
  par->dt = dt; par->g = g; par->Tmax = TtoW * g;
  for (i = 0; i < 3; ++i) {
    par->Qyr[i] = wpr; par->Qyf[i] = wpf; par->Qyr[3 + i] = par->Qyf[3 + i] = ws;
    par->Qdyr[i] = wvr; par->Qdyf[i] = wvf; par->Qdyr[3 + i] = par->Qdyf[3 + i] = wds;
    par->Ibi[i] = (real)1.0 / (real)Ib[i];
  }
  par->R[0] = wthrust; par->R[1] = par->R[2] = wmom;
  /* workspace.c:561, then uprightmpc2.c:116-117 */
  par->set.rho0 = (real)0.1; par->set.sigma = (real)1e-6; par->set.alpha = (real)1.6;
  par->set.scaling_iters = 10; par->set.adaptive_rho_interval = 0; par->set.adaptive_rho_tol = 5;
  par->set.eps_abs = par->set.eps_rel = (real)1e-4;
  par->set.max_iter = maxIter; par->set.check_termination = 0;
  /* workspace as generated: P = I, A = placeholders (template_controllers.py:28-63) */
  memset(P0, 0, sizeof(P0));
  for (i = 0; i < UM_NX; ++i) P0[i * UM_NX + i] = 1;
  build_dense_A(A0, 1, 1, ones3, ones6);
  st->T0 = 0;
  oq_init(&st->qp, &par->set, UM_NX, UM_NC, P0, A0);
  orc_umpc_pattern();
}

/* settings override for the "converged" (Python-path) flavour,
 * template_controllers.py:190-191: eps 1e-4, OSQP defaults otherwise */
void orc_umpc_set_solver(umpc_par_t *par, int max_iter, int check_termination, int adaptive_rho_interval,
                         double eps_abs, double eps_rel) {
  par->set.max_iter = max_iter; par->set.check_termination = check_termination;
  par->set.adaptive_rho_interval = adaptive_rho_interval;
  par->set.eps_abs = (real)eps_abs; par->set.eps_rel = (real)eps_rel;
}

static real A0_times_i(real T0, const real *y, int i) { return i < 3 ? T0 * y[i + 3] : 0; }

/* R0 column-major (uprightmpc2.c:219). Returns 0 like osqp_solve's exitflag. */
int orc_umpc_update(const umpc_par_t *par, umpc_state_t *st, double *uquad_o, double *accdes_o,
                    const double *p0d, const double *R0d, const double *dq0d, const double *pdesd,
                    const double *dpdesd, const double *sdesd, double actualT0) {
  const umpc_pattern_t *pat = orc_umpc_pattern();
  real R0[9], dq0[6], s0[3], ds0[3], Btau[6], y0[6], dy0[6], y1[6], ydes[6], dydes[6];
  real l[UM_NC], u[UM_NC], q[UM_NX], Px[UM_NX], Axd[UM_NADATA], c0[6] = {0, 0, 0, 0, 0, 0};
  real dt = par->dt, w0, w1, w2;
  int i, k, offs;
  oq_t *qp = &st->qp;
  for (i = 0; i < 9; ++i) R0[i] = (real)R0d[i];
  for (i = 0; i < 6; ++i) dq0[i] = (real)dq0d[i];
  c0[2] = -par->g;
  if (actualT0 >= 0) st->T0 = (real)actualT0;
  /* s0 = R e3 ; ds0 = -R e3h w ; Btau = (-R e3h Ibi)[:, :2]   (e3h w = (-w1, w0, 0)) */
  for (i = 0; i < 3; ++i) s0[i] = R0[6 + i];
  w0 = dq0[3]; w1 = dq0[4]; w2 = dq0[5]; (void)w2;
  for (i = 0; i < 3; ++i) ds0[i] = -(R0[i] * (-w1) + R0[3 + i] * w0);
  for (i = 0; i < 3; ++i) {
    /* e3h*Ibi has entries (1,0)=Ibi0 and (0,1)=-Ibi1 */
    Btau[i] = -(R0[3 + i] * par->Ibi[0]);
    Btau[3 + i] = -(R0[i] * (-par->Ibi[1]));
  }
  for (i = 0; i < 3; ++i) {
    y0[i] = (real)p0d[i]; y0[3 + i] = s0[i];
    dy0[i] = dq0[i]; dy0[3 + i] = ds0[i];
    ydes[i] = (real)pdesd[i]; ydes[3 + i] = (real)sdesd[i];
    dydes[i] = (real)dpdesd[i]; dydes[3 + i] = 0;
  }
  /* constraint vectors */
  for (i = 0; i < UM_NC; ++i) l[i] = 0;
  for (i = 0; i < 6; ++i) { y1[i] = y0[i] + dt * dy0[i]; l[i] = -y1[i]; }
  for (k = 0; k < UM_N; ++k)
    for (i = 0; i < 6; ++i) {
      real *pk = &l[UM_NY * (UM_N + k)];
      if (k == 0) pk[i] = -dy0[i] - dt * A0_times_i(st->T0, y0, i) - dt * c0[i];
      else if (k == 1) pk[i] = -dt * A0_times_i(st->T0, y1, i) - dt * c0[i];
      else pk[i] = -dt * c0[i];
    }
  for (i = 0; i < 2 * UM_N * UM_NY; ++i) u[i] = l[i];
  for (k = 0; k < UM_N; ++k) { l[2 * UM_N * UM_NY + k] = -st->T0; u[2 * UM_N * UM_NY + k] = par->Tmax - st->T0; }
  /* constraint matrix data, in Ax_idx order: T0dt x3(N-2), dt x18, s0 x3 copies, Btau x3 copies */
  offs = 0;
  for (i = 0; i < 3 * (UM_N - 2); ++i) Axd[offs++] = dt * st->T0;
  for (i = 0; i < 6 * UM_N; ++i) Axd[offs++] = dt;
  for (k = 0; k < UM_N; ++k) for (i = 0; i < 3; ++i) Axd[offs++] = dt * s0[i];
  for (k = 0; k < UM_N; ++k) for (i = 0; i < 6; ++i) Axd[offs++] = dt * Btau[i];
  /* objective */
  offs = 0;
  for (k = 0; k < UM_N; ++k) for (i = 0; i < 6; ++i, ++offs) {
    Px[offs] = k == UM_N - 1 ? par->Qyf[i] : par->Qyr[i]; q[offs] = -Px[offs] * ydes[i]; }
  for (k = 0; k < UM_N; ++k) for (i = 0; i < 6; ++i, ++offs) {
    Px[offs] = k == UM_N - 1 ? par->Qdyf[i] : par->Qdyr[i]; q[offs] = -Px[offs] * dydes[i]; }
  for (k = 0; k < UM_N; ++k) for (i = 0; i < 3; ++i, ++offs) { Px[offs] = par->R[i]; q[offs] = 0; }
  memcpy(st->l0, l, sizeof(l)); memcpy(st->u0, u, sizeof(u)); memcpy(st->q0, q, sizeof(q));
  memcpy(st->Px0, Px, sizeof(Px)); memcpy(st->Axd, Axd, sizeof(Axd));

  /* OSQP updates in the reference's order (uprightmpc2.c:247-250) */
  oq_update_bounds(qp, &par->set, l, u);
  oq_update_lin_cost(qp, q);
  oq_unscale_data(qp);
  for (i = 0; i < UM_NX; ++i) qp->P[i * OQ_NMAX + i] = Px[i];
  for (i = 0; i < UM_NADATA; ++i) { int e = pat->ax_idx[i]; qp->A[pat->row[e] * OQ_NMAX + pat->col[e]] = Axd[i]; }
  oq_scale_and_factor(qp, &par->set);
  oq_solve(qp, &par->set);

  /* extraction */
  {
    real uq0 = qp->sol_x[2 * UM_NY * UM_N], dy1[6], dq1[6], t0, t1;
    st->T0 += uq0;
    uquad_o[0] = (double)st->T0;
    uquad_o[1] = (double)qp->sol_x[2 * UM_NY * UM_N + 1];
    uquad_o[2] = (double)qp->sol_x[2 * UM_NY * UM_N + 2];
    for (i = 0; i < 6; ++i) dy1[i] = qp->sol_x[UM_NY * UM_N + i];
    for (i = 0; i < 3; ++i) dq1[i] = dy1[i];
    /* e3h * R0' * v : t = R0' v ; e3h t = (-t1, t0, 0) */
    t0 = R0[0] * dy1[3] + R0[1] * dy1[4] + R0[2] * dy1[5];
    t1 = R0[3] * dy1[3] + R0[4] * dy1[4] + R0[5] * dy1[5];
    dq1[3] = -t1; dq1[4] = t0; dq1[5] = 0;
    for (i = 0; i < 6; ++i) accdes_o[i] = (double)((dq1[i] - dq0[i]) / dt);
  }
  return 0;
}

/* accessors: out sizes l39,u39,q45,Px45,Axd48 (vectors()/matrices(), uprightmpc2py.cpp:46-51) */
void orc_umpc_get_qpdata(const umpc_state_t *st, double *l, double *u, double *q, double *Px, double *Axd) {
  int i;
  for (i = 0; i < UM_NC; ++i) { l[i] = st->l0[i]; u[i] = st->u0[i]; }
  for (i = 0; i < UM_NX; ++i) { q[i] = st->q0[i]; Px[i] = st->Px0[i]; }
  for (i = 0; i < UM_NADATA; ++i) Axd[i] = st->Axd[i];
}
void orc_umpc_get_solution(const umpc_state_t *st, double *x, double *y, double *info4) {
  int i;
  for (i = 0; i < UM_NX; ++i) x[i] = st->qp.sol_x[i];
  for (i = 0; i < UM_NC; ++i) y[i] = st->qp.sol_y[i];
  info4[0] = st->qp.iter; info4[1] = st->qp.status; info4[2] = st->qp.pri_res; info4[3] = st->qp.dua_res;
}
void orc_umpc_get_iterates(const umpc_state_t *st, double *x, double *z, double *y, double *D, double *E, double *c) {
  int i;
  for (i = 0; i < UM_NX; ++i) { x[i] = st->qp.x[i]; D[i] = st->qp.D[i]; }
  for (i = 0; i < UM_NC; ++i) { z[i] = st->qp.z[i]; y[i] = st->qp.y[i]; E[i] = st->qp.E[i]; }
  *c = st->qp.c;
}
void orc_umpc_get_ctype(const umpc_state_t *st, int *ctype) { memcpy(ctype, st->qp.ctype, sizeof(int) * UM_NC); }
double orc_umpc_get_T0(const umpc_state_t *st) { return (double)st->T0; }
void orc_umpc_set_T0(umpc_state_t *st, double T0) { st->T0 = (real)T0; }
