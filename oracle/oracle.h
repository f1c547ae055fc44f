/*
 * oracle/oracle.h -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement (plain C) of the robobee3d control hot path, written from the
 * reference's behaviour; used ONLY by tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs as the CHECKER.  The product
 * (robobee3d_b200/) never includes, links or calls anything in oracle/.
 *
 * Parity status: PINNED against the reference's own C code compiled here
 * (oracle/_ref, see oracle/Makefile) by tests/test_oracle_vs_ref.py and against
 * the golden vectors in tests/golden/ that were generated from oracle/_ref by
 * tests/golden/make_golden.py.
 *
 * Compile twice: -DORACLE_F32 (real=float, the C path's arithmetic,
 * template/uprightmpc2/osqp_configure.h DFLOAT) and without (real=double, the
 * Python/desktop-OSQP path's arithmetic, template_controllers.py:190-191).
 * All exported entry points take/return double arrays so ctypes signatures do
 * not depend on the flavour.
 */
#ifndef ORACLE_H
#define ORACLE_H

#ifdef ORACLE_F32
typedef float real;
#define R_SQRT sqrtf
#define R_SIN sinf
#define R_COS cosf
#else
typedef double real;
#define R_SQRT sqrt
#define R_SIN sin
#define R_COS cos
#endif

/* sizes of the shipped instance (template/uprightmpc2/uprightmpc2.h:19-25) */
#define UM_N 3
#define UM_NY 6
#define UM_NU 3
#define UM_NX (UM_N * (2 * UM_NY + UM_NU)) /* 45 */
#define UM_NC (2 * UM_N * UM_NY + UM_N)    /* 39 */
#define UM_NNZA 111
#define UM_NADATA 48
#define UM_NK (UM_NX + UM_NC) /* 84 */

/* OSQP status values (template/uprightmpc2/constants.h:17-31) */
#define ST_SOLVED 1
#define ST_SOLVED_INACCURATE 2
#define ST_MAX_ITER (-2)
#define ST_PRIM_INF (-3)
#define ST_DUAL_INF (-4)
#define ST_NON_CVX (-7)
#define ST_UNSOLVED (-10)

/* ---- generic OSQP 0.6.0-embedded restatement on dense storage ------------ */
#ifndef OQ_NMAX /* the dir-tran oracle (dirtran_oracle.c) compiles the same code with larger caps */
#define OQ_NMAX UM_NX
#define OQ_MMAX UM_NC
#endif
#define OQ_KMAX (OQ_NMAX + OQ_MMAX)

typedef struct {
  int max_iter, check_termination, adaptive_rho_interval, scaling_iters;
  real rho0, sigma, alpha, eps_abs, eps_rel, adaptive_rho_tol;
} oq_set_t;

typedef struct {
  int n, m;
  /* scaled problem data as OSQP holds it between calls */
  real P[OQ_NMAX * OQ_NMAX]; /* full symmetric, row-major */
  real A[OQ_MMAX * OQ_NMAX]; /* row-major m x n; structural zeros are 0 */
  real q[OQ_NMAX], l[OQ_MMAX], u[OQ_MMAX];
  real D[OQ_NMAX], Dinv[OQ_NMAX], E[OQ_MMAX], Einv[OQ_MMAX], c, cinv;
  real rho, rho_vec[OQ_MMAX], rho_inv_vec[OQ_MMAX];
  int ctype[OQ_MMAX];
  /* scaled iterates carried across calls (warm start; never rescaled) */
  real x[OQ_NMAX], z[OQ_MMAX], y[OQ_MMAX];
  /* KKT factor: unit-lower L (row-major) and diagonal */
  real L[OQ_KMAX * OQ_KMAX], Dg[OQ_KMAX];
  /* results of the last solve */
  real sol_x[OQ_NMAX], sol_y[OQ_MMAX];
  int iter, status, rho_updates;
  real pri_res, dua_res;
} oq_t;

void oq_init(oq_t *w, const oq_set_t *s, int n, int m, const real *P0, const real *A0);
int oq_update_bounds(oq_t *w, const oq_set_t *s, const real *l, const real *u);
void oq_update_lin_cost(oq_t *w, const real *q);
void oq_unscale_data(oq_t *w);
void oq_scale_and_factor(oq_t *w, const oq_set_t *s);
void oq_solve(oq_t *w, const oq_set_t *s);

/* ---- UprightMPC2 ---------------------------------------------------------- */
typedef struct {
  real dt, g, Tmax;
  real Qyr[6], Qyf[6], Qdyr[6], Qdyf[6], R[3];
  real Ibi[3];
  oq_set_t set; /* workspace.c:561 + uprightmpc2.c:116-117 overrides */
} umpc_par_t;

typedef struct {
  real T0;
  oq_t qp;
  /* unscaled QP data of the last build (for the data-level parity check,
   * template_controllers.py:318-326) */
  real l0[UM_NC], u0[UM_NC], q0[UM_NX], Px0[UM_NX], Axd[UM_NADATA];
} umpc_state_t;

/* (row, col) of each of the 111 CSC entries of A and the 48 rewritten ones */
typedef struct {
  int row[UM_NNZA], col[UM_NNZA];
  int ax_idx[UM_NADATA];
  real init_val[UM_NNZA]; /* placeholders of template_controllers.py:28-63 */
  int built;
} umpc_pattern_t;

const umpc_pattern_t *orc_umpc_pattern(void);

#endif
