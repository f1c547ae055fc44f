/*
 * include/gqp_b200.h -- C-ABI of the generic batched QP solver of libumpc_b200.so ("gqp")
 *
 * One WARP per QP:  min 1/2 x'Px + q'x  s.t.  l <= Ax <= u, all QPs of a batch sharing one sparsity pattern of A
 * (CSR) and a dense symmetric P.  The algorithm is OSQP 0.6.0's, as the reference calls it through the `osqp` Python
 * package (template/template_controllers.py:190-191,216-217; template/uprightmpc2_withWLQP.py:243-244,283-284;
 * planar/mpc_osqp.py:104-140): Ruiz equilibration (scaling.c:44-156), per-row rho by constraint type
 * (auxil.c:103-146), ADMM with relaxation (auxil.c:164-228), termination test on un-scaled residuals
 * (auxil.c:243-361,684-790), adaptive rho (auxil.c:13-75), warm start kept in scaled coordinates across calls.
 * The linear system is the reduced one, K = P + sigma I + A' diag(rho) A (dense Cholesky in shared memory, triangular
 * solves by the warp), algebraically the quasi-definite KKT system of kkt.c:184-222 with nu eliminated.
 *
 * It is the cooperative (warp-per-QP) mapping of the path: the shapes the thread-per-QP kernels of umpc_b200.h hard-wire
 * (UprightMPC2 N = 3, diagonal P) do not apply to it, so it also runs
 *   - UprightMPC2 for any horizon N (template/uprightmpc2/uprightmpc2.h:19-25 "must be changed along with the autogen code"),
 *   - the fused MPC + WLQP single QP (template/uprightmpc2_withWLQP.py: nx = 49, nc = 43, dense P blocks, eps 1e-7),
 *   - the LTI quadcopter dir-tran QP of planar/mpc_osqp.py (nx = 12, nu = 4, N = 5: n = 92, m = 164).
 * float64 throughout.  No CPU fallback: every entry point needs a CUDA device (UMPC_ERR_CUDA otherwise).
 */
#ifndef GQP_B200_H
#define GQP_B200_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct GqpBatch GqpBatch;

typedef struct {
  int n, m, nnz;                 /* variables, constraints, non-zeros of A */
  int max_iter, check_termination, adaptive_rho_interval, scaling_iters;
  double rho, sigma, alpha, eps_abs, eps_rel, adaptive_rho_tol;
  double infty;                  /* OSQP_INFTY the bounds are clipped to: 1e30 (osqp 0.6.0, constants.h:100), 1e20 (0.5.0) */
} GqpSettings;

/* OSQP's defaults (osqp 0.6.0 constants.h:55-100): rho 0.1, sigma 1e-6, alpha 1.6, eps 1e-3, max_iter 4000,
 * check_termination 25, scaling 10, adaptive_rho_tolerance 5, adaptive_rho_interval 0 (callers set 25..100) */
void gqpDefaults(GqpSettings *s, int n, int m, int nnz);

/* rowptr[m + 1], colidx[nnz]: CSR pattern of A shared by the whole batch (columns increasing inside a row) */
int gqpBatchCreate(GqpBatch **out, int nqp, int device, const GqpSettings *s, const int *rowptr, const int *colidx);
void gqpBatchDestroy(GqpBatch *b);
/* cold start: x = z = y = 0, D = E = c = 1, rho = settings.rho */
int gqpBatchReset(GqpBatch *b);
/* host arrays, any may be NULL (= keep): P[nqp][n*n] dense symmetric row-major, Av[nqp][nnz] in CSR order, q[nqp][n],
 * l[nqp][m], u[nqp][m].  stride0 = 1: one QP's data broadcast to the whole batch (arrays hold a single QP). */
int gqpBatchSetData(GqpBatch *b, const double *P, const double *Av, const double *q, const double *l, const double *u,
                    int broadcast);
/* one OSQP update + solve per QP.  rescale != 0: P / A changed -- re-equilibrate (like osqp_update_P_A);
 * 0: only q / l / u changed -- keep the scaling of the last rescale (like osqp_update_lin_cost / _bounds). */
int gqpBatchSolve(GqpBatch *b, int rescale);
/* x[nqp][n], y[nqp][m], info[nqp][6] = iter, status_val, pri_res, dua_res, rho, rho_updates; any may be NULL */
int gqpBatchGetSolution(GqpBatch *b, double *x, double *y, double *info);
int gqpBatchSync(GqpBatch *b);
float gqpBatchLastKernelMs(GqpBatch *b);
/* device arrays for kernels that build QP data / consume solutions in place: "P", "Av", "q", "l", "u", "x", "y", "info",
 * and "stream" */
void *gqpBatchDevPtr(GqpBatch *b, const char *name);

/* ---- LTI model-predictive loop on top (planar/mpc_osqp.py:117-140) --------------------------------------------
 * x0 <- Ad x0 + Bd u0*, l[:nx] = u[:nx] = -x0, solve again: nsteps closed-loop steps of every QP of the batch on the
 * device.  Ad[nx*nx], Bd[nx*nu] row-major; x0[nqp][nx] in/out; the first-input block of x sits at u_off.  From its first call on
 * the batch keeps every QP's factor between solves (n (n + 1) / 2 + n doubles per QP) and reuses it while only bounds change. */
int gqpLtiRollout(GqpBatch *b, int nsteps, int nx, int nu, int u_off, const double *Ad, const double *Bd, double *x0,
                  double *u_log /* [nsteps][nqp][nu] or NULL */, double *iters_sum /* or NULL */,
                  int rescale_first /* 1: the first solve equilibrates (prob.setup + first solve) */);

#ifdef __cplusplus
}
#endif
#endif
