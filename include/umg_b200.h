/*
 * include/umg_b200.h -- C-ABI of the horizon-generic UprightMPC2 controller of libumpc_b200.so ("umg")
 *
 * The same controller as umpcUpdate (include/umpc_b200.h) for ANY horizon N -- the role OSQP's codegen plays for the
 * reference, whose C build hard-wires N = 3 (template/uprightmpc2/uprightmpc2.h:19-25 "must be changed along with the
 * autogen code"; template_controllers.py:169-258 takes N as an argument) -- and, with withWL = 1, the FUSED MPC + WLQP
 * single QP of template/uprightmpc2_withWLQP.py (x gains the 4 wing-input increments, 4 rate-limit rows, dense P blocks
 * M0t'QwM0t, -M0t'Qw dwdu, dwdu'Qw dwdu: nx = 49, nc = 43 at N = 3; OSQP eps 1e-7, :243-244).
 * Per environment and call: one CUDA thread writes the QP data (l, u, Ax, dense P, q; template_controllers.py:65-143 /
 * uprightmpc2_withWLQP.py:70-191), one warp solves it (gqp_b200.h, float64, converged like the Python path), one thread
 * extracts uquad / accdes and advances the controller state (T0 += dT, u0 += du; :301-308).
 */
#ifndef UMG_B200_H
#define UMG_B200_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct UmgBatch UmgBatch;

typedef struct {
  int N;                        /* horizon (UprightMPC2's N: 3 in the C build, any >= 2 here) */
  double dt, g, TtoWmax, ws, wds, wpr, wpf, wvr, wvf, wthrust, wmom, Ib[3];   /* createMPC, template_controllers.py:260-280 */
  /* OSQP settings of the Python path (template_controllers.py:190-191; eps 1e-7 in uprightmpc2_withWLQP.py:243) */
  int maxIter, checkTermination, adaptiveRhoInterval;
  double epsAbs, epsRel;
  /* fused MPC + WLQP (uprightmpc2_withWLQP.py:218-262) */
  int withWL;
  double umin[4], umax[4], dumax[4], Qw[6], controlRate, mb;
  /* wrench model w(u): 0 = the reference's stand-in w0 = 0, dw/du = [0 (2x4); I4] (uprightmpc2_withWLQP.py:268-271);
   * 1 = the quadratic FunApprox model popts[90] (uprightmpc2/funapprox.c:35-96): w0 = w(u0), dw/du = A1(u0) */
  int wrenchModel;
  double popts[90];
} UmgParams;

void umgDefaults(UmgParams *p, int N, int withWL);
/* sizes of the QP this horizon gives: nx = N (2 ny + nu) (+4), nc = 2 N ny + N (+4), nnz(A) */
void umgSizes(int N, int withWL, int *nx, int *nc, int *nnz);
/* CSR pattern of A (initConstraint, template_controllers.py:28-63 / uprightmpc2_withWLQP.py:23-68); arrays of nc + 1 and nnz */
void umgPattern(int N, int withWL, int *rowptr, int *colidx);

int umgCreate(UmgBatch **out, int nenv, int device, const UmgParams *par);
void umgDestroy(UmgBatch *b);
int umgReset(UmgBatch *b);   /* T0 = 0, u0 = 0, cold solver */
/* controller state: T0[B] (specific thrust), u0[B][4] (wing inputs of the fused WLQP); either may be NULL (= keep) */
int umgSetState(UmgBatch *b, const double *T0, const double *u0);
/* One update per environment, host arrays like umpcBatchUpdate: p0[B][3], R0[B][9] column-major, dq0[B][6], pdes[B][3],
 * dpdes[B][3], sdes[B][3], actualT0[B] or NULL (< 0: keep T0) -> uquad[B][3], accdes[B][6], u_wl[B][4] (withWL; may be NULL) */
int umgUpdate(UmgBatch *b, double *uquad, double *accdes, double *u_wl, const double *p0, const double *R0, const double *dq0,
              const double *pdes, const double *dpdes, const double *sdes, const double *actualT0);
/* the last update's QP and solution: P[B][nx*nx], Av[B][nnz], q[B][nx], l[B][nc], u[B][nc], x[B][nx], y[B][nc],
 * info[B][6] = iter, status, pri_res, dua_res, rho, rho_updates; any may be NULL */
int umgGetQP(UmgBatch *b, double *P, double *Av, double *q, double *l, double *u, double *x, double *y, double *info);
float umgLastSolveMs(UmgBatch *b);

#ifdef __cplusplus
}
#endif
#endif
