/*
 * include/dirtran_b200.h -- C-ABI of the batched planar direct-transcription MPC in libumpc_b200.so
 * (BASELINE config 5).
 *
 * What it replaces in the reference (citations under /root/reference/):
 *   controlutils/py/mpc.py:10-91        LTVMPC(model, N, wx, wu, **settings) / .update(x0, xr, costMode=FINAL)
 *   controlutils/py/ltvsystem.py:96-350 LTVDirTran/LTVSolver: constraint build, updateTrajectory, OSQP update+solve
 *   controlutils/py/csc.py:6-54,102-121 the A pattern and its Ad/Bd block updates
 *   planar/FlappingModels.py:27-129     PlanarThrustStrokeDev(rescale=True): dydt, getLinearDynamics, limits
 *   planar/mpc_thrust_strokedev.py:97-199  runSim(CTRL_MPC): the closed loop (update, int(dt/SIM_DT) Euler substeps)
 *
 * One LTVMPC per environment (own warm start, own adapted rho, own previous input), B environments per call, all
 * float64 like the reference's Python/OSQP path.  x = (y_0..y_N, u_0..u_{N-1}), nx = 6, nu = 2, N in {3,5,10,15,20}.
 * OSQP settings as mpc_thrust_strokedev.py:124: scaling 0, polish off, warm start; the rho adaptation interval is a
 * parameter (OSQP derives it from wall-clock timing of its own setup, which is not reproducible; 25 = the floor that
 * rule has, osqp.c of 0.5.0).
 *
 * Error convention: 0 ok, UMPC_ERR_* of umpc_b200.h otherwise, message via umpcLastError().  Solve quality per
 * environment via dtBatchGetInfo (LTVSolver.solve raises on status != solved, ltvsystem.py:312-316; a batch cannot, so
 * the caller checks status).  No CPU fallback.
 */
#ifndef DIRTRAN_B200_H
#define DIRTRAN_B200_H

#ifdef __cplusplus
extern "C" {
#endif

#define DT_NX 6
#define DT_NU 2

typedef struct {
  int N;                 /* horizon: 3, 5, 10, 15 or 20 */
  double dt;             /* MPC step (model.dt), mpc_thrust_strokedev.py:30 */
  double wx[DT_NX];      /* state weights AFTER the /dt of mpc_thrust_strokedev.py:118 */
  double wu[DT_NU];
  double eps_abs, eps_rel, rho, sigma, alpha;
  int max_iter, check_termination, adaptive_rho_interval; /* adaptive_rho_interval 0: rho fixed */
  double adaptive_rho_tol;
  /* model (FlappingModels.py:29-50,104-112 with rescale=True) */
  double g, ibmb, umin[DT_NU], umax[DT_NU], xmin[DT_NX], xmax[DT_NX]; /* +-inf allowed */
  /* plant (mpc_thrust_strokedev.py:113,189-192) */
  double sim_dt;
  int nsub;
} DtParams;

typedef struct DtBatch DtBatch;

/* the 'hover' parameter set of mpc_thrust_strokedev.py:30 + OSQP defaults; N as given */
void dtBatchDefaults(DtParams *p, int N);
int dtBatchCreate(DtBatch **out, int nenv, int device, const DtParams *par);
int dtBatchDestroy(DtBatch *b);
/* fresh controllers (x = z = y = 0, rho = par.rho, previous input 0) */
int dtBatchReset(DtBatch *b);
/* plant state y[B][6], previous input ctrl[B][2] (NULL: keep), goal xr[B][6] (NULL: keep); host pointers */
int dtBatchSetState(DtBatch *b, const double *y, const double *ctrl, const double *xr);
int dtBatchGetState(DtBatch *b, double *y, double *ctrl);
/* one LTVMPC.update per environment at the stored goal: x0[B][6] in, u[B][2] out (host pointers; copies inside) */
int dtBatchUpdate(DtBatch *b, const double *x0, double *u);
/* nsteps of runSim's loop on the device: update -> nsub Euler substeps; nothing crosses PCIe */
int dtBatchRollout(DtBatch *b, int nsteps);
/* info[B][6] = iter, status_val, pri_res, dua_res, rho, rho_updates of the last solve */
int dtBatchGetInfo(DtBatch *b, double *info);
/* last solution in OSQP's ordering: x[B][n], y[B][m] (either may be NULL); n = (N+1)6 + 2N, m = n + (N+1)6 */
int dtBatchGetSolution(DtBatch *b, double *x, double *y);
/* sum over environments of iterations / solves / non-solved statuses since the last reset: out[3] */
int dtBatchCounters(DtBatch *b, double *out);
/* CUDA-event time of the last kernel launch on the batch's stream, ms */
float dtBatchLastKernelMs(DtBatch *b);
int dtBatchSync(DtBatch *b);

#ifdef __cplusplus
}
#endif
#endif
