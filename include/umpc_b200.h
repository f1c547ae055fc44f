/*
 * include/umpc_b200.h -- C-ABI of libumpc_b200.so
 *
 * B200 (sm_100a) implementation of robobee3d's control hot path behind the
 * reference's own C controller API.  Two groups of entry points:
 *
 *  (1) Drop-in, signature-for-signature, for the reference C controller
 *      (citations under /root/reference/template/):
 *        umpcInit / umpcUpdate / umpcS              uprightmpc2/uprightmpc2.h:45-49
 *        wlConInit / wlConUpdate / wlconS           uprightmpc2/funapprox.h:45-50
 *        wlControllerUpdate                         wlcontroller/c/wlcontroller.h:19
 *      Structs keep the reference's public fields and sizes (caller allocates),
 *      all float32, matrices column-major, units mm / mg / ms.  Unlike the
 *      reference (one process-global OSQP workspace, workspace.c:2620) every
 *      UprightMPC_t / WLCon_t owns its own device state, keyed by its address.
 *      Each call is synchronous: H2D copy, one kernel launch of a batch of 1, D2H copy.
 *
 *  (2) Batched: the same update for B independent environments in one launch,
 *      plus the fused closed-loop rollout (task -> MPC -> [WL] -> clip -> rigid-body
 *      step, template/uprightmpc2.py:87-159 + genqp.py:24-41 + flight_tasks.py).
 *      Host-pointer variants copy H2D/D2H inside the call; *Dev variants take
 *      device pointers and are stream-ordered.
 *
 * No CPU fallback exists: every entry point needs a CUDA device and returns
 * UMPC_ERR_CUDA (or aborts the drop-in calls with a message on stderr) without one.
 */
#ifndef UMPC_B200_H
#define UMPC_B200_H

#ifdef __cplusplus
extern "C" {
#endif

/* ---- (1) reference-compatible single-controller API ------------------------------------ */
#define UMPC_N 3
#define UMPC_NY 6
#define UMPC_NU 3
#ifndef UMPC_NX
#define UMPC_NX (UMPC_N * (2 * UMPC_NY + UMPC_NU)) /* 45 */
#define UMPC_NC (2 * UMPC_N * UMPC_NY + UMPC_N)    /* 39 */
#endif
#define UMPC_nAdata 48

/* same fields, order and sizes as uprightmpc2.h:27-43 */
typedef struct {
  float dt, g, Tmax;
  float Qyr[6], Qyf[6], Qdyr[6], Qdyf[6], R[3];
  float smin[3], smax[3];
  float e3h[3 * 3];
  float e3hIbi[3 * 3];
  float l[UMPC_NC], u[UMPC_NC], q[UMPC_NX];
  float Px_data[UMPC_NX];
  float Ax_data[UMPC_nAdata];
  int Ax_idx[UMPC_nAdata], nAxT0dt, nAxdt;
  float c0[UMPC_NY];
  float T0;
} UprightMPC_t;

void umpcInit(UprightMPC_t *up, float dt, float g, float TtoWmax, float ws, float wds, float wpr, float wpf,
              float wvr, float wvf, float wthrust, float wmom, const float Ib[/* 3 */], int maxIter);

/* returns 0 like osqp_solve's exit flag (uprightmpc2.c:250,271); solve quality via umpcGetInfo */
int umpcUpdate(UprightMPC_t *up, float uquad[/* 3 */], float accdes[/* 6 */], const float p0[/* 3 */],
               const float R0[/* 9 */], const float dq0[/* 6 */], const float pdes[/* 3 */],
               const float dpdes[/* 3 */], const float sdes[/* 3 */], float actualT0);

void umpcS(float uquad_y1[/* 3 */], float accdes_y2[/* 6 */], const float p0_u1[/* 3 */], const float R0_u2[/* 9 */],
           const float dq0_u3[/* 6 */], const float pdes_u4[/* 3 */], const float dpdes_u5[/* 3 */],
           const float sdes_u6[/* 3 */], float dt_u7, float g_u8, float TtoWmax_u9, float ws_u10, float wds_u11,
           float wpr_u12, float wpf_u13, float wvr_u14, float wvf_u15, float wthrust_u16, float wmom_u17,
           const float Ib_u18[/* 3 */], int maxIter_u19, float actualT0_u20);

/* extras the reference exposes only through its global workspace (workspace.info / .solution):
 * info4 = iter, status_val, pri_res, dua_res (types.h:64-89); x[45], y[39] = workspace.solution */
void umpcGetInfo(const UprightMPC_t *up, float info4[/* 4 */]);
void umpcGetSolution(const UprightMPC_t *up, float x[/* 45 */], float y[/* 39 */]);
/* release the device state attached to *up (the reference has nothing to free) */
void umpcFree(UprightMPC_t *up);

#define NDELU 4
typedef struct { /* uprightmpc2/funapprox.h:18-23 */
  int k;
  float a0;
  float a1[NDELU];
  float A2[NDELU * NDELU];
} FunApprox_t;

typedef struct { /* uprightmpc2/funapprox.h:39-43 */
  float u0[NDELU], umin[NDELU], umax[NDELU], dumax[NDELU];
  float Qw[6 * 6];
  FunApprox_t fa[6];
} WLCon_t;

void wlConInit(WLCon_t *wl, const float u0[/* 4 */], const float umin[/* 4 */], const float umax[/* 4 */],
               const float dumax[/* 4 */], const float Qw[/* 6 */], float controlRate, const float popts[/* 90 */]);
void wlConUpdate(WLCon_t *wl, float u1[/* 4 */], float w0[/* 6 */], const float h0[/* 6 */], const float pdotdes[/* 6 */]);
void wlconS(float u1_y1[/* 4 */], float w0_y2[/* 6 */], const float u0init_u1[/* 4 */], const float umin_u2[/* 4 */],
            const float umax_u3[/* 4 */], const float dumax_u4[/* 4 */], const float Qw_u5[/* 6 */],
            float controlRate_u6, const float popts_u7[/* 90 */], const float h0_u8[/* 6 */],
            const float pdotdes_u9[/* 6 */]);

/* true WLQP with the lazily initialised process-global state of wlcontroller.c:17-45 */
void wlControllerUpdate(float *u, const float *u0init, const float *h0, const float *pdotdes, const float *popts,
                        float controlRate);
/* same, also returning w0 = w(u0) like the pybind module (wlcontroller/pybind/wlqppy.cpp:15-19) */
void wlControllerUpdateW(float *u, float *w0, const float *u0init, const float *h0, const float *pdotdes,
                         const float *popts, float controlRate);
/* forget the global WLQP state (tests; the reference needs a fresh process for that) */
void wlControllerReset(void);

/* The WLQP object itself (wlcontroller/c/wlqp.h:21-31), same struct fields and order.  The reference's wlqpUpdate
 * calls the application's wrenchMap / wrenchJacMap (wlqp.h:18-19); host callbacks cannot run on the device, so the
 * quadratic wrench model popts[90] is installed with wlqpSetWrenchModel (or by the first wlControllerUpdate). */
typedef struct {
  float umin[NDELU], umax[NDELU], U0[NDELU];
  float w0[6];
  float Q[6 * 6];
} WLQP_t;
void wlqpInit(WLQP_t *wlqp, float controlRate);
void wlqpUpdate(WLQP_t *wlqp, float *u, const float *u0, const float *h0, const float *pdotdes);
void wlqpSetWrenchModel(const float popts[/* 90 */]);
/* scalar quadratic model (uprightmpc2/funapprox.h:31-35 == wlcontroller/c/funapprox.h:31-35); F and Df evaluate on the device */
void funApproxInit(FunApprox_t *fa, const float popts[/* 15 */]);
float funApproxF(const FunApprox_t *fa, const float *xi);
void funApproxDf(float *Df, const FunApprox_t *fa, const float *xi);

/* ---- (2) batched API -------------------------------------------------------------------- */
#define UMPC_OK 0
#define UMPC_ERR_CUDA 100   /* no device / CUDA runtime error: message via umpcLastError() */
#define UMPC_ERR_ARG 101

typedef struct UmpcBatch UmpcBatch; /* opaque: device SoA state of B environments */

/* controller + plant + WL parameters of a batch; zero-init then fill (umpcBatchDefaults) */
typedef struct {
  /* umpcInit arguments (createMPC defaults, template_controllers.py:260-280) */
  float dt, g, TtoWmax, ws, wds, wpr, wpf, wvr, wvf, wthrust, wmom, Ib[3];
  int maxIter;
  /* controlTest arguments (uprightmpc2.py:87) */
  double dtsim, hlInterval /* <0: None */; /* double: the firing schedule t - thlPrev > hlInterval is evaluated like numpy */
  float taulim, tpert /* <0: None */, mb;
  /* WL stage (robobee_test_controllers.py:71,76): 0 none, 1 WLCon, 2 WLQP */
  int wlMode;
  float wl_u0[4], wl_umin[4], wl_umax[4], wl_dumax[4], wl_Qw[6], wl_controlRate, wl_popts[90];
  int wlqpCheckTermination; /* 25 = C (wlqp.c), 0 = C++ (wlqp.cpp:63-66) */
  /* 1: the task's initialPos has integer dtype, as in controlTest (uprightmpc2.py:94-104 builds p with
   * np.array([0, 0, 0]); flight_tasks' in-place `posdes[i] += x` then truncates pdes toward zero). 0: float. */
  int intInitialPos;
  /* Converged mode = the reference's PYTHON path (template_controllers.py:190-258: osqp on float64, eps_abs = eps_rel
   * = 1e-4, OSQP's defaults otherwise).  checkTermination > 0 selects it: the batch then solves in float64, tests the
   * residuals every checkTermination iterations (OSQP default 25), stops each QP when they pass or at maxIter (OSQP
   * default 4000) and re-estimates rho every adaptiveRhoInterval iterations (0 = never; OSQP's automatic interval
   * works out to 100 for this QP).  0 (default): the C path -- float32, exactly maxIter iterations (uprightmpc2.c:116-117). */
  int checkTermination;
  int adaptiveRhoInterval;
  float epsAbs, epsRel; /* <= 0: 1e-4 */
} UmpcBatchParams;

void umpcBatchDefaults(UmpcBatchParams *par);
const char *umpcLastError(void);

int umpcBatchCreate(UmpcBatch **out, int nenv, int device, const UmpcBatchParams *par);
void umpcBatchDestroy(UmpcBatch *b);
int umpcBatchSize(const UmpcBatch *b);
/* cold-start every controller (x = z = y = 0, T0 = 0, D = E = 1) and zero the metrics */
int umpcBatchReset(UmpcBatch *b);

/* One umpcUpdate per environment.  Host arrays, row e = environment e:
 * p0[B][3], R0[B][9] (column-major 3x3 each), dq0[B][6], pdes[B][3], dpdes[B][3], sdes[B][3],
 * actualT0[B] or NULL (= -1: keep internal T0) -> uquad[B][3], accdes[B][6]. */
int umpcBatchUpdate(UmpcBatch *b, float *uquad, float *accdes, const float *p0, const float *R0, const float *dq0,
                    const float *pdes, const float *dpdes, const float *sdes, const float *actualT0);
/* One closed-loop env-step per environment with HOST buffers: umpcUpdate, torque clip, rigid-body step
 * (the batched twin of uprightmpc2.py:139-151).  Outputs may alias the inputs. */
int umpcBatchEnvStep(UmpcBatch *b, float *uquad, float *accdes, float *p_next, float *R_next, float *dq_next,
                     const float *p0, const float *R0, const float *dq0, const float *pdes, const float *dpdes,
                     const float *sdes, const float *actualT0);
/* same with DEVICE pointers in the same [B][k] layout (arrays of exactly B rows: padded lanes never touch them),
 * asynchronous on the batch's PRIVATE stream: the caller orders it against its own work (umpcBatchSync, or the
 * "stream" handle of umpcBatchDevPtr) */
int umpcBatchUpdateDev(UmpcBatch *b, float *uquad, float *accdes, const float *p0, const float *R0, const float *dq0,
                       const float *pdes, const float *dpdes, const float *sdes, const float *actualT0);
/* same, ordered against `caller_stream` (a cudaStream_t, e.g. torch.cuda.current_stream().cuda_stream): the update
 * starts after everything already queued there and later work there sees its outputs; no host synchronisation */
int umpcBatchUpdateDevOn(UmpcBatch *b, void *caller_stream, float *uquad, float *accdes, const float *p0, const float *R0,
                         const float *dq0, const float *pdes, const float *dpdes, const float *sdes, const float *actualT0);
/* per-environment solver diagnostics of the last update: host arrays, any may be NULL.
 * info[B][4] = iter, status_val, pri_res, dua_res; x[B][45], y[B][39] = OSQP solution;
 * xs[B][45], zs[B][39], ys[B][39] = scaled warm-start iterates; D[B][45], E[B][39] */
int umpcBatchGetSolve(UmpcBatch *b, float *info, float *x, float *y, float *xs, float *zs, float *ys, float *D,
                      float *E);
/* QP data of the last update, qp[B][216] = l[39] u[39] q[45] Px[45] Ax_data[48] */
int umpcBatchGetQPData(UmpcBatch *b, float *qp);

/* Closed loop.  Plant state (host arrays): p[B][3], R[B][9] col-major, dq[B][6];
 * task[B] (0 helix, 1 straightAcc, 2 flip, 3 perch), tp[B][4] task parameters (flight_tasks.py order).
 * SetState also records p as the task's initialPos and rewinds time to 0. */
int umpcBatchSetState(UmpcBatch *b, const float *p, const float *R, const float *dq, const int *task, const float *tp);
int umpcBatchGetState(UmpcBatch *b, float *p, float *R, float *dq, float *uquad);
/* advance every environment nsteps plant steps (MPC fired per hlInterval), asynchronous */
int umpcBatchRollout(UmpcBatch *b, int nsteps);
/* metrics[B][4] = sum |p|^2, sum |tau|^2 (logMetric, uprightmpc2.py:161-175), #solves, #steps */
int umpcBatchGetMetrics(UmpcBatch *b, float *metrics);
/* on-device reduction over the batch: sums8 = sum err, sum eff, sum solves, sum steps,
 * #envs with non-finite state, max err, 0, 0 (double; what ranks all-reduce) */
int umpcBatchReduceMetrics(UmpcBatch *b, double *sums8);
/* optional logging of env `env` during rollouts: y[nt][12] = p, R[:,2], dq ; u[nt][3] */
int umpcBatchSetLog(UmpcBatch *b, int env, int capacity_steps);
int umpcBatchGetLog(UmpcBatch *b, float *y, float *u, int nsteps);
int umpcBatchSync(UmpcBatch *b);
/* device pointers for zero-copy use from torch / other CUDA code.  The state is a tiled struct of arrays
 * float[B/32][ROWS][32] ("slab"): the named pointers ("p", "R", "xs", ...) address the row block of env 0; element i
 * of env e is at ptr[(e / 32) * ROWS * 32 + i * 32 + e % 32] (ROWS: robobee3d_b200/csrc/umpc_dev.cuh).
 * "stage_in"/"stage_out" are the [env][k] staging buffers of the host-buffer calls, "stream" the batch's stream. */
void *umpcBatchDevPtr(UmpcBatch *b, const char *name);
/* per-environment cost weights w[B][8] = (ws, wds, wpr, wpf, wvr, wvf, wthrust, wmom): every env is its own createMPC(**kwargs)
 * (template_controllers.py:260) -- the reference's weight sweep gainTuningSims (uprightmpc2.py:272-303) as ONE batch.
 * NULL restores the batch-wide parameters.  float32 path: TMEM launch shape (128 threads per block) only. */
int umpcBatchSetWeights(UmpcBatch *b, const float *w);
/* CUDA-event time (ms) of the last kernel sequence launched by the Update or Rollout calls */
float umpcBatchLastKernelMs(UmpcBatch *b);
/* CUDA events on the batch's stream around any sequence of calls: Start records, StopMs records + waits */
int umpcBatchTimerStart(UmpcBatch *b);
float umpcBatchTimerStopMs(UmpcBatch *b);
/* Rollouts do not evaluate the per-solve residuals / status / un-scaled solution (GetSolve's info, x, y) by default:
 * nothing on the closed-loop path reads them (the reference never acts on them either, uprightmpc2.c:116-117; the
 * divergence guard of auxil.c:538-561 stays).  The same holds for umpcBatchEnvStep on pinned buffers (a step of a closed
 * loop driven from the host).  on = 1: evaluate them after every solve of a rollout / env-step, like Update does. */
int umpcBatchSetRolloutInfo(UmpcBatch *b, int on);
/* launch configuration knobs (threads per block) for tuning/benchmarks */
int umpcBatchSetLaunch(UmpcBatch *b, int threads_per_block);
/* FP32 FMA-pipe peak of the device, measured with a register-resident FFMA loop (TFLOP/s) */
double umpcMeasureFp32Peak(int device, int iters);
/* FP64 pipe peak, DFMA loop (TFLOP/s): the roofline denominator of the float64 kernels (converged mode, dir-tran, gqp) */
double umpcMeasureFp64Peak(int device, int iters);

#ifdef __cplusplus
}
#endif
#endif
