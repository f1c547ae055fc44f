/*
 * oracle/sim_oracle.c -- TEST INFRASTRUCTURE ONLY (see oracle/oracle.h).
 *
 * Restatement of the closed loop around the controller:
 *   rigid-body step   template/genqp.py:19-41   (quadrotorNLVF / quadrotorNLDyn)
 *   tasks             template/flight_tasks.py:6-49
 *   loop + metric     template/uprightmpc2.py:87-175 (controlTest, logMetric)
 *   MPC -> WL glue    template/robobee_test_controllers.py:162-171, ca6dynamics.py:40-50
 *
 * Built two ways (oracle/Makefile):
 *   - into liboracle_f32/f64.so, driving the restated controller (umpc_oracle.c);
 *   - with -DSIM_USE_REF into oracle/_ref/libref_umpc_f32.so, driving the
 *     reference's own umpcUpdate / wlConUpdate (float controller, double plant,
 *     which is exactly what the reference's Python loop + pybind module does).
 * Time and the hlInterval gate are always evaluated in double, like numpy does.
 */
#include <math.h>
#include <string.h>

#ifdef SIM_USE_REF
#include "uprightmpc2.h"
#include "funapprox.h"
typedef double real;
#define R_SQRT sqrt
#define R_SIN sin
#define R_COS cos
#else
#include "oracle.h"
int orc_umpc_update(const umpc_par_t *par, umpc_state_t *st, double *uquad_o, double *accdes_o,
                    const double *p0d, const double *R0d, const double *dq0d, const double *pdesd,
                    const double *dpdesd, const double *sdesd, double actualT0);
void orc_wlcon_update(void *wl, double *u1, double *w0o, const double *h0, const double *pdotdes);
void orc_wlqp_update(void *wl, double *u1, double *w0o, const double *h0, const double *pdotdes);
#endif

#define PI_D 3.14159265358979323846

/* task ids */
enum { TASK_HELIX = 0, TASK_STRAIGHT = 1, TASK_FLIP = 2, TASK_PERCH = 3 };

static double clipd(double v, double lo, double hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* tp: helix (trajAmp, trajFreq, dz, useY) | straightAcc (tduration, vdes) |
 *     flip (tstart, tend) | perch (tend, trotstart, trotend, vdes) */
/* int_pos: controlTest builds its initial position with np.array([0, 0, 0]) (uprightmpc2.py:94-104), an int64
 * array; flight_tasks.py copies it (np.copy keeps the dtype) and `posdes[i] += x` stores the float sum back
 * into the int array, i.e. truncates toward zero.  dpdes / sdes are float arrays and unaffected. */
static double trunc_if(double v, int int_pos) { return int_pos ? trunc(v) : v; }

void orc_task_q(int task, double t, const double *p_init, const double *tp, int int_pos, double *pdes, double *dpdes,
                double *sdes) {
  int i;
  for (i = 0; i < 3; ++i) { pdes[i] = p_init[i]; dpdes[i] = 0; }
  sdes[0] = 0; sdes[1] = 0; sdes[2] = 1;
  if (task == TASK_HELIX) {
    double amp = tp[0], omg = 2 * PI_D * tp[1] * 1e-3;
    pdes[0] = trunc_if(pdes[0] + amp * sin(omg * t), int_pos); dpdes[0] = amp * omg * cos(omg * t);
    if (tp[3] != 0) { pdes[1] = trunc_if(pdes[1] + amp * (1 - cos(omg * t)), int_pos); dpdes[1] = amp * omg * sin(omg * t); }
    if (amp > 1e-3) { pdes[2] = trunc_if(pdes[2] + tp[2] * t, int_pos); dpdes[2] = tp[2]; }
  } else if (task == TASK_STRAIGHT) {
    dpdes[0] = t < tp[0] ? tp[1] : 0;
    pdes[0] = trunc_if(pdes[0] + tp[1] * clipd(t, 0, tp[0]), int_pos);
  } else if (task == TASK_FLIP) {
    double ph = clipd((t - tp[0]) / tp[1], 0, 1);
    sdes[0] = -sin(ph * 2 * PI_D); sdes[1] = 0; sdes[2] = cos(ph * 2 * PI_D);
  } else if (task == TASK_PERCH) {
    pdes[0] = trunc_if(pdes[0] + tp[3] * clipd(t, 0, tp[0]), int_pos);
    dpdes[0] = t < tp[0] ? tp[3] : 0;
    if (t < tp[0]) {
      double ph = clipd((t - tp[2]) / tp[1], 0, 1);
      sdes[0] = -sin(ph * PI_D); sdes[1] = 0; sdes[2] = cos(ph * PI_D);
    } else { sdes[0] = -1; sdes[1] = 0; sdes[2] = 0; }
  }
}

void orc_task(int task, double t, const double *p_init, const double *tp, double *pdes, double *dpdes, double *sdes) {
  orc_task_q(task, t, p_init, tp, 0, pdes, dpdes, sdes);
}

/* one explicit-Euler step; R column-major; expm(skew(w) dt) by Rodrigues */
static void dyn_step(real *p, real *R, real *dq, const real *u, real dt, const real *Ib, real g) {
  real dv[3], dw[3], Iw[3], E[9], Rn[9], th, a, b, kx, ky, kz;
  int i, j, k;
  for (i = 0; i < 3; ++i) dv[i] = u[0] * R[6 + i];
  dv[2] -= g;
  for (i = 0; i < 3; ++i) Iw[i] = Ib[i] * dq[3 + i];
  dw[0] = (-(dq[4] * Iw[2] - dq[5] * Iw[1]) + u[1]) / Ib[0];
  dw[1] = (-(dq[5] * Iw[0] - dq[3] * Iw[2]) + u[2]) / Ib[1];
  dw[2] = (-(dq[3] * Iw[1] - dq[4] * Iw[0])) / Ib[2];
  for (i = 0; i < 3; ++i) p[i] += dt * dq[i];
  kx = dq[3] * dt; ky = dq[4] * dt; kz = dq[5] * dt;
  th = R_SQRT(kx * kx + ky * ky + kz * kz);
  if (th < (real)1e-4) { a = (real)1.0 - th * th / 6; b = (real)0.5 - th * th / 24; }
  else { a = R_SIN(th) / th; b = ((real)1.0 - R_COS(th)) / (th * th); }
  /* E = I + a K + b K^2, column-major */
  E[0] = 1 - b * (ky * ky + kz * kz); E[4] = 1 - b * (kx * kx + kz * kz); E[8] = 1 - b * (kx * kx + ky * ky);
  E[1] = a * kz + b * kx * ky;  E[3] = -a * kz + b * kx * ky;
  E[2] = -a * ky + b * kx * kz; E[6] = a * ky + b * kx * kz;
  E[5] = a * kx + b * ky * kz;  E[7] = -a * kx + b * ky * kz;
  for (j = 0; j < 3; ++j) for (i = 0; i < 3; ++i) {
    real acc = 0;
    for (k = 0; k < 3; ++k) acc += R[i + 3 * k] * E[k + 3 * j];
    Rn[i + 3 * j] = acc;
  }
  for (i = 0; i < 9; ++i) R[i] = Rn[i];
  for (i = 0; i < 3; ++i) { dq[i] += dt * dv[i]; dq[3 + i] += dt * dw[i]; }
}

void orc_dyn_step(double *p, double *R, double *dq, const double *u, double dt, const double *Ib, double g) {
  real rp[3], rR[9], rdq[6], ru[3], rIb[3]; int i;
  for (i = 0; i < 3; ++i) { rp[i] = (real)p[i]; ru[i] = (real)u[i]; rIb[i] = (real)Ib[i]; }
  for (i = 0; i < 9; ++i) rR[i] = (real)R[i];
  for (i = 0; i < 6; ++i) rdq[i] = (real)dq[i];
  dyn_step(rp, rR, rdq, ru, (real)dt, rIb, (real)g);
  for (i = 0; i < 3; ++i) p[i] = rp[i];
  for (i = 0; i < 9; ++i) R[i] = rR[i];
  for (i = 0; i < 6; ++i) dq[i] = rdq[i];
}

/* number of sim steps: len(np.arange(tend, step=dtsim)) */
int orc_rollout_nt(double tend, double dtsim) { return (int)ceil(tend / dtsim); }

/* MPC firing schedule of controlTest (uprightmpc2.py:118,136,145): fire[ti]=1 when
 * hlInterval is None (<0 here) or tt[ti] - thlPrev > hlInterval, thlPrev starting at 0. */
int orc_schedule(double dtsim, double hlInterval, int nt, unsigned char *fire) {
  int ti, n = 0; double thl = 0;
  for (ti = 0; ti < nt; ++ti) {
    double t = ti * dtsim;
    fire[ti] = (hlInterval < 0 || t - thl > hlInterval) ? 1 : 0;
    if (fire[ti]) { thl = t; ++n; }
  }
  return n;
}

/*
 * Closed-loop flight of ONE env.  ctrl/wl are opaque controller handles:
 *   oracle build : umpc_par_t*, umpc_state_t*, wl_t*   (wlmode 0 none, 1 WLCon, 2 WLQP)
 *   ref build    : UprightMPC_t* (par ignored), WLCon_t* (wlmode 0/1)
 * ic = p[3], R[9 col-major], dq[6].  cfg = tend, dtsim, hlInterval(<0 None), taulim,
 * tpert(<0 None), Ib[3], g, mb, int_pos (integer-dtype initialPos quirk of controlTest).  Outputs: log_y (nt x 12: p, R[:,2], dq), log_u (nt x 3),
 * log_pdes (nt x 3) may be NULL; metrics = err, eff (logMetric), n_solves, sum|u4| (wl) .
 */
void orc_rollout(void *par, void *ctrl, void *wl, int wlmode, const double *ic, const double *cfg,
                 int task, const double *tp, double *log_y, double *log_u, double *log_pdes,
                 double *final_state, double *metrics) {
  double tend = cfg[0], dtsim = cfg[1], hl = cfg[2], taulim = cfg[3], tpert = cfg[4];
  double Ibd[3] = {cfg[5], cfg[6], cfg[7]}, g = cfg[8], mb = cfg[9];
  int int_pos = cfg[10] != 0;
  int nt = orc_rollout_nt(tend, dtsim), ti, i, nsolve = 0;
  real p[3], R[9], dq[6], Ib[3], uq[3] = {0, 0, 0};
  double p_init[3], thl = 0, err = 0, eff = 0, actualT0 = -1.0, u4sum = 0;
  for (i = 0; i < 3; ++i) { p[i] = (real)ic[i]; p_init[i] = ic[i]; Ib[i] = (real)Ibd[i]; }
  for (i = 0; i < 9; ++i) R[i] = (real)ic[3 + i];
  for (i = 0; i < 6; ++i) dq[i] = (real)ic[12 + i];
  for (ti = 0; ti < nt; ++ti) {
    double t = ti * dtsim, pdes[3], dpdes[3], sdes[3];
    orc_task_q(task, t, p_init, tp, int_pos, pdes, dpdes, sdes);
    if (task == TASK_HELIX && tpert >= 0 && t > tpert) { dq[1] += 2; tpert = -1; }
    if (hl < 0 || t - thl > hl) {
      double uqd[3], acc[6], pd[3], Rd[9], dqd[6];
      for (i = 0; i < 3; ++i) pd[i] = p[i];
      for (i = 0; i < 9; ++i) Rd[i] = R[i];
      for (i = 0; i < 6; ++i) dqd[i] = dq[i];
#ifdef SIM_USE_REF
      {
        float fu[3], fa[6], fp[3], fR[9], fdq[6], fpd[3], fdp[3], fs[3];
        for (i = 0; i < 3; ++i) { fp[i] = (float)pd[i]; fpd[i] = (float)pdes[i]; fdp[i] = (float)dpdes[i]; fs[i] = (float)sdes[i]; }
        for (i = 0; i < 9; ++i) fR[i] = (float)Rd[i];
        for (i = 0; i < 6; ++i) fdq[i] = (float)dqd[i];
        umpcUpdate((UprightMPC_t *)ctrl, fu, fa, fp, fR, fdq, fpd, fdp, fs, (float)actualT0);
        for (i = 0; i < 3; ++i) uqd[i] = fu[i];
        for (i = 0; i < 6; ++i) acc[i] = fa[i];
      }
      (void)par;
#else
      orc_umpc_update((const umpc_par_t *)par, (umpc_state_t *)ctrl, uqd, acc, pd, Rd, dqd, pdes, dpdes, sdes, actualT0);
#endif
      if (wlmode) {
        /* h0 = [R'(0,0,mb g); 0], pdotdes = M0 accdes, actualT0 = w0[2]/mb */
        double h0[6], pdd[6], u4[4], w0[6];
        for (i = 0; i < 3; ++i) { h0[i] = Rd[3 * i + 2] * mb * g; h0[3 + i] = 0; pdd[i] = mb * acc[i]; pdd[3 + i] = Ibd[i] * acc[3 + i]; }
#ifdef SIM_USE_REF
        {
          float fh[6], fpd6[6], fu4[4], fw0[6];
          for (i = 0; i < 6; ++i) { fh[i] = (float)h0[i]; fpd6[i] = (float)pdd[i]; }
          wlConUpdate((WLCon_t *)wl, fu4, fw0, fh, fpd6);
          for (i = 0; i < 4; ++i) u4[i] = fu4[i];
          for (i = 0; i < 6; ++i) w0[i] = fw0[i];
        }
#else
        if (wlmode == 1) orc_wlcon_update(wl, u4, w0, h0, pdd); else orc_wlqp_update(wl, u4, w0, h0, pdd);
#endif
        actualT0 = w0[2] / mb;
        for (i = 0; i < 4; ++i) u4sum += fabs(u4[i]);
      }
      for (i = 0; i < 3; ++i) uq[i] = (real)uqd[i];
      thl = t; ++nsolve;
    }
    for (i = 1; i < 3; ++i) uq[i] = uq[i] < (real)-taulim ? (real)-taulim : (uq[i] > (real)taulim ? (real)taulim : uq[i]);
    dyn_step(p, R, dq, uq, (real)dtsim, Ib, (real)g);
    if (log_y) {
      for (i = 0; i < 3; ++i) { log_y[ti * 12 + i] = p[i]; log_y[ti * 12 + 3 + i] = R[6 + i]; }
      for (i = 0; i < 6; ++i) log_y[ti * 12 + 6 + i] = dq[i];
    }
    if (log_u) for (i = 0; i < 3; ++i) log_u[ti * 3 + i] = uq[i];
    if (log_pdes) for (i = 0; i < 3; ++i) log_pdes[ti * 3 + i] = pdes[i];
    err += (double)p[0] * p[0] + (double)p[1] * p[1] + (double)p[2] * p[2];
    eff += (double)uq[1] * uq[1] + (double)uq[2] * uq[2];
  }
  if (final_state) {
    for (i = 0; i < 3; ++i) final_state[i] = p[i];
    for (i = 0; i < 9; ++i) final_state[3 + i] = R[i];
    for (i = 0; i < 6; ++i) final_state[12 + i] = dq[i];
  }
  metrics[0] = err / nt; metrics[1] = eff / nt; metrics[2] = nsolve; metrics[3] = u4sum;
}

/* ---- batched CPU driver for bench.py's cpu_baseline / --impl reference legs -----------------------
 * nenv environments advanced nsteps plant steps each, hover task, MPC every step (hlInterval=None),
 * one core.  ref build: the reference's umpcUpdate with its process-global OSQP workspace, so each
 * env's solver state is swapped in/out with ref_ctx_load/save (oracle/ref_glue.c) around the call.
 * oracle build: one umpc_state_t per env.  ics: nenv x 18 (p, R col-major, dq), updated in place.
 * Returns the number of env-steps done. */
#ifdef SIM_USE_REF
int ref_ctx_floats(void);
void ref_ctx_save(float *buf);
void ref_ctx_load(const float *buf);
#include <stdlib.h>
long orc_batch_steps(int nenv, int nsteps, double *ics, const double *args11, const double *Ib, int maxIter,
                     double dtsim, double taulim) {
  int nf = ref_ctx_floats(), e, k, i;
  float *ctx = (float *)malloc((size_t)nf * nenv * sizeof(float));
  UprightMPC_t *ups = (UprightMPC_t *)malloc(sizeof(UprightMPC_t) * nenv);
  float fIb[3] = {(float)Ib[0], (float)Ib[1], (float)Ib[2]};
  long done = 0;
  for (e = 0; e < nenv; ++e) {
    umpcInit(&ups[e], (float)args11[0], (float)args11[1], (float)args11[2], (float)args11[3], (float)args11[4],
             (float)args11[5], (float)args11[6], (float)args11[7], (float)args11[8], (float)args11[9],
             (float)args11[10], fIb, maxIter);
    ref_ctx_save(ctx + (size_t)e * nf); /* all envs start from the generated workspace + warm start 0 */
  }
  for (k = 0; k < nsteps; ++k)
    for (e = 0; e < nenv; ++e) {
      double *s = ics + (size_t)e * 18;
      float fu[3], fa[6], fp[3], fR[9], fdq[6], fpd[3] = {0, 0, 0}, fdp[3] = {0, 0, 0}, fs[3] = {0, 0, 1};
      real p[3], R[9], dq[6], uq[3], rIb[3];
      for (i = 0; i < 3; ++i) { fp[i] = (float)s[i]; p[i] = s[i]; rIb[i] = Ib[i]; }
      for (i = 0; i < 9; ++i) { fR[i] = (float)s[3 + i]; R[i] = s[3 + i]; }
      for (i = 0; i < 6; ++i) { fdq[i] = (float)s[12 + i]; dq[i] = s[12 + i]; }
      ref_ctx_load(ctx + (size_t)e * nf);
      umpcUpdate(&ups[e], fu, fa, fp, fR, fdq, fpd, fdp, fs, -1.0f);
      ref_ctx_save(ctx + (size_t)e * nf);
      for (i = 0; i < 3; ++i) uq[i] = fu[i];
      for (i = 1; i < 3; ++i) uq[i] = uq[i] < -taulim ? -taulim : (uq[i] > taulim ? taulim : uq[i]);
      dyn_step(p, R, dq, uq, dtsim, rIb, args11[1]);
      for (i = 0; i < 3; ++i) s[i] = p[i];
      for (i = 0; i < 9; ++i) s[3 + i] = R[i];
      for (i = 0; i < 6; ++i) s[12 + i] = dq[i];
      ++done;
    }
  free(ctx); free(ups);
  return done;
}
#else
#include <stdlib.h>
void orc_umpc_init(umpc_par_t *par, umpc_state_t *st, const double *a11, const double *Ib, int maxIter);
long orc_batch_steps(int nenv, int nsteps, double *ics, const double *args11, const double *Ib, int maxIter,
                     double dtsim, double taulim) {
  umpc_par_t par;
  umpc_state_t *sts = (umpc_state_t *)malloc(sizeof(umpc_state_t) * nenv);
  int e, k, i;
  long done = 0;
  double pd[3] = {0, 0, 0}, sd[3] = {0, 0, 1};
  for (e = 0; e < nenv; ++e) orc_umpc_init(&par, &sts[e], args11, Ib, maxIter);
  for (k = 0; k < nsteps; ++k)
    for (e = 0; e < nenv; ++e) {
      double *s = ics + (size_t)e * 18, uqd[3], acc[6];
      real p[3], R[9], dq[6], uq[3], rIb[3];
      orc_umpc_update(&par, &sts[e], uqd, acc, s, s + 3, s + 12, pd, pd, sd, -1.0);
      for (i = 0; i < 3; ++i) { p[i] = (real)s[i]; rIb[i] = (real)Ib[i]; uq[i] = (real)uqd[i]; }
      for (i = 0; i < 9; ++i) R[i] = (real)s[3 + i];
      for (i = 0; i < 6; ++i) dq[i] = (real)s[12 + i];
      for (i = 1; i < 3; ++i) uq[i] = uq[i] < (real)-taulim ? (real)-taulim : (uq[i] > (real)taulim ? (real)taulim : uq[i]);
      dyn_step(p, R, dq, uq, (real)dtsim, rIb, (real)args11[1]);
      for (i = 0; i < 3; ++i) s[i] = p[i];
      for (i = 0; i < 9; ++i) s[3 + i] = R[i];
      for (i = 0; i < 6; ++i) s[12 + i] = dq[i];
      ++done;
    }
  free(sts);
  return done;
}
#endif
