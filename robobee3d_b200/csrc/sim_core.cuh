// robobee3d_b200/csrc/sim_core.cuh
//
// Plant step, reference trajectories and loop bookkeeping of the closed-loop flight,
// one environment per thread.  Reference (under template/):
//   quadrotorNLVF / quadrotorNLDyn   genqp.py:24-41   (explicit Euler, R+ = R expm(skew(w) dt))
//   helix / straightAcc / flip / perch   flight_tasks.py:6-49
//   controlTest loop, torque clip, perturbation, logMetric   uprightmpc2.py:87-175
//   MPC -> WL glue (h0, M0 accdes, actualT0)   robobee_test_controllers.py:162-171, ca6dynamics.py:40-50
#pragma once
#include <math.h>

#ifndef UMPC_HD
#if defined(__CUDACC__)
#define UMPC_HD __host__ __device__ __forceinline__
#else
#define UMPC_HD inline
#endif
#endif

enum { SIM_TASK_HELIX = 0, SIM_TASK_STRAIGHT = 1, SIM_TASK_FLIP = 2, SIM_TASK_PERCH = 3 };

struct SimParams {
  float dtsim;       // plant step (ms); controlTest default 0.2
  float hl_interval; // < 0: MPC every step (hlInterval=None); else fire when t - thlPrev > hl_interval
  float taulim;      // torque clip (uprightmpc2.py:148-149), default 100
  float tpert;       // < 0: none; helix task only: dq[1] += 2 once t > tpert (uprightmpc2.py:131-133)
  float Ib[3];       // body inertia (genqp.py:22)
  float g;           // 9.81e-3 mm/ms^2
  float mb;          // body mass for the WL glue (ca6dynamics.py:5-10)
  int wl_mode;       // 0: none, 1: WLCon projected gradient, 2: true WLQP
  int int_pos;       // 1: initialPos has integer dtype as in controlTest (see sim_task)
};

UMPC_HD float sim_clip(float v, float lo, float hi) { return v < lo ? lo : (v > hi ? hi : v); }

// (pdes, dpdes, sdes) of the task at time t (ms).  tp[4]:
//   helix (trajAmp, trajFreq, dz, useY) | straightAcc (tduration, vdes) |
//   flip (tstart, tend) | perch (tend, trotstart, trotend, vdes)   -- flight_tasks.py argument order
// Evaluated in double like numpy does (it runs once per MPC update, next to ~9e4 float instructions).
// int_pos reproduces controlTest's integer-dtype initialPos (uprightmpc2.py:94-104: p = np.array([0, 0, 0]) is an
// int64 array, flight_tasks copies it and `posdes[i] += x` then stores the float sum back into the int array,
// i.e. truncates it toward zero); dpdes and sdes are float arrays in the reference and are not affected.
UMPC_HD double sim_trunc_if(double v, int int_pos) { return int_pos ? trunc(v) : v; }

UMPC_HD void sim_task(int task, double t, const float p_init[3], const float tp[4], int int_pos, float pdes[3],
                      float dpdes[3], float sdes[3]) {
  const double PI_D = 3.14159265358979323846;
  double pd[3], dpd[3] = {0.0, 0.0, 0.0}, sd[3] = {0.0, 0.0, 1.0};
#pragma unroll
  for (int i = 0; i < 3; ++i) pd[i] = (double)p_init[i];
  if (task == SIM_TASK_HELIX) {
    const double amp = (double)tp[0], omg = 2.0 * PI_D * (double)tp[1] * 1e-3;
    const double sn = sin(omg * t), cs = cos(omg * t);
    pd[0] = sim_trunc_if(pd[0] + amp * sn, int_pos); dpd[0] = amp * omg * cs;
    if (tp[3] != 0.0f) { pd[1] = sim_trunc_if(pd[1] + amp * (1.0 - cs), int_pos); dpd[1] = amp * omg * sn; }
    if (amp > 1e-3) { pd[2] = sim_trunc_if(pd[2] + (double)tp[2] * t, int_pos); dpd[2] = (double)tp[2]; }
  } else if (task == SIM_TASK_STRAIGHT) {
    const double td = (double)tp[0], vd = (double)tp[1];
    dpd[0] = t < td ? vd : 0.0;
    pd[0] = sim_trunc_if(pd[0] + vd * (t < 0.0 ? 0.0 : (t > td ? td : t)), int_pos);
  } else if (task == SIM_TASK_FLIP) {
    double ph = (t - (double)tp[0]) / (double)tp[1];
    ph = ph < 0.0 ? 0.0 : (ph > 1.0 ? 1.0 : ph);
    sd[0] = -sin(ph * 2.0 * PI_D); sd[1] = 0.0; sd[2] = cos(ph * 2.0 * PI_D);
  } else if (task == SIM_TASK_PERCH) {
    const double te = (double)tp[0], vd = (double)tp[3];
    pd[0] = sim_trunc_if(pd[0] + vd * (t < 0.0 ? 0.0 : (t > te ? te : t)), int_pos);
    dpd[0] = t < te ? vd : 0.0;
    if (t < te) {
      double ph = (t - (double)tp[2]) / (double)tp[1];
      ph = ph < 0.0 ? 0.0 : (ph > 1.0 ? 1.0 : ph);
      sd[0] = -sin(ph * PI_D); sd[1] = 0.0; sd[2] = cos(ph * PI_D);
    } else {
      sd[0] = -1.0; sd[1] = 0.0; sd[2] = 0.0;
    }
  }
#pragma unroll
  for (int i = 0; i < 3; ++i) { pdes[i] = (float)pd[i]; dpdes[i] = (float)dpd[i]; sdes[i] = (float)sd[i]; }
}

// One explicit-Euler step of the rigid body (genqp.py:24-41).  R column-major,
// dq = (v_world, omega_body), u = (thrust/mass, tau_x, tau_y).  expm(skew(w) dt) in
// closed form (Rodrigues), which is what scipy's expm returns for a skew argument.
UMPC_HD void sim_dyn_step(float p[3], float R[9], float dq[6], const float u[3], float dt, const float Ib[3], float g) {
  float dv[3], dw[3], Iw[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) { dv[i] = u[0] * R[6 + i]; Iw[i] = Ib[i] * dq[3 + i]; }
  dv[2] -= g;
  dw[0] = (-(dq[4] * Iw[2] - dq[5] * Iw[1]) + u[1]) / Ib[0];
  dw[1] = (-(dq[5] * Iw[0] - dq[3] * Iw[2]) + u[2]) / Ib[1];
  dw[2] = (-(dq[3] * Iw[1] - dq[4] * Iw[0])) / Ib[2];
#pragma unroll
  for (int i = 0; i < 3; ++i) p[i] += dt * dq[i];
  const float kx = dq[3] * dt, ky = dq[4] * dt, kz = dq[5] * dt;
  const float th2 = kx * kx + ky * ky + kz * kz;
  float a, b;
  if (th2 < 1e-8f) {
    a = 1.0f - th2 * (1.0f / 6.0f); b = 0.5f - th2 * (1.0f / 24.0f);
  } else {
    const float th = sqrtf(th2);
    a = sinf(th) / th; b = (1.0f - cosf(th)) / th2;
  }
  float E[9];
  E[0] = 1.0f - b * (ky * ky + kz * kz); E[4] = 1.0f - b * (kx * kx + kz * kz); E[8] = 1.0f - b * (kx * kx + ky * ky);
  E[1] = a * kz + b * kx * ky;  E[3] = -a * kz + b * kx * ky;
  E[2] = -a * ky + b * kx * kz; E[6] = a * ky + b * kx * kz;
  E[5] = a * kx + b * ky * kz;  E[7] = -a * kx + b * ky * kz;
  float Rn[9];
#pragma unroll
  for (int j = 0; j < 3; ++j)
#pragma unroll
    for (int i = 0; i < 3; ++i) Rn[i + 3 * j] = R[i] * E[3 * j] + R[i + 3] * E[1 + 3 * j] + R[i + 6] * E[2 + 3 * j];
#pragma unroll
  for (int i = 0; i < 9; ++i) R[i] = Rn[i];
#pragma unroll
  for (int i = 0; i < 3; ++i) { dq[i] += dt * dv[i]; dq[3 + i] += dt * dw[i]; }
}
