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
};

UMPC_HD float sim_clip(float v, float lo, float hi) { return v < lo ? lo : (v > hi ? hi : v); }

// (pdes, dpdes, sdes) of the task at time t (ms).  tp[4]:
//   helix (trajAmp, trajFreq, dz, useY) | straightAcc (tduration, vdes) |
//   flip (tstart, tend) | perch (tend, trotstart, trotend, vdes)   -- flight_tasks.py argument order
UMPC_HD void sim_task(int task, float t, const float p_init[3], const float tp[4], float pdes[3], float dpdes[3],
                      float sdes[3]) {
  const float PI_F = 3.14159265358979323846f;
#pragma unroll
  for (int i = 0; i < 3; ++i) { pdes[i] = p_init[i]; dpdes[i] = 0.0f; }
  sdes[0] = 0.0f; sdes[1] = 0.0f; sdes[2] = 1.0f;
  if (task == SIM_TASK_HELIX) {
    const float amp = tp[0], omg = 2.0f * PI_F * tp[1] * 1e-3f;
    float sn, cs;
#if defined(__CUDA_ARCH__)
    sincosf(omg * t, &sn, &cs);
#else
    sn = sinf(omg * t); cs = cosf(omg * t);
#endif
    pdes[0] += amp * sn; dpdes[0] = amp * omg * cs;
    if (tp[3] != 0.0f) { pdes[1] += amp * (1.0f - cs); dpdes[1] = amp * omg * sn; }
    if (amp > 1e-3f) { pdes[2] += tp[2] * t; dpdes[2] = tp[2]; }
  } else if (task == SIM_TASK_STRAIGHT) {
    dpdes[0] = t < tp[0] ? tp[1] : 0.0f;
    pdes[0] += tp[1] * sim_clip(t, 0.0f, tp[0]);
  } else if (task == SIM_TASK_FLIP) {
    const float ph = sim_clip((t - tp[0]) / tp[1], 0.0f, 1.0f) * 2.0f * PI_F;
    sdes[0] = -sinf(ph); sdes[1] = 0.0f; sdes[2] = cosf(ph);
  } else if (task == SIM_TASK_PERCH) {
    pdes[0] += tp[3] * sim_clip(t, 0.0f, tp[0]);
    dpdes[0] = t < tp[0] ? tp[3] : 0.0f;
    if (t < tp[0]) {
      const float ph = sim_clip((t - tp[2]) / tp[1], 0.0f, 1.0f) * PI_F;
      sdes[0] = -sinf(ph); sdes[1] = 0.0f; sdes[2] = cosf(ph);
    } else {
      sdes[0] = -1.0f; sdes[1] = 0.0f; sdes[2] = 0.0f;
    }
  }
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
