// robobee3d_b200/csrc/umpc_dev.cuh -- device-side pieces shared by the two translation units of libumpc_b200.so:
//   umpc_kernels.cu  the C path: float32, exactly maxIter ADMM iterations, scratch in tensor / shared memory
//   umpc_conv.cu     the Python path: float64, iterate until the residual test passes, rho adapted on the way
// Both instantiate the SAME kernel templates below with a SOLVER policy that owns the per-QP scratch and runs one
// controller update for environment e; everything around the solve (inputs, task generator, WL stage, plant step,
// metric sums) is common float32 code.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "sim_core.cuh"
#include "wl_core.cuh"

// ---- global SoA state -------------------------------------------------------------------------
struct DevState {
  int B;      // row stride = number of env slots, padded to a multiple of 256 (whole CTAs only)
  int nreal;  // environments the caller asked for
  // plant
  float *p, *R, *dq, *uq, *acc;
  // controller
  float *T0, *xs, *zs, *ys, *D, *E, *solx, *soly, *info, *qpdata;
  // tasks / loop
  int* task;
  float *tp, *pinit, *actualT0, *metrics;
  // WL stage
  float *wl_u0, *wlqp;  // wlqp: 25 floats per env (WlqpState)
  // converged mode only: controller state and per-QP scratch in double, rows of B (layout: umpc_conv.cu)
  double* c64;
  // optional per-environment cost weights, 8 rows of B: ws, wds, wpr, wpf, wvr, wvf, wthrust, wmom (the arguments the
  // reference's weight sweep varies per run, uprightmpc2.py:272-303); nullptr: the batch-wide parameters
  float* wts;
};

// fills the weight vectors of a parameter struct (umpcInit, uprightmpc2.c:19-63) from the per-env rows
template <class P>
__device__ __forceinline__ void load_env_weights(P& par, const DevState& s, int e) {
  const size_t B = s.B;
  const float ws = s.wts[0 * B + e], wds = s.wts[1 * B + e], wpr = s.wts[2 * B + e], wpf = s.wts[3 * B + e];
  const float wvr = s.wts[4 * B + e], wvf = s.wts[5 * B + e], wth = s.wts[6 * B + e], wmo = s.wts[7 * B + e];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    par.Qyr[i] = wpr; par.Qyf[i] = wpf; par.Qyr[3 + i] = ws; par.Qyf[3 + i] = ws;
    par.Qdyr[i] = wvr; par.Qdyf[i] = wvf; par.Qdyr[3 + i] = wds; par.Qdyf[3 + i] = wds;
  }
  par.R[0] = wth; par.R[1] = wmo; par.R[2] = wmo;
}

template <class MPC>
struct KernelParamsT {
  MPC mpc;
  SimParams sim;
};

struct LogBuf {
  int env, cap;
  float *y, *u;  // [cap][12], [cap][3]
};

__device__ __forceinline__ void store_info(const DevState& s, int e, int iter, int status, float pri_res, float dua_res) {
  s.info[(size_t)0 * s.B + e] = (float)iter;
  s.info[(size_t)1 * s.B + e] = (float)status;
  s.info[(size_t)2 * s.B + e] = pri_res;
  s.info[(size_t)3 * s.B + e] = dua_res;
}

__device__ __forceinline__ void load_wlqp(const DevState& s, int e, WlqpState& w) {
  const float* b = s.wlqp + e;
  const size_t B = s.B;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    w.xs[i] = b[(0 + i) * B]; w.zs[i] = b[(4 + i) * B]; w.ys[i] = b[(8 + i) * B];
    w.D[i] = b[(12 + i) * B]; w.E[i] = b[(16 + i) * B]; w.q0[i] = b[(21 + i) * B];
  }
  w.c = b[20 * B];
}
__device__ __forceinline__ void store_wlqp(const DevState& s, int e, const WlqpState& w) {
  float* b = s.wlqp + e;
  const size_t B = s.B;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    b[(0 + i) * B] = w.xs[i]; b[(4 + i) * B] = w.zs[i]; b[(8 + i) * B] = w.ys[i];
    b[(12 + i) * B] = w.D[i]; b[(16 + i) * B] = w.E[i]; b[(21 + i) * B] = w.q0[i];
  }
  b[20 * B] = w.c;
}

// WL stage after an MPC update (accController, robobee_test_controllers.py:165-171)
static __device__ __noinline__ void wl_stage(const SimParams& sim, const WlParams& wlp, const DevState& s, int e,
                                             const float R[9], const float acc[6]) {
  float h0[6], pdd[6], u0[4], w0[6];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    h0[i] = R[3 * i + 2] * sim.mb * sim.g; h0[3 + i] = 0.0f;
    pdd[i] = sim.mb * acc[i]; pdd[3 + i] = sim.Ib[i] * acc[3 + i];
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) u0[i] = s.wl_u0[(size_t)i * s.B + e];
  if (sim.wl_mode == 1) {
    wl_con_update(wlp, u0, w0, h0, pdd);
  } else {
    WlqpState w;
    load_wlqp(s, e, w);
    wlqp_update(wlp, w, u0, w0, h0, pdd, nullptr, nullptr);
    store_wlqp(s, e, w);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) s.wl_u0[(size_t)i * s.B + e] = u0[i];
  s.actualT0[e] = w0[2] / sim.mb;
}

// A SOLVER policy provides
//   typedef ... Params;                       controller parameters in its own scalar type
//   static constexpr int kMinBlocks;          second argument of __launch_bounds__
//   static uint32_t acquire<TPB>(uint32_t* slot), release<TPB>(const uint32_t* slot)
//                                             per-CTA scratch set-up / tear-down; every thread of the CTA calls them
//   void init(uint32_t scratch, const DevState&, int e)
//   void qpdata(const Params&, const DevState&, int e, p, R, dq, pdes, dpdes, sdes, actualT0)   (diagnostics)
//   void update(const Params&, const DevState&, int e, uquad[3], accdes[6], p, R, dq, pdes, dpdes, sdes, actualT0)
//                                             one umpcUpdate (uprightmpc2.c:209-271) of env e, INCLUDING the T0 and
//                                             solve-info rows of DevState; the inputs are dead afterwards (the
//                                             caller re-reads what it still needs from memory)

// ---- kernel 1: one umpcUpdate per environment, inputs in [B][k] (the C-ABI's layout) ---------
template <int TPB, class SOLVER>
__global__ void __launch_bounds__(TPB, SOLVER::kMinBlocks)
umpc_update_kernel(const __grid_constant__ KernelParamsT<typename SOLVER::Params> kp, const __grid_constant__ DevState s, float* __restrict__ uquad,
                   float* __restrict__ accdes, const float* __restrict__ p0, const float* __restrict__ R0,
                   const float* __restrict__ dq0, const float* __restrict__ pdes, const float* __restrict__ dpdes,
                   const float* __restrict__ sdes, const float* __restrict__ actualT0, int want_qpdata,
                   float* __restrict__ p_next, float* __restrict__ R_next, float* __restrict__ dq_next, int env0) {
  const int e = env0 + blockIdx.x * blockDim.x + threadIdx.x;  // env0: first env of this launch (chunked host-buffer calls)
  __shared__ uint32_t tmem_slot;
  const uint32_t scratch_base = SOLVER::template acquire<TPB>(&tmem_slot);  // every warp of the CTA takes part
  if (e < s.B) {  // B is a multiple of 256: whole CTAs (the body contains CTA-wide barriers)
  float p[3], R[9], dq[6], pd[3], dpd[3], sd[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    p[i] = p0[(size_t)e * 3 + i]; pd[i] = pdes[(size_t)e * 3 + i];
    dpd[i] = dpdes[(size_t)e * 3 + i]; sd[i] = sdes[(size_t)e * 3 + i];
  }
#pragma unroll
  for (int i = 0; i < 9; ++i) R[i] = R0[(size_t)e * 9 + i];
#pragma unroll
  for (int i = 0; i < 6; ++i) dq[i] = dq0[(size_t)e * 6 + i];
  const float aT0 = actualT0 ? actualT0[e] : -1.0f;
  SOLVER sol;
  sol.init(scratch_base, s, e);
  if (want_qpdata) sol.qpdata(kp.mpc, s, e, p, R, dq, pd, dpd, sd, aT0);
  float uq[3], acc[6];
  sol.update(kp.mpc, s, e, uq, acc, p, R, dq, pd, dpd, sd, aT0);
#pragma unroll
  for (int i = 0; i < 3; ++i) { uquad[(size_t)e * 3 + i] = uq[i]; s.uq[(size_t)i * s.B + e] = uq[i]; }
#pragma unroll
  for (int i = 0; i < 6; ++i) { accdes[(size_t)e * 6 + i] = acc[i]; s.acc[(size_t)i * s.B + e] = acc[i]; }
  if (p_next) {
    // the two lines of controlTest that follow mdl.update (uprightmpc2.py:148-151): clip, plant step
    // (inputs re-read: they were not kept in registers over the solve)
#pragma unroll
    for (int i = 0; i < 3; ++i) p[i] = *(volatile const float*)(p0 + (size_t)e * 3 + i);
#pragma unroll
    for (int i = 0; i < 9; ++i) R[i] = *(volatile const float*)(R0 + (size_t)e * 9 + i);
#pragma unroll
    for (int i = 0; i < 6; ++i) dq[i] = *(volatile const float*)(dq0 + (size_t)e * 6 + i);
    uq[1] = sim_clip(uq[1], -kp.sim.taulim, kp.sim.taulim);
    uq[2] = sim_clip(uq[2], -kp.sim.taulim, kp.sim.taulim);
    sim_dyn_step(p, R, dq, uq, kp.sim.dtsim, kp.sim.Ib, kp.sim.g);
#pragma unroll
    for (int i = 0; i < 3; ++i) p_next[(size_t)e * 3 + i] = p[i];
#pragma unroll
    for (int i = 0; i < 9; ++i) R_next[(size_t)e * 9 + i] = R[i];
#pragma unroll
    for (int i = 0; i < 6; ++i) dq_next[(size_t)e * 6 + i] = dq[i];
  }
  }
  SOLVER::template release<TPB>(&tmem_slot);
}

// ---- kernel 2: fused closed loop (uprightmpc2.py:87-159) -------------------------------------
// Per plant step: task -> [umpcUpdate when the host-computed schedule fires] -> [WL stage] ->
// torque clip -> rigid-body step -> metric sums.  Plant state stays in registers for the
// whole launch; the warm start is re-read from HBM at each solve (1 KB per ~5e4 flop).
template <int TPB, class SOLVER>
__global__ void __launch_bounds__(TPB, SOLVER::kMinBlocks)
umpc_rollout_kernel(const __grid_constant__ KernelParamsT<typename SOLVER::Params> kp, const __grid_constant__ DevState s,
                    const __grid_constant__ WlParams wlp, int step0, int nsteps, const uint8_t* __restrict__ fire,
                    const double* __restrict__ tvals, LogBuf lg) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  __shared__ uint32_t tmem_slot;
  const uint32_t scratch_base = SOLVER::template acquire<TPB>(&tmem_slot);
  if (e < s.B) {
  const size_t B = s.B;
  float p[3], R[9], dq[6], uq[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) { p[i] = s.p[i * B + e]; uq[i] = s.uq[i * B + e]; }
#pragma unroll
  for (int i = 0; i < 9; ++i) R[i] = s.R[i * B + e];
#pragma unroll
  for (int i = 0; i < 6; ++i) dq[i] = s.dq[i * B + e];
  const int task = s.task[e];
  float err = s.metrics[0 * B + e], eff = s.metrics[1 * B + e], nsolve = s.metrics[2 * B + e], nstep = s.metrics[3 * B + e];
  float pert_done = s.metrics[4 * B + e];
  SOLVER sol;
  sol.init(scratch_base, s, e);
  for (int k = 0; k < nsteps; ++k) {
    const double t = tvals[k];
    if (task == SIM_TASK_HELIX && kp.sim.tpert >= 0.0f && t > (double)kp.sim.tpert && pert_done == 0.0f) {
      dq[1] += 2.0f; pert_done = 1.0f;
    }
    if (fire[k]) {
      // the task's reference is consumed by the controller only: evaluate it when the MPC fires
      float pdes[3], dpdes[3], sdes[3];
      {
        float pinit[3], tp[4];
#pragma unroll
        for (int i = 0; i < 3; ++i) pinit[i] = s.pinit[i * B + e];
#pragma unroll
        for (int i = 0; i < 4; ++i) tp[i] = s.tp[i * B + e];
        sim_task(task, t, pinit, tp, kp.sim.int_pos, pdes, dpdes, sdes);
      }
      float acc[6];
      const float aT0 = kp.sim.wl_mode ? s.actualT0[e] : -1.0f;
      // plant state and metric sums through memory across the solve: they are not touched by it, and the
      // ADMM loop needs every register
#pragma unroll
      for (int i = 0; i < 3; ++i) s.p[i * B + e] = p[i];
#pragma unroll
      for (int i = 0; i < 9; ++i) s.R[i * B + e] = R[i];
#pragma unroll
      for (int i = 0; i < 6; ++i) s.dq[i * B + e] = dq[i];
      s.metrics[0 * B + e] = err; s.metrics[1 * B + e] = eff; s.metrics[2 * B + e] = nsolve; s.metrics[3 * B + e] = nstep;
      sol.update(kp.mpc, s, e, uq, acc, p, R, dq, pdes, dpdes, sdes, aT0);
#pragma unroll
      for (int i = 0; i < 6; ++i) s.acc[i * B + e] = acc[i];
#pragma unroll
      for (int i = 0; i < 3; ++i) p[i] = *(volatile float*)(s.p + i * B + e);
#pragma unroll
      for (int i = 0; i < 9; ++i) R[i] = *(volatile float*)(s.R + i * B + e);
#pragma unroll
      for (int i = 0; i < 6; ++i) dq[i] = *(volatile float*)(s.dq + i * B + e);
      err = *(volatile float*)(s.metrics + 0 * B + e); eff = *(volatile float*)(s.metrics + 1 * B + e);
      nsolve = *(volatile float*)(s.metrics + 2 * B + e); nstep = *(volatile float*)(s.metrics + 3 * B + e);
      if (kp.sim.wl_mode) wl_stage(kp.sim, wlp, s, e, R, acc);
      nsolve += 1.0f;
    }
    uq[1] = sim_clip(uq[1], -kp.sim.taulim, kp.sim.taulim);
    uq[2] = sim_clip(uq[2], -kp.sim.taulim, kp.sim.taulim);
    sim_dyn_step(p, R, dq, uq, kp.sim.dtsim, kp.sim.Ib, kp.sim.g);
    err += p[0] * p[0] + p[1] * p[1] + p[2] * p[2];
    eff += uq[1] * uq[1] + uq[2] * uq[2];
    nstep += 1.0f;
    if (e == lg.env && step0 + k < lg.cap) {
      float* y = lg.y + (size_t)(step0 + k) * 12;
      float* u = lg.u + (size_t)(step0 + k) * 3;
#pragma unroll
      for (int i = 0; i < 3; ++i) { y[i] = p[i]; y[3 + i] = R[6 + i]; u[i] = uq[i]; }
#pragma unroll
      for (int i = 0; i < 6; ++i) y[6 + i] = dq[i];
    }
  }
#pragma unroll
  for (int i = 0; i < 3; ++i) { s.p[i * B + e] = p[i]; s.uq[i * B + e] = uq[i]; }
#pragma unroll
  for (int i = 0; i < 9; ++i) s.R[i * B + e] = R[i];
#pragma unroll
  for (int i = 0; i < 6; ++i) s.dq[i * B + e] = dq[i];
  s.metrics[0 * B + e] = err; s.metrics[1 * B + e] = eff; s.metrics[2 * B + e] = nsolve; s.metrics[3 * B + e] = nstep;
  s.metrics[4 * B + e] = pert_done;
  }
  SOLVER::template release<TPB>(&tmem_slot);
}
