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

// ---- global state: struct of arrays, tiled by warp ------------------------------------------------
// One slab float[B/32][UMPC_ROWS][32]: the 32 envs of a warp own one contiguous tile (46 KB) in which row r of
// lane l sits at r*32 + l.  A warp's access to one component is one 128-byte line as in a plain [row][env] layout,
// but every row of an env is at a COMPILE-TIME offset from ONE per-thread base pointer (env_base), so a load/store is
// a single instruction with an immediate offset.  With a run-time row stride every access cost a 64-bit add-with-carry
// pair plus a live address register pair: ~2,300 of the ~10,000 once-per-solve SASS instructions and a good part of
// the spills (profiles/r01c_phase_and_occupancy_sweeps.txt).
enum {
  ROW_P = 0, ROW_R = ROW_P + 3, ROW_DQ = ROW_R + 9, ROW_UQ = ROW_DQ + 6, ROW_ACC = ROW_UQ + 3,          // plant
  ROW_T0 = ROW_ACC + 6, ROW_XS = ROW_T0 + 1, ROW_ZS = ROW_XS + 45, ROW_YS = ROW_ZS + 39,                // controller:
  ROW_D = ROW_YS + 39, ROW_E = ROW_D + 45, ROW_SOLX = ROW_E + 39, ROW_SOLY = ROW_SOLX + 45,             //  warm start, scaling,
  ROW_INFO = ROW_SOLY + 39,                                                                              //  solution, info
  ROW_TASK = ROW_INFO + 4, ROW_TP = ROW_TASK + 1, ROW_PINIT = ROW_TP + 4, ROW_AT0 = ROW_PINIT + 3,      // task / loop
  ROW_METRICS = ROW_AT0 + 1,  // err, eff (float sums), nsolve, nstep (int32 bit patterns), pert_done, Kahan terms of err, eff
  ROW_WLU0 = ROW_METRICS + 7, ROW_WLQP = ROW_WLU0 + 4,                                                   // WL stage (25 = WlqpState)
  UMPC_ROWS = ROW_WLQP + 25
};
#define ROWF(r) ((r) * 32)  // float offset of row r inside a tile

struct DevState {
  int B;      // number of env slots, padded to a multiple of 256 (whole CTAs only)
  int nreal;  // environments the caller asked for
  float* slab;
  // HOST-side handles = address of the row block of ENV 0 (slab + ROWF(ROW_x)); env e's element i is at
  // handle[tile_off(e) + 32 i].  Device code goes through env_base() instead.
  float *p, *R, *dq, *uq, *acc;
  float *T0, *xs, *zs, *ys, *D, *E, *solx, *soly, *info;
  int* task;
  float *tp, *pinit, *actualT0, *metrics;
  float *wl_u0, *wlqp;
  float* qpdata;  // diagnostics, plain [216][B]; allocated on request
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

__host__ __device__ __forceinline__ size_t tile_off(int e) { return (size_t)(e >> 5) * (UMPC_ROWS * 32) + (size_t)(e & 31); }
__device__ __forceinline__ float* env_base(const DevState& s, int e) { return s.slab + tile_off(e); }

template <class MPC>
struct KernelParamsT {
  MPC mpc;
  SimParams sim;
};

#define UMPC_MAX_CHUNKED_TILES 512   // tiles per CTA up to which a launch's steps are cut into chunks (umpc_rollout_kernel)

struct LogBuf {
  int env, cap;
  float *y, *u;  // [cap][12], [cap][3]
};

__device__ __forceinline__ void store_info(float* eb, int iter, int status, float pri_res, float dua_res) {
  eb[ROWF(ROW_INFO + 0)] = (float)iter;
  eb[ROWF(ROW_INFO + 1)] = (float)status;
  eb[ROWF(ROW_INFO + 2)] = pri_res;
  eb[ROWF(ROW_INFO + 3)] = dua_res;
}

__device__ __forceinline__ void load_wlqp(const float* eb, WlqpState& w) {
  const float* b = eb + ROWF(ROW_WLQP);
  constexpr int B = 32;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    w.xs[i] = b[(0 + i) * B]; w.zs[i] = b[(4 + i) * B]; w.ys[i] = b[(8 + i) * B];
    w.D[i] = b[(12 + i) * B]; w.E[i] = b[(16 + i) * B]; w.q0[i] = b[(21 + i) * B];
  }
  w.c = b[20 * B];
}
__device__ __forceinline__ void store_wlqp(float* eb, const WlqpState& w) {
  float* b = eb + ROWF(ROW_WLQP);
  constexpr int B = 32;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    b[(0 + i) * B] = w.xs[i]; b[(4 + i) * B] = w.zs[i]; b[(8 + i) * B] = w.ys[i];
    b[(12 + i) * B] = w.D[i]; b[(16 + i) * B] = w.E[i]; b[(21 + i) * B] = w.q0[i];
  }
  b[20 * B] = w.c;
}

// WL stage after an MPC update (accController, robobee_test_controllers.py:165-171)
static __device__ __noinline__ void wl_stage(const SimParams& sim, const WlParams& wlp, float* eb, const float R[9],
                                             const float acc[6]) {
  float h0[6], pdd[6], u0[4], w0[6];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    h0[i] = R[3 * i + 2] * sim.mb * sim.g; h0[3 + i] = 0.0f;
    pdd[i] = sim.mb * acc[i]; pdd[3 + i] = sim.Ib[i] * acc[3 + i];
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) u0[i] = eb[ROWF(ROW_WLU0 + i)];
  if (sim.wl_mode == 1) {
    wl_con_update(wlp, u0, w0, h0, pdd);
  } else {
    WlqpState w;
    load_wlqp(eb, w);
    wlqp_update(wlp, w, u0, w0, h0, pdd, nullptr, nullptr);
    store_wlqp(eb, w);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) eb[ROWF(ROW_WLU0 + i)] = u0[i];
  eb[ROWF(ROW_AT0)] = w0[2] / sim.mb;
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
  // No bounds BRANCH: B and every launch's env range are multiples of 256, so all threads of the grid own an env slot
  // in the slab (a thread-dependent branch around the body would also keep ptxas from using the uniform datapath
  // inside it).  The caller's [B][k] arrays, however, hold only the nreal environments it asked for: padded lanes read
  // the LAST real env's inputs (in bounds, discarded) and their output stores are predicated off.
  const bool real = e < s.nreal;
  const size_t ei = (size_t)(real ? e : s.nreal - 1);
  {
  float p[3], R[9], dq[6], pd[3], dpd[3], sd[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    p[i] = p0[ei * 3 + i]; pd[i] = pdes[ei * 3 + i];
    dpd[i] = dpdes[ei * 3 + i]; sd[i] = sdes[ei * 3 + i];
  }
#pragma unroll
  for (int i = 0; i < 9; ++i) R[i] = R0[ei * 9 + i];
#pragma unroll
  for (int i = 0; i < 6; ++i) dq[i] = dq0[ei * 6 + i];
  const float aT0 = actualT0 ? actualT0[ei] : -1.0f;
  float* const eb = env_base(s, e);
  SOLVER sol;
  sol.init(scratch_base, s, e);
  if (want_qpdata && real) sol.qpdata(kp.mpc, s, e, p, R, dq, pd, dpd, sd, aT0);
  float uq[3], acc[6];
  sol.update(kp.mpc, s, e, uq, acc, p, R, dq, pd, dpd, sd, aT0);
#pragma unroll
  for (int i = 0; i < 3; ++i) { if (real) uquad[ei * 3 + i] = uq[i]; eb[ROWF(ROW_UQ + i)] = uq[i]; }
#pragma unroll
  for (int i = 0; i < 6; ++i) { if (real) accdes[ei * 6 + i] = acc[i]; eb[ROWF(ROW_ACC + i)] = acc[i]; }
  if (p_next) {
    // the two lines of controlTest that follow mdl.update (uprightmpc2.py:148-151): clip, plant step
    // (inputs re-read: they were not kept in registers over the solve)
#pragma unroll
    for (int i = 0; i < 3; ++i) p[i] = *(volatile const float*)(p0 + ei * 3 + i);
#pragma unroll
    for (int i = 0; i < 9; ++i) R[i] = *(volatile const float*)(R0 + ei * 9 + i);
#pragma unroll
    for (int i = 0; i < 6; ++i) dq[i] = *(volatile const float*)(dq0 + ei * 6 + i);
    uq[1] = sim_clip(uq[1], -kp.sim.taulim, kp.sim.taulim);
    uq[2] = sim_clip(uq[2], -kp.sim.taulim, kp.sim.taulim);
    sim_dyn_step(p, R, dq, uq, kp.sim.dtsim, kp.sim.Ib, kp.sim.g);
    if (real) {
#pragma unroll
      for (int i = 0; i < 3; ++i) p_next[ei * 3 + i] = p[i];
#pragma unroll
      for (int i = 0; i < 9; ++i) R_next[ei * 9 + i] = R[i];
#pragma unroll
      for (int i = 0; i < 6; ++i) dq_next[ei * 6 + i] = dq[i];
    }
  }
  }
  SOLVER::template release<TPB>(&tmem_slot);
}

// ---- kernel 1b: the same env-step with its inputs and outputs in PINNED HOST memory (zero-copy) ----------------
// The host-buffer entry points (umpcBatchUpdate / umpcBatchEnvStep) hand this kernel the caller's own [B][k] arrays,
// mapped into the device address space: each CTA pulls its environments' slices over PCIe with coalesced 128-bit loads
// into shared memory, solves, and pushes its results back with coalesced 128-bit stores.  No staging copies, no chunk
// pipeline: transfer and compute overlap CTA by CTA (while one CTA waits for its inputs the others solve; result
// stores are posted writes).  in: p0[3] R0[9] dq0[6] pdes[3] dpdes[3] sdes[3]; out: uquad[3] accdes[6] and, for the
// env-step, p_next[3] R_next[9] dq_next[6].
struct HostIO {
  const float* in[6];
  const float* aT0;  // [B] or null
  float* out[5];     // out[2..4] null: update only
  int nenv;          // environments the host arrays hold (the grid covers the padded slots as well)
  unsigned* pulled;  // device counter (zero at launch): CTAs whose input pull has completed
  int window;        // CTAs allowed to pull at the same time (0: no ordering)
};

// one array's slice, phase 1: every thread issues its (at most ceil(k/4)) 128-bit loads from host memory into registers
template <int TPB, int K>
__device__ __forceinline__ void hostio_load(float4 (&v)[(K + 3) / 4], const float* __restrict__ src, int nvalid_env) {
  const int nv = nvalid_env * K;
#pragma unroll
  for (int j = 0; j < (K + 3) / 4; ++j) {
    const int i = (threadIdx.x + j * TPB) * 4;
    v[j] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i < TPB * K) {
      if (i + 3 < nv) v[j] = *reinterpret_cast<const float4*>(src + i);
      else if (i < nv) { v[j].x = src[i]; if (i + 1 < nv) v[j].y = src[i + 1]; if (i + 2 < nv) v[j].z = src[i + 2]; }
    }
  }
}
// phase 2: registers -> shared memory
template <int TPB, int K>
__device__ __forceinline__ void hostio_stash(float* __restrict__ dst, const float4 (&v)[(K + 3) / 4]) {
#pragma unroll
  for (int j = 0; j < (K + 3) / 4; ++j) {
    const int i = (threadIdx.x + j * TPB) * 4;
    if (i < TPB * K) *reinterpret_cast<float4*>(dst + i) = v[j];
  }
}
template <int TPB>
__device__ __forceinline__ void hostio_push(float* __restrict__ dst, const float* __restrict__ src, int k, int nvalid_env) {
  const int n = TPB * k, nv = nvalid_env * k;
  for (int i = threadIdx.x * 4; i < n; i += TPB * 4) {
    const float4 v = *reinterpret_cast<const float4*>(src + i);
    if (i + 3 < nv) *reinterpret_cast<float4*>(dst + i) = v;
    else if (i < nv) { dst[i] = v.x; if (i + 1 < nv) dst[i + 1] = v.y; if (i + 2 < nv) dst[i + 2] = v.z; }
  }
}

template <int TPB, class SOLVER>
__global__ void __launch_bounds__(TPB, SOLVER::kMinBlocks)
umpc_hostio_kernel(const __grid_constant__ KernelParamsT<typename SOLVER::Params> kp, const __grid_constant__ DevState s,
                   const __grid_constant__ HostIO io, int want_qpdata) {
  // slices of the CTA's TPB environments, [env][k] like the host arrays: p R dq pdes dpdes sdes | uq acc p' R' dq'
  constexpr int OFF_IN[6] = {0, 3, 12, 18, 21, 24};
  constexpr int OFF_OUT[5] = {0, 3, 9, 12, 21}, K_OUT[5] = {3, 6, 3, 9, 6};
  __shared__ __align__(16) float sin_[27 * TPB];
  __shared__ __align__(16) float sout_[27 * TPB];
  __shared__ uint32_t tmem_slot;
  const int cta0 = blockIdx.x * TPB, t = threadIdx.x, e = cta0 + t;
  const int nvalid = min(max(io.nenv - cta0, 0), TPB);
  const uint32_t scratch_base = SOLVER::template acquire<TPB>(&tmem_slot);
  // Ordered pulls: the link serves all outstanding requests round-robin, so CTAs that all pull at once all get their
  // data at the same time (the whole first wave of 296 CTAs after ~90 us, SMs idle until then, and again for the second
  // wave).  Letting only `window` CTAs pull at a time, in block order, hands the link's bandwidth to a few CTAs at a
  // time: CTA i has its inputs after ~i x 0.3 us and solves while the later ones are still being served.  Blocks are
  // dispatched in index order, so a CTA only ever waits for CTAs that are already resident (or done).
  if (io.window) {
    if (t == 0) {
      // (the order is a performance hint only: give up after ~50 ms rather than ever hang on a scheduling surprise)
      for (int spin = 0; spin < 250000 && (int)(*(volatile unsigned*)io.pulled) + io.window <= (int)blockIdx.x; ++spin) __nanosleep(200);
    }
    __syncthreads();
  }
  {  // all loads of the CTA's six slices in flight before the first is consumed: one PCIe round trip, not seven
    float4 v0[1], v1[3], v2[2], v3[1], v4[1], v5[1];
    hostio_load<TPB, 3>(v0, io.in[0] + (size_t)cta0 * 3, nvalid); hostio_load<TPB, 9>(v1, io.in[1] + (size_t)cta0 * 9, nvalid);
    hostio_load<TPB, 6>(v2, io.in[2] + (size_t)cta0 * 6, nvalid); hostio_load<TPB, 3>(v3, io.in[3] + (size_t)cta0 * 3, nvalid);
    hostio_load<TPB, 3>(v4, io.in[4] + (size_t)cta0 * 3, nvalid); hostio_load<TPB, 3>(v5, io.in[5] + (size_t)cta0 * 3, nvalid);
    hostio_stash<TPB, 3>(sin_ + OFF_IN[0] * TPB, v0); hostio_stash<TPB, 9>(sin_ + OFF_IN[1] * TPB, v1);
    hostio_stash<TPB, 6>(sin_ + OFF_IN[2] * TPB, v2); hostio_stash<TPB, 3>(sin_ + OFF_IN[3] * TPB, v3);
    hostio_stash<TPB, 3>(sin_ + OFF_IN[4] * TPB, v4); hostio_stash<TPB, 3>(sin_ + OFF_IN[5] * TPB, v5);
  }
  const float aT0 = (io.aT0 && t < nvalid) ? io.aT0[e] : -1.0f;
  __syncthreads();
  if (io.window && t == 0) atomicAdd(io.pulled, 1u);
  {
  const volatile float* sp = sin_ + OFF_IN[0] * TPB + t * 3; const volatile float* sR = sin_ + OFF_IN[1] * TPB + t * 9;
  const volatile float* sdq = sin_ + OFF_IN[2] * TPB + t * 6;
  float p[3], R[9], dq[6], pd[3], dpd[3], sd[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    p[i] = sp[i]; pd[i] = sin_[OFF_IN[3] * TPB + t * 3 + i];
    dpd[i] = sin_[OFF_IN[4] * TPB + t * 3 + i]; sd[i] = sin_[OFF_IN[5] * TPB + t * 3 + i];
  }
#pragma unroll
  for (int i = 0; i < 9; ++i) R[i] = sR[i];
#pragma unroll
  for (int i = 0; i < 6; ++i) dq[i] = sdq[i];
  float* const eb = env_base(s, e);
  SOLVER sol;
  sol.init(scratch_base, s, e);
  if (want_qpdata) sol.qpdata(kp.mpc, s, e, p, R, dq, pd, dpd, sd, aT0);
  float uq[3], acc[6];
  sol.update(kp.mpc, s, e, uq, acc, p, R, dq, pd, dpd, sd, aT0);
#pragma unroll
  for (int i = 0; i < 3; ++i) { sout_[OFF_OUT[0] * TPB + t * 3 + i] = uq[i]; eb[ROWF(ROW_UQ + i)] = uq[i]; }
#pragma unroll
  for (int i = 0; i < 6; ++i) { sout_[OFF_OUT[1] * TPB + t * 6 + i] = acc[i]; eb[ROWF(ROW_ACC + i)] = acc[i]; }
  if (io.out[2]) {
    // the two lines of controlTest that follow mdl.update (uprightmpc2.py:148-151): clip, plant step
    // (inputs re-read from shared memory: they were not kept in registers over the solve)
#pragma unroll
    for (int i = 0; i < 3; ++i) p[i] = sp[i];
#pragma unroll
    for (int i = 0; i < 9; ++i) R[i] = sR[i];
#pragma unroll
    for (int i = 0; i < 6; ++i) dq[i] = sdq[i];
    uq[1] = sim_clip(uq[1], -kp.sim.taulim, kp.sim.taulim);
    uq[2] = sim_clip(uq[2], -kp.sim.taulim, kp.sim.taulim);
    sim_dyn_step(p, R, dq, uq, kp.sim.dtsim, kp.sim.Ib, kp.sim.g);
#pragma unroll
    for (int i = 0; i < 3; ++i) sout_[OFF_OUT[2] * TPB + t * 3 + i] = p[i];
#pragma unroll
    for (int i = 0; i < 9; ++i) sout_[OFF_OUT[3] * TPB + t * 9 + i] = R[i];
#pragma unroll
    for (int i = 0; i < 6; ++i) sout_[OFF_OUT[4] * TPB + t * 6 + i] = dq[i];
  }
  }
  __syncthreads();
#pragma unroll
  for (int a = 0; a < 5; ++a)
    if (io.out[a]) hostio_push<TPB>(io.out[a] + (size_t)cta0 * K_OUT[a], sout_ + OFF_OUT[a] * TPB, K_OUT[a], nvalid);
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
                    const double* __restrict__ tvals, LogBuf lg, int nch) {
  __shared__ uint32_t tmem_slot;
  const uint32_t scratch_base = SOLVER::template acquire<TPB>(&tmem_slot);
  // Warp-tiles (32 envs, one slab tile) are dealt to warps STATICALLY: tile t belongs to CTA t % gridDim.x, and inside the
  // CTA to warp (t / gridDim.x) % (TPB / 32); a warp walks its tiles one after the other.  With one 8-warp CTA per SM
  // (the default rollout shape) every SM gets ceil or floor of ntiles / nSM tiles -- 65,536 envs = 2,048 tiles = 13 or
  // 14 per SM -- instead of whole CTAs queueing behind each other (512 CTAs on 296 slots: 1.73 waves).  A grid of
  // ntiles / (TPB / 32) CTAs degenerates to one tile per warp.
  // The steps of a launch are cut into nch chunks and the work items (tile j, chunk c), numbered c * ntile_cta + j, are
  // handed to the warps of the CTA in that order, each warp taking the next number when it is free (a shared-memory
  // counter): no warp slot idles while others work through a second tile (14 tiles on 8 slots left 2 of 8 slots idle for
  // half the launch: 17 % of all warp samples were `barrier` in profiles/r02f).  Chunk c of a tile starts when chunk c - 1,
  // done by another warp of the same CTA, has published the tile's state (global memory + a shared-memory counter).  Items
  // are taken in increasing number and an item only waits for a lower one, so the lowest unfinished item never waits: no
  // deadlock.  (A static round-robin deal was measured too: its last two warps have no slack on their predecessor, and a
  // one-step difference in chunk length cost 16 %: profiles/r02_rollout_kernel_experiments.txt.)
  constexpr int WPC = TPB / 32;
  const int ntiles = s.B >> 5, warp = threadIdx.x >> 5;
  const int ntile_cta = ntiles > (int)blockIdx.x ? (ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  __shared__ int tile_done[UMPC_MAX_CHUNKED_TILES];
  __shared__ int next_item;
  // SOLVER::kDynamicItems: hand the items out through the counter also when nch == 1 (solvers whose solve time differs
  // from tile to tile -- converged mode: 25, 50, 75 .. iterations -- would leave warp slots idle with a static deal)
  constexpr bool kDyn = SOLVER::kDynamicItems;
  if (nch > 1 || kDyn) {
    for (int i = threadIdx.x; i < ntile_cta; i += TPB) tile_done[i] = 0;
    if (threadIdx.x == 0) next_item = WPC;   // items 0 .. WPC-1 are the warps' first ones
    __syncthreads();
  }
  for (int item = warp; item < ntile_cta * nch;) {
  const int chunk = item / ntile_cta, slot = item - chunk * ntile_cta;
  // equal chunks, the remainder goes to the last one: items of different length in flight at the same time let the warps
  // drift out of phase, and warps in different phases of the solve (Ruiz loop 21 KB + ADMM loop 15 KB of SASS) do not fit
  // the 32 KB instruction cache together -- measured 0.190 ms per env-step against 0.158 in step (profiles/r02_*)
  const int clen = nsteps / nch, k0 = chunk * clen, k1 = chunk == nch - 1 ? nsteps : k0 + clen;
  if (chunk > 0) {
    while (*(volatile int*)&tile_done[slot] < chunk) __nanosleep(100);
    __syncwarp();
  }
  const int e = ((blockIdx.x + gridDim.x * slot) << 5) + (threadIdx.x & 31);
  float* const eb = env_base(s, e);
  float p[3], R[9], dq[6], uq[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) { p[i] = eb[ROWF(ROW_P + i)]; uq[i] = eb[ROWF(ROW_UQ + i)]; }
#pragma unroll
  for (int i = 0; i < 9; ++i) R[i] = eb[ROWF(ROW_R + i)];
#pragma unroll
  for (int i = 0; i < 6; ++i) dq[i] = eb[ROWF(ROW_DQ + i)];
  const int task = ((const int*)eb)[ROWF(ROW_TASK)];
  // logMetric sums (uprightmpc2.py:161-175): float32 with Kahan compensation (a 5,000-step flight adds 5,000 terms of
  // very different size), solve / step counters as integers
  float err = eb[ROWF(ROW_METRICS + 0)], eff = eb[ROWF(ROW_METRICS + 1)];
  float cerr = eb[ROWF(ROW_METRICS + 5)], ceff = eb[ROWF(ROW_METRICS + 6)];
  int nsolve = ((const int*)eb)[ROWF(ROW_METRICS + 2)], nstep = ((const int*)eb)[ROWF(ROW_METRICS + 3)];
  float pert_done = eb[ROWF(ROW_METRICS + 4)];
  SOLVER sol;
  sol.init(scratch_base, s, e);
  for (int k = k0; k < k1; ++k) {
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
        for (int i = 0; i < 3; ++i) pinit[i] = eb[ROWF(ROW_PINIT + i)];
#pragma unroll
        for (int i = 0; i < 4; ++i) tp[i] = eb[ROWF(ROW_TP + i)];
        sim_task(task, t, pinit, tp, kp.sim.int_pos, pdes, dpdes, sdes);
      }
      float acc[6];
      const float aT0 = kp.sim.wl_mode ? eb[ROWF(ROW_AT0)] : -1.0f;
      // plant state and metric sums through memory across the solve: they are not touched by it, and the
      // ADMM loop needs every register
#pragma unroll
      for (int i = 0; i < 3; ++i) eb[ROWF(ROW_P + i)] = p[i];
#pragma unroll
      for (int i = 0; i < 9; ++i) eb[ROWF(ROW_R + i)] = R[i];
#pragma unroll
      for (int i = 0; i < 6; ++i) eb[ROWF(ROW_DQ + i)] = dq[i];
      eb[ROWF(ROW_METRICS + 0)] = err; eb[ROWF(ROW_METRICS + 1)] = eff; eb[ROWF(ROW_METRICS + 5)] = cerr; eb[ROWF(ROW_METRICS + 6)] = ceff;
      ((int*)eb)[ROWF(ROW_METRICS + 2)] = nsolve; ((int*)eb)[ROWF(ROW_METRICS + 3)] = nstep;
      sol.update(kp.mpc, s, e, uq, acc, p, R, dq, pdes, dpdes, sdes, aT0);
#pragma unroll
      for (int i = 0; i < 6; ++i) eb[ROWF(ROW_ACC + i)] = acc[i];
#pragma unroll
      for (int i = 0; i < 3; ++i) p[i] = *(volatile float*)(eb + ROWF(ROW_P + i));
#pragma unroll
      for (int i = 0; i < 9; ++i) R[i] = *(volatile float*)(eb + ROWF(ROW_R + i));
#pragma unroll
      for (int i = 0; i < 6; ++i) dq[i] = *(volatile float*)(eb + ROWF(ROW_DQ + i));
      err = *(volatile float*)(eb + ROWF(ROW_METRICS + 0)); eff = *(volatile float*)(eb + ROWF(ROW_METRICS + 1));
      cerr = *(volatile float*)(eb + ROWF(ROW_METRICS + 5)); ceff = *(volatile float*)(eb + ROWF(ROW_METRICS + 6));
      nsolve = *(volatile int*)((int*)eb + ROWF(ROW_METRICS + 2)); nstep = *(volatile int*)((int*)eb + ROWF(ROW_METRICS + 3));
      if (kp.sim.wl_mode) wl_stage(kp.sim, wlp, eb, R, acc);
      nsolve += 1;
    }
    uq[1] = sim_clip(uq[1], -kp.sim.taulim, kp.sim.taulim);
    uq[2] = sim_clip(uq[2], -kp.sim.taulim, kp.sim.taulim);
    sim_dyn_step(p, R, dq, uq, kp.sim.dtsim, kp.sim.Ib, kp.sim.g);
    {
      const float ye = (p[0] * p[0] + p[1] * p[1] + p[2] * p[2]) - cerr, te = __fadd_rn(err, ye);
      cerr = __fsub_rn(__fsub_rn(te, err), ye); err = te;
      const float yf = (uq[1] * uq[1] + uq[2] * uq[2]) - ceff, tf = __fadd_rn(eff, yf);
      ceff = __fsub_rn(__fsub_rn(tf, eff), yf); eff = tf;
    }
    nstep += 1;
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
  for (int i = 0; i < 3; ++i) { eb[ROWF(ROW_P + i)] = p[i]; eb[ROWF(ROW_UQ + i)] = uq[i]; }
#pragma unroll
  for (int i = 0; i < 9; ++i) eb[ROWF(ROW_R + i)] = R[i];
#pragma unroll
  for (int i = 0; i < 6; ++i) eb[ROWF(ROW_DQ + i)] = dq[i];
  eb[ROWF(ROW_METRICS + 0)] = err; eb[ROWF(ROW_METRICS + 1)] = eff; eb[ROWF(ROW_METRICS + 5)] = cerr; eb[ROWF(ROW_METRICS + 6)] = ceff;
  ((int*)eb)[ROWF(ROW_METRICS + 2)] = nsolve; ((int*)eb)[ROWF(ROW_METRICS + 3)] = nstep;
  eb[ROWF(ROW_METRICS + 4)] = pert_done;
  if (nch > 1 || kDyn) {   // publish the tile's state to the warp that runs its next chunk
    __threadfence_block();
    __syncwarp();
    if ((threadIdx.x & 31) == 0) *(volatile int*)&tile_done[slot] = chunk + 1;
    int nxt = 0;
    if ((threadIdx.x & 31) == 0) nxt = atomicAdd(&next_item, 1);
    item = __shfl_sync(0xffffffffu, nxt, 0);
  } else {
    item += WPC;
  }
  }
  SOLVER::template release<TPB>(&tmem_slot);
}
