// robobee3d_b200/csrc/umpc_conv.cu -- CONVERGED MODE of the controller update (sm_100a): the reference's Python path.
//
// template_controllers.UprightMPC2 (template/template_controllers.py:190-258) hands the same QP to the osqp package
// in float64 with eps_abs = eps_rel = 1e-4: ADMM until the residual test passes (checked every 25 iterations,
// auxil.c:684-790), rho re-estimated and the KKT matrix refactored on the way (auxil.c:13-75).  float32 cannot reach
// that tolerance on this QP (the float32 restatement of OSQP itself runs into max_iter), so this
// translation unit instantiates the SAME device code as the C path (umpc_core.cuh + umpc_gen.cuh) with
// ureal = double and CONV = true, inside a namespace of its own so that both builds live in one library.
//
// One thread per environment as in umpc_kernels.cu.  The per-QP scratch (256 words) does not fit tensor memory in
// double for a useful number of threads, so it lives in HBM as rows [word][env] (coalesced across a warp; at
// 65,536 envs the 134 MB working set sits mostly in the 126 MB L2).  Iteration counts differ per environment:
// a warp runs until its slowest lane has converged, lanes that finished earlier commit their solution and idle.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>

#include "umpc_conv.h"

#define UMPC_REAL double
namespace umpc64 {
#include "umpc_core.cuh"
}  // namespace umpc64

// ---- rows of the double slab DevState::c64, tiled by warp like the float slab: double[B/32][C64_ROWS][32] ---------
enum {
  C64_T0 = 0,
  C64_RHO = 1,
  C64_XS = 2,
  C64_ZS = C64_XS + UMPC_NX,
  C64_YS = C64_ZS + UMPC_NC,
  C64_D = C64_YS + UMPC_NC,
  C64_E = C64_D + UMPC_NX,
  C64_STASH = C64_E + UMPC_NC,          // 28 parked inputs
  C64_VF = C64_STASH + 28,              // first-iteration fix-up vector (36)
  C64_SCR = C64_VF + UMPC_NEQ,          // factor in stream order + sigma/D^2 pairs
  C64_ROWS = C64_SCR + UMPC_GEN_NWORDS
};

#define ROWD(r) ((r) * 32)  // double offset of row r inside a tile
__host__ __device__ __forceinline__ size_t tile_off64(int e) { return (size_t)(e >> 5) * (C64_ROWS * 32) + (size_t)(e & 31); }

struct GlobStore64 {
  double* w;   // this env's scratch rows (immediate offsets from here)
  double* vf;
  __device__ __forceinline__ void init(double* c64, int e) {
    double* cb = c64 + tile_off64(e);
    w = cb + ROWD(C64_SCR); vf = cb + ROWD(C64_VF);
  }
  __device__ __forceinline__ void begin_iteration(int) {}
  __device__ __forceinline__ void issue(int c, double (&v)[UMPC_GEN_CHUNK]) const {
#pragma unroll
    for (int i = 0; i < UMPC_GEN_CHUNK; ++i) v[i] = __ldcg(w + ROWD(c * UMPC_GEN_CHUNK + i));
  }
  __device__ __forceinline__ void wait(double (&)[UMPC_GEN_CHUNK]) const {}
  __device__ __forceinline__ void waitc(int, double (&)[UMPC_GEN_CHUNK]) const {}
  __device__ __forceinline__ void setW(int e, double v) { __stcg(w + ROWD(e), v); }
  __device__ __forceinline__ void setChunk(int c, const double (&v)[UMPC_GEN_CHUNK]) {
#pragma unroll
    for (int i = 0; i < UMPC_GEN_CHUNK; ++i) __stcg(w + ROWD(c * UMPC_GEN_CHUNK + i), v[i]);
  }
  __device__ __forceinline__ void end_writes() {}
  __device__ __forceinline__ void sync() const {}
  __device__ __forceinline__ bool all(bool p) const { return __all_sync(0xffffffffu, p); }
  __device__ __forceinline__ double Vf(int i) const { return __ldcg(vf + ROWD(i)); }
  __device__ __forceinline__ void setVf(int i, double v) { __stcg(vf + ROWD(i), v); }
};

// controller state in double; the float rows of DevState are kept as mirrors for the getters of the C-ABI
struct GlobIO64 {
  float* eb;  // float mirrors: the env's rows of the tiled slab (umpc_dev.cuh)
  double* c;  // the env's rows of the double slab
  __device__ __forceinline__ GlobIO64(const DevState& s_, int e_) : eb(env_base(s_, e_)), c(s_.c64 + tile_off64(e_)) {}
  __device__ __forceinline__ void prefetch_warm_start() const {}
  __device__ __forceinline__ void stash(int k, double v) { c[ROWD(C64_STASH + k)] = v; }
  __device__ __forceinline__ double unstash(int k) const { return *(volatile const double*)(c + ROWD(C64_STASH + k)); }
  __device__ __forceinline__ double ld_rho() const { return c[ROWD(C64_RHO)]; }
  __device__ __forceinline__ void st_rho(double v) { c[ROWD(C64_RHO)] = v; }
  __device__ __forceinline__ double ld_xs(int j) const { return c[ROWD(C64_XS + j)]; }
  __device__ __forceinline__ double ld_zs(int i) const { return c[ROWD(C64_ZS + i)]; }
  __device__ __forceinline__ double ld_ys(int i) const { return c[ROWD(C64_YS + i)]; }
  __device__ __forceinline__ double ld_D(int j) const { return c[ROWD(C64_D + j)]; }
  __device__ __forceinline__ double ld_E(int i) const { return c[ROWD(C64_E + i)]; }
  __device__ __forceinline__ void st_xs(int j, double v) { c[ROWD(C64_XS + j)] = v; eb[ROWF(ROW_XS + j)] = (float)v; }
  __device__ __forceinline__ void st_zs(int i, double v) { c[ROWD(C64_ZS + i)] = v; eb[ROWF(ROW_ZS + i)] = (float)v; }
  __device__ __forceinline__ void st_ys(int i, double v) { c[ROWD(C64_YS + i)] = v; eb[ROWF(ROW_YS + i)] = (float)v; }
  __device__ __forceinline__ void st_D(int j, double v) { c[ROWD(C64_D + j)] = v; eb[ROWF(ROW_D + j)] = (float)v; }
  __device__ __forceinline__ void st_E(int i, double v) { c[ROWD(C64_E + i)] = v; eb[ROWF(ROW_E + i)] = (float)v; }
  __device__ __forceinline__ void st_solx(int j, double v) { eb[ROWF(ROW_SOLX + j)] = (float)v; }
  __device__ __forceinline__ void st_soly(int i, double v) { eb[ROWF(ROW_SOLY + i)] = (float)v; }
};

// SOLVER policy of the kernel templates in umpc_dev.cuh
struct SolverF64 {
  typedef umpc64::UmpcParams Params;
  static constexpr int kMinBlocks = 1;
  GlobStore64 st;
  template <int TPB>
  static __device__ __forceinline__ uint32_t acquire(uint32_t*) { return 0u; }
  template <int TPB>
  static __device__ __forceinline__ void release(const uint32_t*) {}
  __device__ __forceinline__ void init(uint32_t, const DevState& s, int e) { st.init(s.c64, e); }
  __device__ __forceinline__ void qpdata(const Params& par, const DevState& s, int e, const float p[3], const float R[9],
                                         const float dq[6], const float pd[3], const float dpd[3], const float sd[3], float aT0) {
    double p_[3], R_[9], dq_[6], pd_[3], dpd_[3], sd_[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) { p_[i] = p[i]; pd_[i] = pd[i]; dpd_[i] = dpd[i]; sd_[i] = sd[i]; }
#pragma unroll
    for (int i = 0; i < 9; ++i) R_[i] = R[i];
#pragma unroll
    for (int i = 0; i < 6; ++i) dq_[i] = dq[i];
    umpc64::UmpcQP qp;
    umpc64::umpc_build(par, qp, p_, R_, dq_, pd_, dpd_, sd_, aT0 >= 0.0f ? (double)aT0 : s.c64[tile_off64(e) + ROWD(C64_T0)]);
    double out[216];
    umpc64::umpc_qpdata(par, qp, out);
    for (int k = 0; k < 216; ++k) s.qpdata[(size_t)k * s.B + e] = (float)out[k];
  }
  __device__ __forceinline__ void update(const Params& par, const DevState& s, int e, float uq[3], float acc[6], const float p[3],
                                         const float R[9], const float dq[6], const float pd[3], const float dpd[3],
                                         const float sd[3], float aT0) {
    double p_[3], R_[9], dq_[6], pd_[3], dpd_[3], sd_[3], uq_[3], acc_[6];
#pragma unroll
    for (int i = 0; i < 3; ++i) { p_[i] = p[i]; pd_[i] = pd[i]; dpd_[i] = dpd[i]; sd_[i] = sd[i]; }
#pragma unroll
    for (int i = 0; i < 9; ++i) R_[i] = R[i];
#pragma unroll
    for (int i = 0; i < 6; ++i) dq_[i] = dq[i];
    GlobIO64 io(s, e);
    double T0 = io.c[ROWD(C64_T0)];
    umpc64::UmpcInfo info;
    if (s.wts) {  // per-environment weights (warp-uniform branch)
      Params pe = par;
      load_env_weights(pe, s, e);
      umpc64::umpc_update<true>(pe, st, io, uq_, acc_, p_, R_, dq_, pd_, dpd_, sd_, (double)aT0, T0, info);
    } else
    umpc64::umpc_update<true>(par, st, io, uq_, acc_, p_, R_, dq_, pd_, dpd_, sd_, (double)aT0, T0, info);
    io.c[ROWD(C64_T0)] = T0;
    io.eb[ROWF(ROW_T0)] = (float)T0;
    store_info(io.eb, info.iter, info.status, (float)info.pri_res, (float)info.dua_res);
#pragma unroll
    for (int i = 0; i < 3; ++i) uq[i] = (float)uq_[i];
#pragma unroll
    for (int i = 0; i < 6; ++i) acc[i] = (float)acc_[i];
  }
};

// ================================ host side ======================================================================
typedef KernelParamsT<umpc64::UmpcParams> KernelParams64;

struct UmpcConv {
  umpc64::UmpcParams mpc;
  double* slab = nullptr;
  int Bp = 0;
  int tpb = 64;  // measured: 64 slightly ahead of 128 (UMPC_CONV_TPB to override)
};

int umpc_conv_create(UmpcConv** out, int Bp, const UmpcBatchParams* bp, DevState* s) {
  UmpcConv* c = new UmpcConv();
  c->Bp = Bp;
  umpc64::UmpcParams& m = c->mpc;
  memset(&m, 0, sizeof(m));
  // the Python controller is built from the same numbers in double (template_controllers.py:260-280); the C-ABI
  // carries them as float, so round-trip through the shortest decimal that reproduces the float (1e-2f -> 0.01)
  auto d = [](float v) { char buf[32]; snprintf(buf, sizeof(buf), "%.7g", (double)v); return atof(buf); };
  m.dt = d(bp->dt); m.g = d(bp->g); m.Tmax = d(bp->TtoWmax) * d(bp->g);
  for (int i = 0; i < 3; ++i) {
    m.Qyr[i] = d(bp->wpr); m.Qyf[i] = d(bp->wpf); m.Qyr[3 + i] = m.Qyf[3 + i] = d(bp->ws);
    m.Qdyr[i] = d(bp->wvr); m.Qdyf[i] = d(bp->wvf); m.Qdyr[3 + i] = m.Qdyf[3 + i] = d(bp->wds);
    m.Ibi[i] = 1.0 / d(bp->Ib[i]);
  }
  m.R[0] = d(bp->wthrust); m.R[1] = m.R[2] = d(bp->wmom);
  // osqp defaults (constants.h:55-95) with the Python path's eps (template_controllers.py:191)
  m.max_iter = bp->maxIter; m.rho = 0.1; m.sigma = 1e-6; m.alpha = 1.6;
  m.eps_abs = bp->epsAbs > 0.0f ? d(bp->epsAbs) : 1e-4; m.eps_rel = bp->epsRel > 0.0f ? d(bp->epsRel) : 1e-4;
  m.scaling_iters = 10;
  m.check_termination = bp->checkTermination; m.adaptive_rho_interval = bp->adaptiveRhoInterval;
  m.adaptive_rho_tol = 5.0;
  cudaError_t e = cudaMalloc(&c->slab, (size_t)C64_ROWS * Bp * sizeof(double));
  if (e != cudaSuccess) { delete c; return (int)e; }
  if (const char* t = getenv("UMPC_CONV_TPB")) {  // tuning knob
    const int v = atoi(t);
    if (v == 64 || v == 128) c->tpb = v;
  }
  s->c64 = c->slab;
  *out = c;
  return 0;
}

__global__ void fill_rows64_kernel(double* __restrict__ slab, int nslots, int row0, int k, double v) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= nslots) return;
  double* cb = slab + tile_off64(e);
  for (int i = 0; i < k; ++i) cb[ROWD(row0 + i)] = v;
}

int umpc_conv_reset(UmpcConv* c, const DevState* s, cudaStream_t stream) {
  const size_t P = c->Bp;
  cudaError_t e = cudaMemsetAsync(c->slab, 0, (size_t)C64_ROWS * P * sizeof(double), stream);
  if (e != cudaSuccess) return (int)e;
  const int grid = (c->Bp + 255) / 256;
  fill_rows64_kernel<<<grid, 256, 0, stream>>>(c->slab, c->Bp, C64_D, UMPC_NX + UMPC_NC, 1.0);  // D and E are adjacent
  fill_rows64_kernel<<<grid, 256, 0, stream>>>(c->slab, c->Bp, C64_RHO, 1, c->mpc.rho);
  if ((e = cudaGetLastError()) != cudaSuccess) return (int)e;
  (void)s;
  return (int)cudaStreamSynchronize(stream);
}

void umpc_conv_destroy(UmpcConv* c) {
  if (!c) return;
  cudaFree(c->slab);
  delete c;
}

int umpc_conv_launch_update(UmpcConv* c, const SimParams* sim, const DevState* s, float* uquad, float* accdes, const float* p0,
                            const float* R0, const float* dq0, const float* pdes, const float* dpdes, const float* sdes,
                            const float* aT0, int want_qpdata, float* pn, float* Rn, float* dqn, int env0, int nenv,
                            cudaStream_t stream) {
  KernelParams64 kp;
  kp.mpc = c->mpc; kp.sim = *sim;
  const int grid = (nenv + c->tpb - 1) / c->tpb;
#define UMPC_CONV_UPDATE(T)                                                                                              \
  case T:                                                                                                                \
    umpc_update_kernel<T, SolverF64><<<grid, T, 0, stream>>>(kp, *s, uquad, accdes, p0, R0, dq0, pdes, dpdes, sdes, aT0,  \
                                                             want_qpdata, pn, Rn, dqn, env0);                            \
    break;
  switch (c->tpb) { UMPC_CONV_UPDATE(64) UMPC_CONV_UPDATE(128) }
  return (int)cudaGetLastError();
}

int umpc_conv_launch_rollout(UmpcConv* c, const SimParams* sim, const DevState* s, const WlParams* wlp, int step0, int nsteps,
                             const uint8_t* fire, const double* tvals, LogBuf lg, cudaStream_t stream) {
  KernelParams64 kp;
  kp.mpc = c->mpc; kp.sim = *sim;
  const int grid = (s->B + c->tpb - 1) / c->tpb;
#define UMPC_CONV_ROLLOUT(T)                                                                                              \
  case T:                                                                                                                 \
    umpc_rollout_kernel<T, SolverF64><<<grid, T, 0, stream>>>(kp, *s, *wlp, step0, nsteps, fire, tvals, lg, 1);             \
    break;
  switch (c->tpb) { UMPC_CONV_ROLLOUT(64) UMPC_CONV_ROLLOUT(128) }
  return (int)cudaGetLastError();
}
