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
#define UMPC_SCALE_FLOAT  // the ten Ruiz passes in float (umpc_core.cuh `sreal`); host A/B: tests/hostcheck libhostcheck64_sf.so
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

// ---- on-chip store (default): the 256-word scratch in TENSOR MEMORY, the loop vectors in SHARED MEMORY -------------------
// One 4-warp CTA per SM owns all of tensor memory (128 lanes x 512 columns x 32 bit): thread i of warp w reads "word e
// of my own QP" as columns 2e, 2e+1 of lane 32 (w % 4) + i with a warp-uniform address, 256 doubles per thread = the whole
// 256 KB.  A chunk of 16 doubles is one tcgen05.ld.32x32b.x32; the load lands in 32 raw registers that are in/out
// operands of the wait (nothing can be scheduled above it) and are paired into doubles after it -- exactly one load
// is in flight at any time in the generated solve (issue / wait alternate), so one raw buffer serves both chunk arrays.
// In double the ADMM loop's state would be 252 registers (xs 45 + w 36 + r 45 doubles): xs and w -- touched element by
// element, twice per iteration -- and the first-iteration fix-up vector live in shared memory as [word][thread]
// (conflict-free 8-byte accesses, immediate offsets from one per-thread pointer); r, which the triangular solves
// chain through, stays in registers.  117 doubles per thread = 117 KB per CTA: also what keeps a second CTA (which
// could not get tensor memory anyway) off the SM.
// The ADMM loop reads the factor twice per iteration (31 chunks = 127 KB per warp): a warp gets ~23 B/clk out of tensor
// memory however many loads it keeps in flight (tests/microbench/tmem_bw.cu), i.e. >= 5.5k cycles per iteration with all
// chunks there.  The chunks whose bit is set in CONV_SMEM_MASK live in the spare shared memory instead ([slot][pair]
// [thread][2] doubles: a chunk = 8 conflict-free LDS.128), a second, independent read path with a fifth of the latency;
// alternating chunks, so that a tensor-memory load is in flight while a shared-memory chunk is consumed.
#define CONV_TPB 128
#ifndef CONV_SMEM_MASK
#define CONV_SMEM_MASK 0x0AAAu   // chunks 1, 3, .., 11 of the 14 factor chunks
#endif
__host__ __device__ constexpr int conv_popc(unsigned v) { int n = 0; while (v) { n += v & 1u; v >>= 1; } return n; }
enum { CONV_SV_XS = 0, CONV_SV_VF = UMPC_NX + UMPC_NEQ, CONV_SV_WORDS = UMPC_NX + 2 * UMPC_NEQ,
       CONV_SMEM_CHUNKS = conv_popc(CONV_SMEM_MASK) };
extern __shared__ double g_smem64[];

struct TmemStore64 {
  typedef void has_vec;
  static __host__ __device__ constexpr bool in_smem(int c) { return (CONV_SMEM_MASK >> c) & 1u; }
  static __host__ __device__ constexpr int slot(int c) { return conv_popc(CONV_SMEM_MASK & ((1u << c) - 1u)); }
  // this thread's pair p of shared-memory chunk slot q
  __device__ __forceinline__ double2* spair(int q, int p) const { return (double2*)(sf + ((size_t)(q * (UMPC_GEN_CHUNK / 2) + p) * CONV_TPB) * 2); }
  template <int N, int OFF>
  struct Vec {
    double* b;
    __device__ __forceinline__ explicit Vec(const TmemStore64& st) : b(st.sv + OFF * CONV_TPB) {}
    __device__ __forceinline__ double& operator[](int i) const { return b[i * CONV_TPB]; }
  };
  uint32_t tbase;  // TMEM address of this thread's lane block, column 0
  double* sv;      // this thread's column of the CTA's shared vectors
  double* sf;      // this thread's pair column of the shared-memory factor chunks
  uint32_t raw[32];
  __device__ __forceinline__ void init(uint32_t tbase_) {
    tbase = tbase_; sv = g_smem64 + threadIdx.x;
    sf = g_smem64 + (size_t)CONV_SV_WORDS * CONV_TPB + 2 * threadIdx.x;
  }
  __device__ __forceinline__ void begin_iteration(int) {}
  __device__ __forceinline__ void issue(int c, double (&v)[UMPC_GEN_CHUNK]) {
    if (in_smem(c)) {
#pragma unroll
      for (int p = 0; p < UMPC_GEN_CHUNK / 2; ++p) { const double2 t = *spair(slot(c), p); v[2 * p] = t.x; v[2 * p + 1] = t.y; }
      return;
    }
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                 : "=r"(raw[0]), "=r"(raw[1]), "=r"(raw[2]), "=r"(raw[3]), "=r"(raw[4]), "=r"(raw[5]), "=r"(raw[6]), "=r"(raw[7]), "=r"(raw[8]), "=r"(raw[9]), "=r"(raw[10]), "=r"(raw[11]), "=r"(raw[12]), "=r"(raw[13]), "=r"(raw[14]), "=r"(raw[15]), "=r"(raw[16]), "=r"(raw[17]), "=r"(raw[18]), "=r"(raw[19]), "=r"(raw[20]), "=r"(raw[21]), "=r"(raw[22]), "=r"(raw[23]), "=r"(raw[24]), "=r"(raw[25]), "=r"(raw[26]), "=r"(raw[27]), "=r"(raw[28]), "=r"(raw[29]), "=r"(raw[30]), "=r"(raw[31])
                 : "r"(tbase + (uint32_t)c * (2 * UMPC_GEN_CHUNK)));
  }
  __device__ __forceinline__ void waitc(int c, double (&v)[UMPC_GEN_CHUNK]) {
    if (in_smem(c)) return;
    asm volatile("tcgen05.wait::ld.sync.aligned;" : "+r"(raw[0]), "+r"(raw[1]), "+r"(raw[2]), "+r"(raw[3]), "+r"(raw[4]), "+r"(raw[5]), "+r"(raw[6]), "+r"(raw[7]), "+r"(raw[8]), "+r"(raw[9]), "+r"(raw[10]), "+r"(raw[11]), "+r"(raw[12]), "+r"(raw[13]), "+r"(raw[14]), "+r"(raw[15]), "+r"(raw[16]), "+r"(raw[17]), "+r"(raw[18]), "+r"(raw[19]), "+r"(raw[20]), "+r"(raw[21]), "+r"(raw[22]), "+r"(raw[23]), "+r"(raw[24]), "+r"(raw[25]), "+r"(raw[26]), "+r"(raw[27]), "+r"(raw[28]), "+r"(raw[29]), "+r"(raw[30]), "+r"(raw[31]) :: "memory");
#pragma unroll
    for (int i = 0; i < UMPC_GEN_CHUNK; ++i) v[i] = __hiloint2double((int)raw[2 * i + 1], (int)raw[2 * i]);
  }
  __device__ __forceinline__ void setW(int e, double v) {
    if (in_smem(e / UMPC_GEN_CHUNK)) {
      const int r = e % UMPC_GEN_CHUNK;
      ((double*)spair(slot(e / UMPC_GEN_CHUNK), r / 2))[r & 1] = v;
      return;
    }
    asm volatile("tcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {%1, %2};" ::"r"(tbase + 2u * (uint32_t)e), "r"(__double2loint(v)),
                 "r"(__double2hiint(v)) : "memory");
  }
  __device__ __forceinline__ void setChunk(int c, const double (&v)[UMPC_GEN_CHUNK]) {
    if (in_smem(c)) {
#pragma unroll
      for (int p = 0; p < UMPC_GEN_CHUNK / 2; ++p) *spair(slot(c), p) = make_double2(v[2 * p], v[2 * p + 1]);
      return;
    }
    int u[32];
#pragma unroll
    for (int i = 0; i < UMPC_GEN_CHUNK; ++i) { u[2 * i] = __double2loint(v[i]); u[2 * i + 1] = __double2hiint(v[i]); }
    asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(tbase + (uint32_t)c * (2 * UMPC_GEN_CHUNK)), "r"(u[0]), "r"(u[1]), "r"(u[2]), "r"(u[3]), "r"(u[4]), "r"(u[5]), "r"(u[6]), "r"(u[7]), "r"(u[8]), "r"(u[9]), "r"(u[10]), "r"(u[11]), "r"(u[12]), "r"(u[13]), "r"(u[14]), "r"(u[15]), "r"(u[16]), "r"(u[17]), "r"(u[18]), "r"(u[19]), "r"(u[20]), "r"(u[21]), "r"(u[22]), "r"(u[23]), "r"(u[24]), "r"(u[25]), "r"(u[26]), "r"(u[27]), "r"(u[28]), "r"(u[29]), "r"(u[30]), "r"(u[31]) : "memory");
  }
  __device__ __forceinline__ void end_writes() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
  __device__ __forceinline__ void sync() const {}
  __device__ __forceinline__ bool all(bool p) const { return __all_sync(0xffffffffu, p); }
  __device__ __forceinline__ double Vf(int i) const { return sv[(CONV_SV_VF + i) * CONV_TPB]; }
  __device__ __forceinline__ void setVf(int i, double v) { sv[(CONV_SV_VF + i) * CONV_TPB] = v; }
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

// all of tensor memory for one 128-thread CTA: 512 columns, warp w owns lanes 32 (w % 4) ..
static __device__ __forceinline__ uint32_t tmem64_acquire(uint32_t* slot) {
  if ((threadIdx.x >> 5) == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((uint32_t)__cvta_generic_to_shared(slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  return *slot + ((uint32_t)(32 * ((threadIdx.x >> 5) & 3)) << 16);
}
static __device__ __forceinline__ void tmem64_release(const uint32_t* slot) {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if ((threadIdx.x >> 5) == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(*slot) : "memory");
}

template <class ST> struct Store64Traits;
template <> struct Store64Traits<GlobStore64> {
  template <int TPB> static __device__ __forceinline__ uint32_t acquire(uint32_t*) { return 0u; }
  template <int TPB> static __device__ __forceinline__ void release(const uint32_t*) {}
  static __device__ __forceinline__ void init(GlobStore64& st, uint32_t, const DevState& s, int e) { st.init(s.c64, e); }
};
template <> struct Store64Traits<TmemStore64> {
  template <int TPB> static __device__ __forceinline__ uint32_t acquire(uint32_t* slot) {
    static_assert(TPB == CONV_TPB, "the tensor-memory store is laid out for 128-thread CTAs");
    return tmem64_acquire(slot);
  }
  template <int TPB> static __device__ __forceinline__ void release(const uint32_t* slot) { tmem64_release(slot); }
  static __device__ __forceinline__ void init(TmemStore64& st, uint32_t base, const DevState&, int) { st.init(base); }
};

// SOLVER policy of the kernel templates in umpc_dev.cuh
template <class ST64>
struct SolverF64T {
  typedef umpc64::UmpcParams Params;
  static constexpr int kMinBlocks = 1;
  static constexpr bool kDynamicItems = true;   // 25, 50, 75 .. iterations per tile: warps take their next tile when free
  ST64 st;
  template <int TPB>
  static __device__ __forceinline__ uint32_t acquire(uint32_t* slot) { return Store64Traits<ST64>::template acquire<TPB>(slot); }
  template <int TPB>
  static __device__ __forceinline__ void release(const uint32_t* slot) { Store64Traits<ST64>::template release<TPB>(slot); }
  __device__ __forceinline__ void init(uint32_t base, const DevState& s, int e) { Store64Traits<ST64>::init(st, base, s, e); }
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
typedef SolverF64T<GlobStore64> SolverF64;      // scratch streamed from L2 (the round-1 kernel; UMPC_CONV_STORE=glob)
typedef SolverF64T<TmemStore64> SolverF64Tmem;  // on-chip (default)

// ================================ host side ======================================================================
typedef KernelParamsT<umpc64::UmpcParams> KernelParams64;

struct UmpcConv {
  umpc64::UmpcParams mpc;
  double* slab = nullptr;
  int Bp = 0;
  int tpb = 64;  // streamed store only: 64 slightly ahead of 128 (UMPC_CONV_TPB to override)
  bool on_chip = true;  // TmemStore64 (one 128-thread CTA per SM); UMPC_CONV_STORE=glob selects the streamed store
  bool attr_set = false;
  int nsm = 148;
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
  if (const char* t = getenv("UMPC_CONV_STORE")) c->on_chip = strcmp(t, "glob") != 0;
  { int dev = 0; cudaGetDevice(&dev); int n = 0; if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0) c->nsm = n; }
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

static constexpr size_t kConvSmem = (size_t)(CONV_SV_WORDS + CONV_SMEM_CHUNKS * UMPC_GEN_CHUNK) * sizeof(double) * CONV_TPB;
static_assert(kConvSmem + 4096 <= 227 * 1024, "shared memory of the converged-mode CTA");

// opt-in to the 117 KB of dynamic shared memory, once per process (the attribute is per function, not per device context
// here: one device per process)
static int conv_set_attrs(UmpcConv* c) {
  if (c->attr_set) return 0;
  cudaError_t e = cudaFuncSetAttribute(umpc_update_kernel<CONV_TPB, SolverF64Tmem>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kConvSmem);
  if (e == cudaSuccess)
    e = cudaFuncSetAttribute(umpc_rollout_kernel<CONV_TPB, SolverF64Tmem>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kConvSmem);
  c->attr_set = e == cudaSuccess;
  return (int)e;
}

int umpc_conv_launch_update(UmpcConv* c, const SimParams* sim, const DevState* s, float* uquad, float* accdes, const float* p0,
                            const float* R0, const float* dq0, const float* pdes, const float* dpdes, const float* sdes,
                            const float* aT0, int want_qpdata, float* pn, float* Rn, float* dqn, int env0, int nenv,
                            cudaStream_t stream) {
  KernelParams64 kp;
  kp.mpc = c->mpc; kp.sim = *sim;
  if (c->on_chip) {
    if (int rc = conv_set_attrs(c)) return rc;
    umpc_update_kernel<CONV_TPB, SolverF64Tmem><<<(nenv + CONV_TPB - 1) / CONV_TPB, CONV_TPB, kConvSmem, stream>>>(
        kp, *s, uquad, accdes, p0, R0, dq0, pdes, dpdes, sdes, aT0, want_qpdata, pn, Rn, dqn, env0);
    return (int)cudaGetLastError();
  }
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
  if (c->on_chip) {
    if (int rc = conv_set_attrs(c)) return rc;
    // persistent: one CTA per SM, its warps take the CTA's tiles (t = blockIdx + k gridDim) one after the other
    const int ntiles = s->B / 32;
    int grid = c->nsm < ntiles ? c->nsm : ntiles;
    if ((ntiles + grid - 1) / grid > UMPC_MAX_CHUNKED_TILES) grid = (ntiles + UMPC_MAX_CHUNKED_TILES - 1) / UMPC_MAX_CHUNKED_TILES;
    // Whole tiles per item by default.  Cutting a multi-step launch into (tile, chunk) items (UMPC_CONV_CHUNKS) was measured
    // and is SLOWER here (25 steps per launch: 1 chunk 1.27 ms per step, 5 chunks 1.41, 25 chunks 1.41; one step per launch
    // 1.185 -- profiles/r02C_conv_steps_per_launch.txt): launch boundaries keep the four warps of an SM in the same phase of
    // the solve, and warps in different phases of this 690 KB kernel do not share the instruction cache.
    int nch = 1;
    if (const char* ev = getenv("UMPC_CONV_CHUNKS")) nch = atoi(ev);
    if (nch < 1 || nch > nsteps) nch = 1;
    umpc_rollout_kernel<CONV_TPB, SolverF64Tmem><<<grid, CONV_TPB, kConvSmem, stream>>>(kp, *s, *wlp, step0, nsteps, fire, tvals, lg, nch);
    return (int)cudaGetLastError();
  }
  const int grid = (s->B + c->tpb - 1) / c->tpb;
#define UMPC_CONV_ROLLOUT(T)                                                                                              \
  case T:                                                                                                                 \
    umpc_rollout_kernel<T, SolverF64><<<grid, T, 0, stream>>>(kp, *s, *wlp, step0, nsteps, fire, tvals, lg, 1);             \
    break;
  switch (c->tpb) { UMPC_CONV_ROLLOUT(64) UMPC_CONV_ROLLOUT(128) }
  return (int)cudaGetLastError();
}
