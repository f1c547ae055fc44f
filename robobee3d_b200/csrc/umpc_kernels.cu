// robobee3d_b200/csrc/umpc_kernels.cu -- CUDA kernels (sm_100a) + C-ABI of libumpc_b200.so.
//
// Mapping: ONE THREAD PER ENVIRONMENT.  Every QP has the same sparsity, so the whole
// update is straight-line code with compile-time indices (umpc_gen.cuh); the ADMM
// vectors live in registers, the per-QP factor of K (167 + 45 values) and the sigma/D^2
// pairs live in a 256-word per-thread scratch in TENSOR MEMORY (tcgen05.ld/st; a
// shared-memory store with the same interface is the second launch shape), and the warm
// start / plant state live in HBM as a struct of arrays tiled by warp (umpc_dev.cuh): a
// warp's access to one component is one 128-byte line at an immediate offset from one
// base pointer.  There is no inter-thread communication at all; the kernels are bound by
// the FP32 pipe / instruction issue, not by HBM (DESIGN.md has the byte/flop accounting).
// The kernel templates themselves are in umpc_dev.cuh (shared with umpc_conv.cu, the
// float64 converged mode); this file holds the float32 solver policy, the scratch stores,
// the WL / reduction kernels and the host side of the C-ABI.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <stdlib.h>
#include <string>
#include <vector>
#include <map>
#include <mutex>

#include "umpc_core.cuh"
#include "umpc_dev.cuh"
#include "umpc_conv.h"
#include "../../include/umpc_b200.h"

// ---- per-QP scratch stores ---------------------------------------------------------------------------
// The scratch of a QP is UMPC_GEN_NWORDS 32-bit words (factor in stream order + sigma/D^2 pairs), requested
// UMPC_GEN_CHUNK = 16 words at a time by the generated solve: issue(c, buf) ... wait(buf).
//
// (1) SmemStore: shared memory laid out [group of 4 words][thread][4]: a chunk is four 128-bit LDS, a quarter
//     warp reads 128 contiguous bytes (conflict free).  One-warp CTAs, 6 per SM (32 KB each).
// (2) TmemStore: Blackwell tensor memory (256 KB per SM, 128 lanes x 512 columns x 32 bit) used as PER-THREAD
//     PRIVATE storage: tcgen05.ld/st.32x32b give thread i of a warp lane (32*(warp%4) + i) and N consecutive
//     columns -- exactly "word e of my own QP" for every lane, with a warp-uniform address.  256 words per
//     thread x 8 warps = the whole TMEM; one tcgen05.ld.x16 per chunk, completion by tcgen05.wait::ld.  It
//     takes the factor traffic off the shared-memory pipe entirely (measured >= 188 B/clk/SM, tests/microbench).
#define SM_WORDS_USED UMPC_GEN_NWORDS
#define SM_FLOATS_PER_THREAD SM_WORDS_USED

template <int TPB>
struct SmemStore {
  static constexpr bool kTmem = false;
  static constexpr uint32_t gstride = TPB * 16u;  // byte stride between groups: compile-time -> immediate offsets
  uint32_t sbase;  // shared-window byte address of this thread's first group
  float* vf;       // global scratch for the first-iteration fix-up (the env's soly rows, dead until the end)
  __device__ __forceinline__ void init(uint32_t sbase_, float* vf_) { sbase = sbase_; vf = vf_; }
  __device__ __forceinline__ void begin_iteration(int) {}
  // volatile: keeps the generated request order (the software pipeline) and re-loads every pass
  __device__ __forceinline__ void issue(int c, float (&v)[UMPC_GEN_CHUNK]) const {
#pragma unroll
    for (int g = 0; g < UMPC_GEN_CHUNK / 4; ++g) {
      asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(v[4 * g]), "=f"(v[4 * g + 1]), "=f"(v[4 * g + 2]), "=f"(v[4 * g + 3])
                     : "r"(sbase + (uint32_t)(c * (UMPC_GEN_CHUNK / 4) + g) * gstride));
    }
  }
  // nothing to wait for (LDS results are scoreboarded); the empty asm pins the buffer to registers at this point
  __device__ __forceinline__ void wait(float (&v)[UMPC_GEN_CHUNK]) const {
    asm volatile("" : "+f"(v[0]), "+f"(v[1]), "+f"(v[2]), "+f"(v[3]), "+f"(v[4]), "+f"(v[5]), "+f"(v[6]), "+f"(v[7]), "+f"(v[8]),
                 "+f"(v[9]), "+f"(v[10]), "+f"(v[11]), "+f"(v[12]), "+f"(v[13]), "+f"(v[14]), "+f"(v[15]));
  }
  __device__ __forceinline__ void waitc(int, float (&v)[UMPC_GEN_CHUNK]) const { wait(v); }
  __device__ __forceinline__ void setW(int e, float v) {
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(sbase + (uint32_t)(e >> 2) * gstride + (uint32_t)(e & 3) * 4u), "f"(v) : "memory");
  }
  __device__ __forceinline__ void setChunk(int c, const float (&v)[UMPC_GEN_CHUNK]) {
#pragma unroll
    for (int g = 0; g < UMPC_GEN_CHUNK / 4; ++g)
      asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(sbase + (uint32_t)(c * (UMPC_GEN_CHUNK / 4) + g) * gstride),
                   "f"(v[4 * g]), "f"(v[4 * g + 1]), "f"(v[4 * g + 2]), "f"(v[4 * g + 3]) : "memory");
  }
  __device__ __forceinline__ void end_writes() {}
  __device__ __forceinline__ void sync() const {}
  __device__ __forceinline__ float Vf(int i) const { return vf[ROWF(i)]; }
  __device__ __forceinline__ void setVf(int i, float v) { vf[ROWF(i)] = v; }
};

struct TmemStore {
  static constexpr bool kTmem = true;
  uint32_t tbase;  // TMEM address of this warp's word 0: lane block 32*(warp%4), column block 256*(warp/4)
  float* vf;
  __device__ __forceinline__ void init(uint32_t tbase_, float* vf_) { tbase = tbase_; vf = vf_; }
  __device__ __forceinline__ void begin_iteration(int) {}
  __device__ __forceinline__ void issue(int c, float (&v)[UMPC_GEN_CHUNK]) const {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]), "=f"(v[8]),
          "=f"(v[9]), "=f"(v[10]), "=f"(v[11]), "=f"(v[12]), "=f"(v[13]), "=f"(v[14]), "=f"(v[15])
        : "r"(tbase + (uint32_t)c * UMPC_GEN_CHUNK));
  }
  // the loaded registers are in/out operands so that no use of them can be scheduled above the wait
  __device__ __forceinline__ void wait(float (&v)[UMPC_GEN_CHUNK]) const {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+f"(v[0]), "+f"(v[1]), "+f"(v[2]), "+f"(v[3]), "+f"(v[4]), "+f"(v[5]), "+f"(v[6]), "+f"(v[7]), "+f"(v[8]),
                   "+f"(v[9]), "+f"(v[10]), "+f"(v[11]), "+f"(v[12]), "+f"(v[13]), "+f"(v[14]), "+f"(v[15])
                 :: "memory");
  }
  __device__ __forceinline__ void waitc(int, float (&v)[UMPC_GEN_CHUNK]) const { wait(v); }
  __device__ __forceinline__ void setW(int e, float v) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(tbase + (uint32_t)e), "f"(v) : "memory");
  }
  __device__ __forceinline__ void setChunk(int c, const float (&v)[UMPC_GEN_CHUNK]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
        ::"r"(tbase + (uint32_t)c * UMPC_GEN_CHUNK), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]), "f"(v[4]), "f"(v[5]), "f"(v[6]),
        "f"(v[7]), "f"(v[8]), "f"(v[9]), "f"(v[10]), "f"(v[11]), "f"(v[12]), "f"(v[13]), "f"(v[14]), "f"(v[15]) : "memory");
  }
  __device__ __forceinline__ void end_writes() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
  // Optional CTA-wide rendezvous at phase boundaries (-DUMPC_PHASE_SYNC): keeps the warps of a CTA on the same
  // straight-line code so that each instruction-cache line of this ~250 KB kernel is fetched once per CTA.  Measured
  // on B200: no gain (0.338 vs 0.327 ms per 65,536 env-steps) -- the once-per-solve code is bound by the LATENCY of
  // the sequential instruction fetch (ncu: ~10 cycles per cold instruction), which a rendezvous does not shorten,
  // while unsynchronised warps overlap it with other warps' hot loops.  Off by default.
#ifdef UMPC_PHASE_SYNC
  __device__ __forceinline__ void sync() const { __syncthreads(); }
#else
  __device__ __forceinline__ void sync() const {}
#endif
  __device__ __forceinline__ float Vf(int i) const { return vf[ROWF(i)]; }
  __device__ __forceinline__ void setVf(int i, float v) { vf[ROWF(i)] = v; }
};


// (3) HybridStore: the round-1 profile put the ADMM loop at 0.66 IPC per scheduler with ONE or TWO warps on it -- the
//     loop is bound by the TMEM read path (31 x 2 KB per warp and pass = ~184 B/clk/SM, tests/microbench/tmem_bw.cu
//     measures ~190), not by issue.  Shared memory is a second, independent pipe (128 B/clk/SM): the chunks whose bit
//     is set in MASK live there ([chunk slot][group of 4 words][thread][4]: a chunk = four conflict-free LDS.128),
//     the others in TMEM, so both pipes stream the factor at the same time.  MASK = 0 is the pure-TMEM store.
__host__ __device__ constexpr int umpc_popc(unsigned v) { int n = 0; while (v) { n += v & 1u; v >>= 1; } return n; }
#ifndef UMPC_SMEM_MASK
#define UMPC_SMEM_MASK 0x0000u
#endif
template <int TPB, unsigned MASK>
struct HybridStore {
  static constexpr bool kTmem = true;
  static constexpr unsigned kMask = MASK;
  static constexpr int kSmemChunks = umpc_popc(MASK);
  static constexpr size_t kSmemBytes = (size_t)kSmemChunks * UMPC_GEN_CHUNK * 4 * TPB;
  static constexpr uint32_t gstride = TPB * 16u;
  uint32_t tbase, sbase;
  float* vf;
  __device__ __forceinline__ void init(uint32_t tbase_, float* vf_);
  __device__ __forceinline__ void begin_iteration(int) {}
  static __host__ __device__ constexpr bool in_smem(int c) { return (MASK >> c) & 1u; }
  static __host__ __device__ constexpr int slot(int c) { return umpc_popc(MASK & ((1u << c) - 1u)); }
  __device__ __forceinline__ void issue(int c, float (&v)[UMPC_GEN_CHUNK]) const {
    if (in_smem(c)) {
#pragma unroll
      for (int g = 0; g < UMPC_GEN_CHUNK / 4; ++g)
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(v[4 * g]), "=f"(v[4 * g + 1]), "=f"(v[4 * g + 2]), "=f"(v[4 * g + 3])
                     : "r"(sbase + (uint32_t)(slot(c) * (UMPC_GEN_CHUNK / 4) + g) * gstride));
    } else {
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
          : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]), "=f"(v[8]),
            "=f"(v[9]), "=f"(v[10]), "=f"(v[11]), "=f"(v[12]), "=f"(v[13]), "=f"(v[14]), "=f"(v[15])
          : "r"(tbase + (uint32_t)c * UMPC_GEN_CHUNK));
    }
  }
  __device__ __forceinline__ void waitc(int c, float (&v)[UMPC_GEN_CHUNK]) const {
    if (in_smem(c))
      asm volatile("" : "+f"(v[0]), "+f"(v[1]), "+f"(v[2]), "+f"(v[3]), "+f"(v[4]), "+f"(v[5]), "+f"(v[6]), "+f"(v[7]), "+f"(v[8]),
                   "+f"(v[9]), "+f"(v[10]), "+f"(v[11]), "+f"(v[12]), "+f"(v[13]), "+f"(v[14]), "+f"(v[15]));
    else
      asm volatile("tcgen05.wait::ld.sync.aligned;"
                   : "+f"(v[0]), "+f"(v[1]), "+f"(v[2]), "+f"(v[3]), "+f"(v[4]), "+f"(v[5]), "+f"(v[6]), "+f"(v[7]), "+f"(v[8]),
                     "+f"(v[9]), "+f"(v[10]), "+f"(v[11]), "+f"(v[12]), "+f"(v[13]), "+f"(v[14]), "+f"(v[15])
                   :: "memory");
  }
  __device__ __forceinline__ void setW(int e, float v) {
    const int c = e / UMPC_GEN_CHUNK, r = e % UMPC_GEN_CHUNK;
    if (in_smem(c))
      asm volatile("st.shared.f32 [%0], %1;" ::"r"(sbase + (uint32_t)(slot(c) * (UMPC_GEN_CHUNK / 4) + r / 4) * gstride + (uint32_t)(r & 3) * 4u), "f"(v) : "memory");
    else
      asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(tbase + (uint32_t)e), "f"(v) : "memory");
  }
  __device__ __forceinline__ void setChunk(int c, const float (&v)[UMPC_GEN_CHUNK]) {
    if (in_smem(c)) {
#pragma unroll
      for (int g = 0; g < UMPC_GEN_CHUNK / 4; ++g)
        asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(sbase + (uint32_t)(slot(c) * (UMPC_GEN_CHUNK / 4) + g) * gstride),
                     "f"(v[4 * g]), "f"(v[4 * g + 1]), "f"(v[4 * g + 2]), "f"(v[4 * g + 3]) : "memory");
    } else {
      asm volatile(
          "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
          ::"r"(tbase + (uint32_t)c * UMPC_GEN_CHUNK), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]), "f"(v[4]), "f"(v[5]), "f"(v[6]),
          "f"(v[7]), "f"(v[8]), "f"(v[9]), "f"(v[10]), "f"(v[11]), "f"(v[12]), "f"(v[13]), "f"(v[14]), "f"(v[15]) : "memory");
    }
  }
  __device__ __forceinline__ void end_writes() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
  __device__ __forceinline__ void sync() const {}
  __device__ __forceinline__ float Vf(int i) const { return vf[ROWF(i)]; }
  __device__ __forceinline__ void setVf(int i, float v) { vf[ROWF(i)] = v; }
};

// TMEM allocation for a CTA of TPB threads: (TPB / 128) * 256 columns, all 128 lanes.  Warp 0 allocates and frees.
template <int TPB>
__device__ __forceinline__ uint32_t tmem_acquire(uint32_t* slot) {
  constexpr int ncols = (TPB / 128) * UMPC_GEN_NWORDS;
  if ((threadIdx.x >> 5) == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(slot)), "n"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const int warp = threadIdx.x >> 5;
  return *slot + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)((warp >> 2) * UMPC_GEN_NWORDS);
}
template <int TPB>
__device__ __forceinline__ void tmem_release(const uint32_t* slot) {
  constexpr int ncols = (TPB / 128) * UMPC_GEN_NWORDS;
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if ((threadIdx.x >> 5) == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(*slot), "n"(ncols) : "memory");
}

struct GlobIO {
  float* eb;  // env_base(): every row of this env at a compile-time offset
  __device__ __forceinline__ explicit GlobIO(float* eb_) : eb(eb_) {}
  // (a prefetch.global.L2 of the 123 warm-start rows here cost ~500 once-per-solve instructions -- address arithmetic
  // plus CCTL -- for a latency that the 123 independent loads issued together later overlap anyway; removed)
  __device__ __forceinline__ void prefetch_warm_start() const {}
  // inputs parked over the ADMM loop (solx is dead until the solution is stored); volatile read-back so that the
  // compiler cannot keep the values in registers instead
  __device__ __forceinline__ void stash(int k, float v) { eb[ROWF(ROW_SOLX + k)] = v; }
  __device__ __forceinline__ float unstash(int k) const { return *(volatile const float*)(eb + ROWF(ROW_SOLX + k)); }
  __device__ __forceinline__ float ld_xs(int j) const { return eb[ROWF(ROW_XS + j)]; }
  __device__ __forceinline__ float ld_zs(int i) const { return eb[ROWF(ROW_ZS + i)]; }
  __device__ __forceinline__ float ld_ys(int i) const { return eb[ROWF(ROW_YS + i)]; }
  __device__ __forceinline__ float ld_D(int j) const { return eb[ROWF(ROW_D + j)]; }
  __device__ __forceinline__ float ld_E(int i) const { return eb[ROWF(ROW_E + i)]; }
  __device__ __forceinline__ void st_xs(int j, float v) { eb[ROWF(ROW_XS + j)] = v; }
  __device__ __forceinline__ void st_zs(int i, float v) { eb[ROWF(ROW_ZS + i)] = v; }
  __device__ __forceinline__ void st_ys(int i, float v) { eb[ROWF(ROW_YS + i)] = v; }
  __device__ __forceinline__ void st_D(int j, float v) { eb[ROWF(ROW_D + j)] = v; }
  __device__ __forceinline__ void st_E(int i, float v) { eb[ROWF(ROW_E + i)] = v; }
  __device__ __forceinline__ void st_solx(int j, float v) { eb[ROWF(ROW_SOLX + j)] = v; }
  __device__ __forceinline__ void st_soly(int i, float v) { eb[ROWF(ROW_SOLY + i)] = v; }
};

#ifndef UMPC_DEFAULT_TPB
#define UMPC_DEFAULT_TPB 128  // TMEM scratch, two 4-warp CTAs per SM
#endif

typedef KernelParamsT<UmpcParams> KernelParams;

extern __shared__ float g_smem[];
template <int TPB, unsigned MASK>
__device__ __forceinline__ void HybridStore<TPB, MASK>::init(uint32_t tbase_, float* vf_) {
  tbase = tbase_; vf = vf_;
  sbase = (uint32_t)__cvta_generic_to_shared(g_smem) + threadIdx.x * 16u;
}
// default scratch store of the TMEM launch shapes
template <int TPB> using DefStore = HybridStore<TPB, UMPC_SMEM_MASK>;

// The C path's SOLVER policy (umpc_dev.cuh): float32, scratch store ST, exactly max_iter iterations.
// PERENV: cost weights per environment (DevState::wts) instead of the batch-wide ones; a separate instantiation so
// that the default kernel keeps its weights as constant-bank operands.
template <int TPB_, class ST, bool PERENV = false, bool INFO = true>
struct SolverF32 {
  typedef UmpcParams Params;
  static constexpr int kMinBlocks = ST::kTmem ? 1 : 224 / TPB_;
  static constexpr bool kDynamicItems = false;  // fixed iteration count: every item of a launch takes the same time
  ST st;
  float* eb;
  template <int TPB>
  static __device__ __forceinline__ uint32_t acquire(uint32_t* slot) {
    if (ST::kTmem) return tmem_acquire<TPB>(slot);
    return (uint32_t)__cvta_generic_to_shared(g_smem) + threadIdx.x * 16u;
  }
  template <int TPB>
  static __device__ __forceinline__ void release(const uint32_t* slot) {
    if (ST::kTmem) tmem_release<TPB>(slot);
  }
  __device__ __forceinline__ void init(uint32_t scratch_base, const DevState& s, int e) {
    eb = env_base(s, e);
    st.init(scratch_base, eb + ROWF(ROW_SOLY));
  }
  __device__ __forceinline__ void qpdata(const Params& par, const DevState& s, int e, const float p[3], const float R[9],
                                         const float dq[6], const float pd[3], const float dpd[3], const float sd[3], float aT0) {
    UmpcQP qp;
    umpc_build(par, qp, p, R, dq, pd, dpd, sd, aT0 >= 0.0f ? aT0 : eb[ROWF(ROW_T0)]);
    float out[216];
    umpc_qpdata(par, qp, out);
    for (int k = 0; k < 216; ++k) s.qpdata[(size_t)k * s.B + e] = out[k];
  }
  __device__ __forceinline__ void update(const Params& par, const DevState& s, int e, float uq[3], float acc[6], const float p[3],
                                         const float R[9], const float dq[6], const float pd[3], const float dpd[3],
                                         const float sd[3], float aT0) {
    GlobIO io(eb);
    float T0 = eb[ROWF(ROW_T0)];
    UmpcInfo info;
    if constexpr (PERENV) {
      Params pe = par;
      load_env_weights(pe, s, e);
      umpc_update<false, INFO>(pe, st, io, uq, acc, p, R, dq, pd, dpd, sd, aT0, T0, info);
    } else {
      umpc_update<false, INFO>(par, st, io, uq, acc, p, R, dq, pd, dpd, sd, aT0, T0, info);
    }
    eb[ROWF(ROW_T0)] = T0;
    store_info(eb, info.iter, info.status, info.pri_res, info.dua_res);
  }
};

// ---- kernel 3: WL stage alone (drop-in wlConUpdate / wlControllerUpdate, batch of B) ---------
__global__ void wl_update_kernel(const __grid_constant__ WlParams wlp, const __grid_constant__ DevState s, int mode,
                                 float* __restrict__ u1, float* __restrict__ w0o, const float* __restrict__ h0,
                                 const float* __restrict__ pdotdes, float* __restrict__ info2) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= s.nreal) return;
  float* const eb = env_base(s, e);
  float h[6], pd[6], u0[4], w0[6];
#pragma unroll
  for (int i = 0; i < 6; ++i) { h[i] = h0[(size_t)e * 6 + i]; pd[i] = pdotdes[(size_t)e * 6 + i]; }
#pragma unroll
  for (int i = 0; i < 4; ++i) u0[i] = eb[ROWF(ROW_WLU0 + i)];
  int iter = 0, status = 0;
  if (mode == 1) {
    wl_con_update(wlp, u0, w0, h, pd);
  } else {
    WlqpState w;
    load_wlqp(eb, w);
    wlqp_update(wlp, w, u0, w0, h, pd, &iter, &status);
    store_wlqp(eb, w);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) { eb[ROWF(ROW_WLU0 + i)] = u0[i]; u1[(size_t)e * 4 + i] = u0[i]; }
#pragma unroll
  for (int i = 0; i < 6; ++i) w0o[(size_t)e * 6 + i] = w0[i];
  if (info2) { info2[(size_t)e * 2] = (float)iter; info2[(size_t)e * 2 + 1] = (float)status; }
}

// ---- kernel 4: metric reduction (what ranks all-reduce at the end of a run) -------------------
__global__ void reduce_metrics_kernel(const __grid_constant__ DevState s, double* __restrict__ out8) {
  double a[6] = {0, 0, 0, 0, 0, 0};
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < s.nreal; e += gridDim.x * blockDim.x) {
    const float* eb = env_base(s, e);
    const float err = eb[ROWF(ROW_METRICS + 0)], eff = eb[ROWF(ROW_METRICS + 1)];
    const int ns = ((const int*)eb)[ROWF(ROW_METRICS + 3)];
    bool bad = !(isfinite(err) && isfinite(eff));
#pragma unroll
    for (int i = 0; i < 3; ++i) bad = bad || !isfinite(eb[ROWF(ROW_P + i)]);
    if (!bad) { a[0] += err; a[1] += eff; a[5] = fmax(a[5], ns > 0 ? (double)err / ns : 0.0); }
    a[2] += ((const int*)eb)[ROWF(ROW_METRICS + 2)]; a[3] += ns; a[4] += bad ? 1.0 : 0.0;
  }
#pragma unroll
  for (int k = 0; k < 6; ++k)
    for (int o = 16; o > 0; o >>= 1) {
      const double v = __shfl_xor_sync(0xffffffffu, a[k], o);
      a[k] = (k == 5) ? fmax(a[k], v) : a[k] + v;
    }
  if ((threadIdx.x & 31) == 0) {
#pragma unroll
    for (int k = 0; k < 5; ++k) atomicAdd(&out8[k], a[k]);
    // max via CAS on the bit pattern (values are non-negative)
    unsigned long long* addr = (unsigned long long*)&out8[5];
    unsigned long long old = *addr, assumed;
    do {
      assumed = old;
      if (__longlong_as_double((long long)assumed) >= a[5]) break;
      old = atomicCAS(addr, assumed, (unsigned long long)__double_as_longlong(a[5]));
    } while (assumed != old);
  }
}

// ---- kernel 4b: set rows [row0, row0+k) of every env to a value (state initialisation) ---------------
__global__ void fill_rows_kernel(float* __restrict__ slab, int nslots, int row0, int k, float v) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= nslots) return;
  float* eb = slab + tile_off(e);
  for (int i = 0; i < k; ++i) eb[ROWF(row0 + i)] = v;
}

// ---- kernel 5: FP32 pipe peak (independent FFMA chains in registers) --------------------------
__global__ void fp32_peak_kernel(float* out, int iters) {
  float a[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = 1.0f + 1e-3f * (threadIdx.x + i);
  const float m = 0.999f, c = 1e-3f;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
      for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], m, c);
  }
  float sacc = 0.0f;
#pragma unroll
  for (int i = 0; i < 16; ++i) sacc += a[i];
  if (sacc == 12345.678f) out[0] = sacc;
}

// ---- kernel 5b: FP64 pipe peak (independent DFMA chains in registers) -----------------------
__global__ void fp64_peak_kernel(double* out, int iters) {
  double a[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = 1.0 + 1e-3 * (threadIdx.x + i);
  const double m = 0.999, c = 1e-3;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = fma(a[i], m, c);
  }
  double sacc = 0.0;
#pragma unroll
  for (int i = 0; i < 8; ++i) sacc += a[i];
  if (sacc == 12345.678) out[0] = sacc;
}

// ================================ host side: C-ABI ============================================
static thread_local std::string g_err;
static int set_err(const char* what, cudaError_t e) {
  char buf[512];
  snprintf(buf, sizeof(buf), "%s: %s", what, cudaGetErrorString(e));
  g_err = buf;
  return UMPC_ERR_CUDA;
}
#define CK(call)                                              \
  do {                                                        \
    cudaError_t e_ = (call);                                  \
    if (e_ != cudaSuccess) return set_err(#call, e_);         \
  } while (0)

extern "C" const char* umpcLastError(void) { return g_err.c_str(); }
void umpc_internal_set_error(const char* msg) { g_err = msg; }  // for the other translation units (dirtran.cu)

struct UmpcBatch {
  int nsm = 148;           // SMs of the device
  bool tpb_user = false;   // launch shape chosen through umpcBatchSetLaunch / UMPC_TPB
  bool rollout_info = false;  // rollouts also evaluate residuals / status / solution rows after every solve
  int B = 0, Bp = 0, device = 0, tpb = 128;  // B: caller's env count, Bp: padded to whole warps (device stride)
  cudaStream_t stream = nullptr;
  cudaStream_t side[3] = {nullptr, nullptr, nullptr};  // chunk pipeline of the host-buffer calls
  cudaEvent_t side_ev = nullptr, join_ev[3] = {nullptr, nullptr, nullptr};
  unsigned* d_pulled = nullptr;         // zero-copy host-buffer calls: ordered-pull counter
  cudaGraphExec_t g_exec = nullptr;     // captured chunk pipeline of the host-buffer calls + what it was captured for
  std::vector<const void*> g_key;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, tev0 = nullptr, tev1 = nullptr;
  cudaEvent_t x_ev0 = nullptr, x_ev1 = nullptr;  // hand-shake with a caller's stream (umpcBatchUpdateDevOn)
  bool timed = false;
  UmpcBatchParams bp;
  KernelParams kp;
  WlParams wlp;
  DevState s;
  UmpcConv* conv = nullptr;  // converged mode (float64 solver state, umpc_conv.cu); nullptr: the C path
  float* slab = nullptr;
  size_t slab_floats = 0;
  // staging ([B][k] layout of the C-ABI)
  float *d_in = nullptr, *d_out = nullptr;  // in: 28 floats/env, out: 9 floats/env
  float* d_next = nullptr;                  // next plant state of umpcBatchEnvStep: 18 floats/env
  // rollout schedule (host-evaluated in double like numpy, uprightmpc2.py:118,136,145)
  long long step = 0;
  double thl = 0.0;
  uint8_t* d_fire = nullptr;
  double* d_t = nullptr;
  int sched_cap = 0;
  double* d_sums = nullptr;
  LogBuf lg{-1, 0, nullptr, nullptr};
  bool have_qpdata = false;
  // batch-of-1 drop-in (umpcUpdate): one pinned, device-mapped block -- in[28] | out[9] | qp[216][256] -- that the update
  // kernel reads and writes directly (no staging copies); s.qpdata points into it
  float *h_single = nullptr, *d_single = nullptr;
};

extern "C" void umpcBatchDefaults(UmpcBatchParams* p) {
  memset(p, 0, sizeof(*p));
  // createMPC defaults (template_controllers.py:260-280)
  p->dt = 5.0f; p->g = 9.81e-3f; p->TtoWmax = 2.0f; p->ws = 1e1f; p->wds = 1e3f; p->wpr = 1.0f; p->wpf = 5.0f;
  p->wvr = 1e3f; p->wvf = 2e3f; p->wthrust = 1e-1f; p->wmom = 1e-2f;
  p->Ib[0] = 3333.0f; p->Ib[1] = 3333.0f; p->Ib[2] = 1000.0f; p->maxIter = 50;
  // controlTest defaults (uprightmpc2.py:87)
  p->dtsim = 0.2; p->hlInterval = -1.0; p->taulim = 100.0f; p->tpert = -1.0f; p->mb = 100.0f;
  p->wlMode = 0;
  // WaypointHover WLCon limits (robobee_test_controllers.py:76); popts must be supplied by the caller
  const float u0[4] = {140.0f, 0.0f, 0.0f, 0.0f}, umin[4] = {90.0f, -0.5f, -0.2f, -0.1f};
  const float umax[4] = {240.0f, 0.5f, 0.2f, 0.1f}, dumax[4] = {5e3f, 10.0f, 10.0f, 10.0f};
  const float Qw[6] = {1.0f, 1.0f, 1.0f, 0.1f, 0.1f, 0.1f};
  memcpy(p->wl_u0, u0, sizeof(u0)); memcpy(p->wl_umin, umin, sizeof(umin)); memcpy(p->wl_umax, umax, sizeof(umax));
  memcpy(p->wl_dumax, dumax, sizeof(dumax)); memcpy(p->wl_Qw, Qw, sizeof(Qw));
  p->wl_controlRate = 1000.0f;
  p->wlqpCheckTermination = 25;
}

static void fill_kernel_params(UmpcBatch* b) {
  const UmpcBatchParams& p = b->bp;
  UmpcParams& m = b->kp.mpc;
  memset(&b->kp, 0, sizeof(b->kp));
  m.dt = p.dt; m.g = p.g; m.Tmax = p.TtoWmax * p.g;
  for (int i = 0; i < 3; ++i) {
    m.Qyr[i] = p.wpr; m.Qyf[i] = p.wpf; m.Qyr[3 + i] = m.Qyf[3 + i] = p.ws;
    m.Qdyr[i] = p.wvr; m.Qdyf[i] = p.wvf; m.Qdyr[3 + i] = m.Qdyf[3 + i] = p.wds;
    m.Ibi[i] = 1.0f / p.Ib[i];
  }
  m.R[0] = p.wthrust; m.R[1] = m.R[2] = p.wmom;
  // generated settings (uprightmpc2/workspace.c:561) + uprightmpc2.c:116-117
  m.max_iter = p.maxIter; m.rho = 0.1f; m.sigma = 1e-6f; m.alpha = 1.6f; m.eps_abs = 1e-4f; m.eps_rel = 1e-4f;
  m.scaling_iters = 10;
  if (const char* dp = getenv("UMPC_DEBUG_PHASE")) m.debug_phase = atoi(dp);  // measurement aid, see umpc_core.cuh
  SimParams& sm = b->kp.sim;
  sm.dtsim = (float)p.dtsim; sm.hl_interval = (float)p.hlInterval; sm.taulim = p.taulim; sm.tpert = p.tpert;
  for (int i = 0; i < 3; ++i) sm.Ib[i] = p.Ib[i];
  sm.g = p.g; sm.mb = p.mb; sm.wl_mode = p.wlMode; sm.int_pos = p.intInitialPos;
  WlParams& w = b->wlp;
  memset(&w, 0, sizeof(w));
  memcpy(w.popts, p.wl_popts, sizeof(w.popts));
  for (int i = 0; i < 4; ++i) {
    w.umin[i] = p.wl_umin[i]; w.umax[i] = p.wl_umax[i];
    w.dumax[i] = p.wl_controlRate > 0 ? p.wl_dumax[i] / p.wl_controlRate : 0.0f;
  }
  for (int i = 0; i < 6; ++i) w.Qw[i] = p.wl_Qw[i];
  // wlqp/workspace.c:84 + osqp_set_default_settings (eps 1e-3, constants.h:62-63) + max_iter 40 (wlqp.c:43-46)
  w.rho = 0.1f; w.sigma = 1e-6f; w.alpha = 1.6f; w.eps_abs = 1e-3f; w.eps_rel = 1e-3f;
  w.max_iter = 40; w.check_termination = p.wlqpCheckTermination; w.scaling_iters = 10;
}

static size_t smem_bytes(int tpb) {
  // UMPC_SMEM_PAD_FLOATS (debug/tuning only): extra floats per thread, to study occupancy sensitivity
  static int pad = -1;
  if (pad < 0) { const char* e = getenv("UMPC_SMEM_PAD_FLOATS"); pad = e ? atoi(e) : 0; }
  return (size_t)(SM_FLOATS_PER_THREAD + pad) * tpb * sizeof(float);
}

extern "C" int umpcBatchSetRolloutInfo(UmpcBatch* b, int on) {
  if (!b) return UMPC_ERR_ARG;
  b->rollout_info = on != 0;
  return UMPC_OK;
}

extern "C" int umpcBatchSetLaunch(UmpcBatch* b, int tpb) {
  // 32, 64: shared-memory scratch (one/two-warp CTAs); 128, 256: tensor-memory scratch (TMEM needs 4-warp multiples)
  if (!b || (tpb != 32 && tpb != 64 && tpb != 128 && tpb != 256)) { g_err = "threads_per_block must be 32, 64 (shared-memory scratch) or 128, 256 (TMEM scratch)"; return UMPC_ERR_ARG; }
  b->tpb = tpb;
  b->tpb_user = true;   // the rollout then uses this shape too (ntiles / (tpb / 32) CTAs) instead of its persistent default
  return UMPC_OK;
}

static int fill_rows(UmpcBatch* b, int row0, int k, float v) {
  fill_rows_kernel<<<(b->Bp + 255) / 256, 256, 0, b->stream>>>(b->slab, b->Bp, row0, k, v);
  CK(cudaGetLastError());
  return UMPC_OK;
}

static int reset_state(UmpcBatch* b) {
  CK(cudaMemsetAsync(b->slab, 0, b->slab_floats * sizeof(float), b->stream));
  // D = E = 1 ; WL u0 ; wlqp D = E = c = 1 ; actualT0 = -1
  int rc;
  if ((rc = fill_rows(b, ROW_D, 45 + 39, 1.0f))) return rc;       // D and E are adjacent
  if ((rc = fill_rows(b, ROW_WLQP + 12, 9, 1.0f))) return rc;
  for (int i = 0; i < 4; ++i)
    if ((rc = fill_rows(b, ROW_WLU0 + i, 1, b->bp.wl_u0[i]))) return rc;
  if ((rc = fill_rows(b, ROW_AT0, 1, -1.0f))) return rc;
  CK(cudaStreamSynchronize(b->stream));
  if (b->conv) CK((cudaError_t)umpc_conv_reset(b->conv, &b->s, b->stream));
  b->step = 0; b->thl = 0.0;
  return UMPC_OK;
}

extern "C" int umpcBatchReset(UmpcBatch* b) {
  if (!b) return UMPC_ERR_ARG;
  CK(cudaSetDevice(b->device));
  return reset_state(b);
}

static int batch_create_body(UmpcBatch* b, int nenv, const UmpcBatchParams* par);
extern "C" void umpcBatchDestroy(UmpcBatch* b);

extern "C" int umpcBatchCreate(UmpcBatch** out, int nenv, int device, const UmpcBatchParams* par) {
  if (!out || nenv <= 0 || !par) { g_err = "umpcBatchCreate: bad argument"; return UMPC_ERR_ARG; }
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    g_err = std::string("umpcBatchCreate: no CUDA device (") + cudaGetErrorString(e) + "); this library has no CPU path";
    return UMPC_ERR_CUDA;
  }
  CK(cudaSetDevice(device));
  UmpcBatch* b = new UmpcBatch();
  b->B = nenv; b->Bp = (nenv + 255) & ~255; b->device = device; b->bp = *par;  // Bp: whole CTAs of the largest launch shape
  const int rc = batch_create_body(b, nenv, par);
  if (rc) { umpcBatchDestroy(b); return rc; }   // a failed create leaves nothing allocated
  *out = b;
  return UMPC_OK;
}

static int batch_create_body(UmpcBatch* b, int nenv, const UmpcBatchParams* par) {
  fill_kernel_params(b);
  CK(cudaStreamCreateWithFlags(&b->stream, cudaStreamNonBlocking));
  CK(cudaEventCreate(&b->ev0));
  CK(cudaEventCreate(&b->ev1));
  // one slab, tiled struct of arrays float[Bp/32][UMPC_ROWS][32] (umpc_dev.cuh)
  const int np = b->Bp;
  b->slab_floats = (size_t)UMPC_ROWS * np;
  CK(cudaMalloc(&b->slab, b->slab_floats * sizeof(float)));
  DevState& s = b->s;
  s.B = np; s.nreal = nenv; s.slab = b->slab;
  auto row = [&](int r) { return b->slab + ROWF(r); };  // host handles: the row block of env 0
  s.p = row(ROW_P); s.R = row(ROW_R); s.dq = row(ROW_DQ); s.uq = row(ROW_UQ); s.acc = row(ROW_ACC);
  s.T0 = row(ROW_T0); s.xs = row(ROW_XS); s.zs = row(ROW_ZS); s.ys = row(ROW_YS); s.D = row(ROW_D); s.E = row(ROW_E);
  s.solx = row(ROW_SOLX); s.soly = row(ROW_SOLY); s.info = row(ROW_INFO);
  s.task = (int*)row(ROW_TASK); s.tp = row(ROW_TP); s.pinit = row(ROW_PINIT); s.actualT0 = row(ROW_AT0); s.metrics = row(ROW_METRICS);
  s.wl_u0 = row(ROW_WLU0); s.wlqp = row(ROW_WLQP);
  s.qpdata = nullptr;
  s.wts = nullptr;
  s.c64 = nullptr;
  if (par->checkTermination > 0) CK((cudaError_t)umpc_conv_create(&b->conv, np, par, &s));
  CK(cudaMalloc(&b->d_in, (size_t)28 * np * sizeof(float)));
  CK(cudaMalloc(&b->d_out, (size_t)10 * np * sizeof(float)));
  CK(cudaMemset(b->d_in, 0, (size_t)28 * np * sizeof(float)));
  CK(cudaMalloc(&b->d_sums, 8 * sizeof(double)));
#define UMPC_SET_SMEM(T)                                                                                                     \
  CK(cudaFuncSetAttribute(umpc_update_kernel<T, SolverF32<T, SmemStore<T>>>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(200 * 1024))); \
  CK(cudaFuncSetAttribute(umpc_rollout_kernel<T, SolverF32<T, SmemStore<T>>>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(200 * 1024)));
  UMPC_SET_SMEM(32) UMPC_SET_SMEM(64)
  // hybrid TMEM + shared-memory scratch of the TMEM launch shapes (HybridStore): opt in to > 48 KB of dynamic shared memory
#define UMPC_SET_HYB(K, T, ...)                                                                                               \
  if (DefStore<T>::kSmemBytes > 48 * 1024)                                                                                      \
    CK(cudaFuncSetAttribute(K<T, SolverF32<T, DefStore<T>, ##__VA_ARGS__>>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DefStore<T>::kSmemBytes));
  UMPC_SET_HYB(umpc_update_kernel, 128) UMPC_SET_HYB(umpc_update_kernel, 256) UMPC_SET_HYB(umpc_update_kernel, 128, true)
  UMPC_SET_HYB(umpc_rollout_kernel, 128) UMPC_SET_HYB(umpc_rollout_kernel, 256) UMPC_SET_HYB(umpc_rollout_kernel, 128, true)
  UMPC_SET_HYB(umpc_hostio_kernel, 128) UMPC_SET_HYB(umpc_hostio_kernel, 128, false, false)
  {
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, b->device));
    b->nsm = prop.multiProcessorCount;
  }
  b->tpb = UMPC_DEFAULT_TPB;
  if (const char* ev = getenv("UMPC_ROLLOUT_INFO")) b->rollout_info = atoi(ev) != 0;
  if (const char* t = getenv("UMPC_TPB")) {  // tuning/testing knob: same values as umpcBatchSetLaunch
    const int v = atoi(t);
    if (v == 32 || v == 64 || v == 128 || v == 256) { b->tpb = v; b->tpb_user = true; }
  }
  return reset_state(b);
}

extern "C" void umpcBatchDestroy(UmpcBatch* b) {
  if (!b) return;
  cudaSetDevice(b->device);
  if (b->stream) cudaStreamSynchronize(b->stream);
  umpc_conv_destroy(b->conv);
  cudaFree(b->s.wts);
  cudaFree(b->slab); cudaFree(b->d_in); cudaFree(b->d_out); cudaFree(b->d_sums); cudaFree(b->d_next);
  cudaFree(b->d_fire); cudaFree(b->d_t); cudaFree(b->lg.y); cudaFree(b->lg.u);
  if (b->h_single) cudaFreeHost(b->h_single); else cudaFree(b->s.qpdata);
  if (b->ev0) cudaEventDestroy(b->ev0);
  if (b->ev1) cudaEventDestroy(b->ev1);
  if (b->x_ev0) { cudaEventDestroy(b->x_ev0); cudaEventDestroy(b->x_ev1); }
  if (b->tev0) { cudaEventDestroy(b->tev0); cudaEventDestroy(b->tev1); }
  if (b->g_exec) cudaGraphExecDestroy(b->g_exec);
  if (b->d_pulled) cudaFree(b->d_pulled);
  for (int i = 0; i < 3; ++i) { if (b->side[i]) cudaStreamDestroy(b->side[i]); if (b->join_ev[i]) cudaEventDestroy(b->join_ev[i]); }
  if (b->side_ev) cudaEventDestroy(b->side_ev);
  if (b->stream) cudaStreamDestroy(b->stream);
  cudaGetLastError();
  delete b;
}

extern "C" int umpcBatchSize(const UmpcBatch* b) { return b ? b->B : 0; }
extern "C" int umpcBatchSync(UmpcBatch* b) {
  if (!b) return UMPC_ERR_ARG;
  CK(cudaStreamSynchronize(b->stream));
  return UMPC_OK;
}
// CUDA events on the batch's own stream bracketing any sequence of calls (bench.py's timed region)
extern "C" int umpcBatchTimerStart(UmpcBatch* b) {
  if (!b) return UMPC_ERR_ARG;
  CK(cudaSetDevice(b->device));
  if (!b->tev0) { CK(cudaEventCreate(&b->tev0)); CK(cudaEventCreate(&b->tev1)); }
  CK(cudaEventRecord(b->tev0, b->stream));
  return UMPC_OK;
}
extern "C" float umpcBatchTimerStopMs(UmpcBatch* b) {
  if (!b || !b->tev0) return -1.0f;
  if (cudaEventRecord(b->tev1, b->stream) != cudaSuccess) return -1.0f;
  if (cudaEventSynchronize(b->tev1) != cudaSuccess) return -1.0f;
  float ms = -1.0f;
  cudaEventElapsedTime(&ms, b->tev0, b->tev1);
  return ms;
}
extern "C" float umpcBatchLastKernelMs(UmpcBatch* b) {
  if (!b || !b->timed) return -1.0f;
  float ms = -1.0f;
  if (cudaEventSynchronize(b->ev1) != cudaSuccess) return -1.0f;
  cudaEventElapsedTime(&ms, b->ev0, b->ev1);
  return ms;
}

static int launch_update(UmpcBatch* b, float* uquad, float* accdes, const float* p0, const float* R0, const float* dq0,
                         const float* pdes, const float* dpdes, const float* sdes, const float* aT0,
                         float* pn = nullptr, float* Rn = nullptr, float* dqn = nullptr, int env0 = 0, int nenv = -1,
                         cudaStream_t stream = nullptr) {
  if (nenv < 0) nenv = b->Bp;        // whole batch (padded env slots included)
  if (!stream) stream = b->stream;
  const int tpb = b->tpb, grid = (nenv + tpb - 1) / tpb;
  if (stream == b->stream) CK(cudaEventRecord(b->ev0, b->stream));
  if (b->conv) {
    CK((cudaError_t)umpc_conv_launch_update(b->conv, &b->kp.sim, &b->s, uquad, accdes, p0, R0, dq0, pdes, dpdes, sdes, aT0,
                                            b->have_qpdata ? 1 : 0, pn, Rn, dqn, env0, nenv, stream));
    if (stream == b->stream) { CK(cudaEventRecord(b->ev1, b->stream)); b->timed = true; }
    return UMPC_OK;
  }
#define UMPC_LAUNCH_UPDATE(T, ST, SMEM)                                                                               \
  case T:                                                                                                             \
    umpc_update_kernel<T, SolverF32<T, ST>><<<grid, T, SMEM, stream>>>(b->kp, b->s, uquad, accdes, p0, R0, dq0, pdes, dpdes, sdes, aT0, \
                                                         b->have_qpdata ? 1 : 0, pn, Rn, dqn, env0);                      \
    break;
  if (b->s.wts) {  // per-env weights: TMEM launch shape only
    umpc_update_kernel<128, SolverF32<128, DefStore<128>, true>><<<(nenv + 127) / 128, 128, DefStore<128>::kSmemBytes, stream>>>(
        b->kp, b->s, uquad, accdes, p0, R0, dq0, pdes, dpdes, sdes, aT0, b->have_qpdata ? 1 : 0, pn, Rn, dqn, env0);
  } else
  switch (tpb) {
    UMPC_LAUNCH_UPDATE(32, SmemStore<32>, smem_bytes(32)) UMPC_LAUNCH_UPDATE(64, SmemStore<64>, smem_bytes(64))
    UMPC_LAUNCH_UPDATE(128, DefStore<128>, DefStore<128>::kSmemBytes) UMPC_LAUNCH_UPDATE(256, DefStore<256>, DefStore<256>::kSmemBytes)
    default: g_err = "unsupported threads_per_block"; return UMPC_ERR_ARG;
  }
  CK(cudaGetLastError());
  if (stream == b->stream) { CK(cudaEventRecord(b->ev1, b->stream)); b->timed = true; }
  return UMPC_OK;
}

// Host-buffer calls: the batch is cut into chunks of whole CTAs and each chunk's H2D copies, kernel and D2H copies
// go to one of three side streams, so that the copy engines (one per direction) and the SMs work on different
// chunks at the same time.  arrays: {host pointer, device staging pointer, floats per env}.
struct HostArr { const float* h_in; float* h_out; float* d; int k; };
// issues the chunk pipeline: fork from the batch's stream, chunks round-robin over the side streams, join back
static int issue_chunks(UmpcBatch* b, HostArr* in, int nin, HostArr* out, int nout, float* duq, float* dacc, const float* const* din,
                        const float* daT, float* dpn, float* dRn, float* ddqn, int nchunks) {
  CK(cudaEventRecord(b->side_ev, b->stream));  // side streams start after whatever is queued on the batch's stream
  const int B = b->B;
  int chunk = (B + nchunks - 1) / nchunks;
  chunk = (chunk + 255) & ~255;
  if (chunk < 4096) chunk = 4096;              // below this a launch does not fill the GPU anyway
  int ci = 0;
  for (int e0 = 0; e0 < B; e0 += chunk, ++ci) {
    cudaStream_t st = b->side[ci % 3];
    if (ci < 3) CK(cudaStreamWaitEvent(st, b->side_ev, 0));
    const int n = e0 + chunk <= B ? chunk : B - e0;
    for (int a = 0; a < nin; ++a)
      if (in[a].h_in) CK(cudaMemcpyAsync(in[a].d + (size_t)e0 * in[a].k, in[a].h_in + (size_t)e0 * in[a].k, (size_t)n * in[a].k * 4, cudaMemcpyHostToDevice, st));
    const int npad = (n + 255) & ~255;  // the last chunk covers the padded env slots of its CTA
    int rc = launch_update(b, duq, dacc, din[0], din[1], din[2], din[3], din[4], din[5], daT, dpn, dRn, ddqn, e0, npad, st);
    if (rc) return rc;
    for (int a = 0; a < nout; ++a)
      CK(cudaMemcpyAsync(out[a].h_out + (size_t)e0 * out[a].k, out[a].d + (size_t)e0 * out[a].k, (size_t)n * out[a].k * 4, cudaMemcpyDeviceToHost, st));
  }
  for (int i = 0; i < 3 && i < ci; ++i) {
    CK(cudaEventRecord(b->join_ev[i], b->side[i]));
    CK(cudaStreamWaitEvent(b->stream, b->join_ev[i], 0));
  }
  return UMPC_OK;
}

static bool host_pinned(const void* p) {
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return at.type == cudaMemoryTypeHost;
}

// When every host buffer is pinned, the whole pipeline (~13 copies + 1 kernel per chunk) is captured ONCE into a CUDA
// graph keyed by the buffer addresses and launch state, and replayed with a single cudaGraphLaunch per call: the
// per-call host cost drops from ~13 driver calls per chunk to one (0.47 -> 0.41 ms per 65,536 env-steps).  Pageable buffers (and UMPC_NO_GRAPH=1) take the direct path.
static int run_chunked(UmpcBatch* b, HostArr* in, int nin, HostArr* out, int nout, float* duq, float* dacc, const float* const* din,
                       const float* daT, float* dpn, float* dRn, float* ddqn) {
  if (!b->side[0]) {
    for (int i = 0; i < 3; ++i) {
      CK(cudaStreamCreateWithFlags(&b->side[i], cudaStreamNonBlocking));
      CK(cudaEventCreateWithFlags(&b->join_ev[i], cudaEventDisableTiming));
    }
    CK(cudaEventCreateWithFlags(&b->side_ev, cudaEventDisableTiming));
  }
  // UMPC_E2E_CHUNKS: tuning knob.  Measured on B200 + PCIe gen5 per 65,536 env-steps under the graph: 1 chunk 0.53 ms,
  // 2 -> 0.41, 3 -> 0.52, 4 -> 0.50, 6 -> 0.58 (each chunk is 12 small DMA transfers: the copy engines, not the host
  // calls, limit finer pipelining); direct path: 1 -> 0.63, 2 -> 0.47-0.54
  static int nchunks_env = -1, no_graph = -1;
  if (nchunks_env < 0) { const char* ev = getenv("UMPC_E2E_CHUNKS"); nchunks_env = ev ? atoi(ev) : 0; if (nchunks_env < 0) nchunks_env = 0; }
  if (no_graph < 0) { const char* ev = getenv("UMPC_NO_GRAPH"); no_graph = (ev && atoi(ev)) ? 1 : 0; }
  bool pinned = true;
  for (int a = 0; a < nin && pinned; ++a) if (in[a].h_in) pinned = host_pinned(in[a].h_in);
  for (int a = 0; a < nout && pinned; ++a) pinned = host_pinned(out[a].h_out);
  // Zero-copy path (default for pinned buffers; UMPC_E2E_ZEROCOPY=0 falls back to the copy pipeline below): ONE launch of
  // umpc_hostio_kernel on the caller's own arrays.  Needs the default launch shape (128 threads, TMEM scratch, shared
  // weights, float32 path) and 16-byte aligned arrays.
  const char* zc_ev = getenv("UMPC_E2E_ZEROCOPY");  // read per call: the tests switch paths inside one process
  const int zc_env = (zc_ev && !atoi(zc_ev)) ? 0 : 1;
  if (pinned && zc_env && b->tpb == 128 && !b->s.wts && !b->conv && nin == 7 && (nout == 2 || nout == 5)) {
    HostIO io;
    memset(&io, 0, sizeof io);
    bool ok = true;
    for (int a = 0; a < 6 && ok; ++a) {
      void* d = nullptr;
      ok = cudaHostGetDevicePointer(&d, (void*)in[a].h_in, 0) == cudaSuccess && ((uintptr_t)d & 15) == 0;
      io.in[a] = (const float*)d;
    }
    if (ok && in[6].h_in) {
      void* d = nullptr;
      ok = cudaHostGetDevicePointer(&d, (void*)in[6].h_in, 0) == cudaSuccess;
      io.aT0 = (const float*)d;
    }
    for (int a = 0; a < nout && ok; ++a) {
      void* d = nullptr;
      ok = cudaHostGetDevicePointer(&d, (void*)out[a].h_out, 0) == cudaSuccess && ((uintptr_t)d & 15) == 0;
      io.out[a] = (float*)d;
    }
    if (!ok) cudaGetLastError();
    else {
      io.nenv = b->B;
      const char* wv = getenv("UMPC_E2E_WINDOW");   // tuning knob; 0 = unordered pulls
      io.window = wv ? atoi(wv) : 32;
      if (io.window < 0) io.window = 0;
      if (!b->d_pulled) CK(cudaMalloc(&b->d_pulled, sizeof(unsigned)));
      io.pulled = b->d_pulled;
      CK(cudaMemsetAsync(b->d_pulled, 0, sizeof(unsigned), b->stream));
      CK(cudaEventRecord(b->ev0, b->stream));
      // an env-step (5 outputs: the plant moves on) is a step of a closed loop: like a rollout it skips the per-solve
      // residual / status / solution block unless umpcBatchSetRolloutInfo(b, 1) asked for it; an update (2 outputs) keeps it
      if (nout == 5 && !b->rollout_info)
        umpc_hostio_kernel<128, SolverF32<128, DefStore<128>, false, false>><<<b->Bp / 128, 128, DefStore<128>::kSmemBytes, b->stream>>>(b->kp, b->s, io, b->have_qpdata ? 1 : 0);
      else
        umpc_hostio_kernel<128, SolverF32<128, DefStore<128>>><<<b->Bp / 128, 128, DefStore<128>::kSmemBytes, b->stream>>>(b->kp, b->s, io, b->have_qpdata ? 1 : 0);
      CK(cudaGetLastError());
      CK(cudaEventRecord(b->ev1, b->stream));
      b->timed = true;
      CK(cudaStreamSynchronize(b->stream));
      return UMPC_OK;
    }
  }
  if (no_graph) pinned = false;
  if (!pinned) {
    int rc = issue_chunks(b, in, nin, out, nout, duq, dacc, din, daT, dpn, dRn, ddqn, nchunks_env ? nchunks_env : 2);
    if (rc) return rc;
    CK(cudaStreamSynchronize(b->stream));
    return UMPC_OK;
  }
  const int nchunks = nchunks_env ? nchunks_env : 2;
  std::vector<const void*> key;
  for (int a = 0; a < nin; ++a) key.push_back(in[a].h_in);
  for (int a = 0; a < nout; ++a) key.push_back(out[a].h_out);
  key.push_back(dpn); key.push_back(daT); key.push_back(b->s.wts); key.push_back(b->s.qpdata); key.push_back(b->conv);
  key.push_back((const void*)(size_t)(b->tpb * 64 + nchunks * 2 + (b->have_qpdata ? 1 : 0)));
  if (!b->g_exec || key != b->g_key) {
    if (b->g_exec) { cudaGraphExecDestroy(b->g_exec); b->g_exec = nullptr; }
    CK(cudaStreamSynchronize(b->stream));
    CK(cudaStreamBeginCapture(b->stream, cudaStreamCaptureModeThreadLocal));
    int rc = issue_chunks(b, in, nin, out, nout, duq, dacc, din, daT, dpn, dRn, ddqn, nchunks);
    cudaGraph_t g = nullptr;
    cudaError_t e = cudaStreamEndCapture(b->stream, &g);
    if (rc) { if (g) cudaGraphDestroy(g); return rc; }
    if (e != cudaSuccess) return set_err("cudaStreamEndCapture", e);
    e = cudaGraphInstantiate(&b->g_exec, g, 0);
    cudaGraphDestroy(g);
    if (e != cudaSuccess) { b->g_exec = nullptr; return set_err("cudaGraphInstantiate", e); }
    b->g_key = key;
  }
  CK(cudaGraphLaunch(b->g_exec, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  return UMPC_OK;
}

extern "C" int umpcBatchUpdateDev(UmpcBatch* b, float* uquad, float* accdes, const float* p0, const float* R0,
                                  const float* dq0, const float* pdes, const float* dpdes, const float* sdes,
                                  const float* actualT0) {
  if (!b) return UMPC_ERR_ARG;
  CK(cudaSetDevice(b->device));
  return launch_update(b, uquad, accdes, p0, R0, dq0, pdes, dpdes, sdes, actualT0);
}

// The same launch ORDERED AGAINST THE CALLER'S STREAM (e.g. torch.cuda.current_stream().cuda_stream): the batch's
// private stream first waits for everything queued on `caller_stream` (so inputs written by preceding work there are
// complete), and `caller_stream` then waits for the update (so later work there sees the outputs).  No host sync.
extern "C" int umpcBatchUpdateDevOn(UmpcBatch* b, void* caller_stream, float* uquad, float* accdes, const float* p0,
                                    const float* R0, const float* dq0, const float* pdes, const float* dpdes,
                                    const float* sdes, const float* actualT0) {
  if (!b) return UMPC_ERR_ARG;
  CK(cudaSetDevice(b->device));
  cudaStream_t cs = (cudaStream_t)caller_stream;
  if (!b->x_ev0) {
    CK(cudaEventCreateWithFlags(&b->x_ev0, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&b->x_ev1, cudaEventDisableTiming));
  }
  CK(cudaEventRecord(b->x_ev0, cs));
  CK(cudaStreamWaitEvent(b->stream, b->x_ev0, 0));
  int rc = launch_update(b, uquad, accdes, p0, R0, dq0, pdes, dpdes, sdes, actualT0);
  if (rc) return rc;
  CK(cudaEventRecord(b->x_ev1, b->stream));
  CK(cudaStreamWaitEvent(cs, b->x_ev1, 0));
  return UMPC_OK;
}

extern "C" int umpcBatchUpdate(UmpcBatch* b, float* uquad, float* accdes, const float* p0, const float* R0,
                               const float* dq0, const float* pdes, const float* dpdes, const float* sdes,
                               const float* actualT0) {
  if (!b || !uquad || !accdes || !p0 || !R0 || !dq0 || !pdes || !dpdes || !sdes) { g_err = "umpcBatchUpdate: NULL argument"; return UMPC_ERR_ARG; }
  CK(cudaSetDevice(b->device));
  const size_t P = b->Bp;
  float* d = b->d_in;
  float *dp = d, *dR = dp + 3 * P, *ddq = dR + 9 * P, *dpd = ddq + 6 * P, *ddp = dpd + 3 * P, *dsd = ddp + 3 * P, *daT = dsd + 3 * P;
  float *duq = b->d_out, *dacc = duq + 3 * P;
  HostArr in[7] = {{p0, nullptr, dp, 3}, {R0, nullptr, dR, 9}, {dq0, nullptr, ddq, 6}, {pdes, nullptr, dpd, 3},
                   {dpdes, nullptr, ddp, 3}, {sdes, nullptr, dsd, 3}, {actualT0, nullptr, daT, 1}};
  HostArr out[2] = {{nullptr, uquad, duq, 3}, {nullptr, accdes, dacc, 6}};
  const float* din[6] = {dp, dR, ddq, dpd, ddp, dsd};
  return run_chunked(b, in, 7, out, 2, duq, dacc, din, actualT0 ? daT : nullptr, nullptr, nullptr, nullptr);
}

// umpcUpdate + torque clip + quadrotorNLDyn for B environments with HOST buffers: the batched twin of
//   uquad, accdes = mdl.update(p, Rb, dq, pdes, dpdes, sdes); p, Rb, dq = quadrotorNLDyn(p, Rb, dq, uquad, dtsim)
// (uprightmpc2.py:139-151).  p_next/R_next/dq_next may alias p0/R0/dq0.
extern "C" int umpcBatchEnvStep(UmpcBatch* b, float* uquad, float* accdes, float* p_next, float* R_next, float* dq_next,
                                const float* p0, const float* R0, const float* dq0, const float* pdes, const float* dpdes,
                                const float* sdes, const float* actualT0) {
  if (!b || !uquad || !accdes || !p_next || !R_next || !dq_next || !p0 || !R0 || !dq0 || !pdes || !dpdes || !sdes) {
    g_err = "umpcBatchEnvStep: NULL argument";
    return UMPC_ERR_ARG;
  }
  CK(cudaSetDevice(b->device));
  const size_t P = b->Bp;
  if (!b->d_next) CK(cudaMalloc(&b->d_next, (size_t)18 * P * sizeof(float)));
  float* d = b->d_in;
  float *dp = d, *dR = dp + 3 * P, *ddq = dR + 9 * P, *dpd = ddq + 6 * P, *ddp = dpd + 3 * P, *dsd = ddp + 3 * P, *daT = dsd + 3 * P;
  float *duq = b->d_out, *dacc = duq + 3 * P;
  float *dpn = b->d_next, *dRn = dpn + 3 * P, *ddqn = dRn + 9 * P;
  HostArr in[7] = {{p0, nullptr, dp, 3}, {R0, nullptr, dR, 9}, {dq0, nullptr, ddq, 6}, {pdes, nullptr, dpd, 3},
                   {dpdes, nullptr, ddp, 3}, {sdes, nullptr, dsd, 3}, {actualT0, nullptr, daT, 1}};
  HostArr out[5] = {{nullptr, uquad, duq, 3}, {nullptr, accdes, dacc, 6}, {nullptr, p_next, dpn, 3}, {nullptr, R_next, dRn, 9},
                    {nullptr, dq_next, ddqn, 6}};
  const float* din[6] = {dp, dR, ddq, dpd, ddp, dsd};
  return run_chunked(b, in, 7, out, 5, duq, dacc, din, actualT0 ? daT : nullptr, dpn, dRn, ddqn);
}

// tiled SoA on device <-> [B][k] on host (diagnostics / state exchange; not on the hot path).  `dev` is a host handle
// of DevState (row block of env 0): the k rows of every tile are gathered with one strided 2-D copy.
static int get_rows(UmpcBatch* b, const void* dev, int k, void* host_) {
  if (!host_) return UMPC_OK;
  float* host = (float*)host_;
  const size_t B = b->B, nt = b->Bp / 32;
  std::vector<float> tmp(nt * k * 32);
  CK(cudaMemcpy2DAsync(tmp.data(), (size_t)k * 128, dev, (size_t)UMPC_ROWS * 128, (size_t)k * 128, nt, cudaMemcpyDeviceToHost, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  for (size_t e = 0; e < B; ++e)
    for (int i = 0; i < k; ++i) host[e * k + i] = tmp[((e >> 5) * k + i) * 32 + (e & 31)];
  return UMPC_OK;
}
static int put_rows(UmpcBatch* b, void* dev, int k, const void* host_) {
  const float* host = (const float*)host_;
  const size_t B = b->B, nt = b->Bp / 32;
  std::vector<float> tmp(nt * k * 32, 0.0f);
  for (size_t e = 0; e < B; ++e)
    for (int i = 0; i < k; ++i) tmp[((e >> 5) * k + i) * 32 + (e & 31)] = host[e * k + i];
  CK(cudaMemcpy2DAsync(dev, (size_t)UMPC_ROWS * 128, tmp.data(), (size_t)k * 128, (size_t)k * 128, nt, cudaMemcpyHostToDevice, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  return UMPC_OK;
}
// plain [k][Bp] rows (the optional QP-data dump)
static int get_rows_soa(UmpcBatch* b, const float* dev, int k, float* host) {
  if (!host) return UMPC_OK;
  const size_t B = b->B, P = b->Bp;
  std::vector<float> tmp(P * k);
  CK(cudaMemcpyAsync(tmp.data(), dev, P * k * 4, cudaMemcpyDeviceToHost, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  for (size_t e = 0; e < B; ++e)
    for (int i = 0; i < k; ++i) host[e * k + i] = tmp[(size_t)i * P + e];
  return UMPC_OK;
}

extern "C" int umpcBatchGetSolve(UmpcBatch* b, float* info, float* x, float* y, float* xs, float* zs, float* ys, float* D,
                                 float* E) {
  if (!b) return UMPC_ERR_ARG;
  CK(cudaSetDevice(b->device));
  int rc;
  if ((rc = get_rows(b, b->s.info, 4, info))) return rc;
  if ((rc = get_rows(b, b->s.solx, 45, x))) return rc;
  if ((rc = get_rows(b, b->s.soly, 39, y))) return rc;
  if ((rc = get_rows(b, b->s.xs, 45, xs))) return rc;
  if ((rc = get_rows(b, b->s.zs, 39, zs))) return rc;
  if ((rc = get_rows(b, b->s.ys, 39, ys))) return rc;
  if ((rc = get_rows(b, b->s.D, 45, D))) return rc;
  return get_rows(b, b->s.E, 39, E);
}

extern "C" int umpcBatchGetQPData(UmpcBatch* b, float* qp) {
  if (!b || !qp) return UMPC_ERR_ARG;
  CK(cudaSetDevice(b->device));
  if (!b->have_qpdata) {
    // first request: allocate; data is produced by the NEXT update
    CK(cudaMalloc(&b->s.qpdata, (size_t)216 * b->Bp * sizeof(float)));
    CK(cudaMemset(b->s.qpdata, 0, (size_t)216 * b->Bp * sizeof(float)));
    b->have_qpdata = true;
  }
  return get_rows_soa(b, b->s.qpdata, 216, qp);
}

extern "C" int umpcBatchSetState(UmpcBatch* b, const float* p, const float* R, const float* dq, const int* task,
                                 const float* tp) {
  if (!b || !p || !R || !dq) return UMPC_ERR_ARG;
  CK(cudaSetDevice(b->device));
  int rc;
  if ((rc = put_rows(b, b->s.p, 3, p))) return rc;
  if ((rc = put_rows(b, b->s.pinit, 3, p))) return rc;
  if ((rc = put_rows(b, b->s.R, 9, R))) return rc;
  if ((rc = put_rows(b, b->s.dq, 6, dq))) return rc;
  if (task) { if ((rc = put_rows(b, b->s.task, 1, task))) return rc; }  // 4-byte ints travel as bit patterns
  else if ((rc = fill_rows(b, ROW_TASK, 1, 0.0f))) return rc;
  if (tp) { if ((rc = put_rows(b, b->s.tp, 4, tp))) return rc; }
  else if ((rc = fill_rows(b, ROW_TP, 4, 0.0f))) return rc;
  if ((rc = fill_rows(b, ROW_UQ, 3, 0.0f))) return rc;
  if ((rc = fill_rows(b, ROW_METRICS, 7, 0.0f))) return rc;
  CK(cudaStreamSynchronize(b->stream));
  b->step = 0; b->thl = 0.0;
  return UMPC_OK;
}

// per-environment cost weights [B][8] = (ws, wds, wpr, wpf, wvr, wvf, wthrust, wmom) per env; NULL: back to the
// batch-wide parameters.  What the reference's weight sweep varies per run (uprightmpc2.py:272-303).
extern "C" int umpcBatchSetWeights(UmpcBatch* b, const float* w) {
  if (!b) return UMPC_ERR_ARG;
  CK(cudaSetDevice(b->device));
  CK(cudaStreamSynchronize(b->stream));
  if (!w) { cudaFree(b->s.wts); b->s.wts = nullptr; return UMPC_OK; }
  if (!b->s.wts) CK(cudaMalloc(&b->s.wts, (size_t)8 * b->Bp * sizeof(float)));
  // padded env slots get the batch-wide weights (they run on zero state and are never read back)
  const UmpcBatchParams& p = b->bp;
  const float def[8] = {p.ws, p.wds, p.wpr, p.wpf, p.wvr, p.wvf, p.wthrust, p.wmom};
  const size_t B = b->B, P = b->Bp;
  std::vector<float> tmp(P * 8);
  for (size_t e = 0; e < P; ++e)
    for (int i = 0; i < 8; ++i) tmp[(size_t)i * P + e] = e < B ? w[e * 8 + i] : def[i];
  CK(cudaMemcpy(b->s.wts, tmp.data(), P * 8 * sizeof(float), cudaMemcpyHostToDevice));
  return UMPC_OK;
}

extern "C" int umpcBatchGetState(UmpcBatch* b, float* p, float* R, float* dq, float* uquad) {
  if (!b) return UMPC_ERR_ARG;
  CK(cudaSetDevice(b->device));
  int rc;
  if ((rc = get_rows(b, b->s.p, 3, p))) return rc;
  if ((rc = get_rows(b, b->s.R, 9, R))) return rc;
  if ((rc = get_rows(b, b->s.dq, 6, dq))) return rc;
  return get_rows(b, b->s.uq, 3, uquad);
}

extern "C" int umpcBatchRollout(UmpcBatch* b, int nsteps) {
  if (!b || nsteps <= 0) return UMPC_ERR_ARG;
  CK(cudaSetDevice(b->device));
  if (nsteps > b->sched_cap) {
    // wait for any launch still reading the old schedule buffers
    CK(cudaStreamSynchronize(b->stream));
    cudaFree(b->d_fire); cudaFree(b->d_t);
    b->sched_cap = nsteps < 4096 ? 4096 : nsteps;
    CK(cudaMalloc(&b->d_fire, b->sched_cap));
    CK(cudaMalloc(&b->d_t, (size_t)b->sched_cap * sizeof(double)));
  } else {
    CK(cudaStreamSynchronize(b->stream));  // host staging vectors below are reused per call
  }
  // firing schedule in double, like numpy: t = ti*dtsim ; fire when hlInterval is None or t - thlPrev > hlInterval
  std::vector<uint8_t> fire(nsteps);
  std::vector<double> tv(nsteps);
  const double dts = b->bp.dtsim, hl = b->bp.hlInterval;
  for (int k = 0; k < nsteps; ++k) {
    const double t = (double)(b->step + k) * dts;
    const bool f = hl < 0.0 || (t - b->thl > hl);
    if (f) b->thl = t;
    fire[k] = f ? 1 : 0;
    tv[k] = t;
  }
  CK(cudaMemcpyAsync(b->d_fire, fire.data(), nsteps, cudaMemcpyHostToDevice, b->stream));
  CK(cudaMemcpyAsync(b->d_t, tv.data(), (size_t)nsteps * sizeof(double), cudaMemcpyHostToDevice, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  const int tpb = b->tpb, grid = (b->Bp + tpb - 1) / tpb;
  CK(cudaEventRecord(b->ev0, b->stream));
  if (b->conv) {
    CK((cudaError_t)umpc_conv_launch_rollout(b->conv, &b->kp.sim, &b->s, &b->wlp, (int)b->step, nsteps, b->d_fire, b->d_t, b->lg,
                                             b->stream));
    CK(cudaEventRecord(b->ev1, b->stream));
    b->timed = true;
    b->step += nsteps;
    return UMPC_OK;
  }
#define UMPC_LAUNCH_ROLLOUT(T, ST, SMEM)                                                                              \
  case T:                                                                                                             \
    umpc_rollout_kernel<T, SolverF32<T, ST>><<<grid, T, SMEM, b->stream>>>(b->kp, b->s, b->wlp, (int)b->step, nsteps, b->d_fire, b->d_t, \
                                                             b->lg, 1);                                                 \
    break;
  if (b->s.wts) {
    umpc_rollout_kernel<128, SolverF32<128, DefStore<128>, true>><<<(b->Bp + 127) / 128, 128, DefStore<128>::kSmemBytes, b->stream>>>(
        b->kp, b->s, b->wlp, (int)b->step, nsteps, b->d_fire, b->d_t, b->lg, 1);
  } else if (!b->tpb_user) {
    // default rollout shape: ONE 8-warp CTA per SM (all 512 TMEM columns), warp-tiles dealt statically (umpc_dev.cuh)
    const int ntiles = b->Bp / 32, pgrid = ntiles < b->nsm ? ntiles : b->nsm;
    // steps cut into chunks handed out over the 8 warps (umpc_dev.cuh) when the tiles do not fill the warp slots evenly:
    // the smallest chunk count that makes (tiles per CTA) x chunks a multiple of 8
    static int chunk_env = -1;
    if (chunk_env < 0) { const char* ev = getenv("UMPC_ROLLOUT_CHUNKS"); chunk_env = ev ? atoi(ev) : 0; }
    const int tiles_cta = (ntiles + pgrid - 1) / pgrid;
    int nch = tiles_cta % 8 == 0 ? 1 : (tiles_cta % 4 == 0 ? 2 : (tiles_cta % 2 == 0 ? 4 : 8));
    if (chunk_env > 0) nch = chunk_env;   // tuning knob
    if (nsteps < 2 * nch || tiles_cta > UMPC_MAX_CHUNKED_TILES) nch = 1;
    // Solve info (residuals, status) and the un-scaled solution rows are NOT produced per rollout step unless asked for
    // (umpcBatchSetRolloutInfo): nothing on this path reads them -- the reference never acts on them either
    // (check_termination = 0, uprightmpc2.c:116-117) -- and the once-per-solve code is instruction-fetch bound.
    if (b->rollout_info)
      umpc_rollout_kernel<256, SolverF32<256, DefStore<256>>><<<pgrid, 256, DefStore<256>::kSmemBytes, b->stream>>>(
          b->kp, b->s, b->wlp, (int)b->step, nsteps, b->d_fire, b->d_t, b->lg, nch);
    else
      umpc_rollout_kernel<256, SolverF32<256, DefStore<256>, false, false>><<<pgrid, 256, DefStore<256>::kSmemBytes, b->stream>>>(
          b->kp, b->s, b->wlp, (int)b->step, nsteps, b->d_fire, b->d_t, b->lg, nch);
  } else
  switch (tpb) {
    UMPC_LAUNCH_ROLLOUT(32, SmemStore<32>, smem_bytes(32)) UMPC_LAUNCH_ROLLOUT(64, SmemStore<64>, smem_bytes(64))
    UMPC_LAUNCH_ROLLOUT(128, DefStore<128>, DefStore<128>::kSmemBytes) UMPC_LAUNCH_ROLLOUT(256, DefStore<256>, DefStore<256>::kSmemBytes)
    default: g_err = "unsupported threads_per_block"; return UMPC_ERR_ARG;
  }
  CK(cudaGetLastError());
  CK(cudaEventRecord(b->ev1, b->stream));
  b->timed = true;
  b->step += nsteps;
  return UMPC_OK;
}

extern "C" int umpcBatchGetMetrics(UmpcBatch* b, float* metrics) {
  if (!b || !metrics) return UMPC_ERR_ARG;
  CK(cudaSetDevice(b->device));
  int rc = get_rows(b, b->s.metrics, 4, metrics);
  if (rc) return rc;
  // the solve / step counters are int32 on the device (exact beyond 2^24 steps); this call reports them as floats
  for (int e = 0; e < b->B; ++e)
    for (int k = 2; k < 4; ++k) { int v; memcpy(&v, &metrics[(size_t)e * 4 + k], 4); metrics[(size_t)e * 4 + k] = (float)v; }
  return UMPC_OK;
}

extern "C" int umpcBatchReduceMetrics(UmpcBatch* b, double* sums8) {
  if (!b || !sums8) return UMPC_ERR_ARG;
  CK(cudaSetDevice(b->device));
  CK(cudaMemsetAsync(b->d_sums, 0, 8 * sizeof(double), b->stream));
  int grid = (b->B + 255) / 256;
  if (grid > 148 * 4) grid = 148 * 4;
  reduce_metrics_kernel<<<grid, 256, 0, b->stream>>>(b->s, b->d_sums);
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(sums8, b->d_sums, 8 * sizeof(double), cudaMemcpyDeviceToHost, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  return UMPC_OK;
}

extern "C" int umpcBatchSetLog(UmpcBatch* b, int env, int cap) {
  if (!b || env >= b->B) return UMPC_ERR_ARG;
  CK(cudaSetDevice(b->device));
  CK(cudaStreamSynchronize(b->stream));
  cudaFree(b->lg.y); cudaFree(b->lg.u);
  b->lg = LogBuf{-1, 0, nullptr, nullptr};
  if (env < 0 || cap <= 0) return UMPC_OK;
  CK(cudaMalloc(&b->lg.y, (size_t)cap * 12 * sizeof(float)));
  CK(cudaMalloc(&b->lg.u, (size_t)cap * 3 * sizeof(float)));
  CK(cudaMemset(b->lg.y, 0, (size_t)cap * 12 * sizeof(float)));
  CK(cudaMemset(b->lg.u, 0, (size_t)cap * 3 * sizeof(float)));
  b->lg.env = env; b->lg.cap = cap;
  return UMPC_OK;
}

extern "C" int umpcBatchGetLog(UmpcBatch* b, float* y, float* u, int nsteps) {
  if (!b || b->lg.env < 0 || nsteps > b->lg.cap) return UMPC_ERR_ARG;
  CK(cudaSetDevice(b->device));
  CK(cudaStreamSynchronize(b->stream));
  if (y) CK(cudaMemcpy(y, b->lg.y, (size_t)nsteps * 12 * sizeof(float), cudaMemcpyDeviceToHost));
  if (u) CK(cudaMemcpy(u, b->lg.u, (size_t)nsteps * 3 * sizeof(float), cudaMemcpyDeviceToHost));
  return UMPC_OK;
}

extern "C" void* umpcBatchDevPtr(UmpcBatch* b, const char* name) {
  if (!b || !name) return nullptr;
  const DevState& s = b->s;
  struct { const char* n; void* p; } tab[] = {
      {"p", s.p}, {"R", s.R}, {"dq", s.dq}, {"uq", s.uq}, {"acc", s.acc}, {"T0", s.T0}, {"xs", s.xs}, {"zs", s.zs},
      {"ys", s.ys}, {"D", s.D}, {"E", s.E}, {"solx", s.solx}, {"soly", s.soly}, {"info", s.info}, {"task", s.task},
      {"tp", s.tp}, {"pinit", s.pinit}, {"actualT0", s.actualT0}, {"metrics", s.metrics}, {"wl_u0", s.wl_u0},
      {"wlqp", s.wlqp}, {"slab", s.slab}, {"stage_in", b->d_in}, {"stage_out", b->d_out}, {"stream", (void*)b->stream}};
  for (auto& t : tab)
    if (!strcmp(t.n, name)) return t.p;
  return nullptr;
}

extern "C" double umpcMeasureFp32Peak(int device, int iters) {
  if (cudaSetDevice(device) != cudaSuccess) return -1.0;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return -1.0;
  float* d = nullptr;
  if (cudaMalloc(&d, 4) != cudaSuccess) return -1.0;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int grid = prop.multiProcessorCount * 8, tpb = 256;
  fp32_peak_kernel<<<grid, tpb>>>(d, 16);  // warm-up
  double best = 0.0;
  for (int rep = 0; rep < 5; ++rep) {
    cudaEventRecord(e0);
    fp32_peak_kernel<<<grid, tpb>>>(d, iters);
    cudaEventRecord(e1);
    if (cudaEventSynchronize(e1) != cudaSuccess) { best = -1.0; break; }
    float ms = 0.0f;
    cudaEventElapsedTime(&ms, e0, e1);
    const double flop = 2.0 * 16 * 8 * (double)iters * grid * tpb;
    const double tf = flop / (ms * 1e-3) * 1e-12;
    if (tf > best) best = tf;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d);
  return best;
}

extern "C" double umpcMeasureFp64Peak(int device, int iters) {
  if (cudaSetDevice(device) != cudaSuccess) return -1.0;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return -1.0;
  double* d = nullptr;
  if (cudaMalloc(&d, 8) != cudaSuccess) return -1.0;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int grid = prop.multiProcessorCount * 8, tpb = 256;
  fp64_peak_kernel<<<grid, tpb>>>(d, 16);
  double best = 0.0;
  for (int rep = 0; rep < 3; ++rep) {
    cudaEventRecord(e0);
    fp64_peak_kernel<<<grid, tpb>>>(d, iters);
    cudaEventRecord(e1);
    if (cudaEventSynchronize(e1) != cudaSuccess) { best = -1.0; break; }
    float ms = 0.0f;
    cudaEventElapsedTime(&ms, e0, e1);
    const double tf = 2.0 * 8 * 8 * (double)iters * grid * tpb / (ms * 1e-3) * 1e-12;
    if (tf > best) best = tf;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d);
  return best;
}

// ======================= reference-compatible single-controller API ==========================
// The reference keeps ONE process-global OSQP workspace (workspace.c:2620); here each
// UprightMPC_t owns a batch of 1 on the current device, found by the struct's address.
static std::mutex g_mu;
static std::map<const void*, UmpcBatch*> g_single;

static UmpcBatch* find_single(const void* key) {
  std::lock_guard<std::mutex> lk(g_mu);
  auto it = g_single.find(key);
  return it == g_single.end() ? nullptr : it->second;
}

extern "C" void umpcFree(UprightMPC_t* up) {
  UmpcBatch* b = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_single.find(up);
    if (it != g_single.end()) { b = it->second; g_single.erase(it); }
  }
  umpcBatchDestroy(b);
}

// CSC position of entry (row, col) of A (pattern of initConstraint, template_controllers.py:28-63),
// counted column by column.  Used to fill the public Ax_idx field (uprightmpc2.h:38).
static int a_has(int r, int c) {
  const int N = UMPC_N, ny = UMPC_NY, nu = UMPC_NU;
  for (int k = 0; k < N; ++k) {
    for (int i = 0; i < ny; ++i) {
      const int rk = k * ny + i, rd = N * ny + k * ny + i;
      if (r == rk && (c == k * ny + i || c == N * ny + k * ny + i || (k > 0 && c == (k - 1) * ny + i))) return 1;
      if (r == rd) {
        if (c == N * ny + k * ny + i || (k > 0 && c == N * ny + (k - 1) * ny + i)) return 1;
        if (i < 3 && c == 2 * N * ny + k * nu) return 1;
        if (i >= 3 && (c == 2 * N * ny + k * nu + 1 || c == 2 * N * ny + k * nu + 2)) return 1;
        if (k > 1 && i < 3 && c == (k - 2) * ny + 3 + i) return 1;
      }
    }
    if (r == 2 * N * ny + k && c == 2 * N * ny + nu * k) return 1;
  }
  return 0;
}
static int a_csc_index(int r, int c) {
  int n = 0;
  for (int cc = 0; cc < UMPC_NX; ++cc)
    for (int rr = 0; rr < UMPC_NC; ++rr) {
      if (rr == r && cc == c) return a_has(r, c) ? n : -1;
      n += a_has(rr, cc);
    }
  return -1;
}

extern "C" void umpcInit(UprightMPC_t* up, float dt, float g, float TtoWmax, float ws, float wds, float wpr, float wpf,
                         float wvr, float wvf, float wthrust, float wmom, const float Ib[3], int maxIter) {
  umpcFree(up);
  memset(up, 0, sizeof(*up));
  up->dt = dt; up->g = g; up->Tmax = TtoWmax * g;
  for (int i = 0; i < 3; ++i) {
    up->Qyr[i] = wpr; up->Qyf[i] = wpf; up->Qyr[3 + i] = up->Qyf[3 + i] = ws;
    up->Qdyr[i] = wvr; up->Qdyf[i] = wvf; up->Qdyr[3 + i] = up->Qdyf[3 + i] = wds;
  }
  up->R[0] = wthrust; up->R[1] = up->R[2] = wmom;
  up->c0[2] = -g;
  up->e3h[1] = 1.0f; up->e3h[3] = -1.0f;            // skew(e3), column-major
  up->e3hIbi[1] = 1.0f / Ib[0]; up->e3hIbi[3] = -1.0f / Ib[1];  // e3h * diag(1/Ib)
  // Ax_idx: positions of the time-varying entries in A.x, in Ax_data order (uprightmpc2.c:160-179)
  int o = 0;
  const int N = UMPC_N, ny = UMPC_NY, nu = UMPC_NU;
  for (int k = 2; k < N; ++k)
    for (int i = 0; i < 3; ++i) up->Ax_idx[o++] = a_csc_index(N * ny + k * ny + i, (k - 2) * ny + 3 + i);
  up->nAxT0dt = o;
  for (int k = 0; k < N; ++k)
    for (int i = 0; i < ny; ++i) up->Ax_idx[o++] = a_csc_index(k * ny + i, N * ny + k * ny + i);
  up->nAxdt = o;
  for (int k = 0; k < N; ++k)
    for (int i = 0; i < 3; ++i) up->Ax_idx[o++] = a_csc_index(N * ny + k * ny + i, 2 * N * ny + k * nu);
  for (int k = 0; k < N; ++k)
    for (int j = 0; j < 2; ++j)
      for (int i = 0; i < 3; ++i) up->Ax_idx[o++] = a_csc_index(N * ny + k * ny + 3 + i, 2 * N * ny + k * nu + 1 + j);
  UmpcBatchParams bp;
  umpcBatchDefaults(&bp);
  bp.dt = dt; bp.g = g; bp.TtoWmax = TtoWmax; bp.ws = ws; bp.wds = wds; bp.wpr = wpr; bp.wpf = wpf; bp.wvr = wvr;
  bp.wvf = wvf; bp.wthrust = wthrust; bp.wmom = wmom; bp.maxIter = maxIter;
  for (int i = 0; i < 3; ++i) bp.Ib[i] = Ib[i];
  int dev = 0;
  cudaGetDevice(&dev);
  UmpcBatch* b = nullptr;
  if (umpcBatchCreate(&b, 1, dev, &bp) != UMPC_OK) {
    fprintf(stderr, "umpcInit (libumpc_b200): %s\n", umpcLastError());
    return;
  }
  // pinned + mapped I/O block: the update kernel reads the 28 inputs from it and writes uquad/accdes and the public QP
  // vectors (UprightMPC_t.l/u/q/Px_data/Ax_data, uprightmpc2.h:36-40) straight into it
  const size_t nfl = 28 + 9 + (size_t)216 * b->Bp;
  if (cudaHostAlloc(&b->h_single, nfl * sizeof(float), cudaHostAllocMapped) != cudaSuccess ||
      cudaHostGetDevicePointer((void**)&b->d_single, b->h_single, 0) != cudaSuccess) {
    fprintf(stderr, "umpcInit (libumpc_b200): pinned I/O block: %s\n", cudaGetErrorString(cudaGetLastError()));
    umpcBatchDestroy(b);
    return;
  }
  memset(b->h_single, 0, nfl * sizeof(float));
  b->s.qpdata = b->d_single + 37;
  b->have_qpdata = true;
  std::lock_guard<std::mutex> lk(g_mu);
  g_single[up] = b;
}

extern "C" int umpcUpdate(UprightMPC_t* up, float uquad[3], float accdes[6], const float p0[3], const float R0[9],
                          const float dq0[6], const float pdes[3], const float dpdes[3], const float sdes[3],
                          float actualT0) {
  UmpcBatch* b = find_single(up);
  if (!b) {
    fprintf(stderr, "umpcUpdate (libumpc_b200): controller not initialised on a CUDA device (no CPU path)\n");
    for (int i = 0; i < 3; ++i) uquad[i] = nanf("");
    for (int i = 0; i < 6; ++i) accdes[i] = nanf("");
    return UMPC_ERR_CUDA;
  }
  if (cudaSetDevice(b->device) != cudaSuccess) return UMPC_ERR_CUDA;
  // The reference re-reads the public fields of *up on every call (updateObjective / umpcUpdateConstraint,
  // uprightmpc2.c:126-207): a caller may edit weights, dt, Tmax or T0 between calls.
  UmpcParams& m = b->kp.mpc;
  m.dt = up->dt; m.g = up->g; m.Tmax = up->Tmax;
  for (int i = 0; i < 6; ++i) { m.Qyr[i] = up->Qyr[i]; m.Qyf[i] = up->Qyf[i]; m.Qdyr[i] = up->Qdyr[i]; m.Qdyf[i] = up->Qdyf[i]; }
  for (int i = 0; i < 3; ++i) m.R[i] = up->R[i];
  m.Ibi[0] = up->e3hIbi[1]; m.Ibi[1] = -up->e3hIbi[3];   // e3h * diag(1/Ib), column-major (uprightmpc2.c:48-55)
  float* hin = b->h_single;
  memcpy(hin, p0, 12); memcpy(hin + 3, R0, 36); memcpy(hin + 12, dq0, 24); memcpy(hin + 18, pdes, 12);
  memcpy(hin + 21, dpdes, 12); memcpy(hin + 24, sdes, 12);
  // T0: `if (actualT0 >= 0) T0 = actualT0` (uprightmpc2.c:215-216); otherwise the struct's own T0 (a public field the
  // caller may have edited, uprightmpc2.h:42) -- both reach the kernel through its actualT0 input
  hin[27] = actualT0 >= 0.0f ? actualT0 : up->T0;
  if (!(hin[27] >= 0.0f)) {  // negative / NaN T0 cannot travel that way: write the slab row instead
    if (cudaMemcpy(b->s.T0, &up->T0, sizeof(float), cudaMemcpyHostToDevice) != cudaSuccess) return UMPC_ERR_CUDA;
    hin[27] = -1.0f;
  }
  const float* d = b->d_single;
  float* dout = b->d_single + 28;
  int rc = launch_update(b, dout, dout + 3, d, d + 3, d + 12, d + 18, d + 21, d + 24, d + 27, nullptr, nullptr, nullptr, 0, 128);
  if (!rc && cudaStreamSynchronize(b->stream) != cudaSuccess) rc = set_err("umpcUpdate", cudaGetLastError());
  if (rc) { fprintf(stderr, "umpcUpdate (libumpc_b200): %s\n", umpcLastError()); return rc; }
  const float* ho = b->h_single + 28;
  memcpy(uquad, ho, 12); memcpy(accdes, ho + 3, 24);
  const float* qp = b->h_single + 37;  // [216][Bp] rows, env 0
  const size_t P = b->Bp;
  for (int i = 0; i < 39; ++i) { up->l[i] = qp[(size_t)i * P]; up->u[i] = qp[(size_t)(39 + i) * P]; }
  for (int i = 0; i < 45; ++i) { up->q[i] = qp[(size_t)(78 + i) * P]; up->Px_data[i] = qp[(size_t)(123 + i) * P]; }
  for (int i = 0; i < 48; ++i) up->Ax_data[i] = qp[(size_t)(168 + i) * P];
  up->T0 = uquad[0];  // T0 += dT; uquad[0] = T0 (uprightmpc2.c:253-257)
  return 0;
}

static UprightMPC_t g_up_s;
static int g_up_inited = 0;
extern "C" void umpcS(float uquad[3], float accdes[6], const float p0[3], const float R0[9], const float dq0[6],
                      const float pdes[3], const float dpdes[3], const float sdes[3], float dt, float g, float TtoWmax,
                      float ws, float wds, float wpr, float wpf, float wvr, float wvf, float wthrust, float wmom,
                      const float Ib[3], int maxIter, float actualT0) {
  if (!g_up_inited) {  // lazy singleton like uprightmpc2.c:275-283
    umpcInit(&g_up_s, dt, g, TtoWmax, ws, wds, wpr, wpf, wvr, wvf, wthrust, wmom, Ib, maxIter);
    g_up_inited = 1;
  }
  umpcUpdate(&g_up_s, uquad, accdes, p0, R0, dq0, pdes, dpdes, sdes, actualT0);
}

extern "C" void umpcGetInfo(const UprightMPC_t* up, float info4[4]) {
  UmpcBatch* b = find_single(up);
  if (!b) { for (int i = 0; i < 4; ++i) info4[i] = nanf(""); return; }
  umpcBatchGetSolve(b, info4, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr);
}
extern "C" void umpcGetSolution(const UprightMPC_t* up, float x[45], float y[39]) {
  UmpcBatch* b = find_single(up);
  if (!b) return;
  umpcBatchGetSolve(b, nullptr, x, y, nullptr, nullptr, nullptr, nullptr, nullptr);
}

// ---- WL stage, single instance ------------------------------------------------------------------
static UmpcBatch* g_wl_batch = nullptr;  // batch of 1 used as device scratch for the WL drop-ins

static int wl_single(const WlParams& wp, int mode, float u0[4], WlqpState* qs, const float h0[6], const float pdotdes[6],
                     float w0[6]) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_wl_batch) {
    UmpcBatchParams bp;
    umpcBatchDefaults(&bp);
    int dev = 0;
    cudaGetDevice(&dev);
    if (umpcBatchCreate(&g_wl_batch, 1, dev, &bp) != UMPC_OK) {
      fprintf(stderr, "libumpc_b200 WL stage: %s\n", umpcLastError());
      g_wl_batch = nullptr;
      return UMPC_ERR_CUDA;
    }
  }
  UmpcBatch* b = g_wl_batch;
  CK(cudaSetDevice(b->device));
  float in[12], out[10];
  memcpy(in, h0, 24); memcpy(in + 6, pdotdes, 24);
  CK(cudaMemcpyAsync(b->d_in, in, sizeof(in), cudaMemcpyHostToDevice, b->stream));
  const size_t pitch = 32 * sizeof(float);  // tiled SoA: element i of env 0 sits at 32 i (umpc_dev.cuh)
  CK(cudaMemcpy2DAsync(b->s.wl_u0, pitch, u0, 4, 4, 4, cudaMemcpyHostToDevice, b->stream));
  if (qs) {
    float q[25];
    memcpy(q, qs->xs, 16); memcpy(q + 4, qs->zs, 16); memcpy(q + 8, qs->ys, 16); memcpy(q + 12, qs->D, 16);
    memcpy(q + 16, qs->E, 16); q[20] = qs->c; memcpy(q + 21, qs->q0, 16);
    CK(cudaMemcpy2DAsync(b->s.wlqp, pitch, q, 4, 4, 25, cudaMemcpyHostToDevice, b->stream));
  }
  wl_update_kernel<<<1, 32, 0, b->stream>>>(wp, b->s, mode, b->d_out, b->d_out + 4, b->d_in, b->d_in + 6, nullptr);
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(out, b->d_out, sizeof(out), cudaMemcpyDeviceToHost, b->stream));
  float q[25];
  if (qs) CK(cudaMemcpy2DAsync(q, 4, b->s.wlqp, pitch, 4, 25, cudaMemcpyDeviceToHost, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  memcpy(u0, out, 16); memcpy(w0, out + 4, 24);
  if (qs) {
    memcpy(qs->xs, q, 16); memcpy(qs->zs, q + 4, 16); memcpy(qs->ys, q + 8, 16); memcpy(qs->D, q + 12, 16);
    memcpy(qs->E, q + 16, 16); qs->c = q[20]; memcpy(qs->q0, q + 21, 16);
  }
  return UMPC_OK;
}

extern "C" void wlConInit(WLCon_t* wl, const float u0[4], const float umin[4], const float umax[4], const float dumax[4],
                          const float Qw[6], float controlRate, const float popts[90]) {
  memset(wl, 0, sizeof(*wl));
  for (int i = 0; i < NDELU; ++i) {
    wl->u0[i] = u0[i]; wl->umin[i] = umin[i]; wl->umax[i] = umax[i]; wl->dumax[i] = dumax[i] / controlRate;
  }
  for (int i = 0; i < 6; ++i) {
    wl->Qw[i + 6 * i] = Qw[i];
    FunApprox_t& f = wl->fa[i];
    const float* po = popts + 15 * i;
    f.k = NDELU; f.a0 = po[0];
    int kk = 1 + NDELU;
    for (int r = 0; r < NDELU; ++r) f.a1[r] = po[1 + r];
    for (int r = 0; r < NDELU; ++r)
      for (int c = r; c < NDELU; ++c) { f.A2[r + NDELU * c] = f.A2[c + NDELU * r] = po[kk++]; }
  }
}

static void wlparams_from_struct(const WLCon_t* wl, WlParams& wp) {
  memset(&wp, 0, sizeof(wp));
  for (int i = 0; i < 6; ++i) {
    const FunApprox_t& f = wl->fa[i];
    float* po = wp.popts + 15 * i;
    po[0] = f.a0;
    int kk = 1 + NDELU;
    for (int r = 0; r < NDELU; ++r) po[1 + r] = f.a1[r];
    for (int r = 0; r < NDELU; ++r)
      for (int c = r; c < NDELU; ++c) po[kk++] = f.A2[r + NDELU * c];
    wp.Qw[i] = wl->Qw[i + 6 * i];
  }
  for (int i = 0; i < NDELU; ++i) { wp.umin[i] = wl->umin[i]; wp.umax[i] = wl->umax[i]; wp.dumax[i] = wl->dumax[i]; }
}

extern "C" void wlConUpdate(WLCon_t* wl, float u1[4], float w0[6], const float h0[6], const float pdotdes[6]) {
  WlParams wp;
  wlparams_from_struct(wl, wp);
  if (wl_single(wp, 1, wl->u0, nullptr, h0, pdotdes, w0) != UMPC_OK) {
    for (int i = 0; i < 4; ++i) u1[i] = nanf("");
    return;
  }
  for (int i = 0; i < 4; ++i) u1[i] = wl->u0[i];
}

static WLCon_t g_wl_s;
static int g_wl_inited = 0;
extern "C" void wlconS(float u1[4], float w0[6], const float u0init[4], const float umin[4], const float umax[4],
                       const float dumax[4], const float Qw[6], float controlRate, const float popts[90], const float h0[6],
                       const float pdotdes[6]) {
  if (!g_wl_inited) {  // funapprox.c:168-177
    wlConInit(&g_wl_s, u0init, umin, umax, dumax, Qw, controlRate, popts);
    g_wl_inited = 1;
  }
  wlConUpdate(&g_wl_s, u1, w0, h0, pdotdes);
}

// true WLQP with the global lazily-initialised state of wlcontroller.c:17-45
static float g_u0[4] = {-1.0f, 0.0f, 0.0f, 0.0f};
static WlqpState g_wlqp_state;
static WlParams g_wlqp_par;
static int g_wlqp_check = 25;
static int g_wlqp_have_model = 0;

extern "C" void wlControllerReset(void) { g_u0[0] = -1.0f; g_u0[1] = g_u0[2] = g_u0[3] = 0.0f; }
extern "C" void wlControllerSetCheckTermination(int k) { g_wlqp_check = k; }

extern "C" void wlControllerUpdateW(float* u, float* w0, const float* u0init, const float* h0, const float* pdotdes,
                                    const float* popts, float controlRate) {
  if (g_u0[0] < 0.0f) {
    UmpcBatchParams bp;
    umpcBatchDefaults(&bp);  // wlqpInit limits (wlqp.c:17-46) == WaypointHover limits
    UmpcBatch tmp;
    tmp.bp = bp;
    memcpy(tmp.bp.wl_popts, popts, sizeof(tmp.bp.wl_popts));
    tmp.bp.wl_controlRate = controlRate;
    tmp.bp.wlqpCheckTermination = g_wlqp_check;
    fill_kernel_params(&tmp);
    g_wlqp_par = tmp.wlp;
    g_wlqp_have_model = 1;
    wlqp_state_init(g_wlqp_state);
    for (int i = 0; i < 4; ++i) g_u0[i] = u0init[i];
  }
  float w0l[6];
  if (wl_single(g_wlqp_par, 2, g_u0, &g_wlqp_state, h0, pdotdes, w0l) != UMPC_OK) {
    for (int i = 0; i < 4; ++i) u[i] = nanf("");
    return;
  }
  for (int i = 0; i < 4; ++i) u[i] = g_u0[i];
  if (w0) memcpy(w0, w0l, sizeof(w0l));
}

extern "C" void wlControllerUpdate(float* u, const float* u0init, const float* h0, const float* pdotdes, const float* popts,
                                   float controlRate) {
  wlControllerUpdateW(u, nullptr, u0init, h0, pdotdes, popts, controlRate);
}

// ---- wlqpInit / wlqpUpdate (wlcontroller/c/wlqp.h:21-31) and funApproxInit/F/Df (uprightmpc2/funapprox.h:31-35) ----
// A C caller that links the reference's WLQP object directly binds these.  The reference's wlqpUpdate calls the
// application-provided wrenchMap / wrenchJacMap (wlqp.h:18-19; wlcontroller.c defines them from its global fa[]):
// host callbacks cannot run on the device, so the quadratic wrench model (popts[90]) is the one installed by
// wlqpSetWrenchModel, or by the first wlControllerUpdate of the process (wlcontroller.c:21-33).  The OSQP warm start
// is process-global like the reference's generated workspace (wlqp/workspace.c).
extern "C" void wlqpSetWrenchModel(const float popts[90]) {
  memcpy(g_wlqp_par.popts, popts, sizeof(g_wlqp_par.popts));
  g_wlqp_have_model = 1;
}

extern "C" void wlqpInit(WLQP_t* wlqp, float controlRate) {
  memset(wlqp, 0, sizeof(*wlqp));
  for (int i = 0; i < 6; ++i) wlqp->Q[i + 6 * i] = i < 3 ? 1.0f : 0.1f;   // wlqp.c:19-24
  const float umin[4] = {90.0f, -0.5f, -0.2f, -0.1f}, umax[4] = {240.0f, 0.5f, 0.2f, 0.1f};
  const float dumax[4] = {5.0e3f, 10.0f, 10.0f, 10.0f};                  // wlqp.c:27-41
  for (int i = 0; i < 4; ++i) { wlqp->umin[i] = umin[i]; wlqp->umax[i] = umax[i]; wlqp->U0[i] = dumax[i] / controlRate; }
  wlqp_state_init(g_wlqp_state);  // osqp_set_default_settings + max_iter 40 on the global workspace (wlqp.c:43-46)
}

extern "C" void wlqpUpdate(WLQP_t* wlqp, float* u, const float* u0, const float* h0, const float* pdotdes) {
  WlParams wp = g_wlqp_par;  // popts of the installed wrench model
  for (int i = 0; i < 4; ++i) { wp.umin[i] = wlqp->umin[i]; wp.umax[i] = wlqp->umax[i]; wp.dumax[i] = wlqp->U0[i]; }
  for (int i = 0; i < 6; ++i) wp.Qw[i] = wlqp->Q[i + 6 * i];
  wp.rho = 0.1f; wp.sigma = 1e-6f; wp.alpha = 1.6f; wp.eps_abs = 1e-3f; wp.eps_rel = 1e-3f;
  wp.max_iter = 40; wp.check_termination = g_wlqp_check; wp.scaling_iters = 10;
  float uu[4] = {u0[0], u0[1], u0[2], u0[3]};
  if (!g_wlqp_have_model || wl_single(wp, 2, uu, &g_wlqp_state, h0, pdotdes, wlqp->w0) != UMPC_OK) {
    if (!g_wlqp_have_model) fprintf(stderr, "wlqpUpdate (libumpc_b200): no wrench model -- call wlqpSetWrenchModel(popts) first\n");
    for (int i = 0; i < 4; ++i) u[i] = nanf("");
    return;
  }
  for (int i = 0; i < 4; ++i) u[i] = uu[i];
}

extern "C" void funApproxInit(FunApprox_t* fa, const float popts[/* 15 */]) {
  fa->k = NDELU; fa->a0 = popts[0];
  for (int r = 0; r < NDELU; ++r) fa->a1[r] = popts[1 + r];
  int kk = 1 + NDELU;
  for (int r = 0; r < NDELU; ++r)
    for (int c = r; c < NDELU; ++c) fa->A2[r + NDELU * c] = fa->A2[c + NDELU * r] = popts[kk++];
}

// one quadratic model evaluated on the device (the library has no host arithmetic path): out5 = f, df[4]
__global__ void funapprox_kernel(FunApprox_t fa, const float* __restrict__ xi, float* __restrict__ out5) {
  if (threadIdx.x || blockIdx.x) return;
  float x[4], Ax[4], f = fa.a0, quad = 0.0f;
  for (int i = 0; i < 4; ++i) x[i] = xi[i];
  for (int r = 0; r < 4; ++r) {
    float acc = 0.0f;
    for (int c = 0; c < 4; ++c) acc += fa.A2[r + 4 * c] * x[c];
    Ax[r] = acc;
    f += x[r] * fa.a1[r];
  }
  for (int r = 0; r < 4; ++r) { quad += x[r] * Ax[r]; out5[1 + r] = fa.a1[r] + Ax[r]; }
  out5[0] = f + 0.5f * quad;
}

static int funapprox_eval(const FunApprox_t* fa, const float* xi, float out5[5]) {
  static float* d = nullptr;  // 4 in + 5 out
  std::lock_guard<std::mutex> lk(g_mu);
  if (!d) CK(cudaMalloc(&d, 9 * sizeof(float)));
  CK(cudaMemcpy(d, xi, 4 * sizeof(float), cudaMemcpyHostToDevice));
  funapprox_kernel<<<1, 32>>>(*fa, d, d + 4);
  CK(cudaGetLastError());
  CK(cudaMemcpy(out5, d + 4, 5 * sizeof(float), cudaMemcpyDeviceToHost));
  return UMPC_OK;
}

extern "C" float funApproxF(const FunApprox_t* fa, const float* xi) {
  float o[5];
  if (funapprox_eval(fa, xi, o) != UMPC_OK) { fprintf(stderr, "funApproxF (libumpc_b200): %s\n", umpcLastError()); return nanf(""); }
  return o[0];
}

extern "C" void funApproxDf(float* Df, const FunApprox_t* fa, const float* xi) {
  float o[5];
  if (funapprox_eval(fa, xi, o) != UMPC_OK) { fprintf(stderr, "funApproxDf (libumpc_b200): %s\n", umpcLastError()); for (int i = 0; i < 4; ++i) Df[i] = nanf(""); return; }
  for (int i = 0; i < 4; ++i) Df[i] = o[1 + i];
}
