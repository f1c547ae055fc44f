// robobee3d_b200/csrc/dirtran.cu -- batched planar direct-transcription MPC (BASELINE config 5)
//
// One thread per environment runs the reference's whole per-step chain in float64:
//   linearise PlanarThrustStrokeDev at (y, previous input)      planar/FlappingModels.py:51-102
//   QP data of LTVDirTran.updateTrajectory / QOFTrajectoryError  controlutils/py/ltvsystem.py:181-232, :44-77
//   OSQP (scaling 0, warm start, termination test every check_termination iterations, rho adapted and kept)
//                                                                ltvsystem.py:281-319 -> osqp 0.5.0 (not vendored; the
//                                                                algorithm as restated and pinned in oracle/)
//   u_0 -> plant: int(dt/SIM_DT) Euler substeps of dydt         planar/mpc_thrust_strokedev.py:189-192
//
// B200-first design:
//  * x = (y_0..y_N, u_0..u_{N-1}) is re-ordered by knot, w_k = (y_k, u_k): the reduced ADMM system
//    H = P + sigma I + A' diag(rho) A is then block tridiagonal with 8x8 blocks (A = [dynamics; I], so the inequality
//    rows only add to H's diagonal and z~ = A x~ needs no second solve).  A block Cholesky of H replaces OSQP's sparse
//    LDL' of the (n+m)^2 KKT matrix: (N+1) 8x8 factors, no fill, no pivoting, no index arrays.  The couplings
//    H_{k+1,k} S_k^-1 are NOT stored: they are applied as the sparse product with [Ad Bd] (7 entries, registers).
//  * One ADMM iteration = one forward sweep over the knots (right-hand side + forward substitution) and one backward
//    sweep (back substitution + z~ = A x~ + the x/z/y updates), each knot's data touched once per sweep.
//  * All per-environment data (iterates, factor, linearisation) is a tiled struct of arrays double[B/32][rows][32] in HBM:
//    lane = environment, so every access of a warp is one 256-byte line.  The iteration is HBM/L2-bandwidth bound
//    (608 B of unique data per knot and iteration against ~0.8 kflop): the roofline bench.py reports for it.
//  * Environments converge independently (own iteration count, own rho).  The rollout is ONE loop over ADMM iterations
//    (mpc_run): a lane that finishes a solve moves its plant and starts its next MPC period in the same trip instead
//    of idling until the slowest lane of its warp is done.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include "../../include/umpc_b200.h"
#include "dirtran_core.cuh"

void umpc_internal_set_error(const char* msg);  // umpc_kernels.cu (thread-local message behind umpcLastError)

using namespace dtcore;

#ifndef DT_ROLLOUT_MIN_BLOCKS
#define DT_ROLLOUT_MIN_BLOCKS 2   // 255 registers; 3 (168 registers, ~1 KB of spills per thread) measured 1.5x slower: profiles/r01j_dt_min_blocks3.txt
#endif

namespace {

__device__ __forceinline__ Env env_of(double* slab, int rows, int e) {
  Env v;
  v.base = slab + (size_t)(e >> 5) * rows * LANES + (e & 31);
  return v;
}

// kernel 1: one LTVMPC.update per environment from caller-supplied states (x0 [B][6] in, u [B][2] out)
template <bool PF>
__global__ void __launch_bounds__(128) dt_update_kernel(const __grid_constant__ DevPar P, double* __restrict__ slab, int rows, int nenv,
                                                        const double* __restrict__ x0, double* __restrict__ uout) {
  const int id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= nenv) return;
  const Env e = env_of(slab, rows, id);
  double y[NX];
#pragma unroll
  for (int i = 0; i < NX; ++i) { y[i] = x0[(size_t)id * NX + i]; e.h(R_Y + i) = y[i]; }
  double u0 = e.h(R_CTRL), u1 = e.h(R_CTRL + 1);
  mpc_update<PF>(P, e, y, u0, u1);
  uout[(size_t)id * NU] = u0; uout[(size_t)id * NU + 1] = u1;
}

// kernel 2: nsteps of the closed loop (mpc_thrust_strokedev.py:146-199) without leaving the device
template <bool PF>
__global__ void __launch_bounds__(128, DT_ROLLOUT_MIN_BLOCKS) dt_rollout_kernel(const __grid_constant__ DevPar P, double* __restrict__ slab, int rows, int nenv,
                                                         int nsteps) {
  const int id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= nenv) return;
  const Env e = env_of(slab, rows, id);
  double y[NX];
#pragma unroll
  for (int i = 0; i < NX; ++i) y[i] = e.h(R_Y + i);
  double u0 = e.h(R_CTRL), u1 = e.h(R_CTRL + 1);
  mpc_run<true, PF>(P, e, y, u0, u1, nsteps);
#pragma unroll
  for (int i = 0; i < NX; ++i) e.h(R_Y + i) = y[i];
}

__global__ void dt_fill_row_kernel(double* slab, int rows, int nenv_padded, int row, double v) {
  const int id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id < nenv_padded) env_of(slab, rows, id).h(row) = v;
}

// [B][k] host-layout array <-> k consecutive header/knot rows
__global__ void dt_scatter_rows_kernel(double* slab, int rows, int nenv, int row0, int k, const double* __restrict__ src) {
  const int id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= nenv) return;
  const Env e = env_of(slab, rows, id);
  for (int i = 0; i < k; ++i) e.h(row0 + i) = src[(size_t)id * k + i];
}
__global__ void dt_gather_rows_kernel(const double* slab, int rows, int nenv, int row0, int k, double* __restrict__ dst) {
  const int id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= nenv) return;
  const Env e = env_of(const_cast<double*>(slab), rows, id);
  for (int i = 0; i < k; ++i) dst[(size_t)id * k + i] = e.h(row0 + i);
}
// solution in OSQP's ordering: x = (y_0..y_N, u_0..u_{N-1}); y = (dynamics rows, state-bound rows, input-bound rows)
__global__ void dt_solution_kernel(const double* slab, int rows, int nenv, int N, double* __restrict__ x, double* __restrict__ yv) {
  const int id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= nenv) return;
  const Env e = env_of(const_cast<double*>(slab), rows, id);
  const int n = (N + 1) * NX + N * NU, m = n + (N + 1) * NX;
  for (int k = 0; k <= N; ++k) {
    const double* kb = e.knot(k);
    for (int i = 0; i < NX; ++i) {
      if (x) x[(size_t)id * n + k * NX + i] = kb[(K_X + i) * LANES];
      if (yv) {
        yv[(size_t)id * m + k * NX + i] = kb[(K_YE + i) * LANES];
        yv[(size_t)id * m + (N + 1) * NX + k * NX + i] = kb[(K_YI + i) * LANES];
      }
    }
    if (k < N)
      for (int i = 0; i < NU; ++i) {
        if (x) x[(size_t)id * n + (N + 1) * NX + k * NU + i] = kb[(K_X + NX + i) * LANES];
        if (yv) yv[(size_t)id * m + 2 * (N + 1) * NX + k * NU + i] = kb[(K_YI + NX + i) * LANES];
      }
  }
}
__global__ void dt_counters_kernel(const double* slab, int rows, int nenv, double* out3) {
  double a[3] = {0, 0, 0};
  for (int id = blockIdx.x * blockDim.x + threadIdx.x; id < nenv; id += gridDim.x * blockDim.x) {
    const Env e = env_of(const_cast<double*>(slab), rows, id);
    for (int i = 0; i < 3; ++i) a[i] += e.h(R_CNT + i);
  }
  for (int i = 0; i < 3; ++i) {
    for (int o = 16; o; o >>= 1) a[i] += __shfl_xor_sync(0xffffffffu, a[i], o);
    if ((threadIdx.x & 31) == 0) atomicAdd(out3 + i, a[i]);
  }
}

}  // namespace

struct DtBatch {
  int B = 0, Bp = 0, device = 0, rows = 0;
  DtParams par;
  DevPar dp;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  bool timed = false;
  double* slab = nullptr;
  size_t slab_doubles = 0;
  double *d_io = nullptr;  // staging: 8 doubles per env
  double* d_big = nullptr; // staging for solutions (allocated on demand)
  size_t big_doubles = 0;
  double* d_cnt = nullptr;
};

static bool dt_prefetch(const DtBatch* b);

static int dt_err(const char* what, cudaError_t e) {
  char buf[256];
  snprintf(buf, sizeof buf, "%s: %s", what, cudaGetErrorString(e));
  umpc_internal_set_error(buf);
  return UMPC_ERR_CUDA;
}
#define DCK(call)                                        \
  do {                                                   \
    cudaError_t e_ = (call);                             \
    if (e_ != cudaSuccess) return dt_err(#call, e_);     \
  } while (0)

extern "C" void dtBatchDefaults(DtParams* p, int N) {
  memset(p, 0, sizeof(*p));
  p->N = N;
  p->dt = 0.03;  // 'hover' set, mpc_thrust_strokedev.py:30
  const double wx[DT_NX] = {100, 100, 1, 10, 10, 0.1};
  for (int i = 0; i < DT_NX; ++i) p->wx[i] = wx[i] / p->dt;  // :118
  p->wu[0] = p->wu[1] = 0.01;
  p->eps_abs = p->eps_rel = 1e-5;
  p->rho = 0.1; p->sigma = 1e-6; p->alpha = 1.6;  // OSQP defaults
  p->max_iter = 4000; p->check_termination = 25; p->adaptive_rho_interval = 25; p->adaptive_rho_tol = 5.0;
  p->g = 9.81; p->ibmb = 1.0 / 12.0 * 0.1 * 0.1;  // FlappingModels.py:33,45,49
  p->umin[0] = -p->g; p->umin[1] = -0.05 * 0.1; p->umax[0] = 3 * p->g; p->umax[1] = 0.05 * 0.1;  // :104-107
  for (int i = 0; i < DT_NX; ++i) { p->xmin[i] = -INFINITY; p->xmax[i] = INFINITY; }
  p->xmin[2] = -10 * M_PI; p->xmax[2] = 10 * M_PI;  // :111
  p->sim_dt = 0.0005;
  p->nsub = (int)(p->dt / p->sim_dt);  // :190
}

static int dt_fill(DtBatch* b, int row, double v) {
  dt_fill_row_kernel<<<(b->Bp + 255) / 256, 256, 0, b->stream>>>(b->slab, b->rows, b->Bp, row, v);
  DCK(cudaGetLastError());
  return UMPC_OK;
}

extern "C" int dtBatchReset(DtBatch* b) {
  if (!b) return UMPC_ERR_ARG;
  DCK(cudaSetDevice(b->device));
  DCK(cudaMemsetAsync(b->slab, 0, b->slab_doubles * sizeof(double), b->stream));
  return dt_fill(b, R_RHO, b->par.rho);
}

static int dt_create_alloc(DtBatch* b);
extern "C" int dtBatchDestroy(DtBatch* b);

extern "C" int dtBatchCreate(DtBatch** out, int nenv, int device, const DtParams* par) {
  if (!out || nenv <= 0 || !par) { umpc_internal_set_error("dtBatchCreate: bad argument"); return UMPC_ERR_ARG; }
  *out = nullptr;
  if (par->N != 3 && par->N != 5 && par->N != 10 && par->N != 15 && par->N != 20) {
    umpc_internal_set_error("dtBatchCreate: N must be 3, 5, 10, 15 or 20");
    return UMPC_ERR_ARG;
  }
  if (par->max_iter <= 0 || par->nsub < 0 || !(par->rho > 0) || !(par->dt > 0)) {
    umpc_internal_set_error("dtBatchCreate: bad parameter (max_iter, nsub, rho, dt)");
    return UMPC_ERR_ARG;
  }
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    umpc_internal_set_error((std::string("dtBatchCreate: no CUDA device (") + cudaGetErrorString(e) + "); this library has no CPU path").c_str());
    return UMPC_ERR_CUDA;
  }
  DCK(cudaSetDevice(device));
  DtBatch* b = new DtBatch();
  b->B = nenv; b->Bp = (nenv + 31) & ~31; b->device = device; b->par = *par;
  b->rows = R_HDR + K_STRIDE * (par->N + 1);
  DevPar& d = b->dp;
  memset(&d, 0, sizeof(d));
  d.N = par->N; d.max_iter = par->max_iter; d.check_termination = par->check_termination;
  d.adaptive_rho_interval = par->adaptive_rho_interval; d.nsub = par->nsub;
  d.dt = par->dt; d.sim_dt = par->sim_dt; d.sigma = par->sigma; d.alpha = par->alpha; d.eps_abs = par->eps_abs; d.eps_rel = par->eps_rel;
  d.rho0 = par->rho; d.adaptive_rho_tol = par->adaptive_rho_tol; d.g = par->g; d.ibmb_inv = 1.0 / par->ibmb;
  const double OSQP_INFTY = 1e20, MIN_SCALING = 1e-4;  // osqp 0.5.0 constants.h; the Python wrapper clips +-inf to +-OSQP_INFTY
  for (int j = 0; j < NW; ++j) {
    double lo = j < NX ? par->xmin[j] : par->umin[j - NX], hi = j < NX ? par->xmax[j] : par->umax[j - NX];
    lo = fmax(lo, -OSQP_INFTY); hi = fmin(hi, OSQP_INFTY);
    if (lo > hi) { delete b; umpc_internal_set_error("dtBatchCreate: lower bound above upper bound"); return UMPC_ERR_ARG; }
    d.lo[j] = lo; d.hi[j] = hi;
    if (lo < -OSQP_INFTY * MIN_SCALING && hi > OSQP_INFTY * MIN_SCALING) d.loose |= 1u << j;  // auxil.c set_rho_vec
    if (j < NX) d.wx[j] = par->wx[j]; else d.wu[j - NX] = par->wu[j - NX];
  }
  const int rc = dt_create_alloc(b);
  if (rc != UMPC_OK) { dtBatchDestroy(b); return rc; }  // a failed create leaves nothing allocated
  *out = b;
  return UMPC_OK;
}

static int dt_create_alloc(DtBatch* b) {
  DCK(cudaStreamCreateWithFlags(&b->stream, cudaStreamNonBlocking));
  DCK(cudaEventCreate(&b->ev0));
  DCK(cudaEventCreate(&b->ev1));
  b->slab_doubles = (size_t)b->rows * b->Bp;
  DCK(cudaMalloc(&b->slab, b->slab_doubles * sizeof(double)));
  DCK(cudaMalloc(&b->d_io, (size_t)8 * b->Bp * sizeof(double)));
  DCK(cudaMalloc(&b->d_cnt, 3 * sizeof(double)));
  int rc = dtBatchReset(b);
  if (rc != UMPC_OK) return rc;
  DCK(cudaStreamSynchronize(b->stream));
  return UMPC_OK;
}

extern "C" int dtBatchDestroy(DtBatch* b) {
  if (!b) return UMPC_OK;
  cudaSetDevice(b->device);
  if (b->stream) cudaStreamSynchronize(b->stream);
  cudaFree(b->slab); cudaFree(b->d_io); cudaFree(b->d_big); cudaFree(b->d_cnt);
  if (b->ev0) cudaEventDestroy(b->ev0);
  if (b->ev1) cudaEventDestroy(b->ev1);
  if (b->stream) cudaStreamDestroy(b->stream);
  cudaGetLastError();
  delete b;
  return UMPC_OK;
}

static int dt_put(DtBatch* b, const double* host, int row0, int k) {
  DCK(cudaMemcpyAsync(b->d_io, host, (size_t)b->B * k * sizeof(double), cudaMemcpyHostToDevice, b->stream));
  dt_scatter_rows_kernel<<<(b->B + 127) / 128, 128, 0, b->stream>>>(b->slab, b->rows, b->B, row0, k, b->d_io);
  DCK(cudaGetLastError());
  DCK(cudaStreamSynchronize(b->stream));
  return UMPC_OK;
}
static int dt_get(DtBatch* b, double* host, int row0, int k) {
  dt_gather_rows_kernel<<<(b->B + 127) / 128, 128, 0, b->stream>>>(b->slab, b->rows, b->B, row0, k, b->d_io);
  DCK(cudaGetLastError());
  DCK(cudaMemcpyAsync(host, b->d_io, (size_t)b->B * k * sizeof(double), cudaMemcpyDeviceToHost, b->stream));
  DCK(cudaStreamSynchronize(b->stream));
  return UMPC_OK;
}

extern "C" int dtBatchSetState(DtBatch* b, const double* y, const double* ctrl, const double* xr) {
  if (!b) return UMPC_ERR_ARG;
  DCK(cudaSetDevice(b->device));
  int rc = UMPC_OK;
  if (y && (rc = dt_put(b, y, R_Y, NX)) != UMPC_OK) return rc;
  if (ctrl && (rc = dt_put(b, ctrl, R_CTRL, NU)) != UMPC_OK) return rc;
  if (xr && (rc = dt_put(b, xr, R_XR, NX)) != UMPC_OK) return rc;
  return rc;
}

extern "C" int dtBatchGetState(DtBatch* b, double* y, double* ctrl) {
  if (!b) return UMPC_ERR_ARG;
  DCK(cudaSetDevice(b->device));
  int rc = UMPC_OK;
  if (y && (rc = dt_get(b, y, R_Y, NX)) != UMPC_OK) return rc;
  if (ctrl && (rc = dt_get(b, ctrl, R_CTRL, NU)) != UMPC_OK) return rc;
  return rc;
}

extern "C" int dtBatchUpdate(DtBatch* b, const double* x0, double* u) {
  if (!b || !x0 || !u) { umpc_internal_set_error("dtBatchUpdate: NULL argument"); return UMPC_ERR_ARG; }
  DCK(cudaSetDevice(b->device));
  double* d_x0 = b->d_io;
  double* d_u = b->d_io + (size_t)NX * b->Bp;
  DCK(cudaMemcpyAsync(d_x0, x0, (size_t)b->B * NX * sizeof(double), cudaMemcpyHostToDevice, b->stream));
  DCK(cudaEventRecord(b->ev0, b->stream));
  if (dt_prefetch(b)) dt_update_kernel<true><<<(b->B + 127) / 128, 128, 0, b->stream>>>(b->dp, b->slab, b->rows, b->B, d_x0, d_u);
  else dt_update_kernel<false><<<(b->B + 127) / 128, 128, 0, b->stream>>>(b->dp, b->slab, b->rows, b->B, d_x0, d_u);
  DCK(cudaGetLastError());
  DCK(cudaEventRecord(b->ev1, b->stream));
  b->timed = true;
  DCK(cudaMemcpyAsync(u, d_u, (size_t)b->B * NU * sizeof(double), cudaMemcpyDeviceToHost, b->stream));
  DCK(cudaStreamSynchronize(b->stream));
  return UMPC_OK;
}

// Launch shape of the rollout: threads per block and a dynamic shared-memory reservation whose only purpose is to cap
// the resident threads per SM (the kernel uses no shared memory).  Fewer resident environments = a smaller working set
// per ADMM iteration; once it fits the 126 MB L2 the iteration stops paying HBM for its second sweep.
// Tuning knobs DT_TPB / DT_SMEM_KB override the per-horizon default.
// next-knot L2 prefetch (dirtran_core.cuh prefetch_knot): on for horizons above 5 (measured: N = 3, 5 lose, N >= 10 gain 13 %); DT_PREFETCH=0/1 overrides
static bool dt_prefetch(const DtBatch* b) {
  const char* e = getenv("DT_PREFETCH");
  return e ? atoi(e) != 0 : b->par.N > 5;
}

static void dt_launch_shape(const DtBatch* b, int* tpb, int* smem) {
  *tpb = 128; *smem = 0;
  const char* e = getenv("DT_TPB");
  if (e && (atoi(e) == 32 || atoi(e) == 64 || atoi(e) == 128)) *tpb = atoi(e);
  e = getenv("DT_SMEM_KB");
  if (e && atoi(e) >= 0 && atoi(e) <= 220) *smem = atoi(e) * 1024;
}

extern "C" int dtBatchRollout(DtBatch* b, int nsteps) {
  if (!b || nsteps < 0) return UMPC_ERR_ARG;
  DCK(cudaSetDevice(b->device));
  int tpb, smem;
  dt_launch_shape(b, &tpb, &smem);
  const bool pf = dt_prefetch(b);
  if (smem > 48 * 1024)
    DCK(cudaFuncSetAttribute(pf ? dt_rollout_kernel<true> : dt_rollout_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  DCK(cudaEventRecord(b->ev0, b->stream));
  if (pf) dt_rollout_kernel<true><<<(b->B + tpb - 1) / tpb, tpb, smem, b->stream>>>(b->dp, b->slab, b->rows, b->B, nsteps);
  else dt_rollout_kernel<false><<<(b->B + tpb - 1) / tpb, tpb, smem, b->stream>>>(b->dp, b->slab, b->rows, b->B, nsteps);
  DCK(cudaGetLastError());
  DCK(cudaEventRecord(b->ev1, b->stream));
  b->timed = true;
  return UMPC_OK;
}

extern "C" int dtBatchGetInfo(DtBatch* b, double* info) {
  if (!b || !info) return UMPC_ERR_ARG;
  DCK(cudaSetDevice(b->device));
  // iter, status, pri, dua from R_INFO..+3; rho from R_RHO; rho_updates from R_INFO+4: gather rows 8..13 and reorder on the host
  std::vector<double> tmp((size_t)b->B * 6);
  int rc = dt_get(b, tmp.data(), R_RHO, 6);
  if (rc != UMPC_OK) return rc;
  for (int i = 0; i < b->B; ++i) {
    const double* t = &tmp[(size_t)i * 6];
    double* o = info + (size_t)i * 6;
    o[0] = t[1]; o[1] = t[2]; o[2] = t[3]; o[3] = t[4]; o[4] = t[0]; o[5] = t[5];
  }
  return UMPC_OK;
}

extern "C" int dtBatchGetSolution(DtBatch* b, double* x, double* y) {
  if (!b) return UMPC_ERR_ARG;
  DCK(cudaSetDevice(b->device));
  const int N = b->par.N, n = (N + 1) * NX + N * NU, m = n + (N + 1) * NX;
  const size_t need = (size_t)b->B * (n + m);
  if (b->big_doubles < need) {
    cudaFree(b->d_big); b->d_big = nullptr; b->big_doubles = 0;
    DCK(cudaMalloc(&b->d_big, need * sizeof(double)));
    b->big_doubles = need;
  }
  double* dx = b->d_big;
  double* dy = b->d_big + (size_t)b->B * n;
  dt_solution_kernel<<<(b->B + 127) / 128, 128, 0, b->stream>>>(b->slab, b->rows, b->B, N, x ? dx : nullptr, y ? dy : nullptr);
  DCK(cudaGetLastError());
  if (x) DCK(cudaMemcpyAsync(x, dx, (size_t)b->B * n * sizeof(double), cudaMemcpyDeviceToHost, b->stream));
  if (y) DCK(cudaMemcpyAsync(y, dy, (size_t)b->B * m * sizeof(double), cudaMemcpyDeviceToHost, b->stream));
  DCK(cudaStreamSynchronize(b->stream));
  return UMPC_OK;
}

extern "C" int dtBatchCounters(DtBatch* b, double* out) {
  if (!b || !out) return UMPC_ERR_ARG;
  DCK(cudaSetDevice(b->device));
  DCK(cudaMemsetAsync(b->d_cnt, 0, 3 * sizeof(double), b->stream));
  dt_counters_kernel<<<296, 256, 0, b->stream>>>(b->slab, b->rows, b->B, b->d_cnt);
  DCK(cudaGetLastError());
  DCK(cudaMemcpyAsync(out, b->d_cnt, 3 * sizeof(double), cudaMemcpyDeviceToHost, b->stream));
  DCK(cudaStreamSynchronize(b->stream));
  return UMPC_OK;
}

extern "C" float dtBatchLastKernelMs(DtBatch* b) {
  if (!b || !b->timed) return -1.0f;
  float ms = -1.0f;
  if (cudaEventSynchronize(b->ev1) != cudaSuccess) return -1.0f;
  cudaEventElapsedTime(&ms, b->ev0, b->ev1);
  return ms;
}

extern "C" int dtBatchSync(DtBatch* b) {
  if (!b) return UMPC_ERR_ARG;
  DCK(cudaStreamSynchronize(b->stream));
  return UMPC_OK;
}
