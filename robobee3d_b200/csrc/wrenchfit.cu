// robobee3d_b200/csrc/wrenchfit.cu -- wrench-map fitting: the popts[90] the WL stage consumes.
//
// Reference path, template/sdab_num_wrenchmap.py:
//   wrenchFromKinematics   :165-214   stroke-averaged wrench of both wings from 9 kinematics features per trial
//   FunApprox.f            :144-152   w_i(u) = a0 + a1'u + 1/2 u'A2 u, 15 parameters per wrench component
//   the fit                :262-266   curve_fit(fa.f, xdata, ws[:, i], p0 = ones) for i in 0..5 -> popts (6, 15)
// The model is LINEAR in its parameters, so Levenberg-Marquardt from p0 converges to  p0 + pinv(Phi) (w - Phi p0)  --
// the least-squares solution closest to p0 (every LM step lies in the row space of the Jacobian Phi).  The reference's
// own bench data is rank deficient (rank 10 of 15, singular values 1.8e5 .. 1.5e-2, then 1e-16), so the normal
// equations are out (condition number squared): one CTA per dataset runs a ONE-SIDED JACOBI SVD of Phi (M x 15) in
// shared memory -- column pairs orthogonalised by plane rotations, dot products by warp-shuffle reductions -- and
// applies the pseudo-inverse to the six right-hand sides.  Batched over F datasets (sweeps of trials, re-fits).
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <mutex>
#include <string>

#include "../../include/umpc_b200.h"
#include "../../include/wrenchfit_b200.h"

void umpc_internal_set_error(const char* msg);

namespace wfit {

enum { NP = 15, NU4 = 4 };

// par = Aw, AR, r2h, ycp, R, RHO, CD0, CDmax, CLmax, kaerox, strokex, wbias[6]
__device__ inline void iwrencha(const double* par, double f2, double Phim, double Phid, double Psi1, double Psi2, double alpha, double w[6]) {
  const double CD0 = par[6], CDmax = par[7], CLmax = par[8];
  const double kaero = 0.5 * par[5] * par[0] * par[0] * par[1] * par[2] * par[2] * par[9];
  const double ycp = par[3], R = par[4];
  const double c2Psi1 = cos(2 * Psi1), s2Psi1 = sin(2 * Psi1), c2Psi2 = cos(2 * Psi2), s2Psi2 = sin(2 * Psi2);
  const double cPhid = cos(Phid), sPhid = sin(Phid), cPhim = cos(Phim), sPhim = sin(Phim);
  (void)cPhid;
  const double Phid2 = Phid * Phid;
  const double den = (-2 + alpha) * alpha;
  const double dragc = (-2 + alpha) * c2Psi1 * (CD0 - CDmax) + alpha * c2Psi2 * (CD0 - CDmax) - 2 * (-1 + alpha) * (CD0 + CDmax);
  const double liftc = (-2 + alpha) * s2Psi1 + alpha * s2Psi2;
  w[0] = (-4 * dragc * cPhim * f2 * kaero * Phid * sPhid) / den;
  w[1] = (-4 * dragc * f2 * kaero * Phid * sPhid * sPhim) / den;
  w[2] = (8 * CLmax * f2 * kaero * Phid2 * liftc) / den;
  w[3] = (8 * CLmax * f2 * kaero * Phid * liftc * (Phid * R + cPhim * sPhid * ycp)) / den;
  w[4] = (8 * CLmax * f2 * kaero * Phid * liftc * sPhid * sPhim * ycp) / den;
  w[5] = (4 * dragc * f2 * kaero * Phid * (cPhim * R * sPhid + Phid * ycp)) / den;
}

// sdab_num_wrenchmap.py:165-214, one trial per thread
__global__ void wrench_kernel(int M, const double* __restrict__ kins, const double* __restrict__ freq, const double* __restrict__ par,
                              double* __restrict__ ws) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= M) return;
  const double* k = kins + (size_t)r * 9;
  const double f2 = freq[r] * freq[r], strokex = par[10];
  double wl[6], wr[6];
  const double dalpha = 2 * k[8] - 1.0;
  // remapKins: Phim = (q1max - q1nmin)/2, Phid = (q1max + q1nmin)/2 * strokex, Psi1 = q2max, Psi2 = -q2nmin
  iwrencha(par, f2, 0.5 * (k[0] - k[2]), 0.5 * (k[0] + k[2]) * strokex, k[1], -k[3], 1 - dalpha, wl);
  iwrencha(par, f2, 0.5 * (k[4] - k[6]), 0.5 * (k[4] + k[6]) * strokex, k[5], -k[7], 1 + dalpha, wr);
  const double sym[6] = {1, -1, 1, -1, 1, -1};
  for (int i = 0; i < 6; ++i) ws[(size_t)r * 6 + i] = wl[i] + sym[i] * wr[i] + par[11 + i];
}

__device__ __forceinline__ double block_sum(double v, double* red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[w] = v;
  __syncthreads();
  double s = 0;
  for (int i = 0; i < nw; ++i) s += red[i];
  return s;
}

// basis of FunApprox.f: [1, u0..u3, (i <= j): (i == j ? 1/2 : 1) u_i u_j in triu row-major order]
__device__ inline void basis(const double* u, double* phi) {
  phi[0] = 1.0;
  for (int i = 0; i < 4; ++i) phi[1 + i] = u[i];
  int k = 5;
  for (int i = 0; i < 4; ++i)
    for (int j = i; j < 4; ++j) phi[k++] = (i == j ? 0.5 : 1.0) * u[i] * u[j];
}

// one CTA per dataset: popts = p0 + pinv(Phi) (W - Phi p0), p0 = ones
__global__ void __launch_bounds__(128) fit_kernel(int M, const double* __restrict__ xdata, const double* __restrict__ ws,
                                                  double* __restrict__ popts, double* __restrict__ fit, double rcond, int* __restrict__ rank_out) {
  extern __shared__ double sm[];
  double* A = sm;                 // M x 15, column-major: A[j * M + r] -- becomes U diag(sigma)
  double* V = A + (size_t)M * NP; // 15 x 15, column-major
  double* red = V + NP * NP;      // 8
  double* sig = red + 8;          // 15
  double* coef = sig + NP;        // 15 x 6
  const int f = blockIdx.x, t = threadIdx.x, nt = blockDim.x;
  const double* X = xdata + (size_t)f * M * 4;
  const double* W = ws + (size_t)f * M * 6;
  for (int r = t; r < M; r += nt) {
    double phi[NP];
    basis(X + (size_t)r * 4, phi);
    for (int j = 0; j < NP; ++j) A[(size_t)j * M + r] = phi[j];
  }
  for (int i = t; i < NP * NP; i += nt) V[i] = (i / NP == i % NP) ? 1.0 : 0.0;
  __syncthreads();
  // one-sided Jacobi (Hestenes): rotate column pairs until mutually orthogonal
  for (int sweep = 0; sweep < 30; ++sweep) {
    double off = 0.0;
    for (int p = 0; p < NP - 1; ++p)
      for (int q = p + 1; q < NP; ++q) {
        double a = 0, b = 0, c = 0;
        for (int r = t; r < M; r += nt) { const double x = A[(size_t)p * M + r], y = A[(size_t)q * M + r]; a += x * x; b += y * y; c += x * y; }
        a = block_sum(a, red); b = block_sum(b, red); c = block_sum(c, red);
        if (fabs(c) <= 1e-15 * sqrt(a * b) || c == 0.0) continue;     // uniform across the CTA: same reduced values
        off = fmax(off, fabs(c) / sqrt(a * b));
        const double zeta = (b - a) / (2.0 * c);
        const double tt = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
        const double cs = 1.0 / sqrt(1.0 + tt * tt), sn = cs * tt;
        for (int r = t; r < M; r += nt) {
          const double x = A[(size_t)p * M + r], y = A[(size_t)q * M + r];
          A[(size_t)p * M + r] = cs * x - sn * y; A[(size_t)q * M + r] = sn * x + cs * y;
        }
        for (int r = t; r < NP; r += nt) {
          const double x = V[p * NP + r], y = V[q * NP + r];
          V[p * NP + r] = cs * x - sn * y; V[q * NP + r] = sn * x + cs * y;
        }
        __syncthreads();
      }
    if (off < 1e-14) break;
  }
  // singular values; coef[j][i] = (U_j . r_i) / sigma_j  for the significant ones, r_i = W[:, i] - Phi * ones
  double smax = 0.0;
  for (int j = 0; j < NP; ++j) {
    double a = 0;
    for (int r = t; r < M; r += nt) { const double x = A[(size_t)j * M + r]; a += x * x; }
    a = block_sum(a, red);
    if (t == 0) sig[j] = sqrt(a);
    smax = fmax(smax, sqrt(a));
  }
  __syncthreads();
  int rank = 0;
  for (int j = 0; j < NP; ++j) {
    const bool keep = sig[j] > rcond * smax;
    rank += keep;
    for (int i = 0; i < 6; ++i) {
      double a = 0;
      if (keep)
        for (int r = t; r < M; r += nt) {
          double phi[NP], p1 = 0;
          basis(X + (size_t)r * 4, phi);
          for (int k = 0; k < NP; ++k) p1 += phi[k];
          a += A[(size_t)j * M + r] * (W[(size_t)r * 6 + i] - p1);
        }
      a = block_sum(a, red);
      if (t == 0) coef[j * 6 + i] = keep ? a / (sig[j] * sig[j]) : 0.0;   // U_j = A_j / sigma_j, then another 1 / sigma_j
    }
  }
  __syncthreads();
  if (t == 0 && rank_out) rank_out[f] = rank;
  // popts[i][k] = 1 + sum_j V[k][j] coef[j][i]   (V column-major: V[j * NP + k] = V_kj)
  for (int idx = t; idx < 6 * NP; idx += nt) {
    const int i = idx / NP, k = idx % NP;
    double v = 1.0;
    for (int j = 0; j < NP; ++j) v += V[j * NP + k] * coef[j * 6 + i];
    popts[(size_t)f * 90 + idx] = v;
  }
  __syncthreads();
  if (fit)
    for (int r = t; r < M; r += nt) {
      double phi[NP];
      basis(X + (size_t)r * 4, phi);
      for (int i = 0; i < 6; ++i) {
        double v = 0;
        for (int k = 0; k < NP; ++k) v += phi[k] * popts[(size_t)f * 90 + i * NP + k];
        fit[((size_t)f * M + r) * 6 + i] = v;
      }
    }
}


// ---- allocation-free variant (wfitFitNoAlloc) ------------------------------------------------------------------
// The same fit with the working set (design matrix / U diag(sigma), V, sigma, coefficients) in a file-scope device
// array instead of dynamic shared memory: no opt-in attribute, no allocation, any M up to the array's size.
// Every GPU that loads the library holds the array until exit (WFIT_NOALLOC_SCRATCH_DOUBLES * 8 bytes = 2.05 MiB).
enum { SMALL_STRIDE = 352 };   // per dataset: V 225 | sigma 15 | coef 90, padded to a multiple of 256 bytes
static __device__ __align__(256) double g_noalloc_scratch[WFIT_NOALLOC_MAX_TRIALS * NP + WFIT_NOALLOC_MAX_DATASETS * SMALL_STRIDE];

// one CTA per dataset; `scratch` is the DEVICE address of g_noalloc_scratch (cudaGetSymbolAddress, host side below)
__global__ void __launch_bounds__(128) fit_noalloc_kernel(int F, int M, const double* __restrict__ xdata, const double* __restrict__ ws,
                                                          double* __restrict__ popts, double rcond, int* __restrict__ rank_out,
                                                          double* __restrict__ scratch) {
  __shared__ double red[8];
  const int f = blockIdx.x, t = threadIdx.x, nt = blockDim.x;
  double* A = scratch + (size_t)f * M * NP;                                   // M x 15, column-major
  double* V = scratch + (size_t)F * M * NP + (size_t)f * SMALL_STRIDE;        // 15 x 15, column-major
  double* sig = V + NP * NP;
  double* coef = sig + NP;
  const double* X = xdata + (size_t)f * M * 4;
  const double* W = ws + (size_t)f * M * 6;
  // the array keeps the last call's contents: every word read below is written first (A, V = I here; sigma, coef later)
  for (int r = t; r < M; r += nt) {
    double phi[NP];
    basis(X + (size_t)r * 4, phi);
    for (int j = 0; j < NP; ++j) A[(size_t)j * M + r] = phi[j];
  }
  for (int i = t; i < SMALL_STRIDE; i += nt) V[i] = (i < NP * NP && i / NP == i % NP) ? 1.0 : 0.0;   // V = I, sigma = coef = 0
  __syncthreads();
  for (int sweep = 0; sweep < 30; ++sweep) {
    double off = 0.0;
    for (int p = 0; p < NP - 1; ++p)
      for (int q = p + 1; q < NP; ++q) {
        double a = 0, b = 0, c = 0;
        for (int r = t; r < M; r += nt) { const double x = A[(size_t)p * M + r], y = A[(size_t)q * M + r]; a += x * x; b += y * y; c += x * y; }
        a = block_sum(a, red); b = block_sum(b, red); c = block_sum(c, red);
        if (fabs(c) <= 1e-15 * sqrt(a * b) || c == 0.0) continue;
        off = fmax(off, fabs(c) / sqrt(a * b));
        const double zeta = (b - a) / (2.0 * c);
        const double tt = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
        const double cs = 1.0 / sqrt(1.0 + tt * tt), sn = cs * tt;
        for (int r = t; r < M; r += nt) {
          const double x = A[(size_t)p * M + r], y = A[(size_t)q * M + r];
          A[(size_t)p * M + r] = cs * x - sn * y; A[(size_t)q * M + r] = sn * x + cs * y;
        }
        for (int r = t; r < NP; r += nt) {
          const double x = V[p * NP + r], y = V[q * NP + r];
          V[p * NP + r] = cs * x - sn * y; V[q * NP + r] = sn * x + cs * y;
        }
        __syncthreads();
      }
    if (off < 1e-14) break;
  }
  double smax = 0.0;
  for (int j = 0; j < NP; ++j) {
    double a = 0;
    for (int r = t; r < M; r += nt) { const double x = A[(size_t)j * M + r]; a += x * x; }
    a = block_sum(a, red);
    if (t == 0) sig[j] = sqrt(a);
    smax = fmax(smax, sqrt(a));
  }
  __syncthreads();
  int rank = 0;
  for (int j = 0; j < NP; ++j) {
    const bool keep = sig[j] > rcond * smax;
    rank += keep;
    for (int i = 0; i < 6; ++i) {
      double a = 0;
      if (keep)
        for (int r = t; r < M; r += nt) {
          double phi[NP], p1 = 0;
          basis(X + (size_t)r * 4, phi);
          for (int k = 0; k < NP; ++k) p1 += phi[k];
          a += A[(size_t)j * M + r] * (W[(size_t)r * 6 + i] - p1);
        }
      a = block_sum(a, red);
      if (t == 0) coef[j * 6 + i] = keep ? a / (sig[j] * sig[j]) : 0.0;
    }
  }
  __syncthreads();
  if (t == 0 && rank_out) rank_out[f] = rank;
  for (int idx = t; idx < 6 * NP; idx += nt) {
    const int i = idx / NP, k = idx % NP;
    double v = 1.0;
    for (int j = 0; j < NP; ++j) v += V[j * NP + k] * coef[j * 6 + i];
    popts[(size_t)f * 90 + idx] = v;
  }
}

}  // namespace wfit

#define WCK(call)                                                                         \
  do {                                                                                    \
    cudaError_t e_ = (call);                                                              \
    if (e_ != cudaSuccess) {                                                              \
      umpc_internal_set_error((std::string(#call) + ": " + cudaGetErrorString(e_)).c_str()); \
      return UMPC_ERR_CUDA;                                                               \
    }                                                                                     \
  } while (0)

// device buffer / event that frees itself: the entry points below return early on any CUDA error
namespace {
struct DevBuf {
  void* p = nullptr;
  ~DevBuf() { if (p) cudaFree(p); }
  cudaError_t alloc(size_t bytes) { return cudaMalloc(&p, bytes); }
  template <class T> T* as() const { return (T*)p; }
};
struct EvPair {
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  ~EvPair() { if (e0) cudaEventDestroy(e0); if (e1) cudaEventDestroy(e1); }
};
}  // namespace

static int need_device(const char* who) {
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    umpc_internal_set_error((std::string(who) + ": no CUDA device (" + cudaGetErrorString(e) + "); this library has no CPU path").c_str());
    return UMPC_ERR_CUDA;
  }
  return UMPC_OK;
}

extern "C" int wfitWrenchFromKinematics(int M, const double* kins, const double* freq, const double par[17], double* ws) {
  if (M <= 0 || !kins || !freq || !par || !ws) { umpc_internal_set_error("wfitWrenchFromKinematics: bad argument"); return UMPC_ERR_ARG; }
  int rc = need_device("wfitWrenchFromKinematics");
  if (rc) return rc;
  DevBuf bk, bf, bp, bw;
  WCK(bk.alloc((size_t)M * 9 * 8)); WCK(bf.alloc((size_t)M * 8)); WCK(bp.alloc(17 * 8)); WCK(bw.alloc((size_t)M * 6 * 8));
  double *dk = bk.as<double>(), *df = bf.as<double>(), *dp = bp.as<double>(), *dw = bw.as<double>();
  WCK(cudaMemcpy(dk, kins, (size_t)M * 9 * 8, cudaMemcpyHostToDevice)); WCK(cudaMemcpy(df, freq, (size_t)M * 8, cudaMemcpyHostToDevice));
  WCK(cudaMemcpy(dp, par, 17 * 8, cudaMemcpyHostToDevice));
  wfit::wrench_kernel<<<(M + 127) / 128, 128>>>(M, dk, df, dp, dw);
  WCK(cudaGetLastError());
  WCK(cudaMemcpy(ws, dw, (size_t)M * 6 * 8, cudaMemcpyDeviceToHost));
  return UMPC_OK;
}

extern "C" int wfitFit(int F, int M, const double* xdata, const double* ws, double* popts, double* fit, int* rank, float* kernel_ms) {
  if (F <= 0 || M <= 0 || !xdata || !ws || !popts) { umpc_internal_set_error("wfitFit: bad argument"); return UMPC_ERR_ARG; }
  int rc = need_device("wfitFit");
  if (rc) return rc;
  const size_t smem = ((size_t)M * wfit::NP + wfit::NP * wfit::NP + 8 + wfit::NP + wfit::NP * 6) * sizeof(double);
  if (smem > 200 * 1024) { umpc_internal_set_error("wfitFit: more than ~1,600 trials per dataset do not fit one CTA's shared memory"); return UMPC_ERR_ARG; }
  if (smem > 48 * 1024) WCK(cudaFuncSetAttribute(wfit::fit_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  DevBuf bx, bw, bp, br, bfit;
  WCK(bx.alloc((size_t)F * M * 4 * 8)); WCK(bw.alloc((size_t)F * M * 6 * 8)); WCK(bp.alloc((size_t)F * 90 * 8)); WCK(br.alloc((size_t)F * 4));
  if (fit) WCK(bfit.alloc((size_t)F * M * 6 * 8));
  double *dx = bx.as<double>(), *dw = bw.as<double>(), *dp = bp.as<double>(), *dfit = bfit.as<double>();
  int* dr = br.as<int>();
  WCK(cudaMemcpy(dx, xdata, (size_t)F * M * 4 * 8, cudaMemcpyHostToDevice)); WCK(cudaMemcpy(dw, ws, (size_t)F * M * 6 * 8, cudaMemcpyHostToDevice));
  EvPair ev;
  WCK(cudaEventCreate(&ev.e0)); WCK(cudaEventCreate(&ev.e1));
  WCK(cudaEventRecord(ev.e0));
  wfit::fit_kernel<<<F, 128, smem>>>(M, dx, dw, dp, dfit, 1e-10, dr);
  WCK(cudaEventRecord(ev.e1));
  WCK(cudaGetLastError());
  WCK(cudaMemcpy(popts, dp, (size_t)F * 90 * 8, cudaMemcpyDeviceToHost));
  if (fit) WCK(cudaMemcpy(fit, dfit, (size_t)F * M * 6 * 8, cudaMemcpyDeviceToHost));
  if (rank) WCK(cudaMemcpy(rank, dr, (size_t)F * 4, cudaMemcpyDeviceToHost));
  if (kernel_ms) WCK(cudaEventElapsedTime(kernel_ms, ev.e0, ev.e1));
  return UMPC_OK;
}

// ---- wfitFitNoAlloc: no counterpart in the reference; for callers that cannot allocate (see include/wrenchfit_b200.h) ----
namespace {
enum { kMaxDevices = 64 };
std::mutex g_noalloc_mu;
double* g_noalloc_ptr[kMaxDevices];   // device address of wfit::g_noalloc_scratch on each device, 0 = not set up yet
}  // namespace

// Set-up, once per device, with that device current: in host code `wfit::g_noalloc_scratch` is only a host-side stand-in
// for the symbol, never a device address -- the kernel gets the address cudaGetSymbolAddress returns for THIS device.
static int noalloc_scratch(double** out) {
  int dev = -1;
  WCK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) { umpc_internal_set_error("wfitFitNoAlloc: device ordinal out of range"); return UMPC_ERR_ARG; }
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_ptr[dev]) {
    double* ptr = nullptr;
    WCK(cudaGetSymbolAddress((void**)&ptr, wfit::g_noalloc_scratch));
    if (!ptr) { umpc_internal_set_error("wfitFitNoAlloc: cudaGetSymbolAddress returned a null address"); return UMPC_ERR_CUDA; }
    g_noalloc_ptr[dev] = ptr;
  }
  *out = g_noalloc_ptr[dev];
  return UMPC_OK;
}

extern "C" int wfitFitNoAlloc(int F, int M, const double* d_xdata, const double* d_ws, double* d_popts, int* d_rank, void* stream) {
  if (F <= 0 || M <= 0 || !d_xdata || !d_ws || !d_popts) { umpc_internal_set_error("wfitFitNoAlloc: bad argument"); return UMPC_ERR_ARG; }
  if (F > WFIT_NOALLOC_MAX_DATASETS || (long long)F * M > WFIT_NOALLOC_MAX_TRIALS) {
    char msg[256];
    snprintf(msg, sizeof msg, "wfitFitNoAlloc: F = %d datasets x M = %d trials exceeds the static scratch (at most %d datasets and %d trials "
             "in all per call); use wfitFit for larger problems", F, M, (int)WFIT_NOALLOC_MAX_DATASETS, (int)WFIT_NOALLOC_MAX_TRIALS);
    umpc_internal_set_error(msg);
    return UMPC_ERR_ARG;
  }
  int rc = need_device("wfitFitNoAlloc");
  if (rc) return rc;
  double* scratch = nullptr;
  rc = noalloc_scratch(&scratch);
  if (rc) return rc;
  wfit::fit_noalloc_kernel<<<F, 128, 0, (cudaStream_t)stream>>>(F, M, d_xdata, d_ws, d_popts, 1e-10, d_rank, scratch);
  WCK(cudaGetLastError());
  return UMPC_OK;
}
