// robobee3d_b200/csrc/gqp.cu -- generic batched QP solver, ONE WARP PER QP (include/gqp_b200.h).
//
// OSQP 0.6.0's algorithm (the reference calls it through the `osqp` package: template_controllers.py:190-191,216-217,
// uprightmpc2_withWLQP.py:243-244,283-284, planar/mpc_osqp.py:104-140) on a dense symmetric P and a sparse A whose CSR
// pattern is shared by the batch:
//   update_bounds / rho_vec by constraint type   osqp.c:784-832, auxil.c:103-146
//   Ruiz equilibration + cost scaling            scaling.c:44-156
//   ADMM step                                    auxil.c:164-228, proj.c:4-14
//   residuals, termination, status               auxil.c:243-361,567-633,684-790
//   adaptive rho                                 auxil.c:13-75, osqp.c:1270-1330
//   solution un-scaling, NaN + cold start        scaling.c:177-192, auxil.c:527-562
// Mapping: the warp keeps the packed Cholesky factor of the reduced system K = P_s + sigma I + A_s' diag(rho) A_s
// (n(n+1)/2 doubles), the scaled A values and all ADMM vectors in SHARED memory; vectors are strided over the lanes,
// norms are warp-shuffle reductions, the triangular solves are column sweeps with one __syncwarp per column.  The
// scaled problem is never materialised in global memory: P_s = c D P D and A_s = E A D are applied on the fly from the
// un-scaled data, so "re-scaling" costs no global writes.  Warm start (scaled x, z, y -- NOT rescaled when D, E change,
// like the reference's workspace), D, E, c, rho and the constraint types persist per QP in global memory.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <string>
#include <vector>

#include "../../include/umpc_b200.h"
#include "../../include/gqp_b200.h"

void umpc_internal_set_error(const char* msg);  // umpc_kernels.cu

namespace gqp {

// 1 (default): the factor is inverted once per factorisation and every ADMM iteration solves with two triangular
// matrix-vector products; 0: two substitution sweeps with the pivot travelling by shuffle (A/B: build with -DGQP_EXPLICIT_INVERSE=0)
#ifndef GQP_EXPLICIT_INVERSE
#define GQP_EXPLICIT_INVERSE 1
#endif

// OSQP constants (constants.h:69-100)
#define GQ_RHO_MIN 1e-6
#define GQ_RHO_MAX 1e6
#define GQ_RHO_EQ_OVER_RHO_INEQ 1e3
#define GQ_RHO_TOL 1e-4
#define GQ_MIN_SCALING 1e-4
#define GQ_MAX_SCALING 1e4
#define GQ_ST_SOLVED 1
#define GQ_ST_SOLVED_INACCURATE 2
#define GQ_ST_MAX_ITER (-2)
#define GQ_ST_NON_CVX (-7)
#define GQ_ST_UNSOLVED (-10)

struct Dev {
  GqpSettings s;
  int nqp;
  const int *rowptr, *colidx;           // CSR pattern of A
  const int *colptr, *cscrow, *cscpos;  // CSC view: for column j the entries cscpos[k] (index into the CSR values), row cscrow[k]
  // per-QP data (un-scaled), [nqp][...]
  double *P, *Av, *q, *l, *u;
  // per-QP persistent solver state
  double *xs, *zs, *ys, *D, *E, *c, *rho;
  int* ctype;
  // results
  double *x, *y, *info;
  // factor kept between solves (LTI loop: only the bounds change): [nqp][n (n + 1) / 2 + n] and a validity flag per QP; NULL = off
  double* Lc;
  int* lvalid;
};

__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double limit_scaling(double v) {
  v = v < GQ_MIN_SCALING ? 1.0 : v;
  return v > GQ_MAX_SCALING ? GQ_MAX_SCALING : v;
}
__device__ __forceinline__ int tri(int i, int j) { return (i * (i + 1)) / 2 + j; }  // packed lower triangle, i >= j

// shared-memory map of one warp (doubles)
struct Smem {
  double *L, *As, *x, *xt, *wv, *qs, *D, *invd, *tmp, *z, *y, *ls, *us, *rv, *zt, *E;
  // the batch's pattern of A (CSR + CSC view), copied in at kernel start: every ADMM iteration walks it twice, and from
  // global memory each walk was a chain of dependent L1 hits
  int *rowptr, *colidx, *colptr, *cscrow, *cscpos;
  __device__ void carve(double* base, int n, int m, int nnz) {
    L = base; base += (size_t)n * (n + 1) / 2;
    As = base; base += nnz;
    x = base; base += n; xt = base; base += n; wv = base; base += n; qs = base; base += n;
    D = base; base += n; invd = base; base += n; tmp = base; base += n;
    z = base; base += m; y = base; base += m; ls = base; base += m; us = base; base += m;
    rv = base; base += m; zt = base; base += m; E = base; base += m;
    int* ib = (int*)base;
    rowptr = ib; ib += m + 1; colptr = ib; ib += n + 1; colidx = ib; ib += nnz; cscrow = ib; ib += nnz; cscpos = ib;
  }
};
__host__ __device__ inline size_t smem_doubles(int n, int m, int nnz) {
  return (size_t)n * (n + 1) / 2 + nnz + 7 * (size_t)n + 7 * (size_t)m + ((size_t)(m + 1) + (n + 1) + 3 * (size_t)nnz + 1) / 2;
}

// Loop over "my" entries lane, lane + 32, .. < n with a WARP-UNIFORM trip count and the bound test inside.  Written as
// `for (j = lane; j < n; j += 32)` the lanes that run one trip fewer leave the loop as a group of their own, and the
// groups were seen not to re-converge: every later warp-collective (the shuffles of the triangular solves, __syncwarp)
// then took its divergent path and the kernel executed 1.7x the instructions (ncu r02x).
#define LANE_LOOP(i, n) for (int i##_b = 0, i = lane; i##_b < (n); i##_b += 32, i += 32) if (i < (n))

// K = c D P D + sigma I + A_s' diag(rho) A_s (packed lower) and its Cholesky factor, in place in sm.L
__device__ void form_and_factor(const Dev& d, const Smem& sm, const double* __restrict__ P, double c, int lane) {
  const int n = d.s.n, m = d.s.m;
  for (int i = 0; i < n; ++i) {
    const double cdi = c * sm.D[i];
    LANE_LOOP(j, i + 1) sm.L[tri(i, j)] = cdi * sm.D[j] * P[(size_t)i * n + j] + (i == j ? d.s.sigma : 0.0);
  }
  __syncwarp();
  // + A_s' diag(rho) A_s: rows one after the other, the row's entry pairs (a >= b) spread over the lanes -- within one row
  // every pair is a different entry of K, so no atomics (the shared-memory double atomicAdd is a compare-and-swap loop,
  // and its retries split the warp)
  for (int r = 0; r < m; ++r) {
    const int r0 = sm.rowptr[r], cnt = sm.rowptr[r + 1] - r0, npair = (cnt * (cnt + 1)) >> 1;
    const double rho = sm.rv[r];
    for (int pb = 0; pb < npair; pb += 32) {
      const int p = pb + lane;
      if (p < npair) {
        int a = (int)((sqrtf(8.0f * (float)p + 1.0f) - 1.0f) * 0.5f);   // pair p = (a, b), b <= a; branch-free decode
        a += (((a + 1) * (a + 2)) >> 1) <= p;
        a -= ((a * (a + 1)) >> 1) > p;
        const int b = p - ((a * (a + 1)) >> 1);
        sm.L[tri(sm.colidx[r0 + a], sm.colidx[r0 + b])] += rho * sm.As[r0 + a] * sm.As[r0 + b];   // columns increase along a row
      }
    }
    __syncwarp();
  }
  for (int k = 0; k < n; ++k) {
    const double dk = sqrt(sm.L[tri(k, k)]);
    const double inv = 1.0 / dk;
    __syncwarp();
    if (lane == 0) { sm.L[tri(k, k)] = dk; sm.invd[k] = inv; }
    for (int ib = k + 1; ib < n; ib += 32) {
      const int i = ib + lane;
      if (i < n) sm.L[tri(i, k)] *= inv;
    }
    __syncwarp();
    for (int ib = k + 1; ib < n; ib += 32) {
      const int i = ib + lane;
      if (i < n) {
        const double lik = sm.L[tri(i, k)];
        double* row = sm.L + tri(i, 0);
        for (int j = k + 1; j <= i; ++j) row[j] -= lik * sm.L[tri(j, k)];
      }
    }
    __syncwarp();
  }
#if GQP_EXPLICIT_INVERSE
  // L <- L^-1 in place, row by row: Linv[i][j] = -(1 / L[i][i]) sum_{k = j}^{i-1} L[i][k] Linv[k][j].  Row i needs its own
  // original entries (parked in sm.tmp) and the rows above it, which are already inverted; the entries of a row are
  // independent, one lane each.  The ADMM loop then solves with two triangular MATRIX-VECTOR products (gqp_inv_solve)
  // instead of two substitution sweeps: the same n^2 multiply-adds without the chain of n dependent pivot steps.
  if (n <= 128) {
    for (int i = 0; i < n; ++i) {
      const double* rowc = sm.L + tri(i, 0);
      LANE_LOOP(k, i) sm.tmp[k] = rowc[k];
      __syncwarp();
      const double di = sm.invd[i];
      for (int jb = 0; jb < i; jb += 32) {
        const int j = jb + lane;
        if (j < i) {
          double acc = 0.0;
          int a = tri(j, j);                                  // Linv[k][j], k = j .. i-1: tri(k, j) = tri(k-1, j) + k
          for (int k = j; k < i; ++k) { acc += sm.tmp[k] * sm.L[a]; a += k + 1; }
          sm.L[tri(i, j)] = -di * acc;
        }
      }
      if (lane == 0) sm.L[tri(i, i)] = di;
      __syncwarp();
    }
  }
#endif
}

// sm.xt <- K^-1 sm.xt
__device__ __forceinline__ void chol_solve(const Smem& sm, int n, int lane) {
  for (int k = 0; k < n; ++k) {
    __syncwarp();
    const double wk = sm.xt[k] * sm.invd[k];
    if (lane == 0) sm.wv[k] = wk;
    for (int i = k + 1 + lane; i < n; i += 32) sm.xt[i] -= sm.L[tri(i, k)] * wk;
  }
  for (int k = n - 1; k >= 0; --k) {
    __syncwarp();
    const double xk = sm.wv[k] * sm.invd[k];
    if (lane == 0) sm.xt[k] = xk;
    const double* row = sm.L + tri(k, 0);
    LANE_LOOP(j, k) sm.wv[j] -= row[j] * xk;
  }
  __syncwarp();
}

// The same two sweeps with the vector in REGISTERS (lane l owns entries l, l + 32, ..: NR = ceil(n / 32) of them): the
// pivot entry travels by shuffle, each lane updates its own entries.  Same operations in the same order as chol_solve
// (bit-identical results), a tenth of the instructions: the shared-memory version spends ~60 instructions per column on
// loop control, index arithmetic and the round trip of the vector (ncu r02v: its two inner loops were 70 % of all
// executed instructions of the kernel).
template <int NR>
__device__ __forceinline__ void chol_solve_reg(const Smem& sm, int n, int lane) {
  double v[NR];
  int base[NR];
#pragma unroll
  for (int r = 0; r < NR; ++r) {
    const int i = lane + 32 * r;
    v[r] = i < n ? sm.xt[i] : 0.0;
    base[r] = (i * (i + 1)) / 2;
  }
#pragma unroll
  for (int s = 0; s < NR; ++s) {            // forward: L w = b, column by column
    const int kend = min(32, n - 32 * s);
    for (int kk = 0; kk < kend; ++kk) {
      const int k = 32 * s + kk;
      const double wk = __shfl_sync(0xffffffffu, v[s], kk) * sm.invd[k];
      if (lane == kk) v[s] = wk;
#pragma unroll
      for (int r = s; r < NR; ++r) {
        const int i = lane + 32 * r;
        if (i > k && i < n) v[r] -= sm.L[base[r] + k] * wk;
      }
    }
  }
#pragma unroll
  for (int s = NR - 1; s >= 0; --s) {       // backward: L' x = w, row k of L holds column k of L'
    const int kend = min(32, n - 32 * s);
    for (int kk = kend - 1; kk >= 0; --kk) {
      const int k = 32 * s + kk;
      const double xk = __shfl_sync(0xffffffffu, v[s], kk) * sm.invd[k];
      if (lane == kk) v[s] = xk;
      const double* row = sm.L + (k * (k + 1)) / 2;
#pragma unroll
      for (int r = 0; r <= s; ++r) {
        const int j = lane + 32 * r;
        if (j < k) v[r] -= row[j] * xk;
      }
    }
  }
#pragma unroll
  for (int r = 0; r < NR; ++r) {
    const int i = lane + 32 * r;
    if (i < n) sm.xt[i] = v[r];
  }
  __syncwarp();
}
// sm.xt <- K^-1 sm.xt with sm.L = L^-1 (packed lower): y = Linv b (lane i: row i), x = Linv' y (lane j: column j, conflict-free)
template <int NR>
__device__ __forceinline__ void gqp_inv_solve(const Smem& sm, int n, int lane) {
  double acc[NR];
  int base[NR];
#pragma unroll
  for (int r = 0; r < NR; ++r) { const int i = lane + 32 * r; acc[r] = 0.0; base[r] = (i * (i + 1)) / 2; }
  for (int j = 0; j < n; ++j) {
    const double bj = sm.xt[j];                             // broadcast
#pragma unroll
    for (int r = 0; r < NR; ++r) {
      const int i = lane + 32 * r;
      if (j <= i && i < n) acc[r] += sm.L[base[r] + j] * bj;
    }
  }
#pragma unroll
  for (int r = 0; r < NR; ++r) { const int i = lane + 32 * r; if (i < n) sm.wv[i] = acc[r]; acc[r] = 0.0; }
  __syncwarp();
  int rowbase = 0;                                          // tri(i, 0)
  for (int i = 0; i < n; ++i) {
    const double yi = sm.wv[i];
#pragma unroll
    for (int r = 0; r < NR; ++r) {
      const int j = lane + 32 * r;
      if (j <= i) acc[r] += sm.L[rowbase + j] * yi;
    }
    rowbase += i + 1;
  }
#pragma unroll
  for (int r = 0; r < NR; ++r) { const int j = lane + 32 * r; if (j < n) sm.xt[j] = acc[r]; }
  __syncwarp();
}
__device__ __forceinline__ void chol_solve_any(const Smem& sm, int n, int lane) {
  __syncwarp();
#if GQP_EXPLICIT_INVERSE
  switch ((n + 31) >> 5) {
    case 1: gqp_inv_solve<1>(sm, n, lane); return;
    case 2: gqp_inv_solve<2>(sm, n, lane); return;
    case 3: gqp_inv_solve<3>(sm, n, lane); return;
    case 4: gqp_inv_solve<4>(sm, n, lane); return;
    default: break;
  }
#endif
  switch ((n + 31) >> 5) {
    case 1: chol_solve_reg<1>(sm, n, lane); break;
    case 2: chol_solve_reg<2>(sm, n, lane); break;
    case 3: chol_solve_reg<3>(sm, n, lane); break;
    case 4: chol_solve_reg<4>(sm, n, lane); break;
    default: chol_solve(sm, n, lane);
  }
}

// per-row rho from the constraint type (set_rho_vec / update_rho_vec, auxil.c:77-146)
__device__ __forceinline__ double rho_of(int type, double rho) {
  return type == -1 ? GQ_RHO_MIN : (type == 1 ? GQ_RHO_EQ_OVER_RHO_INEQ * rho : rho);
}

// un-scaled residual norms and the scaled ones adaptive rho needs (update_info + check_termination, auxil.c:243-361,684-790;
// compute_rho_estimate, auxil.c:13-75).  Uses sm.tmp / sm.zt as scratch.
struct Resid {
  double pri, dua, eps_pri_n, eps_dua_n;  // un-scaled residuals and the norms the relative tolerances multiply
  double pri_s, pn_s, dua_s, dn_s;         // scaled quantities of the rho estimate
};
__device__ Resid residuals(const Dev& d, const Smem& sm, const double* __restrict__ P, double c, int lane) {
  const int n = d.s.n, m = d.s.m;
  Resid r;
  double pri = 0, nz = 0, nAx = 0, pri_s = 0, pn_s = 0;
  LANE_LOOP(i, m) {
    double ax = 0;
    for (int a = sm.rowptr[i]; a < sm.rowptr[i + 1]; ++a) ax += sm.As[a] * sm.x[sm.colidx[a]];
    const double einv = 1.0 / sm.E[i], rp = ax - sm.z[i];
    pri = fmax(pri, fabs(einv * rp)); nz = fmax(nz, fabs(einv * sm.z[i])); nAx = fmax(nAx, fabs(einv * ax));
    pri_s = fmax(pri_s, fabs(rp)); pn_s = fmax(pn_s, fmax(fabs(sm.z[i]), fabs(ax)));
  }
  // tmp = D .* x  (for P_s x = c D (P (D x)))
  LANE_LOOP(j, n) sm.tmp[j] = sm.D[j] * sm.x[j];
  __syncwarp();
  double dua = 0, nq = 0, nAty = 0, nPx = 0, dua_s = 0, dn_s = 0;
  LANE_LOOP(j, n) {
    double px = 0, aty = 0;
    const double* prow = P + (size_t)j * n;
    for (int i = 0; i < n; ++i) px += prow[i] * sm.tmp[i];
    px *= c * sm.D[j];
    for (int k = sm.colptr[j]; k < sm.colptr[j + 1]; ++k) aty += sm.As[sm.cscpos[k]] * sm.y[sm.cscrow[k]];
    const double dinv = 1.0 / sm.D[j], rd = sm.qs[j] + px + aty;
    dua = fmax(dua, fabs(dinv * rd)); nq = fmax(nq, fabs(dinv * sm.qs[j])); nAty = fmax(nAty, fabs(dinv * aty)); nPx = fmax(nPx, fabs(dinv * px));
    dua_s = fmax(dua_s, fabs(rd)); dn_s = fmax(dn_s, fmax(fabs(sm.qs[j]), fmax(fabs(aty), fabs(px))));
  }
  const double cinv = 1.0 / c;
  r.pri = warp_max(pri); r.dua = cinv * warp_max(dua);
  r.eps_pri_n = fmax(warp_max(nz), warp_max(nAx));
  r.eps_dua_n = cinv * fmax(fmax(warp_max(nq), warp_max(nAty)), warp_max(nPx));
  r.pri_s = warp_max(pri_s); r.pn_s = warp_max(pn_s); r.dua_s = warp_max(dua_s); r.dn_s = warp_max(dn_s);
  return r;
}

__global__ void __launch_bounds__(32) gqp_solve_kernel(const __grid_constant__ Dev d, int rescale) {
  extern __shared__ double smem_d[];
  const int qp = blockIdx.x, lane = threadIdx.x;
  if (qp >= d.nqp) return;
  const int n = d.s.n, m = d.s.m, nnz = d.s.nnz;
  Smem sm;
  sm.carve(smem_d, n, m, nnz);
  LANE_LOOP(i, m + 1) sm.rowptr[i] = d.rowptr[i];
  LANE_LOOP(i, n + 1) sm.colptr[i] = d.colptr[i];
  LANE_LOOP(i, nnz) { sm.colidx[i] = d.colidx[i]; sm.cscrow[i] = d.cscrow[i]; sm.cscpos[i] = d.cscpos[i]; }
  __syncwarp();
  const double* P = d.P + (size_t)qp * n * n;
  const double* Av = d.Av + (size_t)qp * nnz;
  const double* q0 = d.q + (size_t)qp * n;
  const double* l0 = d.l + (size_t)qp * m;
  const double* u0 = d.u + (size_t)qp * m;
  double* gD = d.D + (size_t)qp * n; double* gE = d.E + (size_t)qp * m;
  int* gct = d.ctype + (size_t)qp * m;
  double c = d.c[qp], rho = d.rho[qp];
  const double INF = d.s.infty;
  // ---- osqp_update_bounds with the OLD scaling: constraint types (auxil.c:103-146) ---------------------------------
  int types_changed = 0;
  LANE_LOOP(i, m) {
    const double eo = gE[i];
    const double lo = fmax(l0[i], -INF), hi = fmin(u0[i], INF);
    const double ls = eo * lo, us = eo * hi;
    int type;
    if (ls < -INF * GQ_MIN_SCALING && us > INF * GQ_MIN_SCALING) type = -1;
    else if (us - ls < GQ_RHO_TOL) type = 1;
    else type = 0;
    types_changed |= gct[i] != type;
    gct[i] = type;
    sm.E[i] = eo;
  }
  types_changed = __any_sync(0xffffffffu, types_changed);
  LANE_LOOP(j, n) sm.D[j] = gD[j];
  __syncwarp();
  // ---- osqp_update_P_A: re-equilibrate (scale_data, scaling.c:44-156) ------------------------------------------------
  if (rescale) {
    LANE_LOOP(j, n) sm.D[j] = 1.0;
    LANE_LOOP(i, m) sm.E[i] = 1.0;
    c = 1.0;
    __syncwarp();
    // |P| (lower triangle) parked in the factor's space during the passes
    for (int i = 0; i < n; ++i)
      LANE_LOOP(j, i + 1) sm.L[tri(i, j)] = fabs(P[(size_t)i * n + j]);
    __syncwarp();
    for (int it = 0; it < d.s.scaling_iters; ++it) {
      // column norms of [P_s; A_s], row norms of A_s with the current D, E, c
      LANE_LOOP(j, n) {
        double v = 0;
        for (int i = 0; i < n; ++i) v = fmax(v, sm.L[i >= j ? tri(i, j) : tri(j, i)] * sm.D[i]);
        v *= c * sm.D[j];
        double va = 0;
        for (int k = sm.colptr[j]; k < sm.colptr[j + 1]; ++k) va = fmax(va, fabs(Av[sm.cscpos[k]]) * sm.E[sm.cscrow[k]]);
        sm.tmp[j] = 1.0 / sqrt(limit_scaling(fmax(v, va * sm.D[j])));
      }
      LANE_LOOP(i, m) {
        double v = 0;
        for (int a = sm.rowptr[i]; a < sm.rowptr[i + 1]; ++a) v = fmax(v, fabs(Av[a]) * sm.D[sm.colidx[a]]);
        sm.zt[i] = 1.0 / sqrt(limit_scaling(v * sm.E[i]));
      }
      __syncwarp();
      LANE_LOOP(j, n) sm.D[j] *= sm.tmp[j];
      LANE_LOOP(i, m) sm.E[i] *= sm.zt[i];
      __syncwarp();
      // cost normalisation (scaling.c:100-131): mean column norm of P_s, |q_s|_inf
      double csum = 0, nq = 0;
      LANE_LOOP(j, n) {
        double v = 0;
        for (int i = 0; i < n; ++i) v = fmax(v, sm.L[i >= j ? tri(i, j) : tri(j, i)] * sm.D[i]);
        csum += c * sm.D[j] * v;
        nq = fmax(nq, fabs(c * sm.D[j] * q0[j]));
      }
      csum = warp_sum(csum) / (double)n;
      nq = limit_scaling(warp_max(nq));
      double ct = limit_scaling(fmax(csum, nq));
      c *= 1.0 / ct;
      __syncwarp();
    }
    LANE_LOOP(j, n) gD[j] = sm.D[j];
    LANE_LOOP(i, m) gE[i] = sm.E[i];
  }
  // ---- scaled data ---------------------------------------------------------------------------------------------------
  LANE_LOOP(i, m) {
    const double e = sm.E[i];
    sm.ls[i] = e * fmax(l0[i], -INF); sm.us[i] = e * fmin(u0[i], INF);
    sm.rv[i] = rho_of(gct[i], rho);
    for (int a = sm.rowptr[i]; a < sm.rowptr[i + 1]; ++a) sm.As[a] = e * Av[a] * sm.D[sm.colidx[a]];
    sm.z[i] = d.zs[(size_t)qp * m + i]; sm.y[i] = d.ys[(size_t)qp * m + i];
  }
  LANE_LOOP(j, n) { sm.qs[j] = c * sm.D[j] * q0[j]; sm.x[j] = d.xs[(size_t)qp * n + j]; }
  __syncwarp();
  // The factor depends on P, A, the scaling, rho and the constraint types only: when the caller says that just q / l / u
  // changed (rescale = 0: prob.update(l=, u=), the LTI loop of mpc_osqp.py:137-140) and neither a type nor rho moved, the
  // factor of the previous solve is loaded back instead of being formed, factorised and inverted again (OSQP itself only
  // refactors on update_P_A / update_rho).
  const int ntri = (n * (n + 1)) / 2;
  double* Lq = d.Lc ? d.Lc + (size_t)qp * (ntri + n) : nullptr;
  bool factor_dirty = true;
  if (Lq && !rescale && !types_changed && d.lvalid[qp]) {
    LANE_LOOP(k, ntri) sm.L[k] = Lq[k];
    LANE_LOOP(k, n) sm.invd[k] = Lq[ntri + k];
    __syncwarp();
    factor_dirty = false;
  } else {
    form_and_factor(d, sm, P, c, lane);
  }
  // ---- ADMM (osqp.c:352-500) -----------------------------------------------------------------------------------------
  const double alpha = d.s.alpha, sigma = d.s.sigma;
  int iter = 0, status = GQ_ST_UNSOLVED, rho_updates = 0;
  bool checked = false;
  Resid res;
  for (iter = 1; iter <= d.s.max_iter; ++iter) {
    // (loops over the lanes' entries have WARP-UNIFORM trip counts with the bound test inside: a `for (j = lane; j < n; ..)`
    // leaves the lanes that ran one trip fewer in a group of their own, and the groups do not re-converge on their own --
    // the shuffles of the triangular solves then ran once per group through the divergent-shuffle slow path, ncu r02x)
    for (int jb = 0; jb < n; jb += 32) {
      const int j = jb + lane;
      if (j < n) {
        double v = sigma * sm.x[j] - sm.qs[j];
        for (int k = sm.colptr[j]; k < sm.colptr[j + 1]; ++k) {
          const int r = sm.cscrow[k];
          v += sm.As[sm.cscpos[k]] * (sm.rv[r] * sm.z[r] - sm.y[r]);
        }
        sm.xt[j] = v;
      }
    }
    chol_solve_any(sm, n, lane);
    for (int ib = 0; ib < m; ib += 32) {
      const int i = ib + lane;
      if (i < m) {
        double zt = 0;
        for (int a = sm.rowptr[i]; a < sm.rowptr[i + 1]; ++a) zt += sm.As[a] * sm.xt[sm.colidx[a]];
        const double rv = sm.rv[i], zi = sm.z[i], yi = sm.y[i];
        const double zr = alpha * zt + (1.0 - alpha) * zi;
        const double zn = fmin(fmax(zr + yi / rv, sm.ls[i]), sm.us[i]);
        sm.y[i] = yi + rv * (zr - zn);
        sm.z[i] = zn;
      }
    }
    for (int jb = 0; jb < n; jb += 32) {
      const int j = jb + lane;
      if (j < n) sm.x[j] = alpha * sm.xt[j] + (1.0 - alpha) * sm.x[j];
    }
    __syncwarp();
    checked = d.s.check_termination && (iter % d.s.check_termination == 0);
    const bool adapt = d.s.adaptive_rho_interval && (iter % d.s.adaptive_rho_interval == 0);
    if (checked || adapt) res = residuals(d, sm, P, c, lane);
    if (checked) {
      if (res.pri > INF || res.dua > INF) { status = GQ_ST_NON_CVX; break; }
      if (res.pri < d.s.eps_abs + d.s.eps_rel * res.eps_pri_n && res.dua < d.s.eps_abs + d.s.eps_rel * res.eps_dua_n) { status = GQ_ST_SOLVED; break; }
    }
    if (adapt) {
      const double pr = res.pri_s / (res.pn_s + 1e-10), dr = res.dua_s / (res.dn_s + 1e-10);
      double est = rho * sqrt(pr / (dr + 1e-10));
      est = fmin(fmax(est, GQ_RHO_MIN), GQ_RHO_MAX);
      if (est > rho * d.s.adaptive_rho_tol || est < rho / d.s.adaptive_rho_tol) {
        rho = est; ++rho_updates;
        LANE_LOOP(i, m) sm.rv[i] = rho_of(gct[i], rho);
        __syncwarp();
        form_and_factor(d, sm, P, c, lane);
        factor_dirty = true;
      }
    }
  }
  if (status == GQ_ST_UNSOLVED) {
    iter = d.s.max_iter;
    if (!checked) res = residuals(d, sm, P, c, lane);
    if (res.pri > INF || res.dua > INF) status = GQ_ST_NON_CVX;
    else if (res.pri < d.s.eps_abs + d.s.eps_rel * res.eps_pri_n && res.dua < d.s.eps_abs + d.s.eps_rel * res.eps_dua_n) status = GQ_ST_SOLVED;
    else if (res.pri < 10 * d.s.eps_abs + 10 * d.s.eps_rel * res.eps_pri_n && res.dua < 10 * d.s.eps_abs + 10 * d.s.eps_rel * res.eps_dua_n)
      status = GQ_ST_SOLVED_INACCURATE;
    else status = GQ_ST_MAX_ITER;
  }
  // ---- store_solution (auxil.c:527-562) + persistent state -----------------------------------------------------------
  const bool bad = status == GQ_ST_NON_CVX;
  const double cinv = 1.0 / c;
  LANE_LOOP(j, n) {
    d.x[(size_t)qp * n + j] = bad ? nan("") : sm.D[j] * sm.x[j];
    d.xs[(size_t)qp * n + j] = bad ? 0.0 : sm.x[j];
  }
  LANE_LOOP(i, m) {
    d.y[(size_t)qp * m + i] = bad ? nan("") : sm.E[i] * sm.y[i] * cinv;
    d.zs[(size_t)qp * m + i] = bad ? 0.0 : sm.z[i];
    d.ys[(size_t)qp * m + i] = bad ? 0.0 : sm.y[i];
  }
  if (Lq && factor_dirty) {
    LANE_LOOP(k, ntri) Lq[k] = sm.L[k];
    LANE_LOOP(k, n) Lq[ntri + k] = sm.invd[k];
    if (lane == 0) d.lvalid[qp] = 1;
  }
  if (lane == 0) {
    d.c[qp] = c; d.rho[qp] = rho;
    double* inf = d.info + (size_t)qp * 6;
    inf[0] = iter; inf[1] = status; inf[2] = res.pri; inf[3] = res.dua; inf[4] = rho; inf[5] = rho_updates;
  }
}

// LTI closed loop (planar/mpc_osqp.py:131-140): x0 <- Ad x0 + Bd u0*, l[:nx] = u[:nx] = -x0
__global__ void gqp_lti_advance_kernel(const __grid_constant__ Dev d, int nx, int nu, int u_off, const double* __restrict__ Ad,
                                       const double* __restrict__ Bd, double* __restrict__ x0, double* __restrict__ ulog,
                                       double* __restrict__ iters_sum) {
  const int qp = blockIdx.x * blockDim.x + threadIdx.x;
  if (qp >= d.nqp) return;
  const double* xs = d.x + (size_t)qp * d.s.n;
  double* x = x0 + (size_t)qp * nx;
  double xn[32];
  for (int i = 0; i < nx; ++i) {
    double v = 0;
    for (int j = 0; j < nx; ++j) v += Ad[i * nx + j] * x[j];
    for (int j = 0; j < nu; ++j) v += Bd[i * nu + j] * xs[u_off + j];
    xn[i] = v;
  }
  for (int i = 0; i < nx; ++i) {
    x[i] = xn[i];
    d.l[(size_t)qp * d.s.m + i] = -xn[i];
    d.u[(size_t)qp * d.s.m + i] = -xn[i];
  }
  if (ulog) for (int j = 0; j < nu; ++j) ulog[(size_t)qp * nu + j] = xs[u_off + j];
  if (iters_sum) atomicAdd(iters_sum, d.info[(size_t)qp * 6]);
}

}  // namespace gqp

// ================================ host side: C-ABI ============================================
using gqp::Dev;

#define GCK(call)                                                                         \
  do {                                                                                    \
    cudaError_t e_ = (call);                                                              \
    if (e_ != cudaSuccess) {                                                              \
      umpc_internal_set_error((std::string(#call) + ": " + cudaGetErrorString(e_)).c_str()); \
      return UMPC_ERR_CUDA;                                                               \
    }                                                                                     \
  } while (0)

struct GqpBatch {
  int nqp = 0, device = 0;
  Dev d;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  bool timed = false;
  size_t smem = 0;
  std::vector<void*> allocs;
};

extern "C" void gqpDefaults(GqpSettings* s, int n, int m, int nnz) {
  memset(s, 0, sizeof(*s));
  s->n = n; s->m = m; s->nnz = nnz;
  s->max_iter = 4000; s->check_termination = 25; s->adaptive_rho_interval = 100; s->scaling_iters = 10;
  s->rho = 0.1; s->sigma = 1e-6; s->alpha = 1.6; s->eps_abs = 1e-3; s->eps_rel = 1e-3; s->adaptive_rho_tol = 5.0;
  s->infty = 1e30;
}

template <class T>
static int galloc(GqpBatch* b, T** p, size_t count) {
  GCK(cudaMalloc((void**)p, count * sizeof(T)));
  b->allocs.push_back((void*)*p);
  GCK(cudaMemsetAsync(*p, 0, count * sizeof(T), b->stream));
  return UMPC_OK;
}

extern "C" void gqpBatchDestroy(GqpBatch* b) {
  if (!b) return;
  cudaSetDevice(b->device);
  if (b->stream) cudaStreamSynchronize(b->stream);
  for (void* p : b->allocs) cudaFree(p);
  if (b->ev0) cudaEventDestroy(b->ev0);
  if (b->ev1) cudaEventDestroy(b->ev1);
  if (b->stream) cudaStreamDestroy(b->stream);
  cudaGetLastError();
  delete b;
}

__global__ void gqp_fill_kernel(double* p, size_t n, double v) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}

static int gqp_fill(GqpBatch* b, double* p, size_t n, double v) {
  gqp_fill_kernel<<<(unsigned)((n + 255) / 256 > 1184 ? 1184 : (n + 255) / 256), 256, 0, b->stream>>>(p, n, v);
  GCK(cudaGetLastError());
  return UMPC_OK;
}

extern "C" int gqpBatchReset(GqpBatch* b) {
  if (!b) return UMPC_ERR_ARG;
  GCK(cudaSetDevice(b->device));
  const Dev& d = b->d;
  const size_t B = b->nqp;
  GCK(cudaMemsetAsync(d.xs, 0, B * d.s.n * 8, b->stream));
  GCK(cudaMemsetAsync(d.zs, 0, B * d.s.m * 8, b->stream));
  GCK(cudaMemsetAsync(d.ys, 0, B * d.s.m * 8, b->stream));
  GCK(cudaMemsetAsync(d.ctype, 0, B * d.s.m * 4, b->stream));
  if (d.lvalid) GCK(cudaMemsetAsync(d.lvalid, 0, B * 4, b->stream));
  int rc;
  if ((rc = gqp_fill(b, d.D, B * d.s.n, 1.0))) return rc;
  if ((rc = gqp_fill(b, d.E, B * d.s.m, 1.0))) return rc;
  if ((rc = gqp_fill(b, d.c, B, 1.0))) return rc;
  if ((rc = gqp_fill(b, d.rho, B, d.s.rho))) return rc;
  GCK(cudaStreamSynchronize(b->stream));
  return UMPC_OK;
}

static int gqp_create_body(GqpBatch* b, const GqpSettings* s, const int* rowptr, const int* colidx) {
  Dev& d = b->d;
  memset(&d, 0, sizeof(d));
  d.s = *s; d.nqp = b->nqp;
  const int n = s->n, m = s->m, nnz = s->nnz;
  GCK(cudaStreamCreateWithFlags(&b->stream, cudaStreamNonBlocking));
  GCK(cudaEventCreate(&b->ev0));
  GCK(cudaEventCreate(&b->ev1));
  // CSC view of the pattern
  std::vector<int> colptr(n + 1, 0), cscrow(nnz), cscpos(nnz);
  for (int r = 0; r < m; ++r)
    for (int a = rowptr[r]; a < rowptr[r + 1]; ++a) {
      if (colidx[a] < 0 || colidx[a] >= n || (a > rowptr[r] && colidx[a] <= colidx[a - 1])) {
        umpc_internal_set_error("gqpBatchCreate: bad CSR pattern (column out of range or not increasing inside a row)");
        return UMPC_ERR_ARG;
      }
      colptr[colidx[a] + 1]++;
    }
  for (int j = 0; j < n; ++j) colptr[j + 1] += colptr[j];
  {
    std::vector<int> fillp(colptr.begin(), colptr.end() - 1);
    for (int r = 0; r < m; ++r)
      for (int a = rowptr[r]; a < rowptr[r + 1]; ++a) { const int k = fillp[colidx[a]]++; cscrow[k] = r; cscpos[k] = a; }
  }
  int *d_rowptr, *d_colidx, *d_colptr, *d_cscrow, *d_cscpos;
  int rc;
  if ((rc = galloc(b, &d_rowptr, m + 1)) || (rc = galloc(b, &d_colidx, nnz)) || (rc = galloc(b, &d_colptr, n + 1)) ||
      (rc = galloc(b, &d_cscrow, nnz)) || (rc = galloc(b, &d_cscpos, nnz))) return rc;
  GCK(cudaMemcpyAsync(d_rowptr, rowptr, (m + 1) * 4, cudaMemcpyHostToDevice, b->stream));
  GCK(cudaMemcpyAsync(d_colidx, colidx, nnz * 4, cudaMemcpyHostToDevice, b->stream));
  GCK(cudaMemcpyAsync(d_colptr, colptr.data(), (n + 1) * 4, cudaMemcpyHostToDevice, b->stream));
  GCK(cudaMemcpyAsync(d_cscrow, cscrow.data(), nnz * 4, cudaMemcpyHostToDevice, b->stream));
  GCK(cudaMemcpyAsync(d_cscpos, cscpos.data(), nnz * 4, cudaMemcpyHostToDevice, b->stream));
  GCK(cudaStreamSynchronize(b->stream));  // the host vectors go out of scope
  d.rowptr = d_rowptr; d.colidx = d_colidx; d.colptr = d_colptr; d.cscrow = d_cscrow; d.cscpos = d_cscpos;
  const size_t B = b->nqp;
  if ((rc = galloc(b, &d.P, B * n * n)) || (rc = galloc(b, &d.Av, B * nnz)) || (rc = galloc(b, &d.q, B * n)) ||
      (rc = galloc(b, &d.l, B * m)) || (rc = galloc(b, &d.u, B * m)) || (rc = galloc(b, &d.xs, B * n)) ||
      (rc = galloc(b, &d.zs, B * m)) || (rc = galloc(b, &d.ys, B * m)) || (rc = galloc(b, &d.D, B * n)) ||
      (rc = galloc(b, &d.E, B * m)) || (rc = galloc(b, &d.c, B)) || (rc = galloc(b, &d.rho, B)) ||
      (rc = galloc(b, &d.ctype, B * m)) || (rc = galloc(b, &d.x, B * n)) || (rc = galloc(b, &d.y, B * m)) ||
      (rc = galloc(b, &d.info, B * 6))) return rc;
  b->smem = gqp::smem_doubles(n, m, nnz) * sizeof(double);
  if (b->smem > 227 * 1024) {
    umpc_internal_set_error("gqpBatchCreate: QP too large for one warp's shared memory (n(n+1)/2 + nnz + 7n + 7m doubles <= 227 KB)");
    return UMPC_ERR_ARG;
  }
  if (b->smem > 48 * 1024) GCK(cudaFuncSetAttribute(gqp::gqp_solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)b->smem));
  return gqpBatchReset(b);
}

extern "C" int gqpBatchCreate(GqpBatch** out, int nqp, int device, const GqpSettings* s, const int* rowptr, const int* colidx) {
  if (!out || nqp <= 0 || !s || !rowptr || !colidx || s->n <= 0 || s->m <= 0 || s->nnz <= 0 || rowptr[s->m] != s->nnz) {
    umpc_internal_set_error("gqpBatchCreate: bad argument");
    return UMPC_ERR_ARG;
  }
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    umpc_internal_set_error((std::string("gqpBatchCreate: no CUDA device (") + cudaGetErrorString(e) + "); this library has no CPU path").c_str());
    return UMPC_ERR_CUDA;
  }
  GCK(cudaSetDevice(device));
  GqpBatch* b = new GqpBatch();
  b->nqp = nqp; b->device = device;
  const int rc = gqp_create_body(b, s, rowptr, colidx);
  if (rc) { gqpBatchDestroy(b); return rc; }
  *out = b;
  return UMPC_OK;
}

__global__ void gqp_bcast_kernel(double* dst, size_t per, size_t nqp) {  // dst[0..per) replicated to every QP
  const size_t tot = per * nqp;
  for (size_t i = per + blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < tot; i += (size_t)gridDim.x * blockDim.x) dst[i] = dst[i % per];
}

extern "C" int gqpBatchSetData(GqpBatch* b, const double* P, const double* Av, const double* q, const double* l, const double* u,
                               int broadcast) {
  if (!b) return UMPC_ERR_ARG;
  GCK(cudaSetDevice(b->device));
  const Dev& d = b->d;
  struct { const double* h; double* dv; size_t per; } arr[5] = {{P, d.P, (size_t)d.s.n * d.s.n}, {Av, d.Av, (size_t)d.s.nnz},
                                                                 {q, d.q, (size_t)d.s.n}, {l, d.l, (size_t)d.s.m}, {u, d.u, (size_t)d.s.m}};
  if ((P || Av) && d.lvalid) GCK(cudaMemsetAsync(d.lvalid, 0, (size_t)b->nqp * 4, b->stream));   // kept factors are stale
  for (auto& a : arr) {
    if (!a.h) continue;
    GCK(cudaMemcpyAsync(a.dv, a.h, a.per * (broadcast ? 1 : (size_t)b->nqp) * 8, cudaMemcpyHostToDevice, b->stream));
    if (broadcast && b->nqp > 1) {
      gqp_bcast_kernel<<<592, 256, 0, b->stream>>>(a.dv, a.per, (size_t)b->nqp);
      GCK(cudaGetLastError());
    }
  }
  GCK(cudaStreamSynchronize(b->stream));
  return UMPC_OK;
}

static int gqp_launch(GqpBatch* b, int rescale) {
  gqp::gqp_solve_kernel<<<b->nqp, 32, b->smem, b->stream>>>(b->d, rescale);
  GCK(cudaGetLastError());
  return UMPC_OK;
}

extern "C" int gqpBatchSolve(GqpBatch* b, int rescale) {
  if (!b) return UMPC_ERR_ARG;
  GCK(cudaSetDevice(b->device));
  GCK(cudaEventRecord(b->ev0, b->stream));
  int rc = gqp_launch(b, rescale);
  if (rc) return rc;
  GCK(cudaEventRecord(b->ev1, b->stream));
  b->timed = true;
  return UMPC_OK;
}

extern "C" int gqpBatchSync(GqpBatch* b) {
  if (!b) return UMPC_ERR_ARG;
  GCK(cudaStreamSynchronize(b->stream));
  return UMPC_OK;
}

extern "C" float gqpBatchLastKernelMs(GqpBatch* b) {
  if (!b || !b->timed) return -1.0f;
  float ms = -1.0f;
  if (cudaEventSynchronize(b->ev1) != cudaSuccess) return -1.0f;
  cudaEventElapsedTime(&ms, b->ev0, b->ev1);
  return ms;
}

extern "C" int gqpBatchGetSolution(GqpBatch* b, double* x, double* y, double* info) {
  if (!b) return UMPC_ERR_ARG;
  GCK(cudaSetDevice(b->device));
  const Dev& d = b->d;
  const size_t B = b->nqp;
  if (x) GCK(cudaMemcpyAsync(x, d.x, B * d.s.n * 8, cudaMemcpyDeviceToHost, b->stream));
  if (y) GCK(cudaMemcpyAsync(y, d.y, B * d.s.m * 8, cudaMemcpyDeviceToHost, b->stream));
  if (info) GCK(cudaMemcpyAsync(info, d.info, B * 6 * 8, cudaMemcpyDeviceToHost, b->stream));
  GCK(cudaStreamSynchronize(b->stream));
  return UMPC_OK;
}

extern "C" void* gqpBatchDevPtr(GqpBatch* b, const char* name) {
  if (!b || !name) return nullptr;
  const Dev& d = b->d;
  struct { const char* n; void* p; } tab[] = {{"P", d.P}, {"Av", d.Av}, {"q", d.q}, {"l", d.l}, {"u", d.u}, {"x", d.x}, {"y", d.y},
                                              {"info", d.info}, {"stream", (void*)b->stream}};
  for (auto& t : tab)
    if (!strcmp(t.n, name)) return t.p;
  return nullptr;
}

extern "C" int gqpLtiRollout(GqpBatch* b, int nsteps, int nx, int nu, int u_off, const double* Ad, const double* Bd, double* x0,
                             double* u_log, double* iters_sum, int rescale_first) {
  if (!b || nsteps <= 0 || nx <= 0 || nx > 32 || nu <= 0 || !Ad || !Bd || !x0 || u_off < 0 || u_off + nu > b->d.s.n || nx > b->d.s.m) {
    umpc_internal_set_error("gqpLtiRollout: bad argument (nx <= 32)");
    return UMPC_ERR_ARG;
  }
  GCK(cudaSetDevice(b->device));
  const size_t B = b->nqp;
  if (!b->d.Lc) {   // closed loop = many solves of the same P, A: keep every QP's factor between them
    const size_t per = (size_t)b->d.s.n * (b->d.s.n + 1) / 2 + b->d.s.n;
    int rc0;
    if ((rc0 = galloc(b, &b->d.lvalid, B)) || (rc0 = galloc(b, &b->d.Lc, B * per))) { b->d.Lc = nullptr; return rc0; }
  }
  struct Tmp {   // freed on every return path
    double* p = nullptr;
    ~Tmp() { if (p) cudaFree(p); }
  } tAd, tBd, tx0, tul, tit;
  GCK(cudaMalloc(&tAd.p, (size_t)nx * nx * 8)); GCK(cudaMalloc(&tBd.p, (size_t)nx * nu * 8)); GCK(cudaMalloc(&tx0.p, B * nx * 8));
  GCK(cudaMalloc(&tit.p, 8));
  if (u_log) GCK(cudaMalloc(&tul.p, (size_t)nsteps * B * nu * 8));
  double *dAd = tAd.p, *dBd = tBd.p, *dx0 = tx0.p, *dul = tul.p, *dit = tit.p;
  GCK(cudaMemcpyAsync(dAd, Ad, (size_t)nx * nx * 8, cudaMemcpyHostToDevice, b->stream));
  GCK(cudaMemcpyAsync(dBd, Bd, (size_t)nx * nu * 8, cudaMemcpyHostToDevice, b->stream));
  GCK(cudaMemcpyAsync(dx0, x0, B * nx * 8, cudaMemcpyHostToDevice, b->stream));
  GCK(cudaMemsetAsync(dit, 0, 8, b->stream));
  GCK(cudaEventRecord(b->ev0, b->stream));
  int rc = UMPC_OK;
  for (int k = 0; k < nsteps && !rc; ++k) {
    rc = gqp_launch(b, k == 0 && rescale_first);   // later steps: only the bounds changed, prob.update(l=l, u=u) (mpc_osqp.py:137-140)
    if (rc) break;
    gqp::gqp_lti_advance_kernel<<<(unsigned)((B + 127) / 128), 128, 0, b->stream>>>(b->d, nx, nu, u_off, dAd, dBd, dx0,
                                                                                  dul ? dul + (size_t)k * B * nu : nullptr, dit);
    if (cudaGetLastError() != cudaSuccess) rc = UMPC_ERR_CUDA;
  }
  cudaEventRecord(b->ev1, b->stream);
  b->timed = true;
  if (!rc) {
    cudaMemcpyAsync(x0, dx0, B * nx * 8, cudaMemcpyDeviceToHost, b->stream);
    if (u_log) cudaMemcpyAsync(u_log, dul, (size_t)nsteps * B * nu * 8, cudaMemcpyDeviceToHost, b->stream);
    if (iters_sum) cudaMemcpyAsync(iters_sum, dit, 8, cudaMemcpyDeviceToHost, b->stream);
  }
  if (cudaStreamSynchronize(b->stream) != cudaSuccess && !rc) rc = UMPC_ERR_CUDA;
  if (cudaGetLastError() != cudaSuccess && !rc) rc = UMPC_ERR_CUDA;
  return rc;
}
