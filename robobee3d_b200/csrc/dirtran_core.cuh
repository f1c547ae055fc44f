// robobee3d_b200/csrc/dirtran_core.cuh -- per-environment algorithm of the batched planar dir-tran MPC (see dirtran.cu
// for the design).  Plain functions on one environment's slab column so that the same source also compiles for the
// host: tests/hostcheck builds it with g++ and checks it against the oracle on the CPU box (test infrastructure only; the
// product launches it from dirtran.cu's kernels and has no CPU path).
#pragma once
#include <math.h>
#include <stddef.h>
#include "../../include/dirtran_b200.h"

#ifdef __CUDACC__
#define DT_FN __device__ __forceinline__
#define DT_FN_NOINLINE __device__ __noinline__
#define DT_HD __host__ __device__
#define DT_SINCOS(a, s, c) sincos(a, s, c)
#define DT_PREFETCH_L2(p) asm volatile("prefetch.global.L2 [%0];" ::"l"(p))
#else
#define DT_FN inline
#define DT_FN_NOINLINE inline
#define DT_HD
#define DT_SINCOS(a, s, c) (*(s) = sin(a), *(c) = cos(a))
#define DT_PREFETCH_L2(p) ((void)(p))
#endif

namespace dtcore {


constexpr int NX = DT_NX, NU = DT_NU, NW = NX + NU;
// header rows of an environment's slab column
constexpr int R_Y = 0, R_CTRL = 6, R_RHO = 8, R_INFO = 9 /* iter,status,pri,dua,rho_updates */, R_LIN = 14 /* 12 */,
              R_XR = 26, R_CNT = 32 /* iters, solves, unsolved */, R_STEPS = 35 /* steps left in the running call */,
              R_Y0 = 36 /* 6: initial state of the last solve */, R_HDR = 44;
// rows of knot k (stride 128)
// (z of the equality rows is not stored: after an iteration it IS the bound, -fd for the dynamics rows of every knot and
// -y0 for the initial-state rows; until the first iteration of a solve it is the previous solve's bound -- header rows)
constexpr int K_X = 0, K_ZI = 8, K_YI = 16, K_YE = 24, K_SCR = 30, K_LD = 38, K_STRIDE = 74;
constexpr int LANES = 32;

constexpr double RHO_MIN = 1e-6, RHO_MAX = 1e6, RHO_EQ_OVER_INEQ = 1e3;  // osqp constants.h
constexpr int ST_SOLVED = 1, ST_SOLVED_INACCURATE = 2, ST_MAX_ITER = -2, ST_UNSOLVED = -10;

struct DevPar {
  int N, max_iter, check_termination, adaptive_rho_interval, nsub;
  double dt, sim_dt, sigma, alpha, eps_abs, eps_rel, rho0, adaptive_rho_tol, g, ibmb_inv;
  double wx[NX], wu[NU], lo[NW], hi[NW];
  unsigned loose;  // bit j: inequality row j of a knot has no finite bound (rho_vec = RHO_MIN)
};

DT_HD constexpr int tri(int i, int j) { return i * (i + 1) / 2 + j; }

struct Env {
  double* base;
  DT_FN double& h(int r) const { return base[r * LANES]; }
  DT_FN double* knot(int k) const { return base + (size_t)(R_HDR + k * K_STRIDE) * LANES; }
};

// the linearisation, FlappingModels.py:51-102: Ad = I + dt dfdy, Bd = dt dfdu
struct Lin {
  double dt, a32, a42, b30, b40, b50, b51;
  // out = Ad x + Bd u
  DT_FN void mul(const double (&w)[NW], double (&o)[NX]) const {
    o[0] = w[0] + dt * w[3]; o[1] = w[1] + dt * w[4]; o[2] = w[2] + dt * w[5];
    o[3] = w[3] + a32 * w[2] + b30 * w[6];
    o[4] = w[4] + a42 * w[2] + b40 * w[6];
    o[5] = w[5] + b50 * w[6] + b51 * w[7];
  }
  // o += [Ad Bd]' t
  DT_FN void mulT_add(const double (&t)[NX], double (&o)[NW]) const {
    o[0] += t[0]; o[1] += t[1]; o[2] += t[2] + a32 * t[3] + a42 * t[4];
    o[3] += dt * t[0] + t[3]; o[4] += dt * t[1] + t[4]; o[5] += dt * t[2] + t[5];
    o[6] += b30 * t[3] + b40 * t[4] + b50 * t[5];
    o[7] += b51 * t[5];
  }
  // dense [Ad Bd], row-major 6x8 (factorisation only)
  DT_FN void dense(double (&M)[NX][NW]) const {
#pragma unroll
    for (int i = 0; i < NX; ++i)
#pragma unroll
      for (int j = 0; j < NW; ++j) M[i][j] = (i == j) ? 1.0 : 0.0;
    M[0][3] = dt; M[1][4] = dt; M[2][5] = dt;
    M[3][2] = a32; M[4][2] = a42;
    M[3][6] = b30; M[4][6] = b40; M[5][6] = b50; M[5][7] = b51;
  }
};

DT_FN void dydt(const DevPar& P, const double (&y)[NX], double u0, double u1, double (&f)[NX]) {
  double s, c;
  DT_SINCOS(y[2], &s, &c);
  const double thrust = P.g + u0;
  f[0] = y[3]; f[1] = y[4]; f[2] = y[5];
  f[3] = -thrust * s;
  f[4] = -P.g + thrust * c;
  f[5] = thrust * u1 * P.ibmb_inv;
}

// linearise at (y, u): fills lin and fd; stores both in the header rows
DT_FN void linearise(const DevPar& P, const Env& e, const double (&y)[NX], double u0, double u1, Lin& L,
                                          double (&fd)[NX]) {
  double s, c;
  DT_SINCOS(y[2], &s, &c);
  const double thrust = P.g + u0;
  const double d32 = -thrust * c, d42 = -thrust * s;               // dfdy[3][2], dfdy[4][2]
  const double e30 = -s, e40 = c, e50 = u1 * P.ibmb_inv, e51 = thrust * P.ibmb_inv;  // dfdu
  double f[NX];
  f[0] = y[3]; f[1] = y[4]; f[2] = y[5];
  f[3] = -thrust * s; f[4] = -P.g + thrust * c; f[5] = thrust * u1 * P.ibmb_inv;
  // fd = dt (f - dfdy y - dfdu u), same order of subtraction as the row loops of the reference's matrix products
  fd[0] = P.dt * (f[0] - y[3]); fd[1] = P.dt * (f[1] - y[4]); fd[2] = P.dt * (f[2] - y[5]);
  fd[3] = P.dt * (f[3] - d32 * y[2] - e30 * u0);
  fd[4] = P.dt * (f[4] - d42 * y[2] - e40 * u0);
  fd[5] = P.dt * (f[5] - (e50 * u0 + e51 * u1));
  L.dt = P.dt; L.a32 = P.dt * d32; L.a42 = P.dt * d42;
  L.b30 = P.dt * e30; L.b40 = P.dt * e40; L.b50 = P.dt * e50; L.b51 = P.dt * e51;
  e.h(R_LIN + 0) = L.a32; e.h(R_LIN + 1) = L.a42; e.h(R_LIN + 2) = L.b30; e.h(R_LIN + 3) = L.b40;
  e.h(R_LIN + 4) = L.b50; e.h(R_LIN + 5) = L.b51;
#pragma unroll
  for (int i = 0; i < NX; ++i) e.h(R_LIN + 6 + i) = fd[i];
}

DT_FN bool is_loose(const DevPar& P, int j) { return (P.loose >> j) & 1u; }
DT_FN double rho_of_row(const DevPar& P, int j, double rho) { return is_loose(P, j) ? RHO_MIN : rho; }

// Block Cholesky of H = P + sigma I + A' diag(rho_vec) A in knot order; writes LD_k = chol(S_k) (lower, packed, reciprocal
// diagonal) of every knot, S_0 = H_00, S_{k+1} = H_{k+1,k+1} - LO_k LO_k'.  The coupling LO_k = H_{k+1,k} L_k^-T
// (6x8, H_{k+1,k} = -rho_e [Ad Bd]) lives in registers only: the iteration applies H_{k+1,k} S_k^-1 as a sparse product
// with [Ad Bd] around two triangular solves with LD_k, which it has loaded anyway (admm_iteration).
DT_FN_NOINLINE void factor(const DevPar& P, const Env& e, const Lin& L, double rho) {
  const double rho_e = RHO_EQ_OVER_INEQ * rho;
  double M[NX][NW];
  L.dense(M);
  double Hd[NW * (NW + 1) / 2];  // the diagonal block of a knot k < N
#pragma unroll
  for (int i = 0; i < NW; ++i)
#pragma unroll
    for (int j = 0; j <= i; ++j) {
      double g = 0.0;
#pragma unroll
      for (int r = 0; r < NX; ++r) g += M[r][i] * M[r][j];
      if (i == j && i < NX) g += 1.0;  // the -I of the knot's own dynamics (or initial-state) row
      g *= rho_e;
      if (i == j) g += (i < NX ? P.wx[i] : P.wu[i - NX]) + P.sigma + rho_of_row(P, i, rho);
      Hd[tri(i, j)] = g;
    }
  double S[NW * (NW + 1) / 2];
#pragma unroll
  for (int t = 0; t < NW * (NW + 1) / 2; ++t) S[t] = Hd[t];
  for (int k = 0; k <= P.N; ++k) {
    double* kb = e.knot(k);
    if (k == P.N) {  // last knot: no input, no outgoing dynamics row
#pragma unroll
      for (int i = 0; i < NW; ++i)
#pragma unroll
        for (int j = 0; j <= i; ++j) {
          if (i >= NX) S[tri(i, j)] = (i == j) ? 1.0 : 0.0;
          else if (i == j) S[tri(i, j)] += (P.wx[i] + P.sigma + rho_of_row(P, i, rho) + rho_e) - Hd[tri(i, j)];
          else S[tri(i, j)] -= Hd[tri(i, j)];
        }
    }
    // in-place Cholesky, diagonal kept as its reciprocal
#pragma unroll
    for (int j = 0; j < NW; ++j) {
      double d = S[tri(j, j)];
#pragma unroll
      for (int c = 0; c < j; ++c) d -= S[tri(j, c)] * S[tri(j, c)];
      const double r = 1.0 / sqrt(d);
      S[tri(j, j)] = r;
#pragma unroll
      for (int i = j + 1; i < NW; ++i) {
        double v = S[tri(i, j)];
#pragma unroll
        for (int c = 0; c < j; ++c) v -= S[tri(i, c)] * S[tri(j, c)];
        S[tri(i, j)] = v * r;
      }
    }
#pragma unroll
    for (int t = 0; t < NW * (NW + 1) / 2; ++t) kb[(K_LD + t) * LANES] = S[t];
    if (k == P.N) break;
    // LO = H_{k+1,k} L^-T with H_{k+1,k} = -rho_e [Ad Bd]; next S = Hd - LO LO' on the state rows
    double LO[NX][NW];
#pragma unroll
    for (int r = 0; r < NX; ++r) {
#pragma unroll
      for (int j = 0; j < NW; ++j) {
        double v = -rho_e * M[r][j];
#pragma unroll
        for (int c = 0; c < j; ++c) v -= S[tri(j, c)] * LO[r][c];
        LO[r][j] = v * S[tri(j, j)];
      }
    }
#pragma unroll
    for (int i = 0; i < NW; ++i)
#pragma unroll
      for (int j = 0; j <= i; ++j) {
        double v = Hd[tri(i, j)];
        if (i < NX) {
#pragma unroll
          for (int c = 0; c < NW; ++c) v -= LO[i][c] * LO[j][c];
        }
        S[tri(i, j)] = v;
      }
  }
}

// The sweeps are latency-bound once their bytes are down (two warps per scheduler at 255 registers, each knot = loads then
// a dependent chain of ~300 DFMA): while knot k is being solved, the rows knot k +- 1 will read are requested into L2.
// One prefetch instruction of a warp covers one 256-byte row.  Compile-time switch (PF): horizons whose resident working
// set mostly sits in L2 (N = 3, 5) lose 7-15 % to the extra requests, N >= 10 gains 13 % (profiles/r01j_dt_prefetch.txt).
DT_FN void prefetch_knot(const DevPar& P, const double* kb, bool scratch) {
#pragma unroll
  for (int r = 0; r < K_STRIDE; ++r) {
    if (r >= K_SCR && r < K_LD && !scratch) continue;
    if (r >= K_ZI && r < K_YE && is_loose(P, (r - K_ZI) % NW)) continue;
    DT_PREFETCH_L2(kb + r * LANES);
  }
}

// forward then backward substitution with one knot's packed Cholesky factor: v <- S^-1 v
DT_FN void chol_solve(const double (&Lk)[NW * (NW + 1) / 2], double (&v)[NW]) {
#pragma unroll
  for (int j = 0; j < NW; ++j) {
    double a = v[j];
#pragma unroll
    for (int c = 0; c < j; ++c) a -= Lk[tri(j, c)] * v[c];
    v[j] = a * Lk[tri(j, j)];
  }
#pragma unroll
  for (int j = NW - 1; j >= 0; --j) {
    double a = v[j];
#pragma unroll
    for (int c = j + 1; c < NW; ++c) a -= Lk[tri(c, j)] * v[c];
    v[j] = a * Lk[tri(j, j)];
  }
}

// One ADMM iteration (osqp auxil.c update_xz_tilde / update_x / update_z / update_y in the reduced form).
// Block elimination of the tridiagonal H with t_k = S_k^-1 b~_k as the only stored intermediate:
//   forward   b~_k = b_k + rho_e [Ad Bd] t_{k-1} (state rows),  t_k = S_k^-1 b~_k
//   backward  x~_k = t_k + rho_e S_k^-1 [Ad Bd]' x~_{k+1}
// Per knot and iteration: 36 factor words read per sweep, the 36 iterate words read twice and written once, t written
// and read once (196 doubles).
// zd / z0: z of the dynamics / initial-state rows as OSQP holds it on entry (see K_* above).
template <bool PF>
DT_FN void admm_iteration(const DevPar& P, const Env& e, const Lin& L, const double (&fd)[NX],
                                               const double (&y0)[NX], const double (&q)[NX], double rho,
                                               const double (&zd)[NX], const double (&z0)[NX]) {
  const double rho_e = RHO_EQ_OVER_INEQ * rho, alpha = P.alpha, oma = 1.0 - P.alpha;
  double rin[NW], rin_inv[NW];
#pragma unroll
  for (int j = 0; j < NW; ++j) { rin[j] = rho_of_row(P, j, rho); rin_inv[j] = 1.0 / rin[j]; }
  const int N = P.N;
  // ---- forward sweep: rhs = sigma x - q + A'(rho z - y), block forward elimination
  double te[NX], t[NW];
  {
    const double* kb = e.knot(0);
#pragma unroll
    for (int i = 0; i < NX; ++i) te[i] = rho_e * z0[i] - kb[(K_YE + i) * LANES];
  }
  for (int k = 0; k <= N; ++k) {
    double* kb = e.knot(k);
    const bool last = (k == N);
    if (PF && !last) prefetch_knot(P, e.knot(k + 1), false);
    double Lk[NW * (NW + 1) / 2];
#pragma unroll
    for (int i = 0; i < NW * (NW + 1) / 2; ++i) Lk[i] = kb[(K_LD + i) * LANES];
    double b[NW];
#pragma unroll
    for (int j = 0; j < NW; ++j) {
      const double xj = kb[(K_X + j) * LANES];
      // a bound row without finite bounds ("loose"): z = x and y = 0 exactly, in OSQP's arithmetic too, so neither is stored
      const double zj = is_loose(P, j) ? xj : kb[(K_ZI + j) * LANES], yj = is_loose(P, j) ? 0.0 : kb[(K_YI + j) * LANES];
      b[j] = P.sigma * xj + (rin[j] * zj - yj);
    }
#pragma unroll
    for (int i = 0; i < NX; ++i) b[i] -= q[i] + te[i];
    if (!last) {
      const double* kn = e.knot(k + 1);
#pragma unroll
      for (int i = 0; i < NX; ++i) te[i] = rho_e * zd[i] - kn[(K_YE + i) * LANES];
      L.mulT_add(te, b);
    } else {
      b[NX] = 0.0; b[NX + 1] = 0.0;
    }
    if (k > 0) {
      double mt[NX];
      L.mul(t, mt);
#pragma unroll
      for (int r = 0; r < NX; ++r) b[r] += rho_e * mt[r];
    }
    chol_solve(Lk, b);
#pragma unroll
    for (int j = 0; j < NW; ++j) { t[j] = b[j]; kb[(K_SCR + j) * LANES] = b[j]; }
  }
  // ---- backward sweep: block back substitution, z~ = A x~, x/z/y updates
  double xn[NX];  // x~ of knot k+1 (state part)
#pragma unroll
  for (int i = 0; i < NX; ++i) xn[i] = 0.0;
  for (int k = N; k >= 0; --k) {
    double* kb = e.knot(k);
    const bool last = (k == N);
    if (PF && k > 0) prefetch_knot(P, e.knot(k - 1), true);
    double xt[NW];
#pragma unroll
    for (int j = 0; j < NW; ++j) xt[j] = kb[(K_SCR + j) * LANES];
    if (!last) {
      double Lk[NW * (NW + 1) / 2];
#pragma unroll
      for (int i = 0; i < NW * (NW + 1) / 2; ++i) Lk[i] = kb[(K_LD + i) * LANES];
      double w[NW];
#pragma unroll
      for (int j = 0; j < NW; ++j) w[j] = 0.0;
      L.mulT_add(xn, w);
#pragma unroll
      for (int j = 0; j < NW; ++j) w[j] *= rho_e;
      chol_solve(Lk, w);
#pragma unroll
      for (int j = 0; j < NW; ++j) xt[j] += w[j];
    }
    if (!last) {
      // dynamics rows of knot k+1: z~ = Ad x~_k + Bd u~_k - x~_{k+1}; l = u = -fd
      double* kn = e.knot(k + 1);
      double zt[NX];
      L.mul(xt, zt);
#pragma unroll
      for (int i = 0; i < NX; ++i) {
        const double zp = zd[i], yp = kn[(K_YE + i) * LANES];
        const double v = alpha * (zt[i] - xn[i]) + oma * zp;
        const double zn = -fd[i];  // = min(max(v + y/rho, l), u) with l = u = -fd
        kn[(K_YE + i) * LANES] = yp + rho_e * (v - zn);
      }
    }
    // bound rows of this knot (z~ = x~) and the x update
#pragma unroll
    for (int j = 0; j < NW; ++j) {
      const bool lj = is_loose(P, j);
      const double xp = kb[(K_X + j) * LANES];
      const double zp = lj ? xp : kb[(K_ZI + j) * LANES], yp = lj ? 0.0 : kb[(K_YI + j) * LANES];
      const double v = alpha * xt[j] + oma * zp;
      double zn = fmin(fmax(v + rin_inv[j] * yp, P.lo[j]), P.hi[j]);
      double xv = alpha * xt[j] + oma * xp;
      double yn = yp + rin[j] * (v - zn);
      if (last && j >= NX) { zn = 0.0; xv = 0.0; yn = 0.0; }
      kb[(K_X + j) * LANES] = xv;
      if (!lj) { kb[(K_ZI + j) * LANES] = zn; kb[(K_YI + j) * LANES] = yn; }
    }
#pragma unroll
    for (int i = 0; i < NX; ++i) xn[i] = xt[i];
  }
  {  // initial-state rows: z~ = -x~_0, l = u = -y0
    double* kb = e.knot(0);
#pragma unroll
    for (int i = 0; i < NX; ++i) {
      const double zp = z0[i], yp = kb[(K_YE + i) * LANES];
      const double v = alpha * (-xn[i]) + oma * zp;
      const double zn = -y0[i];
      kb[(K_YE + i) * LANES] = yp + rho_e * (v - zn);
    }
  }
}

struct Resid { double pri, dua, nprim, ndual; };

// osqp auxil.c compute_pri_res / compute_dua_res and the norms of compute_pri_tol / compute_dua_tol (scaling off)
DT_FN Resid residuals(const DevPar& P, const Env& e, const Lin& L, const double (&q)[NX], const double (&zd)[NX],
                      const double (&z0)[NX]) {
  const int N = P.N;
  double pri = 0.0, nz = 0.0, nAx = 0.0, dua = 0.0, nPx = 0.0, nAty = 0.0, nq = 0.0;
#pragma unroll
  for (int i = 0; i < NX; ++i) nq = fmax(nq, fabs(q[i]));
  double xn[NX], yen[NX], zen[NX];
#pragma unroll
  for (int i = 0; i < NX; ++i) { xn[i] = 0.0; yen[i] = 0.0; zen[i] = 0.0; }
  for (int k = N; k >= 0; --k) {
    const double* kb = e.knot(k);
    const bool last = (k == N);
    double x[NW], aty[NW], ye[NX], ze[NX];
#pragma unroll
    for (int j = 0; j < NW; ++j) {
      x[j] = kb[(K_X + j) * LANES];
      const double zi = is_loose(P, j) ? x[j] : kb[(K_ZI + j) * LANES];
      aty[j] = is_loose(P, j) ? 0.0 : kb[(K_YI + j) * LANES];
      if (!(last && j >= NX)) { pri = fmax(pri, fabs(x[j] - zi)); nz = fmax(nz, fabs(zi)); nAx = fmax(nAx, fabs(x[j])); }
    }
#pragma unroll
    for (int i = 0; i < NX; ++i) { ye[i] = kb[(K_YE + i) * LANES]; ze[i] = (k == 0) ? z0[i] : zd[i]; aty[i] -= ye[i]; }
    if (!last) {
      double ax[NX];
      L.mul(x, ax);
#pragma unroll
      for (int i = 0; i < NX; ++i) {
        const double a = ax[i] - xn[i];
        pri = fmax(pri, fabs(a - zen[i])); nz = fmax(nz, fabs(zen[i])); nAx = fmax(nAx, fabs(a));
      }
      L.mulT_add(yen, aty);
    }
#pragma unroll
    for (int j = 0; j < NW; ++j) {
      if (last && j >= NX) continue;
      const double px = (j < NX ? P.wx[j] : P.wu[j - NX]) * x[j];
      const double qq = j < NX ? q[j] : 0.0;
      dua = fmax(dua, fabs(px + qq + aty[j])); nPx = fmax(nPx, fabs(px)); nAty = fmax(nAty, fabs(aty[j]));
    }
#pragma unroll
    for (int i = 0; i < NX; ++i) { xn[i] = x[i]; yen[i] = ye[i]; zen[i] = ze[i]; }
  }
#pragma unroll
  for (int i = 0; i < NX; ++i) {  // initial-state rows: Ax = -x_0
    pri = fmax(pri, fabs(-xn[i] - zen[i])); nz = fmax(nz, fabs(zen[i])); nAx = fmax(nAx, fabs(xn[i]));
  }
  Resid r;
  r.pri = pri; r.dua = dua; r.nprim = fmax(nz, nAx); r.ndual = fmax(nq, fmax(nAty, nPx));
  return r;
}

DT_FN bool converged(const DevPar& P, const Resid& r, double f) {
  return r.pri < f * (P.eps_abs + P.eps_rel * r.nprim) && r.dua < f * (P.eps_abs + P.eps_rel * r.ndual);
}

// nsteps x (one LTVMPC.update [+ the plant's Euler substeps when PLANT]) for this thread's environment, as ONE loop over
// ADMM iterations: a lane that finishes a solve starts its next step in the same trip instead of idling until the
// slowest lane of its warp is done.  Solves end and rho adapts only at trips that are multiples of check_termination /
// adaptive_rho_interval, so the lanes of a warp keep taking the residual branch together however far apart their step
// counters drift.  Returns the last u_0 in (u0, u1), the plant state in y.
template <bool PLANT, bool PF = false>
DT_FN void mpc_run(const DevPar& P, const Env& e, double (&y)[NX], double& u0, double& u1, int nsteps_uniform) {
  if (nsteps_uniform <= 0) return;
  // The per-lane step budget goes through the environment's header (volatile read-back).  Kept in a plain local
  // initialised from the kernel parameter, nvcc 12.9 / ptxas put it in a UNIFORM register (UIADD3 UR5, UR5, -1;
  // BRA.U in the SASS of dt_rollout_kernel) although it is decremented under a lane-dependent condition: the first
  // lanes to finish a step then decremented it for the whole warp (measured: 37 of 96 solves done).  A loaded value
  // is per-lane for the compiler; iter and rho_updates restart from a zero derived from it for the same reason.
  e.h(R_STEPS) = (double)nsteps_uniform;
  int nsteps = (int)*(const volatile double*)&e.h(R_STEPS);
  const int lane_zero = nsteps - nsteps_uniform;
  Lin L;
  double fd[NX], q[NX];
#pragma unroll
  for (int i = 0; i < NX; ++i) q[i] = -P.wx[i] * e.h(R_XR + i);
  double rho = e.h(R_RHO);
  // z of the equality rows as the warm start holds it: the bounds of the previous solve (0 for a fresh controller)
  double zd[NX], z0[NX], y0[NX];
#pragma unroll
  for (int i = 0; i < NX; ++i) { zd[i] = 0.0 - e.h(R_LIN + 6 + i); z0[i] = 0.0 - e.h(R_Y0 + i); y0[i] = y[i]; }
  linearise(P, e, y, u0, u1, L, fd);
  factor(P, e, L, rho);
  int iter = lane_zero, rho_updates = lane_zero;
  while (true) {
    ++iter;
    admm_iteration<PF>(P, e, L, fd, y0, q, rho, zd, z0);
#pragma unroll
    for (int i = 0; i < NX; ++i) { zd[i] = -fd[i]; z0[i] = -y0[i]; }
    const bool at_max = iter >= P.max_iter;
    const bool checked = P.check_termination && (iter % P.check_termination == 0);
    const bool adapt = P.adaptive_rho_interval && (iter % P.adaptive_rho_interval == 0);
    if (!(checked || adapt || at_max)) continue;
    const Resid r = residuals(P, e, L, q, zd, z0);
    // order of osqp.c's loop: termination test, then the rho adaptation, then (out of iterations) the final tests
    int status = (checked && converged(P, r, 1.0)) ? ST_SOLVED : ST_UNSOLVED;
    bool refactor = false;
    if (status == ST_UNSOLVED && adapt) {  // osqp auxil.c compute_rho_estimate / adapt_rho
      const double pn = r.pri / (r.nprim + 1e-10), dn = r.dua / (r.ndual + 1e-10);
      double est = rho * sqrt(pn / (dn + 1e-10));
      est = fmin(fmax(est, RHO_MIN), RHO_MAX);
      if (est > rho * P.adaptive_rho_tol || est < rho / P.adaptive_rho_tol) {
        rho = est;
        ++rho_updates;
        refactor = true;
      }
    }
    if (status == ST_UNSOLVED && at_max)
      status = converged(P, r, 1.0) ? ST_SOLVED : (converged(P, r, 10.0) ? ST_SOLVED_INACCURATE : ST_MAX_ITER);
    if (status != ST_UNSOLVED) {  // this solve is over: record it, move the plant, open the next step's solve
      e.h(R_INFO + 0) = (double)iter; e.h(R_INFO + 1) = (double)status; e.h(R_INFO + 2) = r.pri; e.h(R_INFO + 3) = r.dua;
      e.h(R_INFO + 4) = (double)rho_updates;
      e.h(R_CNT + 0) += (double)iter; e.h(R_CNT + 1) += 1.0; e.h(R_CNT + 2) += (status == ST_SOLVED) ? 0.0 : 1.0;
      const double* k0 = e.knot(0);
      u0 = k0[(K_X + NX) * LANES]; u1 = k0[(K_X + NX + 1) * LANES];  // mpc.py:89
      if (PLANT) {
        for (int t = 0; t < P.nsub; ++t) {
          double f[NX];
          dydt(P, y, u0, u1, f);
#pragma unroll
          for (int i = 0; i < NX; ++i) y[i] += f[i] * P.sim_dt;
        }
      }
      if (--nsteps == 0) break;
#pragma unroll
      for (int i = 0; i < NX; ++i) y0[i] = y[i];
      linearise(P, e, y, u0, u1, L, fd);
      iter = lane_zero; rho_updates = lane_zero;
      refactor = true;
    }
    if (refactor) factor(P, e, L, rho);
  }
  e.h(R_RHO) = rho;
  e.h(R_CTRL) = u0; e.h(R_CTRL + 1) = u1;
#pragma unroll
  for (int i = 0; i < NX; ++i) e.h(R_Y0 + i) = y0[i];
}

// one LTVMPC.update for this thread's environment; returns u_0 in (u0, u1) and updates ctrl
template <bool PF = false>
DT_FN void mpc_update(const DevPar& P, const Env& e, const double (&y)[NX], double& u0, double& u1) {
  double yy[NX];
#pragma unroll
  for (int i = 0; i < NX; ++i) yy[i] = y[i];
  mpc_run<false, PF>(P, e, yy, u0, u1, 1);
}


}  // namespace dtcore
