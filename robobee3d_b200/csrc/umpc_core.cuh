// robobee3d_b200/csrc/umpc_core.cuh
//
// UprightMPC2 controller update for ONE environment, written so that a CUDA
// thread owns one QP (all indices compile-time, vectors in registers, the
// per-QP factor in a caller-provided store -- shared memory in the kernels).
//
// What it computes (reference file:line, all under template/):
//   template projection        uprightmpc2/uprightmpc2.c:214-241
//   l,u,Ax data                uprightmpc2/uprightmpc2.c:126-180  (template_controllers.py:65-125)
//   Px,q                       uprightmpc2/uprightmpc2.c:182-207  (template_controllers.py:127-143)
//   constraint types / rho     uprightmpc2/auxil.c:103-146
//   Ruiz equilibration x10     uprightmpc2/scaling.c:44-156
//   ADMM iterations            uprightmpc2/osqp.c:288-641, auxil.c:164-228, proj.c:4-14
//   residuals / status         uprightmpc2/auxil.c:243-361,567-633,684-790
//   solution un-scaling        uprightmpc2/scaling.c:177-192
//   uquad / accdes extraction  uprightmpc2/uprightmpc2.c:252-269
//
// Formulation.  OSQP iterates on the Ruiz-scaled problem (x_s = D^-1 x,
// z_s = E z, y_s = c E^-1 y ...).  We carry the SAME iterates in un-scaled
// ("hatted") coordinates  xh = D x_s, zh = E^-1 z_s, yh = E y_s, in which the
// scaled ADMM step becomes an ADMM step on the UNSCALED matrix A0 with a
// per-row penalty W = rho .* E^2, a per-variable regulariser sigma / D^2 and the
// cost multiplied by c:
//     K xt = sigma D^-2 xh - c q0 + A0'(W zh - yh),   K = c P0 + sigma D^-2 + A0' W A0
//     zt = A0 xt ; xh+ = a xt + (1-a) xh ; zh+ = clip(a zt + (1-a) zh + yh/W, l0, u0)
//     yh+ = yh + W (a zt + (1-a) zh - zh+)
// which is algebraically identical to auxil.c:164-228 + qdldl_interface.c:343-369
// (eliminate nu from the quasi-definite KKT system) and needs neither the
// scaled copies of P and A nor the 84x84 factor.  The warm start is kept in
// HBM in OSQP's own scaled coordinates and, like the reference (no
// osqp_warm_start call in uprightmpc2.c), is NOT rescaled when D, E change.
#pragma once
#include <math.h>

#if defined(__CUDACC__)
#define UMPC_HD __host__ __device__ __forceinline__
#else
#define UMPC_HD inline
#endif

// Scalar type of the whole update.  float: the C path (uprightmpc2.c is float32 throughout).  double: the Python
// path (osqp on float64 numpy) -- a translation unit that wants it defines UMPC_REAL double and includes this header
// inside a namespace of its own (umpc_conv.cu), so both builds can live in one library.
#ifndef UMPC_REAL
#define UMPC_REAL float
#define UMPC_REAL_IS_FLOAT 1
#endif
typedef UMPC_REAL ureal;

#ifndef UMPC_NX  // also in include/umpc_b200.h
#define UMPC_NX 45
#define UMPC_NC 39
#endif
#define UMPC_NEQ 36

// OSQP status values (uprightmpc2/constants.h:17-31)
#define UMPC_ST_SOLVED 1
#define UMPC_ST_SOLVED_INACCURATE 2
#define UMPC_ST_MAX_ITER (-2)
#define UMPC_ST_NON_CVX (-7)
#define UMPC_ST_UNSOLVED (-10)
#define UMPC_ST_NOT_EVALUATED 0   // ours: update ran with want_info = false (no residuals, no status)

// Controller parameters (what umpcInit derives, uprightmpc2.c:19-63,116-117) plus the
// OSQP settings baked into the generated workspace (uprightmpc2/workspace.c:561).
struct UmpcParams {
  ureal dt, g, Tmax;
  ureal Qyr[6], Qyf[6], Qdyr[6], Qdyf[6], R[3];
  ureal Ibi[3];
  int max_iter;
  ureal rho, sigma, alpha, eps_abs, eps_rel;
  int scaling_iters;
  // converged mode (the Python path's OSQP settings, template_controllers.py:190-191): 0 = the C path (uprightmpc2.c:116-117)
  int check_termination;      // test the residuals every this many iterations and stop when below eps
  int adaptive_rho_interval;  // re-estimate rho every this many iterations (auxil.c:13-75), refactor when it moves x5
  ureal adaptive_rho_tol;
  int debug_phase;  // measurement aid (UMPC_DEBUG_PHASE): 1 = stop after set-up, 2 = stop after the ADMM loop; 0 = normal
};

// unscaled A0 values that change per update (Ax_data, uprightmpc2.c:160-179)
struct UmpcAVals {
  ureal dt, dtT0, ds[3], db[6];
  ureal abs_dt, abs_dtT0, abs_ds0, abs_ds1, abs_ds2, abs_db0, abs_db1, abs_db2, abs_db3, abs_db4, abs_db5;
};

// reciprocal / reciprocal square root: MUFU approximations (1-2 ulp) on the device instead of the
// IEEE division + sqrt slow paths; the host build (tests/hostcheck) and the double build keep the exact forms.
#if defined(__CUDA_ARCH__) && defined(UMPC_REAL_IS_FLOAT)
#define UMPC_RCP(x) __fdividef(1.0f, (x))
#define UMPC_RSQRT(x) rsqrtf(x)
#elif defined(UMPC_REAL_IS_FLOAT)
#define UMPC_RCP(x) (1.0f / (x))
#define UMPC_RSQRT(x) (1.0f / sqrtf(x))
#elif defined(__CUDA_ARCH__)
// double on the device: rsqrt() is MUFU.RSQ64H + Newton steps (~1 ulp), a sixth of the instructions of sqrt + divide --
// the 84 of them per Ruiz pass were two thirds of the pass
// 1 / x for NORMAL x (scalings, rho-weighted diagonals: 1e-10 .. 1e10): MUFU.RCP64H (20 bits) + two Newton steps = 1 ulp
// in 5 instructions; the IEEE division it replaces is ~25 with a slow-path call, 190 of them in the once-per-check code
__device__ __forceinline__ double umpc_drcp(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  r = fma(r, fma(-x, r, 1.0), r);
  return fma(r, fma(-x, r, 1.0), r);
}
#define UMPC_RCP(x) umpc_drcp(x)
#define UMPC_RSQRT(x) rsqrt(x)
#else
#define UMPC_RCP(x) (1.0 / (x))
#define UMPC_RSQRT(x) (1.0 / sqrt(x))
#endif
#ifdef UMPC_REAL_IS_FLOAT
UMPC_HD ureal umpc_fmax(ureal a, ureal b) { return fmaxf(a, b); }
UMPC_HD ureal umpc_fmin(ureal a, ureal b) { return fminf(a, b); }
UMPC_HD ureal umpc_fabs(ureal a) { return fabsf(a); }
UMPC_HD ureal umpc_sqrt(ureal a) { return sqrtf(a); }
UMPC_HD ureal umpc_nan() { return nanf(""); }
#else
UMPC_HD ureal umpc_fmax(ureal a, ureal b) { return fmax(a, b); }
UMPC_HD ureal umpc_fmin(ureal a, ureal b) { return fmin(a, b); }
UMPC_HD ureal umpc_fabs(ureal a) { return fabs(a); }
UMPC_HD ureal umpc_sqrt(ureal a) { return sqrt(a); }
UMPC_HD ureal umpc_nan() { return nan(""); }
#endif

// ---- arithmetic of the Ruiz equilibration ("sreal") ---------------------------------------------------------------
// Any positive D, E, c is a valid equilibration (the termination test runs on un-scaled residuals), so the ten passes
// need not run in the width of the iterates: -DUMPC_SCALE_FLOAT computes them in float also in the double build (the
// GPU's converged mode: the passes in double were 30 % of the kernel); the results are widened once and used
// consistently from there on.  Default: the width of the iterates, like OSQP.
#if defined(UMPC_REAL_IS_FLOAT) || defined(UMPC_SCALE_FLOAT)
typedef float sreal;
#else
typedef double sreal;
#endif
UMPC_HD float umpc_tmax(float a, float b) { return fmaxf(a, b); }
UMPC_HD double umpc_tmax(double a, double b) { return fmax(a, b); }
UMPC_HD float umpc_tmin(float a, float b) { return fminf(a, b); }
UMPC_HD double umpc_tmin(double a, double b) { return fmin(a, b); }
UMPC_HD float umpc_tabs(float a) { return fabsf(a); }
UMPC_HD double umpc_tabs(double a) { return fabs(a); }
#if defined(__CUDA_ARCH__)
UMPC_HD float umpc_trsqrt(float a) { return rsqrtf(a); }
UMPC_HD double umpc_trsqrt(double a) { return rsqrt(a); }
UMPC_HD float umpc_trcp(float a) { return __fdividef(1.0f, a); }
#else
UMPC_HD float umpc_trsqrt(float a) { return 1.0f / sqrtf(a); }
UMPC_HD double umpc_trsqrt(double a) { return 1.0 / sqrt(a); }
UMPC_HD float umpc_trcp(float a) { return 1.0f / a; }
#endif
UMPC_HD double umpc_trcp(double a) { return 1.0 / a; }
template <class T> UMPC_HD T umpc_tlimit(T v) {   // MIN_SCALING / MAX_SCALING (constants.h:97-98), see umpc_limit_scaling
  v = v < (T)1e-4f ? (T)1.0f : v;
  return umpc_tmin(v, (T)1e4f);
}

#include "umpc_gen.cuh"

// scratch words behind the (UMPC_NX + 1) / 2 sigma/D^2 pairs: the 9 loop-invariant input-column constants (kpu)
#define UMPC_KPU_POS (UMPC_GEN_OFF_SG + (UMPC_NX + 1) / 2)

UMPC_HD int umpc_imin(int a, int b) { return a < b ? a : b; }

UMPC_HD ureal umpc_limit_scaling(ureal v) {
  v = v < 1e-4f ? 1.0f : v;   // MIN_SCALING (constants.h:98)
  return umpc_fmin(v, 1e4f);  // MAX_SCALING; one min instead of compare + select (differs from the reference's ternary
                              // only for a NaN norm, where every later quantity is NaN through the factor anyway)
}

// sigma / D_j^2 pairs share one scratch word.  float build: two bf16 halves (8 significant bits, full float range);
// double build: two floats.  The term is ~1e-6 of the diagonal of K, and it multiplies xh and (xh - xt) only.
#if defined(UMPC_SG_FULL) && !defined(__CUDA_ARCH__)
// A/B aid of the HOST build only (tests/hostcheck, -DUMPC_SG_FULL): the pairs keep their full precision in a side table and
// the scratch word carries the table index -- shows what the bf16 packing of the device build costs (tests/test_host_side.py)
static thread_local ureal umpc_sg_tab[2048][2];
static thread_local int umpc_sg_n = 0;
UMPC_HD ureal umpc_pack2(ureal a, ureal b) {
  const int i = umpc_sg_n++ & 2047;
  umpc_sg_tab[i][0] = a; umpc_sg_tab[i][1] = b;
  return (ureal)i;
}
UMPC_HD void umpc_unpack2(ureal w, ureal& a, ureal& b) { a = umpc_sg_tab[(int)w][0]; b = umpc_sg_tab[(int)w][1]; }
#elif defined(UMPC_REAL_IS_FLOAT)
UMPC_HD ureal umpc_pack2(ureal a, ureal b) {
  union { float f; unsigned u; } ua, ub, r;
  ua.f = a; ub.f = b;
  r.u = (ua.u >> 16) | (ub.u & 0xffff0000u);
  return r.f;
}
// The high half is used as it lies (the other value's bits below it perturb it by < 1 bf16 ulp, which is the
// precision the pair has anyway): one instruction per pair in the hot loop instead of two.
UMPC_HD void umpc_unpack2(ureal w, ureal& a, ureal& b) {
  union { float f; unsigned u; } uw, ua;
  uw.f = w;
  ua.u = uw.u << 16;
  a = ua.f; b = w;
}
#else
UMPC_HD ureal umpc_pack2(ureal a, ureal b) {
  union { double d; float f[2]; } r;
  r.f[0] = (float)a; r.f[1] = (float)b;
  return r.d;
}
UMPC_HD void umpc_unpack2(ureal w, ureal& a, ureal& b) {
  union { double d; float f[2]; } r;
  r.d = w;
  a = (ureal)r.f[0]; b = (ureal)r.f[1];
}
#endif

// diagonal of P0 (updateObjective, uprightmpc2.c:182-207)
UMPC_HD ureal umpc_px(const UmpcParams& par, int j) {
  if (j < 12) return par.Qyr[j % 6];
  if (j < 18) return par.Qyf[j % 6];
  if (j < 30) return par.Qdyr[j % 6];
  if (j < 36) return par.Qdyf[j % 6];
  return par.R[(j - 36) % 3];
}

// QP data of one update in unscaled form: A0 values, bounds, references.
struct UmpcQP {
  UmpcAVals a;
  ureal lo[UMPC_NC];  // l (== u on rows 0..35)
  ureal hi_thr;       // u on rows 36..38
  ureal yd[6], dyd[6];
};

// template projection (uprightmpc2.c:219-241) + umpcUpdateConstraint (uprightmpc2.c:126-180)
UMPC_HD void umpc_build(const UmpcParams& par, UmpcQP& qp, const ureal p0[3], const ureal R0[9], const ureal dq0[6],
                        const ureal pdes[3], const ureal dpdes[3], const ureal sdes[3], ureal T0) {
  const ureal dt = par.dt;
  ureal s0[3], ds0[3], y0[6], dy0[6], y1[6];
  UmpcAVals& a = qp.a;
  const ureal w0 = dq0[3], w1 = dq0[4];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    s0[i] = R0[6 + i];
    // e3h*w = (-w1, w0, 0); ds0 = -R0 * (e3h w)
    ds0[i] = -(R0[i] * (-w1) + R0[3 + i] * w0);
    // Btau = (-R0 e3h Ibi)[:, :2]; e3h Ibi has (1,0)=Ibi0, (0,1)=-Ibi1
    a.db[i] = dt * (-(R0[3 + i] * par.Ibi[0]));
    a.db[3 + i] = dt * (R0[i] * par.Ibi[1]);
    a.ds[i] = dt * s0[i];
    y0[i] = p0[i]; y0[3 + i] = s0[i];
    dy0[i] = dq0[i]; dy0[3 + i] = ds0[i];
    qp.yd[i] = pdes[i]; qp.yd[3 + i] = sdes[i];
    qp.dyd[i] = dpdes[i]; qp.dyd[3 + i] = 0.0f;
  }
  a.dt = dt; a.dtT0 = dt * T0;
  a.abs_dt = umpc_fabs(a.dt); a.abs_dtT0 = umpc_fabs(a.dtT0);
  a.abs_ds0 = umpc_fabs(a.ds[0]); a.abs_ds1 = umpc_fabs(a.ds[1]); a.abs_ds2 = umpc_fabs(a.ds[2]);
  a.abs_db0 = umpc_fabs(a.db[0]); a.abs_db1 = umpc_fabs(a.db[1]); a.abs_db2 = umpc_fabs(a.db[2]);
  a.abs_db3 = umpc_fabs(a.db[3]); a.abs_db4 = umpc_fabs(a.db[4]); a.abs_db5 = umpc_fabs(a.db[5]);
#pragma unroll
  for (int i = 0; i < UMPC_NC; ++i) qp.lo[i] = 0.0f;
#pragma unroll
  for (int i = 0; i < 6; ++i) {
    y1[i] = y0[i] + dt * dy0[i];
    qp.lo[i] = -y1[i];
  }
#pragma unroll
  for (int i = 0; i < 6; ++i) {
    const ureal c0 = (i == 2) ? -par.g : 0.0f;
    const ureal A0y0 = i < 3 ? T0 * y0[i + 3] : 0.0f;
    const ureal A0y1 = i < 3 ? T0 * y1[i + 3] : 0.0f;
    qp.lo[18 + i] = -dy0[i] - dt * A0y0 - dt * c0;
    qp.lo[24 + i] = -dt * A0y1 - dt * c0;
    qp.lo[30 + i] = -dt * c0;
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) qp.lo[36 + k] = -T0;
  qp.hi_thr = par.Tmax - T0;
}

// the reference's public QP vectors (UprightMPC_t.l/u/q/Px_data/Ax_data, uprightmpc2.h:36-40),
// out = l[39] u[39] q[45] Px[45] Ax_data[48]  (vectors()/matrices(), py/uprightmpc2py.cpp:46-51)
UMPC_HD void umpc_qpdata(const UmpcParams& par, const UmpcQP& qp, ureal* out) {
  for (int i = 0; i < UMPC_NC; ++i) {
    out[i] = qp.lo[i];
    out[UMPC_NC + i] = i < UMPC_NEQ ? qp.lo[i] : qp.hi_thr;
  }
  ureal* q = out + 2 * UMPC_NC;
  ureal* Px = q + UMPC_NX;
  ureal* Ax = Px + UMPC_NX;
  for (int j = 0; j < UMPC_NX; ++j) {
    Px[j] = umpc_px(par, j);
    q[j] = j < 36 ? -Px[j] * (j < 18 ? qp.yd[j % 6] : qp.dyd[j % 6]) : 0.0f;
  }
  int o = 0;
  for (int i = 0; i < 3; ++i) Ax[o++] = qp.a.dtT0;
  for (int i = 0; i < 18; ++i) Ax[o++] = qp.a.dt;
  for (int k = 0; k < 3; ++k) for (int i = 0; i < 3; ++i) Ax[o++] = qp.a.ds[i];
  for (int k = 0; k < 3; ++k) for (int i = 0; i < 6; ++i) Ax[o++] = qp.a.db[i];
}

struct UmpcInfo {
  int iter, status;
  ureal pri_res, dua_res;
};

// Per-QP scratch the iterations re-read every pass (ST).  The kernels back it with shared memory (and a few
// registers), the host check backs it with arrays.
//   L(n) / Lb(n) n<167 : factor of K (forward / backward read)     Di(k) k<45 : 1/d of the LDL'
//   Sg2(q) q<23 : sigma / D_j^2 in pairs (the store may keep them in reduced precision: the term is ~1e-6 of
//                 the diagonal, and it enters the iteration only through products with (xh - xt) and xh)
//   Vf(j) j<45 : first-iteration fix-up of w (see below), global scratch
// begin_iteration(it) is called at the top of every ADMM pass (the CUDA store ties its loads to `it` so that the
// compiler cannot treat the loop-invariant scratch as 300 registers' worth of hoisted values and spill them).
// Per-env persistent state (IO; global memory, SoA): OSQP's scaled iterates xs/zs/ys (the warm start, never
// rescaled), the last D, E (E of the thrust rows decides their constraint type on the next call: bounds are
// scaled with the OLD E before the re-equilibration, osqp.c:784-832 + auxil.c:103-146), and the solution.
//
// The iteration actually executed ("A-free" form).  For the 36 equality rows (l == u) z_i == l_i exactly after
// the first projection (proj.c:4-14), so only v = W zh - yh matters there, and v enters the right-hand side
// only through g = A_eq' v.  With K_eq = A_eq' W A_eq and b = A_eq' W l the v-update v+ = v - alpha W (A_eq xt - l)
// becomes g+ = g - alpha (K_eq xt - b), and K_eq xt is available WITHOUT multiplying by A: K xt = r (the system
// just solved) and K = G + K_eq + A_3' W_3 A_3 with G diagonal and A_3 the three unit thrust rows, so
//     K_eq xt = r - G xt - A_3' W_3 A_3 xt .
// Shifting the primal iterates by the open-loop roll-out xp (A_eq xp = l, zero inputs; umpc_gen_xp) removes b
// and folds every constant of the right-hand side into the initial value of  w = g + c P0 (ydes - xp) - b :
//     r   = S xh' + w  (+ W_3 z_3 - y_3 on the three thrust columns)            S = sigma D^-2, xh' = xh - xp
//     xt' = K^-1 r                                                               (umpc_gen_solve)
//     w  <- (1 - alpha) w + alpha c P0 xt' - alpha S (xh' - xt')  (+ alpha (W_3 xt' - (W_3 z_3 - y_3)) thrust cols)
//     xh' <- xh' - alpha (xh' - xt') ;  thrust rows: the generic z/y update with clipping
// i.e. per iteration one sparse LDL' solve and ~7 flops per variable: no product with A, no per-row state.
// This is algebraically the iteration of auxil.c:164-228 (same iterates x, and z, y on the thrust rows; the
// equality-row duals are recovered at the end from g by substitution, umpc_gen_v_from_g).  The stale warm
// start z of the FIRST iteration (z_prev != l, and OSQP uses it in xz_tilde and in the relaxation) only changes
// a constant: w is replaced by a second start vector Vf right after the first right-hand side is formed,
//     w0 = kappa + A_eq'(W zh0 - yh0),   wf = kappa + A_eq'((2 - alpha) W l - yh0 - (1 - alpha) W zh0),
//     kappa = c P0 (ydes - xp) - A_eq' W l ,   w1 = wf - alpha w0 + (rest of the update).
UMPC_HD ureal umpc_lo_eq(const UmpcParams& par, int i, ureal stored) {
  // rows 6..17 are 0 and rows 30..35 are -dt*c0 (uprightmpc2.c:139-152): no load needed
  if (i >= 6 && i < 18) return 0.0f;
  if (i >= 30) return i == 32 ? par.dt * par.g : 0.0f;
  return stored;
}

// c * P0_j * ref_j : the constant part of -c q0 (q0 = -P0 .* [ydes x3, dydes x3, 0 x9], uprightmpc2.c:182-207)
UMPC_HD ureal umpc_ref(const ureal (&yd)[6], const ureal (&dyd)[6], int j) {
  return j < 18 ? yd[j % 6] : (j < 36 ? dyd[j % 6] : 0.0f);
}

// Where the two vectors that live across the whole ADMM loop (xs: the shifted iterate, w: the own-column part of
// g + kappa) are kept.  Default: registers.  A store policy that declares `typedef void has_vec` and a member template
// `Vec<N, OFF>` (constructible from the policy, indexable with []) places them elsewhere -- converged mode keeps them
// in shared memory because in double the loop state alone would be 252 registers (umpc_conv.cu).
template <int N>
struct UmpcRegVec {
  ureal v[N];
  template <class ST> UMPC_HD explicit UmpcRegVec(const ST&) {}
  UMPC_HD ureal& operator[](int i) { return v[i]; }
  UMPC_HD const ureal& operator[](int i) const { return v[i]; }
};
template <class ST, int N, int OFF, class = void>
struct UmpcVecOf { typedef UmpcRegVec<N> type; };
template <class ST, int N, int OFF>
struct UmpcVecOf<ST, N, OFF, typename ST::has_vec> { typedef typename ST::template Vec<N, OFF> type; };

template <bool CONV, bool INFO = true, class ST, class IO>
UMPC_HD void umpc_update(const UmpcParams& par, ST& st, IO& io, ureal uquad[3], ureal accdes[6], const ureal p0[3],
                         const ureal R0[9], const ureal dq0[6], const ureal pdes[3], const ureal dpdes[3],
                         const ureal sdes[3], ureal actualT0, ureal& T0, UmpcInfo& info) {
  // INFO = false (C path only, compile time: the once-per-solve code is instruction-fetch bound, every instruction counts): the caller will not read solve info / solution of this update (the reference never acts
  // on it either: check_termination = 0, uprightmpc2.c:116-117) -- the residual norms, status and the solx/soly rows are
  // skipped; the divergence guard (NaN solution + cold start, auxil.c:538-561) then tests the iterates themselves.
  const ureal dt = par.dt;
  if (actualT0 >= 0.0f) T0 = actualT0;
  io.prefetch_warm_start();  // xs/zs/ys are needed only after Ruiz + factor: start pulling them into L2 now
  // The 28 inputs are needed again only after the ADMM loop (bounds for the un-shift, R0/dq0 for accdes): park
  // them in the IO's scratch and re-read them there, so that they do not occupy registers during the loop
  // (the loop's own state is ~170 registers; what is left buys instruction-level parallelism).
#pragma unroll
  for (int i = 0; i < 3; ++i) { io.stash(i, p0[i]); io.stash(18 + i, pdes[i]); io.stash(21 + i, dpdes[i]); io.stash(24 + i, sdes[i]); }
#pragma unroll
  for (int i = 0; i < 9; ++i) io.stash(3 + i, R0[i]);
#pragma unroll
  for (int i = 0; i < 6; ++i) io.stash(12 + i, dq0[i]);
  io.stash(27, T0);
  UmpcAVals a;
  ureal yd[6], dyd[6], lo_thr, hi_thr;
  ureal qp_lo[18];  // run-time equality bounds: rows 0..5 and 18..29
  {
    UmpcQP qp;
    umpc_build(par, qp, p0, R0, dq0, pdes, dpdes, sdes, T0);
    a = qp.a;
    lo_thr = qp.lo[36]; hi_thr = qp.hi_thr;
#pragma unroll
    for (int i = 0; i < 6; ++i) { yd[i] = qp.yd[i]; dyd[i] = qp.dyd[i]; }
#pragma unroll
    for (int i = 0; i < 6; ++i) qp_lo[i] = qp.lo[i];
#pragma unroll
    for (int i = 18; i < 30; ++i) qp_lo[i - 12] = qp.lo[i];
  }
#define UMPC_LO(i) umpc_lo_eq(par, (i), qp_lo[(i) < 6 ? (i) : ((i) >= 18 && (i) < 30 ? (i) - 12 : 0)])
  // ---- constraint type of the thrust rows with the OLD scaling (auxil.c:103-146) --------
  ureal W3[3], Wi3[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const ureal eo = io.ld_E(UMPC_NEQ + k);
    const ureal ls_ = eo * lo_thr, us_ = eo * hi_thr;
    W3[k] = (us_ - ls_ < 1e-4f) ? 1e3f : 1.0f;  // RHO_TOL, RHO_EQ_OVER_RHO_INEQ (constants.h:69-71): factor of rho for now
  }
  // rho is per-controller state: osqp_update_rho writes settings->rho, which later solves inherit (osqp.c:1270-1330)
  ureal rho = par.rho;
  if constexpr (CONV) rho = io.ld_rho();
  const ureal thr_mult[3] = {W3[0], W3[1], W3[2]};
#pragma unroll
  for (int k = 0; k < 3; ++k) W3[k] *= rho;
  const ureal rho_eq = 1e3f * rho;
  const ureal alpha = par.alpha, oma = 1.0f - par.alpha;

  // ---- Ruiz equilibration + cost scaling (scale_data, scaling.c:44-156) -------------------
  ureal c = 1.0f;
  typename UmpcVecOf<ST, UMPC_NX, 0>::type xs(st);
  typename UmpcVecOf<ST, UMPC_NEQ, UMPC_NX>::type w(st);
  ureal z3[3], y3[3];
  {
    sreal Ds[UMPC_NX], Es[UMPC_NC], cs = 1.0f;
#pragma unroll
    for (int j = 0; j < UMPC_NX; ++j) Ds[j] = 1.0f;
#pragma unroll
    for (int i = 0; i < UMPC_NC; ++i) Es[i] = 1.0f;
    for (int it = 0; it < par.scaling_iters; ++it) {
      st.sync();  // keep the warps of a CTA on the same code (shared instruction fetch), see TmemStore::sync
      sreal cn[UMPC_NX], rn[UMPC_NC];
      umpc_gen_norms(cn, rn, Ds, Es, a);
#pragma unroll
      for (int j = 0; j < UMPC_NX; ++j) {
        const sreal pc = cs * (sreal)umpc_px(par, j) * Ds[j] * Ds[j];
        const sreal nj = umpc_tmax(pc, cn[j] * Ds[j]);
        Ds[j] *= umpc_trsqrt(umpc_tlimit(nj));
      }
#pragma unroll
      for (int i = 0; i < UMPC_NC; ++i) Es[i] *= umpc_trsqrt(umpc_tlimit(rn[i] * Es[i]));
      // cost normalisation (scaling.c:100-131)
      sreal psum = 0.0f, nq = 0.0f;
#pragma unroll
      for (int j = 0; j < UMPC_NX; ++j) {
        const sreal px = (sreal)umpc_px(par, j);
        psum += cs * px * Ds[j] * Ds[j];
        if (j < 36) nq = umpc_tmax(nq, umpc_tabs(cs * Ds[j] * px * (sreal)umpc_ref(yd, dyd, j)));
      }
      sreal ct = psum * (1.0f / UMPC_NX);
      ct = umpc_tmax(ct, umpc_tlimit(nq));
      ct = umpc_tlimit(ct);
      cs *= umpc_trcp(ct);
    }
    c = cs;
#if defined(UMPC_REAL_IS_FLOAT) || !defined(UMPC_SCALE_FLOAT)
    sreal (&D)[UMPC_NX] = Ds;
    sreal (&E)[UMPC_NC] = Es;
#else
    ureal D[UMPC_NX], E[UMPC_NC];   // widened once
#pragma unroll
    for (int j = 0; j < UMPC_NX; ++j) D[j] = Ds[j];
#pragma unroll
    for (int i = 0; i < UMPC_NC; ++i) E[i] = Es[i];
#endif
    // ---- K = c P0 + sigma D^-2 + A0' W A0 and its LDL' ----------------------------------
    {
      st.sync();
      ureal G[UMPC_NX], W[UMPC_NC], sgv[UMPC_NX + 1];
      sgv[UMPC_NX] = 0.0f;
#pragma unroll
      for (int j = 0; j < UMPC_NX; ++j) {
        const ureal sg = par.sigma * UMPC_RCP(D[j] * D[j]);
        sgv[j] = sg;
        G[j] = c * umpc_px(par, j) + sg;
        io.st_D(j, D[j]);
      }
#pragma unroll
      for (int i = 0; i < UMPC_NC; ++i) {
        const ureal rho_i = i < UMPC_NEQ ? rho_eq : W3[i - UMPC_NEQ];
        W[i] = rho_i * E[i] * E[i];
        io.st_E(i, E[i]);
      }
#pragma unroll
      for (int cq = 0; cq < (UMPC_GEN_NWORDS - UMPC_GEN_OFF_SG) / UMPC_GEN_CHUNK; ++cq) {
        ureal ck[UMPC_GEN_CHUNK];
#pragma unroll
        for (int u = 0; u < UMPC_GEN_CHUNK; ++u) {
          const int q = cq * UMPC_GEN_CHUNK + u;
          ck[u] = q < (UMPC_NX + 1) / 2 ? umpc_pack2(sgv[q < (UMPC_NX + 1) / 2 ? 2 * q : 0], sgv[q < (UMPC_NX + 1) / 2 ? 2 * q + 1 : 0]) : (ureal)0;
        }
        st.setChunk(UMPC_GEN_OFF_SG / UMPC_GEN_CHUNK + cq, ck);
      }
      umpc_gen_factor(G, W, a, st);
      st.sync();
      st.end_writes();  // the factor is read back by the solve (asynchronous stores of a TMEM-backed scratch)
      // ---- warm start -> hatted, shifted coordinates; the two start vectors of w ---------
      // All global loads are issued before any global store (the compiler cannot move a load across
      // a store that might alias, and a load -> store -> load chain exposes the full memory latency
      // once per element).
      ureal zs0[UMPC_NC], ys0[UMPC_NC];
#pragma unroll
      for (int i = 0; i < UMPC_NC; ++i) { zs0[i] = io.ld_zs(i); ys0[i] = io.ld_ys(i); }
#pragma unroll
      for (int j = 0; j < UMPC_NX; ++j) xs[j] = io.ld_xs(j);
      ureal lo[UMPC_NEQ], xp[UMPC_NX];
#pragma unroll
      for (int i = 0; i < UMPC_NEQ; ++i) lo[i] = UMPC_LO(i);
      umpc_gen_xp(xp, lo, a);
      ureal v0[UMPC_NC], vf[UMPC_NC];  // minus W l: the -b of kappa rides along
#pragma unroll
      for (int i = 0; i < UMPC_NEQ; ++i) {
        const ureal zh = zs0[i] * UMPC_RCP(E[i]), yh = ys0[i] * E[i];
        v0[i] = W[i] * (zh - lo[i]) - yh;
        vf[i] = oma * W[i] * (lo[i] - zh) - yh;   // (2 - alpha) W l - yh - (1 - alpha) W zh - W l
      }
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        v0[UMPC_NEQ + k] = 0.0f; vf[UMPC_NEQ + k] = 0.0f;
        z3[k] = zs0[UMPC_NEQ + k] * UMPC_RCP(E[UMPC_NEQ + k]);
        y3[k] = ys0[UMPC_NEQ + k] * E[UMPC_NEQ + k];
        W3[k] = W[UMPC_NEQ + k];
        Wi3[k] = UMPC_RCP(W3[k]);
      }
      ureal w0[UMPC_NX], wf[UMPC_NX], kapo[UMPC_NEQ];
#pragma unroll
      for (int j = 0; j < UMPC_NX; ++j) {
        const ureal kap = c * umpc_px(par, j) * (umpc_ref(yd, dyd, j) - xp[j]);  // 0 on the input columns
        w0[j] = kap; wf[j] = kap;
        if (j < UMPC_NEQ) kapo[j] = kap;
        xs[j] = xs[j] * D[j] - xp[j];
      }
      st.sync();
      umpc_gen_ATv_add(w0, v0, a);
      umpc_gen_ATv_add(wf, vf, a);
      // constant of the input-column part: w_u = kappa_u + B_u' vhat, vhat = M^-T (w_o - kappa_o), kappa_u = 0
      // It is loop-invariant: parked in the spare words behind the sigma/D^2 pairs of the scratch (same chunk as the
      // pairs of the input columns, so the loop gets it with a load it issues anyway) instead of 9 registers.
      {
        ureal kpu[UMPC_NX - UMPC_NEQ];
        umpc_gen_gu_from_go(kpu, kapo, a);
#pragma unroll
        for (int i = 0; i < UMPC_NX - UMPC_NEQ; ++i) st.setW(UMPC_KPU_POS + i, -kpu[i]);
        st.end_writes();
      }
      // The first pass forms r with w0 but must continue from wf:  w1 = wf - alpha w0 + t  (t = the rest of the
      // update).  Written as the regular update (1 - alpha) X + t of a substitute X = (wf - alpha w0) / (1 - alpha),
      // so that the loop body has no first-iteration special case beyond swapping w (alpha != 1: OSQP's 1.6).
      const ureal ioma = 1.0f / oma;
#pragma unroll
      for (int j = 0; j < UMPC_NEQ; ++j) { w[j] = w0[j]; st.setVf(j, (wf[j] - alpha * w0[j]) * ioma); }
    }
  }
  if (par.debug_phase == 1) {
#pragma unroll
    for (int j = 0; j < UMPC_NX; ++j) io.st_solx(j, xs[j]);
#pragma unroll
    for (int j = 0; j < UMPC_NEQ; ++j) io.st_soly(j, w[j] + (j < 3 ? z3[j] + y3[j] + W3[j] : 0));
#pragma unroll
    for (int i = 0; i < 3; ++i) uquad[i] = 0;
#pragma unroll
    for (int i = 0; i < 6; ++i) accdes[i] = 0;
    return;
  }
  // ---- ADMM (osqp.c:352-500; fixed rho: adaptive_rho_interval = 0 in workspace.c:561) -----
  const ureal ac = alpha * c;
  const int max_iter = par.max_iter < 1 ? 1 : par.max_iter;
  // One SEGMENT of iterations, then the residuals.  C path (check_termination = 0): a single segment of max_iter
  // iterations.  Converged mode: segments of check_termination iterations until the lane's residuals pass
  // (osqp.c:431-470); a lane that has converged commits its solution at that point and idles along until the
  // whole warp is done (the scratch requests are warp-collective, so the trip count must be warp-uniform).
  // CONV is a compile-time flag: the C path's instantiation carries none of this.
  const int chk_every = CONV ? par.check_termination : 0;
  const int rho_every = CONV ? par.adaptive_rho_interval : 0;
  int iter_done = 0;
  bool done = false;
  for (;;) {
  int next_stop = max_iter;  // next iteration count at which the residuals are needed
  if constexpr (CONV) {
    if (chk_every > 0) next_stop = umpc_imin(next_stop, (iter_done / chk_every + 1) * chk_every);
    if (rho_every > 0) next_stop = umpc_imin(next_stop, (iter_done / rho_every + 1) * rho_every);
  }
  const int nseg = next_stop - iter_done;
  for (int it = 0; it < nseg; ++it) {
    st.begin_iteration(it);
    st.sync();
    ureal r[UMPC_NX];
    ureal m3[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) m3[k] = W3[k] * z3[k] - y3[k];
    // input-column part of w, re-derived from the own-column part every pass: g stays exactly in range(A_eq')
    ureal wu[UMPC_NX - UMPC_NEQ];
    umpc_gen_gu_from_go(wu, w, a);
#pragma unroll
    for (int gq = 0; gq < (UMPC_GEN_NWORDS - UMPC_GEN_OFF_SG) / UMPC_GEN_CHUNK; ++gq) {
      ureal wd[UMPC_GEN_CHUNK];
      st.issue(UMPC_GEN_OFF_SG / UMPC_GEN_CHUNK + gq, wd);
      st.waitc(UMPC_GEN_OFF_SG / UMPC_GEN_CHUNK + gq, wd);
#pragma unroll
      for (int u = 0; u < UMPC_GEN_CHUNK; ++u) {
        ureal s2[2];
        if (2 * (UMPC_GEN_CHUNK * gq + u) < UMPC_NX) umpc_unpack2(wd[u], s2[0], s2[1]);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int j = 2 * (UMPC_GEN_CHUNK * gq + u) + h;
          if (j < UMPC_NEQ) r[j < UMPC_NEQ ? j : 0] = s2[h] * xs[j < UMPC_NEQ ? j : 0] + w[j < UMPC_NEQ ? j : 0];
          else if (j < UMPC_NX) {  // input columns: w_u = kpu + B_u' vhat, kpu from the chunk's spare words
            const int i = j >= UMPC_NEQ && j < UMPC_NX ? j - UMPC_NEQ : 0;
            constexpr int kq = UMPC_KPU_POS - UMPC_GEN_OFF_SG - UMPC_GEN_CHUNK * ((UMPC_KPU_POS - UMPC_GEN_OFF_SG) / UMPC_GEN_CHUNK);
            static_assert((UMPC_NX - 1) / 2 / UMPC_GEN_CHUNK == (UMPC_KPU_POS - UMPC_GEN_OFF_SG) / UMPC_GEN_CHUNK, "kpu rides in the chunk of the input columns' pairs");
            r[j < UMPC_NX ? j : 0] = s2[h] * xs[j < UMPC_NX ? j : 0] + (wu[i] + wd[kq + i < UMPC_GEN_CHUNK ? kq + i : 0]);
          }
        }
      }
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) r[36 + 3 * k] += m3[k];
    if (iter_done + it == 0) {
#pragma unroll
      for (int j = 0; j < UMPC_NEQ; ++j) w[j] = st.Vf(j);
    }
    umpc_gen_solve(r, st);
#pragma unroll
    for (int gq = 0; gq < (UMPC_GEN_NWORDS - UMPC_GEN_OFF_SG) / UMPC_GEN_CHUNK; ++gq) {
      ureal wd[UMPC_GEN_CHUNK];
      st.issue(UMPC_GEN_OFF_SG / UMPC_GEN_CHUNK + gq, wd);
      st.waitc(UMPC_GEN_OFF_SG / UMPC_GEN_CHUNK + gq, wd);
#pragma unroll
      for (int u = 0; u < UMPC_GEN_CHUNK; ++u) {
        ureal s2[2];
        if (2 * (UMPC_GEN_CHUNK * gq + u) < UMPC_NX) umpc_unpack2(wd[u], s2[0], s2[1]);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int j = 2 * (UMPC_GEN_CHUNK * gq + u) + h;
          if (j < UMPC_NX) {
            const ureal d = xs[j] - r[j];
            if (j < UMPC_NEQ) {
              const ureal t = (ac * umpc_px(par, j)) * r[j] - alpha * (s2[h] * d);
              w[j < UMPC_NEQ ? j : 0] = oma * w[j < UMPC_NEQ ? j : 0] + t;
            }
            xs[j] -= alpha * d;
          }
        }
      }
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const ureal t = alpha * r[36 + 3 * k] + oma * z3[k];
      const ureal zn = umpc_fmin(umpc_fmax(t + Wi3[k] * y3[k], lo_thr), hi_thr);
      y3[k] += W3[k] * (t - zn);
      z3[k] = zn;
    }
  }
  iter_done += nseg;
  if (par.debug_phase == 2) {
#pragma unroll
    for (int j = 0; j < UMPC_NX; ++j) io.st_solx(j, xs[j]);
#pragma unroll
    for (int j = 0; j < UMPC_NEQ; ++j) io.st_soly(j, w[j] + (j < 3 ? z3[j] + y3[j] : 0));
#pragma unroll
    for (int i = 0; i < 3; ++i) uquad[i] = 0;
#pragma unroll
    for (int i = 0; i < 6; ++i) accdes[i] = 0;
    return;
  }
  // ---- back to un-shifted coordinates; equality-row duals from g = w - kappa ----------------
  st.sync();
  bool again = false;
  {
    ureal zs[UMPC_NC], ys[UMPC_NC], xf[UMPC_NX];  // xf: the iterate in un-shifted coordinates
    ureal qp_lo[18], yd[6], dyd[6], R0b[9], dq0b[6];
    UmpcAVals a;
    {
      ureal p0b[3], pdb[3], dpdb[3], sdb[3];
#pragma unroll
      for (int i = 0; i < 3; ++i) { p0b[i] = io.unstash(i); pdb[i] = io.unstash(18 + i); dpdb[i] = io.unstash(21 + i); sdb[i] = io.unstash(24 + i); }
#pragma unroll
      for (int i = 0; i < 9; ++i) R0b[i] = io.unstash(3 + i);
#pragma unroll
      for (int i = 0; i < 6; ++i) dq0b[i] = io.unstash(12 + i);
      UmpcQP qp;
      umpc_build(par, qp, p0b, R0b, dq0b, pdb, dpdb, sdb, io.unstash(27));
      a = qp.a;
#pragma unroll
      for (int i = 0; i < 6; ++i) { yd[i] = qp.yd[i]; dyd[i] = qp.dyd[i]; qp_lo[i] = qp.lo[i]; }
#pragma unroll
      for (int i = 18; i < 30; ++i) qp_lo[i - 12] = qp.lo[i];
    }
    ureal Dl[UMPC_NX], El[UMPC_NC];  // global loads first, stores after (see the warm-start note)
#pragma unroll
    for (int j = 0; j < UMPC_NX; ++j) Dl[j] = io.ld_D(j);
#pragma unroll
    for (int i = 0; i < UMPC_NC; ++i) El[i] = io.ld_E(i);
    {
      ureal lo[UMPC_NEQ], xp[UMPC_NX], v[UMPC_NEQ];
#pragma unroll
      for (int i = 0; i < UMPC_NEQ; ++i) lo[i] = UMPC_LO(i);
      umpc_gen_xp(xp, lo, a);
      ureal g[UMPC_NX];
#pragma unroll
      for (int j = 0; j < UMPC_NX; ++j) {
        // own columns of g - b = A_eq'(v - W l); the input columns are not read by the substitution
        g[j] = j < UMPC_NEQ ? w[j < UMPC_NEQ ? j : 0] - c * umpc_px(par, j) * (umpc_ref(yd, dyd, j) - xp[j]) : 0.0f;
        xf[j] = xs[j] + xp[j];
      }
      umpc_gen_v_from_g(v, g, a);  // v - W l
#pragma unroll
      for (int i = 0; i < UMPC_NEQ; ++i) { zs[i] = lo[i]; ys[i] = -v[i]; }  // yh = W l - v
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) { zs[UMPC_NEQ + k] = z3[k]; ys[UMPC_NEQ + k] = y3[k]; }
    // ---- residuals, status (update_info + check_termination, auxil.c:567-790) ---------------
    st.sync();
    // un-scaled norms for the termination test; scaled ones (E, D weighted) for the rho estimate (auxil.c:13-75
    // works on the scaled residual vectors left in z_prev / x_prev)
    const bool adapt = rho_every > 0;
    constexpr bool full = CONV || INFO;
    ureal pri = 0.0f, nz = 0.0f, nAx = 0.0f, pri_s = 0.0f, pn_s = 0.0f;
    if constexpr (full) {
      ureal Ax[UMPC_NC];
      umpc_gen_Ax(Ax, xf, a);
#pragma unroll
      for (int i = 0; i < UMPC_NC; ++i) {
        pri = umpc_fmax(pri, umpc_fabs(Ax[i] - zs[i]));
        nz = umpc_fmax(nz, umpc_fabs(zs[i]));
        nAx = umpc_fmax(nAx, umpc_fabs(Ax[i]));
        if (adapt) {
          pri_s = umpc_fmax(pri_s, umpc_fabs(El[i] * (Ax[i] - zs[i])));
          pn_s = umpc_fmax(pn_s, El[i] * umpc_fmax(umpc_fabs(zs[i]), umpc_fabs(Ax[i])));
        }
      }
    }
    ureal dua = 0.0f, nq = 0.0f, nAty = 0.0f, nPx = 0.0f, dua_s = 0.0f, dn_s = 0.0f;
    if constexpr (full) {
      ureal Aty[UMPC_NX];
#pragma unroll
      for (int j = 0; j < UMPC_NX; ++j) Aty[j] = 0.0f;
      umpc_gen_ATv_add(Aty, ys, a);
#pragma unroll
      for (int j = 0; j < UMPC_NX; ++j) {
        const ureal px = c * umpc_px(par, j) * xf[j];
        const ureal qj = -c * umpc_px(par, j) * umpc_ref(yd, dyd, j);
        dua = umpc_fmax(dua, umpc_fabs(px + qj + Aty[j]));
        nq = umpc_fmax(nq, umpc_fabs(qj)); nAty = umpc_fmax(nAty, umpc_fabs(Aty[j])); nPx = umpc_fmax(nPx, umpc_fabs(px));
        if (adapt) {
          dua_s = umpc_fmax(dua_s, umpc_fabs(Dl[j] * (px + qj + Aty[j])));
          dn_s = umpc_fmax(dn_s, Dl[j] * umpc_fmax(umpc_fabs(qj), umpc_fmax(umpc_fabs(Aty[j]), umpc_fabs(px))));
        }
      }
    }
    st.sync();
    const ureal cinv = UMPC_RCP(c);
    dua *= cinv;
    const ureal ep = par.eps_abs + par.eps_rel * umpc_fmax(nz, nAx);
    const ureal ed = par.eps_abs + par.eps_rel * cinv * umpc_fmax(umpc_fmax(nq, nAty), nPx);
    const bool last = iter_done >= max_iter;
    bool check = last;
    if constexpr (CONV) check = last || (chk_every > 0 && iter_done % chk_every == 0);
    int status;
    if constexpr (!full) {
      // no residuals: an iterate that left the float range (or is NaN) is what drives a residual past OSQP_INFTY
      ureal mx = 0.0f;
#pragma unroll
      for (int j = 0; j < UMPC_NX; ++j) mx = umpc_fmax(mx, umpc_fabs(xf[j]));
#pragma unroll
      for (int i = 0; i < UMPC_NC; ++i) mx = umpc_fmax(mx, umpc_fabs(ys[i]));
      ureal sm = 0.0f;  // fmax drops NaNs: a sum keeps them
#pragma unroll
      for (int k = 0; k < 3; ++k) sm += xf[36 + k] + xf[18 + k] + xf[21 + k];
      status = (mx > 1e30f || sm != sm) ? UMPC_ST_NON_CVX : UMPC_ST_NOT_EVALUATED;
      pri = umpc_nan(); dua = umpc_nan();
    } else
    if (pri > 1e30f || dua > 1e30f) status = UMPC_ST_NON_CVX;  // OSQP_INFTY (auxil.c:706-716)
    else if (!check) status = UMPC_ST_UNSOLVED;
    else if (pri < ep && dua < ed) status = UMPC_ST_SOLVED;
    else if (last && pri < 10.0f * ep && dua < 10.0f * ed) status = UMPC_ST_SOLVED_INACCURATE;  // osqp.c:564-568
    else status = last ? UMPC_ST_MAX_ITER : UMPC_ST_UNSOLVED;
    // ---- store_solution (auxil.c:527-562); warm start back to OSQP's scaled coordinates ---
    // committed once per lane: when its test passes, or at the iteration limit
    const bool commit = !done && status != UMPC_ST_UNSOLVED;
    if (commit) {
      info.iter = iter_done; info.pri_res = pri; info.dua_res = dua; info.status = status;
      const bool bad = status == UMPC_ST_NON_CVX;
#pragma unroll
      for (int j = 0; j < UMPC_NX; ++j) {
        if constexpr (full) io.st_solx(j, xf[j]);
        io.st_xs(j, xf[j] * UMPC_RCP(Dl[j]));
      }
#pragma unroll
      for (int i = 0; i < UMPC_NC; ++i) {
        const ureal e = El[i];
        if constexpr (full) io.st_soly(i, ys[i] * cinv);
        io.st_zs(i, zs[i] * e);
        io.st_ys(i, ys[i] * UMPC_RCP(e));
      }
      if constexpr (CONV) io.st_rho(rho);
      // ---- extraction (uprightmpc2.c:252-269) ----------------------------------------------
      T0 += xf[36];
      uquad[0] = T0; uquad[1] = xf[37]; uquad[2] = xf[38];
      const ureal t0 = R0b[0] * xf[21] + R0b[1] * xf[22] + R0b[2] * xf[23];
      const ureal t1 = R0b[3] * xf[21] + R0b[4] * xf[22] + R0b[5] * xf[23];
      const ureal dq1[6] = {xf[18], xf[19], xf[20], -t1, t0, 0.0f};
#pragma unroll
      for (int i = 0; i < 6; ++i) accdes[i] = (dq1[i] - dq0b[i]) / dt;
      if (bad) {
        // residuals beyond OSQP_INFTY: solution NaN, iterates cold-started (auxil.c:538-561).  A branch of its own, after
        // the regular stores, so that the normal path carries no per-element selects.
        const ureal nanv = umpc_nan();
#pragma unroll
        for (int j = 0; j < UMPC_NX; ++j) { io.st_solx(j, nanv); io.st_xs(j, 0.0f); }
#pragma unroll
        for (int i = 0; i < UMPC_NC; ++i) { io.st_soly(i, nanv); io.st_zs(i, 0.0f); io.st_ys(i, 0.0f); }
        T0 = nanv;
#pragma unroll
        for (int i = 0; i < 3; ++i) uquad[i] = nanv;
#pragma unroll
        for (int i = 0; i < 6; ++i) accdes[i] = nanv;
      }
      done = true;
    }
    if constexpr (CONV) again = !st.all(done);
    // ---- adaptive rho (auxil.c:13-75 + osqp_update_rho, osqp.c:1270-1330): per lane; the refactor is executed by
    // the whole warp when any live lane moved (lanes that did not keep their rho: same factor again)
    if constexpr (CONV) if (again && adapt && iter_done % rho_every == 0) {
      const ureal pr = pri_s * UMPC_RCP(pn_s + 1e-10f), dr = dua_s * UMPC_RCP(dn_s + 1e-10f);
      ureal est = rho * umpc_sqrt(pr * UMPC_RCP(dr + 1e-10f));
      est = umpc_fmin(umpc_fmax(est, 1e-6f), 1e6f);  // RHO_MIN, RHO_MAX (constants.h:69-70)
      const bool move = !done && (est > rho * par.adaptive_rho_tol || est < rho * UMPC_RCP(par.adaptive_rho_tol));
      if (!st.all(!move)) {
        if (move) rho = est;
        ureal G[UMPC_NX], W[UMPC_NC];
#pragma unroll
        for (int j = 0; j < UMPC_NX; ++j) G[j] = c * umpc_px(par, j) + par.sigma * UMPC_RCP(Dl[j] * Dl[j]);
#pragma unroll
        for (int i = 0; i < UMPC_NC; ++i) W[i] = (i < UMPC_NEQ ? 1e3f : thr_mult[i < UMPC_NEQ ? 0 : i - UMPC_NEQ]) * rho * El[i] * El[i];
#pragma unroll
        for (int k = 0; k < 3; ++k) { W3[k] = W[UMPC_NEQ + k]; Wi3[k] = UMPC_RCP(W3[k]); }
        umpc_gen_factor(G, W, a, st);
        st.end_writes();
      }
    }
  }
  if (!again) break;
  }
#undef UMPC_LO
}
