// robobee3d_b200/csrc/wl_core.cuh
//
// Wrench-linearisation stage, one environment per thread.  Reference (under template/):
//   FunApprox  w_i(u) = a0 + a1'u + 0.5 u'A2 u, dw_i/du = a1 + A2 u    uprightmpc2/funapprox.c:35-96
//   WLCon      projected-gradient step                                   uprightmpc2/funapprox.c:98-165
//   WLQP       min 0.5 du'P du + q'du, L<=du<=U, P=A1'QA1, q=A1'Q a0     wlcontroller/c/wlqp.c:77-123
//              solved by the vendored OSQP (n=m=4, A=I; wlcontroller/wlqp/workspace.c:42-84),
//              40 iterations, termination test every 25 (C) or never (C++ wlqp.cpp:63-66)
//   stateful wrapper wlControllerUpdate (u0 <- u)                        wlcontroller/c/wlcontroller.c:35-45
#pragma once
#include <math.h>

#ifndef UMPC_HD
#if defined(__CUDACC__)
#define UMPC_HD __host__ __device__ __forceinline__
#else
#define UMPC_HD inline
#endif
#endif

#define WL_NU 4
#define WL_NW 6

// popts[90] = 6 x (a0, a1[4], triu(A2)[10] row-major)  (funApproxInit, funapprox.c:35-52)
struct WlParams {
  float popts[90];
  float umin[4], umax[4], dumax[4];  // dumax already divided by controlRate
  float Qw[6];
  // OSQP settings of the generated 4x4 instance + wlqp.c:43-46 overrides
  float rho, sigma, alpha, eps_abs, eps_rel;
  int max_iter, check_termination, scaling_iters;
};

// w[6] and A1 = dw/du (6x4, A1[i][j]) at u
UMPC_HD void wl_wrench(const WlParams& wp, const float u[4], float w[6], float A1[6][4]) {
#pragma unroll
  for (int i = 0; i < WL_NW; ++i) {
    const float* po = &wp.popts[15 * i];
    float A2[4][4];
    int kk = 5;
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int c = r; c < 4; ++c) { A2[r][c] = po[kk]; A2[c][r] = po[kk]; ++kk; }
    float Au[4], lin = 0.0f, quad = 0.0f;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      float acc = 0.0f;
#pragma unroll
      for (int c = 0; c < 4; ++c) acc += A2[r][c] * u[c];
      Au[r] = acc;
      lin += u[r] * po[1 + r];
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) { quad += u[r] * Au[r]; A1[i][r] = po[1 + r] + Au[r]; }
    w[i] = po[0] + lin + 0.5f * quad;
  }
}

UMPC_HD void wl_common(const WlParams& wp, const float u0[4], const float h0[6], const float pdotdes[6], float w0[6],
                       float A1[6][4], float a0[6], float L[4], float U[4]) {
  wl_wrench(wp, u0, w0, A1);
#pragma unroll
  for (int i = 0; i < WL_NW; ++i) a0[i] = w0[i] - h0[i] - pdotdes[i];
#pragma unroll
  for (int j = 0; j < WL_NU; ++j) {
    L[j] = -wp.dumax[j]; U[j] = wp.dumax[j];
    if (u0[j] < wp.umin[j]) L[j] = 0.0f;       // do not reduce further
    else if (u0[j] > wp.umax[j]) U[j] = 0.0f;  // do not increase further
  }
}

// wlConUpdate: u0 += clip(-1e3 A1' Qw a0, L, U); returns w0 = w(u0 before the step)
UMPC_HD void wl_con_update(const WlParams& wp, float u0[4], float w0[6], const float h0[6], const float pdotdes[6]) {
  float A1[6][4], a0[6], L[4], U[4];
  wl_common(wp, u0, h0, pdotdes, w0, A1, a0, L, U);
#pragma unroll
  for (int j = 0; j < WL_NU; ++j) {
    float acc = 0.0f;
#pragma unroll
    for (int i = 0; i < WL_NW; ++i) acc += A1[i][j] * (wp.Qw[i] * a0[i]);
    float d = -1e3f * acc;
    d = d < L[j] ? L[j] : (d > U[j] ? U[j] : d);
    u0[j] += d;
  }
}

// Persistent OSQP state of one WLQP instance (scaled coordinates, like the reference's
// global workspace): warm start, last scaling, last linear cost.
struct WlqpState {
  float xs[4], zs[4], ys[4], D[4], E[4], c, q0[4];
};

UMPC_HD void wlqp_state_init(WlqpState& s) {
#pragma unroll
  for (int i = 0; i < 4; ++i) { s.xs[i] = s.zs[i] = s.ys[i] = s.q0[i] = 0.0f; s.D[i] = s.E[i] = 1.0f; }
  s.c = 1.0f;
}

UMPC_HD float wl_limit_scaling(float v) {
  v = v < 1e-4f ? 1.0f : v;
  return v > 1e4f ? 1e4f : v;
}

// wlqpUpdate + wlControllerUpdate: u0 <- u0 + argmin.  Same un-scaled ("hatted") ADMM as
// umpc_core.cuh with A0 = I: K = c P + sigma D^-2 + diag(rho E^2), dense 4x4 LDL'.
UMPC_HD void wlqp_update(const WlParams& wp, WlqpState& s, float u0[4], float w0[6], const float h0[6],
                         const float pdotdes[6], int* iter_out, int* status_out) {
  float A1[6][4], a0[6], L[4], U[4], P[4][4], q[4];
  wl_common(wp, u0, h0, pdotdes, w0, A1, a0, L, U);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float acc = 0.0f;
#pragma unroll
      for (int k = 0; k < WL_NW; ++k) acc += A1[k][i] * (wp.Qw[k] * A1[k][j]);
      P[i][j] = acc;
    }
    float acc = 0.0f;
#pragma unroll
    for (int k = 0; k < WL_NW; ++k) acc += A1[k][i] * (wp.Qw[k] * a0[k]);
    q[i] = acc;
  }
  // upper triangle is what osqp_update_P receives (wlqp.c:52-62)
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < i; ++j) P[i][j] = P[j][i];
  // ---- osqp_update_P: re-equilibrate with the PREVIOUS linear cost (wlqp.c:64-68 order) ----
  float D[4] = {1.0f, 1.0f, 1.0f, 1.0f}, E[4] = {1.0f, 1.0f, 1.0f, 1.0f}, c = 1.0f;
  for (int it = 0; it < wp.scaling_iters; ++it) {
    float Dt[4], Et[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float nj = E[j] * D[j];  // A_s = E D (A = I)
#pragma unroll
      for (int i = 0; i < 4; ++i) nj = fmaxf(nj, fabsf(c * P[i][j] * D[i] * D[j]));
      Dt[j] = 1.0f / sqrtf(wl_limit_scaling(nj));
      Et[j] = 1.0f / sqrtf(wl_limit_scaling(E[j] * D[j]));
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) { D[j] *= Dt[j]; E[j] *= Et[j]; }
    float psum = 0.0f, nq = 0.0f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float nj = 0.0f;
#pragma unroll
      for (int i = 0; i < 4; ++i) nj = fmaxf(nj, fabsf(c * P[i][j] * D[i] * D[j]));
      psum += nj;
      nq = fmaxf(nq, fabsf(c * D[j] * s.q0[j]));
    }
    float ct = fmaxf(psum * 0.25f, wl_limit_scaling(nq));
    ct = wl_limit_scaling(ct);
    c *= 1.0f / ct;
  }
  // ---- constraint types with the NEW scaling (osqp_update_bounds comes last) ----------------
  float W[4], Wi[4], K[4][4], Lf[4][4], Dg[4], sg[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float ls_ = E[j] * L[j], us_ = E[j] * U[j];
    const float rho = (us_ - ls_ < 1e-4f) ? 1e3f * wp.rho : wp.rho;
    W[j] = rho * E[j] * E[j];
    Wi[j] = 1.0f / W[j];
    sg[j] = wp.sigma / (D[j] * D[j]);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) K[i][j] = c * P[i][j] + (i == j ? sg[i] + W[i] : 0.0f);
  // dense LDL'
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    float d = K[j][j];
#pragma unroll
    for (int k = 0; k < j; ++k) d -= Lf[j][k] * Lf[j][k] * Dg[k];
    Dg[j] = d;
#pragma unroll
    for (int i = j + 1; i < 4; ++i) {
      float v = K[i][j];
#pragma unroll
      for (int k = 0; k < j; ++k) v -= Lf[i][k] * Lf[j][k] * Dg[k];
      Lf[i][j] = v / d;
    }
  }
  // ---- warm start (stale scaling) -> hatted -------------------------------------------------
  float xh[4], zh[4], yh[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) { xh[j] = s.xs[j] * D[j]; zh[j] = s.zs[j] / E[j]; yh[j] = s.ys[j] * E[j]; }
  const float alpha = wp.alpha, oma = 1.0f - wp.alpha;
  int iter = 0, status = -10;
  float pri = 0.0f, dua = 0.0f, ep = 0.0f, ed = 0.0f;
  const float cinv = 1.0f / c;
  bool checked = false;
  for (iter = 1; iter <= wp.max_iter; ++iter) {
    float r[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) r[j] = sg[j] * xh[j] - c * q[j] + (W[j] * zh[j] - yh[j]);
    // forward / diagonal / backward
#pragma unroll
    for (int i = 1; i < 4; ++i)
#pragma unroll
      for (int k = 0; k < i; ++k) r[i] -= Lf[i][k] * r[k];
#pragma unroll
    for (int i = 0; i < 4; ++i) r[i] /= Dg[i];
#pragma unroll
    for (int i = 2; i >= 0; --i)
#pragma unroll
      for (int k = i + 1; k < 4; ++k) r[i] -= Lf[k][i] * r[k];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float t = alpha * r[j] + oma * zh[j];
      xh[j] = alpha * r[j] + oma * xh[j];
      const float zn = fminf(fmaxf(t + Wi[j] * yh[j], L[j]), U[j]);
      yh[j] += W[j] * (t - zn);
      zh[j] = zn;
    }
    checked = wp.check_termination > 0 && (iter % wp.check_termination == 0);
    if (checked || iter == wp.max_iter) {
      // residuals (auxil.c:243-361) in hatted form: Ax = xh
      float nz = 0.0f, nAx = 0.0f, nq = 0.0f, nAty = 0.0f, nPx = 0.0f;
      pri = 0.0f; dua = 0.0f;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float px = 0.0f;
#pragma unroll
        for (int k = 0; k < 4; ++k) px += c * P[j][k] * xh[k];
        pri = fmaxf(pri, fabsf(xh[j] - zh[j]));
        nz = fmaxf(nz, fabsf(zh[j])); nAx = fmaxf(nAx, fabsf(xh[j]));
        dua = fmaxf(dua, fabsf(px + c * q[j] + yh[j]));
        nq = fmaxf(nq, fabsf(c * q[j])); nAty = fmaxf(nAty, fabsf(yh[j])); nPx = fmaxf(nPx, fabsf(px));
      }
      dua *= cinv;
      ep = wp.eps_abs + wp.eps_rel * fmaxf(nz, nAx);
      ed = wp.eps_abs + wp.eps_rel * cinv * fmaxf(fmaxf(nq, nAty), nPx);
      if (pri < ep && dua < ed) { status = 1; break; }
    }
  }
  if (iter > wp.max_iter) iter = wp.max_iter;
  if (status != 1) status = (pri < 10.0f * ep && dua < 10.0f * ed) ? 2 : -2;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    u0[j] += xh[j];  // solution.x = D x_s = xh ; u = u0 + du ; g_u0 <- u
    s.xs[j] = xh[j] / D[j]; s.zs[j] = zh[j] * E[j]; s.ys[j] = yh[j] / E[j];
    s.D[j] = D[j]; s.E[j] = E[j]; s.q0[j] = q[j];
  }
  s.c = c;
  if (iter_out) *iter_out = iter;
  if (status_out) *status_out = status;
}
