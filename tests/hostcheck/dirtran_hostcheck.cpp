// tests/hostcheck/dirtran_hostcheck.cpp -- TEST INFRASTRUCTURE: host build of robobee3d_b200/csrc/dirtran_core.cuh
// (the exact per-environment code the CUDA kernels of dirtran.cu run), so that the algorithm can be checked against
// oracle/ on the CPU box.  Never part of the product.
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include "../../robobee3d_b200/csrc/dirtran_core.cuh"

using namespace dtcore;

struct HC {
  DevPar P;
  double* slab;
  int rows;
  Env e;
};

extern "C" {

// par: DtParams as the C-ABI declares it
void* dthc_create(const DtParams* par) {
  HC* h = new HC();
  DevPar& d = h->P;
  memset(&d, 0, sizeof(d));
  d.N = par->N; d.max_iter = par->max_iter; d.check_termination = par->check_termination;
  d.adaptive_rho_interval = par->adaptive_rho_interval; d.nsub = par->nsub;
  d.dt = par->dt; d.sim_dt = par->sim_dt; d.sigma = par->sigma; d.alpha = par->alpha; d.eps_abs = par->eps_abs; d.eps_rel = par->eps_rel;
  d.rho0 = par->rho; d.adaptive_rho_tol = par->adaptive_rho_tol; d.g = par->g; d.ibmb_inv = 1.0 / par->ibmb;
  for (int j = 0; j < NW; ++j) {
    double lo = j < NX ? par->xmin[j] : par->umin[j - NX], hi = j < NX ? par->xmax[j] : par->umax[j - NX];
    lo = fmax(lo, -1e20); hi = fmin(hi, 1e20);
    d.lo[j] = lo; d.hi[j] = hi;
    if (lo < -1e16 && hi > 1e16) d.loose |= 1u << j;
    if (j < NX) d.wx[j] = par->wx[j]; else d.wu[j - NX] = par->wu[j - NX];
  }
  h->rows = R_HDR + K_STRIDE * (par->N + 1);
  h->slab = (double*)calloc((size_t)h->rows * LANES, sizeof(double));
  h->e.base = h->slab;
  h->e.h(R_RHO) = par->rho;
  return h;
}
void dthc_destroy(void* p) { HC* h = (HC*)p; free(h->slab); delete h; }
void dthc_set(void* p, const double* ctrl, const double* xr) {
  HC* h = (HC*)p;
  if (ctrl) { h->e.h(R_CTRL) = ctrl[0]; h->e.h(R_CTRL + 1) = ctrl[1]; }
  if (xr) for (int i = 0; i < NX; ++i) h->e.h(R_XR + i) = xr[i];
}
// one update; info: iter,status,pri,dua,rho,rho_updates
void dthc_update(void* p, const double* x0, double* u, double* info) {
  HC* h = (HC*)p;
  double y[NX];
  for (int i = 0; i < NX; ++i) y[i] = x0[i];
  double u0 = h->e.h(R_CTRL), u1 = h->e.h(R_CTRL + 1);
  mpc_update(h->P, h->e, y, u0, u1);
  u[0] = u0; u[1] = u1;
  if (info) {
    for (int i = 0; i < 4; ++i) info[i] = h->e.h(R_INFO + i);
    info[4] = h->e.h(R_RHO); info[5] = h->e.h(R_INFO + 4);
  }
}
// nsteps of the fused loop (update + plant) exactly as dt_rollout_kernel runs it; x in/out, u = last input
void dthc_rollout(void* p, double* x, double* u, int nsteps) {
  HC* h = (HC*)p;
  double y[NX];
  for (int i = 0; i < NX; ++i) y[i] = x[i];
  double u0 = h->e.h(R_CTRL), u1 = h->e.h(R_CTRL + 1);
  mpc_run<true>(h->P, h->e, y, u0, u1, nsteps);
  for (int i = 0; i < NX; ++i) x[i] = y[i];
  u[0] = u0; u[1] = u1;
}
void dthc_plant(void* p, double* x, const double* u) {
  HC* h = (HC*)p;
  double y[NX];
  for (int i = 0; i < NX; ++i) y[i] = x[i];
  for (int t = 0; t < h->P.nsub; ++t) {
    double f[NX];
    dydt(h->P, y, u[0], u[1], f);
    for (int i = 0; i < NX; ++i) y[i] += f[i] * h->P.sim_dt;
  }
  for (int i = 0; i < NX; ++i) x[i] = y[i];
}
// solution in OSQP's ordering
void dthc_solution(void* p, double* x, double* yv) {
  HC* h = (HC*)p;
  const int N = h->P.N, n = (N + 1) * NX + N * NU;
  for (int k = 0; k <= N; ++k) {
    const double* kb = h->e.knot(k);
    for (int i = 0; i < NX; ++i) {
      x[k * NX + i] = kb[(K_X + i) * LANES];
      yv[k * NX + i] = kb[(K_YE + i) * LANES];
      yv[(N + 1) * NX + k * NX + i] = kb[(K_YI + i) * LANES];
    }
    if (k < N) for (int i = 0; i < NU; ++i) {
      x[(N + 1) * NX + k * NU + i] = kb[(K_X + NX + i) * LANES];
      yv[2 * (N + 1) * NX + k * NU + i] = kb[(K_YI + NX + i) * LANES];
    }
  }
  (void)n;
}
}
