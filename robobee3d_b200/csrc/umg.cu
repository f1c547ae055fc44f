// robobee3d_b200/csrc/umg.cu -- UprightMPC2 for any horizon N, optionally fused with the WLQP (include/umg_b200.h).
//
//   QP variables     x = [ y_2..y_{N+1} (N ny) | dy_1..dy_N (N ny) | u_0..u_{N-1} (N nu) | du (4, fused only) ]
//   A pattern        initConstraint        template_controllers.py:28-63, uprightmpc2_withWLQP.py:23-68
//   l, u, A values   updateConstraint      template_controllers.py:65-125, uprightmpc2_withWLQP.py:70-133
//   P, q             updateObjective       template_controllers.py:127-143, uprightmpc2_withWLQP.py:145-191
//   projection, state, outputs             template_controllers.py:226-258, uprightmpc2_withWLQP.py:264-324
// One thread per environment builds / extracts; the solve is gqp.cu (one warp per QP, float64).
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <string>
#include <vector>

#include "../../include/umpc_b200.h"
#include "../../include/gqp_b200.h"
#include "../../include/umg_b200.h"

void umpc_internal_set_error(const char* msg);

namespace umg {

enum { NY = 6, NU = 3 };
// kinds of A entries
enum { K_P1 = 0, K_M1 = 1, K_DT = 2, K_DTT0 = 3, K_S = 4, K_B = 5 };
struct Ent { int col, kind, idx; };

// the structural non-zeros of row r, columns increasing (both the host pattern and the device builder walk this)
__host__ __device__ inline int row_entries(int N, int wl, int r, Ent* e) {
  const int n1 = N * NY, n2 = 2 * N * NY, n3 = n2 + N * NU;
  int c = 0;
  if (r < n1) {
    const int k = r / NY, i = r % NY;
    if (k > 0) e[c++] = Ent{(k - 1) * NY + i, K_P1, 0};
    e[c++] = Ent{k * NY + i, K_M1, 0};
    e[c++] = Ent{n1 + k * NY + i, K_DT, 0};
  } else if (r < n2) {
    const int rr = r - n1, k = rr / NY, i = rr % NY;
    if (k > 1 && i < 3) e[c++] = Ent{(k - 2) * NY + 3 + i, K_DTT0, 0};
    if (k > 0) e[c++] = Ent{n1 + (k - 1) * NY + i, K_P1, 0};
    e[c++] = Ent{n1 + k * NY + i, K_M1, 0};
    if (i < 3) e[c++] = Ent{n2 + k * NU, K_S, i};
    else { e[c++] = Ent{n2 + k * NU + 1, K_B, i - 3}; e[c++] = Ent{n2 + k * NU + 2, K_B, i - 3 + 3}; }
  } else if (r < n2 + N) {
    e[c++] = Ent{n2 + 3 * (r - n2), K_P1, 0};
  } else if (wl) {
    e[c++] = Ent{n3 + (r - n2 - N), K_P1, 0};
  }
  return c;
}

struct Par {
  int N, wl, nx, nc, nnz, wrench;
  double dt, g, Tmax, Qyr[6], Qyf[6], Qdyr[6], Qdyf[6], R[3], Ibi[3], M0[6];
  double umin[4], umax[4], dumax[4], Qw[6];
  double popts[90];
};

struct Io {  // per-env device arrays
  double *in;      // [B][28]: p0 3, R0 9, dq0 6, pdes 3, dpdes 3, sdes 3, actualT0 1
  double *T0, *u0; // [B], [B][4]
  double *out;     // [B][13]: uquad 3, accdes 6, u_wl 4
};

// w(u) and dw/du (6 x 4, row-major A1[i][j]) of the quadratic model (uprightmpc2/funapprox.c:35-96)
__device__ inline void wrench(const Par& p, const double u[4], double w[6], double A1[6][4]) {
  if (p.wrench == 0) {  // the reference's stand-in (uprightmpc2_withWLQP.py:268-271)
    for (int i = 0; i < 6; ++i) { w[i] = 0.0; for (int j = 0; j < 4; ++j) A1[i][j] = (i - 2 == j) ? 1.0 : 0.0; }
    return;
  }
  for (int i = 0; i < 6; ++i) {
    const double* po = p.popts + 15 * i;
    double A2[4][4];
    int kk = 5;
    for (int r = 0; r < 4; ++r) for (int c = r; c < 4; ++c) { A2[r][c] = po[kk]; A2[c][r] = po[kk]; ++kk; }
    double lin = 0, quad = 0;
    for (int r = 0; r < 4; ++r) {
      double acc = 0;
      for (int c = 0; c < 4; ++c) acc += A2[r][c] * u[c];
      A1[i][r] = po[1 + r] + acc;
      lin += u[r] * po[1 + r]; quad += u[r] * acc;
    }
    w[i] = po[0] + lin + 0.5 * quad;
  }
}

__global__ void build_kernel(const __grid_constant__ Par p, Io io, int nenv, double* __restrict__ Pq, double* __restrict__ Av,
                             double* __restrict__ q, double* __restrict__ l, double* __restrict__ u, const int* __restrict__ rowptr) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= nenv) return;
  const int N = p.N, nx = p.nx, nc = p.nc;
  const int n1 = N * NY, n2 = 2 * N * NY, n3 = n2 + N * NU;
  const double* in = io.in + (size_t)e * 28;
  const double *p0 = in, *R0 = in + 3, *dq0 = in + 12, *pdes = in + 18, *dpdes = in + 21, *sdes = in + 24;
  double T0 = io.T0[e];
  if (in[27] >= 0.0) { T0 = in[27]; io.T0[e] = T0; }
  const double dt = p.dt;
  // template projection (template_controllers.py:226-237): s0, ds0 = -R0 e3h w, Btau = (-R0 e3h Ibi)[:, :2]
  double s0[3], ds0[3], Bt[6], y0[6], dy0[6], y1[6], yd[6], dyd[6];
  for (int i = 0; i < 3; ++i) {
    s0[i] = R0[6 + i];
    ds0[i] = -(R0[i] * (-dq0[4]) + R0[3 + i] * dq0[3]);
    Bt[i] = -(R0[3 + i] * p.Ibi[0]);   // column 0
    Bt[3 + i] = R0[i] * p.Ibi[1];      // column 1
    y0[i] = p0[i]; y0[3 + i] = s0[i]; dy0[i] = dq0[i]; dy0[3 + i] = ds0[i];
    yd[i] = pdes[i]; yd[3 + i] = sdes[i]; dyd[i] = dpdes[i]; dyd[3 + i] = 0.0;
  }
  for (int i = 0; i < 6; ++i) y1[i] = y0[i] + dt * dy0[i];
  // ---- l, u (updateConstraint) ------------------------------------------------------------------------
  double* le = l + (size_t)e * nc; double* ue = u + (size_t)e * nc;
  for (int r = 0; r < nc; ++r) le[r] = 0.0;
  for (int i = 0; i < 6; ++i) le[i] = -y1[i];
  for (int k = 0; k < N; ++k)
    for (int i = 0; i < 6; ++i) {
      const double c0 = i == 2 ? -p.g : 0.0;
      double v = -dt * c0;
      if (k == 0) v += -dy0[i] - (i < 3 ? dt * T0 * y0[3 + i] : 0.0);
      else if (k == 1) v += -(i < 3 ? dt * T0 * y1[3 + i] : 0.0);
      le[n1 + k * NY + i] = v;
    }
  for (int r = 0; r < n2; ++r) ue[r] = le[r];
  for (int k = 0; k < N; ++k) { le[n2 + k] = -T0; ue[n2 + k] = p.Tmax - T0; }
  double u0[4] = {0, 0, 0, 0};
  if (p.wl) {
    for (int j = 0; j < 4; ++j) {
      u0[j] = io.u0[(size_t)e * 4 + j];
      double lo = -p.dumax[j], hi = p.dumax[j];
      if (u0[j] < p.umin[j]) lo = 0.0; else if (u0[j] > p.umax[j]) hi = 0.0;
      le[n2 + N + j] = lo; ue[n2 + N + j] = hi;
    }
  }
  // ---- A values in CSR order --------------------------------------------------------------------------
  double* ae = Av + (size_t)e * p.nnz;
  for (int r = 0; r < nc; ++r) {
    Ent en[6];
    const int cnt = row_entries(N, p.wl, r, en);
    for (int a = 0; a < cnt; ++a) {
      double v;
      switch (en[a].kind) {
        case K_P1: v = 1.0; break;
        case K_M1: v = -1.0; break;
        case K_DT: v = dt; break;
        case K_DTT0: v = dt * T0; break;
        case K_S: v = dt * s0[en[a].idx]; break;
        default: v = dt * Bt[en[a].idx]; break;
      }
      ae[rowptr[r] + a] = v;
    }
  }
  // ---- P (dense, symmetric), q (updateObjective) --------------------------------------------------------
  double* Pe = Pq + (size_t)e * nx * nx; double* qe = q + (size_t)e * nx;
  for (int i = 0; i < nx * nx; ++i) Pe[i] = 0.0;
  for (int k = 0; k < N; ++k)
    for (int i = 0; i < 6; ++i) {
      const double wy = k < N - 1 ? p.Qyr[i] : p.Qyf[i], wd = k < N - 1 ? p.Qdyr[i] : p.Qdyf[i];
      Pe[(size_t)(k * NY + i) * nx + k * NY + i] = wy; qe[k * NY + i] = -wy * yd[i];
      Pe[(size_t)(n1 + k * NY + i) * nx + n1 + k * NY + i] = wd; qe[n1 + k * NY + i] = -wd * dyd[i];
    }
  for (int k = 0; k < N; ++k) for (int i = 0; i < 3; ++i) { Pe[(size_t)(n2 + k * NU + i) * nx + n2 + k * NU + i] = p.R[i]; qe[n2 + k * NU + i] = 0.0; }
  if (p.wl) {
    // h0 = [R0'(0,0,mb g); 0], Tm = blkdiag(I3, e3h R0'), M0t = diag(M0) Tm / dt, w0t = w0 - h0 + M0 .* dq0 / dt
    double w0[6], A1[6][4], M0t[6][6], w0t[6];
    wrench(p, u0, w0, A1);
    for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) M0t[i][j] = 0.0;
    for (int i = 0; i < 3; ++i) M0t[i][i] = p.M0[i] / dt;
    // e3h R0' : rows (-R0'[1,:], R0'[0,:], 0) ; R0'[a][b] = R0[b][a] = R0[b + 3 a] (column-major R0)
    for (int j = 0; j < 3; ++j) {
      M0t[3][3 + j] = p.M0[3] * (-R0[j + 3 * 1]) / dt;
      M0t[4][3 + j] = p.M0[4] * (R0[j + 3 * 0]) / dt;
    }
    for (int i = 0; i < 3; ++i) {
      const double h0 = R0[3 * i + 2] * p.M0[0] * p.g;
      w0t[i] = w0[i] - h0 + p.M0[i] * dq0[i] / dt;
      w0t[3 + i] = w0[3 + i] + p.M0[3 + i] * dq0[3 + i] / dt;
    }
    // blocks: (dy1,dy1) += M0t' Qw M0t ; (dy1,du) = -M0t' Qw A1 ; (du,du) = A1' Qw A1 ; q[dy1] += -M0t' Qw w0t ; q[du] = A1' Qw w0t
    for (int a = 0; a < 6; ++a) {
      for (int b = 0; b < 6; ++b) {
        double v = 0;
        for (int i = 0; i < 6; ++i) v += M0t[i][a] * p.Qw[i] * M0t[i][b];
        Pe[(size_t)(n1 + a) * nx + n1 + b] += v;
      }
      for (int j = 0; j < 4; ++j) {
        double v = 0;
        for (int i = 0; i < 6; ++i) v += M0t[i][a] * p.Qw[i] * A1[i][j];
        Pe[(size_t)(n1 + a) * nx + n3 + j] = -v; Pe[(size_t)(n3 + j) * nx + n1 + a] = -v;
      }
      double v = 0;
      for (int i = 0; i < 6; ++i) v += M0t[i][a] * p.Qw[i] * w0t[i];
      qe[n1 + a] += -v;
    }
    for (int a = 0; a < 4; ++a) {
      for (int b = 0; b < 4; ++b) {
        double v = 0;
        for (int i = 0; i < 6; ++i) v += A1[i][a] * p.Qw[i] * A1[i][b];
        Pe[(size_t)(n3 + a) * nx + n3 + b] = v;
      }
      double v = 0;
      for (int i = 0; i < 6; ++i) v += A1[i][a] * p.Qw[i] * w0t[i];
      qe[n3 + a] = v;
    }
  }
}

// update2 / getAccDes (template_controllers.py:239-258, uprightmpc2_withWLQP.py:301-320)
__global__ void extract_kernel(const __grid_constant__ Par p, Io io, int nenv, const double* __restrict__ x) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= nenv) return;
  const int N = p.N, n1 = N * NY, n2 = 2 * N * NY, n3 = n2 + N * NU;
  const double* xe = x + (size_t)e * p.nx;
  const double* in = io.in + (size_t)e * 28;
  const double *R0 = in + 3, *dq0 = in + 12;
  double* out = io.out + (size_t)e * 13;
  const double T0 = io.T0[e] + xe[n2];
  io.T0[e] = T0;
  out[0] = T0; out[1] = xe[n2 + 1]; out[2] = xe[n2 + 2];
  const double* dy1 = xe + n1;
  const double t0 = R0[0] * dy1[3] + R0[1] * dy1[4] + R0[2] * dy1[5];
  const double t1 = R0[3] * dy1[3] + R0[4] * dy1[4] + R0[5] * dy1[5];
  const double dq1[6] = {dy1[0], dy1[1], dy1[2], -t1, t0, 0.0};
  for (int i = 0; i < 6; ++i) out[3 + i] = (dq1[i] - dq0[i]) / p.dt;
  for (int j = 0; j < 4; ++j) {
    double v = 0.0;
    if (p.wl) { v = io.u0[(size_t)e * 4 + j] + xe[n3 + j]; io.u0[(size_t)e * 4 + j] = v; }
    out[9 + j] = v;
  }
}

}  // namespace umg

#define UCK(call)                                                                         \
  do {                                                                                    \
    cudaError_t e_ = (call);                                                              \
    if (e_ != cudaSuccess) {                                                              \
      umpc_internal_set_error((std::string(#call) + ": " + cudaGetErrorString(e_)).c_str()); \
      return UMPC_ERR_CUDA;                                                               \
    }                                                                                     \
  } while (0)

struct UmgBatch {
  int nenv = 0, device = 0;
  umg::Par par;
  umg::Io io{nullptr, nullptr, nullptr, nullptr};
  GqpBatch* qp = nullptr;
  int* d_rowptr = nullptr;
  cudaStream_t stream = nullptr;  // the inner batch's stream
};

extern "C" void umgSizes(int N, int withWL, int* nx, int* nc, int* nnz) {
  const int n = N * (2 * umg::NY + umg::NU) + (withWL ? 4 : 0), m = 2 * N * umg::NY + N + (withWL ? 4 : 0);
  int z = 0;
  umg::Ent e[6];
  for (int r = 0; r < m; ++r) z += umg::row_entries(N, withWL, r, e);
  if (nx) *nx = n;
  if (nc) *nc = m;
  if (nnz) *nnz = z;
}

extern "C" void umgPattern(int N, int withWL, int* rowptr, int* colidx) {
  int m, z = 0;
  umgSizes(N, withWL, nullptr, &m, nullptr);
  umg::Ent e[6];
  for (int r = 0; r < m; ++r) {
    rowptr[r] = z;
    const int c = umg::row_entries(N, withWL, r, e);
    for (int a = 0; a < c; ++a) colidx[z++] = e[a].col;
  }
  rowptr[m] = z;
}

extern "C" void umgDefaults(UmgParams* p, int N, int withWL) {
  memset(p, 0, sizeof(*p));
  p->N = N; p->withWL = withWL;
  p->dt = 5.0; p->g = 9.81e-3; p->TtoWmax = 2.0; p->ws = 1e1; p->wds = 1e3; p->wpr = 1.0; p->wpf = 5.0; p->wvr = 1e3; p->wvf = 2e3;
  p->wthrust = 1e-1; p->wmom = 1e-2; p->Ib[0] = 3333.0; p->Ib[1] = 3333.0; p->Ib[2] = 1000.0;
  p->maxIter = 4000; p->checkTermination = 25; p->adaptiveRhoInterval = 100;
  p->epsAbs = p->epsRel = withWL ? 1e-7 : 1e-4;   // uprightmpc2_withWLQP.py:243 / template_controllers.py:191
  const double umin[4] = {90.0, -0.5, -0.2, -0.1}, umax[4] = {240.0, 0.5, 0.2, 0.1}, dumax[4] = {5e3, 10.0, 10.0, 10.0};
  const double Qw[6] = {1.0, 1.0, 1.0, 0.1, 0.1, 0.1};
  memcpy(p->umin, umin, sizeof umin); memcpy(p->umax, umax, sizeof umax); memcpy(p->dumax, dumax, sizeof dumax); memcpy(p->Qw, Qw, sizeof Qw);
  p->controlRate = 1000.0; p->mb = 100.0; p->wrenchModel = 0;
}

extern "C" void umgDestroy(UmgBatch* b) {
  if (!b) return;
  cudaSetDevice(b->device);
  gqpBatchDestroy(b->qp);
  cudaFree(b->io.in); cudaFree(b->io.T0); cudaFree(b->io.u0); cudaFree(b->io.out); cudaFree(b->d_rowptr);
  cudaGetLastError();
  delete b;
}

extern "C" int umgReset(UmgBatch* b) {
  if (!b) return UMPC_ERR_ARG;
  UCK(cudaSetDevice(b->device));
  UCK(cudaMemsetAsync(b->io.T0, 0, (size_t)b->nenv * 8, b->stream));
  UCK(cudaMemsetAsync(b->io.u0, 0, (size_t)b->nenv * 4 * 8, b->stream));
  UCK(cudaStreamSynchronize(b->stream));
  return gqpBatchReset(b->qp);
}

extern "C" int umgSetState(UmgBatch* b, const double* T0, const double* u0) {
  if (!b) return UMPC_ERR_ARG;
  UCK(cudaSetDevice(b->device));
  if (T0) UCK(cudaMemcpyAsync(b->io.T0, T0, (size_t)b->nenv * 8, cudaMemcpyHostToDevice, b->stream));
  if (u0) UCK(cudaMemcpyAsync(b->io.u0, u0, (size_t)b->nenv * 4 * 8, cudaMemcpyHostToDevice, b->stream));
  UCK(cudaStreamSynchronize(b->stream));
  return UMPC_OK;
}

static int umg_create_body(UmgBatch* b, const UmgParams* up) {
  umg::Par& p = b->par;
  memset(&p, 0, sizeof(p));
  p.N = up->N; p.wl = up->withWL ? 1 : 0; p.wrench = up->wrenchModel;
  umgSizes(p.N, p.wl, &p.nx, &p.nc, &p.nnz);
  p.dt = up->dt; p.g = up->g; p.Tmax = up->TtoWmax * up->g;
  for (int i = 0; i < 3; ++i) {
    p.Qyr[i] = up->wpr; p.Qyf[i] = up->wpf; p.Qyr[3 + i] = p.Qyf[3 + i] = up->ws;
    p.Qdyr[i] = up->wvr; p.Qdyf[i] = up->wvf; p.Qdyr[3 + i] = p.Qdyf[3 + i] = up->wds;
    p.Ibi[i] = 1.0 / up->Ib[i]; p.M0[i] = up->mb; p.M0[3 + i] = up->Ib[i];
  }
  p.R[0] = up->wthrust; p.R[1] = p.R[2] = up->wmom;
  for (int j = 0; j < 4; ++j) { p.umin[j] = up->umin[j]; p.umax[j] = up->umax[j]; p.dumax[j] = up->dumax[j] / up->controlRate; }
  for (int i = 0; i < 6; ++i) p.Qw[i] = up->Qw[i];
  memcpy(p.popts, up->popts, sizeof(p.popts));
  std::vector<int> rowptr(p.nc + 1), colidx(p.nnz);
  umgPattern(p.N, p.wl, rowptr.data(), colidx.data());
  GqpSettings s;
  gqpDefaults(&s, p.nx, p.nc, p.nnz);
  s.max_iter = up->maxIter; s.check_termination = up->checkTermination; s.adaptive_rho_interval = up->adaptiveRhoInterval;
  s.eps_abs = up->epsAbs; s.eps_rel = up->epsRel;
  int rc = gqpBatchCreate(&b->qp, b->nenv, b->device, &s, rowptr.data(), colidx.data());
  if (rc) return rc;
  b->stream = (cudaStream_t)gqpBatchDevPtr(b->qp, "stream");
  const size_t B = b->nenv;
  UCK(cudaMalloc(&b->io.in, B * 28 * 8)); UCK(cudaMalloc(&b->io.T0, B * 8)); UCK(cudaMalloc(&b->io.u0, B * 4 * 8));
  UCK(cudaMalloc(&b->io.out, B * 13 * 8)); UCK(cudaMalloc(&b->d_rowptr, (p.nc + 1) * 4));
  UCK(cudaMemcpy(b->d_rowptr, rowptr.data(), (p.nc + 1) * 4, cudaMemcpyHostToDevice));
  return umgReset(b);
}

extern "C" int umgCreate(UmgBatch** out, int nenv, int device, const UmgParams* par) {
  if (!out || nenv <= 0 || !par || par->N < 2 || par->N > 16) { umpc_internal_set_error("umgCreate: bad argument (2 <= N <= 16)"); return UMPC_ERR_ARG; }
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    umpc_internal_set_error((std::string("umgCreate: no CUDA device (") + cudaGetErrorString(e) + "); this library has no CPU path").c_str());
    return UMPC_ERR_CUDA;
  }
  UCK(cudaSetDevice(device));
  UmgBatch* b = new UmgBatch();
  b->nenv = nenv; b->device = device;
  const int rc = umg_create_body(b, par);
  if (rc) { umgDestroy(b); return rc; }
  *out = b;
  return UMPC_OK;
}

extern "C" int umgUpdate(UmgBatch* b, double* uquad, double* accdes, double* u_wl, const double* p0, const double* R0, const double* dq0,
                         const double* pdes, const double* dpdes, const double* sdes, const double* actualT0) {
  if (!b || !uquad || !accdes || !p0 || !R0 || !dq0 || !pdes || !dpdes || !sdes) { umpc_internal_set_error("umgUpdate: NULL argument"); return UMPC_ERR_ARG; }
  UCK(cudaSetDevice(b->device));
  const size_t B = b->nenv;
  std::vector<double> in(B * 28), out(B * 13);
  for (size_t e = 0; e < B; ++e) {
    double* d = &in[e * 28];
    memcpy(d, p0 + e * 3, 24); memcpy(d + 3, R0 + e * 9, 72); memcpy(d + 12, dq0 + e * 6, 48); memcpy(d + 18, pdes + e * 3, 24);
    memcpy(d + 21, dpdes + e * 3, 24); memcpy(d + 24, sdes + e * 3, 24);
    d[27] = actualT0 ? actualT0[e] : -1.0;
  }
  UCK(cudaMemcpyAsync(b->io.in, in.data(), B * 28 * 8, cudaMemcpyHostToDevice, b->stream));
  const unsigned grid = (unsigned)((B + 63) / 64);
  umg::build_kernel<<<grid, 64, 0, b->stream>>>(b->par, b->io, b->nenv, (double*)gqpBatchDevPtr(b->qp, "P"), (double*)gqpBatchDevPtr(b->qp, "Av"),
                                                (double*)gqpBatchDevPtr(b->qp, "q"), (double*)gqpBatchDevPtr(b->qp, "l"),
                                                (double*)gqpBatchDevPtr(b->qp, "u"), b->d_rowptr);
  UCK(cudaGetLastError());
  int rc = gqpBatchSolve(b->qp, 1);   // model.update(Px, Ax, q, l, u): data changed -> re-equilibrate
  if (rc) return rc;
  umg::extract_kernel<<<grid, 64, 0, b->stream>>>(b->par, b->io, b->nenv, (const double*)gqpBatchDevPtr(b->qp, "x"));
  UCK(cudaGetLastError());
  UCK(cudaMemcpyAsync(out.data(), b->io.out, B * 13 * 8, cudaMemcpyDeviceToHost, b->stream));
  UCK(cudaStreamSynchronize(b->stream));
  for (size_t e = 0; e < B; ++e) {
    memcpy(uquad + e * 3, &out[e * 13], 24); memcpy(accdes + e * 6, &out[e * 13 + 3], 48);
    if (u_wl) memcpy(u_wl + e * 4, &out[e * 13 + 9], 32);
  }
  return UMPC_OK;
}

extern "C" int umgGetQP(UmgBatch* b, double* P, double* Av, double* q, double* l, double* u, double* x, double* y, double* info) {
  if (!b) return UMPC_ERR_ARG;
  UCK(cudaSetDevice(b->device));
  const size_t B = b->nenv;
  const umg::Par& p = b->par;
  struct { double* h; const char* name; size_t per; } arr[5] = {{P, "P", (size_t)p.nx * p.nx}, {Av, "Av", (size_t)p.nnz}, {q, "q", (size_t)p.nx},
                                                                {l, "l", (size_t)p.nc}, {u, "u", (size_t)p.nc}};
  for (auto& a : arr)
    if (a.h) UCK(cudaMemcpyAsync(a.h, gqpBatchDevPtr(b->qp, a.name), B * a.per * 8, cudaMemcpyDeviceToHost, b->stream));
  UCK(cudaStreamSynchronize(b->stream));
  return gqpBatchGetSolution(b->qp, x, y, info);
}

extern "C" float umgLastSolveMs(UmgBatch* b) { return b ? gqpBatchLastKernelMs(b->qp) : -1.0f; }
