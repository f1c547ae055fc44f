// robobee3d_b200/csrc/flapping.cu -- flapping-wing plant, open-loop sweeps (include/flapping_b200.h).
//
// Reference path:
//   aerodynamics / wrench / actuatorModel    template/robobee.py:20-55
//   RobobeeSim.update, openLoop              template/robobee.py:213-300   (torques, aero forces, drive waveform)
//   sweepFile / olAvgWrenchKinFeat           template/sdab_num_wrenchmap_simgen.py:19-55
//   the robot                                urdf/sdab.xacro
// ONE THREAD PER WING (lanes 2k, 2k+1 = left, right wing of simulation k).  On a fixed base the wings do not couple:
// each is a two-joint chain whose joint axes meet in one point O, so both links turn about O and the equations of motion
// are 2 x 2.  Written in LINK frames: w_k = S_k dq, N_k = J_k (S_k ddq + dS_k dq) + w_k x J_k w_k, tau = sum_k S_k' N_k with
// J_k the link's inertia about O in its own frame (constant).  The CPU oracle states the same physics in the world
// frame (oracle/flapping_oracle.py).  Semi-implicit Euler like Bullet's multibody step.  Everything in registers; the
// only global traffic is the optional log of one simulation and the 15 features per simulation at the end.
#include <cuda_runtime.h>
#include <math.h>
#include <string.h>

#include <string>

#include "../../include/umpc_b200.h"
#include "../../include/flapping_b200.h"

void umpc_internal_set_error(const char* msg);

namespace fw {

struct V3 { double x, y, z; };
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ V3 operator*(double s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }
__device__ __forceinline__ double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ V3 cross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
struct Sym3 { double xx, yy, zz, xy, xz, yz; };
__device__ __forceinline__ V3 mul(const Sym3& J, V3 v) {
  return {J.xx * v.x + J.xy * v.y + J.xz * v.z, J.xy * v.x + J.yy * v.y + J.yz * v.z, J.xz * v.x + J.yz * v.y + J.zz * v.z};
}
// link-frame inertia about the joint point: the URDF's inertial frame is the link frame turned by +-90 deg about z
// (urdf/sdab.xacro:77,93: ixx <-> iyy), plus the parallel-axis term of the COM offset c
__device__ __forceinline__ Sym3 inertia_about_O(const double I[3], double m, V3 c) {
  const double cc = dot(c, c);
  return {I[1] + m * (cc - c.x * c.x), I[0] + m * (cc - c.y * c.y), I[2] + m * (cc - c.z * c.z), -m * c.x * c.y, -m * c.x * c.z, -m * c.y * c.z};
}

// robobee.py:20-47; th/dth as the reference's caller passes them (update() hands the right wing -q, -dq: :229-230)
__device__ __forceinline__ void aerodynamics(const FwParams& p, double th0, double th1, double dth0, int sgn, V3& F, V3& pcop) {
  th0 *= sgn; dth0 *= sgn;
  const double AR = p.Aw / (p.cbar * p.cbar);
  double s, c, sh, ch;
  sincos(th0, &s, &c); sincos(th1, &sh, &ch);
  const double cr = p.cbar * p.rcopnondim;
  pcop = {-p.ycp * s - cr * c * sh, p.Roffs + p.ycp * c - cr * s * sh, p.d - cr * ch};
  const double aoa = 0.5 * M_PI - th1;
  const double den = 1.0 / (fabs(dth0) + 1e-4);
  const double Cd = (p.CDmax + p.CD0) / 2 - (p.CDmax - p.CD0) / 2 * cos(2 * aoa), Cl = p.CLmax * sin(2 * aoa);
  const double k = 0.5 * p.RHO * dth0 * dth0 * p.Aw * p.Aw * AR * p.r2h * p.r2h;
  const double sg = dth0 < 0 ? 1.0 : (dth0 > 0 ? -1.0 : 0.0);   // np.sign(-dtheta[0])
  F = {k * Cd * (dth0 * c * den), k * Cd * (dth0 * s * den), k * Cl * sg};
  if (sgn < 0) { F.y = -F.y; pcop.y = -pcop.y; }
}

struct Wing {   // constants of one wing
  int sgn;
  V3 O, c1, c2;
  Sym3 J1, J2;
  double m1g, m2g;   // weight along z
};

__device__ __forceinline__ Wing make_wing(const FwParams& p, int sgn) {
  Wing w;
  w.sgn = sgn;
  w.O = {0.0, sgn * p.Roffs, p.d};
  w.c1 = {p.c_spar[0], sgn * p.c_spar[1], p.c_spar[2]};
  w.c2 = {p.c_mem[0], sgn * p.c_mem[1], p.c_mem[2]};
  w.J1 = inertia_about_O(p.I_spar, p.m_spar, w.c1);
  w.J2 = inertia_about_O(p.I_mem, p.m_mem, w.c2);
  w.m1g = p.m_spar * p.g; w.m2g = p.m_mem * p.g;
  return w;
}

// ddq of one wing: q = (stroke, hinge), joint torques tau, aero force F at pcop (body frame)
__device__ __forceinline__ void wing_accel(const Wing& w, double q1, double q2, double dq1, double dq2, double tau1, double tau2, V3 F,
                                           V3 pcop, double& ddq1, double& ddq2) {
  const double sg = (double)w.sgn;
  double s1, c1, s2, c2;
  sincos(sg * q1, &s1, &c1);   // stroke: rotation about z by sgn q1
  sincos(sg * q2, &s2, &c2);   // hinge: rotation about y by sgn q2 (spar <- membrane)
  // membrane frame: the stroke axis a1 = (0, 0, sgn) seen from it, u = Ry(sgn q2)' a1;  a2 = (0, sgn, 0)
  const V3 u = {-s2 * sg, 0.0, c2 * sg}, a2 = {0.0, sg, 0.0};
  const V3 w2 = dq1 * u + dq2 * a2;
  const V3 beta = (dq1 * dq2) * cross(u, a2);            // dS_2 dq
  const V3 J2u = mul(w.J2, u), J2a = mul(w.J2, a2);
  const double M11 = w.J1.zz + dot(u, J2u), M12 = dot(u, J2a), M22 = dot(a2, J2a);
  const V3 n2 = mul(w.J2, beta) + cross(w2, mul(w.J2, w2));
  // (the spar's own gyroscopic term vanishes: J1 is diagonal and w1 is along z)
  const double h1 = dot(u, n2), h2 = dot(a2, n2);
  // applied torques about O, world = body frame (fixed base, identity attitude): R1 = Rz(sgn q1), R2 = R1 Ry(sgn q2)
  auto R1 = [&](V3 v) { return V3{c1 * v.x - s1 * v.y, s1 * v.x + c1 * v.y, v.z}; };
  auto Ry = [&](V3 v) { return V3{c2 * v.x + s2 * v.z, v.y, -s2 * v.x + c2 * v.z}; };
  const V3 r1 = R1(w.c1), r2 = R1(Ry(w.c2));
  const V3 T1 = cross(r1, V3{0, 0, w.m1g});
  const V3 T2 = cross(r2, V3{0, 0, w.m2g}) + cross(pcop - w.O, F);
  const V3 ax2 = R1(a2);
  const double Q1 = tau1 + sg * (T1.z + T2.z) - h1, Q2 = tau2 + dot(ax2, T2) - h2;
  const double det = 1.0 / (M11 * M22 - M12 * M12);
  ddq1 = det * (M22 * Q1 - M12 * Q2);
  ddq2 = det * (M11 * Q2 - M12 * Q1);
}

__global__ void aero_kernel(const __grid_constant__ FwParams p, int n, const double* __restrict__ theta, const double* __restrict__ dtheta,
                            int sgn, double* __restrict__ F, double* __restrict__ pcop) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  V3 f, pc;
  aerodynamics(p, theta[2 * i], theta[2 * i + 1], dtheta[2 * i], sgn, f, pc);
  F[3 * i] = f.x; F[3 * i + 1] = f.y; F[3 * i + 2] = f.z;
  pcop[3 * i] = pc.x; pcop[3 * i + 1] = pc.y; pcop[3 * i + 2] = pc.z;
}

// RobobeeSim.openLoop (robobee.py:277-300) + olAvgWrenchKinFeat (sdab_num_wrenchmap_simgen.py:31-47)
__global__ void __launch_bounds__(128) sweep_kernel(const __grid_constant__ FwParams p, int nsim, const double* __restrict__ drive,
                                                    int nsteps, double dt, double* __restrict__ feat, int log_sim, double* __restrict__ qw) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int sim = min(t >> 1, nsim - 1), wi = t & 1, sgn = wi ? -1 : 1;
  const bool live = (t >> 1) < nsim;
  const unsigned pair = 3u << (threadIdx.x & 30);
  const Wing w = make_wing(p, sgn);
  const double* dr = drive + (size_t)sim * 5;
  const double Vamp = dr[wi], uoffs = dr[2], f = dr[3], h2 = wi ? -dr[4] : dr[4];   // tau = [VampL sig(ph, h2), VampR sig(ph, -h2)]
  const double omega = 2.0 * M_PI * f;
  const int npts = (int)(1.0 / (f * dt));
  const int win0 = nsteps - npts;
  double q1 = 0, q2 = 0, dq1 = 0, dq2 = 0, simt = 0;
  double ws[6] = {0, 0, 0, 0, 0, 0};
  double mx1 = -INFINITY, mx2 = -INFINITY, mn1 = INFINITY, mn2 = INFINITY;
  int npos = 0;
  const bool logging = live && sim == log_sim && qw != nullptr;
  for (int k = 0; k < nsteps; ++k) {
    const double ph = omega * simt;
    const double V = Vamp * ((1.0 + p.h3) * sin(ph) + p.h3 * sin(3.0 * ph) + h2 * sin(2.0 * ph) + uoffs);
    V3 F, pc;
    if (sgn > 0) aerodynamics(p, q1, q2, dq1, 1, F, pc);
    else aerodynamics(p, -q1, -q2, -dq1, -1, F, pc);
    const V3 mo = cross(pc, F);   // wrench(F, p) = (F, p x F), robobee.py:49-50
    if (k >= win0) {
      ws[0] += F.x; ws[1] += F.y; ws[2] += F.z; ws[3] += mo.x; ws[4] += mo.y; ws[5] += mo.z;
      mx1 = fmax(mx1, q1); mx2 = fmax(mx2, q2); mn1 = fmin(mn1, q1); mn2 = fmin(mn2, q2);
      npos += dq1 >= 0.0;
    }
    if (logging) {
      double* row = qw + (size_t)k * 14;
      row[2 * wi] = q1; row[2 * wi + 1] = q2; row[4 + 2 * wi] = dq1; row[5 + 2 * wi] = dq2;
      const double own[6] = {F.x, F.y, F.z, mo.x, mo.y, mo.z};
      for (int i = 0; i < 6; ++i) {
        const double other = __shfl_xor_sync(pair, own[i], 1);
        if (wi == 0) row[8 + i] = own[i] + other;
      }
    }
    // joint torques: stroke = actuator force - stroke stiffness (robobee.py:219-222), hinge = PD towards 0 (:141)
    const double tau1 = (1.0 / p.act_T) * (50.0 / 180.0 * V) - p.kstroke * q1;
    const double tau2 = -p.khinge * q2 - p.bhinge * dq2;
    double a1, a2;
    wing_accel(w, q1, q2, dq1, dq2, tau1, tau2, F, pc, a1, a2);
    dq1 += dt * a1; dq2 += dt * a2;
    q1 += dt * dq1; q2 += dt * dq2;
    simt += dt;
  }
  // features: left wing's lane gathers the right wing's
  const double inv = 1.0 / npts;
  double o[4] = {mx1, mx2, -mn1, -mn2};
  double po[4], wsum[6];
  for (int i = 0; i < 4; ++i) po[i] = __shfl_xor_sync(0xffffffffu, o[i], 1);
  for (int i = 0; i < 6; ++i) wsum[i] = ws[i] + __shfl_xor_sync(0xffffffffu, ws[i], 1);
  if (live && wi == 0) {
    double* r = feat + (size_t)sim * 15;
    for (int i = 0; i < 6; ++i) r[i] = wsum[i] * inv;
    for (int i = 0; i < 4; ++i) { r[6 + i] = o[i]; r[10 + i] = po[i]; }
    r[14] = npos * inv;
  }
}

}  // namespace fw

#define FCK(call)                                                                         \
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

static int fw_need_device(const char* who) {
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    umpc_internal_set_error((std::string(who) + ": no CUDA device (" + cudaGetErrorString(e) + "); this library has no CPU path").c_str());
    return UMPC_ERR_CUDA;
  }
  return UMPC_OK;
}

extern "C" void fwDefaults(FwParams* p) {
  memset(p, 0, sizeof(*p));
  p->CD0 = 0.4; p->CDmax = 3.4; p->CLmax = 1.8; p->RHO = 1.225e-3; p->rcopnondim = 0.5;   // robobee.py:7-16
  p->Aw = 54.4; p->r2h = 0.551; p->cbar = 3.2;                                            // wparams, robobee.py:18
  p->ycp = 7.0; p->Roffs = 3.0; p->d = 4.0;                                               // urdf/sdab.xacro: Rwing / 2, RwingOffset, d
  p->khinge = 3.0; p->bhinge = 2.0; p->kstroke = 65.0;                                    // urdf/sdab.xacro:36-38
  p->m_spar = 0.72; p->I_spar[0] = 0.03; p->I_spar[1] = 11.775; p->I_spar[2] = 11.775;     // :27-30
  p->c_spar[0] = 0.0; p->c_spar[1] = 7.0; p->c_spar[2] = 0.0;                             // :77
  p->m_mem = 0.06; p->I_mem[0] = 0.08; p->I_mem[1] = 1.06; p->I_mem[2] = 0.981;            // :31-34
  p->c_mem[0] = 0.0; p->c_mem[1] = 7.0; p->c_mem[2] = -2.0;                               // :93 (-cbar / 2 with the xacro's cbar = 4)
  p->g = -9.81e-3; p->act_T = 2.6666; p->h3 = 0.0;                                        // robobee.py:98, :54
}

extern "C" int fwOpenLoopSteps(double tendMS, double timestep) {
  if (!(timestep > 0.0) || !(tendMS > 0.0)) return 0;
  int n = 0;
  for (double simt = 0.0; simt < tendMS; simt += timestep) ++n;   // robobee.py:287 with :254
  return n;
}

extern "C" int fwAerodynamics(const FwParams* p, int n, const double* theta, const double* dtheta, int lrSign, double* F, double* pcop) {
  if (!p || n <= 0 || !theta || !dtheta || !F || !pcop || (lrSign != 1 && lrSign != -1)) {
    umpc_internal_set_error("fwAerodynamics: bad argument");
    return UMPC_ERR_ARG;
  }
  if (int rc = fw_need_device("fwAerodynamics")) return rc;
  DevBuf bth, bdth, bF, bP;
  FCK(bth.alloc((size_t)n * 16)); FCK(bdth.alloc((size_t)n * 16)); FCK(bF.alloc((size_t)n * 24)); FCK(bP.alloc((size_t)n * 24));
  double *dth = bth.as<double>(), *ddth = bdth.as<double>(), *dF = bF.as<double>(), *dP = bP.as<double>();
  FCK(cudaMemcpy(dth, theta, (size_t)n * 16, cudaMemcpyHostToDevice)); FCK(cudaMemcpy(ddth, dtheta, (size_t)n * 16, cudaMemcpyHostToDevice));
  fw::aero_kernel<<<(n + 127) / 128, 128>>>(*p, n, dth, ddth, lrSign, dF, dP);
  FCK(cudaGetLastError());
  FCK(cudaMemcpy(F, dF, (size_t)n * 24, cudaMemcpyDeviceToHost)); FCK(cudaMemcpy(pcop, dP, (size_t)n * 24, cudaMemcpyDeviceToHost));
  return UMPC_OK;
}

extern "C" int fwOpenLoopSweep(const FwParams* p, int nsim, const double* drive, double tendMS, double timestep, double* feat, int log_sim,
                               double* qw, float* kernel_ms) {
  if (!p || nsim <= 0 || !drive || !feat) { umpc_internal_set_error("fwOpenLoopSweep: bad argument"); return UMPC_ERR_ARG; }
  const int nsteps = fwOpenLoopSteps(tendMS, timestep);
  if (nsteps <= 0) { umpc_internal_set_error("fwOpenLoopSweep: tendMS and timestep must be positive"); return UMPC_ERR_ARG; }
  for (int i = 0; i < nsim; ++i) {
    const double f = drive[(size_t)i * 5 + 3];
    if (!(f > 0.0) || (int)(1.0 / (f * timestep)) < 1 || (int)(1.0 / (f * timestep)) > nsteps) {
      umpc_internal_set_error("fwOpenLoopSweep: every simulation must cover at least one flapping period (f > 0, 1 / f <= tendMS)");
      return UMPC_ERR_ARG;
    }
  }
  if (qw && (log_sim < 0 || log_sim >= nsim)) { umpc_internal_set_error("fwOpenLoopSweep: log_sim out of range"); return UMPC_ERR_ARG; }
  if (int rc = fw_need_device("fwOpenLoopSweep")) return rc;
  DevBuf bd, bf, bq;
  FCK(bd.alloc((size_t)nsim * 5 * 8)); FCK(bf.alloc((size_t)nsim * 15 * 8));
  if (qw) FCK(bq.alloc((size_t)nsteps * 14 * 8));
  double *dd = bd.as<double>(), *df = bf.as<double>(), *dq = bq.as<double>();
  FCK(cudaMemcpy(dd, drive, (size_t)nsim * 5 * 8, cudaMemcpyHostToDevice));
  EvPair ev;
  FCK(cudaEventCreate(&ev.e0)); FCK(cudaEventCreate(&ev.e1));
  FCK(cudaEventRecord(ev.e0));
  fw::sweep_kernel<<<(2 * nsim + 127) / 128, 128>>>(*p, nsim, dd, nsteps, timestep, df, qw ? log_sim : -1, dq);
  FCK(cudaEventRecord(ev.e1));
  FCK(cudaGetLastError());
  FCK(cudaMemcpy(feat, df, (size_t)nsim * 15 * 8, cudaMemcpyDeviceToHost));
  if (qw) FCK(cudaMemcpy(qw, dq, (size_t)nsteps * 14 * 8, cudaMemcpyDeviceToHost));
  if (kernel_ms) FCK(cudaEventElapsedTime(kernel_ms, ev.e0, ev.e1));
  return UMPC_OK;
}
