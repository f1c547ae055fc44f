// tests/hostcheck/hostcheck.cpp -- TEST/DEVELOPMENT AID ONLY, never part of the product library.
// Compiles the device algorithm headers (robobee3d_b200/csrc/*.cuh are __host__ __device__)
// with g++ so the formulation can be compared with the oracle on this GPU-less box before
// GPU time is spent.  The product (libumpc_b200.so) has no CPU path and does not contain this.
#include <string.h>
#include "../../robobee3d_b200/csrc/umpc_core.cuh"

struct HostStore {
  float wds[UMPC_GEN_NWORDS], vf[UMPC_NX];
  void issue(int c, float (&buf)[UMPC_GEN_CHUNK]) const { for (int i = 0; i < UMPC_GEN_CHUNK; ++i) buf[i] = wds[c * UMPC_GEN_CHUNK + i]; }
  void wait(float (&)[UMPC_GEN_CHUNK]) const {}
  void setW(int e, float v) { wds[e] = v; }
  void end_writes() {}
  void sync() const {}
  float Vf(int i) const { return vf[i]; }
  void setVf(int i, float v) { vf[i] = v; }
  void begin_iteration(int) {}
};

struct HostIO {
  float xs[UMPC_NX], zs[UMPC_NC], ys[UMPC_NC], D[UMPC_NX], E[UMPC_NC], solx[UMPC_NX], soly[UMPC_NC];
  float scratch[28];
  void prefetch_warm_start() const {}
  void stash(int k, float v) { scratch[k] = v; }
  float unstash(int k) const { return scratch[k]; }
  float ld_xs(int j) const { return xs[j]; }
  float ld_zs(int i) const { return zs[i]; }
  float ld_ys(int i) const { return ys[i]; }
  float ld_D(int j) const { return D[j]; }
  float ld_E(int i) const { return E[i]; }
  void st_xs(int j, float v) { xs[j] = v; }
  void st_zs(int i, float v) { zs[i] = v; }
  void st_ys(int i, float v) { ys[i] = v; }
  void st_D(int j, float v) { D[j] = v; }
  void st_E(int i, float v) { E[i] = v; }
  void st_solx(int j, float v) { solx[j] = v; }
  void st_soly(int i, float v) { soly[i] = v; }
};

struct HcState {
  float T0;
  UmpcInfo info;
  HostStore st;
  HostIO io;
};

extern "C" {
int hc_sizeof_state() { return (int)sizeof(HcState); }
int hc_sizeof_params() { return (int)sizeof(UmpcParams); }
void hc_init(HcState* s) {
  memset(s, 0, sizeof(*s));
  for (int i = 0; i < UMPC_NX; ++i) s->io.D[i] = 1.0f;
  for (int i = 0; i < UMPC_NC; ++i) s->io.E[i] = 1.0f;
}
int hc_update(const UmpcParams* par, HcState* s, float* uquad, float* accdes, const float* p0, const float* R0,
              const float* dq0, const float* pdes, const float* dpdes, const float* sdes, float actualT0) {
  umpc_update(*par, s->st, s->io, uquad, accdes, p0, R0, dq0, pdes, dpdes, sdes, actualT0, s->T0, s->info);
  return 0;
}
void hc_get(const HcState* s, float* solx, float* soly, float* xs, float* zs, float* ys, float* D, float* E, float* info4) {
  memcpy(solx, s->io.solx, sizeof(s->io.solx)); memcpy(soly, s->io.soly, sizeof(s->io.soly));
  memcpy(xs, s->io.xs, sizeof(s->io.xs)); memcpy(zs, s->io.zs, sizeof(s->io.zs)); memcpy(ys, s->io.ys, sizeof(s->io.ys));
  memcpy(D, s->io.D, sizeof(s->io.D)); memcpy(E, s->io.E, sizeof(s->io.E));
  info4[0] = (float)s->info.iter; info4[1] = (float)s->info.status; info4[2] = s->info.pri_res; info4[3] = s->info.dua_res;
}
void hc_qpdata(const UmpcParams* par, const float* p0, const float* R0, const float* dq0, const float* pdes,
               const float* dpdes, const float* sdes, float T0, float* out216) {
  UmpcQP qp;
  umpc_build(*par, qp, p0, R0, dq0, pdes, dpdes, sdes, T0);
  umpc_qpdata(*par, qp, out216);
}
}
