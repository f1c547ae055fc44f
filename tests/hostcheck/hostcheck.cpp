// tests/hostcheck/hostcheck.cpp -- TEST/DEVELOPMENT AID ONLY, never part of the product library.
// Compiles the device algorithm headers (robobee3d_b200/csrc/*.cuh are __host__ __device__)
// with g++ so the formulation can be compared with the oracle on this GPU-less box before
// GPU time is spent.  Built twice: float (libhostcheck.so) and -DUMPC_REAL=double (libhostcheck64.so).  The product (libumpc_b200.so) has no CPU path and does not contain this.
#include <string.h>
#include "../../robobee3d_b200/csrc/umpc_core.cuh"

struct HostStore {
  ureal wds[UMPC_GEN_NWORDS], vf[UMPC_NX];
  void issue(int c, ureal (&buf)[UMPC_GEN_CHUNK]) const { for (int i = 0; i < UMPC_GEN_CHUNK; ++i) buf[i] = wds[c * UMPC_GEN_CHUNK + i]; }
  void wait(ureal (&)[UMPC_GEN_CHUNK]) const {}
  void waitc(int, ureal (&)[UMPC_GEN_CHUNK]) const {}
  void setW(int e, ureal v) { wds[e] = v; }
  void setChunk(int c, const ureal (&v)[UMPC_GEN_CHUNK]) { for (int i = 0; i < UMPC_GEN_CHUNK; ++i) wds[c * UMPC_GEN_CHUNK + i] = v[i]; }
  void end_writes() {}
  void sync() const {}
  ureal Vf(int i) const { return vf[i]; }
  void setVf(int i, ureal v) { vf[i] = v; }
  void begin_iteration(int) {}
  bool all(bool p) const { return p; }
};

struct HostIO {
  ureal xs[UMPC_NX], zs[UMPC_NC], ys[UMPC_NC], D[UMPC_NX], E[UMPC_NC], solx[UMPC_NX], soly[UMPC_NC];
  ureal scratch[28], rho;
  ureal ld_rho() const { return rho; }
  void st_rho(ureal v) { rho = v; }
  void prefetch_warm_start() const {}
  void stash(int k, ureal v) { scratch[k] = v; }
  ureal unstash(int k) const { return scratch[k]; }
  ureal ld_xs(int j) const { return xs[j]; }
  ureal ld_zs(int i) const { return zs[i]; }
  ureal ld_ys(int i) const { return ys[i]; }
  ureal ld_D(int j) const { return D[j]; }
  ureal ld_E(int i) const { return E[i]; }
  void st_xs(int j, ureal v) { xs[j] = v; }
  void st_zs(int i, ureal v) { zs[i] = v; }
  void st_ys(int i, ureal v) { ys[i] = v; }
  void st_D(int j, ureal v) { D[j] = v; }
  void st_E(int i, ureal v) { E[i] = v; }
  void st_solx(int j, ureal v) { solx[j] = v; }
  void st_soly(int i, ureal v) { soly[i] = v; }
};

struct HcState {
  ureal T0;
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
int hc_update(const UmpcParams* par, HcState* s, ureal* uquad, ureal* accdes, const ureal* p0, const ureal* R0,
              const ureal* dq0, const ureal* pdes, const ureal* dpdes, const ureal* sdes, ureal actualT0) {
  if (par->check_termination > 0) {
    if (s->io.rho == 0) s->io.rho = par->rho;
    umpc_update<true>(*par, s->st, s->io, uquad, accdes, p0, R0, dq0, pdes, dpdes, sdes, actualT0, s->T0, s->info);
  } else {
    umpc_update<false>(*par, s->st, s->io, uquad, accdes, p0, R0, dq0, pdes, dpdes, sdes, actualT0, s->T0, s->info);
  }
  return 0;
}
void hc_get(const HcState* s, ureal* solx, ureal* soly, ureal* xs, ureal* zs, ureal* ys, ureal* D, ureal* E, ureal* info4 /* [5]: iter, status, pri_res, dua_res, rho */) {
  memcpy(solx, s->io.solx, sizeof(s->io.solx)); memcpy(soly, s->io.soly, sizeof(s->io.soly));
  memcpy(xs, s->io.xs, sizeof(s->io.xs)); memcpy(zs, s->io.zs, sizeof(s->io.zs)); memcpy(ys, s->io.ys, sizeof(s->io.ys));
  memcpy(D, s->io.D, sizeof(s->io.D)); memcpy(E, s->io.E, sizeof(s->io.E));
  info4[4] = s->io.rho;
  info4[0] = (ureal)s->info.iter; info4[1] = (ureal)s->info.status; info4[2] = s->info.pri_res; info4[3] = s->info.dua_res;
}
void hc_qpdata(const UmpcParams* par, const ureal* p0, const ureal* R0, const ureal* dq0, const ureal* pdes,
               const ureal* dpdes, const ureal* sdes, ureal T0, ureal* out216) {
  UmpcQP qp;
  umpc_build(*par, qp, p0, R0, dq0, pdes, dpdes, sdes, T0);
  umpc_qpdata(*par, qp, out216);
}
}
