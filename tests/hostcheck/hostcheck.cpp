// tests/hostcheck/hostcheck.cpp -- TEST/DEVELOPMENT AID ONLY, never part of the product library.
// Compiles the device algorithm headers (robobee3d_b200/csrc/*.cuh are __host__ __device__)
// with g++ so the formulation can be compared with the oracle on this GPU-less box before
// GPU time is spent.  The product (libumpc_b200.so) has no CPU path and does not contain this.
#include <string.h>
#include "../../robobee3d_b200/csrc/umpc_core.cuh"

struct HostStore {
  float l[UMPC_GEN_NNZL], di[UMPC_NX], w[UMPC_NC], wi[UMPC_NC], sg[UMPC_NX], lo[UMPC_NC], hi[3], d[UMPC_NX], e[UMPC_NC];
  float& L(int n) { return l[n]; }
  float L(int n) const { return l[n]; }
  float& Di(int k) { return di[k]; }
  float Di(int k) const { return di[k]; }
  float& W(int i) { return w[i]; }
  float& Wi(int i) { return wi[i]; }
  float& Sg(int j) { return sg[j]; }
  float& Lo(int i) { return lo[i]; }
  float& Hi(int i) { return hi[i]; }
  float& Dsc(int j) { return d[j]; }
  float& Esc(int i) { return e[i]; }
};

struct HcState {
  float T0, xs[UMPC_NX], zs[UMPC_NC], ys[UMPC_NC], Eold[3], solx[UMPC_NX], soly[UMPC_NC];
  UmpcInfo info;
  HostStore st;
};

extern "C" {
int hc_sizeof_state() { return (int)sizeof(HcState); }
int hc_sizeof_params() { return (int)sizeof(UmpcParams); }
void hc_init(HcState* s) {
  memset(s, 0, sizeof(*s));
  s->Eold[0] = s->Eold[1] = s->Eold[2] = 1.0f;
}
int hc_update(const UmpcParams* par, HcState* s, float* uquad, float* accdes, const float* p0, const float* R0,
              const float* dq0, const float* pdes, const float* dpdes, const float* sdes, float actualT0) {
  umpc_update(*par, s->st, uquad, accdes, p0, R0, dq0, pdes, dpdes, sdes, actualT0, s->T0, s->xs, s->zs, s->ys,
              s->Eold, s->solx, s->soly, s->info);
  return 0;
}
void hc_get(const HcState* s, float* solx, float* soly, float* xs, float* zs, float* ys, float* D, float* E, float* info4) {
  memcpy(solx, s->solx, sizeof(s->solx)); memcpy(soly, s->soly, sizeof(s->soly));
  memcpy(xs, s->xs, sizeof(s->xs)); memcpy(zs, s->zs, sizeof(s->zs)); memcpy(ys, s->ys, sizeof(s->ys));
  memcpy(D, s->st.d, sizeof(s->st.d)); memcpy(E, s->st.e, sizeof(s->st.e));
  info4[0] = (float)s->info.iter; info4[1] = (float)s->info.status; info4[2] = s->info.pri_res; info4[3] = s->info.dua_res;
}
void hc_qpdata(const UmpcParams* par, const float* p0, const float* R0, const float* dq0, const float* pdes,
               const float* dpdes, const float* sdes, float T0, float* out216) {
  UmpcQP qp;
  umpc_build(*par, qp, p0, R0, dq0, pdes, dpdes, sdes, T0);
  umpc_qpdata(*par, qp, out216);
}
}
