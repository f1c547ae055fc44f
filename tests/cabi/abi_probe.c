/* tests/cabi/abi_probe.c -- TEST INFRASTRUCTURE.  Prints sizeof / offsetof of every struct of the drop-in boundary.
 * Compiled twice by tests/test_host_side.py::test_abi_layout_equals_reference_headers:
 *   -DPROBE_REF_UMPC  -I/root/reference/template/uprightmpc2     the reference's own uprightmpc2.h + funapprox.h
 *   -DPROBE_REF_WLQP  -I/root/reference/template/wlcontroller/c  the reference's own wlqp.h + funapprox.h
 *   (neither)         -I include                                  ours (include/umpc_b200.h)
 * and the printed tables are compared line by line: the layouts are checked by the compiler, not by formula. */
#include <stddef.h>
#include <stdio.h>
#if defined(PROBE_REF_UMPC)
#include "uprightmpc2.h"
#include "funapprox.h"
#define HAVE_UMPC 1
#elif defined(PROBE_REF_WLQP)
#include "wlqp.h"
#include "funapprox.h"
#define HAVE_WLQP 1
#else
#include "umpc_b200.h"
#define HAVE_UMPC 1
#define HAVE_WLQP 1
#endif
#define SZ(T) printf(#T " sizeof %zu\n", sizeof(T))
#define OFF(T, f) printf(#T "." #f " offset %zu size %zu\n", offsetof(T, f), sizeof(((T *)0)->f))
int main(int argc, char **argv) {
  int which = argc > 1 ? argv[1][0] : 'a';
#ifdef HAVE_UMPC
  if (which == 'a' || which == 'u') {
    SZ(UprightMPC_t);
    OFF(UprightMPC_t, dt); OFF(UprightMPC_t, g); OFF(UprightMPC_t, Tmax); OFF(UprightMPC_t, Qyr); OFF(UprightMPC_t, Qyf);
    OFF(UprightMPC_t, Qdyr); OFF(UprightMPC_t, Qdyf); OFF(UprightMPC_t, R); OFF(UprightMPC_t, smin); OFF(UprightMPC_t, smax);
    OFF(UprightMPC_t, e3h); OFF(UprightMPC_t, e3hIbi); OFF(UprightMPC_t, l); OFF(UprightMPC_t, u); OFF(UprightMPC_t, q);
    OFF(UprightMPC_t, Px_data); OFF(UprightMPC_t, Ax_data); OFF(UprightMPC_t, Ax_idx); OFF(UprightMPC_t, nAxT0dt);
    OFF(UprightMPC_t, nAxdt); OFF(UprightMPC_t, c0); OFF(UprightMPC_t, T0);
    SZ(WLCon_t);
    OFF(WLCon_t, u0); OFF(WLCon_t, umin); OFF(WLCon_t, umax); OFF(WLCon_t, dumax); OFF(WLCon_t, Qw); OFF(WLCon_t, fa);
    printf("UMPC_N %d UMPC_NY %d UMPC_NU %d UMPC_NX %d UMPC_NC %d UMPC_nAdata %d NDELU %d\n", UMPC_N, UMPC_NY, UMPC_NU, UMPC_NX,
           UMPC_NC, UMPC_nAdata, NDELU);
  }
#endif
#ifdef HAVE_WLQP
  if (which == 'a' || which == 'w') {
    SZ(WLQP_t);
    OFF(WLQP_t, umin); OFF(WLQP_t, umax); OFF(WLQP_t, U0); OFF(WLQP_t, w0); OFF(WLQP_t, Q);
  }
#endif
  if (which == 'a' || which == 'f') {
    SZ(FunApprox_t);
    OFF(FunApprox_t, k); OFF(FunApprox_t, a0); OFF(FunApprox_t, a1); OFF(FunApprox_t, A2);
  }
  return 0;
}
