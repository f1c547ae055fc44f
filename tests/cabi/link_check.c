/* tests/cabi/link_check.c -- TEST INFRASTRUCTURE.  A plain C program that LINKS libumpc_b200.so through include/umpc_b200.h
 * the way a C caller of the reference would (template/uprightmpc2/uprightmpc2.h:45-49; g4bee/app/loop_update.cpp:22-60):
 * umpcInit + a few umpcUpdate calls, wlConInit/wlConUpdate, prints the outputs.  Built on the CPU box (link check), run on
 * the GPU box (tests/test_gpu_parity.py::test_c_program_links_and_runs). */
#include <stdio.h>
#include <string.h>
#include "umpc_b200.h"

int main(void) {
  static UprightMPC_t up;
  const float Ib[3] = {3333.0f, 3333.0f, 1000.0f};
  float uquad[3], accdes[6];
  const float p0[3] = {1.0f, -2.0f, 3.0f}, R0[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, dq0[6] = {0.1f, 0, 0, 0, 0, 0};
  const float pdes[3] = {0, 0, 0}, dpdes[3] = {0, 0, 0}, sdes[3] = {0, 0, 1};
  int k, rc = 0;
  umpcInit(&up, 5.0f, 9.81e-3f, 2.0f, 1e1f, 1e3f, 1.0f, 5.0f, 1e3f, 2e3f, 1e-1f, 1e-2f, Ib, 50);
  for (k = 0; k < 3; ++k) {
    rc |= umpcUpdate(&up, uquad, accdes, p0, R0, dq0, pdes, dpdes, sdes, k == 0 ? 9.81e-3f : -1.0f);
    printf("umpcUpdate %d rc %d uquad %.9g %.9g %.9g accdes %.9g %.9g %.9g T0 %.9g l0 %.9g\n", k, rc, uquad[0], uquad[1], uquad[2],
           accdes[0], accdes[1], accdes[2], up.T0, up.l[0]);
  }
  umpcFree(&up);
  return rc;
}
