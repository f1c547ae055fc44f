/*
 * oracle/ref_planar_glue.c -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Drives the reference's own generated instance of the planar dir-tran MPC QP (planar/code: OSQP 0.5.0, EMBEDDED=1,
 * float32, n=46 m=82, rho=5.694, 50 iterations; planar/code/include/workspace.h) from our tests.  Compiled TOGETHER WITH
 * the unmodified reference sources where they lie (oracle/Makefile -> oracle/_ref/libref_planar_f32.so).  The generated
 * header DEFINES the workspace, so this is the one translation unit that includes it (as planar/mcuqp/main.cpp does);
 * like that file it is compiled as C++ (workspace.h:1071 uses the enum name linsys_solver_type as a type).
 */
#include <string.h>
#include "osqp.h"
#include "workspace.h"

extern "C" {

void rp_dims(int *nm) { nm[0] = (int)workspace.data->n; nm[1] = (int)workspace.data->m; }

/* dense row-major copies of the instance's data + its per-row rho and the settings that matter */
void rp_data(double *P, double *A, double *q, double *l, double *u, double *rho_vec, double *st) {
  const int n = (int)workspace.data->n, m = (int)workspace.data->m;
  int j, k;
  memset(P, 0, sizeof(double) * n * n);
  memset(A, 0, sizeof(double) * m * n);
  for (j = 0; j < n; ++j) {
    for (k = workspace.data->P->p[j]; k < workspace.data->P->p[j + 1]; ++k) {
      const int i = (int)workspace.data->P->i[k];
      P[i * n + j] = P[j * n + i] = workspace.data->P->x[k];   /* upper triangle stored */
    }
    for (k = workspace.data->A->p[j]; k < workspace.data->A->p[j + 1]; ++k) A[(int)workspace.data->A->i[k] * n + j] = workspace.data->A->x[k];
    q[j] = workspace.data->q[j];
  }
  for (k = 0; k < m; ++k) { l[k] = workspace.data->l[k]; u[k] = workspace.data->u[k]; rho_vec[k] = workspace.rho_vec[k]; }
  st[0] = workspace.settings->eps_abs; st[1] = workspace.settings->eps_rel; st[2] = workspace.settings->rho;
  st[3] = workspace.settings->sigma; st[4] = workspace.settings->alpha; st[5] = (double)workspace.settings->max_iter;
  st[6] = (double)workspace.settings->check_termination; st[7] = 0.0; st[8] = 5.0;
  st[9] = (double)workspace.settings->scaling;
}

void rp_cold_start(void) {
  const int n = (int)workspace.data->n, m = (int)workspace.data->m;
  memset(workspace.x, 0, sizeof(c_float) * n);
  memset(workspace.y, 0, sizeof(c_float) * m);
  memset(workspace.z, 0, sizeof(c_float) * m);
}

int rp_update(const double *q, const double *l, const double *u) {
  const int n = (int)workspace.data->n, m = (int)workspace.data->m;
  c_float qf[64], lf[128], uf[128];
  int i, rc = 0;
  for (i = 0; i < n; ++i) qf[i] = (c_float)q[i];
  for (i = 0; i < m; ++i) { lf[i] = (c_float)l[i]; uf[i] = (c_float)u[i]; }
  rc |= (int)osqp_update_lin_cost(&workspace, qf);
  rc |= (int)osqp_update_bounds(&workspace, lf, uf);
  return rc;
}

int rp_update_max_iter(int it) { return (int)osqp_update_max_iter(&workspace, it); }

/* info: iter, status_val, pri_res, dua_res */
int rp_solve(double *x, double *y, double *info) {
  const int n = (int)workspace.data->n, m = (int)workspace.data->m;
  int i;
  const int rc = (int)osqp_solve(&workspace);
  for (i = 0; i < n; ++i) x[i] = workspace.solution->x[i];
  for (i = 0; i < m; ++i) y[i] = workspace.solution->y[i];
  info[0] = (double)workspace.info->iter; info[1] = (double)workspace.info->status_val;
  info[2] = workspace.info->pri_res; info[3] = workspace.info->dua_res;
  return rc;
}
}  /* extern C */
