/*
 * oracle/ref_glue.c -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Glue that lets the reference's own C controller be driven from our tests:
 * it is compiled TOGETHER WITH the unmodified reference sources where they lie
 * under /root/reference (see oracle/Makefile) into oracle/_ref/ *.so.
 *
 *  - matMult / lsSolve: the reference declares these in
 *    template/uprightmpc2/matmult.h:37-48 and expects the application to
 *    provide them (its own providers need Eigen or CBLAS, both absent here).
 *  - ref_ctx_*: the reference keeps all solver state in ONE process-global
 *    OSQPWorkspace (template/uprightmpc2/workspace.c:2620).  To run several
 *    independent flights we snapshot/restore that state around each call.
 *  - small accessors for info / solution (workspace.info, workspace.solution).
 *
 * Works for both generated workspaces (uprightmpc2: n=45,m=39; wlqp: n=m=4)
 * because it only walks the OSQPWorkspace pointers.
 */
#include <string.h>
#include "osqp.h"
#include "workspace.h"

#ifndef REF_NO_MATMULT
/* C = alpha * op(A) * op(B), all column-major (matmult.h:20,37). */
void matMult(float *C, const float *A, const float *B, const int m, const int n,
             const int k, const float alpha, int AT, int BT) {
  int i, j, p;
  for (j = 0; j < n; ++j)
    for (i = 0; i < m; ++i) {
      float acc = 0.0f;
      for (p = 0; p < k; ++p) {
        float a = AT ? A[p + i * k] : A[i + p * m];
        float b = BT ? B[j + p * n] : B[p + j * k];
        acc += a * b;
      }
      C[i + j * m] = alpha * acc;
    }
}
/* never called on this path (grep: only declared); kept so the symbol resolves */
void lsSolve(float *x, const float *A, const float *b, const int m, const int n) {
  (void)A; (void)b; (void)m;
  memset(x, 0, (size_t)n * sizeof(float));
}
#endif

typedef struct { c_float *p; int n; } span_t;

static int ctx_spans(span_t *s) {
  OSQPWorkspace *w = &workspace;
  qdldl_solver *ls = (qdldl_solver *)w->linsys_solver;
  int n = (int)w->data->n, m = (int)w->data->m, k = 0;
  int nnzP = (int)w->data->P->p[n], nnzA = (int)w->data->A->p[n];
  int N = n + m;
  int nnzL = (int)ls->L->p[N], nnzK = (int)ls->KKT->p[N];
#define SP(ptr, cnt) do { s[k].p = (c_float *)(ptr); s[k].n = (cnt); ++k; } while (0)
  SP(w->data->P->x, nnzP); SP(w->data->A->x, nnzA);
  SP(w->data->q, n); SP(w->data->l, m); SP(w->data->u, m);
  SP(w->scaling->D, n); SP(w->scaling->Dinv, n);
  SP(w->scaling->E, m); SP(w->scaling->Einv, m);
  SP(w->rho_vec, m); SP(w->rho_inv_vec, m);
  SP(w->x, n); SP(w->y, m); SP(w->z, m); SP(w->x_prev, n); SP(w->z_prev, m);
  SP(ls->L->x, nnzL); SP(ls->Dinv, N); SP(ls->D, N); SP(ls->KKT->x, nnzK);
  SP(ls->rho_inv_vec, m);
#undef SP
  return k;
}

/* number of c_float slots a context snapshot needs */
int ref_ctx_floats(void) {
  span_t s[32]; int k = ctx_spans(s), i, tot = 0;
  for (i = 0; i < k; ++i) tot += s[i].n;
  return tot + (int)workspace.data->m /*constr_type*/ + 3 /*c,cinv,rho*/;
}

void ref_ctx_save(c_float *buf) {
  span_t s[32]; int k = ctx_spans(s), i, j, m = (int)workspace.data->m;
  for (i = 0; i < k; ++i) { memcpy(buf, s[i].p, (size_t)s[i].n * sizeof(c_float)); buf += s[i].n; }
  for (j = 0; j < m; ++j) *buf++ = (c_float)workspace.constr_type[j];
  *buf++ = workspace.scaling->c; *buf++ = workspace.scaling->cinv; *buf++ = workspace.settings->rho;
}

void ref_ctx_load(const c_float *buf) {
  span_t s[32]; int k = ctx_spans(s), i, j, m = (int)workspace.data->m;
  for (i = 0; i < k; ++i) { memcpy(s[i].p, buf, (size_t)s[i].n * sizeof(c_float)); buf += s[i].n; }
  for (j = 0; j < m; ++j) workspace.constr_type[j] = (c_int)(*buf++);
  workspace.scaling->c = *buf++; workspace.scaling->cinv = *buf++; workspace.settings->rho = *buf++;
}

int ref_sizeof_cfloat(void) { return (int)sizeof(c_float); }
int ref_n(void) { return (int)workspace.data->n; }
int ref_m(void) { return (int)workspace.data->m; }

/* info[0..4] = iter, status_val, pri_res, dua_res, obj_val (types.h:64-89) */
void ref_get_info(double *info5) {
  info5[0] = (double)workspace.info->iter;
  info5[1] = (double)workspace.info->status_val;
  info5[2] = (double)workspace.info->pri_res;
  info5[3] = (double)workspace.info->dua_res;
  info5[4] = (double)workspace.info->obj_val;
}

void ref_get_solution(double *x, double *y) {
  int i, n = (int)workspace.data->n, m = (int)workspace.data->m;
  for (i = 0; i < n; ++i) x[i] = (double)workspace.solution->x[i];
  for (i = 0; i < m; ++i) y[i] = (double)workspace.solution->y[i];
}

/* scaled iterates + scaling, for debugging the restatement */
void ref_get_iterates(double *x, double *z, double *y, double *D, double *E, double *c) {
  int i, n = (int)workspace.data->n, m = (int)workspace.data->m;
  for (i = 0; i < n; ++i) { x[i] = workspace.x[i]; D[i] = workspace.scaling->D[i]; }
  for (i = 0; i < m; ++i) { z[i] = workspace.z[i]; y[i] = workspace.y[i]; E[i] = workspace.scaling->E[i]; }
  *c = workspace.scaling->c;
}

/* generic "set data + solve" entry used by the float64 converged build, where
 * uprightmpc2.c itself cannot be used (its API is hard-typed float).  Mirrors
 * the call order of uprightmpc2.c:247-250. */
int ref_qp_update_solve(const c_float *l, const c_float *u, const c_float *q,
                        const c_float *Px, int nPx, const c_float *Ax,
                        const c_int *Ax_idx, int nAx) {
  osqp_update_bounds(&workspace, l, u);
  osqp_update_lin_cost(&workspace, q);
  osqp_update_P_A(&workspace, Px, OSQP_NULL, nPx, Ax, Ax_idx, nAx);
  return (int)osqp_solve(&workspace);
}

/* settings overrides (osqp.h update_* API) */
void ref_set_iters(int max_iter, int check_termination) {
  osqp_update_max_iter(&workspace, max_iter);
  osqp_update_check_termination(&workspace, check_termination);
}
void ref_set_eps(double eps_abs, double eps_rel) {
  osqp_update_eps_abs(&workspace, (c_float)eps_abs);
  osqp_update_eps_rel(&workspace, (c_float)eps_rel);
}
void ref_set_adaptive_rho_interval(int k) { workspace.settings->adaptive_rho_interval = k; }
