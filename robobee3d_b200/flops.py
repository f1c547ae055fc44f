"""Algorithmic work model of one UprightMPC2 update (used by bench.py's roofline).

Two figures, both per QP solve (N=3, n=45, m=39), FMA = 2 flop, max/compare/select = 1:
  reference_algorithm : the count for the reference's own formulation (84x84 sparse KKT, nnz(L)=213,
                        Ruiz on stored P/A) from SURVEY.md section 8(d): 1.6e4 + 1.87e3*iters + 1.1e3
  ours                : what the CUDA kernel actually executes (reduced 45x45 form, nnz(L)=167, un-scaled, shifted
                        coordinates, A-free iteration), counted from the generator's structures.
"""
from .codegen.gen_umpc import build


def reference_algorithm_flops(iters=50):
    return 1.6e4 + 1.87e3 * iters + 1.1e3


def our_flops(iters=50, scaling_iters=10):
    nx, nc, ent, rows, kent, perm, cols = build(3)
    nnzA = len(ent)
    n_pm1 = sum(1 for (_, _, k) in ent if k[0] == "c")       # +-1 entries: add only
    n_coef = nnzA - n_pm1                                     # entries with a run-time coefficient: FMA
    nnzL = sum(len(c) for c in cols)
    # one ADMM iteration (A-free form, umpc_core.cuh): no product with A
    neq = nc - 3
    n_own_off = nnzA - 3 - neq - (n_coef - 3)                 # +-1 off-diagonal entries of the own-column block
    gu = (n_own_off + 3) * 2 + neq + 27 * 2 + 9               # v = M^-T w_o by substitution, w_u = kpu + B_u' v
    rhs = nx * 2 + 3 * 3                                      # r = S xh + w (+ W3 z3 - y3 on the thrust columns)
    solve = nnzL * 2 + (nx + nnzL * 2)                        # forward, then diagonal + backward
    wupd = neq * 7                                            # d, S d, (a c P) r, combine, w update
    xupd = nx * 2 + (nx - neq)                                # xh -= alpha d
    thr = 3 * 9                                               # generic z/y update + clamp on the thrust rows
    it = gu + rhs + solve + wupd + xupd + thr
    # Ruiz pass: norms (coefficient products + max), D/E updates (limit, sqrt, div, mul), cost scaling
    norms = 2 * n_coef + 2 * nnzA + 2 * (nnzA - nx - nc + nx)
    upd = nx * 9 + nc * 6
    cost = nx * 4 + 36 * 5 + 8
    ruiz = scaling_iters * (norms + upd + cost)
    # K assembly + numeric LDL'
    kasm = n_coef + sum(2 * len(v) for v in kent.values())
    fact = nx * 4 + nnzL + sum(len(c) * (len(c) + 1) for c in cols)
    setup = 250 + ruiz + kasm + fact + nx * 8 + nc * 10 + 2 * (n_pm1 + 2 * n_coef) + 150  # QP build, G/W, warm start, xp, w0/wf
    post = 2 * (n_pm1 + 2 * n_coef) + nc * 6 + nx * 12 + 60 + 250 + 100  # rebuild, un-shift, duals, residuals, status, store
    return dict(per_iter=it, setup=setup, post=post, total=setup + iters * it + post, nnzL=nnzL, nnzA=nnzA)


DYN_STEP_FLOPS = 150  # quadrotorNLDyn (genqp.py:24-41), SURVEY 8(d)

# algorithmic HBM bytes per env-step with MPC every step (SURVEY 8(d) accounting): plant state r/w
# (p3 R9 dq6 uq3) + controller state r/w (T0, xs45 zs39 ys39, D45 E39 written, E3 read) + outputs
BYTES_PER_SOLVE = 4 * (2 * 21 + (1 + 45 + 39 + 39) * 2 + 45 + 39 + 3 + 45 + 39 + 4 + 6 + 36 * 2)

if __name__ == "__main__":
    print(our_flops(), reference_algorithm_flops(), BYTES_PER_SOLVE)
