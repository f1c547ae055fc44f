"""Generates tests/golden/dirtran_hover.npz from the CPU oracle (oracle/dirtran_oracle.c, float64):

  * hover_*: the reference's 'hover' run of planar/mpc_thrust_strokedev.py (IC y0[3:6] = (1, 0, 10), goal (-0.03, 0.02,
    0...), N = 5, dt = 0.03, eps 1e-5, 49 MPC steps, max_iter large enough that every solve converges) -- states,
    inputs, iteration counts;
  * mild_*: 16 single updates from small perturbations about the goal for N in {3, 5, 10, 20} -- u_0, iterations;
  * ref_x0: the initial state baked into the reference's own generated instance of this QP (planar/code/include/
    workspace.h, ldata[0:6] = -x0), extracted through oracle/_ref/libref_planar_f32.so: the state the reference's run
    had reached when its author generated the code, after the same 49 steps.

The osqp / autograd / matplotlib packages the reference's Python needs are absent here, so the script cannot import the
reference itself; the oracle's solver is pinned against the reference's compiled instance by tests/test_dirtran_oracle.py.
Run from the repo root:  python tests/golden/make_golden_dirtran.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import dt_util  # noqa: E402
from oracle import binding as OB  # noqa: E402

out = {}
o = OB.OracleDirTran(N=5, max_iter=1000000)
X, U, it, bad = o.rollout(dt_util.Y0_HOVER, dt_util.XR_HOVER, 49)
assert bad == 0
out.update(hover_X=X, hover_U=U, hover_iters=it)
for N in (3, 5, 10, 20):
    ys = dt_util.mild_states(16, seed=N)
    us = np.zeros((16, 2)); its = np.zeros(16); sts = np.zeros(16)
    for e in range(16):
        oo = OB.OracleDirTran(N=N, max_iter=200000)
        us[e], sts[e] = oo.update(ys[e], dt_util.XR_HOVER)
        its[e] = oo.solution()[2][0]
    assert (sts == 1).all(), sts
    out["mild_u_N%d" % N] = us
    out["mild_iters_N%d" % N] = its
if OB.have_ref_planar():
    P, A, q, l, u, rv, st = OB.RefPlanarQP().data()
    out["ref_x0"] = -l[:6]
np.savez(os.path.join(ROOT, "tests", "golden", "dirtran_hover.npz"), **out)
print({k: v.shape for k, v in out.items()})
print("hover iters", out["hover_iters"][:8], "final", out["hover_X"][-1], "ref_x0", out.get("ref_x0"))
