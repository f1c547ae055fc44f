"""tests/golden/make_golden_pyloop.py -- golden trajectories from the reference's OWN PYTHON LOOP.

Imports the reference's Python sources (template/uprightmpc2.py, template_controllers.py, genqp.py,
flight_tasks.py) from /root/reference and runs `controlTest` (uprightmpc2.py:87-159) unmodified.
The reference's third-party / compiled dependencies that are absent from this image are replaced
by thin adapters onto the reference's own C code compiled by oracle/Makefile:

    uprightmpc2py (pybind module, needs Eigen)  -> ctypes onto oracle/_ref/libref_umpc_f32.so
                                                   (uprightmpc2.c + vendored OSQP, float32, 50 iterations)
    osqp (PyPI osqp==0.5.0, requirements.txt:3) -> the vendored OSQP 0.6.0 sources of template/uprightmpc2
                                                   built in float64 (oracle/_ref/libref_umpc_f64.so) with the
                                                   Python path's settings (eps 1e-4, template_controllers.py:191)
    matplotlib / progressbar / plot_helpers     -> inert stubs (plots only)

So the Python loop, flight_tasks, quadrotorNLDyn (scipy expm, float64), the Python QP-data build
(updateConstraint/updateObjective) and createMPC's defaults are the reference's; only the solver
backend is substituted by the reference's own vendored solver.

Run in the build container:  python tests/golden/make_golden_pyloop.py   ->  tests/golden/pyloop_c1.npz
"""
import os
import sys
import types
from unittest import mock

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference/template"

from oracle import binding as B  # noqa: E402


# ---- adapters ------------------------------------------------------------------------------------
class _UprightMPC2C:
    """pybind surface (uprightmpc2/py/uprightmpc2py.cpp:30-51) over the reference's C code."""

    def __init__(self, dt, g, TtoWmax, ws, wds, wpr, wpf, wvr, wvf, wthrust, wmom, Ib, maxIter):
        self.ref = B.RefUMPC(dt=dt, g=g, TtoWmax=TtoWmax, ws=ws, wds=wds, wpr=wpr, wpf=wpf, wvr=wvr, wvf=wvf,
                             wthrust=wthrust, wmom=wmom, Ib=tuple(float(v) for v in Ib), maxIter=int(maxIter))

    def update(self, p0, R0, dq0, pdes, dpdes, sdes, actualT0=-1.0):  # uprightmpc2.py:139 passes 6 arguments
        return self.ref.update(p0, R0, dq0, pdes, dpdes, sdes, actualT0)

    def vectors(self):
        l, u, q, Px, Ax, idx = self.ref.qpdata()
        return l, u, q

    def matrices(self):
        l, u, q, Px, Ax, idx = self.ref.qpdata()
        return Px, Ax, idx


class _WLCon:
    def __init__(self, u0, umin, umax, dumax, Qw, controlRate, popts):
        self.ref = B.RefWLCon(popts, u0=tuple(u0), umin=tuple(umin), umax=tuple(umax), dumax=tuple(dumax), Qw=tuple(Qw),
                              controlRate=float(controlRate))

    def update(self, h0, pdotdes):
        return self.ref.update(h0, pdotdes)


class _OSQP:
    """osqp.OSQP surface used by template_controllers.py:190-191,216-217 over the vendored OSQP (float64)."""
    STATUS = {1: "solved", 2: "solved inaccurate", -2: "maximum iterations reached", -3: "primal infeasible",
              -4: "dual infeasible", -7: "problem non convex", -10: "unsolved"}

    def setup(self, P=None, q=None, A=None, l=None, u=None, eps_rel=1e-3, eps_abs=1e-3, verbose=False, **kw):
        assert A.shape == (39, 45) and A.nnz == 111 and P.shape == (45, 45), "the vendored workspace is the N=3 instance"
        self.qp = B.RefQP64(eps=eps_abs)
        self.nA = A.nnz

    def update(self, Px=None, Ax=None, q=None, l=None, u=None):
        self.data = (l, u, q, Px, Ax)

    def solve(self):
        l, u, q, Px, Ax = self.data
        # full A.x update: Ax_idx = NULL semantics of osqp_update_P_A are expressed as the identity map
        x, y, info = self.qp.solve_full(l, u, q, Px, Ax)
        res = types.SimpleNamespace()
        res.x, res.y = x, y
        res.info = types.SimpleNamespace(status=self.STATUS.get(info["status"], "unsolved"), obj_val=info.get("obj_val", 0.0),
                                         iter=info["iter"], status_val=info["status"])
        return res


def _solve_full(self, l, u, q, Px, Ax):
    import ctypes as C
    idx = np.arange(len(Ax), dtype=np.int32)
    self.lib.ref_qp_update_solve(B._p(B._d(l, 39)), B._p(B._d(u, 39)), B._p(B._d(q, 45)), B._p(B._d(Px, 45)), 45,
                                 B._p(B._d(Ax, len(Ax))), idx.ctypes.data_as(C.POINTER(C.c_int)), len(Ax))
    x, y, info = np.zeros(45), np.zeros(39), np.zeros(5)
    self.lib.ref_get_solution(B._p(x), B._p(y))
    self.lib.ref_get_info(B._p(info))
    return x, y, dict(iter=int(info[0]), status=int(info[1]), pri_res=info[2], dua_res=info[3], obj_val=info[4])


B.RefQP64.solve_full = _solve_full


def install_stubs():
    m = types.ModuleType("uprightmpc2py")
    m.UprightMPC2C, m.WLCon = _UprightMPC2C, _WLCon
    sys.modules["uprightmpc2py"] = m
    o = types.ModuleType("osqp")
    o.OSQP = _OSQP
    sys.modules["osqp"] = o
    for name in ("matplotlib", "matplotlib.pyplot", "matplotlib.collections", "matplotlib.cm", "mpl_toolkits",
                 "mpl_toolkits.mplot3d", "mpl_toolkits.mplot3d.art3d", "progressbar", "plot_helpers", "matplotlib.colors",
                 "matplotlib.patches", "matplotlib.animation", "matplotlib.gridspec"):
        sys.modules[name] = mock.MagicMock()
    sys.path.insert(0, REF)


class PyWithDefaultT0:
    """template_controllers.UprightMPC2.update takes actualT0 positionally (template_controllers.py:252) while
    controlTest calls update with 6 arguments (uprightmpc2.py:139): supply -1 ('keep the internal T0')."""

    def __init__(self, up):
        self.up = up

    def update(self, *a):
        return self.up.update(*a, -1)


def main():
    B.build(ref=True)
    assert B.have_ref()
    install_stubs()
    import flight_tasks  # noqa: the reference's
    import uprightmpc2 as U  # noqa: the reference's
    from template_controllers import createMPC
    out = {}
    # ---- flight_tasks on a time grid (flight_tasks.py:6-49) ----
    tgrid = np.linspace(0.0, 700.0, 141)
    p_init = np.array([3.0, -2.0, 5.0])
    out["task_t"], out["task_pinit"] = tgrid, p_init
    cases = dict(helix=lambda t: flight_tasks.helix(t, p_init, trajAmp=50, trajFreq=1, dz=0.1, useY=False),
                 helixY=lambda t: flight_tasks.helix(t, p_init, trajAmp=80, trajFreq=1, dz=0.15, useY=True),
                 hover=lambda t: flight_tasks.helix(t, p_init, trajAmp=0, trajFreq=0, dz=0.1, useY=False),
                 straight=lambda t: flight_tasks.straightAcc(t, p_init, vdes=2, tduration=500),
                 flip=lambda t: flight_tasks.flip(t, p_init),
                 perch=lambda t: flight_tasks.perch(t, p_init))
    for k, f in cases.items():
        out["task_" + k] = np.array([np.hstack([np.asarray(v, float) for v in f(t)]) for t in tgrid])
    # ---- quadrotorNLDyn single steps (genqp.py:24-41) ----
    from genqp import quadrotorNLDyn
    from scipy.spatial.transform import Rotation
    rng = np.random.default_rng(2)
    dyn_in, dyn_out = [], []
    for i in range(40):
        p, dq, u = rng.normal(0, 10, 3), np.r_[rng.normal(0, 0.3, 3), rng.normal(0, 0.05 if i < 30 else 2.0, 3)], \
            np.r_[rng.uniform(0, 0.02), rng.normal(0, 20, 2)]
        R = Rotation.from_rotvec(rng.normal(0, 0.7, 3)).as_matrix()
        dt = 0.2 if i % 2 == 0 else 5.0
        p2, R2, dq2 = quadrotorNLDyn(p, R, dq, u, dt)
        dyn_in.append(np.r_[p, R.T.ravel(), dq, u, dt]); dyn_out.append(np.r_[p2, R2.T.ravel(), dq2])
    out["dyn_in"], out["dyn_out"] = np.array(dyn_in), np.array(dyn_out)
    # ---- controlTest (uprightmpc2.py:87-159) ----
    runs = dict(
        c_hl5=dict(kind="c", tend=500, kw=dict(hlInterval=5)),                         # config 1: uprightmpc2.py:383
        c_every=dict(kind="c", tend=60, kw=dict()),
        c_helix=dict(kind="c", tend=300, kw=dict(hlInterval=5, trajAmp=50, trajFreq=1)),  # uprightmpc2.py:387
        c_helixpert=dict(kind="c", tend=300, kw=dict(hlInterval=5, trajAmp=50, trajFreq=1, tpert=100)),
        c_flip=dict(kind="c", tend=300, kw=dict(hlInterval=5, flipTask=True)),
        c_perch=dict(kind="c", tend=300, kw=dict(hlInterval=5, perchTraj=True)),
        c_speed=dict(kind="c", tend=300, kw=dict(hlInterval=5, speedTest=True)),
        py_hl5=dict(kind="py", tend=500, kw=dict(hlInterval=5)),
        py_every=dict(kind="py", tend=40, kw=dict()),
    )
    for name, r in runs.items():
        up, upc = createMPC()
        mdl = upc if r["kind"] == "c" else PyWithDefaultT0(up)
        log = U.controlTest(mdl, r["tend"], useMPC=True, showPlots=False, **r["kw"])
        out[name + "_y"] = log["y"][::10].astype(np.float64)
        out[name + "_u"] = log["u"][::10].astype(np.float64)
        out[name + "_pdes"] = log["pdes"][::10].astype(np.float64)
        out[name + "_nt"] = np.array(len(log["t"]))
        out[name + "_final"] = log["y"][-1]
        # logMetric (uprightmpc2.py:161-175)
        out[name + "_metrics"] = np.array([np.mean(np.sum(log["y"][:, :3] ** 2, axis=1)), np.mean(np.sum(log["u"][:, 1:] ** 2, axis=1))])
        print(name, "nt", len(log["t"]), "final p", log["y"][-1, :3], "err/eff", out[name + "_metrics"])
    np.savez_compressed(os.path.join(HERE, "pyloop_c1.npz"), **out)
    print("pyloop_c1.npz", os.path.getsize(os.path.join(HERE, "pyloop_c1.npz")))


if __name__ == "__main__":
    main()
