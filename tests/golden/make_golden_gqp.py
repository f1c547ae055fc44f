"""tests/golden/make_golden_gqp.py -- golden fixtures of the generic-QP rows (SURVEY 8 a17, a18 second instance, f1).

Run in the build container (needs /root/reference):  python tests/golden/make_golden_gqp.py
  fused_wlqp.npz  the reference's OWN functions of template/uprightmpc2_withWLQP.py (initConstraint, updateConstraint,
                  updateObjective, class UprightMPC2; the file has no imports, so it is exec'd with numpy / scipy.sparse /
                  genqp's skew injected and `osqp` replaced by an adapter onto the pinned OSQP restatement in float64 with
                  the file's settings: eps 1e-7) on a sequence of states: QP data (dense P, A, q, l, u) and solutions.
  mpc_osqp_quad.npz  the QP data that planar/mpc_osqp.py:29-103 builds (the file's statements up to the solver set-up,
                  exec'd unmodified with scipy.sparse) and the 15-step closed loop of :117-140 through the same adapter.
  umpc_genN.npz   template_controllers.py's own initConstraint / updateConstraint / updateObjective for N = 5.
"""
import os
import sys
import types

import numpy as np
import scipy.sparse as sp

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
REF = "/root/reference"

from oracle import binding as OB  # noqa: E402
import util  # noqa: E402


def skew(v):  # genqp.py:19
    return np.array([[0, -v[2], v[1]], [v[2], 0, -v[0]], [-v[1], v[0], 0]])


class OsqpAdapter:
    """the osqp.OSQP surface the reference files use, over oracle.binding.OracleQP (float64 restatement of OSQP)"""
    ADAPT = 100

    def __init__(self):
        self.log = []

    def setup(self, P=None, q=None, A=None, l=None, u=None, eps_rel=1e-3, eps_abs=1e-3, verbose=False, warm_start=True, **kw):
        self.Pp = sp.triu(sp.csc_matrix(P)).tocsc()
        self.Ap = sp.csc_matrix(A).copy()
        self.n, self.m = self.Ap.shape[1], self.Ap.shape[0]
        self.q = np.zeros(self.n) if q is None else np.array(q, float)
        self.l = np.array(l, float)
        self.u = np.full(self.m, np.inf) if u is None else np.array(u, float)
        self.o = OB.OracleQP(self.n, self.m, eps=eps_rel, adapt=self.ADAPT)
        self.dirty = True

    def dense(self):
        Pu = self.Pp.toarray()
        return Pu + Pu.T - np.diag(np.diag(Pu)), self.Ap.toarray()

    def update(self, Px=None, Ax=None, q=None, l=None, u=None):
        if Px is not None:
            self.Pp.data[:] = Px; self.dirty = True
        if Ax is not None:
            self.Ap.data[:] = Ax; self.dirty = True
        if q is not None:
            self.q = np.array(q, float)
        if l is not None:
            self.l = np.array(l, float)
        if u is not None:
            self.u = np.array(u, float)

    def solve(self):
        P, A = self.dense()
        x, y, info = self.o.update_solve(P if self.dirty else None, A if self.dirty else None, self.q, self.l, self.u)
        self.log.append(dict(P=P, A=A, q=self.q.copy(), l=self.l.copy(), u=self.u.copy(), x=x, y=y, **info))
        self.dirty = False
        res = types.SimpleNamespace(x=x, y=y, info=types.SimpleNamespace(status={1: "solved", 2: "solved inaccurate"}.get(info["status"], "unsolved"),
                                                                       obj_val=0.0, iter=info["iter"]))
        return res


def fused_wlqp():
    src = open(os.path.join(REF, "template", "uprightmpc2_withWLQP.py")).read()
    cut = src.index("def createMPC") if "def createMPC" in src else len(src)
    ns = dict(np=np, sp=sp, skew=skew, osqp=types.SimpleNamespace(OSQP=OsqpAdapter))
    exec(compile(src[:cut], "uprightmpc2_withWLQP.py", "exec"), ns)
    d = util.load_constants()
    ud, wd = d["umpc_defaults"], d["wl_defaults"]
    mk = lambda: ns["UprightMPC2"](3, ud["dt"], ud["g"], ud["TtoWmax"], ud["ws"], ud["wds"], ud["wpr"], ud["wpf"], ud["wvr"], ud["wvf"],
                                   ud["wthrust"], ud["wmom"], np.array(wd["umin"]), np.array(wd["umax"]), np.array(wd["dumax"]), 100.0,
                                   np.array(ud["Ib"]), np.array(wd["Qw"]), wd["controlRate"])
    nenv, ncall = 12, 4
    p, R, dq = util.config2_states(nenv, seed=31)
    rng = np.random.default_rng(5)
    out = {k: [] for k in ("p", "R", "dq", "pdes", "dpdes", "sdes", "P", "A", "q", "l", "u", "x", "y", "iter", "status", "uquad", "accdes", "u0", "T0")}
    u0_inits = []
    for e in range(nenv):
        c = mk()
        c.u0 = np.array([140.0, 0.0, 0.0, 0.0]) if e % 2 else np.zeros(4)   # also exercise the umin / umax rule (:258-262)
        u0_init = c.u0.copy()
        rec = {k: [] for k in out}
        for k in range(ncall):
            pe = p[e] + 0.5 * k
            pd, dpd = rng.uniform(-20, 20, 3), rng.uniform(-0.05, 0.05, 3)
            sd = np.array([0, 0, 1.0]) if k < 2 else np.array([np.sin(0.1), 0, np.cos(0.1)])
            uq = c.update2(pe, R[e], dq[e], pd, dpd, sd)
            acc = c.getAccDes(R[e], dq[e])
            lg = c.model.log[-1]
            vals = dict(p=pe, R=R[e], dq=dq[e], pdes=pd, dpdes=dpd, sdes=sd, P=lg["P"], A=lg["A"], q=lg["q"], l=lg["l"], u=lg["u"], x=lg["x"],
                        y=lg["y"], iter=lg["iter"], status=lg["status"], uquad=uq, accdes=acc, u0=c.u0.copy(), T0=c.T0)
            for kk in rec:
                rec[kk].append(np.asarray(vals[kk], float))
        for kk in out:
            out[kk].append(np.stack(rec[kk]))
        u0_inits.append(u0_init)
    np.savez_compressed(os.path.join(HERE, "fused_wlqp.npz"), u0_init=np.stack(u0_inits), **{k: np.stack(v) for k, v in out.items()})


def mpc_osqp_quad():
    src = open(os.path.join(REF, "planar", "mpc_osqp.py")).read()
    a, b = src.index("# Discrete time model of a quadcopter"), src.index("# Create an OSQP object")
    ns = dict(np=np, sparse=sp)
    exec(compile(src[a:b], "mpc_osqp.py", "exec"), ns)   # the file's own statements: Ad .. u
    P, A, q, l, u = ns["P"].toarray(), ns["A"].toarray(), ns["q"], ns["l"].copy(), ns["u"].copy()
    Ad, Bd, N, nx, nu = ns["Ad"].toarray(), ns["Bd"].toarray(), ns["N"], ns["nx"], ns["nu"]
    prob = OsqpAdapter()
    OsqpAdapter.ADAPT = 25
    prob.setup(ns["P"], q, ns["A"], l, u, warm_start=True, eps_rel=1e-2, eps_abs=1e-2)   # mpc_osqp.py:109
    x0 = ns["x0"].copy()
    nsim = 15
    X, U, IT = np.zeros((nsim, nx)), np.zeros((nsim, nu)), np.zeros(nsim)
    for i in range(nsim):                                   # mpc_osqp.py:117-140
        res = prob.solve()
        assert res.info.status == "solved"
        ctrl = res.x[-N * nu:-(N - 1) * nu]
        x0 = Ad @ x0 + Bd @ ctrl
        X[i], U[i], IT[i] = x0, ctrl, res.info.iter
        l[:nx] = -x0; u[:nx] = -x0
        prob.update(l=l, u=u)
    OsqpAdapter.ADAPT = 100
    np.savez_compressed(os.path.join(HERE, "mpc_osqp_quad.npz"), P=P, A=A, q=q, l=ns["l"], u=ns["u"], Ad=Ad, Bd=Bd, N=N, X=X, U=U, iters=IT,
                        x_first=prob.log[0]["x"], y_first=prob.log[0]["y"])


def umpc_genN(N=5):
    """template_controllers.py's own QP functions for a horizon the C build does not ship"""
    src = open(os.path.join(REF, "template", "template_controllers.py")).read()
    a, b = src.index("ny = 6"), src.index("class UprightMPC2")
    ns = dict(np=np, sp=sp, skew=skew)
    exec(compile(src[a:b], "template_controllers.py", "exec"), ns)
    d = util.load_constants()["umpc_defaults"]
    nx, nc = N * (2 * 6 + 3), 2 * N * 6 + N
    A, Pm = ns["initConstraint"](N, nx, nc)
    Qyr = np.hstack((np.full(3, d["wpr"]), np.full(3, d["ws"]))); Qyf = np.hstack((np.full(3, d["wpf"]), np.full(3, d["ws"])))
    Qdyr = np.hstack((np.full(3, d["wvr"]), np.full(3, d["wds"]))); Qdyf = np.hstack((np.full(3, d["wvf"]), np.full(3, d["wds"])))
    Rw = np.hstack((d["wthrust"], np.full(2, d["wmom"])))
    p, R, dq = util.config2_states(8, seed=41)
    e3h = skew([0, 0, 1]); Ibi = np.diag(1 / np.array(d["Ib"]))
    out = {k: [] for k in ("p", "R", "dq", "pdes", "dpdes", "sdes", "T0", "P", "A", "q", "l", "u", "x", "y", "iter")}
    rng = np.random.default_rng(9)
    for e in range(8):
        R0 = R[e]; s0 = R0[:, 2].copy(); Btau = (-R0 @ e3h @ Ibi)[:, :2]; ds0 = -R0 @ e3h @ dq[e][3:6]
        y0, dy0 = np.hstack((p[e], s0)), np.hstack((dq[e][:3], ds0))
        pd, dpd, sd = rng.uniform(-20, 20, 3), rng.uniform(-0.05, 0.05, 3), np.array([0, 0, 1.0])
        T0 = 9.81e-3 * (1 + 0.1 * e)
        A, l, u, _ = ns["updateConstraint"](N, A, d["dt"], T0, [s0] * N, [Btau] * N, y0, dy0, d["g"], d["TtoWmax"] * d["g"])
        Px, q = ns["updateObjective"](N, Qyr, Qyf, Qdyr, Qdyf, Rw, np.hstack((pd, sd)), np.hstack((dpd, 0, 0, 0)))
        o = OB.OracleQP(nx, nc, eps=1e-4)
        x, y, info = o.update_solve(np.diag(Px), A.toarray(), q, l, u)
        vals = dict(p=p[e], R=R0, dq=dq[e], pdes=pd, dpdes=dpd, sdes=sd, T0=T0, P=np.diag(Px), A=A.toarray(), q=q, l=l, u=u, x=x, y=y, iter=info["iter"])
        for k in out:
            out[k].append(np.asarray(vals[k], float))
    np.savez_compressed(os.path.join(HERE, "umpc_genN.npz"), N=N, **{k: np.stack(v) for k, v in out.items()})


if __name__ == "__main__":
    OB.build(ref=True)
    fused_wlqp()
    mpc_osqp_quad()
    umpc_genN()
    for f in ("fused_wlqp.npz", "mpc_osqp_quad.npz", "umpc_genN.npz"):
        g = np.load(os.path.join(HERE, f))
        print(f, os.path.getsize(os.path.join(HERE, f)), {k: g[k].shape for k in g.files if g[k].ndim > 1})
