"""ctypes wrappers of the generic warp-per-QP solver (include/gqp_b200.h) and of the horizon-generic / fused-WLQP
UprightMPC2 controller on top of it (include/umg_b200.h).

    BatchQP                 B QPs sharing A's sparsity pattern: set_data / solve / solution (OSQP semantics, float64)
    BatchUprightMPC2G       B UprightMPC2 controllers of ANY horizon N, optionally fused with the WLQP
    UprightMPC2WithWLQP     the reference's archived class of that name (template/uprightmpc2_withWLQP.py:218-324):
                            same constructor arguments and methods, one controller on the GPU
    lti_mpc_osqp            the LTI quadcopter MPC of planar/mpc_osqp.py:29-140 as a BatchQP + device closed loop
"""
import ctypes as C

import numpy as np

from . import _lib

DP = C.POINTER(C.c_double)
IP = C.POINTER(C.c_int)


class GqpSettings(C.Structure):
    _fields_ = [("n", C.c_int), ("m", C.c_int), ("nnz", C.c_int), ("max_iter", C.c_int), ("check_termination", C.c_int),
                ("adaptive_rho_interval", C.c_int), ("scaling_iters", C.c_int), ("rho", C.c_double), ("sigma", C.c_double),
                ("alpha", C.c_double), ("eps_abs", C.c_double), ("eps_rel", C.c_double), ("adaptive_rho_tol", C.c_double),
                ("infty", C.c_double)]


class UmgParams(C.Structure):
    _fields_ = [("N", C.c_int), ("dt", C.c_double), ("g", C.c_double), ("TtoWmax", C.c_double), ("ws", C.c_double),
                ("wds", C.c_double), ("wpr", C.c_double), ("wpf", C.c_double), ("wvr", C.c_double), ("wvf", C.c_double),
                ("wthrust", C.c_double), ("wmom", C.c_double), ("Ib", C.c_double * 3), ("maxIter", C.c_int),
                ("checkTermination", C.c_int), ("adaptiveRhoInterval", C.c_int), ("epsAbs", C.c_double), ("epsRel", C.c_double),
                ("withWL", C.c_int), ("umin", C.c_double * 4), ("umax", C.c_double * 4), ("dumax", C.c_double * 4),
                ("Qw", C.c_double * 6), ("controlRate", C.c_double), ("mb", C.c_double), ("wrenchModel", C.c_int),
                ("popts", C.c_double * 90)]


GQP_EXPORTS = ["gqpDefaults", "gqpBatchCreate", "gqpBatchDestroy", "gqpBatchReset", "gqpBatchSetData", "gqpBatchSolve",
               "gqpBatchGetSolution", "gqpBatchSync", "gqpBatchLastKernelMs", "gqpBatchDevPtr", "gqpLtiRollout"]
UMG_EXPORTS = ["umgDefaults", "umgSizes", "umgPattern", "umgCreate", "umgDestroy", "umgReset", "umgSetState", "umgUpdate", "umgGetQP", "umgLastSolveMs"]

_bound = False


def _bind():
    global _bound
    L = _lib.load()
    if _bound:
        return L
    vp = C.c_void_p
    L.gqpDefaults.argtypes = [C.POINTER(GqpSettings), C.c_int, C.c_int, C.c_int]
    L.gqpDefaults.restype = None
    L.gqpBatchCreate.argtypes = [C.POINTER(vp), C.c_int, C.c_int, C.POINTER(GqpSettings), IP, IP]
    L.gqpBatchDestroy.argtypes = [vp]
    L.gqpBatchDestroy.restype = None
    L.gqpBatchReset.argtypes = [vp]
    L.gqpBatchSetData.argtypes = [vp, DP, DP, DP, DP, DP, C.c_int]
    L.gqpBatchSolve.argtypes = [vp, C.c_int]
    L.gqpBatchGetSolution.argtypes = [vp, DP, DP, DP]
    L.gqpBatchSync.argtypes = [vp]
    L.gqpBatchLastKernelMs.argtypes = [vp]
    L.gqpBatchLastKernelMs.restype = C.c_float
    L.gqpBatchDevPtr.argtypes = [vp, C.c_char_p]
    L.gqpBatchDevPtr.restype = vp
    L.gqpLtiRollout.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, DP, DP, DP, DP, DP, C.c_int]
    L.umgDefaults.argtypes = [C.POINTER(UmgParams), C.c_int, C.c_int]
    L.umgDefaults.restype = None
    L.umgSizes.argtypes = [C.c_int, C.c_int, IP, IP, IP]
    L.umgSizes.restype = None
    L.umgPattern.argtypes = [C.c_int, C.c_int, IP, IP]
    L.umgPattern.restype = None
    L.umgCreate.argtypes = [C.POINTER(vp), C.c_int, C.c_int, C.POINTER(UmgParams)]
    L.umgDestroy.argtypes = [vp]
    L.umgDestroy.restype = None
    L.umgReset.argtypes = [vp]
    L.umgSetState.argtypes = [vp, DP, DP]
    L.umgUpdate.argtypes = [vp] + [DP] * 10
    L.umgGetQP.argtypes = [vp] + [DP] * 8
    L.umgLastSolveMs.argtypes = [vp]
    L.umgLastSolveMs.restype = C.c_float
    _bound = True
    return L


def _d(a, shape=None):
    a = np.ascontiguousarray(np.asarray(a, dtype=np.float64))
    if shape is not None and a.shape != tuple(shape):
        a = np.ascontiguousarray(np.broadcast_to(a, shape))
    return a


def _p(a):
    return None if a is None else a.ctypes.data_as(DP)


def csr_pattern(A):
    """(rowptr, colidx) of a dense / scipy-sparse matrix's structural non-zeros, rows walked with columns increasing"""
    A = np.asarray(A.toarray() if hasattr(A, "toarray") else A)
    rowptr, colidx = [0], []
    for r in range(A.shape[0]):
        cols = np.nonzero(A[r])[0]
        colidx.extend(int(c) for c in cols)
        rowptr.append(len(colidx))
    return np.asarray(rowptr, np.int32), np.asarray(colidx, np.int32)


class BatchQP:
    """B QPs  min 1/2 x'Px + q'x, l <= Ax <= u  with one CSR pattern of A; OSQP's algorithm, one warp per QP."""

    def __init__(self, nqp, n, m, rowptr, colidx, device=0, **settings):
        self.L = _bind()
        self.B, self.n, self.m = int(nqp), int(n), int(m)
        self.rowptr = np.ascontiguousarray(rowptr, np.int32)
        self.colidx = np.ascontiguousarray(colidx, np.int32)
        self.nnz = int(self.rowptr[-1])
        s = GqpSettings()
        self.L.gqpDefaults(C.byref(s), self.n, self.m, self.nnz)
        for k, v in settings.items():
            if not hasattr(s, k):
                raise TypeError("unknown setting %r" % k)
            setattr(s, k, v)
        self.settings = s
        self.h = C.c_void_p()
        _lib.check(self.L.gqpBatchCreate(C.byref(self.h), self.B, int(device), C.byref(s), self.rowptr.ctypes.data_as(IP),
                                         self.colidx.ctypes.data_as(IP)), "gqpBatchCreate")

    def close(self):
        if getattr(self, "h", None):
            self.L.gqpBatchDestroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reset(self):
        _lib.check(self.L.gqpBatchReset(self.h), "gqpBatchReset")

    def csr_values(self, A):
        """values of dense A (m, n) or (B, m, n) in the batch's CSR order"""
        A = np.asarray(A, float)
        rows = np.repeat(np.arange(self.m), np.diff(self.rowptr))
        return A[..., rows, self.colidx]

    def set_data(self, P=None, Av=None, q=None, l=None, u=None, broadcast=False):
        B = 1 if broadcast else self.B
        arr = [None if P is None else _d(P, (B, self.n, self.n)), None if Av is None else _d(Av, (B, self.nnz)),
               None if q is None else _d(q, (B, self.n)), None if l is None else _d(l, (B, self.m)),
               None if u is None else _d(u, (B, self.m))]
        _lib.check(self.L.gqpBatchSetData(self.h, *[_p(a) for a in arr], int(bool(broadcast))), "gqpBatchSetData")

    def solve(self, rescale=True):
        _lib.check(self.L.gqpBatchSolve(self.h, int(bool(rescale))), "gqpBatchSolve")
        _lib.check(self.L.gqpBatchSync(self.h), "gqpBatchSync")

    def solution(self):
        x, y, info = np.empty((self.B, self.n)), np.empty((self.B, self.m)), np.empty((self.B, 6))
        _lib.check(self.L.gqpBatchGetSolution(self.h, _p(x), _p(y), _p(info)), "gqpBatchGetSolution")
        return x, y, dict(iter=info[:, 0].astype(int), status=info[:, 1].astype(int), pri_res=info[:, 2], dua_res=info[:, 3],
                          rho=info[:, 4], rho_updates=info[:, 5].astype(int))

    def last_kernel_ms(self):
        return float(self.L.gqpBatchLastKernelMs(self.h))

    def lti_rollout(self, nsteps, Ad, Bd, x0, u_off, log_u=False, rescale_first=False):
        """planar/mpc_osqp.py:131-140 on the device: solve, x0 <- Ad x0 + Bd u0*, l[:nx] = u[:nx] = -x0, nsteps times"""
        nx, nu = Bd.shape
        x0 = _d(x0, (self.B, nx)).copy()
        ul = np.empty((nsteps, self.B, nu)) if log_u else None
        its = np.zeros(1)
        _lib.check(self.L.gqpLtiRollout(self.h, int(nsteps), nx, nu, int(u_off), _p(_d(Ad)), _p(_d(Bd)), _p(x0), _p(ul), _p(its), int(bool(rescale_first))),
                   "gqpLtiRollout")
        return x0, ul, float(its[0])


def umg_sizes(N, withWL=False):
    L = _bind()
    nx, nc, nnz = C.c_int(), C.c_int(), C.c_int()
    L.umgSizes(int(N), int(bool(withWL)), C.byref(nx), C.byref(nc), C.byref(nnz))
    return nx.value, nc.value, nnz.value


def umg_pattern(N, withWL=False):
    L = _bind()
    nx, nc, nnz = umg_sizes(N, withWL)
    rp, ci = np.zeros(nc + 1, np.int32), np.zeros(nnz, np.int32)
    L.umgPattern(int(N), int(bool(withWL)), rp.ctypes.data_as(IP), ci.ctypes.data_as(IP))
    return rp, ci


class BatchUprightMPC2G:
    """B UprightMPC2 controllers of horizon N (template_controllers.py:169-258), optionally fused with the WLQP
    (uprightmpc2_withWLQP.py); converged float64 solves like the reference's Python path."""

    def __init__(self, nenv, N=3, withWL=False, device=0, **kw):
        self.L = _bind()
        self.B, self.N, self.withWL = int(nenv), int(N), bool(withWL)
        p = UmgParams()
        self.L.umgDefaults(C.byref(p), self.N, int(self.withWL))
        for k, v in kw.items():
            if not hasattr(p, k):
                raise TypeError("unknown parameter %r" % k)
            cur = getattr(p, k)
            if hasattr(cur, "__len__"):
                for i, x in enumerate(np.asarray(v, float).ravel()):
                    cur[i] = x
            else:
                setattr(p, k, v)
        self.params = p
        self.nx, self.nc, self.nnz = umg_sizes(self.N, self.withWL)
        self.h = C.c_void_p()
        _lib.check(self.L.umgCreate(C.byref(self.h), self.B, int(device), C.byref(p)), "umgCreate")

    def close(self):
        if getattr(self, "h", None):
            self.L.umgDestroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reset(self):
        _lib.check(self.L.umgReset(self.h), "umgReset")

    def set_state(self, T0=None, u0=None):
        T0 = None if T0 is None else _d(T0, (self.B,))
        u0 = None if u0 is None else _d(u0, (self.B, 4))
        _lib.check(self.L.umgSetState(self.h, _p(T0), _p(u0)), "umgSetState")

    def update(self, p0, R0, dq0, pdes, dpdes, sdes, actualT0=None):
        """R0 (B,3,3) numpy row-major; returns uquad (B,3), accdes (B,6), u_wl (B,4) (zeros without the WLQP)"""
        B = self.B
        Rc = _d(np.asarray(R0, float).reshape(B, 3, 3).transpose(0, 2, 1).reshape(B, 9))
        uq, acc, uwl = np.empty((B, 3)), np.empty((B, 6)), np.empty((B, 4))
        aT = None if actualT0 is None else _d(actualT0, (B,))
        _lib.check(self.L.umgUpdate(self.h, _p(uq), _p(acc), _p(uwl), _p(_d(p0, (B, 3))), _p(Rc), _p(_d(dq0, (B, 6))),
                                    _p(_d(pdes, (B, 3))), _p(_d(dpdes, (B, 3))), _p(_d(sdes, (B, 3))), _p(aT)), "umgUpdate")
        return uq, acc, uwl

    def qp(self):
        B = self.B
        P, Av, q = np.empty((B, self.nx, self.nx)), np.empty((B, self.nnz)), np.empty((B, self.nx))
        l, u, x, y, info = np.empty((B, self.nc)), np.empty((B, self.nc)), np.empty((B, self.nx)), np.empty((B, self.nc)), np.empty((B, 6))
        _lib.check(self.L.umgGetQP(self.h, _p(P), _p(Av), _p(q), _p(l), _p(u), _p(x), _p(y), _p(info)), "umgGetQP")
        return dict(P=P, Av=Av, q=q, l=l, u=u, x=x, y=y, iter=info[:, 0].astype(int), status=info[:, 1].astype(int),
                    pri_res=info[:, 2], dua_res=info[:, 3], rho=info[:, 4])

    def last_solve_ms(self):
        return float(self.L.umgLastSolveMs(self.h))


class UprightMPC2WithWLQP:
    """template/uprightmpc2_withWLQP.py:218-324, one controller on the GPU (same constructor arguments and methods)."""

    def __init__(self, N, dt, g, TtoWmax, ws, wds, wpr, wpf, wvr, wvf, wthrust, wmom, umin, umax, dumax, mb, Ib, Qw, controlRate,
                 device=0, **kw):
        self.N, self.dt, self.g, self.Tmax = N, dt, g, TtoWmax * g
        self._b = BatchUprightMPC2G(1, N=N, withWL=True, device=device, dt=dt, g=g, TtoWmax=TtoWmax, ws=ws, wds=wds, wpr=wpr, wpf=wpf,
                                    wvr=wvr, wvf=wvf, wthrust=wthrust, wmom=wmom, umin=umin, umax=umax, dumax=dumax, mb=mb,
                                    Ib=np.asarray(Ib, float).ravel(), Qw=Qw, controlRate=controlRate, **kw)
        self.T0 = 0.0
        self.u0 = np.zeros(4)
        self.prevsol = np.zeros(self._b.nx)

    def update2(self, p0, R0, dq0, pdes, dpdes, sdes):
        uq, acc, uwl = self._b.update(np.asarray(p0, float)[None], np.asarray(R0, float)[None], np.asarray(dq0, float)[None],
                                      np.asarray(pdes, float)[None], np.asarray(dpdes, float)[None], np.asarray(sdes, float)[None])
        self.prevsol = self._b.qp()["x"][0]
        self.T0, self.u0, self._acc = float(uq[0, 0]), uwl[0].copy(), acc[0].copy()
        return uq[0].copy()

    def getAccDes(self, R0, dq0):
        return self._acc.copy()

    def update(self, p0, R0, dq0, pdes, dpdes, sdes):
        u = self.update2(p0, R0, dq0, pdes, dpdes, sdes)
        return u, self.getAccDes(R0, dq0)

    def close(self):
        self._b.close()


def mpc_osqp_problem(N=5):
    """The QP of planar/mpc_osqp.py:29-115 (stock OSQP quadcopter MPC: nx = 12, nu = 4): returns dict(P, A, q, l, u, Ad, Bd,
    nx, nu, N, x0, xr) with dense P (n, n) and A (m, n)."""
    Ad = np.array([
        [1., 0., 0., 0., 0., 0., 0.1, 0., 0., 0., 0., 0.], [0., 1., 0., 0., 0., 0., 0., 0.1, 0., 0., 0., 0.],
        [0., 0., 1., 0., 0., 0., 0., 0., 0.1, 0., 0., 0.], [0.0488, 0., 0., 1., 0., 0., 0.0016, 0., 0., 0.0992, 0., 0.],
        [0., -0.0488, 0., 0., 1., 0., 0., -0.0016, 0., 0., 0.0992, 0.], [0., 0., 0., 0., 0., 1., 0., 0., 0., 0., 0., 0.0992],
        [0., 0., 0., 0., 0., 0., 1., 0., 0., 0., 0., 0.], [0., 0., 0., 0., 0., 0., 0., 1., 0., 0., 0., 0.],
        [0., 0., 0., 0., 0., 0., 0., 0., 1., 0., 0., 0.], [0.9734, 0., 0., 0., 0., 0., 0.0488, 0., 0., 0.9846, 0., 0.],
        [0., -0.9734, 0., 0., 0., 0., 0., -0.0488, 0., 0., 0.9846, 0.], [0., 0., 0., 0., 0., 0., 0., 0., 0., 0., 0., 0.9846]])
    Bd = np.array([
        [0., -0.0726, 0., 0.0726], [-0.0726, 0., 0.0726, 0.], [-0.0152, 0.0152, -0.0152, 0.0152], [-0., -0.0006, -0., 0.0006],
        [0.0006, 0., -0.0006, 0.0000], [0.0106, 0.0106, 0.0106, 0.0106], [0, -1.4512, 0., 1.4512], [-1.4512, 0., 1.4512, 0.],
        [-0.3049, 0.3049, -0.3049, 0.3049], [-0., -0.0236, 0., 0.0236], [0.0236, 0., -0.0236, 0.], [0.2107, 0.2107, 0.2107, 0.2107]])
    nx, nu = Bd.shape
    u0 = 10.5916
    umin = np.array([9.6, 9.6, 9.6, 9.6]) - u0
    umax = np.array([13., 13., 13., 13.]) - u0
    xmin = np.array([-np.pi / 6, -np.pi / 6, -np.inf, -np.inf, -np.inf, -1., -np.inf, -np.inf, -np.inf, -np.inf, -np.inf, -np.inf])
    xmax = np.array([np.pi / 6, np.pi / 6, np.inf, np.inf, np.inf, np.inf, np.inf, np.inf, np.inf, np.inf, np.inf, np.inf])
    Q = np.diag([10., 10., 10., 10., 10., 10., 0., 0., 0., 5., 5., 5.])
    QN, R = Q, 0.1 * np.eye(4)
    x0, xr = np.zeros(12), np.array([1., 1., 1., 0., 0., 0., 0., 0., 0., 0., 0., 0.])
    n = (N + 1) * nx + N * nu
    P = np.zeros((n, n))
    for k in range(N):
        P[k * nx:(k + 1) * nx, k * nx:(k + 1) * nx] = Q
    P[N * nx:(N + 1) * nx, N * nx:(N + 1) * nx] = QN
    for k in range(N):
        o = (N + 1) * nx + k * nu
        P[o:o + nu, o:o + nu] = R
    q = np.hstack([np.kron(np.ones(N), -Q @ xr), -QN @ xr, np.zeros(N * nu)])
    Ax = np.kron(np.eye(N + 1), -np.eye(nx)) + np.kron(np.eye(N + 1, k=-1), Ad)
    Bu = np.kron(np.vstack([np.zeros((1, N)), np.eye(N)]), Bd)
    Aeq = np.hstack([Ax, Bu])
    leq = np.hstack([-x0, np.zeros(N * nx)])
    A = np.vstack([Aeq, np.eye(n)])
    l = np.hstack([leq, np.kron(np.ones(N + 1), xmin), np.kron(np.ones(N), umin)])
    u = np.hstack([leq, np.kron(np.ones(N + 1), xmax), np.kron(np.ones(N), umax)])
    return dict(P=P, A=A, q=q, l=l, u=u, Ad=Ad, Bd=Bd, nx=nx, nu=nu, N=N, x0=x0, xr=xr, n=n, m=A.shape[0])
