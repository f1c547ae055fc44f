"""Batched planar direct-transcription MPC on B200 (include/dirtran_b200.h) and drop-in mirrors of the reference's
Python classes for this path:

    LTVMPC(model, N, wx, wu, **settings).update(x0, xr, costMode=FINAL)   controlutils/py/mpc.py:10-91
    PlanarThrustStrokeDev(rescale=True)                                    planar/FlappingModels.py:27-129

No CPU fallback: everything computes in libumpc_b200.so on a CUDA device or raises.
"""
import ctypes as C

import numpy as np

from . import _lib

DP = C.POINTER(C.c_double)
NX, NU = 6, 2
FINAL, TRAJ = 0, 1  # ltvsystem.py:24-25

STATUS = {1: "solved", 2: "solved inaccurate", -2: "maximum iterations reached", -10: "unsolved"}


class DtParams(C.Structure):
    """include/dirtran_b200.h: DtParams"""
    _fields_ = [
        ("N", C.c_int), ("dt", C.c_double), ("wx", C.c_double * NX), ("wu", C.c_double * NU),
        ("eps_abs", C.c_double), ("eps_rel", C.c_double), ("rho", C.c_double), ("sigma", C.c_double), ("alpha", C.c_double),
        ("max_iter", C.c_int), ("check_termination", C.c_int), ("adaptive_rho_interval", C.c_int),
        ("adaptive_rho_tol", C.c_double),
        ("g", C.c_double), ("ibmb", C.c_double), ("umin", C.c_double * NU), ("umax", C.c_double * NU),
        ("xmin", C.c_double * NX), ("xmax", C.c_double * NX),
        ("sim_dt", C.c_double), ("nsub", C.c_int),
    ]


EXPORTS = ["dtBatchDefaults", "dtBatchCreate", "dtBatchDestroy", "dtBatchReset", "dtBatchSetState", "dtBatchGetState",
           "dtBatchUpdate", "dtBatchRollout", "dtBatchGetInfo", "dtBatchGetSolution", "dtBatchCounters",
           "dtBatchLastKernelMs", "dtBatchSync"]


def _bind(L):
    if getattr(L, "_dt_bound", False):
        return L
    L.dtBatchDefaults.argtypes = [C.POINTER(DtParams), C.c_int]; L.dtBatchDefaults.restype = None
    L.dtBatchCreate.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.POINTER(DtParams)]
    L.dtBatchDestroy.argtypes = [C.c_void_p]
    L.dtBatchReset.argtypes = [C.c_void_p]
    L.dtBatchSetState.argtypes = [C.c_void_p, DP, DP, DP]
    L.dtBatchGetState.argtypes = [C.c_void_p, DP, DP]
    L.dtBatchUpdate.argtypes = [C.c_void_p, DP, DP]
    L.dtBatchRollout.argtypes = [C.c_void_p, C.c_int]
    L.dtBatchGetInfo.argtypes = [C.c_void_p, DP]
    L.dtBatchGetSolution.argtypes = [C.c_void_p, DP, DP]
    L.dtBatchCounters.argtypes = [C.c_void_p, DP]
    L.dtBatchLastKernelMs.argtypes = [C.c_void_p]; L.dtBatchLastKernelMs.restype = C.c_float
    L.dtBatchSync.argtypes = [C.c_void_p]
    L.umpcLastError.restype = C.c_char_p
    L._dt_bound = True
    return L


def default_params(N=5, **kw):
    """the 'hover' parameter set of planar/mpc_thrust_strokedev.py:30 with OSQP's defaults; kw overrides fields"""
    L = _bind(_lib.load())
    p = DtParams()
    L.dtBatchDefaults(C.byref(p), N)
    for k, v in kw.items():
        cur = getattr(p, k)
        if hasattr(cur, "__len__"):
            for i, x in enumerate(v):
                cur[i] = x
        else:
            setattr(p, k, v)
    return p


def _d(a, shape):
    a = np.ascontiguousarray(np.asarray(a, dtype=np.float64))
    if a.shape != shape:
        a = np.ascontiguousarray(np.broadcast_to(a, shape))
    return a


def _p(a):
    return a.ctypes.data_as(DP) if a is not None else None


class BatchLTVMPC:
    """B independent LTVMPC controllers + plants on one GPU."""

    def __init__(self, nenv, params=None, device=0, **kw):
        self.L = _bind(_lib.load())
        self.par = params if params is not None else default_params(**kw)
        self.B = int(nenv)
        self.N = self.par.N
        self.n = (self.N + 1) * NX + self.N * NU
        self.m = self.n + (self.N + 1) * NX
        self.h = C.c_void_p()
        self._ck(self.L.dtBatchCreate(C.byref(self.h), self.B, device, C.byref(self.par)))

    def _ck(self, rc):
        if rc != 0:
            raise RuntimeError("libumpc_b200: %s (rc=%d)" % ((self.L.umpcLastError() or b"").decode(), rc))

    def close(self):
        if self.h:
            self.L.dtBatchDestroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reset(self):
        self._ck(self.L.dtBatchReset(self.h))

    def set_state(self, y=None, ctrl=None, xr=None):
        y = _d(y, (self.B, NX)) if y is not None else None
        ctrl = _d(ctrl, (self.B, NU)) if ctrl is not None else None
        xr = _d(xr, (self.B, NX)) if xr is not None else None
        self._ck(self.L.dtBatchSetState(self.h, _p(y), _p(ctrl), _p(xr)))

    def state(self):
        y = np.empty((self.B, NX)); c = np.empty((self.B, NU))
        self._ck(self.L.dtBatchGetState(self.h, _p(y), _p(c)))
        return y, c

    def update(self, x0):
        """one LTVMPC.update per env at the stored goal: x0 (B,6) -> u (B,2)"""
        x0 = _d(x0, (self.B, NX))
        u = np.empty((self.B, NU))
        self._ck(self.L.dtBatchUpdate(self.h, _p(x0), _p(u)))
        return u

    def rollout(self, nsteps, sync=True):
        self._ck(self.L.dtBatchRollout(self.h, int(nsteps)))
        if sync:
            self._ck(self.L.dtBatchSync(self.h))

    def info(self):
        a = np.empty((self.B, 6))
        self._ck(self.L.dtBatchGetInfo(self.h, _p(a)))
        return {"iter": a[:, 0].astype(int), "status": a[:, 1].astype(int), "pri_res": a[:, 2], "dua_res": a[:, 3],
                "rho": a[:, 4], "rho_updates": a[:, 5].astype(int)}

    def solution(self):
        x = np.empty((self.B, self.n)); y = np.empty((self.B, self.m))
        self._ck(self.L.dtBatchGetSolution(self.h, _p(x), _p(y)))
        return x, y

    def counters(self):
        a = np.empty(3)
        self._ck(self.L.dtBatchCounters(self.h, _p(a)))
        return {"iters": a[0], "solves": a[1], "unsolved": a[2]}

    def last_kernel_ms(self):
        return float(self.L.dtBatchLastKernelMs(self.h))


class PlanarThrustStrokeDev:
    """Host mirror of planar/FlappingModels.py:27-129 (constants and limits only; the dynamics run on the device)."""
    mb = 100e-6
    l = 12e-3
    g = 9.81
    ib = 1 / 12. * mb * l**2
    nx, nu = NX, NU
    dt = 0.005
    STROKE_EXTENT = 3e-3
    lscaled = 0.1

    def __init__(self, rescale=False):
        if not rescale:
            raise NotImplementedError("only rescale=True (what planar/mpc_thrust_strokedev.py:56 uses) is built")
        self.rescale = rescale
        self.ibmb = 1. / 12. * self.lscaled**2
        self.y0 = np.zeros(NX)
        self.u0 = np.zeros(NU)

    @property
    def limits(self):
        umin = np.array([-self.g, -0.05 * self.lscaled])
        umax = np.array([3 * self.g, 0.05 * self.lscaled])
        xmin = np.array([-np.inf, -np.inf, -10 * np.pi, -np.inf, -np.inf, -np.inf])
        return umin, umax, xmin, -xmin


class LTVMPC:
    """Drop-in for controlutils/py/mpc.py:10-91 with a PlanarThrustStrokeDev model: a batch of one on the GPU.

    settings: the OSQP keywords the reference passes (mpc_thrust_strokedev.py:124): eps_rel, eps_abs, max_iter, scaling
    (must be 0), polish (must be False), plus adaptive_rho_interval (default 25, see include/dirtran_b200.h)."""

    def __init__(self, model, N, wx, wu, kdamping=0, polyBlocks=None, **settings):
        if kdamping != 0 or polyBlocks is not None:
            raise NotImplementedError("kdamping / polyBlocks are not part of the built path")
        if settings.get("scaling", 0) != 0 or settings.get("polish", False):
            raise NotImplementedError("scaling=0, polish=False only (the reference's settings for this path)")
        umin, umax, xmin, xmax = model.limits
        self.model = model
        self.par = default_params(
            N, dt=model.dt, wx=list(wx), wu=list(wu), eps_abs=settings.get("eps_abs", 1e-3), eps_rel=settings.get("eps_rel", 1e-3),
            max_iter=min(int(settings.get("max_iter", 4000)), 2**31 - 1), rho=settings.get("rho", 0.1),
            sigma=settings.get("sigma", 1e-6), alpha=settings.get("alpha", 1.6),
            check_termination=settings.get("check_termination", 25),
            adaptive_rho_interval=settings.get("adaptive_rho_interval", 25) if settings.get("adaptive_rho", True) else 0,
            adaptive_rho_tol=settings.get("adaptive_rho_tolerance", 5.0),
            g=model.g, ibmb=model.ibmb, umin=list(umin), umax=list(umax), xmin=list(xmin), xmax=list(xmax))
        self.b = BatchLTVMPC(1, self.par)
        self.ctrl = np.zeros(NU)
        self.prevSol = None
        self.niter = np.nan

    def update(self, x0, xr, *args, u0=None, ur=None, trajMode=0, costMode=TRAJ, ugoalCost=False):
        if u0 is not None or ur is not None or trajMode != 0 or np.ndim(x0) != 1:
            raise NotImplementedError("single-point linearisation at the previous input only (mpc_thrust_strokedev.py:171)")
        self.b.set_state(ctrl=self.ctrl[None, :], xr=np.asarray(xr, float)[None, :])
        self.ctrl = self.b.update(np.asarray(x0, float)[None, :])[0]
        i = self.b.info()
        self.niter = int(i["iter"][0])
        if i["status"][0] != 1:
            raise ValueError(STATUS.get(int(i["status"][0]), str(i["status"][0])))  # ltvsystem.py:312-316
        self.prevSol = self.b.solution()[0][0]
        return self.ctrl
