"""oracle/binding.py -- TEST INFRASTRUCTURE ONLY.

ctypes loaders for (a) our CPU restatement (oracle/liboracle_f32.so, _f64.so) and
(b) the reference's own C code compiled by oracle/Makefile (oracle/_ref/*.so).
Imported only by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs.  The product package never imports this module.
"""
import ctypes as C
import os
import shutil
import subprocess
import tempfile
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")
dbl = C.c_double
DP = C.POINTER(C.c_double)

# createMPC defaults, template_controllers.py:260-280 (umpcInit argument order)
UMPC_DEFAULTS = dict(dt=5.0, g=9.81e-3, TtoWmax=2.0, ws=1e1, wds=1e3, wpr=1.0, wpf=5.0,
                     wvr=1e3, wvf=2e3, wthrust=1e-1, wmom=1e-2, Ib=(3333.0, 3333.0, 1000.0), maxIter=50)
ARG_ORDER = ("dt", "g", "TtoWmax", "ws", "wds", "wpr", "wpf", "wvr", "wvf", "wthrust", "wmom")

TASK_HELIX, TASK_STRAIGHT, TASK_FLIP, TASK_PERCH = 0, 1, 2, 3


def build(ref=True):
    """(re)build the oracle libs; the reference libs only if /root/reference exists."""
    subprocess.run(["make", "-s", "-C", HERE, "oracle"] + (["ref"] if ref else []), check=True,
                   stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)


def have_ref():
    return os.path.exists(os.path.join(REF_DIR, "libref_umpc_f32.so"))


def _d(a, n=None):
    a = np.ascontiguousarray(a, dtype=np.float64).ravel()
    assert n is None or a.size == n, (a.size, n)
    return a


def _p(a):
    return a.ctypes.data_as(DP)


def _private_copy(path):
    """dlopen a private copy so each instance gets its own globals."""
    fd, tmp = tempfile.mkstemp(suffix=".so")
    os.close(fd)
    shutil.copyfile(path, tmp)
    lib = C.CDLL(tmp)
    os.unlink(tmp)
    return lib


def umpc_args(**kw):
    p = dict(UMPC_DEFAULTS)
    p.update(kw)
    return p


class OracleUMPC:
    """Restated controller; flavour 'f32' (C path arithmetic) or 'f64'."""

    def __init__(self, flavour="f32", **kw):
        self.lib = C.CDLL(os.path.join(HERE, "liboracle_%s.so" % flavour))
        L = self.lib
        L.orc_umpc_get_T0.restype = dbl
        self.p = umpc_args(**kw)
        self.par = C.create_string_buffer(L.orc_umpc_sizeof_par())
        self.st = C.create_string_buffer(L.orc_umpc_sizeof_state())
        a = _d([self.p[k] for k in ARG_ORDER])
        L.orc_umpc_init(self.par, self.st, _p(a), _p(_d(self.p["Ib"])), int(self.p["maxIter"]))

    def set_converged(self, max_iter=4000, check=25, adapt=100, eps=1e-4):
        self.lib.orc_umpc_set_solver(self.par, max_iter, check, adapt, dbl(eps), dbl(eps))

    def update(self, p0, R0, dq0, pdes, dpdes, sdes, actualT0=-1.0):
        """R0 is a (3,3) matrix (row-major numpy), like the pybind module takes."""
        uq, acc = np.zeros(3), np.zeros(6)
        Rc = _d(np.asarray(R0, dtype=np.float64).T, 9)  # column-major
        self.lib.orc_umpc_update(self.par, self.st, _p(uq), _p(acc), _p(_d(p0, 3)), _p(Rc), _p(_d(dq0, 6)),
                                 _p(_d(pdes, 3)), _p(_d(dpdes, 3)), _p(_d(sdes, 3)), dbl(actualT0))
        return uq, acc

    def qpdata(self):
        l, u, q, Px, Ax = np.zeros(39), np.zeros(39), np.zeros(45), np.zeros(45), np.zeros(48)
        self.lib.orc_umpc_get_qpdata(self.st, _p(l), _p(u), _p(q), _p(Px), _p(Ax))
        return l, u, q, Px, Ax

    def solution(self):
        x, y, info = np.zeros(45), np.zeros(39), np.zeros(4)
        self.lib.orc_umpc_get_solution(self.st, _p(x), _p(y), _p(info))
        return x, y, dict(iter=int(info[0]), status=int(info[1]), pri_res=info[2], dua_res=info[3])

    def iterates(self):
        x, z, y, D, E, c = np.zeros(45), np.zeros(39), np.zeros(39), np.zeros(45), np.zeros(39), np.zeros(1)
        self.lib.orc_umpc_get_iterates(self.st, _p(x), _p(z), _p(y), _p(D), _p(E), _p(c))
        return x, z, y, D, E, c[0]

    def ctype(self):
        ct = np.zeros(39, dtype=np.int32)
        self.lib.orc_umpc_get_ctype(self.st, ct.ctypes.data_as(C.POINTER(C.c_int)))
        return ct

    def pattern(self):
        row, col, idx = np.zeros(111, np.int32), np.zeros(111, np.int32), np.zeros(48, np.int32)
        ip = C.POINTER(C.c_int)
        ok = self.lib.orc_umpc_get_pattern(row.ctypes.data_as(ip), col.ctypes.data_as(ip), idx.ctypes.data_as(ip))
        return ok, row, col, idx

    @property
    def T0(self):
        return self.lib.orc_umpc_get_T0(self.st)


class RefUMPC:
    """The reference's own umpcInit/umpcUpdate (float32, vendored OSQP). One private
    dlopen per instance because the reference keeps solver state in globals."""

    def __init__(self, **kw):
        self.lib = _private_copy(os.path.join(REF_DIR, "libref_umpc_f32.so"))
        self.p = umpc_args(**kw)
        self.up = (C.c_float * 1024)()  # >= sizeof(UprightMPC_t)
        f = C.c_float
        Ib = (f * 3)(*self.p["Ib"])
        self.lib.umpcInit(self.up, *[f(self.p[k]) for k in ARG_ORDER], Ib, int(self.p["maxIter"]))

    @staticmethod
    def _f(a, n):
        a = np.ascontiguousarray(a, dtype=np.float32).ravel()
        assert a.size == n
        return a

    def update(self, p0, R0, dq0, pdes, dpdes, sdes, actualT0=-1.0):
        FP = C.POINTER(C.c_float)
        uq, acc = np.zeros(3, np.float32), np.zeros(6, np.float32)
        args = [self._f(p0, 3), self._f(np.asarray(R0, dtype=np.float64).T, 9), self._f(dq0, 6),
                self._f(pdes, 3), self._f(dpdes, 3), self._f(sdes, 3)]
        self.ret = self.lib.umpcUpdate(self.up, uq.ctypes.data_as(FP), acc.ctypes.data_as(FP),
                                       *[a.ctypes.data_as(FP) for a in args], C.c_float(actualT0))
        return uq.astype(np.float64), acc.astype(np.float64)

    def qpdata(self):
        """(l,u,q,Px,Ax_data,Ax_idx) read out of UprightMPC_t (uprightmpc2.h:27-43)."""
        raw = np.frombuffer(self.up, dtype=np.float32)
        o = 3 + 27 + 6 + 18  # dt,g,Tmax | Qyr,Qyf,Qdyr,Qdyf,R | smin,smax | e3h,e3hIbi
        l, u, q, Px, Ax = raw[o:o + 39], raw[o + 39:o + 78], raw[o + 78:o + 123], raw[o + 123:o + 168], raw[o + 168:o + 216]
        idx = np.frombuffer(self.up, dtype=np.int32)[o + 216:o + 264]
        return [a.astype(np.float64) for a in (l, u, q, Px, Ax)] + [idx.copy()]

    def solution(self):
        x, y, info = np.zeros(45), np.zeros(39), np.zeros(5)
        self.lib.ref_get_solution(_p(x), _p(y))
        self.lib.ref_get_info(_p(info))
        return x, y, dict(iter=int(info[0]), status=int(info[1]), pri_res=info[2], dua_res=info[3])

    def iterates(self):
        x, z, y, D, E, c = np.zeros(45), np.zeros(39), np.zeros(39), np.zeros(45), np.zeros(39), np.zeros(1)
        self.lib.ref_get_iterates(_p(x), _p(z), _p(y), _p(D), _p(E), _p(c))
        return x, z, y, D, E, c[0]


class RefQP64:
    """Flavour B: the vendored OSQP built in float64, run to convergence with the
    Python path's settings (template_controllers.py:190-191)."""

    def __init__(self, max_iter=4000, check=25, adapt=100, eps=1e-4):
        self.lib = _private_copy(os.path.join(REF_DIR, "libref_umpc_f64.so"))
        assert self.lib.ref_sizeof_cfloat() == 8
        self.lib.ref_set_iters(max_iter, check)
        self.lib.ref_set_eps(dbl(eps), dbl(eps))
        self.lib.ref_set_adaptive_rho_interval(adapt)

    def solve(self, l, u, q, Px, Ax, Ax_idx):
        idx = np.ascontiguousarray(Ax_idx, dtype=np.int32)
        self.lib.ref_qp_update_solve(_p(_d(l, 39)), _p(_d(u, 39)), _p(_d(q, 45)), _p(_d(Px, 45)), 45,
                                     _p(_d(Ax, 48)), idx.ctypes.data_as(C.POINTER(C.c_int)), 48)
        x, y, info = np.zeros(45), np.zeros(39), np.zeros(5)
        self.lib.ref_get_solution(_p(x), _p(y))
        self.lib.ref_get_info(_p(info))
        return x, y, dict(iter=int(info[0]), status=int(info[1]), pri_res=info[2], dua_res=info[3])


# WaypointHover constants, robobee_test_controllers.py:71,76
POPTS_WAYPOINT = None  # filled from tests/golden/constants.json by callers that need it
WL_DEFAULTS = dict(u0=(140.0, 0.0, 0.0, 0.0), umin=(90.0, -0.5, -0.2, -0.1), umax=(240.0, 0.5, 0.2, 0.1),
                   dumax=(5e3, 10.0, 10.0, 10.0), Qw=(1.0, 1.0, 1.0, 0.1, 0.1, 0.1), controlRate=1000.0)


class OracleWL:
    """Restated WLCon (mode 'con') / true WLQP (mode 'qp')."""

    def __init__(self, popts, flavour="f32", check_termination=None, **kw):
        self.lib = C.CDLL(os.path.join(HERE, "liboracle_%s.so" % flavour))
        p = dict(WL_DEFAULTS)
        p.update(kw)
        self.wl = C.create_string_buffer(self.lib.orc_wl_sizeof())
        self.lib.orc_wl_init(self.wl, _p(_d(p["u0"], 4)), _p(_d(p["umin"], 4)), _p(_d(p["umax"], 4)),
                             _p(_d(p["dumax"], 4)), _p(_d(p["Qw"], 6)), dbl(p["controlRate"]), _p(_d(popts, 90)))
        if check_termination is not None:
            self.lib.orc_wl_set_check_termination(self.wl, int(check_termination))

    def update(self, h0, pdotdes, mode="con"):
        u1, w0 = np.zeros(4), np.zeros(6)
        fn = self.lib.orc_wlcon_update if mode == "con" else self.lib.orc_wlqp_update
        fn(self.wl, _p(u1), _p(w0), _p(_d(h0, 6)), _p(_d(pdotdes, 6)))
        return u1, w0

    def info(self):
        i4 = np.zeros(4)
        self.lib.orc_wl_get_info(self.wl, _p(i4))
        return dict(iter=int(i4[0]), status=int(i4[1]), pri_res=i4[2], dua_res=i4[3])


class RefWLCon:
    """Reference wlConInit/wlConUpdate (uprightmpc2/funapprox.c), float32."""

    def __init__(self, popts, lib=None, **kw):
        self.lib = lib or _private_copy(os.path.join(REF_DIR, "libref_umpc_f32.so"))
        p = dict(WL_DEFAULTS)
        p.update(kw)
        self.wl = (C.c_float * 512)()
        f = lambda a: (C.c_float * len(a))(*a)
        self.lib.wlConInit(self.wl, f(p["u0"]), f(p["umin"]), f(p["umax"]), f(p["dumax"]), f(p["Qw"]),
                           C.c_float(p["controlRate"]), f(list(popts)))

    def update(self, h0, pdotdes):
        FP = C.POINTER(C.c_float)
        u1, w0 = np.zeros(4, np.float32), np.zeros(6, np.float32)
        h, pd = np.asarray(h0, np.float32), np.asarray(pdotdes, np.float32)
        self.lib.wlConUpdate(self.wl, u1.ctypes.data_as(FP), w0.ctypes.data_as(FP), h.ctypes.data_as(FP), pd.ctypes.data_as(FP))
        return u1.astype(np.float64), w0.astype(np.float64)


class RefWLQP:
    """Reference wlControllerUpdate (wlcontroller/c), float32; global state => private dlopen."""

    def __init__(self, popts, u0=WL_DEFAULTS["u0"], controlRate=1000.0):
        self.lib = _private_copy(os.path.join(REF_DIR, "libref_wlqp_f32.so"))
        self.popts = np.asarray(popts, np.float32)
        self.u0 = np.asarray(u0, np.float32)
        self.rate = controlRate

    def update(self, h0, pdotdes):
        FP = C.POINTER(C.c_float)
        u = np.zeros(4, np.float32)
        h, pd = np.asarray(h0, np.float32), np.asarray(pdotdes, np.float32)
        self.lib.wlControllerUpdate(u.ctypes.data_as(FP), self.u0.ctypes.data_as(FP), h.ctypes.data_as(FP),
                                    pd.ctypes.data_as(FP), self.popts.ctypes.data_as(FP), C.c_float(self.rate))
        return u.astype(np.float64)

    def info(self):
        i5 = np.zeros(5)
        self.lib.ref_get_info(_p(i5))
        return dict(iter=int(i5[0]), status=int(i5[1]), pri_res=i5[2], dua_res=i5[3])


def rollout_cfg(tend=500.0, dtsim=0.2, hlInterval=None, taulim=100.0, tpert=None,
                Ib=(3333.0, 3333.0, 1000.0), g=9.81e-3, mb=100.0, intInitialPos=False):
    """intInitialPos: controlTest's integer-dtype initialPos (uprightmpc2.py:94-104) truncates pdes toward zero."""
    return _d([tend, dtsim, -1.0 if hlInterval is None else hlInterval, taulim,
               -1.0 if tpert is None else tpert, Ib[0], Ib[1], Ib[2], g, mb, 1.0 if intInitialPos else 0.0])


def control_test_ic():
    """IC of controlTest's default branch, uprightmpc2.py:99-103 (scipy 'xyz' extrinsic)."""
    cx, sx, cy, sy = np.cos(0.5), np.sin(0.5), np.cos(-0.5), np.sin(-0.5)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    return np.zeros(3), Ry @ Rx, np.array([0.1, 0, 0, 0, 0, 0.0])


def rollout(lib, par, ctrl, ic, cfg, task=TASK_HELIX, tp=(0.0, 0.0, 0.1, 0.0), wl=None, wlmode=0, log=True):
    """ic = (p, R(3x3 row-major numpy), dq).  Returns dict(y,u,pdes,final,err,eff,nsolve)."""
    p, R, dq = ic
    icv = _d(np.r_[np.asarray(p, float), np.asarray(R, float).T.ravel(), np.asarray(dq, float)], 18)
    lib.orc_rollout_nt.argtypes = [dbl, dbl]
    nt = lib.orc_rollout_nt(dbl(cfg[0]), dbl(cfg[1]))
    y = np.zeros((nt, 12)); u = np.zeros((nt, 3)); pd = np.zeros((nt, 3)); fin = np.zeros(18); met = np.zeros(4)
    nul = C.cast(None, DP)
    lib.orc_rollout(par, ctrl, wl, int(wlmode), _p(icv), _p(cfg), int(task), _p(_d(tp, 4)),
                    _p(y) if log else nul, _p(u) if log else nul, _p(pd) if log else nul, _p(fin), _p(met))
    return dict(y=y, u=u, pdes=pd, final=fin, err=met[0], eff=met[1], nsolve=int(met[2]), u4sum=met[3])


# ---- timed CPU baseline (bench.py cpu_baseline / --impl reference) ---------------------------------
def _cpu_worker(args):
    kind, nenv, nsteps, ics, seed = args
    import time
    if kind == "reference":
        lib = C.CDLL(os.path.join(REF_DIR, "libref_umpc_f32_fast.so"))
    else:
        lib = C.CDLL(os.path.join(HERE, "liboracle_f32.so"))
    lib.orc_batch_steps.restype = C.c_long
    lib.orc_batch_steps.argtypes = [C.c_int, C.c_int, DP, DP, DP, C.c_int, dbl, dbl]
    a = _d([UMPC_DEFAULTS[k] for k in ARG_ORDER])
    Ib = _d(UMPC_DEFAULTS["Ib"])
    ics = np.ascontiguousarray(ics, dtype=np.float64)
    t0 = time.perf_counter()
    n = lib.orc_batch_steps(int(nenv), int(nsteps), _p(ics), _p(a), _p(Ib), int(UMPC_DEFAULTS["maxIter"]), 0.2, 100.0)
    return n, time.perf_counter() - t0, ics


def cpu_baseline_kind():
    return "reference" if os.path.exists(os.path.join(REF_DIR, "libref_umpc_f32_fast.so")) else "port"


def cpu_env_steps(ics18, nsteps, procs, per_worker=False):
    """Advance the envs in ics18 (n x 18: p, R col-major, dq) nsteps closed-loop steps (hover, MPC every
    step) on `procs` host processes (one per core: the reference's OSQP workspace is a process global,
    template/uprightmpc2/workspace.c:2620, so threads cannot share it).  Returns (env_steps, seconds, kind)
    with seconds = wall time of the slowest worker."""
    import multiprocessing as mp
    import time
    kind = cpu_baseline_kind()
    ics18 = np.asarray(ics18, dtype=np.float64).reshape(-1, 18)
    chunks = [c for c in np.array_split(ics18, procs) if len(c)]
    jobs = [(kind, len(c), nsteps, c, i) for i, c in enumerate(chunks)]
    t0 = time.perf_counter()
    if len(jobs) == 1:
        res = [_cpu_worker(jobs[0])]
    else:
        with mp.get_context("fork").Pool(len(jobs)) as pool:
            res = pool.map(_cpu_worker, jobs)
    wall = time.perf_counter() - t0
    done = sum(r[0] for r in res)
    if per_worker:  # + env-steps/s of every worker (one per core)
        return done, max(r[1] for r in res), wall, kind, [r[0] / r[1] for r in res]
    return done, max(r[1] for r in res), wall, kind


# ---- planar dir-tran MPC (config 5): oracle/dirtran_oracle.c, oracle/_ref/libref_planar_f32.so -------------------
def dt_settings(eps=1e-5, rho=0.1, sigma=1e-6, alpha=1.6, max_iter=4000, check=25, adapt=25, tol=5.0):
    return np.array([eps, eps, rho, sigma, alpha, max_iter, check, adapt, tol], dtype=np.float64)


def _dt_lib(flavour):
    L = _private_copy(os.path.join(HERE, "liboracle_dt_%s.so" % flavour))
    L.dto_create.restype = C.c_void_p
    L.dto_create.argtypes = [C.c_int, C.c_double, DP, DP, DP]
    L.dto_destroy.argtypes = [C.c_void_p]
    L.dto_set_ctrl.argtypes = [C.c_void_p, DP]
    L.dto_update.argtypes = [C.c_void_p, DP, DP, DP]
    L.dto_get.argtypes = [C.c_void_p, DP, DP, DP]
    L.dto_qpdata.argtypes = [C.c_void_p, DP, DP, DP, DP, DP]
    L.dto_plant.argtypes = [C.c_void_p, DP, DP, C.c_double, C.c_int]
    L.dto_rollout.argtypes = [C.c_void_p, DP, DP, C.c_int, C.c_double, C.c_int, DP, DP, DP]
    L.dto_raw_solve.argtypes = [C.c_int, C.c_int, DP, DP, DP, DP, DP, DP, DP, DP, DP]
    return L


class OracleDirTran:
    """CPU restatement of LTVMPC(PlanarThrustStrokeDev(rescale=True), N, wx, wu, scaling=0, ...) + the plant."""
    HOVER_WX = np.array([100, 100, 1, 10, 10, 0.1]) / 0.03   # mpc_thrust_strokedev.py:30,118
    HOVER_WU = np.array([0.01, 0.01])

    def __init__(self, N=5, dt=0.03, wx=None, wu=None, flavour="f64", **st):
        self.L = _dt_lib(flavour)
        self.N = N
        self.n = (N + 1) * 6 + N * 2
        self.m = self.n + (N + 1) * 6
        wx = _d(self.HOVER_WX if wx is None else wx); wu = _d(self.HOVER_WU if wu is None else wu)
        self.st = dt_settings(**st)
        self.h = self.L.dto_create(N, dt, _p(wx), _p(wu), _p(self.st))
        assert self.h, "N too large for the oracle's caps"

    def __del__(self):
        try:
            self.L.dto_destroy(self.h)
        except Exception:
            pass

    def set_ctrl(self, c):
        self.L.dto_set_ctrl(self.h, _p(_d(c)))

    def update(self, x0, xr):
        u = np.zeros(2)
        st = self.L.dto_update(self.h, _p(_d(x0)), _p(_d(xr)), _p(u))
        return u, st

    def solution(self):
        x = np.zeros(self.n); y = np.zeros(self.m); info = np.zeros(6)
        self.L.dto_get(self.h, _p(x), _p(y), _p(info))
        return x, y, info

    def qpdata(self):
        Pd = np.zeros(self.n); A = np.zeros((self.m, self.n)); q = np.zeros(self.n); l = np.zeros(self.m); u = np.zeros(self.m)
        self.L.dto_qpdata(self.h, _p(Pd), _p(A), _p(q), _p(l), _p(u))
        return Pd, A, q, l, u

    def plant(self, x, u, sim_dt=0.0005, nsub=60):
        x = _d(x).copy()
        self.L.dto_plant(self.h, _p(x), _p(_d(u)), sim_dt, nsub)
        return x

    def rollout(self, x, xr, nsteps, sim_dt=0.0005, nsub=60):
        x = _d(x).copy()
        X = np.zeros((nsteps, 6)); U = np.zeros((nsteps, 2)); it = np.zeros(nsteps)
        bad = self.L.dto_rollout(self.h, _p(x), _p(_d(xr)), nsteps, sim_dt, nsub, _p(X), _p(U), _p(it))
        return X, U, it, bad


def dt_raw_solve(P, A, q, l, u, st, flavour="f64"):
    L = _dt_lib(flavour)
    m, n = A.shape
    x = np.zeros(n); y = np.zeros(m); info = np.zeros(6)
    L.dto_raw_solve(n, m, _p(_d(P)), _p(_d(A)), _p(_d(q)), _p(_d(l)), _p(_d(u)), _p(_d(st)), _p(x), _p(y), _p(info))
    return x, y, info


class RefPlanarQP:
    """The reference's generated planar dir-tran instance (planar/code), compiled in place."""

    def __init__(self):
        self.L = _private_copy(os.path.join(HERE, "_ref", "libref_planar_f32.so"))
        nm = (C.c_int * 2)()
        self.L.rp_dims(nm)
        self.n, self.m = nm[0], nm[1]

    def data(self):
        n, m = self.n, self.m
        P = np.zeros((n, n)); A = np.zeros((m, n)); q = np.zeros(n); l = np.zeros(m); u = np.zeros(m); rv = np.zeros(m); st = np.zeros(10)
        self.L.rp_data(_p(P), _p(A), _p(q), _p(l), _p(u), _p(rv), _p(st))
        return P, A, q, l, u, rv, st

    def cold_start(self):
        self.L.rp_cold_start()

    def update(self, q, l, u):
        return self.L.rp_update(_p(_d(q)), _p(_d(l)), _p(_d(u)))

    def set_max_iter(self, it):
        return self.L.rp_update_max_iter(int(it))

    def solve(self):
        x = np.zeros(self.n); y = np.zeros(self.m); info = np.zeros(4)
        self.L.rp_solve(_p(x), _p(y), _p(info))
        return x, y, info


def have_ref_planar():
    return os.path.exists(os.path.join(HERE, "_ref", "libref_planar_f32.so"))


# ---- generic OSQP object (oracle/gqp_oracle.c): checker of robobee3d_b200/csrc/gqp.cu -----------------------------------
class OracleQP:
    """setup once, then update(P, A, q, l, u) + solve like the reference's Python controllers call the osqp package
    (template_controllers.py:190-191,216-217; uprightmpc2_withWLQP.py:243-244,283-284; planar/mpc_osqp.py:104-140).
    Dense P (n, n), A (m, n).  P = A = None: only q / l / u change (no re-equilibration)."""

    def __init__(self, n, m, eps=1e-4, rho=0.1, sigma=1e-6, alpha=1.6, max_iter=4000, check=25, adapt=100, tol=5.0, scaling=10,
                 infty=1e30, flavour="f64"):
        self.L = _private_copy(os.path.join(HERE, "liboracle_dt_%s.so" % flavour))
        self.L.gqo_create.restype = C.c_void_p
        self.L.gqo_create.argtypes = [C.c_int, C.c_int, DP]
        self.L.gqo_destroy.argtypes = [C.c_void_p]
        self.L.gqo_update_solve.argtypes = [C.c_void_p, DP, DP, DP, DP, DP, DP, DP, DP]
        self.n, self.m = n, m
        st = _d([eps, eps, rho, sigma, alpha, max_iter, check, adapt, tol, scaling, infty])
        self.h = self.L.gqo_create(n, m, _p(st))
        assert self.h, "QP larger than the oracle's caps"

    def __del__(self):
        try:
            self.L.gqo_destroy(self.h)
        except Exception:
            pass

    def update_solve(self, P, A, q, l, u):
        x, y, info = np.zeros(self.n), np.zeros(self.m), np.zeros(6)
        nul = C.cast(None, DP)
        l = np.clip(_d(l, self.m), -1e300, 1e300); u = np.clip(_d(u, self.m), -1e300, 1e300)
        self.L.gqo_update_solve(self.h, nul if P is None else _p(_d(P, self.n * self.n)), nul if A is None else _p(_d(A, self.m * self.n)),
                                _p(_d(q, self.n)), _p(l), _p(u), _p(x), _p(y), _p(info))
        return x, y, dict(iter=int(info[0]), status=int(info[1]), pri_res=info[2], dua_res=info[3], rho=info[4], rho_updates=int(info[5]))
