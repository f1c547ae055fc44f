"""Drop-in for the reference's pybind module `uprightmpc2py`
(template/uprightmpc2/py/uprightmpc2py.cpp:30-80): same class names, constructor
arguments, method names and return shapes, backed by libumpc_b200.so on a B200.

    import robobee3d_b200.uprightmpc2py as uprightmpc2py        # instead of the pybind build
    up = uprightmpc2py.UprightMPC2C(dt, g, TtoWmax, ws, wds, wpr, wpf, wvr, wvf, wthrust, wmom, Ib, maxIter)
    uquad, accdes = up.update(p0, R0, dq0, pdes, dpdes, sdes, actualT0)

`actualT0` defaults to -1 (keep the internal T0) because template/uprightmpc2.py:139 calls
update with 6 arguments.
"""
import ctypes as C
import numpy as np

from . import _lib


def _fp(a):
    return a.ctypes.data_as(_lib.FP)


def _v(a, n):
    a = np.ascontiguousarray(a, dtype=np.float32).ravel()
    if a.size != n:
        raise ValueError("expected %d values, got %d" % (n, a.size))
    return a


class UprightMPC2C:
    def __init__(self, dt, g, TtoWmax, ws, wds, wpr, wpf, wvr, wvf, wthrust, wmom, Ib, maxIter):
        self._lib = _lib.load()
        self._up = _lib.UprightMPC_t()
        ib = _v(Ib, 3)
        self._lib.umpcInit(C.byref(self._up), dt, g, TtoWmax, ws, wds, wpr, wpf, wvr, wvf, wthrust, wmom, _fp(ib), int(maxIter))
        info = np.zeros(4, np.float32)
        self._lib.umpcGetInfo(C.byref(self._up), _fp(info))
        if np.isnan(info[0]):
            raise RuntimeError("UprightMPC2C: no CUDA device / init failed: %s" % self._lib.umpcLastError().decode())

    def update(self, p0, R0, dq0, pdes, dpdes, sdes, actualT0=-1.0):
        uquad, accdes = np.zeros(3, np.float32), np.zeros(6, np.float32)
        # numpy (3,3) is row-major; the C-ABI wants column-major like Eigen hands it over
        Rc = _v(np.asarray(R0, dtype=np.float32).reshape(3, 3).T, 9)
        rc = self._lib.umpcUpdate(C.byref(self._up), _fp(uquad), _fp(accdes), _fp(_v(p0, 3)), _fp(Rc), _fp(_v(dq0, 6)),
                                  _fp(_v(pdes, 3)), _fp(_v(dpdes, 3)), _fp(_v(sdes, 3)), float(actualT0))
        if rc != 0:
            raise RuntimeError("umpcUpdate failed (%d): %s" % (rc, self._lib.umpcLastError().decode()))
        return uquad, accdes

    def vectors(self):
        u = self._up
        return np.array(u.l, np.float32), np.array(u.u, np.float32), np.array(u.q, np.float32)

    def matrices(self):
        u = self._up
        return np.array(u.Px_data, np.float32), np.array(u.Ax_data, np.float32), np.array(u.Ax_idx, np.int32)

    # extras (the reference keeps these in its global OSQP workspace)
    def info(self):
        i = np.zeros(4, np.float32)
        self._lib.umpcGetInfo(C.byref(self._up), _fp(i))
        return dict(iter=int(i[0]), status=int(i[1]), pri_res=float(i[2]), dua_res=float(i[3]))

    def solution(self):
        x, y = np.zeros(45, np.float32), np.zeros(39, np.float32)
        self._lib.umpcGetSolution(C.byref(self._up), _fp(x), _fp(y))
        return x, y

    @property
    def T0(self):
        return float(self._up.T0)

    def __del__(self):
        try:
            self._lib.umpcFree(C.byref(self._up))
        except Exception:
            pass


class WLCon:
    def __init__(self, u0, umin, umax, dumax, Qw, controlRate, popts):
        self._lib = _lib.load()
        self._wl = _lib.WLCon_t()
        self._lib.wlConInit(C.byref(self._wl), _fp(_v(u0, 4)), _fp(_v(umin, 4)), _fp(_v(umax, 4)), _fp(_v(dumax, 4)),
                            _fp(_v(Qw, 6)), float(controlRate), _fp(_v(popts, 90)))

    def update(self, h0, pdotdes):
        u1, w0 = np.zeros(4, np.float32), np.zeros(6, np.float32)
        self._lib.wlConUpdate(C.byref(self._wl), _fp(u1), _fp(w0), _fp(_v(h0, 6)), _fp(_v(pdotdes, 6)))
        if np.isnan(u1[0]):
            raise RuntimeError("wlConUpdate failed: %s" % self._lib.umpcLastError().decode())
        return u1, w0
