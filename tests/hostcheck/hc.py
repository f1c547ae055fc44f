"""ctypes driver of tests/hostcheck/libhostcheck.so (development aid: the device algorithm compiled for the host)."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def _par_type(real):
    class Par(C.Structure):
        _fields_ = [('dt', real), ('g', real), ('Tmax', real), ('Qyr', real * 6), ('Qyf', real * 6),
                    ('Qdyr', real * 6), ('Qdyf', real * 6), ('R', real * 3), ('Ibi', real * 3),
                    ('max_iter', C.c_int), ('rho', real), ('sigma', real), ('alpha', real), ('eps_abs', real),
                    ('eps_rel', real), ('scaling_iters', C.c_int), ('check_termination', C.c_int),
                    ('adaptive_rho_interval', C.c_int), ('adaptive_rho_tol', real), ('debug_phase', C.c_int)]
    return Par


Par = _par_type(C.c_float)     # mirrors UmpcParams of the float build
Par64 = _par_type(C.c_double)  # ... and of the -DUMPC_REAL=double build


def make_par(d, f64=False):
    npf = np.float64 if f64 else np.float32
    par = (Par64 if f64 else Par)(dt=d['dt'], g=d['g'], Tmax=d['TtoWmax'] * d['g'], max_iter=d['maxIter'], rho=0.1, sigma=1e-6, alpha=1.6,
              eps_abs=1e-4, eps_rel=1e-4, scaling_iters=10, check_termination=0, adaptive_rho_interval=0, adaptive_rho_tol=5.0)
    for i in range(3):
        par.Qyr[i] = d['wpr']; par.Qyf[i] = d['wpf']; par.Qyr[3 + i] = par.Qyf[3 + i] = d['ws']
        par.Qdyr[i] = d['wvr']; par.Qdyf[i] = d['wvf']; par.Qdyr[3 + i] = par.Qdyf[3 + i] = d['wds']
        par.Ibi[i] = npf(1) / npf(d['Ib'][i])
    par.R[0] = d['wthrust']; par.R[1] = par.R[2] = d['wmom']
    return par


class HostUMPC:
    def __init__(self, defaults, f64=False, lib=None):
        self.lib = C.CDLL(os.path.join(HERE, lib or ('libhostcheck64.so' if f64 else 'libhostcheck.so')))
        self.par = make_par(defaults, f64)
        self.np, self.P = (np.float64, C.POINTER(C.c_double)) if f64 else (np.float32, C.POINTER(C.c_float))
        self.creal = C.c_double if f64 else C.c_float
        assert C.sizeof(self.par) == self.lib.hc_sizeof_params()
        self.st = C.create_string_buffer(self.lib.hc_sizeof_state())
        self.lib.hc_init(self.st)

    def set_converged(self, max_iter=4000, check_termination=25, adaptive_rho_interval=100):
        """the Python path's OSQP settings (template_controllers.py:190-191 + OSQP defaults)"""
        self.par.max_iter, self.par.check_termination = max_iter, check_termination
        self.par.adaptive_rho_interval = adaptive_rho_interval

    def update(self, p, R, dq, pdes, dpdes, sdes, aT0=-1.0):
        f = lambda a: np.ascontiguousarray(a, self.np).ravel()
        P = lambda a: a.ctypes.data_as(self.P)
        uq, acc = np.zeros(3, self.np), np.zeros(6, self.np)
        args = [f(p), f(np.asarray(R, float).T), f(dq), f(pdes), f(dpdes), f(sdes)]
        self.lib.hc_update(C.byref(self.par), self.st, P(uq), P(acc), *[P(a) for a in args], self.creal(aT0))
        return uq, acc

    def get(self):
        P = lambda a: a.ctypes.data_as(self.P)
        z = lambda n: np.zeros(n, self.np)
        o = dict(x=z(45), y=z(39), xs=z(45), zs=z(39), ys=z(39), D=z(45), E=z(39), info=z(5))
        self.lib.hc_get(self.st, *[P(o[k]) for k in ('x', 'y', 'xs', 'zs', 'ys', 'D', 'E', 'info')])
        return o
