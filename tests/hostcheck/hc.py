"""ctypes driver of tests/hostcheck/libhostcheck.so (development aid: the device algorithm compiled for the host)."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
FP = C.POINTER(C.c_float)


class Par(C.Structure):
    _fields_ = [('dt', C.c_float), ('g', C.c_float), ('Tmax', C.c_float), ('Qyr', C.c_float * 6), ('Qyf', C.c_float * 6),
                ('Qdyr', C.c_float * 6), ('Qdyf', C.c_float * 6), ('R', C.c_float * 3), ('Ibi', C.c_float * 3),
                ('max_iter', C.c_int), ('rho', C.c_float), ('sigma', C.c_float), ('alpha', C.c_float), ('eps_abs', C.c_float),
                ('eps_rel', C.c_float), ('scaling_iters', C.c_int)]


def make_par(d):
    par = Par(dt=d['dt'], g=d['g'], Tmax=d['TtoWmax'] * d['g'], max_iter=d['maxIter'], rho=0.1, sigma=1e-6, alpha=1.6,
              eps_abs=1e-4, eps_rel=1e-4, scaling_iters=10)
    for i in range(3):
        par.Qyr[i] = d['wpr']; par.Qyf[i] = d['wpf']; par.Qyr[3 + i] = par.Qyf[3 + i] = d['ws']
        par.Qdyr[i] = d['wvr']; par.Qdyf[i] = d['wvf']; par.Qdyr[3 + i] = par.Qdyf[3 + i] = d['wds']
        par.Ibi[i] = np.float32(1) / np.float32(d['Ib'][i])
    par.R[0] = d['wthrust']; par.R[1] = par.R[2] = d['wmom']
    return par


class HostUMPC:
    def __init__(self, defaults):
        self.lib = C.CDLL(os.path.join(HERE, 'libhostcheck.so'))
        self.par = make_par(defaults)
        assert C.sizeof(Par) == self.lib.hc_sizeof_params()
        self.st = C.create_string_buffer(self.lib.hc_sizeof_state())
        self.lib.hc_init(self.st)

    def update(self, p, R, dq, pdes, dpdes, sdes, aT0=-1.0):
        f = lambda a: np.ascontiguousarray(a, np.float32).ravel()
        P = lambda a: a.ctypes.data_as(FP)
        uq, acc = np.zeros(3, np.float32), np.zeros(6, np.float32)
        args = [f(p), f(np.asarray(R, float).T), f(dq), f(pdes), f(dpdes), f(sdes)]
        self.lib.hc_update(C.byref(self.par), self.st, P(uq), P(acc), *[P(a) for a in args], C.c_float(aT0))
        return uq, acc

    def get(self):
        P = lambda a: a.ctypes.data_as(FP)
        o = dict(x=np.zeros(45, np.float32), y=np.zeros(39, np.float32), xs=np.zeros(45, np.float32), zs=np.zeros(39, np.float32),
                 ys=np.zeros(39, np.float32), D=np.zeros(45, np.float32), E=np.zeros(39, np.float32), info=np.zeros(4, np.float32))
        self.lib.hc_get(self.st, *[P(o[k]) for k in ('x', 'y', 'xs', 'zs', 'ys', 'D', 'E', 'info')])
        return o
