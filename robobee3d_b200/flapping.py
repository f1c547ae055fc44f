"""Flapping-wing plant on the GPU (include/flapping_b200.h): the twin of template/robobee.py's `aerodynamics` (:20-47),
`RobobeeSim.openLoop` (:277-300, fixed base) and of template/sdab_num_wrenchmap_simgen.py's `sweepFile` (:19-55), batched
over simulations.

    F, pcop = aerodynamics(theta, dtheta, lrSign)                       # (n, 2), (n, 2) -> (n, 3), (n, 3)
    qw = open_loop(VampL, VampR, uoffs, f, tendMS=100, h2=0)             # (steps, 14) like RobobeeSim.openLoop
    res = sweep(Vmeans, uoffss, fs, udiffs, h2s)                         # rows of numkins.npy: 5 inputs, 6 wrench, 9 features
    sweep_file('numkins.npy', Vmeans, uoffss, fs, udiffs, h2s)           # sweepFile, same file
"""
import ctypes as C
import itertools

import numpy as np

from . import _lib

DP = C.POINTER(C.c_double)
EXPORTS = ["fwDefaults", "fwAerodynamics", "fwOpenLoopSteps", "fwOpenLoopSweep"]


class FwParams(C.Structure):
    _fields_ = [(k, C.c_double) for k in ("CD0", "CDmax", "CLmax", "RHO", "rcopnondim", "Aw", "r2h", "cbar", "ycp", "Roffs", "d",
                                          "khinge", "bhinge", "kstroke")] + [
        ("m_spar", C.c_double), ("I_spar", C.c_double * 3), ("c_spar", C.c_double * 3), ("m_mem", C.c_double),
        ("I_mem", C.c_double * 3), ("c_mem", C.c_double * 3), ("g", C.c_double), ("act_T", C.c_double), ("h3", C.c_double)]


def _bind():
    L = _lib.load()
    L.fwDefaults.argtypes = [C.POINTER(FwParams)]
    L.fwDefaults.restype = None
    L.fwAerodynamics.argtypes = [C.POINTER(FwParams), C.c_int, DP, DP, C.c_int, DP, DP]
    L.fwOpenLoopSteps.argtypes = [C.c_double, C.c_double]
    L.fwOpenLoopSweep.argtypes = [C.POINTER(FwParams), C.c_int, DP, C.c_double, C.c_double, DP, C.c_int, DP, C.POINTER(C.c_float)]
    return L


def default_params(**over):
    """FwParams with the constants of robobee.py:7-18 and urdf/sdab.xacro; keyword overrides by field name"""
    L = _bind()
    p = FwParams()
    L.fwDefaults(C.byref(p))
    for k, v in over.items():
        if not hasattr(p, k):
            raise TypeError("unknown parameter %r" % k)
        setattr(p, k, v)
    return p


def _d(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64))


def aerodynamics(theta, dtheta, lrSign, params=None):
    L = _bind()
    p = params or default_params()
    th, dth = _d(theta).reshape(-1, 2), _d(dtheta).reshape(-1, 2)
    n = len(th)
    F, pc = np.empty((n, 3)), np.empty((n, 3))
    _lib.check(L.fwAerodynamics(C.byref(p), n, th.ctypes.data_as(DP), dth.ctypes.data_as(DP), int(lrSign), F.ctypes.data_as(DP),
                                pc.ctypes.data_as(DP)), "fwAerodynamics")
    return F, pc


def open_loop_steps(tendMS=100.0, timestep=0.1):
    return int(_bind().fwOpenLoopSteps(float(tendMS), float(timestep)))


def open_loop_batch(drive, tendMS=100.0, timestep=0.1, params=None, log_sim=None):
    """drive (B, 5) = VampL, VampR, uoffs, f, h2 -> features (B, 15) [, qw (steps, 14) of simulation log_sim], kernel ms"""
    L = _bind()
    p = params or default_params()
    drive = _d(drive).reshape(-1, 5)
    B = len(drive)
    feat, ms = np.empty((B, 15)), C.c_float()
    qw = None if log_sim is None else np.empty((open_loop_steps(tendMS, timestep), 14))
    _lib.check(L.fwOpenLoopSweep(C.byref(p), B, drive.ctypes.data_as(DP), float(tendMS), float(timestep), feat.ctypes.data_as(DP),
                                 -1 if log_sim is None else int(log_sim), None if qw is None else qw.ctypes.data_as(DP), C.byref(ms)),
               "fwOpenLoopSweep")
    return feat, qw, ms.value


def open_loop(VampL, VampR, uoffs, f, tendMS=100.0, h2=0.0, h3=0.0, timestep=0.1):
    """RobobeeSim.openLoop(VampL, VampR, uoffs, f, tendMS, h2, h3) on a fixed base: rows of q[:4], dq[:4], wrench sum"""
    return open_loop_batch([[VampL, VampR, uoffs, f, h2]], tendMS, timestep, default_params(h3=h3), log_sim=0)[1]


def sweep(Vmeans, uoffss, fs, udiffs, h2s, tendMS=100.0, timestep=0.1, params=None, return_ms=False):
    """sweepFile's `res` (sdab_num_wrenchmap_simgen.py:49-52): one row per grid point, in its loop order"""
    grid = np.array(list(itertools.product(Vmeans, uoffss, fs, udiffs, h2s)), float)
    Vm, uo, f, ud, h2 = grid.T
    drive = np.column_stack([Vm * (1 + ud), Vm * (1 - ud), uo, f, h2])       # singleSim, :18-19
    feat, _, ms = open_loop_batch(drive, tendMS, timestep, params)
    res = np.hstack([grid, feat])
    return (res, ms) if return_ms else res


def sweep_file(fname, Vmeans, uoffss, fs, udiffs, h2s, **kw):
    with open(fname, "wb") as f:
        np.save(f, sweep(Vmeans, uoffss, fs, udiffs, h2s, **kw))
