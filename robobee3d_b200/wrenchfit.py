"""Wrench-map fitting on the GPU (include/wrenchfit_b200.h): the twin of template/sdab_num_wrenchmap.py's
`wrenchFromKinematics` (:165-214) and of its fitting statement (:262-266), batched over datasets.

    ws = wrench_from_kinematics(kins, fs, params, kaerox=1, strokex=1.5)
    popts = fit_wrench_map(xdata, ws)            # (6, 15); np.ravel(popts) is the popts[90] of the WL stage
"""
import ctypes as C

import numpy as np

from . import _lib

DP = C.POINTER(C.c_double)
EXPORTS = ["wfitWrenchFromKinematics", "wfitFit", "wfitFitNoAlloc"]
NOALLOC_MAX_DATASETS, NOALLOC_MAX_TRIALS = 64, 16384   # include/wrenchfit_b200.h
# aero constants of template/robobee.py:7-17 and the wing parameters sdab_num_wrenchmap.py:252-253 uses
DEFAULT_PARAMS = dict(Aw=54.4, AR=4.5, r2h=0.551, ycp=7.5, R=3.0, RHO=1.225e-3, CD0=0.4, CDmax=3.4, CLmax=1.8)


def _bind():
    L = _lib.load()
    L.wfitWrenchFromKinematics.argtypes = [C.c_int, DP, DP, DP, DP]
    L.wfitFit.argtypes = [C.c_int, C.c_int, DP, DP, DP, DP, C.POINTER(C.c_int), C.POINTER(C.c_float)]
    L.wfitFitNoAlloc.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    return L


def fit_wrench_map_noalloc(F, M, d_xdata, d_ws, d_popts, d_rank=None, stream=None):
    """wfitFitNoAlloc: device addresses (ints) in and out, nothing allocated, enqueued on `stream` (a cudaStream_t handle
    as an int, None = default stream) without synchronising.  No counterpart in the reference (see the header)."""
    L = _bind()
    _lib.check(L.wfitFitNoAlloc(int(F), int(M), d_xdata, d_ws, d_popts, d_rank, stream), "wfitFitNoAlloc")


def _d(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64))


def wrench_from_kinematics(kins, freq, params=None, kaerox=1.0, strokex=1.0, wbias=np.zeros(6)):
    L = _bind()
    p = dict(DEFAULT_PARAMS)
    p.update(params or {})
    kins = _d(kins).reshape(-1, 9)
    M = len(kins)
    freq = _d(np.broadcast_to(freq, (M,)))
    par = _d([p["Aw"], p["AR"], p["r2h"], p["ycp"], p["R"], p["RHO"], p["CD0"], p["CDmax"], p["CLmax"], kaerox, strokex, *np.asarray(wbias, float)])
    ws = np.empty((M, 6))
    _lib.check(L.wfitWrenchFromKinematics(M, kins.ctypes.data_as(DP), freq.ctypes.data_as(DP), par.ctypes.data_as(DP), ws.ctypes.data_as(DP)),
               "wfitWrenchFromKinematics")
    return ws


def fit_wrench_map(xdata, ws, return_info=False):
    """xdata (M, 4) / (F, M, 4), ws (M, 6) / (F, M, 6) -> popts (6, 15) / (F, 6, 15)"""
    L = _bind()
    xdata, ws = _d(xdata), _d(ws)
    single = xdata.ndim == 2
    if single:
        xdata, ws = xdata[None], ws[None]
    F, M = xdata.shape[:2]
    popts, fit, rank, ms = np.empty((F, 6, 15)), np.empty((F, M, 6)), np.zeros(F, np.int32), C.c_float()
    _lib.check(L.wfitFit(F, M, xdata.ctypes.data_as(DP), ws.ctypes.data_as(DP), popts.ctypes.data_as(DP), fit.ctypes.data_as(DP),
                         rank.ctypes.data_as(C.POINTER(C.c_int)), C.byref(ms)), "wfitFit")
    if single:
        popts, fit, rank = popts[0], fit[0], rank[0]
    return (popts, dict(fit=fit, rank=rank, kernel_ms=ms.value)) if return_info else popts
