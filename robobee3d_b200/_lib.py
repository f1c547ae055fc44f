"""ctypes loader for libumpc_b200.so (the C-ABI declared in include/umpc_b200.h).

There is deliberately NO fallback: if the shared library is missing or no CUDA device
is present the product raises, it never routes through oracle/ or any CPU path.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# UMPC_B200_LIB: alternative build of the same library (kernel tuning experiments)
LIB_PATH = os.environ.get("UMPC_B200_LIB") or os.path.join(HERE, "libumpc_b200.so")

FP = C.POINTER(C.c_float)
IP = C.POINTER(C.c_int)
DP = C.POINTER(C.c_double)


class UmpcBatchParams(C.Structure):
    """include/umpc_b200.h: UmpcBatchParams"""
    _fields_ = [
        ("dt", C.c_float), ("g", C.c_float), ("TtoWmax", C.c_float), ("ws", C.c_float), ("wds", C.c_float),
        ("wpr", C.c_float), ("wpf", C.c_float), ("wvr", C.c_float), ("wvf", C.c_float), ("wthrust", C.c_float),
        ("wmom", C.c_float), ("Ib", C.c_float * 3), ("maxIter", C.c_int),
        ("dtsim", C.c_double), ("hlInterval", C.c_double), ("taulim", C.c_float), ("tpert", C.c_float), ("mb", C.c_float),
        ("wlMode", C.c_int),
        ("wl_u0", C.c_float * 4), ("wl_umin", C.c_float * 4), ("wl_umax", C.c_float * 4), ("wl_dumax", C.c_float * 4),
        ("wl_Qw", C.c_float * 6), ("wl_controlRate", C.c_float), ("wl_popts", C.c_float * 90),
        ("wlqpCheckTermination", C.c_int), ("intInitialPos", C.c_int),
        ("checkTermination", C.c_int), ("adaptiveRhoInterval", C.c_int), ("epsAbs", C.c_float), ("epsRel", C.c_float),
    ]


class UprightMPC_t(C.Structure):
    """include/umpc_b200.h: UprightMPC_t (== reference uprightmpc2.h:27-43)"""
    _fields_ = [
        ("dt", C.c_float), ("g", C.c_float), ("Tmax", C.c_float),
        ("Qyr", C.c_float * 6), ("Qyf", C.c_float * 6), ("Qdyr", C.c_float * 6), ("Qdyf", C.c_float * 6), ("R", C.c_float * 3),
        ("smin", C.c_float * 3), ("smax", C.c_float * 3), ("e3h", C.c_float * 9), ("e3hIbi", C.c_float * 9),
        ("l", C.c_float * 39), ("u", C.c_float * 39), ("q", C.c_float * 45), ("Px_data", C.c_float * 45),
        ("Ax_data", C.c_float * 48), ("Ax_idx", C.c_int * 48), ("nAxT0dt", C.c_int), ("nAxdt", C.c_int),
        ("c0", C.c_float * 6), ("T0", C.c_float),
    ]


class FunApprox_t(C.Structure):
    _fields_ = [("k", C.c_int), ("a0", C.c_float), ("a1", C.c_float * 4), ("A2", C.c_float * 16)]


class WLCon_t(C.Structure):
    _fields_ = [("u0", C.c_float * 4), ("umin", C.c_float * 4), ("umax", C.c_float * 4), ("dumax", C.c_float * 4),
                ("Qw", C.c_float * 36), ("fa", FunApprox_t * 6)]


class WLQP_t(C.Structure):  # wlcontroller/c/wlqp.h:21-25
    _fields_ = [("umin", C.c_float * 4), ("umax", C.c_float * 4), ("U0", C.c_float * 4), ("w0", C.c_float * 6),
                ("Q", C.c_float * 36)]


# every symbol include/umpc_b200.h declares (tests check the .so exports all of them)
EXPORTS = [
    "umpcInit", "umpcUpdate", "umpcS", "umpcGetInfo", "umpcGetSolution", "umpcFree",
    "wlConInit", "wlConUpdate", "wlconS", "wlControllerUpdate", "wlControllerUpdateW", "wlControllerReset",
    "wlqpInit", "wlqpUpdate", "wlqpSetWrenchModel", "funApproxInit", "funApproxF", "funApproxDf",
    "umpcBatchDefaults", "umpcLastError", "umpcBatchCreate", "umpcBatchDestroy", "umpcBatchSize", "umpcBatchReset",
    "umpcBatchUpdate", "umpcBatchEnvStep", "umpcBatchUpdateDev", "umpcBatchUpdateDevOn", "umpcBatchGetSolve", "umpcBatchGetQPData", "umpcBatchSetState",
    "umpcBatchGetState", "umpcBatchRollout", "umpcBatchGetMetrics", "umpcBatchReduceMetrics", "umpcBatchSetLog",
    "umpcBatchGetLog", "umpcBatchSync", "umpcBatchDevPtr", "umpcBatchSetWeights", "umpcBatchLastKernelMs", "umpcBatchTimerStart", "umpcBatchTimerStopMs", "umpcBatchSetLaunch", "umpcBatchSetRolloutInfo",
    "umpcMeasureFp32Peak", "umpcMeasureFp64Peak",
]

_lib = None


def load():
    """dlopen the product library; raises (no fallback) when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("robobee3d_b200: %s not found -- run `python -m robobee3d_b200.build` "
                           "(needs nvcc); there is no CPU fallback" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    L.umpcLastError.restype = C.c_char_p
    L.umpcBatchCreate.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.POINTER(UmpcBatchParams)]
    L.umpcBatchDestroy.argtypes = [C.c_void_p]
    L.umpcBatchDestroy.restype = None
    L.umpcBatchSize.argtypes = [C.c_void_p]
    L.umpcBatchReset.argtypes = [C.c_void_p]
    vp = C.c_void_p
    L.umpcBatchUpdate.argtypes = [vp] + [vp] * 9
    L.umpcBatchUpdateDev.argtypes = [vp] + [vp] * 9
    L.umpcBatchUpdateDevOn.argtypes = [vp, vp] + [vp] * 9
    L.umpcBatchEnvStep.argtypes = [vp] + [vp] * 12
    L.umpcBatchGetSolve.argtypes = [vp] + [vp] * 8
    L.umpcBatchGetQPData.argtypes = [vp, vp]
    L.umpcBatchSetState.argtypes = [vp] + [vp] * 5
    L.umpcBatchGetState.argtypes = [vp] + [vp] * 4
    L.umpcBatchRollout.argtypes = [vp, C.c_int]
    L.umpcBatchGetMetrics.argtypes = [vp, vp]
    L.umpcBatchReduceMetrics.argtypes = [vp, vp]
    L.umpcBatchSetLog.argtypes = [vp, C.c_int, C.c_int]
    L.umpcBatchGetLog.argtypes = [vp, vp, vp, C.c_int]
    L.umpcBatchSync.argtypes = [vp]
    L.umpcBatchDevPtr.argtypes = [vp, C.c_char_p]
    L.umpcBatchDevPtr.restype = C.c_void_p
    L.umpcBatchLastKernelMs.argtypes = [vp]
    L.umpcBatchLastKernelMs.restype = C.c_float
    L.umpcBatchTimerStart.argtypes = [vp]
    L.umpcBatchTimerStopMs.argtypes = [vp]
    L.umpcBatchTimerStopMs.restype = C.c_float
    L.umpcBatchSetLaunch.argtypes = [vp, C.c_int]
    L.umpcBatchSetRolloutInfo.argtypes = [vp, C.c_int]
    L.umpcBatchSetWeights.argtypes = [vp, vp]
    L.umpcMeasureFp32Peak.argtypes = [C.c_int, C.c_int]
    L.umpcMeasureFp32Peak.restype = C.c_double
    L.umpcMeasureFp64Peak.argtypes = [C.c_int, C.c_int]
    L.umpcMeasureFp64Peak.restype = C.c_double
    f = C.c_float
    L.umpcInit.argtypes = [C.POINTER(UprightMPC_t)] + [f] * 11 + [FP, C.c_int]
    L.umpcInit.restype = None
    L.umpcUpdate.argtypes = [C.POINTER(UprightMPC_t)] + [FP] * 8 + [f]
    L.umpcGetInfo.argtypes = [C.POINTER(UprightMPC_t), FP]
    L.umpcGetInfo.restype = None
    L.umpcGetSolution.argtypes = [C.POINTER(UprightMPC_t), FP, FP]
    L.umpcGetSolution.restype = None
    L.umpcFree.argtypes = [C.POINTER(UprightMPC_t)]
    L.umpcFree.restype = None
    L.umpcS.argtypes = [FP] * 8 + [f] * 11 + [FP, C.c_int, f]
    L.umpcS.restype = None
    L.wlConInit.argtypes = [C.POINTER(WLCon_t)] + [FP] * 5 + [f, FP]
    L.wlConInit.restype = None
    L.wlConUpdate.argtypes = [C.POINTER(WLCon_t)] + [FP] * 4
    L.wlConUpdate.restype = None
    L.wlconS.argtypes = [FP] * 7 + [f] + [FP] * 3
    L.wlconS.restype = None
    L.wlControllerUpdate.argtypes = [FP] * 5 + [f]
    L.wlControllerUpdate.restype = None
    L.wlControllerUpdateW.argtypes = [FP] * 6 + [f]
    L.wlControllerUpdateW.restype = None
    L.wlControllerReset.restype = None
    L.wlqpInit.argtypes = [C.POINTER(WLQP_t), f]
    L.wlqpInit.restype = None
    L.wlqpUpdate.argtypes = [C.POINTER(WLQP_t)] + [FP] * 4
    L.wlqpUpdate.restype = None
    L.wlqpSetWrenchModel.argtypes = [FP]
    L.wlqpSetWrenchModel.restype = None
    L.funApproxInit.argtypes = [C.POINTER(FunApprox_t), FP]
    L.funApproxInit.restype = None
    L.funApproxF.argtypes = [C.POINTER(FunApprox_t), FP]
    L.funApproxF.restype = f
    L.funApproxDf.argtypes = [FP, C.POINTER(FunApprox_t), FP]
    L.funApproxDf.restype = None
    _lib = L
    return L


def check(rc, what="libumpc_b200"):
    if rc != 0:
        raise RuntimeError("%s failed (%d): %s" % (what, rc, load().umpcLastError().decode()))
