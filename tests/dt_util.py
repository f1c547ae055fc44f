"""Shared helpers of the planar dir-tran tests (config 5): scenarios, the host build of the CUDA core, tolerances."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
DP = C.POINTER(C.c_double)

XR_HOVER = np.array([-0.03, 0.02, 0, 0, 0, 0.0])      # goal(), mpc_thrust_strokedev.py:80
Y0_HOVER = np.array([0, 0, 0, 1.0, 0, 10.0])           # :42
# converged solutions of a strictly convex QP agree to O(eps) whatever the iteration path: eps_abs = eps_rel = 1e-5
TOL_U = np.array([2e-3, 2e-6])    # |u - u_oracle| per input (u0 ~ 10, u1 ~ 5e-3): 20 eps relative to the input scale
TOL_X_LOCKSTEP = 1e-6             # when the iteration counts agree the iterates agree to rounding amplified by the loop


def mild_states(B, seed=0):
    """small perturbations about the hover goal: solves of a few hundred iterations at most"""
    r = np.random.default_rng(seed)
    y = np.tile(XR_HOVER, (B, 1))
    y[:, 0:2] += r.uniform(-0.01, 0.01, (B, 2))
    y[:, 2] += r.uniform(-0.05, 0.05, B)
    y[:, 3:5] += r.uniform(-0.05, 0.05, (B, 2))
    y[:, 5] += r.uniform(-0.2, 0.2, B)
    return y


def build_hostcheck():
    out = os.path.join(ROOT, "tests", "hostcheck", "libdthostcheck.so")
    src = os.path.join(ROOT, "tests", "hostcheck", "dirtran_hostcheck.cpp")
    core = os.path.join(ROOT, "robobee3d_b200", "csrc", "dirtran_core.cuh")
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(src), os.path.getmtime(core)):
        subprocess.run(["g++", "-O2", "-fPIC", "-shared", "-ffp-contract=off", "-x", "c++", "-o", out, src], check=True)
    return out


class HostCore:
    """robobee3d_b200/csrc/dirtran_core.cuh compiled for the host (one environment) -- the code the kernels run"""

    def __init__(self, par):
        self.H = C.CDLL(build_hostcheck())
        self.H.dthc_create.restype = C.c_void_p
        self.H.dthc_create.argtypes = [C.c_void_p]
        self.H.dthc_destroy.argtypes = [C.c_void_p]
        self.H.dthc_set.argtypes = [C.c_void_p, DP, DP]
        self.H.dthc_update.argtypes = [C.c_void_p, DP, DP, DP]
        self.H.dthc_plant.argtypes = [C.c_void_p, DP, DP]
        self.H.dthc_rollout.argtypes = [C.c_void_p, DP, DP, C.c_int]
        self.H.dthc_solution.argtypes = [C.c_void_p, DP, DP]
        self.N = par.N
        self.n = (par.N + 1) * 6 + par.N * 2
        self.m = self.n + (par.N + 1) * 6
        self.h = self.H.dthc_create(C.byref(par))

    def __del__(self):
        try:
            self.H.dthc_destroy(self.h)
        except Exception:
            pass

    @staticmethod
    def _p(a):
        return a.ctypes.data_as(DP)

    def set(self, ctrl=None, xr=None):
        c = np.ascontiguousarray(ctrl, dtype=np.float64) if ctrl is not None else None
        x = np.ascontiguousarray(xr, dtype=np.float64) if xr is not None else None
        self.H.dthc_set(self.h, self._p(c) if c is not None else None, self._p(x) if x is not None else None)

    def update(self, x0):
        x0 = np.ascontiguousarray(x0, dtype=np.float64)
        u = np.zeros(2); info = np.zeros(6)
        self.H.dthc_update(self.h, self._p(x0), self._p(u), self._p(info))
        return u, info

    def plant(self, x, u):
        x = np.ascontiguousarray(x, dtype=np.float64).copy()
        self.H.dthc_plant(self.h, self._p(x), self._p(np.ascontiguousarray(u, dtype=np.float64)))
        return x

    def rollout(self, x, nsteps):
        x = np.ascontiguousarray(x, dtype=np.float64).copy()
        u = np.zeros(2)
        self.H.dthc_rollout(self.h, self._p(x), self._p(u), int(nsteps))
        return x, u

    def solution(self):
        x = np.zeros(self.n); y = np.zeros(self.m)
        self.H.dthc_solution(self.h, self._p(x), self._p(y))
        return x, y
