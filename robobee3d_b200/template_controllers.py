"""Drop-in for the PYTHON controller of the reference, `template_controllers.UprightMPC2`
(template/template_controllers.py:169-258) and `createMPC` (:260-280): same constructor arguments, `update`
signature, return shapes and public attributes (`T0`, `prevsol`, `dt`, `g`, `Tmax`, `N`), backed by the
converged-mode CUDA kernel of libumpc_b200.so (float64 ADMM to eps 1e-4 with the osqp package's defaults:
termination test every 25 iterations, adaptive rho) instead of the osqp package.

    from robobee3d_b200.template_controllers import createMPC      # instead of template_controllers
    pyver, cver = createMPC()
    u, accdes = pyver.update(p0, R0, dq0, pdes, dpdes, sdes, actualT0)

One object = one controller on the GPU; thousands of them in lockstep are `robobee3d_b200.batch.BatchUMPC(B,
checkTermination=25, adaptiveRhoInterval=100, maxIter=4000)`.  No CPU fallback: construction raises without a device.
The C-ABI carries float32 inputs and outputs; the solve itself is float64.
"""
import numpy as np

from .batch import BatchUMPC
from .uprightmpc2py import UprightMPC2C

ny, nu = 6, 3
# osqp defaults the reference relies on (osqp 0.5/0.6 constants.h:55-95; eps from template_controllers.py:191)
OSQP_DEFAULTS = dict(maxIter=4000, checkTermination=25, adaptiveRhoInterval=100, epsAbs=1e-4, epsRel=1e-4)
Ib = np.diag([3333.0, 3333.0, 1000.0])  # genqp.py:22


class UprightMPC2:
    def __init__(self, N, dt, g, TtoWmax, ws, wds, wpr, wpf, wvr, wvf, wthrust, wmom, Ib, device=0):
        self.N, self.dt, self.g, self.Tmax = N, dt, g, TtoWmax * g
        ib = tuple(np.asarray(Ib, float).ravel())
        if N == 3:
            # the generated thread-per-QP device code is the N = 3 instance the reference ships (uprightmpc2.h:19-25)
            self._b = BatchUMPC(1, device=device, dt=dt, g=g, TtoWmax=TtoWmax, ws=ws, wds=wds, wpr=wpr, wpf=wpf, wvr=wvr,
                                wvf=wvf, wthrust=wthrust, wmom=wmom, Ib=ib, **OSQP_DEFAULTS)
            self._g = None
        else:
            # any other horizon: the warp-per-QP generic solver (gqp.cu) behind the same QP functions
            from .gqp import BatchUprightMPC2G
            self._g = BatchUprightMPC2G(1, N=N, device=device, dt=dt, g=g, TtoWmax=TtoWmax, ws=ws, wds=wds, wpr=wpr, wpf=wpf,
                                        wvr=wvr, wvf=wvf, wthrust=wthrust, wmom=wmom, Ib=ib, maxIter=OSQP_DEFAULTS["maxIter"],
                                        checkTermination=OSQP_DEFAULTS["checkTermination"],
                                        adaptiveRhoInterval=OSQP_DEFAULTS["adaptiveRhoInterval"], epsAbs=OSQP_DEFAULTS["epsAbs"],
                                        epsRel=OSQP_DEFAULTS["epsRel"])
            self._b = self._g
        self.T0 = 0.0
        self.prevsol = np.zeros(N * (2 * ny + nu))
        self.obj_val = None
        self._setT0 = 0.0  # what the device holds

    def update(self, p0, R0, dq0, pdes, dpdes, sdes, actualT0):
        # an edit of self.T0 between calls is honoured like the reference's attribute (template_controllers.py:239-241)
        a = actualT0 if actualT0 >= 0 else (self.T0 if self.T0 != self._setT0 else -1.0)
        args = (np.asarray(p0, float)[None], np.asarray(R0, float)[None], np.asarray(dq0, float)[None],
                np.asarray(pdes, float)[None], np.asarray(dpdes, float)[None], np.asarray(sdes, float)[None])
        if self._g is not None:
            uq, acc, _ = self._g.update(*args, np.array([a], float))
            s = self._g.qp()
        else:
            uq, acc = self._b.update(*args, np.array([a], np.float32))
            s = self._b.solve_info(full=True)
        self.prevsol = s["x"][0].astype(float)
        self.info = dict(iter=int(s["iter"][0]), status=int(s["status"][0]), pri_res=float(s["pri_res"][0]),
                         dua_res=float(s["dua_res"][0]))
        if self.info["status"] != 1:
            print({2: "solved inaccurate", -2: "maximum iterations reached", -7: "problem non convex"}.get(self.info["status"],
                                                                                                            "unsolved"))
        self.T0 = self._setT0 = float(uq[0, 0])
        return uq[0].astype(float), acc[0].astype(float)

    def update2(self, p0, R0, dq0, pdes, dpdes, sdes):
        return self.update(p0, R0, dq0, pdes, dpdes, sdes, -1.0)[0]

    def getAccDes(self, R0, dq0):
        """desired acceleration from the last solution (template_controllers.py:245-251)"""
        R0, dq0 = np.asarray(R0, float), np.asarray(dq0, float)
        dy1 = self.prevsol[ny * self.N: ny * self.N + ny]
        e3h = np.array([[0.0, -1.0, 0.0], [1.0, 0.0, 0.0], [0.0, 0.0, 0.0]])
        dq1 = np.hstack((dy1[:3], e3h @ R0.T @ dy1[3:6]))
        return (dq1 - dq0) / self.dt

    def close(self):
        self._b.close()


def createMPC(N=3, ws=1e1, wds=1e3, wpr=1, wvr=1e3, wpf=5, wvf=2e3, wthrust=1e-1, wmom=1e-2, TtoWmax=2, popts=None, **kwargs):
    """(pyver, cver) like the reference: the converged float64 controller and the 50-iteration float32 one"""
    dt, g = 5, 9.81e-3
    pyver = UprightMPC2(N, dt, g, TtoWmax, ws, wds, wpr, wpf, wvr, wvf, wthrust, wmom, Ib.diagonal())
    cver = UprightMPC2C(dt, g, TtoWmax, ws, wds, wpr, wpf, wvr, wvf, wthrust, wmom, Ib.diagonal(), 50)
    return pyver, cver
