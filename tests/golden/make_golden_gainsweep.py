"""tests/golden/make_golden_gainsweep.py -- golden of the reference's MPC weight sweep (uprightmpc2.py:272-303).

`gainTuningSims(True, 'wpos', [0.5, 10], [0.5e3, 10e3], None, None)` (uprightmpc2.py:363) is a nested function of the
reference's __main__ block, so it cannot be imported; this script runs ITS BODY's calls on the reference's own
module-level functions: for every grid point  createMPC(wpr=k1, wvr=k2, wpf=5 k1, wvf=2 k2)  (template_controllers.py:260),
controlTest(mdl, 1000, useMPC=True, showPlots=False, taulim=10)  (uprightmpc2.py:87-159) and logMetric (:161-175), with
the solver backends substituted exactly as in make_golden_pyloop.py (osqp -> the vendored OSQP in float64).

As shipped, controlTest(None, ...) picks the PYTHON controller (`mdl, _ = createMPC(**kwargs)`, uprightmpc2.py:89-90) and
then calls mdl.update with 6 arguments (:139) although UprightMPC2.update needs actualT0 (template_controllers.py:253):
every grid point would raise and be recorded as NaN by the sweep's bare except.  The adapter PyWithDefaultT0 supplies
actualT0 = -1, which is what the C controller's pybind default does.

Run in the build container:  python tests/golden/make_golden_gainsweep.py   ->  tests/golden/gain_sweep.npz
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden_pyloop as G  # noqa: E402


def main():
    G.B.build(ref=True)
    G.install_stubs()
    import uprightmpc2 as U  # noqa: the reference's
    from template_controllers import createMPC
    npts, tend = 3, 1000
    k1s, k2s = np.linspace(0.5, 10, num=npts), np.linspace(0.5e3, 10e3, num=npts)
    xv, yv = np.meshgrid(k1s, k2s, indexing="ij")
    costs, efforts = np.zeros_like(xv), np.zeros_like(xv)
    finals = np.zeros(xv.shape + (12,))
    for i in range(npts):
        for j in range(npts):
            kwargs = {"wpr": xv[i, j], "wvr": yv[i, j], "wpf": 5 * xv[i, j], "wvf": 2 * yv[i, j]}
            up, _ = createMPC(**kwargs)
            log = U.controlTest(G.PyWithDefaultT0(up), tend, useMPC=True, showPlots=False, taulim=10)
            finals[i, j] = log["y"][-1]
            costs[i, j], efforts[i, j] = U.logMetric(log)
            print(i, j, kwargs, costs[i, j], efforts[i, j], flush=True)
    np.savez_compressed(os.path.join(HERE, "gain_sweep.npz"), xv=xv, yv=yv, costs=costs, efforts=efforts, finals=finals,
                        tend=np.array(tend), taulim=np.array(10.0))


if __name__ == "__main__":
    main()
