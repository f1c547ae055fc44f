"""Batched twin of the reference's hover-tuning sweep, `gainTuningSims` (template/uprightmpc2.py:272-303): every grid
point of the (wpr, wvr) mesh is one environment of ONE batch -- its own controller weights, the same controlTest flight
-- instead of npts^2 sequential Python runs.  The result dictionary and the .npz it writes have the reference's keys
(`xv, yv, costs, efforts`, uprightmpc2.py:303), so `gainTuningPlots` (uprightmpc2.py:305-352) reads it unchanged.

    from robobee3d_b200.sweeps import gain_tuning_sims
    r = gain_tuning_sims([0.5, 10], [0.5e3, 10e3], npts=10)          # uprightmpc2.py:363

The reference sweep flies the PYTHON controller (controlTest(None, ...) -> createMPC(**kwargs)[0],
uprightmpc2.py:89-90): converged float64 solves, MPC every plant step, tend = 1000 ms, taulim = 10.
`converged=False` flies the 50-iteration float32 C controller with the same weights instead.
"""
import numpy as np

from .batch import BatchUMPC
from .template_controllers import OSQP_DEFAULTS


def control_test_ic():
    """IC of controlTest's default branch (uprightmpc2.py:99-103): p = 0, R = Ry(-0.5) Rx(0.5), dq[0] = 0.1"""
    cx, sx, cy, sy = np.cos(0.5), np.sin(0.5), np.cos(-0.5), np.sin(-0.5)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    return np.zeros(3), Ry @ Rx, np.array([0.1, 0, 0, 0, 0, 0.0])


def gain_tuning_sims(k1range, k2range, npts=10, tend=1000.0, dtsim=0.2, taulim=10.0, hlInterval=None, converged=True,
                     device=0, save="mpc_wpos.npz", chunk=500):
    """costs[i, j], efforts[i, j] = logMetric(controlTest(createMPC(wpr=xv[i,j], wvr=yv[i,j], wpf=5 xv, wvf=2 yv), tend,
    taulim=taulim)) for the mesh xv, yv = meshgrid(linspace(*k1range, npts), linspace(*k2range, npts), indexing='ij')."""
    k1s, k2s = np.linspace(*k1range, num=npts), np.linspace(*k2range, num=npts)
    xv, yv = np.meshgrid(k1s, k2s, indexing="ij")
    B = xv.size
    kw = dict(OSQP_DEFAULTS) if converged else {}
    # the default branch of controlTest builds p with an integer dtype: pdes is truncated
    b = BatchUMPC(B, device=device, dtsim=dtsim, hlInterval=hlInterval, taulim=taulim, intInitialPos=1, **kw)
    try:
        b.set_weights(wpr=xv.ravel(), wvr=yv.ravel(), wpf=5 * xv.ravel(), wvf=2 * yv.ravel())
        p, R, dq = control_test_ic()
        b.set_state(np.tile(p, (B, 1)), np.tile(R, (B, 1, 1)), np.tile(dq, (B, 1)), task=0, tp=(0.0, 0.0, 0.1, 0.0))
        nt = len(np.arange(tend, step=dtsim))
        for k in range(0, nt, chunk):
            b.rollout(min(chunk, nt - k))
        m = b.metrics()
        pf = b.state()[0]
        costs = m["err"].astype(np.float64).reshape(xv.shape)
        efforts = m["eff"].astype(np.float64).reshape(xv.shape)
        bad = ~np.isfinite(pf).all(axis=1).reshape(xv.shape)  # the reference records a failed run as NaN (uprightmpc2.py:300-301)
        costs[bad] = np.nan
        efforts[bad] = np.nan
    finally:
        b.close()
    out = dict(xv=xv, yv=yv, costs=costs, efforts=efforts)
    if save:
        np.savez(save, **out)
    return out
