"""GPU parity tests of CONVERGED MODE (the reference's Python path: osqp on float64, eps 1e-4, termination test every
25 iterations, adaptive rho; template_controllers.py:190-258) through the C-ABI, against
 (a) tests/golden/umpc_conv.npz -- the reference's vendored OSQP compiled in float64 with those settings,
 (b) the oracle's float64 converged flavour (pinned to (a) and to the reference's Python loop by the CPU tests),
 (c) the closed loop of the reference's own Python controller + loop (tests/golden/pyloop_c1.npz, py_hl5).
Tolerance: |x - x*|_inf <= 10 eps_abs + eps_rel |x*|_inf, eps = 1e-4; same for the duals.
Nothing here reads /root/reference."""
import os

import numpy as np
import pytest

import util
from util import GOLDEN

pytestmark = pytest.mark.gpu

EPS = 1e-4
CONV = dict(maxIter=4000, checkTermination=25, adaptiveRhoInterval=100)


def _tol(ref):
    return 10 * EPS + EPS * np.abs(ref).max(axis=-1)


def test_converged_update_matches_float64_osqp(cuda_required):
    from oracle import binding as OB
    from robobee3d_b200.batch import BatchUMPC
    gold = np.load(os.path.join(GOLDEN, "umpc_seq.npz"))
    conv = np.load(os.path.join(GOLDEN, "umpc_conv.npz"))
    B, ncall = gold["p"].shape[:2]
    b = BatchUMPC(B, **CONV)
    oracles = [OB.OracleUMPC("f64") for _ in range(B)]
    for o in oracles:
        o.set_converged()
    for c in range(ncall):
        aT0 = gold["aT0"][:, c].astype(np.float32)
        uq, acc = b.update(gold["p"][:, c], gold["R"][:, c], gold["dq"][:, c], gold["pdes"][:, c], gold["dpdes"][:, c],
                           gold["sdes"][:, c], aT0)
        s = b.solve_info(full=True)
        xo, yo, it, uo, ao = [], [], [], [], []
        for e, o in enumerate(oracles):
            u_, a_ = o.update(gold["p"][e, c], gold["R"][e, c], gold["dq"][e, c], gold["pdes"][e, c], gold["dpdes"][e, c],
                              gold["sdes"][e, c], float(aT0[e]))
            x, y, info = o.solution()
            xo.append(x); yo.append(y); it.append(info["iter"]); uo.append(u_); ao.append(a_)
            assert info["status"] == 1
        xo, yo, it, uo, ao = map(np.asarray, (xo, yo, it, uo, ao))
        assert (s["status"] == 1).all(), "call %d: %s" % (c, s["status"])
        # same algorithm in the same arithmetic width: the iteration counts agree (a QP whose residual sits on the
        # threshold may stop one check earlier or later)
        assert (s["iter"] == it).mean() >= 0.9 and (np.abs(s["iter"] - it) <= 25).all(), (s["iter"], it)
        assert (s["iter"] % 25 == 0).all() and (s["iter"] < 4000).all()
        assert (np.abs(s["x"] - xo).max(axis=1) <= _tol(xo)).all(), np.abs(s["x"] - xo).max()
        assert (np.abs(s["y"] - yo).max(axis=1) <= _tol(yo)).all(), np.abs(s["y"] - yo).max()
        assert (np.abs(uq - uo).max(axis=1) <= _tol(xo)).all()
        assert (np.abs(acc - ao).max(axis=1) <= _tol(xo) / 5.0 + 1e-6 * np.abs(ao).max(axis=1)).all()
        assert (s["pri_res"] <= EPS + EPS * 1e3).all()
        if c == 0:
            xr, yr = conv["x"], conv["y"]
            assert (np.abs(s["x"] - xr).max(axis=1) <= _tol(xr)).all()
            assert (np.abs(s["y"] - yr).max(axis=1) <= _tol(yr)).all()
            assert (s["iter"] == conv["iter"].astype(int)).mean() >= 0.9
    b.close()


def test_converged_mode_is_better_than_50_iterations(cuda_required):
    """size-independent property at a larger batch: the converged solution has residuals below eps for every env, the
    C path's 50 truncated iterations do not -- and the two controllers agree on uquad to the C path's accuracy"""
    from robobee3d_b200.batch import BatchUMPC
    B = 4096
    p, R, dq = util.config2_states(B, seed=21)
    z3 = np.zeros((B, 3)); e3 = np.tile([0, 0, 1.0], (B, 1))
    aT0 = np.full(B, 9.81e-3, np.float32)
    bc, bf = BatchUMPC(B, **CONV), BatchUMPC(B)
    uqc, _ = bc.update(p, R, dq, z3, z3, e3, aT0)
    uqf, _ = bf.update(p, R, dq, z3, z3, e3, aT0)
    sc, sf = bc.solve_info(full=True), bf.solve_info(full=True)
    assert (sc["status"] == 1).all() and (sc["iter"] % 25 == 0).all()
    assert np.isfinite(sc["x"]).all() and np.isfinite(uqc).all()
    assert np.median(sc["pri_res"]) < np.median(sf["pri_res"])
    # warm repeat of the same inputs: already converged -> stops at the first check
    bc.update(p, R, dq, z3, z3, e3, np.full(B, -1.0, np.float32))
    s2 = bc.solve_info()
    assert (s2["status"] == 1).all() and np.median(s2["iter"]) <= np.median(sc["iter"])
    bc.close(); bf.close()


def test_converged_rollout_matches_reference_python_loop(cuda_required):
    """config 1 through the Python path: controlTest(useMPC, hlInterval=5) with template_controllers.UprightMPC2 +
    OSQP (float64) in the reference's own loop vs the fused rollout kernel in converged mode"""
    from robobee3d_b200.batch import BatchUMPC
    gold = np.load(os.path.join(GOLDEN, "pyloop_c1.npz"))
    gy, gm = gold["py_hl5_y"], gold["py_hl5_metrics"]
    nt = 2500
    from oracle import binding as OB
    p0, R0, dq0 = OB.control_test_ic()
    B = 256
    b = BatchUMPC(B, hlInterval=5.0, **CONV)
    b.set_state(np.tile(p0, (B, 1)), np.tile(np.asarray(R0).reshape(1, 3, 3), (B, 1, 1)), np.tile(dq0, (B, 1)), task=0,
                tp=(0.0, 0.0, 0.1, 0.0))
    b.set_log(0, nt)
    b.rollout(nt)
    y, u = b.get_log(nt)
    assert np.abs(y[::10, :3] - gy[:, :3]).max() <= util.TOL_TRAJ_P
    assert np.abs(y[::10, 3:6] - gy[:, 3:6]).max() <= util.TOL_TRAJ_S
    m = b.metrics()
    assert (m["nstep"] == nt).all() and np.ptp(m["nsolve"]) == 0
    assert abs(m["err"][0] / gm[0] - 1) < 1e-2 and abs(m["eff"][0] / gm[1] - 1) < 2e-2
    # every env saw the same inputs -> identical flights
    pf, _, _, _ = b.state()
    assert np.abs(pf - pf[0]).max() == 0.0
    b.close()


def test_python_controller_dropin(cuda_required):
    """robobee3d_b200.template_controllers.UprightMPC2 / createMPC: the reference's Python class surface
    (template_controllers.py:169-280) over the converged-mode kernel, vs the float64 oracle on a warm-started sequence"""
    from oracle import binding as OB
    from robobee3d_b200.template_controllers import createMPC
    gold = np.load(os.path.join(GOLDEN, "umpc_seq.npz"))
    pyver, cver = createMPC()
    o = OB.OracleUMPC("f64")
    o.set_converged()
    e = 5
    for c in range(gold["p"].shape[1]):
        a = (gold["p"][e, c], gold["R"][e, c], gold["dq"][e, c], gold["pdes"][e, c], gold["dpdes"][e, c], gold["sdes"][e, c],
             float(gold["aT0"][e, c]))
        u, acc = pyver.update(*a)
        uo, ao = o.update(*a)
        x, y, info = o.solution()
        tol = 10 * EPS + EPS * np.abs(x).max()
        assert u.shape == (3,) and acc.shape == (6,) and pyver.prevsol.shape == (45,)
        assert np.abs(u - uo).max() <= tol and np.abs(pyver.prevsol - x).max() <= tol
        assert np.abs(acc - pyver.getAccDes(gold["R"][e, c], gold["dq"][e, c])).max() <= 1e-5 * max(1.0, np.abs(acc).max())
        assert pyver.info["status"] == 1 and abs(pyver.T0 - u[0]) < 1e-7
    uc, _ = cver.update(*a[:6], a[6])
    assert np.isfinite(uc).all()
    pyver.close()
