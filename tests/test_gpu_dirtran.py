"""GPU parity tests of the batched planar dir-tran MPC (config 5) through the C-ABI (include/dirtran_b200.h) against
the CPU oracle's golden vectors (tests/golden/dirtran_hover.npz) and the oracle itself."""
import os

import numpy as np
import pytest

import dt_util
from dt_util import GOLDEN, TOL_U, XR_HOVER, Y0_HOVER

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(GOLDEN, "dirtran_hover.npz"))


@pytest.mark.parametrize("N", [3, 5, 10, 20])
def test_update_matches_golden(N):
    from robobee3d_b200.dirtran import BatchLTVMPC
    ys = dt_util.mild_states(16, seed=N)
    b = BatchLTVMPC(16, N=N, max_iter=200000)
    b.set_state(xr=XR_HOVER)
    u = b.update(ys)
    i = b.info()
    assert (i["status"] == 1).all()
    assert (np.abs(u - G["mild_u_N%d" % N]) <= TOL_U).all(), np.abs(u - G["mild_u_N%d" % N]).max(0)
    # same iteration path as the oracle (fused multiply-adds may move a borderline termination test by one check)
    same = i["iter"] == G["mild_iters_N%d" % N]
    assert same.sum() >= 14, (i["iter"], G["mild_iters_N%d" % N])
    assert np.abs(u - G["mild_u_N%d" % N])[same].max() <= 1e-7
    c = b.counters()
    assert c["solves"] == 16 and c["unsolved"] == 0 and c["iters"] == i["iter"].sum()
    b.close()


def test_update_matches_oracle_full_solution():
    from oracle import binding as OB
    from robobee3d_b200.dirtran import BatchLTVMPC
    N = 5
    ys = dt_util.mild_states(8, seed=77)
    b = BatchLTVMPC(8, N=N, max_iter=200000)
    b.set_state(xr=XR_HOVER)
    for s in range(2):      # second call: warm start, previous input as linearisation point, adapted rho kept
        u = b.update(ys)
        x, y = b.solution()
        i = b.info()
        for e in range(8):
            if s == 0:
                o = OB.OracleDirTran(N=N, max_iter=200000)
                ys_o = [] if e == 0 else ys_o
                ys_o.append(o)
            o = ys_o[e]
            uo, st = o.update(ys[e], XR_HOVER)
            xo, yo, io = o.solution()
            assert st == 1 and i["status"][e] == 1
            assert np.abs(x[e] - xo).max() <= 20 * 1e-5 * max(1.0, np.abs(xo).max())
            assert np.abs(y[e] - yo).max() <= 20 * 1e-5 * max(1.0, np.abs(yo).max())
            assert abs(i["rho"][e] - io[4]) <= 1e-6 * io[4] or i["iter"][e] != io[0]
        ys = ys + 1e-3
    b.close()


def test_hover_run_matches_golden():
    """the reference's 'hover' run (solves of up to 63k iterations in the transient), 32 copies, update + plant on the
    device through dtBatchRollout, one step per call so that every state can be compared"""
    from robobee3d_b200.dirtran import BatchLTVMPC
    B = 32
    b = BatchLTVMPC(B, N=5, max_iter=1000000)
    b.set_state(y=Y0_HOVER, xr=XR_HOVER)
    for s in range(12):
        b.rollout(1)
        y, c = b.state()
        i = b.info()
        assert (i["status"] == 1).all()
        assert np.abs(y - y[0]).max() == 0 and np.abs(c - c[0]).max() == 0     # identical envs stay identical
        # every solve is converged to eps = 1e-5 relative to inputs of order 10, and the 10 rad/s spin of the first
        # steps amplifies the difference between two converged iterates; the run contracts onto the goal afterwards
        assert np.abs(y[0] - G["hover_X"][s]).max() <= (3e-2 if s < 10 else 5e-3), s
        if s < 4:
            assert abs(i["iter"][0] - G["hover_iters"][s]) <= 0.05 * G["hover_iters"][s]
    b.rollout(37)
    y, c = b.state()
    assert np.abs(y[0] - G["hover_X"][-1]).max() <= 2e-4
    # the input at rest: two solves converged to eps = 1e-5 (relative to |u0| ~ 10 bounds) differ by O(TOL_U)
    assert (np.abs(c[0] - G["hover_U"][-1]) <= TOL_U).all()
    b.close()


def test_rollout_equals_update_plus_host_plant():
    """dtBatchRollout (everything on the device) == dtBatchUpdate through host buffers + the oracle's plant"""
    from oracle import binding as OB
    from robobee3d_b200.dirtran import BatchLTVMPC
    B, N = 64, 5
    ys = dt_util.mild_states(B, seed=5)
    a = BatchLTVMPC(B, N=N); a.set_state(y=ys, xr=XR_HOVER)
    a.rollout(4)
    ya, ca = a.state()
    b = BatchLTVMPC(B, N=N); b.set_state(xr=XR_HOVER)
    o = OB.OracleDirTran(N=N)
    y = ys.copy()
    for s in range(4):
        u = b.update(y)
        y = np.stack([o.plant(y[e], u[e]) for e in range(B)])
    # device plant (fused multiply-adds) against the oracle's plant, through 4 solves of ~100 iterations each
    assert np.abs(ca - u).max() <= 1e-8
    assert np.abs(ya - y).max() <= 1e-8
    a.close(); b.close()


def test_fused_rollout_is_bit_identical_to_single_steps():
    """dtBatchRollout(k) runs the k steps of a lane as one loop over ADMM iterations (lanes of a warp drift apart in
    step count); it must give exactly what k calls of dtBatchRollout(1) give, solve for solve"""
    from robobee3d_b200.dirtran import BatchLTVMPC
    B, N, k = 96, 10, 6
    ys = dt_util.mild_states(B, seed=21)
    a = BatchLTVMPC(B, N=N); a.set_state(y=ys, xr=XR_HOVER); a.rollout(k)
    b = BatchLTVMPC(B, N=N); b.set_state(y=ys, xr=XR_HOVER)
    for s in range(k):
        b.rollout(1)
    assert np.array_equal(a.state()[0], b.state()[0]) and np.array_equal(a.state()[1], b.state()[1])
    ca, cb = a.counters(), b.counters()
    assert ca == cb and ca["solves"] == B * k
    assert np.array_equal(a.solution()[0], b.solution()[0])
    a.close(); b.close()


def test_ragged_batch_and_reset():
    from robobee3d_b200.dirtran import BatchLTVMPC
    ys = dt_util.mild_states(33, seed=9)      # not a multiple of the warp size
    b = BatchLTVMPC(33, N=3)
    b.set_state(xr=XR_HOVER)
    u1 = b.update(ys)
    b.reset(); b.set_state(xr=XR_HOVER)
    u2 = b.update(ys)
    assert np.array_equal(u1, u2)
    s = BatchLTVMPC(1, N=3); s.set_state(xr=XR_HOVER)
    assert np.array_equal(s.update(ys[32:33])[0], u1[32])
    b.close(); s.close()


def test_ltvmpc_drop_in():
    """the reference's call sequence, mpc_thrust_strokedev.py:124,171"""
    from robobee3d_b200 import dirtran
    model = dirtran.PlanarThrustStrokeDev(rescale=True)
    model.dt = 0.03
    wx = np.array([100, 100, 1, 10, 10, 0.1]) / 0.03
    ltvmpc = dirtran.LTVMPC(model, 5, wx, [0.01, 0.01], verbose=False, polish=False, scaling=0, eps_rel=1e-5, eps_abs=1e-5,
                            max_iter=1000000, kdamping=0)
    ys = dt_util.mild_states(16, seed=5)
    ctrl = ltvmpc.update(ys[3].copy(), XR_HOVER, costMode=dirtran.FINAL)
    assert np.abs(ctrl - G["mild_u_N5"][3]).max() <= 1e-7
    assert ltvmpc.niter == G["mild_iters_N5"][3] and ltvmpc.prevSol.shape == (46,)
    np.testing.assert_array_equal(ltvmpc.prevSol[36:38], ctrl)       # mpc.py:89
    with pytest.raises(ValueError):
        bad = dirtran.LTVMPC(model, 5, wx, [0.01, 0.01], scaling=0, eps_rel=1e-5, eps_abs=1e-5, max_iter=25)
        bad.update(Y0_HOVER.copy(), XR_HOVER, costMode=dirtran.FINAL)


def test_status_max_iter_reported():
    from robobee3d_b200.dirtran import BatchLTVMPC
    b = BatchLTVMPC(4, N=5, max_iter=50)
    b.set_state(xr=XR_HOVER)
    b.update(np.tile(Y0_HOVER, (4, 1)))
    i = b.info()
    assert (i["iter"] == 50).all() and (i["status"] == -2).all()
    assert b.counters()["unsolved"] == 4
    b.close()
