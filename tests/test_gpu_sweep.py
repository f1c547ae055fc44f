"""GPU tests of per-environment cost weights and the batched weight sweep (the reference's gainTuningSims,
uprightmpc2.py:272-303) against tests/golden/gain_sweep.npz -- the reference's own controlTest + createMPC(**kwargs) +
logMetric run per grid point (make_golden_gainsweep.py).  Nothing here reads /root/reference."""
import os

import numpy as np
import pytest

import util
from util import GOLDEN

pytestmark = pytest.mark.gpu


def test_per_env_weights_equal_per_batch_weights(cuda_required):
    """one batch with per-env weights == separate batches created with those weights (float32 C path)"""
    from robobee3d_b200.batch import BatchUMPC
    B = 512
    p, R, dq = util.config2_states(B, seed=3)
    z3 = np.zeros((B, 3)); e3 = np.tile([0, 0, 1.0], (B, 1))
    aT0 = np.full(B, 9.81e-3, np.float32)
    sets = [dict(wpr=0.5, wvr=500.0, wpf=2.5, wvf=1000.0), dict(wpr=5.25, wvr=5250.0, wpf=26.25, wvf=10500.0, ws=20.0, wmom=2e-2)]
    b = BatchUMPC(B)
    sel = (np.arange(B) // 32) % 2   # whole warps alternate between the two weight sets
    per = {k: np.where(sel == 0, sets[0].get(k, getattr(b.params, k)), sets[1].get(k, getattr(b.params, k)))
           for k in ("ws", "wds", "wpr", "wpf", "wvr", "wvf", "wthrust", "wmom")}
    b.set_weights(**per)
    for call in range(2):
        uq, acc = b.update(p, R, dq, z3, z3, e3, aT0 if call == 0 else None)
        x = b.solve_info(full=True)["x"]
        for k, ws in enumerate(sets):
            if call == 0:
                ws["_b"] = BatchUMPC(B, **{k_: v for k_, v in ws.items() if not k_.startswith("_")})
            uqk, acck = ws["_b"].update(p, R, dq, z3, z3, e3, aT0 if call == 0 else None)
            xk = ws["_b"].solve_info(full=True)["x"]
            m = sel == k
            assert np.abs(uq[m] - uqk[m]).max() <= 1e-5 * max(1.0, np.abs(uqk).max())
            assert np.abs(x[m] - xk[m]).max() <= 1e-5 * max(1.0, np.abs(xk).max())
    # back to the batch-wide weights
    b.set_weights()
    b.reset()
    b0 = BatchUMPC(B)
    uq, _ = b.update(p, R, dq, z3, z3, e3, aT0)
    uq0, _ = b0.update(p, R, dq, z3, z3, e3, aT0)
    assert np.array_equal(uq, uq0)
    for ws in sets:
        ws["_b"].close()
    b.close(); b0.close()


def test_gain_sweep_matches_reference(cuda_required, tmp_path):
    from robobee3d_b200.sweeps import gain_tuning_sims
    g = np.load(os.path.join(GOLDEN, "gain_sweep.npz"))
    out_file = str(tmp_path / "mpc_wpos.npz")
    r = gain_tuning_sims([0.5, 10], [0.5e3, 10e3], npts=3, tend=float(g["tend"]), taulim=float(g["taulim"]), save=out_file)
    assert np.array_equal(r["xv"], g["xv"]) and np.array_equal(r["yv"], g["yv"])
    saved = np.load(out_file)
    assert sorted(saved.files) == ["costs", "efforts", "xv", "yv"]   # the schema gainTuningPlots reads (uprightmpc2.py:303)
    stable = g["costs"] < 100.0
    assert stable.sum() == 8
    rc, re_ = r["costs"][stable] / g["costs"][stable] - 1, r["efforts"][stable] / g["efforts"][stable] - 1
    assert np.abs(rc).max() < 2e-2 and np.abs(re_).max() < 3e-2, (rc, re_)
    # the unstable corner (high position weight, low velocity weight) diverges in the reference too
    assert (~np.isfinite(r["costs"][~stable]) | (r["costs"][~stable] > 1e3)).all()


def test_gain_sweep_float32_controller_runs_the_full_grid(cuda_required):
    """10 x 10 grid with the 50-iteration float32 controller (per-env weights on the TMEM kernel): finite where the
    converged sweep is stable, and the cost surface has the reference's shape (falls with the position weight)"""
    from robobee3d_b200.sweeps import gain_tuning_sims
    r = gain_tuning_sims([0.5, 10], [0.5e3, 10e3], npts=10, tend=300.0, converged=False, save=None)
    c = r["costs"]
    assert c.shape == (10, 10) and np.isfinite(c[:, 3:]).all()
    assert (c[0, 3:] > c[5, 3:]).all()
