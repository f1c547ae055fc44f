"""Row f2 on the GPU: flapping-wing plant through the C-ABI (include/flapping_b200.h) against the golden fixture of the
reference's own aerodynamics (pinned) and against the oracle's independently formulated multibody step (unpinned)."""
import os

import numpy as np
import pytest

from util import GOLDEN
from oracle import flapping_oracle as fo

pytestmark = pytest.mark.gpu


def test_aerodynamics_matches_reference_function(cuda_required):
    from robobee3d_b200 import flapping
    g = np.load(os.path.join(GOLDEN, "flapping_aero.npz"))
    for sgn, t in ((1, "L"), (-1, "R")):
        F, pc = flapping.aerodynamics(g["theta"], g["dtheta"], sgn)
        assert np.abs(F - g["F_" + t]).max() <= 1e-13 * max(1.0, np.abs(g["F_" + t]).max())
        assert np.abs(pc - g["pcop_" + t]).max() <= 1e-13 * np.abs(g["pcop_" + t]).max()


def test_open_loop_trajectory_matches_oracle(cuda_required):
    """RobobeeSim.openLoop's rows for an asymmetric drive: 1,001 steps of two coupled-joint wings, link-frame kernel
    against world-frame oracle"""
    from robobee3d_b200 import flapping
    args = (150.0 * 1.1, 150.0 * 0.9, 0.2, 0.17)
    qw = flapping.open_loop(*args, tendMS=100, h2=0.1)
    ref = fo.open_loop([args[0]], [args[1]], [args[2]], [args[3]], 100.0, [0.1])[0]
    assert qw.shape == ref.shape == (1001, 14)
    scale = np.abs(ref).max(0)
    assert (np.abs(qw - ref).max(0) <= 1e-7 * scale + 1e-12).all(), np.abs(qw - ref).max(0) / scale


def test_sweep_matches_oracle_and_feeds_the_wrench_map_fit(cuda_required):
    """sweepFile's grid (here 3 x 3 x 1 x 3 x 3 = 81 simulations in one launch) against the oracle's features; then the
    pipeline of sdab_num_wrenchmap.py on it: kinematics features -> wrench -> fitted popts"""
    from robobee3d_b200 import flapping, wrenchfit
    Vm, uo, fs, ud, h2 = np.linspace(90, 160, 3), np.linspace(-0.5, 0.5, 3), [0.17], np.linspace(-0.2, 0.2, 3), np.linspace(-0.2, 0.2, 3)
    res = flapping.sweep(Vm, uo, fs, ud, h2)
    assert res.shape == (81, 20)
    ref = fo.sweep_features(res[:, :5])
    assert np.abs(res[:, :5] - ref[:, :5]).max() == 0
    scale = np.abs(ref[:, 5:]).max(0)
    assert (np.abs(res[:, 5:] - ref[:, 5:]).max(0) <= 1e-6 * scale + 1e-9).all()
    # batch-composition independence: a sub-grid gives the same rows
    sub = flapping.sweep(Vm[:2], uo[1:2], fs, ud, h2[:1])
    for row in sub:
        k = np.where((np.abs(res[:, :5] - row[:5]).max(1) == 0))[0]
        assert len(k) == 1 and np.array_equal(res[k[0]], row)
    # downstream: unpackDat's columns (sdab_num_wrenchmap.py:100-116) -> xdata (Vmean, uoffs, udiff, h2), averaged wrench
    xdata = res[:, [0, 1, 3, 4]]
    popts, info = wrenchfit.fit_wrench_map(xdata, res[:, 5:11], return_info=True)
    assert np.isfinite(popts).all() and info["rank"] == 15          # a full 3^4 grid identifies all 15 parameters
    assert np.abs(info["fit"][:, 2] - res[:, 7]).max() <= 0.15 * np.abs(res[:, 7]).max()     # lift is close to quadratic in the inputs
