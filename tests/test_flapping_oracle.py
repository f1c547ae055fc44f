"""Row f2 on the CPU: the oracle of the flapping-wing plant against the golden fixture made by the reference's own
functions (tests/golden/make_golden_flapping.py), its multibody step against physics (PARITY UNPINNED for that part:
pybullet is absent), and the C-ABI's exports."""
import os
import re

import numpy as np
import pytest

import util
from util import GOLDEN, ROOT
from oracle import flapping_oracle as fo


def test_oracle_aerodynamics_actuator_filter_match_reference_functions():
    g = np.load(os.path.join(GOLDEN, "flapping_aero.npz"))
    assert np.allclose(g["consts"], [fo.CD0, fo.CDMAX, fo.CLMAX, fo.RHO, fo.RCOPNONDIM, fo.WPARAMS["Aw"], fo.WPARAMS["r1h"],
                                     fo.WPARAMS["r2h"], fo.WPARAMS["cbar"]], rtol=0, atol=0)
    for sgn, t in ((1, "L"), (-1, "R")):
        F, pc = fo.aerodynamics(g["theta"], g["dtheta"], sgn)
        assert np.abs(F - g["F_" + t]).max() <= 1e-15 and np.abs(pc - g["pcop_" + t]).max() <= 1e-14
        assert np.abs(fo.wrench(F, pc) - g["wrench_" + t]).max() <= 1e-13
    assert np.abs(fo.actuator_model(g["act_V"]) - g["act_F"]).max() <= 1e-15
    wf = fo.WingFlapFilter(0.16, 0.1, 6)
    assert len(wf.buf) == int(g["filt_len"][0])
    assert np.abs(np.array([wf.update(x) for x in g["filt_in"]]) - g["filt_out"]).max() <= 1e-14


def test_oracle_wing_dynamics_conserve_energy_and_are_mirror_symmetric():
    """free motion (no gravity, springs, aero, drive): kinetic energy constant to the integrator's order; the right wing
    is the left wing's mirror image"""
    rng = np.random.default_rng(3)
    q, dq = rng.uniform(-1, 1, (6, 2)), rng.uniform(-1, 1, (6, 2))
    z3 = np.zeros((6, 3))
    E0 = fo.kinetic_energy(1, q, dq)
    dt = 2e-4
    for _ in range(5000):
        dd = fo.wing_accel(1, q, dq, np.zeros_like(q), z3, z3, with_gravity=False)[0]
        dq = dq + dt * dd; q = q + dt * dq
    assert np.abs(fo.kinetic_energy(1, q, dq) - E0).max() <= 2e-5 * E0.max()
    # mirror in the x-z plane: force and lever arm reflected in y; the stroke coordinate keeps its value (the right stroke
    # axis is -z), the hinge coordinate changes sign (a rotation about y is its own mirror image, the right hinge axis is -y)
    F, pc = rng.standard_normal((6, 3)), rng.standard_normal((6, 3)) + [0, 8, 4]
    tau = rng.standard_normal((6, 2))
    m, h = np.array([1.0, -1.0, 1.0]), np.array([1.0, -1.0])
    aL = fo.wing_accel(1, q, dq, tau, F, pc)[0]
    aR = fo.wing_accel(-1, q * h, dq * h, tau * h, F * m, pc * m)[0]
    assert np.abs(aL - aR * h).max() <= 1e-10 * np.abs(aL).max()


def test_oracle_open_loop_is_a_flapping_robobee():
    """symmetric drive: both wings flap alike, no side force, lift above the 100 mg robot's weight (0.98 mN), and the row
    layout of RobobeeSim.openLoop / olAvgWrenchKinFeat"""
    sw = fo.open_loop([120.0], [120.0], [0.0], [0.17])
    assert sw.shape == (1, 1001, 14)          # `while simt < 100` with simt += 0.1 in double runs 1001 times
    last = sw[0, -58:]
    assert np.abs(last[:, 0] - last[:, 2]).max() <= 1e-9 and np.abs(last[:, 1] + last[:, 3]).max() <= 1e-9   # mirror image
    feats = fo.sweep_features([[120.0, 0.0, 0.17, 0.0, 0.0], [150.0, 0.2, 0.17, 0.1, 0.1]])
    assert feats.shape == (2, 20)
    w = feats[0, 5:11]
    assert 1.0 < w[2] < 3.0 and abs(w[1]) <= 1e-9 and abs(w[3]) <= 1e-9 and abs(w[5]) <= 1e-9
    assert 0.4 < feats[0, 11] < 0.9 and 0.8 < feats[0, 12] < 1.4 and feats[0, 19] == 0.5
    assert feats[1, 7] > feats[0, 7]          # more voltage, more lift
    assert abs(feats[1, 8]) > 0.5             # a left/right amplitude difference rolls


def test_flapping_header_is_exported_and_refuses_without_device():
    from robobee3d_b200 import flapping
    L = flapping._bind()
    src = re.sub(r"/\*.*?\*/", " ", open(os.path.join(ROOT, "include", "flapping_b200.h")).read(), flags=re.S)
    decl = sorted(set(re.findall(r"^\s*(?:const\s+)?(?:void|int|float|double|char)\s*\*?\s*(\w+)\s*\(", src, flags=re.M)))
    assert decl == sorted(flapping.EXPORTS), decl
    assert all(hasattr(L, s) for s in decl)
    p = flapping.default_params()                       # host arithmetic only
    assert (p.kstroke, p.khinge, p.bhinge, p.m_spar, p.c_mem[2]) == (65.0, 3.0, 2.0, 0.72, -2.0)
    assert flapping.open_loop_steps(100.0, 0.1) == 1001
    if not util.has_cuda():
        with pytest.raises(RuntimeError, match="no CUDA device|CUDA"):
            flapping.aerodynamics([[0.1, 0.2]], [[0.3, 0.4]], 1)
        with pytest.raises(RuntimeError, match="no CUDA device|CUDA"):
            flapping.open_loop(120, 120, 0, 0.17)
