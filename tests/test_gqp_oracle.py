"""CPU tests of the generic-QP rows (SURVEY 8 a17, a18 second instance, f1): the checker oracle.binding.OracleQP is pinned
to the reference's own OSQP (oracle/_ref, vendored 0.6.0 built in float64) and the committed fixtures are what the
reference's own Python functions produce today (tests/golden/make_golden_gqp.py)."""
import os

import numpy as np
import pytest

import util
from util import GOLDEN
from oracle import binding as OB


def _dense_A(gold, e, c):
    o = OB.OracleUMPC("f64")
    ok, row, col, idx = o.pattern()
    A = np.zeros((39, 45))
    A[row, col] = o_init_vals(o)
    A[row[idx], col[idx]] = gold["Ax"][e, c]
    return A


def o_init_vals(o):
    # constant entries of A: -1 on the own columns, +1 couplings, 1 on the thrust rows (template_controllers.py:28-63)
    ok, row, col, idx = o.pattern()
    v = np.zeros(111)
    for k, (r, c) in enumerate(zip(row, col)):
        if r < 18:
            v[k] = -1.0 if c == r else 1.0
        elif r < 36:
            v[k] = -1.0 if c == r else 1.0
        else:
            v[k] = 1.0
    return v


@pytest.mark.skipif(not OB.have_ref(), reason="oracle/_ref (the reference's own OSQP) not built here")
def test_generic_oracle_equals_reference_osqp_f64():
    """OracleQP (dense storage, Ruiz x10, adaptive rho) against the reference's vendored OSQP built in float64 on the
    golden UprightMPC2 QPs: same iterates, same iteration counts"""
    gold = np.load(os.path.join(GOLDEN, "umpc_seq.npz"))
    for e in (0, 7, 23):
        ref = OB.RefQP64()
        o = OB.OracleQP(45, 39, eps=1e-4)
        for c in range(2):
            A = _dense_A(gold, e, c)
            xr, yr, ir = ref.solve(gold["l"][e, c], gold["u"][e, c], gold["q"][e, c], gold["Px"][e, c], gold["Ax"][e, c], gold["Ax_idx"])
            xo, yo, io = o.update_solve(np.diag(gold["Px"][e, c].astype(float)), A, gold["q"][e, c], gold["l"][e, c], gold["u"][e, c])
            assert io["iter"] == ir["iter"] and io["status"] == ir["status"], (e, c, io, ir)
            # same iteration, different factorisation (dense LDL' here, QDLDL with its AMD order there): float64 rounding
            # amplified by the conditioning -- measured 6e-8 relative
            assert np.abs(xo - xr).max() <= 1e-6 * max(1.0, np.abs(xr).max())
            assert np.abs(yo - yr).max() <= 1e-6 * max(1.0, np.abs(yr).max())


@pytest.mark.skipif(not os.path.isdir("/root/reference"), reason="needs the reference sources")
def test_gqp_fixtures_reproduce_from_reference(tmp_path, monkeypatch):
    """the committed fixtures are what the reference's own functions (exec'd from /root/reference) give today"""
    import importlib.util
    spec = importlib.util.spec_from_file_location("mg", os.path.join(GOLDEN, "make_golden_gqp.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    monkeypatch.setattr(mg, "HERE", str(tmp_path))
    mg.mpc_osqp_quad(); mg.umpc_genN()
    for f in ("mpc_osqp_quad.npz", "umpc_genN.npz"):
        a, b = np.load(os.path.join(str(tmp_path), f)), np.load(os.path.join(GOLDEN, f))
        for k in b.files:
            assert np.array_equal(np.nan_to_num(a[k], posinf=9e99, neginf=-9e99), np.nan_to_num(b[k], posinf=9e99, neginf=-9e99)), (f, k)


def test_quad_kat_closed_loop_reaches_the_reference():
    g = np.load(os.path.join(GOLDEN, "mpc_osqp_quad.npz"))
    assert g["P"].shape == (92, 92) and g["A"].shape == (164, 92)       # n = 92, m = 164
    assert (g["iters"] == 25).all()
    assert abs(g["X"][-1][2] - 1.0) < 2e-3                               # z reaches the reference xr[2] = 1 (mpc_osqp.py:80)


def test_product_quad_problem_equals_reference_statements():
    """robobee3d_b200.gqp.mpc_osqp_problem() restates planar/mpc_osqp.py:29-103: identical to what the file's own statements build"""
    from robobee3d_b200.gqp import mpc_osqp_problem
    g = np.load(os.path.join(GOLDEN, "mpc_osqp_quad.npz"))
    m = mpc_osqp_problem()
    f = lambda a: np.nan_to_num(a, posinf=9e99, neginf=-9e99)
    for k in ("P", "A", "q", "l", "u", "Ad", "Bd"):
        assert np.array_equal(f(m[k]), f(g[k])), k
