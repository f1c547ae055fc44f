"""GPU parity tests (run on the B200 box): the CUDA path, called through the C-ABI, against
 (a) golden vectors produced by the reference's own C code (tests/golden/*.npz, make_golden.py),
 (b) the CPU oracle (oracle/liboracle_f32.so, pinned to the reference by test_oracle_vs_ref.py)
     on fresh seeded inputs, and
 (c) size-independent properties at the BASELINE sizes.
Nothing here reads /root/reference."""
import os

import numpy as np
import pytest

import util
from util import GOLDEN, TOL_QPDATA, TOL_X_REL, TOL_Y_REL

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLDEN, "umpc_seq.npz"))


def _close_rel(a, b, tol, what):
    a, b = np.asarray(a, float), np.asarray(b, float)
    scale = np.maximum(1.0, np.abs(b).max(axis=-1, keepdims=True))
    err = (np.abs(a - b) / scale).max()
    assert err <= tol, "%s: rel err %.3e > %.1e" % (what, err, tol)
    return err


def test_library_loads_on_gpu(cuda_required):
    from robobee3d_b200 import _lib
    L = _lib.load()
    assert all(hasattr(L, s) for s in _lib.EXPORTS)


def test_batch_update_matches_reference_golden(cuda_required, gold):
    """3 consecutive umpcUpdate calls per env (cold start, then two warm starts with the stale-scaling
    quirk) reproduce the reference C controller: QP data, solution x / y, uquad, accdes, status."""
    from robobee3d_b200.batch import BatchUMPC
    B, ncall = gold["p"].shape[:2]
    b = BatchUMPC(B)
    b.qpdata()  # switch QP-data output on
    for c in range(ncall):
        aT0 = gold["aT0"][:, c].astype(np.float32)
        uq, acc = b.update(gold["p"][:, c], gold["R"][:, c], gold["dq"][:, c], gold["pdes"][:, c], gold["dpdes"][:, c],
                           gold["sdes"][:, c], aT0)
        l, u, q, Px, Ax = b.qpdata()
        # call 0 has identical inputs -> float32 rounding only; later calls build on T0 += x[36] of the
        # previous (tolerance-level) solution, so T0-dependent entries inherit that tolerance
        tol_data = TOL_QPDATA if c == 0 else 1e-3
        for name, got in (("l", l), ("u", u), ("q", q), ("Px", Px), ("Ax", Ax)):
            _close_rel(got, gold[name][:, c], tol_data, "call %d QP data %s" % (c, name))
        s = b.solve_info(full=True)
        _close_rel(s["x"], gold["x"][:, c], TOL_X_REL, "call %d x" % c)
        _close_rel(s["y"], gold["y"][:, c], TOL_Y_REL, "call %d y" % c)
        tol_sc = 1e-4 if c == 0 else 2e-3  # D, E depend on T0 (dt*T0 entries) and on c through q
        _close_rel(s["D"], gold["D"][:, c], tol_sc, "call %d D" % c)
        _close_rel(s["E"], gold["E"][:, c], tol_sc, "call %d E" % c)
        assert (np.abs(uq - gold["uquad"][:, c]).max(axis=1) <= TOL_X_REL * np.maximum(1.0, np.abs(gold["x"][:, c]).max(axis=1))).all()
        # accdes = (x[18:24] - dq0)/dt: same absolute tolerance as x, divided by dt
        assert np.abs(acc - gold["accdes"][:, c]).max() <= TOL_X_REL * np.abs(gold["x"][:, c]).max() / 5.0
        assert (s["iter"] == 50).all()
        # status is a threshold test on the residuals of a TRUNCATED iterate.  Where the reference says "solved" we must
        # too.  The un-converged ones sit between "solved inaccurate" (10x the tolerance) and "max iter" by their DUAL
        # residual, which after 50 float32 iterations differs by factors between implementations (the reference's own
        # float32 and float64 builds disagree the same way): there, and only there, the two labels may swap -- measured
        # 2 of the 144 golden QPs on the host build of this code.
        ref_st = gold["info"][:, c, 1].astype(int)
        differ = s["status"] != ref_st
        assert not differ[ref_st == 1].any(), "status differs on a QP the reference solved"
        assert set(s["status"][differ]) <= {2, -2} and set(ref_st[differ]) <= {2, -2}
        assert differ.mean() <= 0.05, differ.mean()
        _close_rel(s["pri_res"][:, None], gold["info"][:, c, 2:3], 1e-3, "pri_res")
    b.close()


def test_batch_update_matches_oracle_fresh_inputs(cuda_required):
    """config 2 at 4096 QPs: cold solve + warm repeat against the CPU oracle on a 256-env subset."""
    from oracle import binding as OB
    from robobee3d_b200.batch import BatchUMPC
    B, nsub = 4096, 256
    p, R, dq = util.config2_states(B, seed=7)
    z3 = np.zeros((B, 3)); e3 = np.tile([0, 0, 1.0], (B, 1))
    b = BatchUMPC(B)
    outs = []
    for c in range(2):
        aT0 = np.full(B, 9.81e-3 if c == 0 else -1.0, np.float32)
        uq, acc = b.update(p, R, dq, z3, z3, e3, aT0)
        outs.append((uq.copy(), acc.copy(), b.solve_info(full=True)))
    sub = np.linspace(0, B - 1, nsub).astype(int)
    for e in sub:
        o = OB.OracleUMPC("f32")
        for c in range(2):
            uqo, acco = o.update(p[e], R[e], dq[e], z3[e], z3[e], e3[e], 9.81e-3 if c == 0 else -1.0)
            xo, yo, _ = o.solution()
            _close_rel(outs[c][2]["x"][e], xo, TOL_X_REL, "env %d call %d x" % (e, c))
            # uquad = (T0 + x[36], x[37], x[38]): tolerance relative to |x|_inf like x itself
            assert np.abs(outs[c][0][e] - uqo).max() <= TOL_X_REL * max(1.0, np.abs(xo).max())
    # property at full size: every output finite, thrust rows' z inside the box after projection
    s = outs[1][2]
    assert np.isfinite(s["x"]).all() and np.isfinite(s["y"]).all()
    b.close()


def test_batch_independent_of_batch_composition(cuda_required):
    """env e's result does not depend on who else is in the batch or on the launch shape (shard invariance)."""
    from robobee3d_b200.batch import BatchUMPC
    p, R, dq = util.config2_states(1000, seed=3)
    z3 = np.zeros((1000, 3)); e3 = np.tile([0, 0, 1.0], (1000, 1))
    aT0 = np.full(1000, 9.81e-3, np.float32)
    b1 = BatchUMPC(1000)
    u1, a1 = b1.update(p, R, dq, z3, z3, e3, aT0)
    lo, hi = 333, 777
    b2 = BatchUMPC(hi - lo)
    b2.set_launch(32)
    u2, a2 = b2.update(p[lo:hi], R[lo:hi], dq[lo:hi], z3[lo:hi], z3[lo:hi], e3[lo:hi], aT0[lo:hi])
    assert np.array_equal(u1[lo:hi], u2) and np.array_equal(a1[lo:hi], a2)
    b1.close(); b2.close()


def test_dropin_single_controller_matches_golden(cuda_required, gold):
    """uprightmpc2py.UprightMPC2C (the pybind module's surface) on env 0's three-call sequence."""
    from robobee3d_b200 import uprightmpc2py
    d = util.load_constants()["umpc_defaults"]
    up = uprightmpc2py.UprightMPC2C(d["dt"], d["g"], d["TtoWmax"], d["ws"], d["wds"], d["wpr"], d["wpf"], d["wvr"], d["wvf"],
                                    d["wthrust"], d["wmom"], d["Ib"], d["maxIter"])
    for e in (0, 5):
        up = uprightmpc2py.UprightMPC2C(d["dt"], d["g"], d["TtoWmax"], d["ws"], d["wds"], d["wpr"], d["wpf"], d["wvr"],
                                        d["wvf"], d["wthrust"], d["wmom"], d["Ib"], d["maxIter"])
        for c in range(gold["p"].shape[1]):
            uq, acc = up.update(gold["p"][e, c], gold["R"][e, c], gold["dq"][e, c], gold["pdes"][e, c], gold["dpdes"][e, c],
                                gold["sdes"][e, c], float(gold["aT0"][e, c]))
            assert np.abs(uq - gold["uquad"][e, c]).max() <= TOL_X_REL * max(1.0, np.abs(gold["x"][e, c]).max())
            l, u, q = up.vectors()
            Px, Ax, idx = up.matrices()
            tol_data = TOL_QPDATA if c == 0 else 1e-3
            _close_rel(l, gold["l"][e, c], tol_data, "l"); _close_rel(u, gold["u"][e, c], tol_data, "u")
            _close_rel(q, gold["q"][e, c], tol_data, "q"); _close_rel(Ax, gold["Ax"][e, c], tol_data, "Ax")
            assert np.array_equal(idx, gold["Ax_idx"])
            x, y = up.solution()
            _close_rel(x, gold["x"][e, c], TOL_X_REL, "x")
            assert up.info()["iter"] == 50


@pytest.mark.parametrize("zerocopy", ["1", "0"])
def test_host_buffer_calls_pinned_paths_equal_pageable_path(cuda_required, zerocopy, monkeypatch):
    """umpcBatchUpdate / umpcBatchEnvStep with PINNED host buffers run ONE zero-copy kernel on the caller's arrays
    (umpc_hostio_kernel; default) or, with UMPC_E2E_ZEROCOPY=0, replay a captured CUDA graph of the chunk pipeline;
    with pageable numpy buffers they issue the copies directly.  Same solver code, same results -- bit for bit -- over
    several calls (in-place state buffers, graph replay), after a change of buffers, and for ragged batch sizes."""
    import torch
    from robobee3d_b200.batch import BatchUMPC
    monkeypatch.setenv("UMPC_E2E_ZEROCOPY", zerocopy)
    for B in (20000, 4097, 130):
        p, R, dq = util.config2_states(B, seed=5)
        Rc = np.ascontiguousarray(R.transpose(0, 2, 1).reshape(B, 9), np.float32)
        f = lambda a: np.ascontiguousarray(a, np.float32)
        z3, e3 = f(np.zeros((B, 3))), f(np.tile([0, 0, 1.0], (B, 1)))
        pin = lambda a: torch.from_numpy(np.array(a, np.float32, copy=True)).pin_memory()
        bg, bd = BatchUMPC(B), BatchUMPC(B)
        # closed loop through the env-step call: state buffers are overwritten in place
        sp, sR, sdq = pin(p), pin(Rc), pin(dq)
        uq, acc = pin(np.zeros((B, 3))), pin(np.zeros((B, 6)))
        hpd, hdp, hsd = pin(z3), pin(z3), pin(e3)
        np_p, np_R, np_dq = f(p).copy(), Rc.copy(), f(dq).copy()
        np_uq, np_acc = np.zeros((B, 3), np.float32), np.zeros((B, 6), np.float32)
        for step in range(4):
            bg.env_step_raw(uq, acc, sp, sR, sdq, hpd, hdp, hsd)
            bd.env_step_raw(np_uq, np_acc, np_p, np_R, np_dq, z3, z3, e3)
            assert np.array_equal(uq.numpy(), np_uq) and np.array_equal(acc.numpy(), np_acc), (B, step)
            assert np.array_equal(sp.numpy(), np_p) and np.array_equal(sR.numpy(), np_R) and np.array_equal(sdq.numpy(), np_dq)
        # different buffers -> the graph is captured again; update call (no plant step), with and without actualT0
        uq2, acc2 = pin(np.zeros((B, 3))), pin(np.zeros((B, 6)))
        aT0 = np.full(B, 9.81e-3, np.float32)
        for a_pin, a_np in ((pin(aT0), aT0), (None, None)):
            bg.update_raw(sp, sR, sdq, hpd, hdp, hsd, a_pin, uq2, acc2)
            bd.update_raw(np_p, np_R, np_dq, z3, z3, e3, a_np, np_uq, np_acc)
            assert np.array_equal(uq2.numpy(), np_uq) and np.array_equal(acc2.numpy(), np_acc)
        bg.close(); bd.close()


@pytest.mark.parametrize("B", [130, 4097])
def test_update_dev_ragged_batch_stays_in_bounds(cuda_required, B):
    """umpcBatchUpdateDev on device arrays of EXACTLY B rows (B not a multiple of 256): the padded lanes of the launch
    must neither read nor write past them.  The arrays sit inside guard-filled allocations; results equal update()."""
    import torch
    from robobee3d_b200.batch import BatchUMPC
    p, R, dq = util.config2_states(B, seed=21)
    Rc = np.ascontiguousarray(R.transpose(0, 2, 1).reshape(B, 9), np.float32)
    z3 = np.zeros((B, 3), np.float32); e3 = np.tile(np.float32([0, 0, 1.0]), (B, 1))
    aT0 = np.full(B, 9.81e-3, np.float32)
    G = 1024  # guard elements on both sides
    def guarded(a, fill):
        a = np.ascontiguousarray(a, np.float32)
        t = torch.full((a.size + 2 * G,), fill, dtype=torch.float32, device="cuda")
        t[G:G + a.size] = torch.from_numpy(a.ravel()).cuda()
        return t, t[G:G + a.size].view(*a.shape)
    ins = [guarded(a, 7.5e8) for a in (p, Rc, dq, z3, z3, e3, aT0)]
    outs = [guarded(np.zeros((B, k), np.float32), -3.25) for k in (3, 6)]
    b1, b2 = BatchUMPC(B), BatchUMPC(B)
    for call in range(2):
        # ordered against torch's current stream: inputs written just above by torch ops, outputs read right after
        b1.update_dev(outs[0][1], outs[1][1], *[v for _, v in ins])
        uq_d, acc_d = outs[0][1].cpu().numpy(), outs[1][1].cpu().numpy()
        uq, acc = b2.update(p, R, dq, z3, z3, e3, aT0)
        assert np.array_equal(uq_d, uq) and np.array_equal(acc_d, acc), (B, call)
        for full, view in outs:
            f = full.cpu().numpy()
            assert (f[:G] == -3.25).all() and (f[G + view.numel():] == -3.25).all(), "write past the caller's array"
    b1.close(); b2.close()


def test_wlqp_and_funapprox_exports(cuda_required):
    """wlqpInit / wlqpUpdate (wlcontroller/c/wlqp.h:29-31) and funApproxInit/F/Df (funapprox.h:31-35) as a C caller
    binds them: same results as wlControllerUpdate on the scripted demand, F/Df equal to the closed form."""
    import ctypes as C
    from robobee3d_b200 import _lib
    L = _lib.load()
    c = util.load_constants()
    g = np.load(os.path.join(GOLDEN, "wl_seq.npz"))
    FP = C.POINTER(C.c_float)
    f32 = lambda a: np.ascontiguousarray(a, np.float32)
    fp = lambda a: a.ctypes.data_as(FP)
    po = f32(c["popts_waypoint"])
    # FunApprox
    rng = np.random.default_rng(0)
    for i in range(6):
        fa = _lib.FunApprox_t()
        L.funApproxInit(C.byref(fa), fp(po[15 * i:15 * i + 15].copy()))
        a0, a1 = po[15 * i], po[15 * i + 1:15 * i + 5]
        A2 = np.zeros((4, 4)); A2[np.triu_indices(4)] = po[15 * i + 5:15 * i + 15]; A2 = A2 + A2.T - np.diag(np.diag(A2))
        assert np.allclose(np.array(fa.A2).reshape(4, 4), A2.astype(np.float32))
        x = f32(rng.uniform(-1, 1, 4) * [100, 0.5, 0.2, 0.1] + [140, 0, 0, 0])
        fv = L.funApproxF(C.byref(fa), fp(x))
        df = np.zeros(4, np.float32)
        L.funApproxDf(fp(df), C.byref(fa), fp(x))
        ref_f = a0 + a1 @ x + 0.5 * x @ A2 @ x
        assert abs(fv - ref_f) <= 1e-5 * max(1.0, abs(ref_f))
        assert np.abs(df - (a1 + A2 @ x)).max() <= 1e-5 * max(1.0, np.abs(a1 + A2 @ x).max())
    # WLQP object: u0 carried by the caller (wlControllerUpdate does exactly that, wlcontroller.c:35-45)
    L.wlqpSetWrenchModel(fp(po))
    q = _lib.WLQP_t()
    L.wlqpInit(C.byref(q), C.c_float(1000.0))
    assert abs(q.U0[0] - 5.0) < 1e-6 and q.Q[0] == 1.0 and abs(q.Q[35] - 0.1) < 1e-7 and q.umin[0] == 90.0
    u0 = f32(c["wl_defaults"]["u0"]); h0 = f32(g["h0"])
    n = 120
    scale = np.abs(g["popts_waypoint_qp_u"][:n]).max(axis=0)
    for i in range(1, n + 1):
        u = np.zeros(4, np.float32)
        L.wlqpUpdate(C.byref(q), fp(u), fp(u0), fp(h0), fp(f32(g["pdotdes"][i])))
        assert (np.abs(u - g["popts_waypoint_qp_u"][i - 1]) / scale).max() < 2e-3, i
        u0 = u.copy()
    assert np.isfinite(np.array(q.w0)).all() and abs(q.w0[2]) > 0


def test_dropin_rereads_public_fields(cuda_required, gold):
    """the reference re-reads UprightMPC_t's weights / dt / Tmax / T0 on every umpcUpdate (uprightmpc2.c:126-207):
    editing them between calls must act like a controller built with the new values."""
    import ctypes as C
    from robobee3d_b200 import _lib
    L = _lib.load()
    d = util.load_constants()["umpc_defaults"]
    FP = C.POINTER(C.c_float)
    f32 = lambda a: np.ascontiguousarray(a, np.float32)
    fp = lambda a: a.ctypes.data_as(FP)
    def mk(wpr):
        up = _lib.UprightMPC_t()
        L.umpcInit(C.byref(up), d["dt"], d["g"], d["TtoWmax"], d["ws"], d["wds"], wpr, d["wpf"], d["wvr"], d["wvf"],
                   d["wthrust"], d["wmom"], fp(f32(d["Ib"])), d["maxIter"])
        return up
    def call(up, e, T0=None):
        if T0 is not None:
            up.T0 = T0
        uq, acc = np.zeros(3, np.float32), np.zeros(6, np.float32)
        L.umpcUpdate(C.byref(up), fp(uq), fp(acc), fp(f32(gold["p"][e, 0])), fp(f32(gold["R"][e, 0].T)), fp(f32(gold["dq"][e, 0])),
                     fp(f32(gold["pdes"][e, 0])), fp(f32(gold["dpdes"][e, 0])), fp(f32(gold["sdes"][e, 0])), C.c_float(-1.0))
        return uq, acc
    a, bq = mk(d["wpr"]), mk(3.0)
    for i in range(3):
        a.Qyr[i] = 3.0          # edit the running controller's position weight
    ua, _ = call(a, 1, T0=9.81e-3)
    ub, _ = call(bq, 1, T0=9.81e-3)
    assert np.array_equal(ua, ub)
    assert a.Px_data[0] == 3.0 and abs(a.T0 - ua[0]) == 0.0
    u0, _ = call(mk(d["wpr"]), 1, T0=9.81e-3)
    assert not np.array_equal(u0, ua)
    for up in (a, bq):
        L.umpcFree(C.byref(up))


def test_c_program_links_and_runs(cuda_required, tmp_path, gold):
    """tests/cabi/link_check.c (plain C, links libumpc_b200.so through include/umpc_b200.h) runs three umpcUpdate calls;
    its outputs equal the ctypes drop-in on the same inputs."""
    import subprocess
    from robobee3d_b200 import uprightmpc2py
    root = util.ROOT
    libdir = os.path.join(root, "robobee3d_b200")
    exe = str(tmp_path / "link_check")
    subprocess.run(["gcc", "-O1", "-I" + os.path.join(root, "include"), os.path.join(root, "tests", "cabi", "link_check.c"),
                    "-L" + libdir, "-lumpc_b200", "-Wl,-rpath," + libdir, "-o", exe], check=True)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    lines = [l.split() for l in r.stdout.splitlines() if l.startswith("umpcUpdate")]
    assert len(lines) == 3
    d = util.load_constants()["umpc_defaults"]
    up = uprightmpc2py.UprightMPC2C(d["dt"], d["g"], d["TtoWmax"], d["ws"], d["wds"], d["wpr"], d["wpf"], d["wvr"], d["wvf"],
                                    d["wthrust"], d["wmom"], d["Ib"], d["maxIter"])
    for k, l in enumerate(lines):
        uq, acc = up.update([1.0, -2.0, 3.0], np.eye(3), [0.1, 0, 0, 0, 0, 0], [0, 0, 0], [0, 0, 0], [0, 0, 1.0], 9.81e-3 if k == 0 else -1.0)
        got = np.array([float(v) for v in l[5:8]])
        assert np.array_equal(got.astype(np.float32), uq.astype(np.float32)), (k, got, uq)


def test_env_step_on_pinned_buffers_follows_the_rollout_info_switch(cuda_required):
    """umpcBatchEnvStep on pinned host buffers (one zero-copy kernel) is a closed-loop step: like a rollout it does not
    evaluate the per-solve residuals / status unless umpcBatchSetRolloutInfo(b, 1) asks for them -- outputs are the same
    bits either way, and equal to the pageable path's"""
    import torch
    from robobee3d_b200.batch import BatchUMPC
    B = 1000
    p, R, dq = util.config2_states(B, seed=11)
    Rc = np.ascontiguousarray(R.transpose(0, 2, 1).reshape(B, 9), np.float32)
    f = lambda a: np.ascontiguousarray(a, np.float32)
    pin = lambda a: torch.from_numpy(np.array(a, np.float32, copy=True)).pin_memory()
    z3, e3 = f(np.zeros((B, 3))), f(np.tile([0, 0, 1.0], (B, 1)))
    outs = []
    for info_on in (0, 1):
        b = BatchUMPC(B)
        b.set_rollout_info(info_on)
        sp, sR, sdq = pin(p), pin(Rc), pin(dq)
        uq, acc = pin(np.zeros((B, 3))), pin(np.zeros((B, 6)))
        for _ in range(3):
            b.env_step_raw(uq, acc, sp, sR, sdq, pin(z3), pin(z3), pin(e3))
        st = b.solve_info()["status"]
        assert (st == (1 if info_on else 0)).all() or (info_on and np.isin(st, (1, 2, -2)).all()), (info_on, np.unique(st))
        outs.append((uq.numpy().copy(), acc.numpy().copy(), sp.numpy().copy(), sR.numpy().copy(), sdq.numpy().copy()))
        b.close()
    for a, c in zip(*outs):
        assert np.array_equal(a, c)
