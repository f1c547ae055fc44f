"""tests/golden/make_golden.py -- generate the committed golden fixtures FROM THE REFERENCE ITSELF.

Run in the build container (needs /root/reference and oracle/_ref built by oracle/Makefile):
    python tests/golden/make_golden.py
Outputs (small, committed):
    constants.json   numeric constants parsed out of the reference's Python sources
    umpc_seq.npz     reference C controller (oracle/_ref/libref_umpc_f32.so, flavour A: float32, 50 fixed
                     iterations) on config-2 random states: 3 consecutive calls per env (cold, warm, warm)
    umpc_conv.npz    the same QPs solved by the reference's vendored OSQP built in float64 and run to
                     convergence with the Python path's settings (flavour B, libref_umpc_f64.so)
    rollout_c1.npz   config 1: controlTest hover flight, reference C controller + double plant
    cascade_wlqp.npz config 4 with the true WLQP: reference umpcUpdate -> reference wlControllerUpdate -> actualT0 feedback
    wl_seq.npz       WLCon (uprightmpc2/funapprox.c) and true WLQP (wlcontroller/c) on the scripted demand of
                     template/wlcontroller/testwlqppy.py:15-40, plus the cascade of config 4
The GPU box has no /root/reference: tests only read these files.
"""
import ast
import json
import os
import re
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
REF = "/root/reference/template"

from oracle import binding as B  # noqa: E402
import util  # noqa: E402


def parse_list(path, prefix):
    """first un-commented `popts = [...]` literal (may span lines) in a reference Python source"""
    src = open(path).read()
    for m in re.finditer(r"^[ \t]*" + re.escape(prefix), src, flags=re.M):
        i = src.index("[", m.start())
        j = src.index("]", i)
        return [float(v) for v in ast.literal_eval(src[i:j + 1])]
    raise RuntimeError("not found in " + path)


def constants():
    c = dict(
        popts_waypoint=parse_list(os.path.join(REF, "robobee_test_controllers.py"), "popts = ["),
        popts_testwlqp=parse_list(os.path.join(REF, "wlcontroller", "testwlqppy.py"), "popts = ["),
        source={"popts_waypoint": "template/robobee_test_controllers.py:71",
                "popts_testwlqp": "template/wlcontroller/testwlqppy.py:7",
                "umpc_defaults": "template/template_controllers.py:260-280",
                "wl_defaults": "template/robobee_test_controllers.py:76"},
        umpc_defaults={k: (list(v) if isinstance(v, tuple) else v) for k, v in B.UMPC_DEFAULTS.items()},
        wl_defaults={k: (list(v) if isinstance(v, tuple) else v) for k, v in B.WL_DEFAULTS.items()},
    )
    assert len(c["popts_waypoint"]) == 90 and len(c["popts_testwlqp"]) == 90
    with open(os.path.join(HERE, "constants.json"), "w") as f:
        json.dump(c, f, indent=1)
    return c


def umpc_seq(nenv=48, ncalls=3):
    out = {k: [] for k in ("p", "R", "dq", "pdes", "dpdes", "sdes", "aT0", "uquad", "accdes", "x", "y", "info",
                           "l", "u", "q", "Px", "Ax", "xs", "zs", "ys", "D", "E", "c")}
    idx = None
    states = [util.config2_states(nenv, seed=s) for s in range(ncalls)]
    rng = np.random.default_rng(123)
    pdes = rng.uniform(-20, 20, (ncalls, nenv, 3))
    dpdes = rng.uniform(-0.05, 0.05, (ncalls, nenv, 3))
    conv = {k: [] for k in ("x", "y", "iter", "status")}
    for e in range(nenv):
        ref = B.RefUMPC()
        q64 = B.RefQP64()
        rec = {k: [] for k in out}
        for c in range(ncalls):
            p, R, dq = states[c][0][e], states[c][1][e], states[c][2][e]
            # call 0: the config-2 setup (pdes=0, sdes=e3, actualT0=g); later calls: moving reference, internal T0
            pd = np.zeros(3) if c == 0 else pdes[c, e]
            dpd = np.zeros(3) if c == 0 else dpdes[c, e]
            sd = np.array([0, 0, 1.0]) if c < 2 else np.array([np.sin(0.2), 0, np.cos(0.2)])
            aT0 = 9.81e-3 if c == 0 else -1.0
            uq, acc = ref.update(p, R, dq, pd, dpd, sd, aT0)
            x, y, info = ref.solution()
            l, u, q, Px, Ax, idx = ref.qpdata()
            xs, zs, ys, D, E, cc = ref.iterates()
            vals = dict(p=p, R=R, dq=dq, pdes=pd, dpdes=dpd, sdes=sd, aT0=aT0, uquad=uq, accdes=acc, x=x, y=y,
                        info=[info["iter"], info["status"], info["pri_res"], info["dua_res"]],
                        l=l, u=u, q=q, Px=Px, Ax=Ax, xs=xs, zs=zs, ys=ys, D=D, E=E, c=cc)
            for k in rec:
                rec[k].append(np.asarray(vals[k], dtype=np.float64))
            if c == 0:
                xc, yc, ic = q64.solve(l, u, q, Px, Ax, idx)
                conv["x"].append(xc); conv["y"].append(yc); conv["iter"].append(ic["iter"]); conv["status"].append(ic["status"])
        for k in out:
            out[k].append(np.stack(rec[k]))
    np.savez_compressed(os.path.join(HERE, "umpc_seq.npz"), Ax_idx=idx,
                        **{k: np.stack(v).astype(np.float32 if k not in ("p", "R", "dq", "pdes", "dpdes", "sdes") else np.float64)
                           for k, v in out.items()})
    np.savez_compressed(os.path.join(HERE, "umpc_conv.npz"), **{k: np.asarray(v) for k, v in conv.items()})


def rollout_c1():
    res = {}
    for name, hl, tend in (("hl5", 5.0, 500.0), ("every", None, 100.0)):
        ref = B.RefUMPC()
        cfg = B.rollout_cfg(tend=tend, hlInterval=hl)
        r = B.rollout(ref.lib, None, ref.up, B.control_test_ic(), cfg)
        # keep every 10th sample to stay small
        res[name + "_y"] = r["y"][::10].astype(np.float32)
        res[name + "_u"] = r["u"][::10].astype(np.float32)
        res[name + "_metrics"] = np.array([r["err"], r["eff"], r["nsolve"]])
        res[name + "_final"] = r["final"]
    # tasks: helix with perturbation, flip, perch, straightAcc (flight_tasks.py), MPC every 5 ms, 300 ms
    tasks = dict(helix=(B.TASK_HELIX, (50.0, 1.0, 0.1, 1.0)), helixpert=(B.TASK_HELIX, (50.0, 1.0, 0.1, 1.0)), flip=(B.TASK_FLIP, (100.0, 150.0, 0, 0)),
                 perch=(B.TASK_PERCH, (200.0, 100.0, 150.0, 0.05)), straight=(B.TASK_STRAIGHT, (150.0, 0.1, 0, 0)))
    for name, (tid, tp) in tasks.items():
        ref = B.RefUMPC()
        cfg = B.rollout_cfg(tend=300.0, hlInterval=5.0, tpert=100.0 if name == "helixpert" else None)
        r = B.rollout(ref.lib, None, ref.up, B.control_test_ic(), cfg, task=tid, tp=tp)
        res[name + "_y"] = r["y"][::10].astype(np.float32)
        res[name + "_u"] = r["u"][::10].astype(np.float32)
        res[name + "_metrics"] = np.array([r["err"], r["eff"], r["nsolve"]])
        res[name + "_tp"] = np.asarray(tp, float)
    np.savez_compressed(os.path.join(HERE, "rollout_c1.npz"), **res)


def wl_seq(c):
    mb, g = 100.0, 9.81e-3
    h0 = np.array([0, 0, mb * g, 0, 0, 0.0])
    Nt = 500
    pd = np.zeros((Nt, 6)); pd[:, 2] = 1.0
    pd[101:, 4] = 1; pd[201:, 4] = -1; pd[301:, 4] = 2; pd[401:, 4] = -2
    res = dict(h0=h0, pdotdes=pd)
    for pname in ("popts_testwlqp", "popts_waypoint"):
        po = c[pname]
        con = B.RefWLCon(po)
        qp = B.RefWLQP(po)
        Ucon, Wcon, Uqp, Iqp = [], [], [], []
        for i in range(1, Nt):
            u1, w0 = con.update(h0, pd[i])
            Ucon.append(u1); Wcon.append(w0)
            Uqp.append(qp.update(h0, pd[i]))
            inf = qp.info()
            Iqp.append([inf["iter"], inf["status"]])
        res[pname + "_con_u"] = np.array(Ucon, np.float32)
        res[pname + "_con_w0"] = np.array(Wcon, np.float32)
        res[pname + "_qp_u"] = np.array(Uqp, np.float32)
        res[pname + "_qp_info"] = np.array(Iqp, np.int32)
    # cascade (config 4): reference umpcUpdate + wlConUpdate + actualT0 feedback, hover from the tilted IC
    ref = B.RefUMPC()
    con = B.RefWLCon(c["popts_waypoint"], lib=ref.lib)
    cfg = B.rollout_cfg(tend=200.0, hlInterval=1.0)
    r = B.rollout(ref.lib, None, ref.up, B.control_test_ic(), cfg, wl=con.wl, wlmode=1)
    res["cascade_y"] = r["y"][::10].astype(np.float32)
    res["cascade_u"] = r["u"][::10].astype(np.float32)
    res["cascade_metrics"] = np.array([r["err"], r["eff"], r["nsolve"], r["u4sum"]])
    np.savez_compressed(os.path.join(HERE, "wl_seq.npz"), **res)


def cascade_wlqp(c):
    """config 4 with the TRUE WLQP: the loop of robobee_test_controllers.py:162-171 --
    umpcUpdate (reference C, libref_umpc_f32.so) -> h0 = [R'(0,0,mb g);0], pdotdes = M0 accdes -> wlControllerUpdate
    (reference C, libref_wlqp_f32.so: wlcontroller/c + its own OSQP workspace) -> actualT0 = w0[2]/mb fed back,
    w0 = wlqp.w0 = w(u0 before the step) (wlqp.c:87) -- around the double plant of the oracle (quadrotorNLDyn).
    The two reference libraries each carry their own global `workspace`, so the loop runs here in Python."""
    import ctypes as C
    mb, g, Ib = 100.0, 9.81e-3, np.array([3333.0, 3333.0, 1000.0])
    res = {}
    for name, hl, tend, task, tp in (("hover", 1.0, 200.0, B.TASK_HELIX, (0.0, 0.0, 0.1, 0.0)),
                                     ("helix", 5.0, 300.0, B.TASK_HELIX, (50.0, 1.0, 0.1, 1.0))):
        ref = B.RefUMPC()
        wl = B.RefWLQP(c["popts_waypoint"])
        w0v = (C.c_float * 58).in_dll(wl.lib, "wlqp")      # WLQP_t: umin[4] umax[4] U0[4] w0[6] Q[36] (wlqp.h:21-25)
        orc = C.CDLL(os.path.join(ROOT, "oracle", "liboracle_f64.so"))
        orc.orc_dyn_step.argtypes = [B.DP, B.DP, B.DP, B.DP, C.c_double, B.DP, C.c_double]
        orc.orc_task_q.argtypes = [C.c_int, C.c_double, B.DP, B.DP, C.c_int, B.DP, B.DP, B.DP]
        p, R, dq = B.control_test_ic()
        p = p.copy(); Rc = np.ascontiguousarray(R.T).ravel().copy(); dq = dq.copy()   # column-major R
        pinit = p.copy()
        nt = int(np.ceil(tend / 0.2))
        uq = np.zeros(3); aT0 = -1.0; thl = 0.0; nsolve = 0
        Y = np.zeros((nt, 12)); U = np.zeros((nt, 3)); U4 = []
        err = eff = 0.0
        tpa = np.asarray(tp, float)
        for ti in range(nt):
            t = ti * 0.2
            if t - thl > hl:
                pdes, dpdes, sdes = np.zeros(3), np.zeros(3), np.zeros(3)
                orc.orc_task_q(int(task), t, B._p(pinit), B._p(tpa), 0, B._p(pdes), B._p(dpdes), B._p(sdes))
                Rm = Rc.reshape(3, 3).T
                uqn, acc = ref.update(p, Rm, dq, pdes, dpdes, sdes, aT0)
                h0 = np.r_[Rm.T @ np.array([0, 0, mb * g]), 0, 0, 0]
                pdd = np.r_[mb * acc[:3], Ib * acc[3:]]
                u4 = wl.update(h0, pdd)
                aT0 = float(w0v[14]) / mb                                  # w0[2] / mb
                U4.append(np.r_[u4, [w0v[12 + i] for i in range(6)]])
                uq = uqn.copy(); thl = t; nsolve += 1
            uq[1:] = np.clip(uq[1:], -100.0, 100.0)
            orc.orc_dyn_step(B._p(p), B._p(Rc), B._p(dq), B._p(uq), 0.2, B._p(Ib), g)
            Y[ti, :3] = p; Y[ti, 3:6] = Rc[6:9]; Y[ti, 6:] = dq; U[ti] = uq
            err += p @ p; eff += uq[1] ** 2 + uq[2] ** 2
        res[name + "_y"] = Y[::10].astype(np.float32)
        res[name + "_u"] = U[::10].astype(np.float32)
        res[name + "_u4w0"] = np.array(U4, np.float32)
        res[name + "_metrics"] = np.array([err / nt, eff / nt, nsolve])
        res[name + "_cfg"] = np.array([hl, tend, task] + list(tp))
    np.savez_compressed(os.path.join(HERE, "cascade_wlqp.npz"), **res)


if __name__ == "__main__":
    B.build(ref=True)
    assert B.have_ref(), "oracle/_ref not built (needs /root/reference)"
    c = constants()
    umpc_seq()
    rollout_c1()
    wl_seq(c)
    cascade_wlqp(c)
    for f in sorted(os.listdir(HERE)):
        print(f, os.path.getsize(os.path.join(HERE, f)))
