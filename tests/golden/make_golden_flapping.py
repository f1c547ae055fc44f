"""tests/golden/make_golden_flapping.py -- golden fixture of the flapping-wing plant row.

Run in the build container (needs /root/reference):  python tests/golden/make_golden_flapping.py
The reference's OWN functions of template/robobee.py (`aerodynamics` :20-47, `wrench` :49-50, `actuatorModel` :52-55,
`WingFlapFilter` :57-73; exec'd from the file between "# Good parameters" and "class RobobeeSim" -- the module itself
imports pybullet, which is absent) on seeded inputs -> flapping_aero.npz.  The multibody step of the plant is pybullet's
and has no fixture: that part of the row is "parity unpinned" (oracle/flapping_oracle.py)."""
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/template/robobee.py"


def ref_namespace():
    src = open(REF).read()
    a, b = src.index("# Good parameters"), src.index("class RobobeeSim")
    ns = dict(np=np)
    exec(compile(src[a:b], "robobee.py", "exec"), ns)
    return ns


def main():
    ns = ref_namespace()
    rng = np.random.default_rng(20260211)
    n = 512
    theta = np.column_stack([rng.uniform(-1.2, 1.2, n), rng.uniform(-1.3, 1.3, n)])
    dtheta = np.column_stack([rng.uniform(-1.5, 1.5, n), rng.uniform(-2.0, 2.0, n)])
    dtheta[:8, 0] = [0.0, 1e-6, -1e-6, 1e-4, -1e-4, 0.0, 1e-3, -1e-3]      # around the regularised sign change of the drag direction
    params = dict(ycp=7.0, Roffs=3.0, d=4.0)   # what getInfoFromURDF reads out of urdf/sdab.xacro (:160-171): 0.5 Rwing, RwingOffset, d
    out = dict(theta=theta, dtheta=dtheta, params=np.array([params["ycp"], params["Roffs"], params["d"]]),
               consts=np.array([ns["CD0"], ns["CDmax"], ns["CLmax"], ns["RHO"], ns["rcopnondim"], ns["wparams"]["Aw"],
                                ns["wparams"]["r1h"], ns["wparams"]["r2h"], ns["wparams"]["cbar"]]))
    for sgn, tag in ((1, "L"), (-1, "R")):
        F, P, W = [], [], []
        for i in range(n):
            # the reference mutates its arguments (theta[0] *= lrSign): hand it copies, like RobobeeSim.update does for the
            # right wing (-self.q[2:4] is a new array) -- the fixture stores what the CALLER passed
            f, pc = ns["aerodynamics"](theta[i].copy(), dtheta[i].copy(), sgn, params)
            F.append(f); P.append(pc); W.append(ns["wrench"](f, pc))
        out["F_" + tag], out["pcop_" + tag], out["wrench_" + tag] = np.array(F), np.array(P), np.array(W)
    V = rng.uniform(-200, 200, 32)
    out["act_V"] = V
    out["act_F"] = np.array([ns["actuatorModel"](v, 0.3, -0.2) for v in V])
    # WingFlapFilter(freq, timestep, n): moving average over one flapping period
    wf = ns["WingFlapFilter"](0.16, 0.1, 6)
    xs = rng.standard_normal((200, 6))
    out["filt_in"] = xs
    out["filt_out"] = np.array([wf.update(x) for x in xs])
    out["filt_len"] = np.array([wf.buf.shape[0]])
    np.savez_compressed(os.path.join(HERE, "flapping_aero.npz"), **out)
    print("wrote flapping_aero.npz", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
