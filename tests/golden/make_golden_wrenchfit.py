"""tests/golden/make_golden_wrenchfit.py -- golden fixture of the wrench-map fitting row.

Run in the build container (needs /root/reference):  python tests/golden/make_golden_wrenchfit.py
The reference's OWN functions of template/sdab_num_wrenchmap.py (loadEmpiricalData, unpackDat, FunApprox,
wrenchFromKinematics; exec'd from the file with numpy standing in for autograd.numpy and the aero constants of
template/robobee.py:7-17) on the reference's own bench data (template/data/*.csv), then the file's own fitting statement
(:263, scipy curve_fit from p0 = ones) -> wrenchfit.npz: xdata (M,4), kins (M,9), fs (M,), ws (M,6), popts (6,15)."""
import os
import sys
import types

import numpy as np
from scipy.optimize import curve_fit

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/template"


def ref_namespace():
    rb = open(os.path.join(REF, "robobee.py")).read()
    a, b = rb.index("# Good parameters"), rb.index("def aerodynamics")
    rns = dict(np=np)
    exec(compile(rb[a:b], "robobee.py", "exec"), rns)
    robobee = types.SimpleNamespace(**{k: v for k, v in rns.items() if not k.startswith("__")})
    src = open(os.path.join(REF, "sdab_num_wrenchmap.py")).read()
    a, b = src.index("# Load empirical data"), src.index('if __name__ == "__main__":')
    import itertools
    ns = dict(np=np, itertools=itertools, robobee=robobee, curve_fit=curve_fit)
    body = src[a:b]
    # the plotting helper needs matplotlib: cut it (it is not on the path)
    c0, c1 = body.index("def splineContour"), body.index("class FunApprox")
    body = body[:c0] + body[c1:]
    c0, c1 = body.index("def wrenchCompare"), body.index("fa = FunApprox(4)")
    body = body[:c0] + body[c1:]
    exec(compile(body, "sdab_num_wrenchmap.py", "exec"), ns)
    return ns, robobee


def main():
    ns, robobee = ref_namespace()
    out = {}
    files = sorted(f for f in os.listdir(os.path.join(REF, "data")) if f.endswith(".csv"))
    for fi, f in enumerate(files):
        try:
            dat = ns["loadEmpiricalData"](os.path.join(REF, "data", f))
        except Exception as e:
            print("skip", f, type(e).__name__, e)
            continue
        Vmeans, uoffss, fs, udiffs, h2s, ws0, kins = ns["unpackDat"](dat)
        params = dict(robobee.wparams)
        params.update({'ycp': 7.5, 'AR': 4.5, 'R': 3})                       # sdab_num_wrenchmap.py:252-253
        ws = ns["wrenchFromKinematics"](kins, fs, params, kaerox=1, strokex=1.5)   # :255
        xdata = np.vstack((Vmeans, uoffss, udiffs, h2s)).T                   # :262
        fa = ns["fa"]
        popts = np.vstack([curve_fit(fa.f, xdata, ws[:, i], p0=np.ones(fa.nparams()))[0] for i in range(6)])   # :266
        fit = np.vstack([fa.f(xdata, *popts[i]) for i in range(6)]).T
        tag = "d%d" % fi
        out.update({tag + "_xdata": xdata, tag + "_kins": kins, tag + "_fs": fs, tag + "_ws": ws, tag + "_popts": popts, tag + "_fit": fit})
        Phi = np.array([[1, *x, *[(0.5 if i == j else 1.0) * x[i] * x[j] for i in range(4) for j in range(i, 4)]] for x in xdata])
        print(f, "rows", len(xdata), "rank", np.linalg.matrix_rank(Phi), "resid", np.abs(fit - ws).max(), "|ws|", np.abs(ws).max())
        out[tag + "_name"] = np.array(f)
    out["params"] = np.array([robobee.wparams['Aw'], 4.5, robobee.wparams['r2h'], 7.5, 3.0, robobee.RHO, robobee.CD0, robobee.CDmax, robobee.CLmax, 1.0, 1.5])
    np.savez_compressed(os.path.join(HERE, "wrenchfit.npz"), **out)


if __name__ == "__main__":
    main()
