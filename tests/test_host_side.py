"""CPU tests of the host side (no GPU, no compute calls into the CUDA library):
  * the C-ABI library loads and exports every symbol include/umpc_b200.h declares;
  * the product fails loudly (no CPU fallback) when no CUDA device is present;
  * the generated device header is what the generator emits today;
  * the device algorithm headers (robobee3d_b200/csrc/*.cuh are __host__ __device__), compiled for the host by
    tests/hostcheck, reproduce the reference golden vectors -- a development aid that validates the kernel's
    FORMULATION (reduced 45x45 system, un-scaled coordinates) before GPU time is spent.  The real parity
    tests are the -m gpu ones, through the C-ABI on the device.
"""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

import util
from util import GOLDEN, ROOT, TOL_X_REL, TOL_Y_REL, TOL_QPDATA

HEADER = os.path.join(ROOT, "include", "umpc_b200.h")


def _declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", " ", src, flags=re.S)
    names = re.findall(r"^\s*(?:const\s+)?(?:void|int|float|double|char)\s*\*?\s*(\w+)\s*\(", src, flags=re.M)
    return sorted(set(names))


def test_header_symbols_are_exported():
    from robobee3d_b200 import _lib
    decl = _declared_symbols()
    assert len(decl) >= 35, decl
    L = _lib.load()
    missing = [s for s in decl if not hasattr(L, s)]
    assert not missing, "libumpc_b200.so lacks %s" % missing
    assert sorted(set(_lib.EXPORTS)) == decl, "robobee3d_b200/_lib.py EXPORTS out of sync with include/umpc_b200.h"
    # struct layouts of the drop-in API (uprightmpc2.h:27-43, funapprox.h:18-43): sizes as the reference compiles them
    assert C.sizeof(_lib.UprightMPC_t) == 4 * (3 + 27 + 6 + 18 + 39 + 39 + 45 + 45 + 48 + 48 + 2 + 6 + 1)
    assert C.sizeof(_lib.FunApprox_t) == 4 * (1 + 1 + 4 + 16) and C.sizeof(_lib.WLCon_t) == 4 * (16 + 36) + 6 * 88


def test_only_cuda_runtime_dependencies():
    """the product library links the CUDA runtime and libc/libstdc++ only -- not torch, not the oracle"""
    from robobee3d_b200 import _lib
    out = subprocess.run(["ldd", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in out and "torch" not in out
    nm = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "orc_" not in nm and "osqp_" not in nm


@pytest.mark.skipif(util.has_cuda(), reason="checks the no-GPU behaviour")
def test_no_cpu_fallback_without_device():
    from robobee3d_b200 import _lib, uprightmpc2py
    from robobee3d_b200.batch import BatchUMPC
    with pytest.raises(RuntimeError, match="no CUDA device|CUDA"):
        BatchUMPC(4)
    d = util.load_constants()["umpc_defaults"]
    with pytest.raises(RuntimeError):
        uprightmpc2py.UprightMPC2C(d["dt"], d["g"], d["TtoWmax"], d["ws"], d["wds"], d["wpr"], d["wpf"], d["wvr"], d["wvf"],
                                   d["wthrust"], d["wmom"], d["Ib"], d["maxIter"])
    from robobee3d_b200 import template_controllers
    with pytest.raises(RuntimeError):
        template_controllers.createMPC()
    # raw C-ABI: error code + message, outputs poisoned with NaN
    L = _lib.load()
    up = _lib.UprightMPC_t()
    uq, acc = (C.c_float * 3)(), (C.c_float * 6)()
    z3, R = (C.c_float * 3)(), (C.c_float * 9)(1, 0, 0, 0, 1, 0, 0, 0, 1)
    z6 = (C.c_float * 6)()
    rc = L.umpcUpdate(C.byref(up), uq, acc, z3, R, z6, z3, z3, z3, C.c_float(-1.0))
    assert rc == 100 and np.isnan(uq[0]) and np.isnan(acc[0])


def test_product_package_never_imports_oracle():
    pkg = os.path.join(ROOT, "robobee3d_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, flags=re.M), f
                assert "liboracle" not in src and "libref_" not in src and "binding" not in src, f


def test_generated_header_is_current():
    out = subprocess.run([sys.executable, "-m", "robobee3d_b200.codegen.gen_umpc"], capture_output=True, text=True, cwd=ROOT,
                         check=True).stdout
    cur = open(os.path.join(ROOT, "robobee3d_b200", "csrc", "umpc_gen.cuh")).read()
    assert out == cur, "robobee3d_b200/csrc/umpc_gen.cuh is stale: re-run python -m robobee3d_b200.codegen.gen_umpc"


def test_symbolic_structure():
    """A has the reference's 111 non-zeros of which 48 change per update (uprightmpc2.h:25, workspace.c:154-427);
    the reduced K's factor has the nnz the kernel's shared-memory layout assumes."""
    from robobee3d_b200.codegen.gen_umpc import build
    nx, nc, ent, rows, kent, perm, cols = build(3)
    assert (nx, nc, len(ent)) == (45, 39, 111)
    assert sum(1 for (_, _, k) in ent if k[0] != "c") == 48
    assert sorted(perm) == list(range(45))
    hdr = open(os.path.join(ROOT, "robobee3d_b200", "csrc", "umpc_gen.cuh")).read()
    nnzl = int(re.search(r"#define UMPC_GEN_NNZL (\d+)", hdr).group(1))
    assert nnzl == sum(len(c) for c in cols)


def test_device_formulation_on_host_matches_reference_golden():
    sys.path.insert(0, os.path.join(ROOT, "tests", "hostcheck"))
    import hc
    gold = np.load(os.path.join(GOLDEN, "umpc_seq.npz"))
    d = util.load_constants()["umpc_defaults"]
    B, ncall = gold["p"].shape[:2]
    worst = 0.0
    for e in range(B):
        h = hc.HostUMPC(d)
        for c in range(ncall):
            uq, acc = h.update(gold["p"][e, c], gold["R"][e, c], gold["dq"][e, c], gold["pdes"][e, c], gold["dpdes"][e, c],
                               gold["sdes"][e, c], float(gold["aT0"][e, c]))
            o = h.get()
            ex, ey = util.relerr(o["x"], gold["x"][e, c]), util.relerr(o["y"], gold["y"][e, c])
            worst = max(worst, ex)
            assert ex <= TOL_X_REL and ey <= TOL_Y_REL, (e, c, ex, ey)
            assert util.relerr(o["D"], gold["D"][e, c]) <= (1e-4 if c == 0 else 2e-3)
            assert util.relerr(o["E"], gold["E"][e, c]) <= (1e-4 if c == 0 else 2e-3)
            # warm start kept in OSQP's scaled coordinates like the reference's workspace.x/z/y
            assert util.relerr(o["xs"], gold["xs"][e, c]) <= 2 * TOL_X_REL
            assert int(o["info"][0]) == 50
            assert np.abs(uq - gold["uquad"][e, c]).max() <= TOL_X_REL * max(1.0, np.abs(gold["x"][e, c]).max())
    print("hostcheck worst rel err vs reference C: %.2e" % worst)


@pytest.mark.parametrize("lib", ["libhostcheck64.so", "libhostcheck64_sf.so"])
def test_converged_formulation_on_host_matches_oracle_f64(lib):
    """converged mode (termination test every 25 iterations, rho re-estimated every 100, osqp.c:431-500) of the device
    code, compiled for the host in double, against the oracle's float64 OSQP restatement and the golden solutions of
    the reference's vendored OSQP in float64: same iteration counts, solutions within 10 eps_abs + eps_rel |x*|.
    `_sf` = the build the GPU runs: Ruiz equilibration in float (-DUMPC_SCALE_FLOAT), everything else in double."""
    sys.path.insert(0, os.path.join(ROOT, "tests", "hostcheck"))
    import hc
    from oracle import binding as OB
    gold = np.load(os.path.join(GOLDEN, "umpc_seq.npz"))
    conv = np.load(os.path.join(GOLDEN, "umpc_conv.npz"))
    d = util.load_constants()["umpc_defaults"]
    B, ncall = gold["p"].shape[:2]
    for e in range(0, B, 3):
        h = hc.HostUMPC(d, f64=True, lib=lib)
        h.set_converged()
        o = OB.OracleUMPC("f64")
        o.set_converged()
        for c in range(ncall):
            a = (gold["p"][e, c], gold["R"][e, c], gold["dq"][e, c], gold["pdes"][e, c], gold["dpdes"][e, c], gold["sdes"][e, c],
                 float(gold["aT0"][e, c]))
            uq, acc = h.update(*a)
            uo, ao = o.update(*a)
            g = h.get()
            x, y, info = o.solution()
            assert int(g["info"][0]) == info["iter"] and int(g["info"][1]) == info["status"] == 1, (e, c)
            assert np.abs(g["x"] - x).max() <= 10 * 1e-4 + 1e-4 * np.abs(x).max()
            assert np.abs(g["y"] - y).max() <= 10 * 1e-4 + 1e-4 * np.abs(y).max()
            tol = 10 * 1e-4 + 1e-4 * np.abs(x).max()
            assert np.abs(uq - uo).max() <= tol and np.abs(acc - ao).max() <= tol / d["dt"]
            if c == 0:
                xr = conv["x"][e]
                assert int(g["info"][0]) == int(conv["iter"][e])
                assert np.abs(g["x"] - xr).max() <= 10 * 1e-4 + 1e-4 * np.abs(xr).max()


def test_dirtran_bench_shards_reproduce_the_same_envs():
    """bench.py --workload dirtran: a rank's initial conditions are keyed by the global env id, so N ranks of B envs hold
    exactly the envs one rank of N*B would"""
    import bench
    whole, xr = bench._dt_mild_states(0, 96)
    for lo, hi in ((0, 32), (32, 64), (64, 96)):
        part, xr2 = bench._dt_mild_states(lo, hi)
        assert np.array_equal(part, whole[lo:hi]) and np.array_equal(xr, xr2)
    assert np.abs(whole - xr).max() <= 0.2 and np.abs(whole - xr).max() > 0.05


def _cc(args, out):
    import subprocess
    subprocess.run(["gcc", "-O1"] + args + ["-o", out], check=True)
    return out


def test_abi_layout_equals_reference_headers(tmp_path):
    """Compiler-checked boundary: tests/cabi/abi_probe.c compiled against the REFERENCE's own headers
    (template/uprightmpc2/{uprightmpc2,funapprox}.h, template/wlcontroller/c/{wlqp,funapprox}.h) and against
    include/umpc_b200.h must print the same sizeof/offsetof tables.  Where /root/reference is absent (GPU box) the
    reference-side tables are the committed tests/golden/abi_layout_ref_*.txt, generated by this same probe."""
    import subprocess
    src = os.path.join(ROOT, "tests", "cabi", "abi_probe.c")
    ours = _cc(["-I" + os.path.join(ROOT, "include"), src], str(tmp_path / "ours"))
    run = lambda exe, w: subprocess.run([exe, w], capture_output=True, text=True, check=True).stdout
    refs = {}
    for tag, define, inc in (("umpc", "PROBE_REF_UMPC", "/root/reference/template/uprightmpc2"),
                             ("wlqp", "PROBE_REF_WLQP", "/root/reference/template/wlcontroller/c")):
        gold = open(os.path.join(ROOT, "tests", "golden", "abi_layout_ref_%s.txt" % tag)).read()
        if os.path.isdir(inc):
            exe = _cc(["-D" + define, "-I" + inc, src], str(tmp_path / ("ref_" + tag)))
            live = run(exe, "a")
            assert live == gold, "committed reference layout table is stale"
        refs[tag] = gold
    assert run(ours, "u") + run(ours, "f") == refs["umpc"]
    assert run(ours, "w") + run(ours, "f") == refs["wlqp"]
    assert "UprightMPC_t sizeof 1308" in refs["umpc"]


def test_c_program_links_against_the_library(tmp_path):
    """a plain C translation unit including include/umpc_b200.h links libumpc_b200.so (no ctypes in between)"""
    import subprocess
    libdir = os.path.join(ROOT, "robobee3d_b200")
    exe = _cc(["-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "cabi", "link_check.c"), "-L" + libdir,
               "-lumpc_b200", "-Wl,-rpath," + libdir], str(tmp_path / "link_check"))
    out = subprocess.run(["ldd", exe], capture_output=True, text=True).stdout
    assert "libumpc_b200.so" in out


def test_gqp_and_umg_headers_are_exported_and_refuse_without_device():
    """include/gqp_b200.h and include/umg_b200.h: every declared symbol is exported and listed by the Python wrapper; the
    pattern / size helpers (host arithmetic of the QP's structure, no compute) agree with the reference's shapes"""
    from robobee3d_b200 import _lib, gqp
    L = gqp._bind()
    for hdr, lst in (("gqp_b200.h", gqp.GQP_EXPORTS), ("umg_b200.h", gqp.UMG_EXPORTS)):
        src = re.sub(r"/\*.*?\*/", " ", open(os.path.join(ROOT, "include", hdr)).read(), flags=re.S)
        decl = sorted(set(re.findall(r"^\s*(?:const\s+)?(?:void|int|float|double|char)\s*\*?\s*(\w+)\s*\(", src, flags=re.M)))
        assert decl == sorted(lst), (hdr, decl, sorted(lst))
        assert all(hasattr(L, s) for s in decl)
    # shipped instance N = 3: nx 45, nc 39, nnz 111 (uprightmpc2.h:20-25, workspace.c); fused: 49 / 43 (uprightmpc2.h.withWLQP:20-51)
    assert gqp.umg_sizes(3) == (45, 39, 111) and gqp.umg_sizes(3, True) == (49, 43, 115)
    rp, ci = gqp.umg_pattern(3)
    gold = np.load(os.path.join(GOLDEN, "umpc_genN.npz"))
    rp5, ci5 = gqp.umg_pattern(int(gold["N"]))
    A = gold["A"][0]
    rows = np.repeat(np.arange(A.shape[0]), np.diff(rp5))
    off = A.copy(); off[rows, ci5] = 0
    assert not off.any() and (A[rows, ci5] != 0).sum() >= len(ci5) - 3     # pattern of template_controllers.initConstraint at N = 5
    if not util.has_cuda():
        with pytest.raises(RuntimeError, match="no CUDA device|CUDA"):
            gqp.BatchQP(2, 45, 39, rp, ci)
        with pytest.raises(RuntimeError, match="no CUDA device|CUDA"):
            gqp.BatchUprightMPC2G(2, N=5)


def test_bf16_sigma_pairs_ab_against_full_precision():
    """A/B of the two arithmetic shortcuts of the float32 path on the host build of the device code: the sigma / D^2 pairs
    packed as bf16 halves (umpc_pack2) against the same code with the pairs at full precision (-DUMPC_SG_FULL), over the
    48 x 3 golden QPs.  The term is ~1e-6 of K's diagonal: the packing must not move x beyond float32 noise."""
    sys.path.insert(0, os.path.join(ROOT, "tests", "hostcheck"))
    import hc
    gold = np.load(os.path.join(GOLDEN, "umpc_seq.npz"))
    d = util.load_constants()["umpc_defaults"]
    B, ncall = gold["p"].shape[:2]
    worst = werr_a = werr_b = 0.0
    for e in range(B):
        ha, hb = hc.HostUMPC(d), hc.HostUMPC(d, lib="libhostcheck_sgfull.so")
        for c in range(ncall):
            a = (gold["p"][e, c], gold["R"][e, c], gold["dq"][e, c], gold["pdes"][e, c], gold["dpdes"][e, c], gold["sdes"][e, c], float(gold["aT0"][e, c]))
            ha.update(*a); hb.update(*a)
            xa, xb = ha.get()["x"], hb.get()["x"]
            worst = max(worst, util.relerr(xa, xb))
            werr_a = max(werr_a, util.relerr(xa, gold["x"][e, c])); werr_b = max(werr_b, util.relerr(xb, gold["x"][e, c]))
    print("bf16 pairs vs full precision: max rel diff of x %.2e; error vs reference C: packed %.2e, full %.2e" % (worst, werr_a, werr_b))
    assert worst <= 5e-5           # the packing's own effect (measured 2.2e-5): an order below the float32 noise floor of the 50-iteration iterate
    assert werr_a <= max(1.15 * werr_b, werr_b + 2e-4)   # and it does not make parity with the reference worse


def test_wrenchfit_header_is_exported_and_refuses_without_device():
    """include/wrenchfit_b200.h: every declared symbol is exported and bound; no CPU fallback behind it"""
    from robobee3d_b200 import wrenchfit
    L = wrenchfit._bind()
    src = re.sub(r"/\*.*?\*/", " ", open(os.path.join(ROOT, "include", "wrenchfit_b200.h")).read(), flags=re.S)
    decl = sorted(set(re.findall(r"^\s*(?:const\s+)?(?:void|int|float|double|char)\s*\*?\s*(\w+)\s*\(", src, flags=re.M)))
    assert decl == sorted(wrenchfit.EXPORTS), decl
    assert all(hasattr(L, s) for s in decl)
    g = np.load(os.path.join(GOLDEN, "wrenchfit.npz"))
    assert g["d0_popts"].shape == (6, 15) and g["d0_xdata"].shape[1] == 4 and g["d0_kins"].shape[1] == 9
    if not util.has_cuda():
        with pytest.raises(RuntimeError, match="no CUDA device|CUDA"):
            wrenchfit.fit_wrench_map(g["d0_xdata"], g["d0_ws"])
        with pytest.raises(RuntimeError, match="no CUDA device|CUDA"):
            wrenchfit.wrench_from_kinematics(g["d0_kins"], g["d0_fs"])
