import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    # oracle libs: built here when missing (gcc only); oracle/_ref only where /root/reference exists
    from oracle import binding
    need = [os.path.join(ROOT, "oracle", "liboracle_f32.so"), os.path.join(ROOT, "oracle", "liboracle_f64.so")]
    if not all(os.path.exists(p) for p in need) or (os.path.isdir("/root/reference") and not binding.have_ref()):
        binding.build(ref=os.path.isdir("/root/reference"))
    hc = os.path.join(ROOT, "tests", "hostcheck")
    srcs = [os.path.join(hc, "hostcheck.cpp")] + [os.path.join(ROOT, "robobee3d_b200", "csrc", f)
                                                   for f in ("umpc_core.cuh", "umpc_gen.cuh", "sim_core.cuh", "wl_core.cuh")]
    for name, defs in (("libhostcheck.so", []), ("libhostcheck64.so", ["-DUMPC_REAL=double"]), ("libhostcheck_sgfull.so", ["-DUMPC_SG_FULL"]),
                       ("libhostcheck64_sf.so", ["-DUMPC_REAL=double", "-DUMPC_SCALE_FLOAT"])):
        so = os.path.join(hc, name)
        if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
            subprocess.run(["g++", "-O2", "-fPIC", "-shared", "-ffp-contract=off"] + defs + ["-o", so, srcs[0]], check=True)


def has_cuda():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.fixture(scope="session")
def cuda_required():
    if not has_cuda():
        pytest.skip("no CUDA device")
