"""Build libumpc_b200.so in-tree with nvcc for sm_100a (no JIT cache).  The .so is git-ignored (the history stays
source-only) but not gpurun-ignored: a built library travels to the GPU box with the working-tree snapshot."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "csrc", "umpc_kernels.cu")
SRC_CONV = os.path.join(HERE, "csrc", "umpc_conv.cu")  # converged mode: the same device code in double
SRC_DT = os.path.join(HERE, "csrc", "dirtran.cu")      # planar dir-tran MPC (config 5)
SRC_UMG = os.path.join(HERE, "csrc", "umg.cu")        # any-N UprightMPC2 / fused MPC+WLQP QP builder on top of gqp
SRC_WFIT = os.path.join(HERE, "csrc", "wrenchfit.cu")  # wrench-map fitting (popts): kinematics -> wrench, batched SVD least squares
SRC_FW = os.path.join(HERE, "csrc", "flapping.cu")     # flapping-wing plant: open-loop sweeps (row f2)
SRC_GQP = os.path.join(HERE, "csrc", "gqp.cu")        # generic warp-per-QP OSQP (fused MPC+WLQP, any-N UprightMPC2, LTI quadcopter)
OUT = os.path.join(HERE, "libumpc_b200.so")
DEPS = [SRC, SRC_CONV, SRC_DT, SRC_GQP, SRC_UMG, SRC_WFIT, SRC_FW] + [os.path.join(HERE, "csrc", f) for f in ("umpc_core.cuh", "umpc_gen.cuh", "sim_core.cuh", "wl_core.cuh",
                                                                   "umpc_dev.cuh", "umpc_conv.h", "dirtran_core.cuh")] + [
    os.path.join(HERE, "..", "include", "umpc_b200.h"), os.path.join(HERE, "..", "include", "dirtran_b200.h"), os.path.join(HERE, "..", "include", "gqp_b200.h"), os.path.join(HERE, "..", "include", "umg_b200.h"), os.path.join(HERE, "..", "include", "wrenchfit_b200.h"), os.path.join(HERE, "..", "include", "flapping_b200.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-shared",
              "-Xcompiler", "-fPIC", "-Xptxas", "-v", "--threads", "6"]


def stale():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    return any(os.path.getmtime(d) > t for d in DEPS)


def build(force=False, verbose=False, defines=(), out=None):
    """defines/out: tuning variants (e.g. defines=("UMPC_LREG_GROUPS=8",), out="/path/libumpc_b200_v.so")"""
    if out is None and not force and not stale():
        return OUT
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    final = out or OUT
    tmp = final + ".partial"     # linked beside the target and renamed: a reader (or a snapshot) never sees half a library
    cmd = [nvcc] + NVCC_FLAGS + ["-D" + d for d in defines] + ["-o", tmp, SRC, SRC_CONV, SRC_DT, SRC_GQP, SRC_UMG, SRC_WFIT, SRC_FW]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode == 0:
        os.replace(tmp, final)
    log = r.stdout + r.stderr
    if out is None:
        with open(os.path.join(HERE, "csrc", "ptxas_v.log"), "w") as f:
            f.write(log)
    if verbose or r.returncode:
        sys.stderr.write(log)
    if r.returncode:
        raise RuntimeError("nvcc failed for libumpc_b200.so")
    return out or OUT


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
