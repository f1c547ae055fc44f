"""wfitFitNoAlloc (include/wrenchfit_b200.h): the allocation-free variant of the wrench-map fit, checked against the
existing entry point wfitFit through the C-ABI.  Collected last on purpose.

The entry point has no counterpart in the reference; it exists for callers that cannot call cudaMalloc.  Device buffers
for its inputs and outputs are the CALLER's: here torch tensors, all created before the calls under test."""
import ctypes as C
import os

import numpy as np
import pytest

from util import GOLDEN, ROOT

TOL = 1e-5      # the tolerance of test_wrench_map_fit_matches_reference_curve_fit, relative to the parameters' scale


def _datasets():
    g = np.load(os.path.join(GOLDEN, "wrenchfit.npz"))
    tags = sorted({k.split("_")[0] for k in g.files if k.startswith("d")})
    return g, tags


@pytest.mark.gpu
def test_no_alloc_fit_matches_wfitFit_on_two_calls_in_a_row(cuda_required):
    import torch
    from robobee3d_b200 import wrenchfit
    g, tags = _datasets()
    L = wrenchfit._bind()
    # call 1: every dataset of the fixture in one launch (F = 4, M = 120); call 2: ONE dataset with fewer trials and
    # perturbed wrenches, so the shared scratch still holds call 1's contents beyond (and under) what call 2 rewrites
    X1 = np.array([g[t + "_xdata"] for t in tags]); W1 = np.array([g[t + "_ws"] for t in tags])
    rng = np.random.default_rng(7)
    X2 = g[tags[1] + "_xdata"][None, :97].copy()
    W2 = g[tags[1] + "_ws"][None, :97] * (1.0 + 0.05 * rng.standard_normal((1, 97, 6)))
    cases = [(X1, W1), (X2, W2)]
    want = [wrenchfit.fit_wrench_map(X, W, return_info=True) for X, W in cases]      # the existing entry point (wfitFit)
    dev = [(torch.from_numpy(X).cuda(), torch.from_numpy(W).cuda(), torch.full((len(X), 90), float("nan"), dtype=torch.float64, device="cuda"),
            torch.full((len(X),), -1, dtype=torch.int32, device="cuda")) for X, W in cases]
    stream = torch.cuda.current_stream().cuda_stream
    torch.cuda.synchronize()
    mem0 = torch.cuda.memory_allocated()
    for dx, dw, dp, dr in dev:       # two calls in a row on one stream, no synchronisation between them
        rc = L.wfitFitNoAlloc(dx.shape[0], dx.shape[1], dx.data_ptr(), dw.data_ptr(), dp.data_ptr(), dr.data_ptr(), stream)
        assert rc == 0, L.umpcLastError().decode()
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == mem0
    for (dx, dw, dp, dr), (popts, info) in zip(dev, want):
        got = dp.cpu().numpy().reshape(popts.shape)
        assert np.isfinite(got).all()
        assert np.abs(got - popts).max() <= TOL * np.abs(popts).max()
        assert (dr.cpu().numpy() == np.atleast_1d(info["rank"])).all()
    # and against the reference's own curve_fit answer (golden), like the existing entry point's test
    got1 = dev[0][2].cpu().numpy().reshape(len(tags), 6, 15)
    for i, t in enumerate(tags):
        assert np.abs(got1[i] - g[t + "_popts"]).max() <= TOL * np.abs(g[t + "_popts"]).max()


@pytest.mark.gpu
def test_no_alloc_fit_rejects_sizes_beyond_the_static_scratch(cuda_required):
    import torch
    from robobee3d_b200 import wrenchfit
    L = wrenchfit._bind()
    d = torch.zeros(16, dtype=torch.float64, device="cuda")      # never touched: the checks come before any launch
    for F, M in ((wrenchfit.NOALLOC_MAX_DATASETS + 1, 4), (1, wrenchfit.NOALLOC_MAX_TRIALS + 1), (64, 257)):
        with pytest.raises(RuntimeError, match="exceeds the static scratch"):
            wrenchfit.fit_wrench_map_noalloc(F, M, d.data_ptr(), d.data_ptr(), d.data_ptr())
    with pytest.raises(RuntimeError, match="bad argument"):
        wrenchfit.fit_wrench_map_noalloc(1, 16, None, d.data_ptr(), d.data_ptr())
    torch.cuda.synchronize()


def test_no_alloc_entry_point_source_calls_no_allocator():
    """static check of the entry point and its set-up helper: no allocating runtime call between their braces"""
    src = open(os.path.join(ROOT, "robobee3d_b200", "csrc", "wrenchfit.cu")).read()
    body = src[src.index("static int noalloc_scratch("):]
    assert 'extern "C" int wfitFitNoAlloc(' in body and "cudaGetSymbolAddress((void**)&ptr, wfit::g_noalloc_scratch)" in body
    for banned in ("cudaMalloc", "cudaFree", "cudaHostAlloc", "cudaMemcpy", "cudaEventCreate", "cudaStreamCreate", "new ", "malloc("):
        assert banned not in body, banned
    assert "static __device__ __align__(256) double g_noalloc_scratch[" in src
