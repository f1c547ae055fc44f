"""Config-5 sweep on one GPU: B x N, mild initial conditions about the hover goal (each env its own), a bounded number of
closed-loop MPC steps per point; prints env-steps/s (= QP solves/s), QP iterations/s and the HBM traffic model."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import dt_util  # noqa: E402
from robobee3d_b200.dirtran import BatchLTVMPC  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--B", default="1024,4096,16384,65536,262144")
ap.add_argument("--N", default="3,5,10,15,20")
ap.add_argument("--steps", type=int, default=10)
ap.add_argument("--max-gb", type=float, default=60.0)
a = ap.parse_args()
for N in [int(x) for x in a.N.split(",")]:
    for B in [int(x) for x in a.B.split(",")]:
        rows = 44 + 74 * (N + 1)
        gb = rows * 8 * B / 1e9
        if gb > a.max_gb:
            print(json.dumps({"N": N, "B": B, "skipped": "slab %.1f GB" % gb})); continue
        b = BatchLTVMPC(B, N=N)
        b.set_state(y=dt_util.mild_states(B, seed=1), xr=dt_util.XR_HOVER)
        b.rollout(2)                      # warm-up: first solves from a cold controller
        c0 = b.counters()
        t0 = time.time()
        b.rollout(a.steps)
        ms = b.last_kernel_ms()
        c1 = b.counters()
        its = c1["iters"] - c0["iters"]
        # bytes per knot and iteration of the two sweeps: 140 doubles read + 38 written
        traffic = its * (N + 1) * 178 * 8
        print(json.dumps({"N": N, "B": B, "steps": a.steps, "ms": round(ms, 3), "env_steps_per_s": B * a.steps / ms * 1e3,
                          "qp_iters_per_s": its / ms * 1e3, "iters_per_solve": its / (B * a.steps),
                          "unsolved": c1["unsolved"] - c0["unsolved"], "model_GBps": traffic / ms / 1e6,
                          "slab_GB": round(gb, 2)}), flush=True)
        b.close()
