#!/usr/bin/env python
"""bench.py -- closed-loop UprightMPC2 Monte Carlo on B200 (BASELINE.json configs[2]).

Workload (per GPU): 65,536 independent flights (config-3 perturbed ICs, hover task), MPC every plant step
(hlInterval=None, controlTest's default), so one env-step = one umpcUpdate-equivalent QP solve (Ruiz x10,
factor, 50 fixed-rho ADMM iterations, float32 -- the reference C path) + torque clip + rigid-body step.
A "step" = one lockstep env-step of all envs of the rank = one launch of umpc_rollout_kernel.

  value : env-steps/s (= QP solves/s), state resident in HBM, CUDA events on the launching stream.
  e2e   : the same env-step through the C-ABI with HOST (pinned) buffers, umpcBatchEnvStep: H2D of
          (p,R,dq,pdes,dpdes,sdes), kernel, D2H of (uquad,accdes,p,R,dq) inside the timed region.
  --impl reference : the reference's own C controller (oracle/_ref, compiled from /root/reference sources)
          driven by the same loop on all host cores, one process per core.

Launch: python bench.py [--gpus N --steps K --warmup W]; for N>1 via torchrun (one rank per GPU).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

NENV_PER_GPU = 65536
# plant steps per launch of the fused rollout kernel = per bench "step": 125 x 0.2 ms = 25 ms of flight, so the default
# K = 20 steps are exactly ONE config-3 flight (tend = 500 ms, 2,500 plant steps, uprightmpc2.py:87) of every env
STEPS_PER_LAUNCH = 125
METRIC = "closed-loop UprightMPC2 env-steps/s (MPC every step: = QP solves/s)"
UNIT = "env-steps/s"


def workload_config(n_gpus, nenv, spl=STEPS_PER_LAUNCH):
    return {"workload": "BASELINE configs[2]: %d-env closed-loop Monte Carlo per GPU (quadrotorNLDyn plant + UprightMPC2 "
                        "N=3, faithful C path: 10 Ruiz passes + refactor + 50 fixed-rho float32 ADMM iterations per "
                        "env-step), perturbed controlTest ICs, hover task, hlInterval=None, dtsim=0.2 ms; one bench step = "
                        "%d plant steps (each with its MPC solve) of every env in ONE launch of the fused rollout kernel "
                        "(20 steps = one 500 ms flight)" % (nenv, spl),
            "envs_per_gpu": nenv, "envs_total": nenv * n_gpus, "mpc_every_step": True, "max_iter": 50,
            "plant_steps_per_step": spl,
            "sharding": "contiguous env ranges per rank, no data-path collective; one all-reduce of 8 doubles at the end",
            "l2_policy": "per-step working set %.0f MB/GPU (SoA state of all envs) streams through L2 each step; "
                         "inputs are not re-used between steps except by the same env" % (nenv * 2172 / 1e6)}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, gpu_index):
        self.rows = []
        self.stop = False
        self.idx = gpu_index
        self.t = None

    def _run(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        while not self.stop:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def start(self):
        self.t = threading.Thread(target=self._run, daemon=True)
        self.t.start()

    def finish(self):
        self.stop = True
        if self.t:
            self.t.join(timeout=6)
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(float(r[0]) for r in self.rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for j, n in enumerate(names) if any(r[3 + j].lower().startswith("active") for r in self.rows)]
        out = {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons,
               "samples": len(self.rows), "power_w_max": max(float(r[2]) for r in self.rows)}
        rec = os.environ.get("BENCH_CLOCK_RECORD")   # full record of the sampler (one row per 0.1 s) for sustained runs
        if rec:
            with open(rec, "w") as f:
                f.write("sm_mhz,sm_max_mhz,power_w,hw_slowdown,hw_thermal_slowdown,sw_thermal_slowdown,sw_power_cap\n")
                for r in self.rows:
                    f.write(",".join(r) + "\n")
            out["sm_mhz_min"], out["sm_mhz_max_seen"] = sm[0], sm[-1]
        return out


def make_ics(nenv, seed):
    import util
    p, R, dq = util.config3_ics(nenv, seed=seed)
    return p, R, dq


def fast_ics(nenv, seed):
    """config-3 ICs vectorised (65k rotations through the python loop of tests/util.py would take a minute)."""
    from robobee3d_b200.sweeps import control_test_ic
    rng = np.random.default_rng(seed)
    p0, R0, dq0 = control_test_ic()
    p = p0[None] + rng.normal(0, 5.0, (nenv, 3))
    dq = np.tile(dq0, (nenv, 1))
    dq[:, :3] += rng.normal(0, 0.05, (nenv, 3))
    w = rng.normal(0, 0.1, (nenv, 3))
    th = np.linalg.norm(w, axis=1)
    K = np.zeros((nenv, 3, 3))
    K[:, 0, 1], K[:, 0, 2], K[:, 1, 0], K[:, 1, 2], K[:, 2, 0], K[:, 2, 1] = -w[:, 2], w[:, 1], w[:, 2], -w[:, 0], -w[:, 1], w[:, 0]
    a = (np.sin(th) / th)[:, None, None]
    b = ((1 - np.cos(th)) / th ** 2)[:, None, None]
    Ex = np.eye(3)[None] + a * K + b * (K @ K)
    R = R0[None] @ Ex
    return p, R, dq


def run_reference(args):
    """CPU arm: the reference's C controller on all host cores (rank 0 only)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import binding as OB
    if not os.path.exists(os.path.join(ROOT, "oracle", "liboracle_f32.so")):
        OB.build(ref=os.path.isdir("/root/reference"))
    cores = os.cpu_count() or 1
    spl = args.steps_per_launch or STEPS_PER_LAUNCH
    # bounded sample of the workload per step: 64 envs per core, each advanced the step's `spl` plant steps (~0.8 s of work
    # per core and step at ~100 us per umpcUpdate)
    nenv = min(NENV_PER_GPU, cores * 64)
    ics = _cpu_ics(nenv)
    # warm-up + timed: each "step" advances the sampled envs `spl` closed-loop plant steps on all cores
    if args.warmup > 0:
        OB.cpu_env_steps(ics[: cores * 8], max(1, args.warmup), cores)
    done, tmax, wall, kind, per_core = OB.cpu_env_steps(ics, args.steps * spl, cores, per_worker=True)
    value = done / tmax
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * tmax / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args.gpus, NENV_PER_GPU, spl),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
                             "per_core": {"min": min(per_core), "median": sorted(per_core)[len(per_core) // 2], "max": max(per_core)},
                             "seconds_per_worker": tmax,
                             "sample": "%d of %d envs x %d plant steps (%d steps of %d), one process per core (%s)" % (
                                 nenv, NENV_PER_GPU, args.steps * spl, args.steps, spl,
                                 "reference C sources compiled -O3, oracle/_ref" if kind == "reference" else "oracle port")},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))
    return 0


def _cpu_ics(nenv):
    p, R, dq = fast_ics(nenv, seed=0)
    return np.concatenate([p, R.transpose(0, 2, 1).reshape(-1, 9), dq], axis=1)


def _init_nccl(dist, torch, local):
    """process group + communicator set-up with file descriptor 1 pointed at stderr: the box exports NCCL_DEBUG=VERSION
    and NCCL writes its version banner to stdout, which belongs to the one JSON line"""
    sys.stdout.flush()
    saved = os.dup(1)
    os.dup2(2, 1)
    try:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist.barrier()
        torch.cuda.synchronize()
    finally:
        sys.stdout.flush()
        os.dup2(saved, 1)
        os.close(saved)


def _bind_to_gpu_numa_node(local):
    """pin this rank to the CPU cores next to its GPU (NVML's affinity mask) BEFORE it allocates pinned host buffers, so
    that first touch puts them on the GPU's own NUMA node: the zero-copy e2e leg of 8 ranks then does not cross sockets"""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        pynvml.nvmlDeviceSetCpuAffinity(h)
        return sorted(os.sched_getaffinity(0))
    except Exception:
        return None


def _dt_mild_states(lo, hi, seed=1):
    """config-5 initial conditions about the hover goal, keyed by the global env id (shard-invariant)"""
    xr = np.array([-0.03, 0.02, 0, 0, 0, 0.0])
    y = np.tile(xr, (hi - lo, 1))
    r = np.random.default_rng(seed)
    blk = r.uniform(-1.0, 1.0, (hi, 6))[lo:hi]
    y += blk * np.array([0.01, 0.01, 0.05, 0.05, 0.05, 0.2])
    return y, xr


def _dt_cpu_worker(job):
    ys, N, nsteps = job
    from oracle import binding as OB
    t0 = time.perf_counter()
    its = 0.0
    for y in ys:
        o = OB.OracleDirTran(N=N)
        X, U, it, bad = o.rollout(y, np.array([-0.03, 0.02, 0, 0, 0, 0.0]), nsteps)
        its += float(it.sum())
    return time.perf_counter() - t0, its


def _np_dyn_step(p, R, dq, u, dt, Ib=np.array([3333.0, 3333.0, 1000.0]), g=9.81e-3):
    """quadrotorNLDyn (genqp.py:24-41) in numpy for the config-1 loop below: explicit Euler, R <- R expm(skew(w) dt)"""
    dv = u[0] * R[:, 2] - np.array([0, 0, g])
    w = dq[3:]
    dw = (-np.cross(w, Ib * w) + np.array([u[1], u[2], 0.0])) / Ib
    k = w * dt
    th = np.linalg.norm(k)
    K = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    a, b = (1 - th * th / 6, 0.5 - th * th / 24) if th < 1e-4 else (np.sin(th) / th, (1 - np.cos(th)) / th ** 2)
    return p + dt * dq[:3], R @ (np.eye(3) + a * K + b * (K @ K)), np.r_[dq[:3] + dt * dv, w + dt * dw]


def run_hostlink(args):
    """--workload hostlink: the host link's own ceiling for the e2e call's traffic.  Every rank copies exactly the bytes one
    `umpcBatchEnvStep` call moves (108 B per env each way, pinned buffers) with plain asynchronous copies -- host->device and
    device->host on two streams at once, no kernel -- and reports bytes/s per rank and in all, and the env-steps/s a call
    that did NOTHING but move its bytes could reach.  VERDICT r01 item 8: the measurement to hold the 8-GPU e2e figure against."""
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    _bind_to_gpu_numa_node(local)
    nenv = args.envs
    nbytes = 108 * nenv
    hin = torch.empty(nbytes, dtype=torch.uint8).pin_memory(); hout = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    din = torch.empty(nbytes, dtype=torch.uint8, device="cuda"); dout = torch.zeros(nbytes, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    calls = 125

    def step():
        for _ in range(calls):
            with torch.cuda.stream(s1):
                din.copy_(hin, non_blocking=True)
            with torch.cuda.stream(s2):
                hout.copy_(dout, non_blocking=True)
    for _ in range(max(args.warmup, 3)):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    s1.wait_stream(torch.cuda.current_stream()); s2.wait_stream(torch.cuda.current_stream())
    for _ in range(args.steps):
        step()
    torch.cuda.current_stream().wait_stream(s1); torch.cuda.current_stream().wait_stream(s2)
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) * 1e-3], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.barrier()
    sec = float(t[0])
    if rank == 0:
        per_dir = nbytes * calls * args.steps / sec
        print(json.dumps({"metric": "host link ceiling for the e2e call's bytes (plain async copies, both directions at once)",
                          "value": 2 * per_dir * world * 1e-9, "unit": "GB/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
                          "ms_per_step": 1e3 * sec / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                          "dtype": "u8", "data": "synthetic",
                          "config": {"workload": "hostlink: %d envs per GPU x 108 B each way per call, %d calls per step, pinned host buffers" % (nenv, calls)},
                          "gb_per_s_per_rank_each_way": per_dir * 1e-9, "ms_per_call": 1e3 * sec / (args.steps * calls),
                          "env_steps_per_s_if_a_call_only_moved_its_bytes": nenv * calls * args.steps * world / sec,
                          "gpu_launches": 0}))
    if world > 1:
        dist.destroy_process_group()
    return 0


def run_flapsweep(args):
    """--workload flapsweep (row f2): sweepFile's open-loop wing simulations (sdab_num_wrenchmap_simgen.py:19-75), the
    reference's 8 x 8 x 1 x 8 x 8 grid (4,096 simulations x 1,001 steps of 0.1 ms) or a finer grid with --envs (rounded to k^4),
    all in ONE launch.  value = simulation steps/s (one step of one two-wing simulation); e2e = the same through host arrays
    (grid in, numkins rows out); cpu_baseline = the numpy oracle (vectorised over simulations) on a sample."""
    from robobee3d_b200 import flapping
    k = 8 if args.envs == NENV_PER_GPU else max(2, int(round(args.envs ** 0.25)))
    grids = (np.linspace(90, 160, k), np.linspace(-0.5, 0.5, k), [0.17], np.linspace(-0.2, 0.2, k), np.linspace(-0.2, 0.2, k))
    nsim, nst = k ** 4, flapping.open_loop_steps(100.0, 0.1)
    clocks = ClockSampler(0)
    for _ in range(max(args.warmup, 3)):
        flapping.sweep(*grids)
    clocks.start()
    ms, wall = [], []
    for _ in range(args.steps):
        t0 = time.perf_counter()
        res, m = flapping.sweep(*grids, return_ms=True)
        wall.append(time.perf_counter() - t0); ms.append(m)
    ck = clocks.finish()
    assert np.isfinite(res).all()
    kms = float(np.mean(ms))
    L = flapping._bind()
    peak64 = L.umpcMeasureFp64Peak(0, 1 << 16)
    fl = 2 * 721.0 * nsim * nst      # executed DFMA(2) + DMUL + DADD per wing and step in the kernel's SASS, sincos polynomials included
    line = {"metric": "open-loop flapping-wing simulation steps/s (sweepFile: two wings, quasi-steady aero, fixed base)",
            "value": nsim * nst / (kms * 1e-3), "unit": "sim-steps/s", "n_gpus": 1, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": kms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "sdab_num_wrenchmap_simgen.py grid %d^4 = %d simulations x %d steps of 0.1 ms, one launch; inputs are KB: "
                                   "nothing to flush between launches" % (k, nsim, nst)},
            "gpu_launches": args.steps, "clocks": ck,
            "roofline": {"bound": "fp64", "achieved": fl / (kms * 1e-3) * 1e-12, "peak": peak64, "unit": "TFLOP/s",
                         "frac": fl / (kms * 1e-3) * 1e-12 / peak64, "traffic": None, "kernel": "fw::sweep_kernel",
                         "note": "%d threads (one per wing) on 148 SMs: the reference's grid does not fill the GPU; latency-bound "
                                 "dependent chain of ~1,250 instructions per step" % (2 * nsim)},
            "e2e": {"value": nsim * nst / float(np.median(wall)), "unit": "sim-steps/s", "h2d_bytes_per_step": 40 * nsim,
                    "d2h_bytes_per_step": 120 * nsim, "call": "flapping.sweep (fwOpenLoopSweep with host arrays)"}}
    if not args.no_cpu:
        from oracle import flapping_oracle as fo
        ns = min(nsim, 1024)
        t0 = time.perf_counter()
        fo.sweep_features(res[:ns, :5])
        tc = time.perf_counter() - t0
        line["cpu_baseline"] = {"value": ns * nst / tc, "unit": "sim-steps/s", "cores": 1, "kind": "port",
                                "sample": "%d of the %d simulations, numpy oracle vectorised over simulations (the reference steps pybullet, absent here)" % (ns, nsim)}
    print(json.dumps(line))
    return 0


def run_small(args):
    """BASELINE configs[0] and configs[1] (extra lines, not the driver's headline):
    c1     : ONE controlTest flight (tend 500 ms, hlInterval 5 ms: ~96 controller calls) through the drop-in
             uprightmpc2py.UprightMPC2C -- microseconds per update call of a batch of 1, next to the reference C call.
    qp4096 : 4,096 independent UprightMPC2 QPs from random upright states, solve only: cold call + warm repeats."""
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product has no CPU path")
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from robobee3d_b200 import build as _build
    _build.build()
    clocks = ClockSampler(0)
    clocks.start()
    if args.workload == "c1":
        from robobee3d_b200 import uprightmpc2py
        from robobee3d_b200.sweeps import control_test_ic
        from robobee3d_b200.batch import DEFAULTS as D
        mk = lambda: uprightmpc2py.UprightMPC2C(D["dt"], D["g"], D["TtoWmax"], D["ws"], D["wds"], D["wpr"], D["wpf"], D["wvr"],
                                                D["wvf"], D["wthrust"], D["wmom"], D["Ib"], D["maxIter"])

        def flight(ctrl_update):
            p, R, dq = control_test_ic()
            uq, thl, calls, tsum = np.zeros(3), 0.0, [], 0.0
            pdes, dpdes, sdes = np.zeros(3), np.zeros(3), np.array([0, 0, 1.0])
            for ti in range(2500):
                t = ti * 0.2
                if t - thl > 5.0:
                    t0 = time.perf_counter()
                    uq, _ = ctrl_update(p, R, dq, pdes, dpdes, sdes)
                    calls.append(time.perf_counter() - t0)
                    uq = np.array(uq, float); thl = t
                uq[1:] = np.clip(uq[1:], -100, 100)
                p, R, dq = _np_dyn_step(p, R, dq, uq, 0.2)
            return np.array(calls), p
        up = mk()
        flight(up.update)                     # warm-up flight (first launch, module load)
        per_flight = []
        for _ in range(max(1, args.steps)):
            up = mk()
            calls, pf = flight(up.update)
            per_flight.append(calls)
        calls = np.concatenate(per_flight)
        line = {"metric": "microseconds per umpcUpdate of ONE controller (drop-in, batch of 1)", "value": 1e6 * float(np.median(calls)),
                "unit": "us/update", "n_gpus": 1, "steps": args.steps, "warmup": 1, "ms_per_step": 1e3 * float(calls.sum()) / max(1, args.steps),
                "higher_is_better": False, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": "BASELINE configs[0]: template/uprightmpc2.py controlTest(upc, 500, hlInterval=5), hover from the tilted IC, "
                                       "%d controller calls per flight through uprightmpc2py.UprightMPC2C (ctypes -> umpcUpdate -> one 128-thread "
                                       "CTA reading/writing a pinned mapped block)" % len(per_flight[0])},
                "us_mean": 1e6 * float(calls.mean()), "us_p95": 1e6 * float(np.percentile(calls, 95)), "final_pos_norm_mm": float(np.linalg.norm(pf)),
                "gpu_launches": int(len(calls)),
                "e2e": {"value": 1e6 * float(np.median(calls)), "unit": "us/update", "h2d_bytes_per_step": 28 * 4, "d2h_bytes_per_step": (9 + 216) * 4,
                        "note": "the call IS end to end: host arrays in, host arrays + UprightMPC_t.l/u/q/Px/Ax out"},
                "roofline": {"bound": "latency", "achieved": None, "peak": None, "unit": "us", "frac": None, "traffic": None,
                             "note": "a batch of 1 is the serial chain of one solve (~1.8e5 SM cycles) + launch + PCIe round trip; throughput "
                                     "rooflines do not apply"}}
        if not args.no_cpu:
            from oracle import binding as OB
            if OB.have_ref():
                ref = OB.RefUMPC()
                flight(lambda *a: ref.update(*a, -1.0))
                ref = OB.RefUMPC()
                rc, _ = flight(lambda *a: ref.update(*a, -1.0))
                line["cpu_baseline"] = {"value": 1e6 * float(np.median(rc)), "unit": "us/update", "cores": 1, "kind": "reference",
                                        "sample": "the same flight through the reference C umpcUpdate (oracle/_ref, -O2) via ctypes"}
    else:
        import util
        from robobee3d_b200.batch import BatchUMPC
        B = args.envs if args.envs != NENV_PER_GPU else 4096
        p, R, dq = util.config2_states(B, seed=0)
        Rc = np.ascontiguousarray(R.transpose(0, 2, 1).reshape(B, 9), np.float32)
        dev = lambda a: torch.from_numpy(np.ascontiguousarray(a, np.float32)).cuda()
        ins = [dev(p), dev(Rc), dev(dq), dev(np.zeros((B, 3))), dev(np.zeros((B, 3))), dev(np.tile([0, 0, 1.0], (B, 1)))]
        aT0 = dev(np.full(B, 9.81e-3))
        uq, acc = torch.empty(B, 3, device="cuda"), torch.empty(B, 6, device="cuda")
        b = BatchUMPC(B)
        cold, warm = [], []
        for rep in range(max(3, args.warmup) + args.steps):
            b.reset()
            b.update_dev(uq, acc, *ins, aT0, stream=None); b.sync()
            c = b.last_kernel_ms()
            b.update_dev(uq, acc, *ins, None, stream=None); b.sync()
            w = b.last_kernel_ms()
            if rep >= max(3, args.warmup):
                cold.append(c); warm.append(w)
        # e2e: the same two calls with host buffers (umpcBatchUpdate)
        z3 = np.zeros((B, 3), np.float32); e3 = np.tile(np.float32([0, 0, 1.0]), (B, 1))
        t0 = time.perf_counter()
        for _ in range(args.steps):
            b.update(p, R, dq, z3, z3, e3, np.full(B, 9.81e-3, np.float32))
        te = (time.perf_counter() - t0) / args.steps
        ms = float(np.median(warm))
        line = {"metric": "UprightMPC2 QP solves/s (solve only, 4,096 independent QPs)", "value": B / (ms * 1e-3), "unit": "QP-solves/s",
                "n_gpus": 1, "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": "BASELINE configs[1]: %d independent UprightMPC2 QPs from random upright states (SURVEY 8d config 2), QP solve "
                                       "only (umpcBatchUpdateDev, device-resident inputs): cold start, then one warm repeat; value = warm" % B,
                           "l2_policy": "whole working set (%.1f MB) is L2 resident; every call re-reads its own state once" % (B * 1.5e-3)},
                "cold_ms": float(np.median(cold)), "warm_ms": ms, "cold_solves_per_s": B / (float(np.median(cold)) * 1e-3),
                "gpu_launches": 2 * args.steps,
                "e2e": {"value": B / te, "unit": "QP-solves/s", "h2d_bytes_per_step": 28 * 4 * B, "d2h_bytes_per_step": 9 * 4 * B,
                        "call": "umpcBatchUpdate with pageable host arrays"},
                "roofline": {"bound": "latency", "achieved": None, "peak": None, "unit": "TFLOP/s", "frac": None, "traffic": None,
                             "note": "4,096 QPs = 128 warps on 148 SMs x 8 warp slots: the launch is one solve's serial latency; "
                                     "the GPU is ~11%% occupied -- see the 65,536-env line for the throughput regime"}}
        if not args.no_cpu:
            from oracle import binding as OB
            cores = os.cpu_count() or 1
            ics = np.concatenate([p, R.transpose(0, 2, 1).reshape(-1, 9), dq], axis=1)[: cores * 64]
            done, tmax, _, kind = OB.cpu_env_steps(ics, 50, cores)
            line["cpu_baseline"] = {"value": done / tmax, "unit": "QP-solves/s", "cores": cores, "kind": kind,
                                    "sample": "%d of the QPs x 50 warm closed-loop solves, one process per core" % len(ics)}
        b.close()
    line["clocks"] = clocks.finish()
    print(json.dumps(line))
    return 0


def gqp_flops(n, m, nnz, iters, scaling_iters=10, refactors=1):
    """executed-algorithm flops of one gqp solve (gqp.cu): Ruiz passes, K formation + dense Cholesky, ADMM iterations"""
    ruiz = scaling_iters * (4 * n * n + 6 * nnz + 8 * n + 4 * m)
    factor = refactors * (n * n + n ** 3 / 3.0 + 6 * nnz)
    it = 2 * n * n + 8 * nnz + 12 * m + 6 * n
    return ruiz + factor + iters * it


def run_gqp(args):
    """Extra lines (not the driver's headline) for the rows that run on the generic warp-per-QP solver:
    quad       : BASELINE configs[4]'s namesake -- the LTI quadcopter MPC of planar/mpc_osqp.py (n = 92, m = 164), B
                 closed loops of 15 steps each on the device (gqpLtiRollout); value = QP solves/s
    fused-wlqp : BASELINE configs[3]'s single-QP variant -- uprightmpc2_withWLQP (nx = 49, nc = 43, eps 1e-7): one
                 umgUpdate of B controllers per step (host buffers: the call IS the e2e number; value = the solve kernel alone)
    umpc-n     : UprightMPC2 at horizon --horizon through the same path (f1)"""
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product has no CPU path")
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import util
    from robobee3d_b200 import build as _build, _lib
    from robobee3d_b200 import gqp as G
    _build.build()
    L = _lib.load()
    clocks = ClockSampler(0)
    clocks.start()
    B = args.envs if args.envs != NENV_PER_GPU else (16384 if args.workload == "quad" else 32768)
    if args.workload == "quad":
        m = G.mpc_osqp_problem()
        rp, ci = G.csr_pattern(m["A"])
        b = G.BatchQP(B, m["n"], m["m"], rp, ci, eps_abs=1e-2, eps_rel=1e-2, adaptive_rho_interval=25)
        rng = np.random.default_rng(0)
        x0 = np.tile(m["x0"], (B, 1)); x0[:, 2] = rng.uniform(-0.5, 0.5, B); x0[:, :2] = rng.uniform(-0.1, 0.1, (B, 2))
        u_off, nst = m["n"] - m["N"] * m["nu"], 15
        ms, its = [], []
        for rep in range(max(3, args.warmup) + args.steps):
            b.reset()
            l, u = np.tile(m["l"], (B, 1)), np.tile(m["u"], (B, 1)); l[:, :12] = -x0; u[:, :12] = -x0
            b.set_data(P=m["P"], Av=b.csr_values(m["A"]), q=m["q"], broadcast=True)
            b.set_data(l=l, u=u)
            xf, _, it = b.lti_rollout(nst, m["Ad"], m["Bd"], x0, u_off, rescale_first=True)
            if rep >= max(3, args.warmup):
                ms.append(b.last_kernel_ms()); its.append(it)
        assert np.isfinite(xf).all()
        msm, itm = float(np.median(ms)), float(np.median(its)) / (B * nst)
        n_, m_, nnz = m["n"], m["m"], len(ci)
        fl = B * nst * gqp_flops(n_, m_, nnz, itm, refactors=1.0 / nst, scaling_iters=10.0 / nst)
        name, unit, solves = "LTI quadcopter MPC (planar/mpc_osqp.py) QP solves/s", "QP-solves/s", B * nst
        cfgw = "planar/mpc_osqp.py:29-140: n = 92, m = 164, eps 1e-2, warm start; %d closed loops x %d steps per launch sequence, initial states perturbed per loop" % (B, nst)
        e2e = None
        oracle_job = ("quad", m)
    else:
        wl = args.workload == "fused-wlqp"
        N = 3 if wl else args.horizon
        p, R, dq = util.config2_states(B, seed=0)
        z3, e3 = np.zeros((B, 3)), np.tile([0, 0, 1.0], (B, 1))
        b = G.BatchUprightMPC2G(B, N=N, withWL=wl)
        ms, wall, its = [], [], []
        for rep in range(max(3, args.warmup) + args.steps):
            b.reset()
            t0 = time.perf_counter()
            uq, acc, uwl = b.update(p, R, dq, z3, z3, e3, np.full(B, 9.81e-3))
            t1 = time.perf_counter() - t0
            if rep >= max(3, args.warmup):
                ms.append(b.last_solve_ms()); wall.append(t1); its.append(float(b.qp()["iter"].mean()) if rep == max(3, args.warmup) else its[-1])
        assert np.isfinite(uq).all()
        msm, itm = float(np.median(ms)), its[0]
        n_, m_, nnz = b.nx, b.nc, b.nnz
        fl = B * gqp_flops(n_, m_, nnz, itm, refactors=1 + itm / 400.0)
        name = ("fused UprightMPC2 + WLQP single QP (uprightmpc2_withWLQP) solves/s" if wl else "UprightMPC2 N=%d QP solves/s (generic horizon)" % N)
        unit, solves = "QP-solves/s", B
        cfgw = ("template/uprightmpc2_withWLQP.py: nx = 49, nc = 43, dense P blocks, eps 1e-7, max_iter 4000; " if wl else
                "template_controllers.UprightMPC2(N=%d): nx = %d, nc = %d, eps 1e-4; " % (N, n_, m_)) + "%d controllers, cold start, config-2 random upright states" % B
        e2e = {"value": B / float(np.median(wall)), "unit": unit, "h2d_bytes_per_step": 28 * 8 * B, "d2h_bytes_per_step": 13 * 8 * B,
               "call": "umgUpdate with host arrays (build kernel + warp-per-QP solve + extract kernel inside)"}
        oracle_job = None
    peak64 = L.umpcMeasureFp64Peak(0, 1 << 16)
    ach = fl / (msm * 1e-3) * 1e-12
    line = {"metric": name, "value": solves / (msm * 1e-3), "unit": unit, "n_gpus": 1, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": msm, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": cfgw, "l2_policy": "every QP's data is read once per solve; state (<= 100 MB) is L2 resident"},
            "iters_per_solve": itm, "gpu_launches": args.steps * (15 if args.workload == "quad" else 3),
            "roofline": {"bound": "fp64", "achieved": ach, "peak": peak64, "unit": "TFLOP/s", "frac": ach / peak64, "traffic": None,
                         "kernel": "gqp_solve_kernel",
                         "note": "one warp per QP, factor and vectors in shared memory: the column sweeps of the triangular solves are "
                                 "__syncwarp-latency bound, not pipe bound; peak = DFMA microbenchmark in this process"}}
    if e2e:
        line["e2e"] = e2e
    if not args.no_cpu and oracle_job:
        from oracle import binding as OB
        mm = oracle_job[1]
        t0 = time.perf_counter(); nq = 0
        for e in range(8):
            o = OB.OracleQP(mm["n"], mm["m"], eps=1e-2, adapt=25)
            l, u = mm["l"].copy(), mm["u"].copy(); xx = x0[e].copy()
            l[:12] = -xx; u[:12] = -xx
            xs, _, _ = o.update_solve(mm["P"], mm["A"], mm["q"], l, u)
            for k in range(14):
                xx = mm["Ad"] @ xx + mm["Bd"] @ xs[u_off:u_off + 4]; l[:12] = -xx; u[:12] = -xx
                xs, _, _ = o.update_solve(None, None, mm["q"], l, u)
            nq += 15
        dtc = time.perf_counter() - t0
        line["cpu_baseline"] = {"value": nq / dtc, "unit": unit, "cores": 1, "kind": "port",
                                "sample": "8 of the loops x 15 steps through the oracle's dense OSQP restatement on one core (the reference's "
                                          "solver here is the un-vendored osqp wheel)"}
    line["clocks"] = clocks.finish()
    print(json.dumps(line))
    b.close()
    return 0


def run_dirtran(args):
    """BASELINE configs[4]: planar flapping dir-tran MPC (mpc_thrust_strokedev.py runSim(CTRL_MPC)), B envs x horizon N per
    GPU.  A step = `--dt-steps` closed-loop MPC periods of every env in ONE launch of dt_rollout_kernel (update -> 60 Euler
    substeps per period); value = env-steps/s = QP solves/s.  Extra measurement, not the driver's headline line."""
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product has no CPU path")
    torch.cuda.set_device(local)
    if world > 1:
        _init_nccl(dist, torch, local)
    from robobee3d_b200 import build as _build
    from robobee3d_b200.dirtran import BatchLTVMPC
    if rank == 0:
        _build.build()
    if world > 1:
        dist.barrier()
    nenv, N, S = args.envs, args.horizon, args.dt_steps
    ys, xr = _dt_mild_states(rank * nenv, (rank + 1) * nenv)
    b = BatchLTVMPC(nenv, N=N, device=local)
    b.set_state(y=ys, xr=xr)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        b.rollout(S)
    barrier()
    c0 = b.counters()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    dev_ms = 0.0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        b.rollout(S)                       # synchronous; CUDA events on the library's stream bracket the launch
        dev_ms += b.last_kernel_ms()
    barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.finish() if rank == 0 else None
    c1 = b.counters()
    its = c1["iters"] - c0["iters"]
    # guards the uniform-register workaround of mpc_run: every lane must have done all its MPC periods
    assert c1["solves"] - c0["solves"] == nenv * S * args.steps, ("dt_rollout_kernel lost solves", c1["solves"] - c0["solves"])
    t = torch.tensor([dev_ms * 1e-3, wall], dtype=torch.float64, device="cuda")
    cnt = torch.tensor([its, c1["unsolved"] - c0["unsolved"]], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    dev_s, wall = float(t[0]), float(t[1])
    # second regime: the TRANSIENT -- the same envs with COLD controllers (no warm start, rho at its default): the first
    # MPC periods after a reset take ~1.8x the iterations of the at-rest regime above; every timed launch restarts cold
    ytr = ys
    tr_ms, tr_its = 0.0, 0.0
    for _ in range(args.steps):
        b.reset(); b.set_state(y=ytr, xr=xr)
        ca = b.counters()
        b.rollout(S)
        tr_ms += b.last_kernel_ms()
        tr_its += b.counters()["iters"] - ca["iters"]
    ttr = torch.tensor([tr_ms * 1e-3, tr_its], dtype=torch.float64, device="cuda")
    if world > 1:
        m0 = ttr[0:1].clone(); dist.all_reduce(m0, op=dist.ReduceOp.MAX); dist.all_reduce(ttr, op=dist.ReduceOp.SUM); ttr[0] = m0[0]
    b.reset(); b.set_state(y=ys, xr=xr)
    # e2e: LTVMPC.update through the C-ABI with host buffers (x0 in, u out), the call mpc_thrust_strokedev.py:171 makes
    e2e = None
    if not args.no_e2e:
        yh, _ = b.state()
        b.update(yh)
        barrier()
        t0 = time.perf_counter()
        ne = 3
        for _ in range(ne):
            u = b.update(yh)
        barrier()
        te = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        assert np.isfinite(u).all()
        e2e = {"value": nenv * world * ne / float(te[0]), "unit": UNIT, "h2d_bytes_per_step": 48 * nenv, "d2h_bytes_per_step": 16 * nenv,
               "call": "dtBatchUpdate (host x0[B][6] in, u[B][2] out; one LTVMPC.update per env at rest near the goal)"}
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        its_all = float(cnt[0])
        bytes_per_it = (N + 1) * 76 * 8           # unique words per knot and iteration (factor 36 R, iterates 20 R + 20 W)
        ach = (its / args.steps) * bytes_per_it / (dev_s / args.steps) * 1e-9   # this rank's launch
        line = {"metric": "closed-loop planar dir-tran MPC env-steps/s (= QP solves/s), envs at rest about the hover goal (value) and in the "
                          "transient right after a cold start of every controller (value_transient)", "value": nenv * world * S * args.steps / dev_s,
                "value_at_rest": nenv * world * S * args.steps / dev_s, "value_transient": nenv * world * S * args.steps / float(ttr[0]),
                "iters_per_solve_transient": float(ttr[1]) / (nenv * world * S * args.steps),
                "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dev_s / args.steps,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": "BASELINE configs[4]: planar mpc_osqp flapping MPC (PlanarThrustStrokeDev, LTVMPC N=%d, OSQP eps 1e-5, "
                                       "check 25, adaptive rho 25), %d envs per GPU about the hover goal, %d MPC periods per launch" % (N, nenv, S),
                           "envs_per_gpu": nenv, "horizon": N, "mpc_periods_per_launch": S,
                           "l2_policy": "slab of %.2f GB per GPU streams through L2 every ADMM iteration" % ((44 + 74 * (N + 1)) * 8 * nenv / 1e9)},
                "qp_iters_per_s": its_all / dev_s, "iters_per_solve": its_all / (nenv * world * S * args.steps), "unsolved": float(cnt[1]),
                "wall_ms_per_step": 1e3 * wall / args.steps, "gpu_launches": args.steps, "clocks": clocks,
                "roofline": {"bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak,
                             "traffic": DT_TRAFFIC.get((nenv, N, S)), "kernel": "dt_rollout_kernel",
                             "peak_source": "MEASURED_PEAKS.json" if "hbm_gbs" in peaks else "fallback",
                             "note": "algorithmic bytes = ADMM iterations of the launch x (N+1) knots x 608 B (unique words: 36 factor "
                                     "words read, 20 iterate words read and written -- x 8, z/y of the 3 finitely bounded rows, y of the "
                                     "6 dynamics rows); the two sweeps ISSUE 148 words (1,184 B) per knot "
                                     "and iteration, of which L2 serves a part: `traffic` is what reached DRAM"}}
        if e2e:
            line["e2e"] = e2e
        if world == 1 and not args.no_cpu:
            import multiprocessing as mp
            cores = os.cpu_count() or 1
            per, cs = 16, 20
            jobs = [(ys[i * per:(i + 1) * per], N, cs) for i in range(cores)]
            t0 = time.perf_counter()
            with mp.get_context("spawn").Pool(cores) as pool:
                res = pool.map(_dt_cpu_worker, jobs)
            tmax = max(r[0] for r in res)
            line["cpu_baseline"] = {"value": cores * per * cs / tmax, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": "%d of the %d envs x %d MPC periods, one process per host core (oracle port of "
                                              "LTVMPC + OSQP, dense LDL')" % (cores * per, nenv, cs)}
        print(json.dumps(line))
    b.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


# dram__bytes_read.sum + dram__bytes_write.sum of one umpc_rollout_kernel launch from the ncu --set full capture under profiles/
# (key: envs, plant steps per launch); from profiles/r02m_ncu_full_umpc_rollout_kernel.txt (69.2 MB read + 1.498 GB written;
# the L2 write-back of the slab's state rows every step, not algorithmic traffic)
ROLLOUT_TRAFFIC = {(65536, 125): 1.568e9}
# dram__bytes_read.sum + dram__bytes_write.sum of one dt_rollout_kernel launch from the ncu --set full captures under
# profiles/ (key: envs, horizon, MPC periods per launch)
DT_TRAFFIC = {(65536, 10, 10): 152.38e9}   # profiles/r01j_ncu_full_dt_rollout_kernel.txt: 110.69 GB read + 41.68 GB written


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--envs", type=int, default=NENV_PER_GPU, help="envs per GPU (default: the BASELINE config)")
    ap.add_argument("--tpb", type=int, default=0, help="threads per block override (tuning)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--hl-interval", type=float, default=None,
                    help="MPC period in ms (controlTest's hlInterval; config 1 uses 5): a step then advances every env by one "
                         "MPC period (26 plant steps at 5 ms) in one launch.  Extra measurement; the headline line uses None")
    ap.add_argument("--steps-per-launch", type=int, default=0,
                    help="plant steps (each with an MPC solve when hlInterval is None) advanced per launch of the rollout kernel")
    ap.add_argument("--horizon", type=int, default=10, help="--workload dirtran: knots N (3, 5, 10, 15, 20)")
    ap.add_argument("--dt-steps", type=int, default=10, help="--workload dirtran: MPC periods per launch")
    ap.add_argument("--workload", default="mc", choices=["mc", "cascade-wlcon", "cascade-wlqp", "dirtran", "c1", "qp4096", "quad", "fused-wlqp", "umpc-n", "hostlink", "flapsweep"],
                    help="mc: BASELINE configs[2] (the headline line); cascade-*: configs[3], UprightMPC2 -> WL stage -> "
                         "actualT0 feedback per env-step (extra measurement, not the driver's line)")
    ap.add_argument("--converged", action="store_true",
                    help="converged mode (the reference's Python path: float64 ADMM to eps 1e-4, check every 25, adaptive rho "
                         "every 100, max 4000).  Extra measurement, not the driver's line")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        return run_reference(args)
    if args.workload == "dirtran":
        return run_dirtran(args)
    if args.workload in ("c1", "qp4096"):
        return run_small(args)
    if args.workload == "hostlink":
        return run_hostlink(args)
    if args.workload == "flapsweep":
        return run_flapsweep(args)
    if args.workload in ("quad", "fused-wlqp", "umpc-n"):
        return run_gqp(args)

    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        _bind_to_gpu_numa_node(local)
        _init_nccl(dist, torch, local)
    from robobee3d_b200 import build as _build
    from robobee3d_b200.batch import BatchUMPC, gather_metrics, shard_range
    from robobee3d_b200 import flops as F
    if rank == 0:
        _build.build()
    if world > 1:
        dist.barrier()
    nenv = args.envs
    # global env ids [rank*nenv, (rank+1)*nenv): seed keyed by the global range -> shard-invariant ICs
    p, R, dq = fast_ics(nenv * world, seed=0)
    lo, hi = shard_range(nenv * world, rank, world)
    p, R, dq = p[lo:hi], R[lo:hi], dq[lo:hi]
    kw = {}
    if args.workload != "mc":
        import util
        kw = dict(wlMode=1 if args.workload == "cascade-wlcon" else 2, wl_popts=util.load_constants()["popts_waypoint"])
        args.no_e2e = True  # the host-buffer env-step call has no WL stage
    if args.converged:
        kw.update(maxIter=4000, checkTermination=25, adaptiveRhoInterval=100)
        args.no_e2e = True
    spl = 1 if args.converged else STEPS_PER_LAUNCH  # plant steps per launch (converged mode: one step per launch is its fastest shape, profiles/r02C_*)
    if args.hl_interval is not None:
        kw["hlInterval"] = args.hl_interval
        spl = int(np.floor(args.hl_interval / 0.2 + 1e-9)) + 1  # t - thlPrev > hlInterval on a 0.2 ms grid
        args.no_e2e = True
    if args.steps_per_launch:
        spl = args.steps_per_launch
    b = BatchUMPC(nenv, device=local, **kw)
    if args.tpb:
        b.set_launch(args.tpb)
    b.set_state(p, R, dq, task=0, tp=(0.0, 0.0, 0.1, 0.0))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident leg ------------------------------------------------------------------
    for _ in range(args.warmup):
        b.rollout(spl, sync=False)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    # CUDA events recorded on the stream the kernels are launched on (the library's own stream)
    t0 = time.perf_counter()
    b.timer_start()
    for _ in range(args.steps):
        b.rollout(spl, sync=False)
    dev_s = b.timer_stop_ms() * 1e-3
    barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.finish() if rank == 0 else None
    t = torch.tensor([dev_s, wall], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_s, wall = float(t[0]), float(t[1])
    sums = gather_metrics(b.reduce_metrics())

    # ---- e2e leg: host buffers through the C-ABI -------------------------------------------------
    e2e = None
    if not args.no_e2e:
        pin = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).pin_memory()
        hp, hR, hdq = pin(p), pin(R.transpose(0, 2, 1).reshape(nenv, 9)), pin(dq)
        hpd, hdp, hsd = pin(np.zeros((nenv, 3))), pin(np.zeros((nenv, 3))), pin(np.tile([0, 0, 1.0], (nenv, 1)))
        huq, hacc = pin(np.zeros((nenv, 3))), pin(np.zeros((nenv, 6)))
        b2 = BatchUMPC(nenv, device=local)
        if args.tpb:
            b2.set_launch(args.tpb)
        # one e2e "step" = the same `spl` env-steps of every env, each through umpcBatchEnvStep with host buffers
        ne2e = spl if args.hl_interval is None else 1
        for _ in range(args.warmup):
            b2.env_step_raw(huq, hacc, hp, hR, hdq, hpd, hdp, hsd)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps * ne2e):
            b2.env_step_raw(huq, hacc, hp, hR, hdq, hpd, hdp, hsd)
        barrier()
        te = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        assert torch.isfinite(hp).all()
        e2e = {"value": nenv * world * args.steps * ne2e / float(te[0]), "unit": UNIT, "h2d_bytes_per_step": 27 * 4 * nenv * ne2e,
               "d2h_bytes_per_step": 27 * 4 * nenv * ne2e, "ms_per_step": 1e3 * float(te[0]) / args.steps,
               "calls_per_step": ne2e, "pull_window_ctas": int(os.environ.get("UMPC_E2E_WINDOW", "32")), "ms_per_call": 1e3 * float(te[0]) / (args.steps * ne2e), "gpu_launches": args.steps * ne2e,
               "call": "umpcBatchEnvStep (pinned host buffers read and written in place by one zero-copy kernel; umpcUpdate + clip + "
                       "quadrotorNLDyn per env)"}
        # second host-buffer call of the API, the one a caller of the reference's controlTest makes: initial conditions in,
        # `spl` closed-loop steps on the device, final states + logMetric sums out (reported beside, not instead of, the above)
        if args.hl_interval is None and not args.tpb:
            from robobee3d_b200.batch import _vp
            pz = lambda k: torch.zeros((nenv, k), dtype=torch.float32).pin_memory().numpy()   # pinned host arrays
            gp, gR, gdq, guq, gm = pz(3), pz(9), pz(6), pz(3), pz(4)
            sp, sR, sdq = pz(3), pz(9), pz(6)
            sp[:], sR[:], sdq[:] = hp.numpy(), hR.numpy(), hdq.numpy()

            def flight():
                L = b2.lib
                L.umpcBatchSetState(b2.h, _vp(sp), _vp(sR), _vp(sdq), None, None)
                L.umpcBatchRollout(b2.h, spl)
                L.umpcBatchGetState(b2.h, _vp(gp), _vp(gR), _vp(gdq), _vp(guq))
                L.umpcBatchGetMetrics(b2.h, _vp(gm))
            for _ in range(args.warmup):
                flight()
            barrier()
            t0 = time.perf_counter()
            for _ in range(args.steps):
                flight()
            barrier()
            tf = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
            if world > 1:
                dist.all_reduce(tf, op=dist.ReduceOp.MAX)
            assert np.isfinite(gp).all()
            e2e["flight_call"] = {"value": nenv * world * args.steps * spl / float(tf[0]), "unit": UNIT, "ms_per_step": 1e3 * float(tf[0]) / args.steps,
                                  "h2d_bytes_per_step": 18 * 4 * nenv, "d2h_bytes_per_step": (21 + 4) * 4 * nenv,
                                  "call": "umpcBatchSetState (host ICs) + umpcBatchRollout(%d) + umpcBatchGetState + umpcBatchGetMetrics (pinned host "
                                          "arrays): the controlTest-shaped call, plant on the device" % spl}
        b2.close()

    mean_iter = float(b.solve_info()["iter"].mean()) if args.converged else 50.0
    if rank == 0:
        fl = F.our_flops(mean_iter)
        if args.converged:   # the residual block runs after every 25 iterations, not once
            fl = dict(fl, total=fl["setup"] + mean_iter * fl["per_iter"] + max(mean_iter / 25.0, 1.0) * fl["post"])
        solves_per_launch = spl if args.hl_interval is None else 1
        per_launch_flops = (solves_per_launch * fl["total"] + spl * F.DYN_STEP_FLOPS) * nenv
        launch_s = dev_s / args.steps
        # FFMA microbenchmark in this process, right after the timed legs (same thermal / clock state): a long run
        # (sustained, ~0.25 s: the denominator for a step that itself runs for tens of ms) and a short one (burst)
        peak_burst = b.lib.umpcMeasureFp32Peak(local, 4096)
        peak_tf = b.lib.umpcMeasureFp32Peak(local, 1 << 18)
        if args.converged:   # float64 kernel: the DFMA pipe is the roof
            peak_burst = peak_tf = b.lib.umpcMeasureFp64Peak(local, 1 << 16)
        achieved_tf = per_launch_flops / launch_s * 1e-12
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        roofline = {"bound": "fp64" if args.converged else "fp32", "achieved": achieved_tf, "peak": peak_tf, "unit": "TFLOP/s", "frac": achieved_tf / peak_tf,
                    # dram__bytes_read.sum + dram__bytes_write.sum of this kernel from the ncu --set full capture summarised in
                    # profiles/r01e_ncu_full_umpc_rollout_kernel_tmem.txt (42.4 + 35.3 MB per launch of 65,536 envs; the
                    # algorithmic figure is 2,172 B x 65,536 = 142 MB -- part of the 96 MB state slab stays in the 126 MB L2
                    # from one launch to the next); only known for that launch shape
                    "traffic": ROLLOUT_TRAFFIC.get((nenv, spl)) if (args.workload == "mc" and not args.converged and args.hl_interval is None) else None,
                    "peak_burst": peak_burst,
                    "kernel": "umpc_rollout_kernel<128, SolverF64T<TmemStore64>>" if args.converged else "umpc_rollout_kernel",
                    "note": ("FP64 pipe roof (DFMA microbenchmark in this process); achieved = executed-algorithm flops at the lanes' MEAN "
                             "iteration count (%.1f; a warp runs to its slowest lane's count) x envs / CUDA-event launch time; the loop is bound "
                             "by the per-warp tensor-memory read rate, the rest by instruction fetch at 4 warps per SM (DESIGN 4.1)" % mean_iter)
                            if args.converged else "FP32 CUDA-core pipe bound (arithmetic intensity %.0f flop/B vs ridge ~11): peak = FFMA "
                            "microbenchmark run in this process (MEASURED_PEAKS.json has no FP32 figure); achieved = "
                            "executed-algorithm flops %d/solve (FMA=2; reference formulation would be %d) x envs / "
                            "CUDA-event launch time" % (fl["total"] / F.BYTES_PER_SOLVE, fl["total"],
                                                        F.reference_algorithm_flops(50)),
                    "hbm": {"achieved": F.BYTES_PER_SOLVE * nenv / launch_s * 1e-9, "peak": hbm_peak, "unit": "GB/s",
                            "frac": F.BYTES_PER_SOLVE * nenv / launch_s * 1e-9 / hbm_peak,
                            "peak_source": "MEASURED_PEAKS.json" if "hbm_gbs" in peaks else "fallback"},
                    "ref_algorithm_equiv_tflops": F.reference_algorithm_flops(50) * nenv / launch_s * 1e-12}
        line = {"metric": METRIC, "value": spl * nenv * world * args.steps / dev_s, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": 1e3 * dev_s / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64" if args.converged else "f32", "data": "synthetic",
                "config": dict(workload_config(world, nenv, spl), **({} if args.workload == "mc" else {
                    "workload": "BASELINE configs[3] shape: %d-env cascade per GPU, UprightMPC2 -> %s -> actualT0 feedback every "
                                "env-step (robobee_test_controllers.py:162-171), WaypointHover constants" % (
                                    nenv, "WLCon" if args.workload == "cascade-wlcon" else "true WLQP (40 iterations)")})),
                "plant_steps_per_launch": spl, "qp_solves_per_s": (spl if args.hl_interval is None else 1) * nenv * world * args.steps / dev_s, "wall_ms_per_step": 1e3 * wall / args.steps,
                "gpu_launches": args.steps, "clocks": clocks, "roofline": roofline,
                "tracking": {"mean_err_mm2": sums[0] / max(sums[3], 1), "mean_eff": sums[1] / max(sums[3], 1),
                             "env_steps_total": sums[3], "non_finite_envs": sums[4]}}
        if e2e:
            line["e2e"] = e2e
        if args.converged:
            line["converged_mode"] = {"mean_iter_last_step": mean_iter, "max_iter": 4000, "check_termination": 25,
                                      "adaptive_rho_interval": 100, "eps": 1e-4,
                                      "note": "float64 kernel (umpc_conv.cu): scratch in tensor memory, loop vectors in shared "
                                              "memory (UMPC_CONV_STORE=glob: the round-1 kernel, scratch streamed from L2)"}
        if world == 1 and not args.no_cpu:
            from oracle import binding as OB
            cores = os.cpu_count() or 1
            ns = min(nenv, cores * 64)
            steps_cpu = 800                      # ~2.2 s per worker at ~27k env-steps/s per core (>= 2 s per worker)
            done, tmax, _, kind, per_core = OB.cpu_env_steps(_cpu_ics(ns), steps_cpu, cores, per_worker=True)
            line["cpu_baseline"] = {"value": done / tmax, "unit": UNIT, "cores": cores, "kind": kind,
                                    "per_core": {"min": min(per_core), "median": sorted(per_core)[len(per_core) // 2], "max": max(per_core)},
                                    "seconds_per_worker": tmax,
                                    "sample": "%d of the %d envs x %d plant steps, one process per host core" % (ns, nenv, steps_cpu)}
        print(json.dumps(line))
    b.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
