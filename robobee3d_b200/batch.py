"""Batched UprightMPC2 controller + closed-loop rollout on one B200 (ctypes over the C-ABI).

One BatchUMPC holds the struct-of-arrays device state of B independent environments
(plant state, per-env controller state T0 / warm start / scaling, WL state, metrics).
`update` is the batched twin of the reference's UprightMPC2C.update
(template/uprightmpc2/py/uprightmpc2py.cpp:38-43); `rollout` is the batched twin of
controlTest (template/uprightmpc2.py:87-159).
"""
import ctypes as C
import numpy as np

from . import _lib

TASK_HELIX, TASK_STRAIGHT, TASK_FLIP, TASK_PERCH = 0, 1, 2, 3

# createMPC defaults (template_controllers.py:260-280) and controlTest defaults (uprightmpc2.py:87)
DEFAULTS = dict(dt=5.0, g=9.81e-3, TtoWmax=2.0, ws=1e1, wds=1e3, wpr=1.0, wpf=5.0, wvr=1e3, wvf=2e3,
                wthrust=1e-1, wmom=1e-2, Ib=(3333.0, 3333.0, 1000.0), maxIter=50,
                dtsim=0.2, hlInterval=None, taulim=100.0, tpert=None, mb=100.0,
                wlMode=0, wl_u0=(140.0, 0.0, 0.0, 0.0), wl_umin=(90.0, -0.5, -0.2, -0.1),
                wl_umax=(240.0, 0.5, 0.2, 0.1), wl_dumax=(5e3, 10.0, 10.0, 10.0),
                wl_Qw=(1.0, 1.0, 1.0, 0.1, 0.1, 0.1), wl_controlRate=1000.0, wl_popts=None,
                wlqpCheckTermination=25, intInitialPos=0,
                checkTermination=0, adaptiveRhoInterval=0, epsAbs=1e-4, epsRel=1e-4)


def _f32(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float32)
    if shape is not None:
        a = a.reshape(shape)
    return a


def _vp(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


def make_params(**kw):
    d = dict(DEFAULTS)
    for k in kw:
        if k not in d:
            raise TypeError("unknown parameter %r" % k)
    d.update(kw)
    p = _lib.UmpcBatchParams()
    _lib.load().umpcBatchDefaults(C.byref(p))
    for k in ("dt", "g", "TtoWmax", "ws", "wds", "wpr", "wpf", "wvr", "wvf", "wthrust", "wmom", "maxIter",
              "dtsim", "taulim", "mb", "wlMode", "wl_controlRate", "wlqpCheckTermination", "intInitialPos",
              "checkTermination", "adaptiveRhoInterval", "epsAbs", "epsRel"):
        setattr(p, k, d[k])
    p.hlInterval = -1.0 if d["hlInterval"] is None else float(d["hlInterval"])
    p.tpert = -1.0 if d["tpert"] is None else float(d["tpert"])
    for k, n in (("Ib", 3), ("wl_u0", 4), ("wl_umin", 4), ("wl_umax", 4), ("wl_dumax", 4), ("wl_Qw", 6)):
        for i in range(n):
            getattr(p, k)[i] = d[k][i]
    if d["wl_popts"] is not None:
        po = np.asarray(d["wl_popts"], dtype=np.float32).ravel()
        assert po.size == 90
        for i in range(90):
            p.wl_popts[i] = po[i]
    elif d["wlMode"]:
        raise ValueError("wlMode != 0 needs wl_popts[90]")
    return p


class BatchUMPC:
    def __init__(self, nenv, device=0, **kw):
        self.lib = _lib.load()
        self.B = int(nenv)
        self.device = int(device)
        self.params = make_params(**kw)
        h = C.c_void_p()
        _lib.check(self.lib.umpcBatchCreate(C.byref(h), self.B, self.device, C.byref(self.params)), "umpcBatchCreate")
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.lib.umpcBatchDestroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reset(self):
        _lib.check(self.lib.umpcBatchReset(self.h), "umpcBatchReset")

    WEIGHT_NAMES = ("ws", "wds", "wpr", "wpf", "wvr", "wvf", "wthrust", "wmom")

    def set_weights(self, **kw):
        """Per-environment cost weights: every env becomes its own createMPC(**kwargs) (template_controllers.py:260).
        Each keyword is a scalar or an array of B values; weights not given keep the batch-wide value.
        set_weights() with no arguments restores the batch-wide parameters."""
        if not kw:
            _lib.check(self.lib.umpcBatchSetWeights(self.h, None), "umpcBatchSetWeights")
            return
        bad = [k for k in kw if k not in self.WEIGHT_NAMES]
        if bad:
            raise TypeError("unknown weight(s) %s" % bad)
        w = np.empty((self.B, 8), np.float32)
        for i, k in enumerate(self.WEIGHT_NAMES):
            w[:, i] = np.broadcast_to(np.asarray(kw.get(k, getattr(self.params, k)), np.float32), (self.B,))
        _lib.check(self.lib.umpcBatchSetWeights(self.h, _vp(w)), "umpcBatchSetWeights")

    def set_rollout_info(self, on=True):
        """rollouts also evaluate residuals / status / solution rows after every solve (off by default, see the header)"""
        _lib.check(self.lib.umpcBatchSetRolloutInfo(self.h, int(bool(on))), "umpcBatchSetRolloutInfo")

    def set_launch(self, threads_per_block):
        _lib.check(self.lib.umpcBatchSetLaunch(self.h, int(threads_per_block)), "umpcBatchSetLaunch")

    # ---- one controller update per environment -------------------------------------------------
    def update(self, p0, R0, dq0, pdes, dpdes, sdes, actualT0=None):
        """R0: (B,3,3) rotation matrices (numpy row-major, like the pybind module takes one (3,3));
        returns (uquad (B,3), accdes (B,6))."""
        B = self.B
        Rc = _f32(np.asarray(R0, dtype=np.float32).reshape(B, 3, 3).transpose(0, 2, 1), (B, 9))
        return self.update_raw(_f32(p0, (B, 3)), Rc, _f32(dq0, (B, 6)), _f32(pdes, (B, 3)), _f32(dpdes, (B, 3)),
                               _f32(sdes, (B, 3)), None if actualT0 is None else _f32(actualT0, (B,)))

    def update_raw(self, p0, R0c, dq0, pdes, dpdes, sdes, actualT0=None, uquad=None, accdes=None):
        """Same, with float32 C-contiguous arrays already in the C-ABI layout (R0c column-major per env).
        Any object exposing .ctypes.data or torch's .data_ptr() on HOST memory works."""
        uquad = np.empty((self.B, 3), np.float32) if uquad is None else uquad
        accdes = np.empty((self.B, 6), np.float32) if accdes is None else accdes
        ptr = lambda a: None if a is None else C.c_void_p(a.data_ptr() if hasattr(a, "data_ptr") else a.ctypes.data)
        _lib.check(self.lib.umpcBatchUpdate(self.h, ptr(uquad), ptr(accdes), ptr(p0), ptr(R0c), ptr(dq0), ptr(pdes),
                                            ptr(dpdes), ptr(sdes), ptr(actualT0)), "umpcBatchUpdate")
        return uquad, accdes

    def env_step_raw(self, uquad, accdes, p, Rc, dq, pdes, dpdes, sdes, actualT0=None):
        """One closed-loop env-step with HOST buffers in the C-ABI layout (numpy or pinned torch tensors):
        controller update, torque clip, rigid-body step; p, Rc, dq are overwritten with the next state."""
        ptr = lambda a: None if a is None else C.c_void_p(a.data_ptr() if hasattr(a, "data_ptr") else a.ctypes.data)
        _lib.check(self.lib.umpcBatchEnvStep(self.h, ptr(uquad), ptr(accdes), ptr(p), ptr(Rc), ptr(dq), ptr(p), ptr(Rc), ptr(dq),
                                             ptr(pdes), ptr(dpdes), ptr(sdes), ptr(actualT0)), "umpcBatchEnvStep")

    def update_dev(self, uquad, accdes, p0, R0c, dq0, pdes, dpdes, sdes, actualT0=None, stream="current"):
        """torch CUDA tensors (float32, contiguous, [B,k] -- exactly B rows, no padding needed).  Ordering contract:
        the launch goes to the batch's private stream, hand-shaken with `stream` by events (no host sync): it starts after
        everything already queued on `stream`, and work queued on `stream` afterwards sees its outputs.
        stream="current" (default): torch.cuda.current_stream() of the batch's device; an int / cudaStream_t handle:
        that stream; None: no hand-shake (the caller synchronises, e.g. with sync())."""
        ptr = lambda a: None if a is None else C.c_void_p(a.data_ptr())
        if isinstance(stream, str) and stream == "current":
            import torch
            stream = torch.cuda.current_stream(self.device).cuda_stream
        if stream is None:
            _lib.check(self.lib.umpcBatchUpdateDev(self.h, ptr(uquad), ptr(accdes), ptr(p0), ptr(R0c), ptr(dq0), ptr(pdes),
                                                   ptr(dpdes), ptr(sdes), ptr(actualT0)), "umpcBatchUpdateDev")
        else:
            _lib.check(self.lib.umpcBatchUpdateDevOn(self.h, C.c_void_p(int(stream)), ptr(uquad), ptr(accdes), ptr(p0), ptr(R0c),
                                                     ptr(dq0), ptr(pdes), ptr(dpdes), ptr(sdes), ptr(actualT0)), "umpcBatchUpdateDevOn")

    def solve_info(self, full=False):
        B = self.B
        info = np.empty((B, 4), np.float32)
        if not full:
            _lib.check(self.lib.umpcBatchGetSolve(self.h, _vp(info), None, None, None, None, None, None, None))
            return dict(iter=info[:, 0].astype(int), status=info[:, 1].astype(int), pri_res=info[:, 2], dua_res=info[:, 3])
        x, y = np.empty((B, 45), np.float32), np.empty((B, 39), np.float32)
        xs, zs, ys = np.empty((B, 45), np.float32), np.empty((B, 39), np.float32), np.empty((B, 39), np.float32)
        D, E = np.empty((B, 45), np.float32), np.empty((B, 39), np.float32)
        _lib.check(self.lib.umpcBatchGetSolve(self.h, _vp(info), _vp(x), _vp(y), _vp(xs), _vp(zs), _vp(ys), _vp(D), _vp(E)))
        return dict(iter=info[:, 0].astype(int), status=info[:, 1].astype(int), pri_res=info[:, 2], dua_res=info[:, 3],
                    x=x, y=y, xs=xs, zs=zs, ys=ys, D=D, E=E)

    def qpdata(self):
        """(l, u, q, Px, Ax_data) of the LAST update, each (B, k). The first call only switches the
        output on (returns zeros); data is produced by updates after it."""
        qp = np.empty((self.B, 216), np.float32)
        _lib.check(self.lib.umpcBatchGetQPData(self.h, _vp(qp)))
        return qp[:, :39], qp[:, 39:78], qp[:, 78:123], qp[:, 123:168], qp[:, 168:216]

    # ---- closed loop ---------------------------------------------------------------------------
    def set_state(self, p, R, dq, task=None, tp=None):
        B = self.B
        Rc = _f32(np.asarray(R, dtype=np.float32).reshape(B, 3, 3).transpose(0, 2, 1), (B, 9))
        t = None if task is None else np.ascontiguousarray(np.broadcast_to(np.asarray(task, np.int32), (B,)))
        tpa = None if tp is None else _f32(np.broadcast_to(np.asarray(tp, np.float32), (B, 4)))
        pa, dqa = _f32(p, (B, 3)), _f32(dq, (B, 6))  # keep the temporaries alive across the call
        _lib.check(self.lib.umpcBatchSetState(self.h, _vp(pa), _vp(Rc), _vp(dqa), _vp(t), _vp(tpa)), "umpcBatchSetState")

    def state(self):
        B = self.B
        p, R, dq, uq = (np.empty((B, 3), np.float32), np.empty((B, 9), np.float32), np.empty((B, 6), np.float32),
                        np.empty((B, 3), np.float32))
        _lib.check(self.lib.umpcBatchGetState(self.h, _vp(p), _vp(R), _vp(dq), _vp(uq)))
        return p, R.reshape(B, 3, 3).transpose(0, 2, 1).copy(), dq, uq

    def rollout(self, nsteps, sync=True):
        _lib.check(self.lib.umpcBatchRollout(self.h, int(nsteps)), "umpcBatchRollout")
        if sync:
            self.sync()

    def sync(self):
        _lib.check(self.lib.umpcBatchSync(self.h), "umpcBatchSync")

    def timer_start(self):
        _lib.check(self.lib.umpcBatchTimerStart(self.h), "umpcBatchTimerStart")

    def timer_stop_ms(self):
        return float(self.lib.umpcBatchTimerStopMs(self.h))

    def last_kernel_ms(self):
        return float(self.lib.umpcBatchLastKernelMs(self.h))

    def metrics(self):
        """per env: err = mean |p|^2, eff = mean |tau|^2 (logMetric, uprightmpc2.py:161-175), nsolve, nstep"""
        m = np.empty((self.B, 4), np.float32)
        _lib.check(self.lib.umpcBatchGetMetrics(self.h, _vp(m)))
        n = np.maximum(m[:, 3], 1.0)
        return dict(err=m[:, 0] / n, eff=m[:, 1] / n, nsolve=m[:, 2].astype(np.int64), nstep=m[:, 3].astype(np.int64))

    def reduce_metrics(self):
        """on-device sums over this shard: [sum err, sum eff, sum solves, sum steps, n_bad, max mean-err, 0, 0]"""
        s = np.zeros(8, np.float64)
        _lib.check(self.lib.umpcBatchReduceMetrics(self.h, _vp(s)))
        return s

    def set_log(self, env, capacity):
        _lib.check(self.lib.umpcBatchSetLog(self.h, int(env), int(capacity)))

    def get_log(self, nsteps):
        y, u = np.empty((nsteps, 12), np.float32), np.empty((nsteps, 3), np.float32)
        _lib.check(self.lib.umpcBatchGetLog(self.h, _vp(y), _vp(u), int(nsteps)))
        return y, u

    def dev_ptr(self, name):
        return self.lib.umpcBatchDevPtr(self.h, name.encode())


def gather_metrics(local_sums, group=None):
    """Aggregate tracking-error statistics over ranks (the ONLY collective of the path: envs are
    independent).  Uses torch.distributed (NCCL on GPUs, gloo in CPU tests) when it is
    initialised, else returns the local sums.  Max entry (index 5) is reduced with MAX."""
    import torch
    import torch.distributed as dist
    s = np.asarray(local_sums, np.float64)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return s.copy()
    dev = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
    t = torch.tensor(s, dtype=torch.float64, device=dev)
    mx = t[5:6].clone()
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(mx, op=dist.ReduceOp.MAX, group=group)
    t[5] = mx[0]
    return t.cpu().numpy()


def shard_range(nenv_total, rank, world):
    """contiguous env-id range of a rank"""
    per = (nenv_total + world - 1) // world
    lo = min(rank * per, nenv_total)
    return lo, min(lo + per, nenv_total)
