"""Shared test helpers: seeded input generators of the BASELINE configs and tolerances."""
import json
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")

# Faithful mode (50 fixed-rho float32 ADMM iterations, like the C path): our kernel re-associates the
# arithmetic (reduced KKT form, un-scaled coordinates), so agreement with the reference C code is at
# float32 round-off amplified by the QP's conditioning, NOT bit-exact.  Noise floor, measured on the 48x3
# golden QPs (tests/golden/umpc_seq.npz): the reference C code (float32) against the SAME algorithm in
# float64 (oracle f64) differs by up to 1.2e-3 of |x|_inf (median 3e-5) -- i.e. the reference's own
# float32 rounding -- and our kernel's arithmetic gives up to 2.3e-3 (median 6e-5) on the same worst env.
TOL_X_REL = 2.5e-3    # |x - x_ref|_inf <= TOL_X_REL * max(1, |x_ref|_inf)   per QP: 2x the reference's own f32-vs-f64 gap (1.2e-3)
TOL_Y_REL = 2.5e-3    # same for the duals
TOL_QPDATA = 2e-6     # QP data l,u,q,Px,Ax: relative (float32 rounding)
TOL_TRAJ_P = 0.5      # mm, closed-loop position over the 500 ms horizon
TOL_TRAJ_S = 1e-2     # body z-axis components


def rot_from(ax, ang, yaw):
    k = np.array([np.cos(ax), np.sin(ax), 0.0])
    K = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    Rt = np.eye(3) + np.sin(ang) * K + (1 - np.cos(ang)) * K @ K
    cz, sz = np.cos(yaw), np.sin(yaw)
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return Rt @ Rz


def config2_states(B, seed=0):
    """random upright states. Returns p (B,3), R (B,3,3), dq (B,6)."""
    rng = np.random.default_rng(seed)
    ax = rng.uniform(0, 2 * np.pi, B)
    ang = rng.uniform(0, 0.6, B)
    yaw = rng.uniform(-np.pi, np.pi, B)
    R = np.stack([rot_from(ax[i], ang[i], yaw[i]) for i in range(B)])
    p = rng.uniform(-50, 50, (B, 3))
    dq = np.concatenate([rng.uniform(-0.2, 0.2, (B, 3)), rng.uniform(-0.01, 0.01, (B, 3))], axis=1)
    return p, R, dq


def config3_ics(B, seed=0):
    """controlTest IC (uprightmpc2.py:94-103) perturbed per env."""
    from oracle.binding import control_test_ic
    rng = np.random.default_rng(seed)
    p0, R0, dq0 = control_test_ic()
    p = p0[None] + rng.normal(0, 5.0, (B, 3))
    dq = np.tile(dq0, (B, 1))
    dq[:, :3] += rng.normal(0, 0.05, (B, 3))
    R = np.empty((B, 3, 3))
    for i in range(B):
        w = rng.normal(0, 0.1, 3)
        th = np.linalg.norm(w)
        K = np.array([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]])
        Ex = np.eye(3) + (np.sin(th) / th) * K + ((1 - np.cos(th)) / th ** 2) * K @ K if th > 1e-12 else np.eye(3)
        R[i] = R0 @ Ex
    return p, R, dq


def load_constants():
    with open(os.path.join(GOLDEN, "constants.json")) as f:
        return json.load(f)


def relerr(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    return np.abs(a - b).max() / max(1.0, np.abs(b).max())


def has_cuda():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False
