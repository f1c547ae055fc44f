"""oracle/flapping_oracle.py -- TEST INFRASTRUCTURE ONLY (never imported by robobee3d_b200/): CPU restatement of the
flapping-wing plant row, the open-loop sweep that produces the wrench-map training data.

Reference path:
    aerodynamics, wrench, actuatorModel      template/robobee.py:20-55        (restated below, PINNED: tests/golden/flapping_aero.npz)
    WingFlapFilter                           template/robobee.py:57-73        (restated, pinned)
    RobobeeSim.update / openLoop             template/robobee.py:213-300      (torques, aero, waveform restated)
    sweepFile / olAvgWrenchKinFeat           template/sdab_num_wrenchmap_simgen.py:19-55
    the robot                                urdf/sdab.xacro (masses, inertias, joint axes, khinge 3, bhinge 2, kstroke 65)

PARITY UNPINNED for the multibody step: the reference integrates the wings with pybullet (p.stepSimulation, fixed base:
sdab_num_wrenchmap_simgen.py:16), which is absent here and on the GPU box.  The step below is this repo's own statement of
the same physics -- each wing a two-joint chain (stroke about z, hinge about the spar) turning about one fixed point,
explicit joint torques, semi-implicit Euler like Bullet's multibody (v += a dt, q += v dt) -- written in WORLD-frame
Newton-Euler form, independently of the link-frame form the CUDA kernel uses (robobee3d_b200/csrc/flapping.cu); tests
check the two against each other and against energy conservation.  Differences from the reference that are known:
a sweep's sims each start from rest (the reference's `reset` does not reset pybullet's joints, so sim k starts where
k - 1 ended; only the last period is used either way).
"""
import numpy as np

# ---- constants of template/robobee.py:7-18 and urdf/sdab.xacro ------------------------------------------------------
CD0, CDMAX, CLMAX, RHO, RCOPNONDIM = 0.4, 3.4, 1.8, 1.225e-3, 0.5
WPARAMS = dict(Aw=54.4, r1h=0.49, r2h=0.551, cbar=3.2)
URDF = dict(d=4.0, Roffs=3.0, ycp=7.0, khinge=3.0, bhinge=2.0, kstroke=65.0,
            m_spar=0.72, I_spar=(0.03, 11.775, 11.775), c_spar=(0.0, 7.0, 0.0),
            m_mem=0.06, I_mem=(0.08, 1.06, 0.981), c_mem=(0.0, 7.0, -2.0), g=-9.81e-3, act_T=2.6666)


def aerodynamics(theta, dtheta, lrSign, params=URDF):
    """robobee.py:20-47, vectorised over the leading axis; theta, dtheta (B, 2) as the CALLER passes them (not mutated)"""
    theta, dtheta = np.array(theta, float), np.array(dtheta, float)
    th0, dth0 = theta[..., 0] * lrSign, dtheta[..., 0] * lrSign
    th1 = theta[..., 1]
    cbar, Aw, r2h = WPARAMS["cbar"], WPARAMS["Aw"], WPARAMS["r2h"]
    AR = Aw / cbar ** 2
    ycp = params["ycp"]
    s, c, sh, ch = np.sin(th0), np.cos(th0), np.sin(th1), np.cos(th1)
    z = np.zeros_like(s)
    pcop = np.stack([z, z + params["Roffs"], z + params["d"]], -1) + np.stack(
        [-ycp * s - cbar * RCOPNONDIM * c * sh, ycp * c - cbar * RCOPNONDIM * s * sh, -cbar * RCOPNONDIM * ch], -1)
    aoa = 0.5 * np.pi - th1
    eD = np.stack([dth0 * c, dth0 * s, z], -1) / (np.abs(dth0) + 1e-4)[..., None]
    eL = np.stack([z, z, z + 1.0], -1)
    Cd = (CDMAX + CD0) / 2 - (CDMAX - CD0) / 2 * np.cos(2 * aoa)
    Cl = CLMAX * np.sin(2 * aoa)
    F = (0.5 * RHO * dth0 ** 2 * Aw ** 2 * AR * r2h ** 2)[..., None] * (Cd[..., None] * eD + (Cl * np.sign(-dth0))[..., None] * eL)
    if lrSign < 0:
        F = F * np.array([1.0, -1.0, 1.0]); pcop = pcop * np.array([1.0, -1.0, 1.0])
    return F, pcop


def wrench(F, p):
    """robobee.py:49-50"""
    return np.concatenate([F, np.cross(p, F)], -1)


def actuator_model(V):
    """robobee.py:52-55"""
    return 1.0 / URDF["act_T"] * (50.0 / 180.0 * V)


class WingFlapFilter:
    """robobee.py:57-73: moving average over one flapping period"""

    def __init__(self, freq, timestep, n):
        self.buf = np.zeros((int(1 / (freq * timestep)), n)); self.bufi = 0; self.first = True

    def update(self, x):
        if self.first:
            self.first = False
            self.buf[:] = x
        else:
            self.buf[self.bufi] = x
        self.bufi = (self.bufi + 1) % len(self.buf)
        return self.buf.mean(0)


# ---- the wing as a two-joint chain about the fixed point O = (0, s Roffs, d) ----------------------------------------
def _rot(axis, ang):
    """Rodrigues, batched: axis (3,), ang (B,) -> (B, 3, 3)"""
    a = np.asarray(axis, float)
    K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    s, c = np.sin(ang)[:, None, None], np.cos(ang)[:, None, None]
    return np.eye(3) + s * K + (1 - c) * (K @ K)


def _inertia_about_O(I_diag_inertial, m, c):
    """link-frame inertia about the joint point: the URDF's inertial frame is the link frame turned by +-90 deg about z
    (urdf/sdab.xacro:77,93), which swaps ixx and iyy; then the parallel-axis term"""
    ixx, iyy, izz = I_diag_inertial
    c = np.asarray(c, float)
    return np.diag([iyy, ixx, izz]) + m * (c @ c * np.eye(3) - np.outer(c, c))


def wing_accel(sgn, q, dq, tau, F, pcop, with_gravity=True):
    """q, dq, tau (B, 2) of one wing (stroke, hinge), aero force F at pcop (B, 3), body frame -> ddq (B, 2).
    World-frame Newton-Euler about O: N_k = I_k alpha_k + w_k x I_k w_k, tau_i = axis_i . sum_{k >= i} N_k, evaluated for
    ddq = 0, e_1, e_2 to read off the bias and the mass matrix."""
    B = len(q)
    a1, a2 = np.array([0, 0, float(sgn)]), np.array([0, float(sgn), 0])
    O = np.array([0.0, sgn * URDF["Roffs"], URDF["d"]])
    c1, c2 = np.array(URDF["c_spar"]) * [1, sgn, 1], np.array(URDF["c_mem"]) * [1, sgn, 1]
    J1, J2 = _inertia_about_O(URDF["I_spar"], URDF["m_spar"], c1), _inertia_about_O(URDF["I_mem"], URDF["m_mem"], c2)
    R1 = _rot(a1, q[:, 0]); R2 = R1 @ _rot(a2, q[:, 1])
    I1 = R1 @ J1 @ R1.transpose(0, 2, 1); I2 = R2 @ J2 @ R2.transpose(0, 2, 1)
    ax1 = np.tile(a1, (B, 1)); ax2 = R1 @ a2
    w1 = ax1 * dq[:, :1]; w2 = w1 + ax2 * dq[:, 1:2]

    def invdyn(ddq):
        al1 = ax1 * ddq[:, :1]
        al2 = al1 + ax2 * ddq[:, 1:2] + np.cross(w1, ax2) * dq[:, 1:2]
        N1 = np.einsum("bij,bj->bi", I1, al1) + np.cross(w1, np.einsum("bij,bj->bi", I1, w1))
        N2 = np.einsum("bij,bj->bi", I2, al2) + np.cross(w2, np.einsum("bij,bj->bi", I2, w2))
        return np.stack([np.einsum("bi,bi->b", ax1, N1 + N2), np.einsum("bi,bi->b", ax2, N2)], -1)
    z = np.zeros((B, 2))
    h = invdyn(z)
    M = np.stack([invdyn(np.tile([1.0, 0.0], (B, 1))) - h, invdyn(np.tile([0.0, 1.0], (B, 1))) - h], -1)   # columns
    g = np.array([0, 0, URDF["g"] if with_gravity else 0.0])
    T1 = np.cross(np.einsum("bij,j->bi", R1, c1), URDF["m_spar"] * g)
    T2 = np.cross(np.einsum("bij,j->bi", R2, c2), URDF["m_mem"] * g) + np.cross(pcop - O, F)
    Q = tau + np.stack([np.einsum("bi,bi->b", ax1, T1 + T2), np.einsum("bi,bi->b", ax2, T2)], -1)
    return np.linalg.solve(M, (Q - h)[..., None])[..., 0], M


def kinetic_energy(sgn, q, dq):
    M = wing_accel(sgn, q, dq, np.zeros_like(q), np.zeros((len(q), 3)), np.zeros((len(q), 3)))[1]
    return 0.5 * np.einsum("bi,bij,bj->b", dq, M, dq)


def step(q, dq, V, dt, aero=True, with_gravity=True, passive=True):
    """RobobeeSim.update (robobee.py:213-243) for B sims at once: q, dq (B, 4) = (l stroke, l hinge, r stroke, r hinge),
    V (B, 2) the two drive voltages.  Returns the new (q, dq) and the wrench sum of both wings at the OLD state."""
    qn, dqn = q.copy(), dq.copy()
    wsum = 0.0
    for w, sgn in ((0, 1), (1, -1)):
        qs, dqs = q[:, 2 * w:2 * w + 2], dq[:, 2 * w:2 * w + 2]
        F, pc = aerodynamics(qs if sgn > 0 else -qs, dqs if sgn > 0 else -dqs, sgn)       # :229-230
        wsum = wsum + wrench(F, pc)
        if not aero:
            F = F * 0
        tau = np.zeros_like(qs)
        tau[:, 0] = actuator_model(V[:, w]) - URDF["kstroke"] * qs[:, 0]                   # :219-222
        if passive:
            tau[:, 1] = -URDF["khinge"] * qs[:, 1] - URDF["bhinge"] * dqs[:, 1]              # :141 (PD_CONTROL towards 0)
        ddq = wing_accel(sgn, qs, dqs, tau, F, pc, with_gravity)[0]
        dqn[:, 2 * w:2 * w + 2] = dqs + dt * ddq
        qn[:, 2 * w:2 * w + 2] = qs + dt * dqn[:, 2 * w:2 * w + 2]
    return qn, dqn, wsum


def open_loop(VampL, VampR, uoffs, f, tendMS=100.0, h2=0.0, h3=0.0, dt=0.1):
    """RobobeeSim.openLoop (robobee.py:277-300) for B parameter sets at once -> qw (B, nsteps, 14)"""
    VampL, VampR, uoffs, f, h2 = (np.atleast_1d(np.asarray(a, float)) for a in (VampL, VampR, uoffs, f, h2))
    B = len(VampL)
    q, dq = np.zeros((B, 4)), np.zeros((B, 4))
    omega = 2 * np.pi * f
    sig = lambda ph, hh: (1 + h3) * np.sin(ph) + h3 * np.sin(3 * ph) + hh * np.sin(2 * ph) + uoffs
    simt, rows = 0.0, []
    while simt < tendMS:
        ph = omega * simt
        V = np.stack([VampL * sig(ph, h2), VampR * sig(ph, -h2)], -1)
        qn, dqn, wsum = step(q, dq, V, dt)
        rows.append(np.concatenate([q, dq, wsum], -1))
        q, dq = qn, dqn
        simt += dt
    return np.stack(rows, 1)


def sweep_features(inputs, tendMS=100.0, dt=0.1):
    """olAvgWrenchKinFeat (sdab_num_wrenchmap_simgen.py:31-47): inputs (B, 5) = Vmean, uoffs, f, udiff, h2 -> (B, 20) rows
    of numkins.npy: the inputs, the average wrench and the 9 kinematics features of the last period"""
    inputs = np.asarray(inputs, float)
    Vmean, uoffs, f, udiff, h2 = inputs.T
    sw = open_loop(Vmean * (1 + udiff), Vmean * (1 - udiff), uoffs, f, tendMS, h2, dt=dt)
    out = []
    for b in range(len(inputs)):
        n = int(1 / (f[b] * dt))
        s = sw[b, -n:]
        qw, dqw = s[:, :4], s[:, 4:8]
        kins = np.hstack([np.hstack((qw[:, 2 * i:2 * i + 2].max(0), -qw[:, 2 * i:2 * i + 2].min(0))) for i in range(2)])
        out.append(np.hstack((inputs[b], s[:, -6:].mean(0), kins, (dqw[:, 0] >= 0).sum() / n)))
    return np.array(out)
