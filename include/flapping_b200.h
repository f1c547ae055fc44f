/*
 * include/flapping_b200.h -- C-ABI of the flapping-wing plant of libumpc_b200.so
 *
 * The open-loop wing simulations that produce the wrench-map training data (template/sdab_num_wrenchmap_simgen.py:19-55:
 * `sweepFile` runs `RobobeeSim.openLoop` for every point of a (Vmean, uoffs, f, udiff, h2) grid on a FIXED base and
 * keeps the average wrench and nine kinematics features of the last flapping period), batched: one CUDA thread per wing,
 * all simulations of a sweep in one launch.  Quasi-steady aerodynamics, actuator model and drive waveform are the
 * reference's own (template/robobee.py:20-55, :277-300); the multibody step the reference leaves to pybullet is stated
 * here for the robot of urdf/sdab.xacro: each wing a two-joint chain (stroke, hinge) about one fixed point, explicit
 * joint torques (stroke: actuator force - kstroke q; hinge: -khinge q - bhinge dq, robobee.py:141,219-222), semi-implicit
 * Euler (PARITY UNPINNED for that step: pybullet is absent; checked against oracle/flapping_oracle.py and energy conservation).
 * float64, host arrays in and out, no CPU path (UMPC_ERR_CUDA without a device).
 */
#ifndef FLAPPING_B200_H
#define FLAPPING_B200_H
#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
  /* aerodynamics (robobee.py:7-18): CD0, CDmax, CLmax, RHO [mg/mm^3], rcopnondim, wparams Aw, r2h, cbar */
  double CD0, CDmax, CLmax, RHO, rcopnondim, Aw, r2h, cbar;
  /* what getInfoFromURDF reads (robobee.py:146-186): ycp = Rwing / 2, Roffs, d, khinge, bhinge, kstroke */
  double ycp, Roffs, d, khinge, bhinge, kstroke;
  /* links of urdf/sdab.xacro:27-34,77,93: mass, inertia diag (ixx iyy izz in the URDF's inertial frame), COM in the link frame (left wing) */
  double m_spar, I_spar[3], c_spar[3], m_mem, I_mem[3], c_mem[3];
  double g;      /* gravity along z [mm/ms^2] (robobee.py:98) */
  double act_T;  /* actuatorModel's transmission ratio (robobee.py:54) */
  double h3;     /* third-harmonic share of the drive waveform (openLoop's h3, default 0) */
} FwParams;

void fwDefaults(FwParams *p);

/* aerodynamics(theta, dtheta, lrSign, params) (robobee.py:20-47) for n wings: theta[n][2], dtheta[n][2] as the caller of
 * the reference passes them -> F[n][3], pcop[n][3] in the body frame */
int fwAerodynamics(const FwParams *p, int n, const double *theta, const double *dtheta, int lrSign, double *F, double *pcop);

/* number of rows RobobeeSim.openLoop records: its `while self.simt < tendMS` with simt accumulated in double */
int fwOpenLoopSteps(double tendMS, double timestep);

/* nsim open-loop simulations from rest: drive[nsim][5] = VampL, VampR, uoffs, f [kHz], h2 (openLoop's arguments) ->
 * feat[nsim][15] = average wrench of both wings over the last period (6), kinematics features (9: per wing max stroke,
 * max hinge, -min stroke, -min hinge; share of the period with positive left stroke velocity) --
 * sdab_num_wrenchmap_simgen.py:31-47.  qw (may be NULL): the rows openLoop returns for simulation log_sim,
 * [steps][14] = q[4], dq[4], wrench sum [6]. */
int fwOpenLoopSweep(const FwParams *p, int nsim, const double *drive, double tendMS, double timestep, double *feat,
                    int log_sim, double *qw, float *kernel_ms);

#ifdef __cplusplus
}
#endif
#endif
