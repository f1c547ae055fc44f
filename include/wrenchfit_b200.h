/*
 * include/wrenchfit_b200.h -- C-ABI of the wrench-map fitting of libumpc_b200.so
 *
 * Produces the popts[90] that wlConInit / wlControllerUpdate / the fused WLQP consume (include/umpc_b200.h), the way
 * template/sdab_num_wrenchmap.py does: stroke-averaged wrenches from kinematics features (wrenchFromKinematics, :165-214),
 * then six 15-parameter quadratic models w_i(u) = a0 + a1'u + 1/2 u'A2 u fitted to them (FunApprox.f :144-152; curve_fit
 * from p0 = ones, :262-266).  Host arrays in and out, all arithmetic on the device (no CPU path), float64.
 */
#ifndef WRENCHFIT_B200_H
#define WRENCHFIT_B200_H
#ifdef __cplusplus
extern "C" {
#endif

/* kins[M][9] = left (q1max, q2max, -q1min, -q2min), right (same), alpha; freq[M] in kHz;
 * par[17] = Aw, AR, r2h, ycp, R, RHO, CD0, CDmax, CLmax, kaerox, strokex, wbias[6]  ->  ws[M][6] */
int wfitWrenchFromKinematics(int M, const double *kins, const double *freq, const double par[17], double *ws);
/* F independent datasets of M trials each: xdata[F][M][4] = (Vmean, uoffs, udiff, h2), ws[F][M][6] -> popts[F][90]
 * ("popts row major", sdab_num_wrenchmap.py:267-268), optionally the fitted values fit[F][M][6], the numerical rank of
 * each design matrix and the kernel time.  Rank-deficient designs (the reference's own data has rank 10 of 15) get the
 * least-squares solution closest to p0 = ones, which is what Levenberg-Marquardt from p0 converges to. */
int wfitFit(int F, int M, const double *xdata, const double *ws, double *popts, double *fit, int *rank, float *kernel_ms);

#ifdef __cplusplus
}
#endif
#endif
