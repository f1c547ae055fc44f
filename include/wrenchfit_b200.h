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

/* wfitFitNoAlloc has NO counterpart in the reference: it was added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The same fit as
 * wfitFit (same kernel arithmetic, rcond 1e-10), but it allocates nothing -- no cudaMalloc / cudaFree / host staging: the
 * working set (design matrix, V, sigma, coefficients) lives in one file-scope device array of the library, which every
 * GPU that loads libumpc_b200.so holds until exit (2.05 MiB).  LIMITS: F <= WFIT_NOALLOC_MAX_DATASETS and
 * F * M <= WFIT_NOALLOC_MAX_TRIALS per call; beyond them the call returns UMPC_ERR_ARG with a message (umpcLastError).
 * d_xdata[F][M][4], d_ws[F][M][6], d_popts[F][90], d_rank[F] (or NULL) are DEVICE pointers of the CURRENT device, owned by
 * the caller; the kernel is enqueued on `stream` (a cudaStream_t, NULL = default stream) and the call returns without
 * synchronising.  The first call on each device looks the array's device address up once (cudaGetSymbolAddress): make
 * that call before starting a graph capture.  Every call on a device SHARES the array, which keeps the last call's
 * contents: the kernel rewrites every word it reads, but two calls that overlap on the GPU (different streams, no
 * dependency between them) RACE on it -- serialise them. */
#define WFIT_NOALLOC_MAX_DATASETS 64
#define WFIT_NOALLOC_MAX_TRIALS 16384
int wfitFitNoAlloc(int F, int M, const double *d_xdata, const double *d_ws, double *d_popts, int *d_rank, void *stream);

#ifdef __cplusplus
}
#endif
#endif
