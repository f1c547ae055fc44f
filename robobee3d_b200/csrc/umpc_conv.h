// robobee3d_b200/csrc/umpc_conv.h -- internal interface between the two translation units of libumpc_b200.so.
// umpc_conv.cu holds the CONVERGED-MODE controller: the reference's Python path (template_controllers.py:190-258,
// osqp on float64 with eps 1e-4, termination test every 25 iterations, adaptive rho) as a float64 CUDA kernel.
// Not part of the C-ABI (include/umpc_b200.h); selected per batch by UmpcBatchParams.checkTermination > 0.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "umpc_dev.cuh"
#include "../../include/umpc_b200.h"

struct UmpcConv;  // double-precision controller parameters + the double slab DevState::c64 points into

// Allocates the double slab for Bp env slots, fills the parameters from bp, sets s->c64.  Returns a cudaError_t.
int umpc_conv_create(UmpcConv** out, int Bp, const UmpcBatchParams* bp, DevState* s);
// warm start = 0, D = E = 1, rho = 0.1, T0 = 0 (a fresh OSQP workspace)
int umpc_conv_reset(UmpcConv* c, const DevState* s, cudaStream_t stream);
void umpc_conv_destroy(UmpcConv* c);
// same arguments as the float kernels' launches (umpc_kernels.cu: launch_update / umpcBatchRollout)
int umpc_conv_launch_update(UmpcConv* c, const SimParams* sim, const DevState* s, float* uquad, float* accdes, const float* p0,
                            const float* R0, const float* dq0, const float* pdes, const float* dpdes, const float* sdes,
                            const float* aT0, int want_qpdata, float* pn, float* Rn, float* dqn, int env0, int nenv,
                            cudaStream_t stream);
int umpc_conv_launch_rollout(UmpcConv* c, const SimParams* sim, const DevState* s, const WlParams* wlp, int step0, int nsteps,
                             const uint8_t* fire, const double* tvals, LogBuf lg, cudaStream_t stream);
