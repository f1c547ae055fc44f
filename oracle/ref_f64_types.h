/* oracle/ref_f64_types.h -- TEST INFRASTRUCTURE.
 * Force-included when building the float64 ("converged", SURVEY 8c flavour B)
 * variant of the reference's vendored OSQP: stands in for qdldl_types.h with
 * double values; DFLOAT stays undefined so glob_opts.h:81-85 picks double. */
#pragma once
typedef int    QDLDL_int;
typedef double QDLDL_float;
typedef int    QDLDL_bool;
