/*
 * Host-side planning maths of the ES gridder (internal header).
 * Restates, value for value, the reference's
 *   src/ska-sdp-func/grid_data/sdp_gridder_uvw_es_fft.cpp:305-424   (plan scalars)
 *   src/ska-sdp-func/grid_data/sdp_gridder_uvw_es_fft_utils.cpp:13-218,488-536
 */
#ifndef ES_PLAN_H_
#define ES_PLAN_H_

#include "ska-sdp-func/grid_data/sdp_gridder_uvw_es_fft_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

#define ES_QUADRATURE_BOUND 32 /* es_fft_utils.h:13 */

/* Smallest 11-smooth integer >= n (n <= 12 returned as is). */
int es_good_size_complex(int n);

/* KernelDB selection: smallest grid that meets epsilon; ties -> larger support. */
void es_select_kernel(double epsilon, int image_size, int is_double,
        int* grid_size, int* support, double* beta_over_support);

/* Positive nodes / weights of the 2p-point Gauss-Legendre rule, ES kernel at
 * the nodes, and the normalised 1-D correction table [image_size/2 + 1].
 * Arrays qk/qn/qw hold ES_QUADRATURE_BOUND doubles (zero padded).
 * Returns C(0) = support * sum(qk * qw). */
double es_correction_tables(int image_size, int grid_size, int support,
        double beta, double* qk, double* qn, double* qw, double* conv_corr);

/* Everything in sdp_EsPlanInfo from the user-facing arguments. */
void es_plan_scalars(int num_rows, int num_chan, int image_size, int is_double,
        double pixel_size, double epsilon, double min_abs_w, double max_abs_w,
        int do_wstacking, sdp_EsPlanInfo* info);

#ifdef __cplusplus
}
#endif

#endif
