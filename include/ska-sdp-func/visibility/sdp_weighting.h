/*
 * Visibility weighting (uniform / Briggs) on the uv grid: the step upstream of
 * the w-stacking gridder, which consumes its `weight[rows, channels]`.
 * Drop-in for the reference's
 *     src/ska-sdp-func/visibility/sdp_weighting.h:59-80
 * (same prototypes, argument meaning, error codes and check order,
 * sdp_weighting.cpp:282-336,512-568).
 *
 * Array shapes, slowest to fastest varying:
 *   uvw            [num_times, num_baselines, 3]   real, metres (fp64)
 *   freq_hz        [num_channels]                  real, Hz (fp64)
 *   weight_grid_uv [grid_size, grid_size, num_pols] real, zero-initialised by
 *                  the caller; on return the sum of the input weights per uv
 *                  cell and polarisation
 *   input_weights / output_weights
 *                  [num_times, num_baselines, num_channels, num_pols], real,
 *                  fp64 or fp32 (the grid has the same type)
 * A visibility lands in cell (iu, iv) = floor(u f / (c max_abs_uv) * G/2) + G/2
 * (sdp_weighting.cpp:44-52); visibilities outside the grid are skipped and
 * their output weight is left untouched.
 *   uniform: output = 1 / grid[cell]                      (:205-206)
 *   Briggs:  output = input / (1 + f2 grid[cell]),
 *            f2 = (5 / 10^robust)^2 / (sum grid[cell]^2 / sum grid[cell]), the
 *            sums running over the visibilities (:136-144,266-269)
 *
 * This implementation runs on the GPU only: SDP_MEM_GPU arrays are processed
 * in place on the legacy default stream; SDP_MEM_CPU arrays are staged
 * through the current device (extension: the reference has a host loop for
 * them). All arrays of a call must be in the same location.
 */
#ifndef SKA_SDP_PROC_FUNC_WEIGHTING_H_
#define SKA_SDP_PROC_FUNC_WEIGHTING_H_

#include "ska-sdp-func/utility/sdp_mem.h"

#ifdef __cplusplus
extern "C" {
#endif

void sdp_weighting_briggs(
        const sdp_Mem* uvw,
        const sdp_Mem* freq_hz,
        double max_abs_uv,
        const double robust_param,
        sdp_Mem* weight_grid_uv,
        sdp_Mem* input_weights,
        sdp_Mem* output_weights,
        sdp_Error* status
);

void sdp_weighting_uniform(
        const sdp_Mem* uvw,
        const sdp_Mem* freq_hz,
        double max_abs_uv,
        sdp_Mem* weight_grid_uv,
        sdp_Mem* input_weights,
        sdp_Mem* output_weights,
        sdp_Error* status
);

/*
 * Briggs weighting that allocates nothing. NO COUNTERPART IN THE REFERENCE:
 * added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it). Same arithmetic and output_weights as sdp_weighting_briggs,
 * but the uv weight grid is not an argument: it is kept, with the two Briggs
 * sums, in a file-scope device array of the library (8 MiB, held by every GPU
 * that loads the library until exit).
 *  - limit: grid_size * grid_size * num_pols <= 1,048,576 cells
 *    (e.g. 1024^2 x 1 or 512^2 x 4); larger sizes return
 *    SDP_ERR_INVALID_ARGUMENT;
 *  - device (SDP_MEM_GPU) arrays only (SDP_ERR_MEM_LOCATION otherwise):
 *    staging host arrays would allocate;
 *  - calls no cudaMalloc / cudaFree and does not synchronise; work goes to
 *    the legacy default stream;
 *  - the array is shared by all calls and cleared at the start of each:
 *    calls that overlap on the GPU (other streams or host threads) race.
 */
void sdp_weighting_briggs_no_alloc(
        const sdp_Mem* uvw,
        const sdp_Mem* freq_hz,
        double max_abs_uv,
        const double robust_param,
        int64_t grid_size,
        const sdp_Mem* input_weights,
        sdp_Mem* output_weights,
        sdp_Error* status
);

#ifdef __cplusplus
}
#endif

#endif
