/*
 * C launchers of the uv-grid weighting kernels (es_weighting.cu): plain device
 * pointers and sizes; work is enqueued on `stream`; the return value is a
 * cudaError_t as int. num_rows = num_times * num_baselines.
 */
#ifndef ES_WEIGHTING_H_
#define ES_WEIGHTING_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* grid[cell] += input; output = 1 / grid[cell] (sdp_weighting.cpp:20-75,150-211) */
int es_weight_uniform(int64_t num_rows, int num_chan, int num_pols,
        int64_t grid_size, const double* uvw, const double* freq_hz,
        double max_abs_uv, void* grid, const void* input, void* output,
        int is_double, void* stream);

/* grid[cell] += input; sums = (sum grid[cell], sum grid[cell]^2) over the
 * visibilities (device doubles[2], zeroed here); output = input /
 * (1 + numerator / (sums[1] / sums[0]) * grid[cell]) (:78-144,214-279) */
int es_weight_briggs(int64_t num_rows, int num_chan, int num_pols,
        int64_t grid_size, const double* uvw, const double* freq_hz,
        double max_abs_uv, double numerator, void* grid, const void* input,
        void* output, double* sums, int is_double, void* stream);

/* Cells (grid_size^2 x num_pols) of the file-scope device array the
 * allocation-free variant keeps its weight grid in (8 bytes each: 8 MiB). */
#define ES_WT_STATIC_CELLS (1024 * 1024)
int64_t es_weight_static_max_cells(void);

/* es_weight_briggs with the grid and the sums in that static array (cleared
 * per call; shared by all calls: overlapping calls race). No cudaMalloc /
 * cudaFree; cudaErrorInvalidValue when grid_size^2 x num_pols exceeds
 * ES_WT_STATIC_CELLS. */
int es_weight_briggs_static(int64_t num_rows, int num_chan, int num_pols,
        int64_t grid_size, const double* uvw, const double* freq_hz,
        double max_abs_uv, double numerator, const void* input, void* output,
        int is_double, void* stream);

#ifdef __cplusplus
}
#endif

#endif
