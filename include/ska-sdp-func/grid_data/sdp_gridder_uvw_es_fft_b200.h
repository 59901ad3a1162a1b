/*
 * B200 extensions to the sdp_GridderUvwEsFft plan. Nothing here exists in the
 * reference; the drop-in ABI (sdp_gridder_uvw_es_fft.h) does not depend on it.
 *
 *  - plan introspection (so tests can compare every planning scalar and table
 *    with the reference's es_fft.cpp:301-424 / es_fft_utils.cpp without a GPU);
 *  - w-plane ownership for multi-GPU runs: every rank builds the SAME global
 *    lattice and grids / degrids only the planes it owns, so partial dirty
 *    images sum (one NCCL reduce) and partial predictions add;
 *  - per-call statistics for the bench harness (kernel launch counts, device
 *    times by stage).
 */
#ifndef SDP_B200_GRIDDER_UVW_ES_FFT_EXT_H_
#define SDP_B200_GRIDDER_UVW_ES_FFT_EXT_H_

#include "ska-sdp-func/grid_data/sdp_gridder_uvw_es_fft.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Planning scalars exactly as the reference derives them (es_fft.cpp:305-424). */
typedef struct sdp_EsPlanInfo
{
    int32_t num_rows, num_chan, image_size, grid_size, support;
    int32_t num_total_w_grids, do_wstacking, is_double;
    double epsilon, pixel_size, beta, uv_scale;
    double min_abs_w, max_abs_w, min_plane_w, max_plane_w;
    double w_scale, inv_w_scale, inv_w_range, conv_corr_norm_factor;
} sdp_EsPlanInfo;

/* Host-only planning (no device touched): fills info and, when non-NULL, the
 * tables quad_kernel/quad_nodes/quad_weights [32 each] and
 * conv_corr [image_size/2 + 1], all in double. */
void sdp_gridder_uvw_es_fft_plan_info(
        int32_t num_rows, int32_t num_chan, int32_t image_size,
        int32_t is_double, double pixel_size_rad, double epsilon,
        double min_abs_w, double max_abs_w, int32_t do_w_stacking,
        sdp_EsPlanInfo* info, double* quad_kernel, double* quad_nodes,
        double* quad_weights, double* conv_corr, sdp_Error* status);

/* The scalars of an existing plan. */
void sdp_gridder_uvw_es_fft_get_info(const sdp_GridderUvwEsFft* plan,
        sdp_EsPlanInfo* info);

/* Restrict subsequent grid / degrid calls to w-planes [begin, end) of the
 * global lattice (default: all). With apply_correction == 0 the final grid
 * correction of sdp_grid_uvw_es_fft is skipped (apply it once after the
 * image reduce with sdp_gridder_uvw_es_fft_correct_image). */
void sdp_gridder_uvw_es_fft_set_plane_range(sdp_GridderUvwEsFft* plan,
        int32_t plane_begin, int32_t plane_end, int32_t apply_correction,
        sdp_Error* status);

/* Grid correction (kernels.cu:690-769 equivalent) on a device or host image. */
void sdp_gridder_uvw_es_fft_correct_image(sdp_GridderUvwEsFft* plan,
        sdp_Mem* dirty_image, sdp_Error* status);

/* Number of visibilities (row, channel pairs) whose first w-plane is p, for
 * p in [0, num_total_w_grids): the histogram a caller needs to balance plane
 * ownership across ranks. counts is int64 [num_total_w_grids] on the host. */
void sdp_gridder_uvw_es_fft_plane_histogram(sdp_GridderUvwEsFft* plan,
        const sdp_Mem* uvw, const sdp_Mem* freq_hz, int64_t* counts,
        sdp_Error* status);

/* Statistics of the last grid / degrid call on this plan. */
typedef struct sdp_EsCallStats
{
    int64_t kernel_launches;   /* our own kernels */
    int64_t fft_execs;         /* cufftExec calls */
    int64_t planes_processed;
    int64_t num_entries;       /* (visibility, tile) records after the presort */
    int64_t h2d_bytes, d2h_bytes;
    double ms_total, ms_presort, ms_scatter_gather, ms_fft, ms_screen;
    int32_t plane_batch;       /* w-planes resident per FFT batch */
    int32_t timing_enabled;
    int32_t pruned_fft;        /* 1: in-shared-memory pruned line FFTs with the
                                * w-screen fused (ms_fft = pass 1, ms_screen =
                                * pass 2 + screen); 0: cuFFT 2-D plans */
    int32_t reserved_;
    /* pruned mode, timing enabled: (64 x 64 tile, plane) and (tile row,
     * plane) pairs that held visibilities in this call -- what the scatter /
     * gather touched and what the row pass transformed, out of
     * planes_processed x ntile^2 and planes_processed x ntile
     * (ntile = ceil(grid_size / 64)) */
    int64_t occupied_tiles, occupied_tile_rows;
} sdp_EsCallStats;

/* enable != 0: record CUDA-event timings per stage (adds stream syncs at the
 * end of each call). */
void sdp_gridder_uvw_es_fft_enable_timing(sdp_GridderUvwEsFft* plan,
        int32_t enable);
void sdp_gridder_uvw_es_fft_get_stats(const sdp_GridderUvwEsFft* plan,
        sdp_EsCallStats* stats);

/* Device time of every w-plane batch of the last grid / degrid call (timing
 * enabled): batch k covered planes [first_plane[k], first_plane[k] +
 * num_planes[k]) and took ms[k]. Returns the number of batches written (at
 * most capacity). A caller that shards the lattice over several GPUs uses
 * these as measured per-plane costs to balance plane ownership. */
int32_t sdp_gridder_uvw_es_fft_get_batch_times(
        const sdp_GridderUvwEsFft* plan, int32_t capacity,
        int32_t* first_plane, int32_t* num_planes, double* ms);

#ifdef __cplusplus
}
#endif

#endif
