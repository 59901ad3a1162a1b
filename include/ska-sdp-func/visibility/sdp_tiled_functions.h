/*
 * Tiled visibility functions: which uv tiles every visibility's footprint
 * touches -- counted per tile and prefix-summed, then the visibilities (or
 * their indices) bucket-sorted by tile, duplicated where a footprint spans
 * several tiles. The presort that uv-tiled weighting and gridding consume.
 * Drop-in for the reference's
 *     src/ska-sdp-func/visibility/sdp_tiled_functions.h:62-76   sdp_count_and_prefix_sum
 *     src/ska-sdp-func/visibility/sdp_tiled_functions.h:136-153 sdp_bucket_sort
 *     src/ska-sdp-func/visibility/sdp_tiled_functions.h:200-217 sdp_tiled_indexing
 * (same prototypes, argument meaning, data-model checks and error codes,
 * sdp_tiled_functions.cpp:302-363,493-572,712-760; arithmetic of its kernels,
 * sdp_tiled_functions.cu:63-288).
 *
 * Array shapes, slowest to fastest varying:
 *   uvw     [num_times, num_baselines, 3]  real, metres
 *   freqs   [num_channels]                 real, Hz (same precision as uvw)
 *   vis     [num_times, num_baselines, num_channels, num_pols] complex
 *   weights [num_times, num_baselines, num_channels, num_pols] real
 *   tile_offsets        [>= num_tiles + 1] int32
 *   num_points_in_tiles [>= num_tiles]     int32
 *   num_skipped         [>= 1]             int32
 *   sorted_uu / sorted_vv / sorted_vis / sorted_weight / sorted_tile /
 *   sorted_vis_index    [num_visibilites]  (real x4, int32 x2)
 * with num_tiles = ceil(grid_size / tile_size_u) * ceil(grid_size / tile_size_v).
 *
 * Tiles: the grid centre grid_size / 2 is the middle of a tile
 * (top_left = centre - (centre / tile_size) * tile_size - tile_size / 2). A
 * visibility sits at cell round(u f / c * grid_size * cell_size_rad) + centre,
 * is SKIPPED unless [cell - support, cell + support] lies on the grid, and
 * otherwise belongs to every tile in
 *   [floor((rel - support) / tile_size), ceil((rel + support + 1) / tile_size))
 * per axis, rel = cell - top_left, the quotients formed in fp32 as products
 * with 1 / tile_size. The linear tile index is pu + pv * num_tiles_u, the tile
 * code stored in sorted_tile is pv * 32768 + pu.
 *
 * sdp_bucket_sort / sdp_tiled_indexing write entry k of tile t to slot
 * tile_offsets[t] + k and leave tile_offsets[t] advanced by the entries of
 * tile t, as the reference's atomic slot counters do. Differences, all
 * deliberate: (i) inside a tile the entries are in the order of the
 * reference's host loops (time, baseline, channel; pv outer, pu inner) on
 * every run -- the reference's kernels leave them in the arbitrary order of
 * their atomics; (ii) pairs whose linear tile index leaves [0, num_tiles) and
 * slots beyond the sorted arrays are skipped -- the reference indexes out of
 * bounds there (a tile column pu = num_tiles_u, which holds the last half
 * tile of a grid whose size is a multiple of the tile size, keeps the linear
 * index and the tile code the reference forms for it, so counts and offsets
 * are the reference's); (iii) array types and minimum sizes are checked
 * (SDP_ERR_DATA_TYPE / SDP_ERR_INVALID_ARGUMENT); (iv) sdp_bucket_sort reads
 * `vis` as the reference's kernels do -- element i_vis of the array taken as
 * reals of the coordinates' precision -- when sorted_vis is real, and copies
 * the complex visibility of polarisation 0 when sorted_vis is complex.
 *
 * This implementation runs on the GPU only: SDP_MEM_GPU arrays are processed
 * in place on the legacy default stream (each call reads one count back and
 * therefore synchronises that stream once); SDP_MEM_CPU arrays are staged
 * through the current device (the reference has a host loop for them).
 * All arrays of a call must be in the same location.
 */
#ifndef SKA_SDP_PROC_FUNC_TILED_H_
#define SKA_SDP_PROC_FUNC_TILED_H_

#include "ska-sdp-func/utility/sdp_mem.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Visibilities per tile, their exclusive prefix sum and total (returned in
 * *num_visibilites), and the number of skipped visibilities. The three
 * output arrays are cleared first. */
void sdp_count_and_prefix_sum(
        const sdp_Mem* uvw,
        const sdp_Mem* freqs,
        const sdp_Mem* vis,
        const int grid_size,
        const int64_t tile_size_u,
        const int64_t tile_size_v,
        const double cell_size_rad,
        const int64_t support,
        int* num_visibilites,
        sdp_Mem* tile_offsets,
        sdp_Mem* num_points_in_tiles,
        sdp_Mem* num_skipped,
        sdp_Error* status
);

/* Position (in cells from the centre), visibility, weight and tile code of
 * every (visibility, tile) pair, grouped by tile through tile_offsets (as
 * computed by sdp_count_and_prefix_sum with the same geometry). The sorted
 * arrays are cleared first. */
void sdp_bucket_sort(
        const sdp_Mem* uvw,
        const sdp_Mem* freqs,
        const sdp_Mem* vis,
        const sdp_Mem* weights,
        const int grid_size,
        const int64_t tile_size_u,
        const int64_t tile_size_v,
        const double cell_size_rad,
        const int64_t support,
        sdp_Mem* sorted_uu,
        sdp_Mem* sorted_vv,
        sdp_Mem* sorted_weight,
        sdp_Mem* sorted_tile,
        sdp_Mem* sorted_vis,
        sdp_Mem* tile_offsets,
        sdp_Error* status
);

/* As sdp_bucket_sort, with the flat index (time, baseline, channel) of the
 * visibility in place of its value and weight. */
void sdp_tiled_indexing(
        const sdp_Mem* uvw,
        const sdp_Mem* freqs,
        const int grid_size,
        const int64_t tile_size_u,
        const int64_t tile_size_v,
        const double cell_size_rad,
        const int64_t support,
        const int64_t num_channels,
        const int64_t num_baselines,
        const int64_t num_times,
        sdp_Mem* sorted_tile,
        sdp_Mem* sorted_uu,
        sdp_Mem* sorted_vv,
        sdp_Mem* sorted_vis_index,
        sdp_Mem* tile_offsets,
        sdp_Error* status
);

#ifdef __cplusplus
}
#endif

#endif /* include guard */
