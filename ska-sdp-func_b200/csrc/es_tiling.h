/*
 * C launchers of the tiled visibility functions (es_tiling.cu): plain device
 * pointers and sizes; work is enqueued on `stream`; the return value is a
 * cudaError_t as int, or ES_TILE_ERR_TOO_MANY when the (visibility, tile)
 * pairs do not fit the reference's int32 offsets.
 */
#ifndef ES_TILING_H_
#define ES_TILING_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ES_TILE_ERR_TOO_MANY (-2)

/* Tile geometry of sdp_tiled_functions.cpp:333-342, filled by the host. */
typedef struct EsTileGeom
{
    int64_t num_rows;       /* num_times * num_baselines */
    int num_chan;
    int grid_size;
    int64_t support;
    float inv_tile_size_u, inv_tile_size_v;
    int64_t num_tiles_u, num_tiles_v;
    int64_t top_left_u, top_left_v;
    double grid_scale;      /* grid_size * cell_size_rad */
} EsTileGeom;

/* counts[num_tiles] += visibilities per tile, *skipped += off-grid ones (both
 * must be zero on entry), offsets[0 .. num_tiles] = exclusive prefix sum;
 * *num_entries (host) = offsets[num_tiles], read back (one stream sync). */
int es_tile_count(const EsTileGeom* g, const void* uvw, const void* freqs,
        int is_double, int32_t* counts, int32_t* offsets, int32_t* skipped,
        int* num_entries, void* stream);

/* Stable bucket sort of the (visibility, tile) pairs into the caller's
 * tile_offsets (advanced by the entries of each tile on return, as the
 * reference leaves them). Outputs that are NULL are not written:
 * bucket sort passes vis / weights / sorted_vis / sorted_weight, tiled
 * indexing passes sorted_vis_index. vis_complex: sorted_vis is complex and
 * receives element i_vis * num_pols of the complex vis; otherwise element
 * i_vis of vis taken as reals (the reference's indexing). num_out = elements
 * of the sorted arrays (slots beyond it are skipped). */
int es_tile_sort(const EsTileGeom* g, const void* uvw, const void* freqs,
        int is_double, const void* vis, const void* weights, int vis_complex,
        int num_pols, int32_t* tile_offsets, void* sorted_uu, void* sorted_vv,
        void* sorted_vis, void* sorted_weight, int32_t* sorted_tile,
        int32_t* sorted_vis_index, int64_t num_out, void* stream);

/* Frees the per-device scratch the sort keeps between calls. */
void es_tile_release_scratch(void);

#ifdef __cplusplus
}
#endif

#endif
