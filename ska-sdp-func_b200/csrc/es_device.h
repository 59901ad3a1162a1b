/*
 * Thin C-ABI launcher interface between the C host side (es_gridder.c) and
 * the sm_100a kernels (es_kernels.cu). Plain pointers and sizes only.
 * Every function enqueues work on `stream` and returns a cudaError_t as int
 * (0 == cudaSuccess); nothing here synchronises unless stated.
 */
#ifndef ES_DEVICE_H_
#define ES_DEVICE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ES_TILE 64 /* uv tile edge in grid cells; must be >= support + 1 */
#define ES_MAX_DEVICES 64 /* per-device launch configuration caches */

/* Geometry shared by all kernels (doubles: coordinates are always fp64). */
typedef struct EsGeom
{
    int32_t num_rows, num_chan, image_size, grid_size, support;
    int32_t num_planes;    /* global lattice size P */
    int32_t do_wstacking;
    int32_t is_double;     /* data precision: 0 fp32, 1 fp64 */
    int32_t ntile;         /* tiles per axis = ceil(grid_size / ES_TILE) */
    int32_t plane_begin;   /* owned w-plane range [plane_begin, plane_end) */
    int32_t plane_end;
    int32_t pad_;
    double beta, pixel_size, uv_scale, w_scale, inv_w_scale, min_plane_w;
    double conv_corr_norm_factor;
} EsGeom;

typedef void* es_stream_t; /* cudaStream_t */

/* Bytes of one sorted visibility record for the given precision. */
size_t es_record_bytes(int is_double);

/* ---- presort (once per grid / degrid call) ---- */

/* Per visibility (row, chan): number of (tile) entries it produces for the
 * owned plane range -> counts[row * C + chan] (uint8). */
int es_launch_count_entries(const EsGeom* g, const void* uvw, const void* freq,
        uint8_t* counts, es_stream_t stream);

/* Exclusive prefix sum uint8 counts -> uint32 offsets (n + 1 outputs: the
 * last one is the total). temp == NULL: only *temp_bytes is set. */
int es_scan_counts(const uint8_t* counts, uint32_t* offsets, int64_t n,
        void* temp, size_t* temp_bytes, es_stream_t stream);

/* Emit (key, value) = (tile_id * PK + first owned plane - plane_begin,
 * visibility index) at offsets[vis] .. */
int es_launch_emit_entries(const EsGeom* g, const void* uvw, const void* freq,
        const uint32_t* offsets, uint64_t* keys, uint32_t* vals,
        es_stream_t stream);

/* Stable LSD radix sort of the pairs on key bits [0, end_bit).
 * temp == NULL: only *temp_bytes is set. */
int es_sort_entries(const uint64_t* keys_in, uint64_t* keys_out,
        const uint32_t* vals_in, uint32_t* vals_out, int64_t num_entries,
        int end_bit, void* temp, size_t* temp_bytes, es_stream_t stream);

/* tile_start[t] = first sorted entry of tile t (t in [0, ntile^2]); */
int es_launch_tile_bounds(const EsGeom* g, const uint64_t* keys,
        int64_t num_entries, uint32_t* tile_start, es_stream_t stream);

/* Build the sorted records. vis/weight non-NULL: scatter records carrying the
 * weighted (conjugated if w < 0) visibility; NULL: gather records carrying
 * the flip sign. entry_pos non-NULL (degridding): entry_pos[offsets[i] + k] =
 * sorted position of the k-th entry of visibility i. */
int es_launch_build_records(const EsGeom* g, const void* uvw, const void* freq,
        const void* vis, const void* weight, const uint64_t* keys,
        const uint32_t* vals, int64_t num_entries, void* records,
        const uint32_t* offsets, uint32_t* entry_pos, void* kweights,
        es_stream_t stream);

/* Reals per record of the precomputed ES weights (kweights non-NULL above):
 * ku[ns], kv[ns] (with the (-1)^(cell) signs folded in), kw[ns],
 * ns = support + 1 -- plane-independent, so evaluated once per call instead
 * of once per (record, plane) in the fp64 scatter. */
size_t es_kweight_count(const EsGeom* g);

/* ---- per batch of resident w-planes [first_plane, first_plane + nplanes) ---- */

/* dirty += sum_planes Re(plane_centre * w-screen) (after the inverse FFT). */
int es_launch_apply_screen(const EsGeom* g, const void* planes,
        int first_plane, int nplanes, void* dirty, es_stream_t stream);

/* planes = zero-padded(dirty * conj w-screen) (before the forward FFT). */
int es_launch_reverse_screen(const EsGeom* g, const void* dirty,
        int first_plane, int nplanes, void* planes, es_stream_t stream);

/* vis[i] += sum over the entries of visibility i, in entry order, of
 * acc[entry] (imag * flip): one thread per visibility, bitwise reproducible. */
int es_launch_finish_degrid(const EsGeom* g, const void* records,
        const void* acc, const uint32_t* offsets, const uint32_t* entry_pos,
        int64_t num_entries, void* vis, es_stream_t stream);

/* ---- once per call ---- */

/* dirty *= 1 / (C(l) C(m) C(n) C(0)^2); tables are device doubles. */
int es_launch_correct(const EsGeom* g, void* dirty, const double* quad_kernel,
        const double* quad_nodes, const double* quad_weights,
        const double* conv_corr, es_stream_t stream);

/* hist[p] += number of visibilities whose first w-plane is p (int64[P] device,
 * pre-zeroed). */
int es_launch_plane_histogram(const EsGeom* g, const void* uvw,
        const void* freq, unsigned long long* hist, es_stream_t stream);

/* ---- occupancy of the resident planes (es_scatter.cu) ----
 * Per plane of a batch, es_occupancy_words(g) uint32 words:
 *   [tu * nw + (tv >> 5)] bit (tv & 31): tile (tu, tv) has records on the plane
 *   [ntile * nw + (tu >> 5)] bit (tu & 31): tile row tu has an occupied tile
 * with nw = (ntile + 31) / 32. Unoccupied tiles of `planes` and unoccupied
 * tile rows of the FFT intermediate are neither written nor read. */
size_t es_occupancy_words(const EsGeom* g);
/* totals (device, may be NULL): [0] += occupied (tile, plane) pairs,
 * [1] += occupied (tile row, plane) pairs of this batch. */
int es_launch_occupancy(const EsGeom* g, const uint64_t* keys,
        const uint32_t* tile_start, int first_plane, int nplanes,
        uint32_t* occ, unsigned long long* totals, es_stream_t stream);

/* ---- load-balanced work units (es_scatter.cu) ---- */

size_t es_unit_bytes(void);                   /* bytes of one work unit */
size_t es_partial_tile_bytes(int is_double);  /* bytes of one partial tile */

/* Builds the work list of planes [first_plane, first_plane + nplanes):
 * per_plane != 0: one item per (tile, plane) (1: plane-major list, 2:
 * tile-major list), else one per tile for the whole range. Items with more than `split` records become several units (slices).
 * counters: 8 x uint32 on the device ([0] = number of units afterwards);
 * split_done: max_split x uint32 completion counters. Both are zeroed here.
 * skip_empty != 0: items without records produce no unit. */
int es_launch_work_plan(const EsGeom* g, const uint64_t* keys,
        const uint32_t* tile_start, int first_plane, int nplanes,
        int per_plane, uint32_t split, uint32_t max_units, uint32_t max_split,
        uint32_t max_partials, int skip_empty, void* units, uint32_t* counters,
        uint32_t* split_done, es_stream_t stream);

/* Gridding: persistent CTAs execute the units; every cell of the nplanes
 * planes is written exactly once. scratch: max_partials partial tiles.
 * kweights (fp64 only, may be NULL): ES weights per record from
 * es_launch_build_records, es_kweight_count(g) reals per record. */
int es_launch_scatter2(const EsGeom* g, const void* records, const void* units,
        uint32_t* counters, int first_plane, void* planes, void* scratch,
        uint32_t* split_done, const void* kweights, es_stream_t stream);

/* Degridding: persistent CTAs execute the units of a per_plane == 0 work
 * list (one record belongs to exactly one unit). */
int es_launch_gather2(const EsGeom* g, const void* records, const void* units,
        uint32_t* counters, int first_plane, int nplanes, const void* planes,
        void* acc, es_stream_t stream);

/* ---- pruned w-plane FFT with fused w-screen (es_fft.cu) ---- */

typedef struct EsFft EsFft;

/* Plans the in-shared-memory line FFT for this grid / image size. Returns 0
 * and *out on success; -1 (and *out == NULL) when the size is not supported
 * (line does not fit in shared memory, prime factor > 11, or the environment
 * variable SKA_SDP_FUNC_ES_USE_CUFFT is set): the caller then uses cuFFT 2-D
 * plans; any other value is a cudaError_t. */
int es_fft_create(int grid_size, int image_size, int is_double, EsFft** out);
void es_fft_destroy(EsFft* fft);

/* Bytes of the transposed, pruned intermediate strip[plane][N][G]. */
size_t es_fft_strip_bytes(const EsFft* fft, int nplanes);

/* `occ` (all four passes): occupancy words of the batch, see above.
 * Gridding pass 1: planes[q][u][:] -> inverse FFT along v -> strip[q][v'][u]. */
int es_fft_grid_rows(const EsFft* fft, const void* planes, int nplanes,
        void* strip, const uint32_t* occ, es_stream_t stream);

/* Gridding pass 2: strip[q][v'][:] -> inverse FFT along u -> w-screen ->
 * image_t[v'][u'] += sum over the nplanes planes (image_t is the transposed
 * dirty image, N x N reals). */
int es_fft_grid_cols_screen(const EsFft* fft, const EsGeom* g,
        const void* strip, int first_plane, int nplanes, void* image_t,
        const uint32_t* occ, es_stream_t stream);

/* Degridding pass 1: image_t[v'][:] x conj w-screen, zero padded -> forward
 * FFT along u -> strip[q][v'][u]. */
int es_fft_degrid_cols(const EsFft* fft, const EsGeom* g, const void* image_t,
        int first_plane, int nplanes, void* strip, const uint32_t* occ,
        es_stream_t stream);

/* Degridding pass 2: strip[q][:][u] zero padded -> forward FFT along v ->
 * planes[q][u][:]. */
int es_fft_degrid_rows(const EsFft* fft, const void* strip, int nplanes,
        void* planes, const uint32_t* occ, es_stream_t stream);

/* Fills the plan's w-screen table (n - 1 per pixel of one image quadrant,
 * fp64) and chooses the fixed-point format of the phase from the largest
 * |w| of the lattice. Call once per plan that uses w-stacking; synchronises. */
int es_fft_set_screen(EsFft* fft, double pixel_size, double max_abs_w);

/* dst[r][c] = (add == 0) or += src[c][r] x (-1)^(r + c) (/ n(r, c) when
 * wstack != 0), N x N reals: the plane-independent factors of the w-screen
 * are applied by the one transpose per call between the caller's image and
 * the column-major image the line FFTs work on. */
int es_fft_image_transpose(const EsFft* fft, const void* src, void* dst,
        int add, int wstack, es_stream_t stream);

#ifdef __cplusplus
}
#endif

#endif
