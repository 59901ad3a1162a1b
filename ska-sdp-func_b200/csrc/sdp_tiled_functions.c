/*
 * Host side of the tiled visibility functions: the reference's argument
 * contract (sdp_tiled_functions.cpp:302-363, :493-572, :712-760 -- its checks
 * in its order with its error codes, through the data-model checks of
 * sdp_data_model_checks.c:7-43,123-159,162-203), the sizes and types its
 * kernels index without checking, then the sm_100a path of es_tiling.cu.
 * There is no host loop: SDP_MEM_CPU arrays are staged through the current
 * device (the reference's host branch is restated in oracle/tiled_oracle.py).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <cuda_runtime_api.h>

#include "es_tiling.h"
#include "ska-sdp-func/utility/sdp_logging.h"
#include "ska-sdp-func/visibility/sdp_tiled_functions.h"

#define TL_MAX_ARRAYS 12

/* Device views of the arrays of one call; host arrays are copied in, and the
 * outputs among them copied back at the end. */
typedef struct TlStage
{
    int on_host;
    int n;
    void* staged[TL_MAX_ARRAYS];
    void* host[TL_MAX_ARRAYS];
    size_t bytes[TL_MAX_ARRAYS];
    int is_output[TL_MAX_ARRAYS];
} TlStage;

static size_t mem_bytes(const sdp_Mem* m)
{
    return (size_t) sdp_mem_num_elements(m) *
            (size_t) sdp_mem_type_size(sdp_mem_type(m));
}

static void cuda_failed(cudaError_t err, sdp_Error code, sdp_Error* status)
{
    *status = code;
    SDP_LOG_CRITICAL("CUDA error: %s", cudaGetErrorString(err));
    (void) cudaGetLastError();
}

/* Device pointer of `m` (NULL for a NULL array). `copy_in` = 0 for outputs
 * the call clears anyway: they are zeroed on the device instead. */
static void* stage(TlStage* s, const sdp_Mem* m, int is_output, int copy_in,
        sdp_Error* status)
{
    if (*status || !m) return 0;
    void* data = (void*) sdp_mem_data_const(m);
    if (!s->on_host) return data;
    const size_t bytes = mem_bytes(m);
    void* dev = 0;
    cudaError_t err = cudaMalloc(&dev, bytes ? bytes : 16);
    if (err != cudaSuccess)
    {
        cuda_failed(err, SDP_ERR_MEM_ALLOC_FAILURE, status);
        return 0;
    }
    const int k = s->n++;
    s->staged[k] = dev;
    s->host[k] = data;
    s->bytes[k] = bytes;
    s->is_output[k] = is_output;
    err = copy_in ?
            cudaMemcpyAsync(dev, data, bytes, cudaMemcpyHostToDevice, 0) :
            cudaMemsetAsync(dev, 0, bytes, 0);
    if (err != cudaSuccess) cuda_failed(err, SDP_ERR_MEM_COPY_FAILURE, status);
    return dev;
}

static void finish(TlStage* s, sdp_Error* status)
{
    if (!s->on_host) return;
    for (int k = 0; k < s->n && !*status; ++k)
    {
        if (!s->is_output[k] || !s->bytes[k]) continue;
        cudaError_t err = cudaMemcpyAsync(s->host[k], s->staged[k], s->bytes[k],
                cudaMemcpyDeviceToHost, 0);
        if (err != cudaSuccess) cuda_failed(err, SDP_ERR_MEM_COPY_FAILURE, status);
    }
    cudaError_t err = cudaStreamSynchronize(0);
    if (err != cudaSuccess && !*status)
        cuda_failed(err, SDP_ERR_MEM_COPY_FAILURE, status);
    for (int k = 0; k < s->n; ++k) (void) cudaFree(s->staged[k]);
}

/* sdp_data_model_get_vis_metadata, sdp_data_model_checks.c:123-159 */
static void vis_metadata(const sdp_Mem* vis, sdp_MemLocation* location,
        int64_t dims[4], sdp_Error* status)
{
    if (*status) return;
    if (!sdp_mem_is_complex(vis))
    {
        *status = SDP_ERR_DATA_TYPE;
        SDP_LOG_ERROR("The visibility array must be complex");
        return;
    }
    if (sdp_mem_num_dims(vis) != 4)
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("Expected 'vis' to have 4 dimensions");
        return;
    }
    *location = sdp_mem_location(vis);
    for (int i = 0; i < 4; ++i) dims[i] = sdp_mem_shape_dim(vis, i);
    if (dims[3] != 4 && dims[3] != 1)
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("The number of polarisations should be 4 or 1");
    }
}

/* sdp_data_model_check_uvw_at, sdp_data_model_checks.c:7-43 */
static void check_uvw(const sdp_Mem* uvw, sdp_MemLocation location,
        int64_t num_times, int64_t num_baselines, sdp_Error* status)
{
    if (*status) return;
    if (!sdp_mem_is_floating_point(uvw) || sdp_mem_is_complex(uvw))
    {
        *status = SDP_ERR_DATA_TYPE;
        SDP_LOG_ERROR("The uvw array must be real-valued");
        return;
    }
    if (sdp_mem_num_dims(uvw) != 3 ||
            sdp_mem_shape_dim(uvw, 0) != num_times ||
            sdp_mem_shape_dim(uvw, 1) != num_baselines ||
            sdp_mem_shape_dim(uvw, 2) != 3)
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("Expected 'uvw' to have shape [num_times, "
                "num_baselines, 3]");
        return;
    }
    if (sdp_mem_location(uvw) != location)
    {
        *status = SDP_ERR_MEM_LOCATION;
        SDP_LOG_ERROR("Expected 'uvw' to be where the other arrays are");
        return;
    }
    if (!sdp_mem_is_c_contiguous(uvw))
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("Expected 'uvw' to be C contiguous!");
    }
}

/* sdp_data_model_check_weights_at, sdp_data_model_checks.c:162-203 */
static void check_weights(const sdp_Mem* weights, sdp_MemLocation location,
        const int64_t dims[4], sdp_Error* status)
{
    if (*status) return;
    if (sdp_mem_is_complex(weights))
    {
        *status = SDP_ERR_DATA_TYPE;
        SDP_LOG_ERROR("The weights array cannot be complex");
        return;
    }
    int same = sdp_mem_num_dims(weights) == 4;
    for (int i = 0; i < 4 && same; ++i)
        same = sdp_mem_shape_dim(weights, i) == dims[i];
    if (!same)
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("Expected 'weights' to have the shape of 'vis'");
        return;
    }
    if (sdp_mem_location(weights) != location)
    {
        *status = SDP_ERR_MEM_LOCATION;
        SDP_LOG_ERROR("Expected 'weights' to be where 'vis' is");
        return;
    }
    if (!sdp_mem_is_c_contiguous(weights))
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("Expected 'weights' to be C contiguous!");
    }
}

/* What the kernels write without further checks: a writeable, contiguous
 * array of `type` with at least `min_elements` elements, where the inputs
 * are. */
static void check_output(const sdp_Mem* m, const char* name, sdp_MemType type,
        sdp_MemLocation location, int64_t min_elements, sdp_Error* status)
{
    if (*status) return;
    if (!m)
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("'%s' is NULL", name);
        return;
    }
    if (sdp_mem_type(m) != type)
    {
        *status = SDP_ERR_DATA_TYPE;
        SDP_LOG_ERROR("Unsupported data type(s): '%s'", name);
        return;
    }
    if (sdp_mem_location(m) != location)
    {
        *status = SDP_ERR_MEM_LOCATION;
        SDP_LOG_ERROR("Expected '%s' to be where the other arrays are", name);
        return;
    }
    if (sdp_mem_is_read_only(m) || !sdp_mem_is_c_contiguous(m) ||
            sdp_mem_num_elements(m) < min_elements)
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("Expected '%s' to be writeable, C contiguous and to "
                "hold at least %lld elements", name, (long long) min_elements);
    }
}

/* Geometry of .cpp:333-342 plus the limits of the reference's int32 indices. */
static void geometry(int grid_size, int64_t tile_size_u, int64_t tile_size_v,
        double cell_size_rad, int64_t support, const int64_t dims[3],
        EsTileGeom* g, sdp_Error* status)
{
    if (*status) return;
    if (grid_size < 1 || tile_size_u < 1 || tile_size_v < 1 || support < 0 ||
            support > INT32_MAX / 4 || tile_size_u > INT32_MAX / 4 ||
            tile_size_v > INT32_MAX / 4)
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("grid_size, tile_size_u and tile_size_v must be "
                "positive, support must not be negative");
        return;
    }
    const int64_t centre = grid_size / 2;
    g->num_rows = dims[0] * dims[1];
    g->num_chan = (int) dims[2];
    g->grid_size = grid_size;
    g->support = support;
    g->inv_tile_size_u = (float) (1.0 / (double) tile_size_u);
    g->inv_tile_size_v = (float) (1.0 / (double) tile_size_v);
    g->num_tiles_u = (grid_size + tile_size_u - 1) / tile_size_u;
    g->num_tiles_v = (grid_size + tile_size_v - 1) / tile_size_v;
    g->top_left_u = centre - (centre / tile_size_u) * tile_size_u -
            tile_size_u / 2;
    g->top_left_v = centre - (centre / tile_size_v) * tile_size_v -
            tile_size_v / 2;
    g->grid_scale = grid_size * cell_size_rad;
    if (dims[2] > INT32_MAX || g->num_rows > INT32_MAX ||
            g->num_rows * dims[2] > INT32_MAX ||
            g->num_tiles_u * g->num_tiles_v >= INT32_MAX ||
            g->num_tiles_u >= 32768)
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("More visibilities or tiles than the int32 indices "
                "and the 'pv * 32768 + pu' tile code can hold");
    }
}

static void check_freqs(const sdp_Mem* freqs, sdp_MemType type,
        sdp_MemLocation location, int64_t num_channels, sdp_Error* status)
{
    if (*status) return;
    if (sdp_mem_type(freqs) != type)
    {
        *status = SDP_ERR_DATA_TYPE;
        SDP_LOG_ERROR("Unsupported data type(s)");
        return;
    }
    if (sdp_mem_location(freqs) != location)
    {
        *status = SDP_ERR_MEM_LOCATION;
        SDP_LOG_ERROR("Expected 'freqs' to be where the other arrays are");
        return;
    }
    if (sdp_mem_num_dims(freqs) != 1 ||
            sdp_mem_shape_dim(freqs, 0) != num_channels ||
            !sdp_mem_is_c_contiguous(freqs))
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("Expected 'freqs' to have shape [num_channels]");
    }
}

static void launch_failed(int rc, sdp_Error* status)
{
    if (rc == ES_TILE_ERR_TOO_MANY)
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("More (visibility, tile) pairs than an int32 offset "
                "can hold");
    }
    else
    {
        *status = rc == (int) cudaErrorMemoryAllocation ?
                SDP_ERR_MEM_ALLOC_FAILURE : SDP_ERR_RUNTIME;
        SDP_LOG_ERROR("CUDA error: %s", cudaGetErrorString((cudaError_t) rc));
        (void) cudaGetLastError();
    }
}

void sdp_count_and_prefix_sum(const sdp_Mem* uvw, const sdp_Mem* freqs,
        const sdp_Mem* vis, const int grid_size, const int64_t tile_size_u,
        const int64_t tile_size_v, const double cell_size_rad,
        const int64_t support, int* num_visibilites, sdp_Mem* tile_offsets,
        sdp_Mem* num_points_in_tiles, sdp_Mem* num_skipped, sdp_Error* status)
{
    if (*status) return;
    sdp_MemLocation location = SDP_MEM_CPU;
    int64_t dims[4] = {0, 0, 0, 0};
    EsTileGeom g;
    memset(&g, 0, sizeof(g));
    vis_metadata(vis, &location, dims, status);
    check_uvw(uvw, location, dims[0], dims[1], status);
    if (*status) return;
    const sdp_MemType real = sdp_mem_type(uvw);
    check_freqs(freqs, real, location, dims[2], status);
    geometry(grid_size, tile_size_u, tile_size_v, cell_size_rad, support, dims,
            &g, status);
    if (*status) return;
    const int64_t num_tiles = g.num_tiles_u * g.num_tiles_v;
    check_output(tile_offsets, "tile_offsets", SDP_MEM_INT, location,
            num_tiles + 1, status);
    check_output(num_points_in_tiles, "num_points_in_tiles", SDP_MEM_INT,
            location, num_tiles, status);
    check_output(num_skipped, "num_skipped", SDP_MEM_INT, location, 1, status);
    if (!*status && !num_visibilites)
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("'num_visibilites' is NULL");
    }
    if (*status) return;

    TlStage s;
    memset(&s, 0, sizeof(s));
    s.on_host = (location != SDP_MEM_GPU);
    const void* d_uvw = stage(&s, uvw, 0, 1, status);
    const void* d_freqs = stage(&s, freqs, 0, 1, status);
    int32_t* d_offsets = (int32_t*) stage(&s, tile_offsets, 1, 0, status);
    int32_t* d_counts = (int32_t*) stage(&s, num_points_in_tiles, 1, 0, status);
    int32_t* d_skipped = (int32_t*) stage(&s, num_skipped, 1, 0, status);
    if (!*status && !s.on_host)
    {
        /* the three outputs start from zero, whole arrays (.cpp:319-321) */
        sdp_mem_clear_contents(tile_offsets, status);
        sdp_mem_clear_contents(num_points_in_tiles, status);
        sdp_mem_clear_contents(num_skipped, status);
    }
    if (!*status)
    {
        const int rc = es_tile_count(&g, d_uvw, d_freqs,
                real == SDP_MEM_DOUBLE, d_counts, d_offsets, d_skipped,
                num_visibilites, 0);
        if (rc != 0) launch_failed(rc, status);
    }
    finish(&s, status);
}

/* Shared by the bucket sort and the indexing: `vis`, `weights`, `sorted_vis`
 * and `sorted_weight` are NULL for the indexing, `sorted_vis_index` for the
 * sort. `dims` = [num_times, num_baselines, num_channels, num_pols]. */
static void sort_into_tiles(const sdp_Mem* uvw, const sdp_Mem* freqs,
        const sdp_Mem* vis, const sdp_Mem* weights, sdp_MemLocation location,
        const int64_t dims[4], int grid_size, int64_t tile_size_u,
        int64_t tile_size_v, double cell_size_rad, int64_t support,
        sdp_Mem* sorted_uu, sdp_Mem* sorted_vv, sdp_Mem* sorted_weight,
        sdp_Mem* sorted_tile, sdp_Mem* sorted_vis, sdp_Mem* sorted_vis_index,
        sdp_Mem* tile_offsets, sdp_Error* status)
{
    if (*status) return;
    EsTileGeom g;
    memset(&g, 0, sizeof(g));
    const sdp_MemType real = sdp_mem_type(uvw);
    check_freqs(freqs, real, location, dims[2], status);
    if (!*status && weights && sdp_mem_type(weights) != real)
    {
        *status = SDP_ERR_DATA_TYPE;
        SDP_LOG_ERROR("Unsupported data type(s)");
    }
    geometry(grid_size, tile_size_u, tile_size_v, cell_size_rad, support, dims,
            &g, status);
    if (*status) return;
    check_output(tile_offsets, "tile_offsets", SDP_MEM_INT, location,
            g.num_tiles_u * g.num_tiles_v + 1, status);
    check_output(sorted_uu, "sorted_uu", real, location, 0, status);
    check_output(sorted_vv, "sorted_vv", real, location, 0, status);
    check_output(sorted_tile, "sorted_tile", SDP_MEM_INT, location, 0, status);
    if (*status) return;
    int64_t num_out = sdp_mem_num_elements(sorted_uu);
    if (sdp_mem_num_elements(sorted_vv) < num_out)
        num_out = sdp_mem_num_elements(sorted_vv);
    if (sdp_mem_num_elements(sorted_tile) < num_out)
        num_out = sdp_mem_num_elements(sorted_tile);
    int vis_complex = 0;
    if (vis)
    {
        /* the reference's kernels take `vis` and `sorted_vis` as reals of the
         * coordinates' precision (.cu:143,151,203): a real sorted_vis gets
         * element i_vis of that view. A complex sorted_vis (an extension)
         * gets the complex visibility. */
        vis_complex = sdp_mem_is_complex(sorted_vis);
        const sdp_MemType want = vis_complex ?
                (sdp_MemType) (real | SDP_MEM_COMPLEX) : real;
        if ((sdp_MemType) (sdp_mem_type(vis) & ~SDP_MEM_COMPLEX) != real)
        {
            *status = SDP_ERR_DATA_TYPE;
            SDP_LOG_ERROR("Unsupported data type(s): 'vis'");
            return;
        }
        check_output(sorted_vis, "sorted_vis", want, location, 0, status);
        check_output(sorted_weight, "sorted_weight", real, location, 0, status);
        if (*status) return;
        if (sdp_mem_num_elements(sorted_vis) < num_out)
            num_out = sdp_mem_num_elements(sorted_vis);
        if (sdp_mem_num_elements(sorted_weight) < num_out)
            num_out = sdp_mem_num_elements(sorted_weight);
    }
    else
    {
        check_output(sorted_vis_index, "sorted_vis_index", SDP_MEM_INT,
                location, 0, status);
        if (*status) return;
        if (sdp_mem_num_elements(sorted_vis_index) < num_out)
            num_out = sdp_mem_num_elements(sorted_vis_index);
    }

    TlStage s;
    memset(&s, 0, sizeof(s));
    s.on_host = (location != SDP_MEM_GPU);
    const void* d_uvw = stage(&s, uvw, 0, 1, status);
    const void* d_freqs = stage(&s, freqs, 0, 1, status);
    const void* d_vis = stage(&s, vis, 0, 1, status);
    const void* d_weights = stage(&s, weights, 0, 1, status);
    int32_t* d_offsets = (int32_t*) stage(&s, tile_offsets, 1, 1, status);
    void* d_uu = stage(&s, sorted_uu, 1, 0, status);
    void* d_vv = stage(&s, sorted_vv, 1, 0, status);
    void* d_weight = stage(&s, sorted_weight, 1, 0, status);
    int32_t* d_tile = (int32_t*) stage(&s, sorted_tile, 1, 0, status);
    void* d_svis = stage(&s, sorted_vis, 1, 0, status);
    int32_t* d_index = (int32_t*) stage(&s, sorted_vis_index, 1, 0, status);
    if (!*status && !s.on_host)
    {
        /* the sorted arrays start from zero (.cpp:514-518, :733-736) */
        sdp_mem_clear_contents(sorted_uu, status);
        sdp_mem_clear_contents(sorted_vv, status);
        sdp_mem_clear_contents(sorted_tile, status);
        if (sorted_weight) sdp_mem_clear_contents(sorted_weight, status);
        if (sorted_vis) sdp_mem_clear_contents(sorted_vis, status);
        if (sorted_vis_index) sdp_mem_clear_contents(sorted_vis_index, status);
    }
    if (!*status)
    {
        const int rc = es_tile_sort(&g, d_uvw, d_freqs, real == SDP_MEM_DOUBLE,
                d_vis, d_weights, vis_complex, (int) dims[3], d_offsets, d_uu,
                d_vv, d_svis, d_weight, d_tile, d_index, num_out, 0);
        if (rc != 0) launch_failed(rc, status);
    }
    finish(&s, status);
}

void sdp_bucket_sort(const sdp_Mem* uvw, const sdp_Mem* freqs,
        const sdp_Mem* vis, const sdp_Mem* weights, const int grid_size,
        const int64_t tile_size_u, const int64_t tile_size_v,
        const double cell_size_rad, const int64_t support, sdp_Mem* sorted_uu,
        sdp_Mem* sorted_vv, sdp_Mem* sorted_weight, sdp_Mem* sorted_tile,
        sdp_Mem* sorted_vis, sdp_Mem* tile_offsets, sdp_Error* status)
{
    if (*status) return;
    sdp_MemLocation location = SDP_MEM_CPU;
    int64_t dims[4] = {0, 0, 0, 0};
    vis_metadata(vis, &location, dims, status);
    check_uvw(uvw, location, dims[0], dims[1], status);
    check_weights(weights, location, dims, status);
    if (!*status && (!sorted_vis || !sorted_weight))
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("'sorted_vis' / 'sorted_weight' is NULL");
    }
    sort_into_tiles(uvw, freqs, vis, weights, location, dims, grid_size,
            tile_size_u, tile_size_v, cell_size_rad, support, sorted_uu,
            sorted_vv, sorted_weight, sorted_tile, sorted_vis, 0, tile_offsets,
            status);
}

void sdp_tiled_indexing(const sdp_Mem* uvw, const sdp_Mem* freqs,
        const int grid_size, const int64_t tile_size_u,
        const int64_t tile_size_v, const double cell_size_rad,
        const int64_t support, const int64_t num_channels,
        const int64_t num_baselines, const int64_t num_times,
        sdp_Mem* sorted_tile, sdp_Mem* sorted_uu, sdp_Mem* sorted_vv,
        sdp_Mem* sorted_vis_index, sdp_Mem* tile_offsets, sdp_Error* status)
{
    if (*status) return;
    const sdp_MemLocation location = sdp_mem_location(uvw);
    const int64_t dims[4] = {num_times, num_baselines, num_channels, 1};
    check_uvw(uvw, location, num_times, num_baselines, status);
    if (!*status && !sorted_vis_index)
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("'sorted_vis_index' is NULL");
    }
    sort_into_tiles(uvw, freqs, 0, 0, location, dims, grid_size, tile_size_u,
            tile_size_v, cell_size_rad, support, sorted_uu, sorted_vv, 0,
            sorted_tile, 0, sorted_vis_index, tile_offsets, status);
}
