/*
 * Host side of the uv-grid visibility weighting (uniform / Briggs): argument
 * contract of the reference (sdp_weighting.cpp:282-336 and :512-568: same
 * checks in the same order with the same error codes), then the sm_100a
 * kernels of es_weighting.cu. There is no host loop: SDP_MEM_CPU arrays are
 * staged through the current device (the reference's CPU branch, :339-399,
 * is what oracle/weighting_oracle.py restates).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

#include <cuda_runtime_api.h>

#include "es_weighting.h"
#include "ska-sdp-func/utility/sdp_logging.h"
#include "ska-sdp-func/visibility/sdp_weighting.h"

/* sdp_data_model_get_weights_metadata, sdp_data_model_checks.c:205-235 */
static void weights_metadata(const sdp_Mem* weights, sdp_MemType* type,
        sdp_MemLocation* location, int64_t dims[4], sdp_Error* status)
{
    if (*status) return;
    if (sdp_mem_num_dims(weights) != 4)
    {
        *status = SDP_ERR_RUNTIME;
        SDP_LOG_ERROR("The weights array must be 4D");
        return;
    }
    if (sdp_mem_is_complex(weights))
    {
        *status = SDP_ERR_DATA_TYPE;
        SDP_LOG_ERROR("The weights array cannot be complex");
        return;
    }
    *type = sdp_mem_type(weights);
    *location = sdp_mem_location(weights);
    for (int i = 0; i < 4; ++i) dims[i] = sdp_mem_shape_dim(weights, i);
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
                "num_baselines, 3] of the weights");
        return;
    }
    if (sdp_mem_location(uvw) != location)
    {
        *status = SDP_ERR_MEM_LOCATION;
        SDP_LOG_ERROR("Expected 'uvw' to be where the weights are");
        return;
    }
    if (!sdp_mem_is_c_contiguous(uvw))
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("Expected 'uvw' to be C contiguous!");
    }
}

static void check_location(const sdp_Mem* m, sdp_MemLocation location,
        const char* name, sdp_Error* status)
{
    if (*status) return;
    if (sdp_mem_location(m) != location)
    {
        *status = SDP_ERR_MEM_LOCATION;
        SDP_LOG_ERROR("Expected '%s' to be where the weights are", name);
    }
}

static void check_writeable(const sdp_Mem* m, const char* name,
        sdp_Error* status)
{
    if (*status) return;
    if (sdp_mem_is_read_only(m))
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("Expected '%s' not to be read-only!", name);
    }
}

typedef struct WtArrays
{
    const void* uvw;
    const void* freq;
    void* grid;
    const void* input;
    void* output;
    void* staged[5];      /* device copies of host arrays (freed at the end) */
    size_t bytes[5];
    int on_host;
} WtArrays;

static size_t mem_bytes(const sdp_Mem* m)
{
    size_t n = sdp_mem_type(m) == SDP_MEM_FLOAT ? 4 : 8;
    for (int i = 0; i < sdp_mem_num_dims(m); ++i)
        n *= (size_t) sdp_mem_shape_dim(m, i);
    return n;
}

static void release(WtArrays* a)
{
    for (int i = 0; i < 5; ++i)
        if (a->staged[i]) (void) cudaFree(a->staged[i]);
}

/* Device pointers of the five arrays; host arrays are copied to the device. */
static void resolve(const sdp_Mem* uvw, const sdp_Mem* freq, sdp_Mem* grid,
        const sdp_Mem* input, sdp_Mem* output, sdp_MemLocation location,
        WtArrays* a, sdp_Error* status)
{
    const sdp_Mem* m[5] = {uvw, freq, grid, input, output};
    const void* dev[5];
    for (int i = 0; i < 5; ++i) { a->staged[i] = 0; a->bytes[i] = mem_bytes(m[i]); }
    a->on_host = (location != SDP_MEM_GPU);
    for (int i = 0; i < 5 && !*status; ++i)
    {
        dev[i] = sdp_mem_data_const(m[i]);
        if (!a->on_host) continue;
        const size_t bytes = a->bytes[i] ? a->bytes[i] : 16;
        cudaError_t err = cudaMalloc(&a->staged[i], bytes);
        if (err != cudaSuccess)
        {
            a->staged[i] = 0;
            *status = SDP_ERR_MEM_ALLOC_FAILURE;
            SDP_LOG_CRITICAL("GPU memory allocation failure: %s",
                    cudaGetErrorString(err));
            (void) cudaGetLastError();
            break;
        }
        /* output is write-only except for visibilities outside the grid,
         * whose entries keep the caller's values: copied in as well */
        err = cudaMemcpyAsync(a->staged[i], dev[i], a->bytes[i],
                cudaMemcpyHostToDevice, 0);
        if (err != cudaSuccess)
        {
            *status = SDP_ERR_MEM_COPY_FAILURE;
            SDP_LOG_CRITICAL("cudaMemcpy error: %s", cudaGetErrorString(err));
            (void) cudaGetLastError();
            break;
        }
        dev[i] = a->staged[i];
    }
    a->uvw = dev[0];
    a->freq = dev[1];
    a->grid = (void*) dev[2];
    a->input = dev[3];
    a->output = (void*) dev[4];
}

static void copy_back(WtArrays* a, sdp_Mem* grid, sdp_Mem* output,
        sdp_Error* status)
{
    if (!a->on_host || *status) return;
    cudaError_t err = cudaMemcpyAsync(sdp_mem_data(grid), a->staged[2],
            a->bytes[2], cudaMemcpyDeviceToHost, 0);
    if (err == cudaSuccess) err = cudaMemcpyAsync(sdp_mem_data(output),
            a->staged[4], a->bytes[4], cudaMemcpyDeviceToHost, 0);
    if (err == cudaSuccess) err = cudaStreamSynchronize(0);
    if (err != cudaSuccess)
    {
        *status = SDP_ERR_MEM_COPY_FAILURE;
        SDP_LOG_CRITICAL("cudaMemcpy error: %s", cudaGetErrorString(err));
        (void) cudaGetLastError();
    }
}

static void weighting(const sdp_Mem* uvw, const sdp_Mem* freq_hz,
        double max_abs_uv, int briggs, double robust_param,
        sdp_Mem* weights_grid_uv, sdp_Mem* input_weight,
        sdp_Mem* output_weight, sdp_Error* status)
{
    if (*status) return;
    sdp_MemType weights_type = SDP_MEM_VOID;
    sdp_MemLocation location = sdp_mem_location(output_weight);
    int64_t dims[4] = {0, 0, 0, 0};

    /* the reference's checks, in its order (:297-336) */
    weights_metadata(input_weight, &weights_type, &location, dims, status);
    weights_metadata(output_weight, &weights_type, &location, dims, status);
    check_uvw(uvw, location, dims[0], dims[1], status);
    check_location(freq_hz, location, "freq_hz", status);
    check_location(weights_grid_uv, location, "weights_grid_uv", status);
    check_writeable(output_weight, "output_weight", status);
    check_writeable(weights_grid_uv, "weights_grid_uv", status);
    if (!*status && sdp_mem_num_dims(weights_grid_uv) != 3)
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("Expected 'weights_grid_uv' to have 3 dimensions");
    }
    if (*status) return;
    const int64_t grid_size = sdp_mem_shape_dim(weights_grid_uv, 0);
    if (sdp_mem_shape_dim(weights_grid_uv, 1) != grid_size)
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("Expected 'weights_grid_uv' to be square");
        return;
    }
    /* supported combinations (:342-399): fp64 coordinates, weights and grid
     * both fp64 or both fp32 */
    const int is_double = (weights_type == SDP_MEM_DOUBLE);
    if (sdp_mem_type(uvw) != SDP_MEM_DOUBLE ||
            sdp_mem_type(freq_hz) != SDP_MEM_DOUBLE ||
            (weights_type != SDP_MEM_DOUBLE && weights_type != SDP_MEM_FLOAT) ||
            sdp_mem_type(weights_grid_uv) != weights_type ||
            sdp_mem_type(input_weight) != weights_type)
    {
        *status = SDP_ERR_DATA_TYPE;
        SDP_LOG_ERROR("Unsupported data type(s)");
        return;
    }
    /* what the kernels index without further checks */
    if (sdp_mem_num_dims(freq_hz) != 1 ||
            sdp_mem_shape_dim(freq_hz, 0) != dims[2] ||
            sdp_mem_shape_dim(weights_grid_uv, 2) != dims[3] ||
            dims[2] > INT32_MAX || dims[3] > INT32_MAX)
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("freq_hz / weights_grid_uv do not match the weights' "
                "[num_times, num_baselines, num_channels, num_pols]");
        return;
    }
    for (int i = 0; i < 4; ++i)
        if (sdp_mem_shape_dim(input_weight, i) != dims[i])
        {
            *status = SDP_ERR_INVALID_ARGUMENT;
            SDP_LOG_ERROR("input and output weights differ in shape");
            return;
        }

    WtArrays a;
    resolve(uvw, freq_hz, weights_grid_uv, input_weight, output_weight,
            location, &a, status);
    double* sums = 0;
    if (!*status && briggs)
    {
        if (cudaMalloc((void**) &sums, 2 * sizeof(double)) != cudaSuccess)
        {
            *status = SDP_ERR_MEM_ALLOC_FAILURE;
            SDP_LOG_CRITICAL("GPU memory allocation failure");
            (void) cudaGetLastError();
        }
    }
    if (!*status)
    {
        const int64_t num_rows = dims[0] * dims[1];
        int rc;
        if (briggs)
        {
            /* robustness_calc, :142 */
            const double numerator = pow(5.0 * 1 / (pow(10.0, robust_param)),
                    2.0);
            rc = es_weight_briggs(num_rows, (int) dims[2], (int) dims[3],
                    grid_size, (const double*) a.uvw, (const double*) a.freq,
                    max_abs_uv, numerator, a.grid, a.input, a.output, sums,
                    is_double, 0);
        }
        else
        {
            rc = es_weight_uniform(num_rows, (int) dims[2], (int) dims[3],
                    grid_size, (const double*) a.uvw, (const double*) a.freq,
                    max_abs_uv, a.grid, a.input, a.output, is_double, 0);
        }
        if (rc != 0)
        {
            *status = SDP_ERR_RUNTIME;
            SDP_LOG_ERROR("CUDA error: %s",
                    cudaGetErrorString((cudaError_t) rc));
            (void) cudaGetLastError();
        }
    }
    copy_back(&a, weights_grid_uv, output_weight, status);
    if (sums)
    {
        /* the read kernel uses the sums: drained before they are freed */
        if (!a.on_host) (void) cudaStreamSynchronize(0);
        (void) cudaFree(sums);
    }
    release(&a);
}

void sdp_weighting_uniform(const sdp_Mem* uvw, const sdp_Mem* freq_hz,
        double max_abs_uv, sdp_Mem* weight_grid_uv, sdp_Mem* input_weights,
        sdp_Mem* output_weights, sdp_Error* status)
{
    weighting(uvw, freq_hz, max_abs_uv, 0, 0.0, weight_grid_uv, input_weights,
            output_weights, status);
}

void sdp_weighting_briggs(const sdp_Mem* uvw, const sdp_Mem* freq_hz,
        double max_abs_uv, const double robust_param, sdp_Mem* weight_grid_uv,
        sdp_Mem* input_weights, sdp_Mem* output_weights, sdp_Error* status)
{
    weighting(uvw, freq_hz, max_abs_uv, 1, robust_param, weight_grid_uv,
            input_weights, output_weights, status);
}

/* Briggs weighting for callers that cannot allocate (CUDA graph capture,
 * serving loops): no counterpart in the reference; the weight grid is not an
 * argument, it lives with the two sums in a file-scope device array of
 * es_weighting.cu (ES_WT_STATIC_CELLS = 1 Mi cells, 8 MiB per GPU). Device
 * (SDP_MEM_GPU) arrays only -- staging host arrays would allocate. Calls no
 * cudaMalloc / cudaFree and does not synchronise; work is enqueued on the
 * legacy default stream. */
void sdp_weighting_briggs_no_alloc(const sdp_Mem* uvw, const sdp_Mem* freq_hz,
        double max_abs_uv, const double robust_param, int64_t grid_size,
        const sdp_Mem* input_weights, sdp_Mem* output_weights,
        sdp_Error* status)
{
    if (*status) return;
    sdp_MemType weights_type = SDP_MEM_VOID;
    sdp_MemLocation location = sdp_mem_location(output_weights);
    int64_t dims[4] = {0, 0, 0, 0};
    weights_metadata(input_weights, &weights_type, &location, dims, status);
    weights_metadata(output_weights, &weights_type, &location, dims, status);
    if (*status) return;
    if (location != SDP_MEM_GPU)
    {
        *status = SDP_ERR_MEM_LOCATION;
        SDP_LOG_ERROR("sdp_weighting_briggs_no_alloc takes device (SDP_MEM_GPU) "
                "arrays only: staging host arrays would need an allocation");
        return;
    }
    check_uvw(uvw, location, dims[0], dims[1], status);
    check_location(freq_hz, location, "freq_hz", status);
    check_location(input_weights, location, "input_weights", status);
    check_writeable(output_weights, "output_weights", status);
    if (*status) return;
    const int is_double = (weights_type == SDP_MEM_DOUBLE);
    if (sdp_mem_type(uvw) != SDP_MEM_DOUBLE ||
            sdp_mem_type(freq_hz) != SDP_MEM_DOUBLE ||
            (weights_type != SDP_MEM_DOUBLE && weights_type != SDP_MEM_FLOAT) ||
            sdp_mem_type(input_weights) != weights_type)
    {
        *status = SDP_ERR_DATA_TYPE;
        SDP_LOG_ERROR("Unsupported data type(s)");
        return;
    }
    if (sdp_mem_num_dims(freq_hz) != 1 ||
            sdp_mem_shape_dim(freq_hz, 0) != dims[2] ||
            dims[2] > INT32_MAX || dims[3] > INT32_MAX || dims[3] < 1 ||
            !sdp_mem_is_c_contiguous(input_weights) ||
            !sdp_mem_is_c_contiguous(output_weights))
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("freq_hz does not match the weights' [num_times, "
                "num_baselines, num_channels, num_pols], or the weights are "
                "not C contiguous");
        return;
    }
    for (int i = 0; i < 4; ++i)
        if (sdp_mem_shape_dim(input_weights, i) != dims[i])
        {
            *status = SDP_ERR_INVALID_ARGUMENT;
            SDP_LOG_ERROR("input and output weights differ in shape");
            return;
        }
    const int64_t max_cells = es_weight_static_max_cells();
    if (grid_size < 1 || grid_size > max_cells ||
            grid_size * grid_size > max_cells / dims[3])
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("sdp_weighting_briggs_no_alloc: grid_size^2 x num_pols = "
                "%lld x %lld x %lld cells exceeds the static scratch of %lld "
                "cells; use sdp_weighting_briggs with a caller-owned grid",
                (long long) grid_size, (long long) grid_size,
                (long long) dims[3], (long long) max_cells);
        return;
    }
    const double numerator = pow(5.0 * 1 / (pow(10.0, robust_param)), 2.0);
    const int rc = es_weight_briggs_static(dims[0] * dims[1], (int) dims[2],
            (int) dims[3], grid_size, (const double*) sdp_mem_data_const(uvw),
            (const double*) sdp_mem_data_const(freq_hz), max_abs_uv, numerator,
            sdp_mem_data_const(input_weights), sdp_mem_data(output_weights),
            is_double, 0);
    if (rc != 0)
    {
        *status = SDP_ERR_RUNTIME;
        SDP_LOG_ERROR("CUDA error: %s", cudaGetErrorString((cudaError_t) rc));
        (void) cudaGetLastError();
    }
}
