/*
 * Host side of the point-source DFT predict: the reference's argument
 * contract (sdp_dft.cpp:102-147 check_params_v00, :289-344 check_params_v01:
 * its checks in its order with its error codes), the supported type
 * combinations (:162-199, :361-400), then the sm_100a kernels of es_dft.cu.
 * There is no host loop: SDP_MEM_CPU arrays are staged through the current
 * device (the reference's OpenMP branch is restated in oracle/dft_oracle.py).
 */
#include <stdint.h>
#include <string.h>

#include <cuda_runtime_api.h>

#include "es_dft.h"
#include "ska-sdp-func/utility/sdp_logging.h"
#include "ska-sdp-func/visibility/sdp_dft.h"

static void fail(sdp_Error code, const char* what, sdp_Error* status)
{
    if (*status) return;
    *status = code;
    SDP_LOG_ERROR("%s", what);
}

/* sdp_data_model_get_vis_metadata, sdp_data_model_checks.c:123-159 */
static void vis_metadata(const sdp_Mem* vis, sdp_MemLocation* location,
        int64_t dims[4], sdp_Error* status)
{
    if (*status) return;
    if (!sdp_mem_is_complex(vis))
    {
        fail(SDP_ERR_DATA_TYPE, "The visibility array must be complex", status);
        return;
    }
    if (sdp_mem_num_dims(vis) != 4)
    {
        fail(SDP_ERR_INVALID_ARGUMENT, "Expected 'vis' to have 4 dimensions",
                status);
        return;
    }
    *location = sdp_mem_location(vis);
    for (int i = 0; i < 4; ++i) dims[i] = sdp_mem_shape_dim(vis, i);
    if (dims[3] != 4 && dims[3] != 1)
        fail(SDP_ERR_INVALID_ARGUMENT,
                "The number of polarisations should be 4 or 1", status);
}

static void check_contiguous(const sdp_Mem* m, const char* name,
        sdp_Error* status)
{
    if (*status) return;
    if (!sdp_mem_is_c_contiguous(m))
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("Expected '%s' to be C contiguous", name);
    }
}

static void check_writeable(const sdp_Mem* m, const char* name,
        sdp_Error* status)
{
    if (*status) return;
    if (sdp_mem_is_read_only(m))
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("Expected '%s' not to be read-only", name);
    }
}

static void check_location(const sdp_Mem* m, sdp_MemLocation location,
        const char* name, sdp_Error* status)
{
    if (*status) return;
    if (sdp_mem_location(m) != location)
    {
        *status = SDP_ERR_MEM_LOCATION;
        SDP_LOG_ERROR("Expected '%s' to be where 'vis' is", name);
    }
}

static void check_shape(const sdp_Mem* m, int num_dims, const int64_t* shape,
        const char* name, sdp_Error* status)
{
    if (*status) return;
    int same = sdp_mem_num_dims(m) == num_dims;
    for (int i = 0; i < num_dims && same; ++i)
        same = sdp_mem_shape_dim(m, i) == shape[i];
    if (!same)
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("'%s' has not the expected shape", name);
    }
}

/* Device views of the four arrays; host arrays are copied in (vis is
 * write-only: not copied in) and vis copied back at the end. */
typedef struct DftStage
{
    int on_host;
    void* staged[4];
} DftStage;

static size_t mem_bytes(const sdp_Mem* m)
{
    return (size_t) sdp_mem_num_elements(m) *
            (size_t) sdp_mem_type_size(sdp_mem_type(m));
}

static void* stage(DftStage* s, int k, const sdp_Mem* m, int copy_in,
        sdp_Error* status)
{
    if (*status) return 0;
    void* data = (void*) sdp_mem_data_const(m);
    if (!s->on_host) return data;
    const size_t bytes = mem_bytes(m);
    cudaError_t err = cudaMalloc(&s->staged[k], bytes ? bytes : 16);
    if (err != cudaSuccess)
    {
        s->staged[k] = 0;
        *status = SDP_ERR_MEM_ALLOC_FAILURE;
        SDP_LOG_CRITICAL("GPU memory allocation failure: %s",
                cudaGetErrorString(err));
        (void) cudaGetLastError();
        return 0;
    }
    if (copy_in && bytes)
    {
        err = cudaMemcpyAsync(s->staged[k], data, bytes,
                cudaMemcpyHostToDevice, 0);
        if (err != cudaSuccess)
        {
            *status = SDP_ERR_MEM_COPY_FAILURE;
            SDP_LOG_CRITICAL("cudaMemcpy error: %s", cudaGetErrorString(err));
            (void) cudaGetLastError();
        }
    }
    return s->staged[k];
}

static void finish(DftStage* s, sdp_Mem* vis, sdp_Error* status)
{
    if (!s->on_host) return;
    if (!*status && s->staged[3] && mem_bytes(vis))
    {
        cudaError_t err = cudaMemcpyAsync(sdp_mem_data(vis), s->staged[3],
                mem_bytes(vis), cudaMemcpyDeviceToHost, 0);
        if (err == cudaSuccess) err = cudaStreamSynchronize(0);
        if (err != cudaSuccess)
        {
            *status = SDP_ERR_MEM_COPY_FAILURE;
            SDP_LOG_CRITICAL("cudaMemcpy error: %s", cudaGetErrorString(err));
            (void) cudaGetLastError();
        }
    }
    else
    {
        (void) cudaStreamSynchronize(0);
    }
    for (int k = 0; k < 4; ++k)
        if (s->staged[k]) (void) cudaFree(s->staged[k]);
}

/* The type combinations the reference runs (sdp_dft.cpp:162-199): fp64
 * directions, fluxes and coordinates; complex fp64 or fp32 visibilities. */
static int types_supported(const sdp_Mem* source_directions,
        const sdp_Mem* source_fluxes, const sdp_Mem* uvw, const sdp_Mem* vis,
        sdp_Error* status)
{
    if (*status) return 0;
    if (sdp_mem_type(source_directions) != SDP_MEM_DOUBLE ||
            sdp_mem_type(source_fluxes) != SDP_MEM_COMPLEX_DOUBLE ||
            sdp_mem_type(uvw) != SDP_MEM_DOUBLE ||
            (sdp_mem_type(vis) != SDP_MEM_COMPLEX_DOUBLE &&
            sdp_mem_type(vis) != SDP_MEM_COMPLEX_FLOAT))
    {
        *status = SDP_ERR_DATA_TYPE;
        SDP_LOG_ERROR("Unsupported data type(s)");
        return 0;
    }
    return 1;
}

static void run(int linear, const sdp_Mem* source_directions,
        const sdp_Mem* source_fluxes, const sdp_Mem* uvw,
        double channel_start_hz, double channel_step_hz, sdp_Mem* vis,
        sdp_Error* status)
{
    if (!types_supported(source_directions, source_fluxes, uvw, vis, status))
        return;
    const int64_t num_rows = sdp_mem_shape_dim(vis, 0) *
            sdp_mem_shape_dim(vis, 1);
    const int64_t num_chan = sdp_mem_shape_dim(vis, 2);
    const int64_t num_pols = sdp_mem_shape_dim(vis, 3);
    const int64_t num_components = sdp_mem_shape_dim(source_directions, 0);
    if (num_chan > INT32_MAX || num_components > INT32_MAX)
    {
        fail(SDP_ERR_INVALID_ARGUMENT, "Too many channels or components",
                status);
        return;
    }
    DftStage s;
    memset(&s, 0, sizeof(s));
    s.on_host = (sdp_mem_location(vis) != SDP_MEM_GPU);
    const double* d_dir = (const double*) stage(&s, 0, source_directions, 1,
            status);
    const double* d_flux = (const double*) stage(&s, 1, source_fluxes, 1,
            status);
    const double* d_uvw = (const double*) stage(&s, 2, uvw, 1, status);
    void* d_vis = stage(&s, 3, vis, 0, status);
    if (!*status)
    {
        const int vis_double = sdp_mem_type(vis) == SDP_MEM_COMPLEX_DOUBLE;
        const int rc = linear ?
                es_dft_point_v01((int) num_components, (int) num_pols,
                        (int) num_chan, num_rows, d_dir, d_flux, d_uvw,
                        channel_start_hz, channel_step_hz, d_vis, vis_double, 0) :
                es_dft_point_v00((int) num_components, (int) num_pols,
                        (int) num_chan, num_rows, d_dir, d_flux, d_uvw, d_vis,
                        vis_double, 0);
        if (rc != 0)
        {
            *status = SDP_ERR_RUNTIME;
            SDP_LOG_ERROR("CUDA error: %s",
                    cudaGetErrorString((cudaError_t) rc));
            (void) cudaGetLastError();
        }
    }
    finish(&s, vis, status);
}

void sdp_dft_point_v00(const sdp_Mem* source_directions,
        const sdp_Mem* source_fluxes, const sdp_Mem* uvw_lambda, sdp_Mem* vis,
        sdp_Error* status)
{
    if (*status) return;
    /* check_params_v00, sdp_dft.cpp:102-147 */
    sdp_MemLocation location = SDP_MEM_CPU;
    int64_t dims[4] = {0, 0, 0, 0};
    vis_metadata(vis, &location, dims, status);
    check_contiguous(source_directions, "source_directions", status);
    check_contiguous(source_fluxes, "source_fluxes", status);
    check_contiguous(uvw_lambda, "uvw_lambda", status);
    check_contiguous(vis, "vis", status);
    check_writeable(vis, "vis", status);
    if (*status) return;
    if (sdp_mem_location(source_fluxes) != location ||
            sdp_mem_location(source_directions) != location ||
            sdp_mem_location(uvw_lambda) != location)
    {
        fail(SDP_ERR_MEM_LOCATION, "Memory location mismatch", status);
        return;
    }
    if (!sdp_mem_is_complex(source_fluxes))
    {
        fail(SDP_ERR_DATA_TYPE, "Source flux values must be complex", status);
        return;
    }
    const int64_t num_components = sdp_mem_shape_dim(source_directions, 0);
    const int64_t shape_directions[] = {num_components, 3};
    const int64_t shape_fluxes[] = {num_components, dims[2], dims[3]};
    const int64_t shape_uvw[] = {dims[0], dims[1], dims[2], 3};
    check_shape(source_directions, 2, shape_directions, "source_directions",
            status);
    check_shape(source_fluxes, 3, shape_fluxes, "source_fluxes", status);
    check_shape(uvw_lambda, 4, shape_uvw, "uvw_lambda", status);
    run(0, source_directions, source_fluxes, uvw_lambda, 0.0, 0.0, vis, status);
}

void sdp_dft_point_v01(const sdp_Mem* source_directions,
        const sdp_Mem* source_fluxes, const sdp_Mem* uvw,
        const double channel_start_hz, const double channel_step_hz,
        sdp_Mem* vis, sdp_Error* status)
{
    if (*status) return;
    /* check_params_v01, sdp_dft.cpp:289-344 */
    sdp_MemLocation location = SDP_MEM_CPU;
    int64_t dims[4] = {0, 0, 0, 0};
    vis_metadata(vis, &location, dims, status);
    check_writeable(vis, "vis", status);
    check_contiguous(vis, "vis", status);
    /* sdp_data_model_check_uvw_at, sdp_data_model_checks.c:7-43 */
    if (!*status && (!sdp_mem_is_floating_point(uvw) || sdp_mem_is_complex(uvw)))
        fail(SDP_ERR_DATA_TYPE, "The uvw array must be real-valued", status);
    const int64_t shape_uvw[] = {dims[0], dims[1], 3};
    check_shape(uvw, 3, shape_uvw, "uvw", status);
    check_location(uvw, location, "uvw", status);
    check_contiguous(uvw, "uvw", status);
    check_location(source_directions, location, "source_directions", status);
    check_contiguous(source_directions, "source_directions", status);
    if (*status) return;
    const int64_t num_components = sdp_mem_shape_dim(source_directions, 0);
    const int64_t shape_directions[] = {num_components, 3};
    check_shape(source_directions, 2, shape_directions, "source_directions",
            status);
    if (*status) return;
    if (!sdp_mem_is_complex(source_fluxes))
    {
        fail(SDP_ERR_DATA_TYPE, "Source flux values must be complex", status);
        return;
    }
    check_location(source_fluxes, location, "source_fluxes", status);
    check_contiguous(source_fluxes, "source_fluxes", status);
    const int64_t shape_fluxes[] = {num_components, dims[2], dims[3]};
    check_shape(source_fluxes, 3, shape_fluxes, "source_fluxes", status);
    run(1, source_directions, source_fluxes, uvw, channel_start_hz,
            channel_step_hz, vis, status);
}
