/*
 * Host side of the Hogbom CLEAN: the reference's argument contract
 * (sdp_hogbom_clean.cpp:733-853: its checks in its order with its error
 * codes), the CLEAN beam coefficients (.cpp:46-63), then the sm_100a path of
 * es_clean.cu. There is no host loop: SDP_MEM_CPU images are staged through
 * the current device (the reference's host loop is restated in
 * oracle/clean_oracle.py). `cbeam_details` is read by the host, as in the
 * reference's GPU path (.cpp:299-301): it is a host array even when the
 * images are on the device (a device array is copied back first).
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

#include <cuda_runtime_api.h>

#include "es_clean.h"
#include "ska-sdp-func/clean/sdp_hogbom_clean.h"
#include "ska-sdp-func/utility/sdp_logging.h"

static void fail(sdp_Error code, const char* what, sdp_Error* status)
{
    *status = code;
    SDP_LOG_ERROR("%s", what);
}

/* report and leave the (void) entry point */
#define FAIL(code, what) do { fail(code, what, status); return; } while (0)

static void cuda_failed(cudaError_t err, sdp_Error code, sdp_Error* status)
{
    *status = code;
    SDP_LOG_CRITICAL("CUDA error: %s", cudaGetErrorString(err));
    (void) cudaGetLastError();
}

void sdp_hogbom_clean(const sdp_Mem* dirty_img, const sdp_Mem* psf,
        const sdp_Mem* cbeam_details, const double loop_gain,
        const double threshold, const int cycle_limit, sdp_Mem* clean_model,
        sdp_Mem* residual, sdp_Mem* skymodel, sdp_Error* status)
{
    if (*status) return;
    const int64_t n = sdp_mem_shape_dim(dirty_img, 0);
    const int64_t n1 = sdp_mem_shape_dim(dirty_img, 1);
    const int64_t psf_dim = sdp_mem_shape_dim(psf, 0);
    const int64_t psf_dim1 = sdp_mem_shape_dim(psf, 1);
    const sdp_MemLocation location = sdp_mem_location(dirty_img);
    const sdp_MemType type = sdp_mem_type(dirty_img);

    /* .cpp:760-853, in the reference's order */
    if (sdp_mem_is_read_only(skymodel))
        FAIL(SDP_ERR_RUNTIME, "Output is not writable");
    if (sdp_mem_location(dirty_img) != sdp_mem_location(skymodel))
        FAIL(SDP_ERR_MEM_LOCATION, "Memory location mismatch");
    if (type != sdp_mem_type(psf))
        FAIL(SDP_ERR_DATA_TYPE, "The Dirty image and PSF must be of "
                "the same data type");
    if (type != sdp_mem_type(skymodel))
        FAIL(SDP_ERR_DATA_TYPE, "The input and output must be of the "
                "same data type");
    if (sdp_mem_shape_dim(clean_model, 0) != n)
        FAIL(SDP_ERR_RUNTIME, "The CLEAN model and the dirty image "
                "must be the same size");
    if (sdp_mem_shape_dim(residual, 0) != n)
        FAIL(SDP_ERR_RUNTIME, "The residual image and the dirty image "
                "must be the same size");
    if (sdp_mem_shape_dim(skymodel, 0) != n)
        FAIL(SDP_ERR_RUNTIME, "The skymodel image and the dirty image "
                "must be the same size");
    if (sdp_mem_shape_dim(cbeam_details, 0) != 4)
        FAIL(SDP_ERR_RUNTIME, "The array describing the CLEAN beam "
                "must include BMAJ, BMIN, THETA and SIZE");
    if (n != n1)
        FAIL(SDP_ERR_RUNTIME, "Dirty image array must be square shaped");
    if (psf_dim != psf_dim1)
        FAIL(SDP_ERR_RUNTIME, "PSF array must be square shaped");
    if (psf_dim != n * 2)
        FAIL(SDP_ERR_RUNTIME, "PSF array dimensions must be double the "
                "Dirty image array dimensions");
    if (cycle_limit < 1)
        FAIL(SDP_ERR_RUNTIME, "Number of cycles to perform must be > 0");
    if (loop_gain <= 0)
        FAIL(SDP_ERR_RUNTIME, "Loop gain must be > 0");
    if (type != SDP_MEM_DOUBLE && type != SDP_MEM_FLOAT)
        FAIL(SDP_ERR_DATA_TYPE, "Unsupported data type");
    /* what the kernels index without further checks */
    const sdp_Mem* images[] = {dirty_img, psf, clean_model, residual, skymodel};
    for (int k = 0; k < 5; ++k)
    {
        const int64_t want = (k == 1) ? 2 * n : n;
        if (sdp_mem_num_dims(images[k]) != 2 ||
                sdp_mem_shape_dim(images[k], 0) != want ||
                sdp_mem_shape_dim(images[k], 1) != want ||
                !sdp_mem_is_c_contiguous(images[k]) ||
                sdp_mem_type(images[k]) != type)
            FAIL(SDP_ERR_RUNTIME, "The images must be square, C "
                    "contiguous 2-D arrays of one data type");
        if (sdp_mem_location(images[k]) != location)
            FAIL(SDP_ERR_MEM_LOCATION, "Memory location mismatch");
    }
    if (sdp_mem_is_read_only(clean_model) || sdp_mem_is_read_only(residual))
        FAIL(SDP_ERR_RUNTIME, "Output is not writable");
    if (n < 1 || n > 46340 || sdp_mem_type(cbeam_details) != type)
        FAIL(SDP_ERR_RUNTIME, "Image size out of range, or the CLEAN "
                "beam description is not of the images' data type");

    /* CLEAN beam: [sigma_x, sigma_y, theta (degrees), size], .cpp:46-63 */
    double details[4] = {0.0, 0.0, 0.0, 0.0};
    {
        union { double d[4]; float f[4]; } raw;
        const size_t bytes = 4 * (type == SDP_MEM_DOUBLE ? 8 : 4);
        if (sdp_mem_location(cbeam_details) == SDP_MEM_GPU)
        {
            cudaError_t err = cudaMemcpy(&raw, sdp_mem_data_const(cbeam_details),
                    bytes, cudaMemcpyDeviceToHost);
            if (err != cudaSuccess)
            {
                cuda_failed(err, SDP_ERR_MEM_COPY_FAILURE, status);
                return;
            }
        }
        else
        {
            memcpy(&raw, sdp_mem_data_const(cbeam_details), bytes);
        }
        for (int k = 0; k < 4; ++k)
            details[k] = type == SDP_MEM_DOUBLE ? raw.d[k] : (double) raw.f[k];
    }
    const int64_t cbeam_dim = (int64_t) details[3];
    if (cbeam_dim < 1 || cbeam_dim > 32768 || details[0] == 0.0 ||
            details[1] == 0.0)
        FAIL(SDP_ERR_RUNTIME, "The CLEAN beam needs non-zero widths "
                "and a size of at least 1");
    const double theta = (M_PI / 180) * details[2];
    const double sx2 = details[0] * details[0], sy2 = details[1] * details[1];
    const double beam_a = pow(cos(theta), 2) / (2 * sx2) +
            pow(sin(theta), 2) / (2 * sy2);
    const double beam_b = sin(2 * theta) / (4 * sx2) -
            sin(2 * theta) / (4 * sy2);
    const double beam_c = pow(sin(theta), 2) / (2 * sx2) +
            pow(cos(theta), 2) / (2 * sy2);

    /* device views: host images are staged (psf in; model, residual and
     * skymodel out) */
    const int on_host = (location != SDP_MEM_GPU);
    const int is_double = (type == SDP_MEM_DOUBLE);
    const size_t image_bytes = (size_t) (n * n) * (is_double ? 8 : 4);
    const size_t scratch_bytes = es_hogbom_scratch_bytes(n, cycle_limit,
            cbeam_dim, is_double);
    char* block = 0;
    cudaError_t err = cudaMalloc((void**) &block, scratch_bytes +
            (on_host ? 7 * image_bytes + 1024 : 0));
    if (err != cudaSuccess)
    {
        cuda_failed(err, SDP_ERR_MEM_ALLOC_FAILURE, status);
        return;
    }
    char* d_psf = (char*) sdp_mem_data_const(psf);
    char* d_model = (char*) sdp_mem_data(clean_model);
    char* d_residual = (char*) sdp_mem_data(residual);
    char* d_sky = (char*) sdp_mem_data(skymodel);
    const enum cudaMemcpyKind to_dev = on_host ? cudaMemcpyHostToDevice :
            cudaMemcpyDeviceToDevice;
    if (on_host)
    {
        char* at = block + ((scratch_bytes + 255) & ~(size_t) 255);
        d_psf = at;
        d_model = at + 4 * image_bytes;
        d_residual = at + 5 * image_bytes;
        d_sky = at + 6 * image_bytes;
        err = cudaMemcpyAsync(d_psf, sdp_mem_data_const(psf), 4 * image_bytes,
                cudaMemcpyHostToDevice, 0);
    }
    /* residual starts as the dirty image, the model from zero (.cpp:855-858,
     * :172) */
    if (err == cudaSuccess)
        err = cudaMemcpyAsync(d_residual, sdp_mem_data_const(dirty_img),
                image_bytes, to_dev, 0);
    if (err == cudaSuccess) err = cudaMemsetAsync(d_model, 0, image_bytes, 0);
    if (err == cudaSuccess) err = cudaMemsetAsync(block, 0, 256, 0);
    if (err != cudaSuccess)
    {
        cuda_failed(err, SDP_ERR_MEM_COPY_FAILURE, status);
    }
    else
    {
        const int rc = es_hogbom_clean(n, d_psf, loop_gain, threshold,
                cycle_limit, beam_a, beam_b, beam_c, cbeam_dim, d_model,
                d_residual, d_sky, block, is_double, 0);
        if (rc != 0) cuda_failed((cudaError_t) rc, SDP_ERR_RUNTIME, status);
    }
    if (!*status && on_host)
    {
        err = cudaMemcpyAsync(sdp_mem_data(clean_model), d_model, image_bytes,
                cudaMemcpyDeviceToHost, 0);
        if (err == cudaSuccess)
            err = cudaMemcpyAsync(sdp_mem_data(residual), d_residual,
                    image_bytes, cudaMemcpyDeviceToHost, 0);
        if (err == cudaSuccess)
            err = cudaMemcpyAsync(sdp_mem_data(skymodel), d_sky, image_bytes,
                    cudaMemcpyDeviceToHost, 0);
        if (err != cudaSuccess)
            cuda_failed(err, SDP_ERR_MEM_COPY_FAILURE, status);
    }
    /* the scratch is freed here: the work is drained first */
    err = cudaStreamSynchronize(0);
    if (err != cudaSuccess && !*status)
        cuda_failed(err, SDP_ERR_RUNTIME, status);
    (void) cudaFree(block);
}
