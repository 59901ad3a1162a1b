/*
 * sdp_GridderUvwEsFft for B200: C host side (plan object, argument contract,
 * staging, w-plane batch loop, cuFFT) over the thin C-ABI kernel launchers of
 * es_device.h. Replaces the reference's
 *   src/ska-sdp-func/grid_data/sdp_gridder_uvw_es_fft.cpp
 * behind the same four entry points.
 *
 * Schedule per call (reference: es_fft.cpp:578-698 / :832-952, one plane at a
 * time, every visibility re-scanned per plane, memset + atomics):
 *   presort once  -> for each BATCH of B resident w-planes:
 *     grid:   scatter (writes every cell once) -> batched cuFFT inverse ->
 *             w-screen multiply + accumulate into the image (one RMW per batch)
 *     degrid: reverse w-screen (writes every cell once) -> batched cuFFT
 *             forward -> gather into per-record accumulators
 *   -> grid correction / un-permute.
 * There is no CPU fallback: without a CUDA device plan creation fails with
 * SDP_ERR_MEM_ALLOC_FAILURE exactly like the reference does (es_fft.cpp:491-518).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <cuda_runtime_api.h>
#include <cufft.h>

#include "es_device.h"
#include "es_plan.h"
#include "ska-sdp-func/grid_data/sdp_gridder_uvw_es_fft_b200.h"

#define ES_MAX_PLANE_BATCH 24
#define ES_NUM_FFT_PLANS 2

typedef struct EsBuffer
{
    void* ptr;
    size_t bytes;
} EsBuffer;

struct sdp_GridderUvwEsFft
{
    sdp_EsPlanInfo info;
    EsGeom geom;
    double pixsize_x_rad, pixsize_y_rad;
    cudaStream_t stream;
    cudaStream_t copy_stream; /* host staging of the array needed last */
    cudaEvent_t copy_ev;
    int32_t have_copy_stream;

    /* multi-GPU extension */
    int32_t plane_begin, plane_end, apply_correction;

    /* device tables (always double) */
    double* d_quad_kernel;
    double* d_quad_nodes;
    double* d_quad_weights;
    double* d_conv_corr;

    /* resident w-planes and FFT */
    int32_t plane_batch;
    EsBuffer planes;
    cufftHandle fft[ES_NUM_FFT_PLANS];
    int32_t fft_batch[ES_NUM_FFT_PLANS];
    EsBuffer fft_work;

    /* pruned in-shared-memory FFT with fused w-screen (es_fft.cu); NULL:
     * cuFFT 2-D plans + separate w-screen kernels */
    EsFft* pruned_fft;
    EsBuffer strip;   /* plane_batch x N x G complex */
    EsBuffer image_t; /* N x N reals: transposed image accumulator */

    /* presort work space (grow-only) */
    EsBuffer counts, offsets, keys_a, keys_b, vals_a, vals_b, records;
    EsBuffer kweights;    /* fp64 gridding: ES weights per record (es_kweight_count) */
    int have_kweights;    /* of the current call */
    EsBuffer tile_start, sort_temp, acc;

    /* load-balanced work units of the current plane batch */
    EsBuffer units, unit_counters, split_done, partials;
    EsBuffer occ;     /* plane_batch x es_occupancy_words: occupied tiles / tile rows */
    EsBuffer occ_totals; /* 2 x u64: occupied (tile, plane) / (tile row, plane) pairs of the call */

    /* staging for host arrays */
    EsBuffer st_uvw, st_freq, st_vis, st_weight, st_dirty;

    sdp_EsCallStats stats;
    int32_t timing;
    cudaEvent_t ev[8];
    int32_t have_events;
    int32_t stats_pending; /* 1 grid / 2 degrid timings not yet collected */
    cudaEvent_t* stage_ev; /* 4 per plane batch when timing is enabled */
    int32_t stage_cap, stage_used;
    int32_t* batch_first;  /* first plane / plane count of every batch of the */
    int32_t* batch_count;  /* last call (timing enabled), stage_ev[4 k ..] */
    int32_t batch_cap, batch_used;
};

/* ------------------------------------------------------------ utilities */

static int cuda_failed(cudaError_t err, const char* what, sdp_Error* status)
{
    if (err == cudaSuccess) return 0;
    if (!*status) *status = SDP_ERR_RUNTIME;
    SDP_LOG_ERROR("%s failed: %s", what, cudaGetErrorString(err));
    (void) cudaGetLastError();
    return 1;
}

#define ES_CUDA(call) \
    do { if (cuda_failed((cudaError_t) (call), #call, status)) return; } while (0)

static void buffer_free(EsBuffer* b)
{
    if (b->ptr) (void) cudaFree(b->ptr);
    b->ptr = 0;
    b->bytes = 0;
}

/* Grow-only device buffer. */
static void buffer_reserve(EsBuffer* b, size_t bytes, sdp_Error* status)
{
    if (*status || bytes <= b->bytes) return;
    buffer_free(b);
    const cudaError_t err = cudaMalloc(&b->ptr, bytes);
    if (err != cudaSuccess || !b->ptr)
    {
        *status = SDP_ERR_MEM_ALLOC_FAILURE;
        SDP_LOG_CRITICAL("GPU memory allocation failure: %s "
                "(requested %zu bytes)", cudaGetErrorString(err), bytes);
        (void) cudaGetLastError();
        b->ptr = 0;
        return;
    }
    b->bytes = bytes;
}

static size_t real_bytes(const sdp_GridderUvwEsFft* plan)
{
    return plan->info.is_double ? 8 : 4;
}

/* ------------------------------------------------- argument validation */

/* Same checks, same order, same codes as es_fft.cpp:72-243. */
static void check_buffers(const sdp_Mem* uvw, const sdp_Mem* freq_hz,
        const sdp_Mem* vis, const sdp_Mem* weight, const sdp_Mem* dirty_image,
        int do_degridding, sdp_Error* status)
{
    const sdp_MemLocation where = sdp_mem_location(uvw);
    if (where != sdp_mem_location(freq_hz) || where != sdp_mem_location(vis) ||
            where != sdp_mem_location(weight) ||
            where != sdp_mem_location(dirty_image))
    {
        *status = SDP_ERR_MEM_LOCATION;
        SDP_LOG_ERROR("Memory location mismatch.");
        return;
    }
    const struct { const sdp_Mem* mem; int want_complex; const char* msg; }
    kinds[] = {
        {uvw, 0, "uvw values must be real."},
        {freq_hz, 0, "Frequency values must be real."},
        {vis, 1, "Visibility values must be complex."},
        {weight, 0, "Weight values must be real."},
        {dirty_image, 0, "Dirty image must be real"}
    };
    for (size_t k = 0; k < sizeof(kinds) / sizeof(kinds[0]); ++k)
    {
        if ((sdp_mem_is_complex(kinds[k].mem) != 0) != kinds[k].want_complex)
        {
            *status = SDP_ERR_DATA_TYPE;
            SDP_LOG_ERROR("%s", kinds[k].msg);
            return;
        }
    }
    const int64_t rows = sdp_mem_shape_dim(vis, 0);
    const int64_t chans = sdp_mem_shape_dim(vis, 1);
    if (sdp_mem_shape_dim(uvw, 0) != rows)
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("The number of rows in uvw and vis must match.");
        SDP_LOG_ERROR("uvw is %lld by %lld, vis is %lld by %lld",
                (long long) sdp_mem_shape_dim(uvw, 0),
                (long long) sdp_mem_shape_dim(uvw, 1),
                (long long) rows, (long long) chans);
        return;
    }
    if (sdp_mem_shape_dim(uvw, 1) != 3)
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("uvw must be N x 3 (it is %lld by %lld).",
                (long long) sdp_mem_shape_dim(uvw, 0),
                (long long) sdp_mem_shape_dim(uvw, 1));
        return;
    }
    if (sdp_mem_shape_dim(freq_hz, 0) != chans)
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("The number of channels in vis and freq_hz must match "
                "(freq_hz has %lld, vis is %lld by %lld).",
                (long long) sdp_mem_shape_dim(freq_hz, 0),
                (long long) rows, (long long) chans);
        return;
    }
    if (sdp_mem_shape_dim(weight, 0) != rows ||
            sdp_mem_shape_dim(weight, 1) != chans)
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("weight and vis must be the same size "
                "(weight is %lld by %lld, vis is %lld by %lld).",
                (long long) sdp_mem_shape_dim(weight, 0),
                (long long) sdp_mem_shape_dim(weight, 1),
                (long long) rows, (long long) chans);
        return;
    }
    if (sdp_mem_shape_dim(dirty_image, 0) != sdp_mem_shape_dim(dirty_image, 1))
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("Dirty image must be square (it is %lld by %lld).",
                (long long) sdp_mem_shape_dim(dirty_image, 0),
                (long long) sdp_mem_shape_dim(dirty_image, 1));
        return;
    }
    const int dbl = sdp_mem_type(uvw) == SDP_MEM_DOUBLE;
    if (!dbl && sdp_mem_type(uvw) != SDP_MEM_FLOAT)
    {
        *status = SDP_ERR_DATA_TYPE;
        SDP_LOG_ERROR("uvw must be single or double precision.");
        return;
    }
    const sdp_MemType want_real = dbl ? SDP_MEM_DOUBLE : SDP_MEM_FLOAT;
    const sdp_MemType want_cplx = dbl ?
            SDP_MEM_COMPLEX_DOUBLE : SDP_MEM_COMPLEX_FLOAT;
    if (sdp_mem_type(freq_hz) != want_real || sdp_mem_type(vis) != want_cplx ||
            sdp_mem_type(weight) != want_real ||
            sdp_mem_type(dirty_image) != want_real)
    {
        *status = SDP_ERR_DATA_TYPE;
        SDP_LOG_ERROR("All buffers must be the same precision.");
        return;
    }
    if (!sdp_mem_is_c_contiguous(uvw) || !sdp_mem_is_c_contiguous(freq_hz) ||
            !sdp_mem_is_c_contiguous(vis) ||
            !sdp_mem_is_c_contiguous(weight) ||
            !sdp_mem_is_c_contiguous(dirty_image))
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("All input arrays must be C contiguous");
        return;
    }
    if (do_degridding ? sdp_mem_is_read_only(vis) :
            sdp_mem_is_read_only(dirty_image))
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR(do_degridding ? "Visibility data must be writable." :
                "Dirty image must be writable.");
        return;
    }
}

/* es_fft.cpp:246-262 */
static void check_parameters(double pixsize_x_rad, double pixsize_y_rad,
        sdp_Error* status)
{
    if (pixsize_x_rad != pixsize_y_rad)
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("Only square images supported, so pixsize_x_rad and "
                "pixsize_y_rad must be equal (%.12e vs %.12e).",
                pixsize_x_rad, pixsize_y_rad);
    }
}

/* Shapes are frozen at plan creation in the reference (es_fft.cpp:310-313,
 * 560) and later calls are trusted; here a mismatch would run past the ends
 * of the work buffers, so it is rejected. */
static void check_against_plan(const sdp_GridderUvwEsFft* plan,
        const sdp_Mem* vis, const sdp_Mem* dirty_image, sdp_Error* status)
{
    if (*status) return;
    if (sdp_mem_shape_dim(vis, 0) != plan->info.num_rows ||
            sdp_mem_shape_dim(vis, 1) != plan->info.num_chan ||
            sdp_mem_shape_dim(dirty_image, 0) != plan->info.image_size ||
            (sdp_mem_is_double(vis) != 0) != (plan->info.is_double != 0))
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("Array shapes or precision differ from those the plan "
                "was created for.");
    }
}

/* ---------------------------------------------------------------- plan */

void sdp_gridder_uvw_es_fft_plan_info(int32_t num_rows, int32_t num_chan,
        int32_t image_size, int32_t is_double, double pixel_size_rad,
        double epsilon, double min_abs_w, double max_abs_w,
        int32_t do_w_stacking, sdp_EsPlanInfo* info, double* quad_kernel,
        double* quad_nodes, double* quad_weights, double* conv_corr,
        sdp_Error* status)
{
    if (*status || !info) return;
    es_plan_scalars(num_rows, num_chan, image_size, is_double, pixel_size_rad,
            epsilon, min_abs_w, max_abs_w, do_w_stacking, info);
    double qk[ES_QUADRATURE_BOUND], qn[ES_QUADRATURE_BOUND];
    double qw[ES_QUADRATURE_BOUND];
    info->conv_corr_norm_factor = es_correction_tables(image_size,
            info->grid_size, info->support, info->beta, qk, qn, qw, conv_corr);
    if (quad_kernel) memcpy(quad_kernel, qk, sizeof(qk));
    if (quad_nodes) memcpy(quad_nodes, qn, sizeof(qn));
    if (quad_weights) memcpy(quad_weights, qw, sizeof(qw));
}

void sdp_gridder_uvw_es_fft_get_info(const sdp_GridderUvwEsFft* plan,
        sdp_EsPlanInfo* info)
{
    if (plan && info) *info = plan->info;
}

void sdp_gridder_uvw_es_fft_free_plan(sdp_GridderUvwEsFft* plan)
{
    if (!plan) return;
    for (int k = 0; k < ES_NUM_FFT_PLANS; ++k)
        if (plan->fft_batch[k] > 0) (void) cufftDestroy(plan->fft[k]);
    es_fft_destroy(plan->pruned_fft);
    if (plan->d_quad_kernel) (void) cudaFree(plan->d_quad_kernel);
    if (plan->d_quad_nodes) (void) cudaFree(plan->d_quad_nodes);
    if (plan->d_quad_weights) (void) cudaFree(plan->d_quad_weights);
    if (plan->d_conv_corr) (void) cudaFree(plan->d_conv_corr);
    EsBuffer* all[] = {
        &plan->planes, &plan->fft_work, &plan->strip, &plan->image_t,
        &plan->counts, &plan->offsets,
        &plan->keys_a, &plan->keys_b, &plan->vals_a, &plan->vals_b,
        &plan->records, &plan->kweights, &plan->tile_start, &plan->sort_temp, &plan->acc,
        &plan->units, &plan->unit_counters, &plan->split_done, &plan->partials,
        &plan->occ, &plan->occ_totals,
        &plan->st_uvw, &plan->st_freq, &plan->st_vis, &plan->st_weight,
        &plan->st_dirty
    };
    for (size_t k = 0; k < sizeof(all) / sizeof(all[0]); ++k)
        buffer_free(all[k]);
    if (plan->have_copy_stream)
    {
        (void) cudaEventDestroy(plan->copy_ev);
        (void) cudaStreamDestroy(plan->copy_stream);
    }
    if (plan->have_events)
        for (int k = 0; k < 8; ++k) (void) cudaEventDestroy(plan->ev[k]);
    for (int k = 0; k < plan->stage_cap; ++k)
        (void) cudaEventDestroy(plan->stage_ev[k]);
    free(plan->stage_ev);
    free(plan->batch_first);
    free(plan->batch_count);
    free(plan);
}

static double* upload_table(const double* host, size_t count,
        sdp_Error* status)
{
    if (*status) return 0;
    double* dev = 0;
    cudaError_t err = cudaMalloc((void**) &dev, count * sizeof(double));
    if (err != cudaSuccess || !dev)
    {
        *status = SDP_ERR_MEM_ALLOC_FAILURE;
        SDP_LOG_CRITICAL("GPU memory allocation failure: %s "
                "(requested %zu bytes)", cudaGetErrorString(err),
                count * sizeof(double));
        (void) cudaGetLastError();
        return 0;
    }
    err = cudaMemcpy(dev, host, count * sizeof(double),
            cudaMemcpyHostToDevice);
    if (err != cudaSuccess)
    {
        *status = SDP_ERR_MEM_COPY_FAILURE;
        SDP_LOG_CRITICAL("cudaMemcpy error: %s", cudaGetErrorString(err));
        (void) cudaGetLastError();
    }
    return dev;
}

/* How many w-planes to keep resident: as many as fit comfortably (planes +
 * an equally large cuFFT work area), at most ES_MAX_PLANE_BATCH. */
static int choose_plane_batch(const sdp_GridderUvwEsFft* plan)
{
    const char* env = getenv("SKA_SDP_FUNC_ES_PLANE_BATCH");
    int limit = ES_MAX_PLANE_BATCH;
    if (env && atoi(env) > 0) limit = atoi(env);
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess)
    {
        (void) cudaGetLastError();
        return 1;
    }
    const double plane_bytes = 2.0 * (double) real_bytes(plan) *
            (double) plan->info.grid_size * (double) plan->info.grid_size;
    /* per resident plane: cuFFT mode = plane + an equally large work area;
     * pruned mode = plane + its N x G strip */
    const double per_plane = plan->pruned_fft ?
            plane_bytes * (1.0 + (double) plan->info.image_size /
                    (double) plan->info.grid_size) : 2.0 * plane_bytes;
    /* the presort scratch, records and staging copies need room too */
    int fit = (int) (0.55 * (double) free_b / per_plane);
    if (fit < 1) fit = 1;
    int batch = plan->info.num_total_w_grids;
    if (batch > limit) batch = limit;
    if (batch > fit) batch = fit;
    return batch < 1 ? 1 : batch;
}

sdp_GridderUvwEsFft* sdp_gridder_uvw_es_fft_create_plan(
        const sdp_Mem* uvw, const sdp_Mem* freq_hz, const sdp_Mem* vis,
        const sdp_Mem* weight, const sdp_Mem* dirty_image,
        const double pixel_size_x_rad, const double pixel_size_y_rad,
        const double epsilon, const double min_abs_w, const double max_abs_w,
        const int do_w_stacking, sdp_Error* status)
{
    if (*status) return 0;
    check_parameters(pixel_size_x_rad, pixel_size_y_rad, status);
    if (*status) return 0;
    check_buffers(uvw, freq_hz, vis, weight, dirty_image, 0, status);
    if (*status) return 0;
    const int64_t rows = sdp_mem_shape_dim(vis, 0);
    const int64_t chans = sdp_mem_shape_dim(vis, 1);
    if (rows * chans >= (1LL << 30))
    {
        /* The reference's limit is its 32-bit i_vis, 2^31 (kernels.cu:150-
         * 152). Here a visibility expands into up to 4 tile records whose
         * offsets and total are kept in 32 bits, and the sort values carry
         * the entry number in their two top bits: 2^30. */
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("Too many visibilities for one plan (rows x channels "
                "must be below 2^30).");
        return 0;
    }

    sdp_GridderUvwEsFft* plan =
            (sdp_GridderUvwEsFft*) calloc(1, sizeof(sdp_GridderUvwEsFft));
    if (!plan)
    {
        *status = SDP_ERR_MEM_ALLOC_FAILURE;
        return 0;
    }
    plan->pixsize_x_rad = pixel_size_x_rad;
    plan->pixsize_y_rad = pixel_size_y_rad;
    const int is_double = sdp_mem_is_double(vis) ? 1 : 0;
    const int image_size = (int) sdp_mem_shape_dim(dirty_image, 0);
    es_plan_scalars((int) rows, (int) chans, image_size, is_double,
            pixel_size_x_rad, epsilon, min_abs_w, max_abs_w, do_w_stacking,
            &plan->info);
    double qk[ES_QUADRATURE_BOUND], qn[ES_QUADRATURE_BOUND];
    double qw[ES_QUADRATURE_BOUND];
    double* cc = (double*) calloc((size_t) image_size / 2 + 1, sizeof(double));
    if (!cc)
    {
        *status = SDP_ERR_MEM_ALLOC_FAILURE;
        free(plan);
        return 0;
    }
    plan->info.conv_corr_norm_factor = es_correction_tables(image_size,
            plan->info.grid_size, plan->info.support, plan->info.beta,
            qk, qn, qw, cc);
    if (plan->info.grid_size < ES_TILE / 4 ||
            plan->info.support + 1 > 17 || plan->info.num_total_w_grids < 1 ||
            !(plan->info.w_scale > 0.0) || !isfinite(plan->info.w_scale))
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("Cannot build a w-plane lattice for these parameters "
                "(grid %d, support %d, planes %d, w_scale %g).",
                plan->info.grid_size, plan->info.support,
                plan->info.num_total_w_grids, plan->info.w_scale);
        free(cc);
        free(plan);
        return 0;
    }

    EsGeom* g = &plan->geom;
    g->num_rows = plan->info.num_rows;
    g->num_chan = plan->info.num_chan;
    g->image_size = image_size;
    g->grid_size = plan->info.grid_size;
    g->support = plan->info.support;
    g->num_planes = plan->info.num_total_w_grids;
    g->do_wstacking = plan->info.do_wstacking;
    g->is_double = is_double;
    g->ntile = (g->grid_size + ES_TILE - 1) / ES_TILE;
    g->beta = plan->info.beta;
    g->pixel_size = plan->info.pixel_size;
    g->uv_scale = plan->info.uv_scale;
    g->w_scale = plan->info.w_scale;
    g->inv_w_scale = plan->info.inv_w_scale;
    g->min_plane_w = plan->info.min_plane_w;
    g->conv_corr_norm_factor = plan->info.conv_corr_norm_factor;
    plan->plane_begin = 0;
    plan->plane_end = g->num_planes;
    plan->apply_correction = 1;
    plan->stream = 0; /* legacy default stream, like the reference */

    /* Device side: tables and the resident w-planes (the reference allocates
     * its single plane here too, es_fft.cpp:509-518). */
    plan->d_quad_kernel = upload_table(qk, ES_QUADRATURE_BOUND, status);
    plan->d_quad_nodes = upload_table(qn, ES_QUADRATURE_BOUND, status);
    plan->d_quad_weights = upload_table(qw, ES_QUADRATURE_BOUND, status);
    plan->d_conv_corr = upload_table(cc, (size_t) image_size / 2 + 1, status);
    free(cc);
    if (!*status)
    {
        const int rc = es_fft_create(g->grid_size, g->image_size, is_double,
                &plan->pruned_fft);
        int rc2 = 0;
        if (rc == 0 && plan->pruned_fft && g->do_wstacking)
        {
            /* w-screen table of the fused column passes */
            const double w_lo = fabs(g->min_plane_w);
            const double w_hi = fabs(g->min_plane_w +
                    (double) (g->num_planes - 1) * g->inv_w_scale);
            rc2 = es_fft_set_screen(plan->pruned_fft, g->pixel_size,
                    w_lo > w_hi ? w_lo : w_hi);
        }
        if (rc > 0 || rc2 > 0)
        {
            *status = SDP_ERR_MEM_ALLOC_FAILURE;
            SDP_LOG_CRITICAL("GPU memory allocation failure: %s",
                    cudaGetErrorString((cudaError_t) (rc > 0 ? rc : rc2)));
        }
    }
    if (!*status)
    {
        plan->plane_batch = choose_plane_batch(plan);
        const size_t plane_bytes = 2 * real_bytes(plan) *
                (size_t) g->grid_size * (size_t) g->grid_size;
        while (plan->plane_batch >= 1)
        {
            sdp_Error try_status = SDP_SUCCESS;
            buffer_reserve(&plan->planes,
                    plane_bytes * (size_t) plan->plane_batch, &try_status);
            if (plan->pruned_fft)
            {
                buffer_reserve(&plan->strip, es_fft_strip_bytes(
                        plan->pruned_fft, plan->plane_batch), &try_status);
                buffer_reserve(&plan->image_t, real_bytes(plan) *
                        (size_t) image_size * (size_t) image_size,
                        &try_status);
                buffer_reserve(&plan->occ, (size_t) plan->plane_batch *
                        es_occupancy_words(g) * sizeof(uint32_t), &try_status);
                buffer_reserve(&plan->occ_totals, 2 * sizeof(uint64_t),
                        &try_status);
            }
            if (!try_status) break;
            buffer_free(&plan->planes);
            buffer_free(&plan->strip);
            if (plan->plane_batch == 1) { *status = try_status; break; }
            plan->plane_batch /= 2;
        }
    }
    if (*status)
    {
        sdp_gridder_uvw_es_fft_free_plan(plan);
        return 0;
    }
    plan->stats.plane_batch = plan->plane_batch;
    return plan;
}

void sdp_gridder_uvw_es_fft_set_plane_range(sdp_GridderUvwEsFft* plan,
        int32_t plane_begin, int32_t plane_end, int32_t apply_correction,
        sdp_Error* status)
{
    if (*status || !plan) return;
    if (plane_begin < 0 || plane_end > plan->geom.num_planes ||
            plane_begin > plane_end)
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("Invalid w-plane range [%d, %d) for a lattice of %d "
                "planes.", plane_begin, plane_end, plan->geom.num_planes);
        return;
    }
    plan->plane_begin = plane_begin;
    plan->plane_end = plane_end;
    plan->apply_correction = apply_correction;
}

static double event_ms(sdp_GridderUvwEsFft* plan, int from, int to);
static void stage_collect(sdp_GridderUvwEsFft* plan, int degrid);

void sdp_gridder_uvw_es_fft_enable_timing(sdp_GridderUvwEsFft* plan,
        int32_t enable)
{
    if (plan) plan->timing = enable;
}

void sdp_gridder_uvw_es_fft_get_stats(const sdp_GridderUvwEsFft* plan_,
        sdp_EsCallStats* stats)
{
    sdp_GridderUvwEsFft* plan = (sdp_GridderUvwEsFft*) plan_;
    if (!plan || !stats) return;
    if (plan->stats_pending && plan->timing && plan->have_events)
    {
        /* waits for the call whose timings are wanted, nothing else does */
        if (cudaEventSynchronize(plan->ev[3]) == cudaSuccess)
        {
            plan->stats.ms_presort = event_ms(plan, 0, 1);
            plan->stats.ms_total = event_ms(plan, 0, 3);
            stage_collect(plan, plan->stats_pending == 2);
            uint64_t totals[2] = {0, 0};
            if (plan->occ_totals.ptr && cudaMemcpy(totals,
                    plan->occ_totals.ptr, sizeof(totals),
                    cudaMemcpyDeviceToHost) == cudaSuccess)
            {
                plan->stats.occupied_tiles = (int64_t) totals[0];
                plan->stats.occupied_tile_rows = (int64_t) totals[1];
            }
        }
        else
        {
            (void) cudaGetLastError();
        }
    }
    plan->stats_pending = 0;
    *stats = plan->stats;
}

int32_t sdp_gridder_uvw_es_fft_get_batch_times(
        const sdp_GridderUvwEsFft* plan_, int32_t capacity,
        int32_t* first_plane, int32_t* num_planes, double* ms)
{
    sdp_GridderUvwEsFft* plan = (sdp_GridderUvwEsFft*) plan_;
    if (!plan || !plan->timing || !plan->have_events) return 0;
    if (plan->stats_pending)
    {
        sdp_EsCallStats unused;
        sdp_gridder_uvw_es_fft_get_stats(plan, &unused); /* waits for the call */
    }
    int32_t n = 0;
    for (int k = 0; k < plan->batch_used && 4 * k + 3 < plan->stage_used &&
            n < capacity; ++k, ++n)
    {
        float t = 0.0f;
        if (cudaEventElapsedTime(&t, plan->stage_ev[4 * k],
                plan->stage_ev[4 * k + 3]) != cudaSuccess)
            (void) cudaGetLastError();
        if (first_plane) first_plane[n] = plan->batch_first[k];
        if (num_planes) num_planes[n] = plan->batch_count[k];
        if (ms) ms[n] = (double) t;
    }
    return n;
}

/* ----------------------------------------------------------------- FFT */

/* Batched in-place 2-D C2C / Z2Z plan over `batch` contiguous planes, cached;
 * all cached plans share one work area. Same transform as the reference asks
 * of cuFFT (sdp_fft.cpp:362-456), only batched and created once. */
static cufftHandle* fft_plan_for(sdp_GridderUvwEsFft* plan, int batch,
        sdp_Error* status)
{
    if (*status) return 0;
    int slot = -1;
    for (int k = 0; k < ES_NUM_FFT_PLANS; ++k)
        if (plan->fft_batch[k] == batch) return &plan->fft[k];
    for (int k = 0; k < ES_NUM_FFT_PLANS; ++k)
        if (plan->fft_batch[k] == 0) { slot = k; break; }
    if (slot < 0)
    {
        slot = ES_NUM_FFT_PLANS - 1;
        (void) cufftDestroy(plan->fft[slot]);
        plan->fft_batch[slot] = 0;
    }
    cufftHandle h;
    cufftResult rc = cufftCreate(&h);
    if (rc == CUFFT_SUCCESS) rc = cufftSetAutoAllocation(h, 0);
    int n[2] = {plan->geom.grid_size, plan->geom.grid_size};
    size_t work = 0;
    if (rc == CUFFT_SUCCESS)
        rc = cufftMakePlanMany(h, 2, n, 0, 1, 0, 0, 1, 0,
                plan->info.is_double ? CUFFT_Z2Z : CUFFT_C2C, batch, &work);
    if (rc != CUFFT_SUCCESS)
    {
        *status = SDP_ERR_RUNTIME;
        SDP_LOG_ERROR("Error in cufftMakePlanMany (code %d)", (int) rc);
        (void) cufftDestroy(h);
        return 0;
    }
    if (work > plan->fft_work.bytes)
    {
        /* all plans must be re-pointed at the new area */
        (void) cudaStreamSynchronize(plan->stream);
        buffer_reserve(&plan->fft_work, work, status);
        if (*status) { (void) cufftDestroy(h); return 0; }
        for (int k = 0; k < ES_NUM_FFT_PLANS; ++k)
            if (plan->fft_batch[k] > 0)
                (void) cufftSetWorkArea(plan->fft[k], plan->fft_work.ptr);
    }
    rc = cufftSetWorkArea(h, plan->fft_work.ptr);
    if (rc == CUFFT_SUCCESS) rc = cufftSetStream(h, plan->stream);
    if (rc != CUFFT_SUCCESS)
    {
        *status = SDP_ERR_RUNTIME;
        SDP_LOG_ERROR("Error configuring cuFFT plan (code %d)", (int) rc);
        (void) cufftDestroy(h);
        return 0;
    }
    plan->fft[slot] = h;
    plan->fft_batch[slot] = batch;
    return &plan->fft[slot];
}

static void fft_exec(sdp_GridderUvwEsFft* plan, int batch, int forward,
        sdp_Error* status)
{
    cufftHandle* h = fft_plan_for(plan, batch, status);
    if (*status || !h) return;
    cufftResult rc;
    const int dir = forward ? CUFFT_FORWARD : CUFFT_INVERSE;
    if (plan->info.is_double)
        rc = cufftExecZ2Z(*h, (cufftDoubleComplex*) plan->planes.ptr,
                (cufftDoubleComplex*) plan->planes.ptr, dir);
    else
        rc = cufftExecC2C(*h, (cufftComplex*) plan->planes.ptr,
                (cufftComplex*) plan->planes.ptr, dir);
    if (rc != CUFFT_SUCCESS)
    {
        *status = SDP_ERR_RUNTIME;
        SDP_LOG_ERROR("Error in cufftExec (code %d)", (int) rc);
    }
    plan->stats.fft_execs++;
}

/* -------------------------------------------------------------- staging */

typedef struct EsArrays
{
    const void* uvw;
    const void* freq;
    void* vis;
    const void* weight;
    void* dirty;
    int staged;
} EsArrays;

static size_t mem_bytes(const sdp_Mem* m)
{
    int64_t n = 1;
    for (int32_t d = 0; d < sdp_mem_num_dims(m); ++d)
        n *= sdp_mem_shape_dim(m, d);
    return (size_t) (n * sdp_mem_type_size(sdp_mem_type(m)));
}

static void stage_in(sdp_GridderUvwEsFft* plan, EsBuffer* b,
        const sdp_Mem* m, cudaStream_t stream, sdp_Error* status)
{
    const size_t bytes = mem_bytes(m);
    buffer_reserve(b, bytes > 0 ? bytes : 16, status);
    if (*status || bytes == 0) return;
    const cudaError_t err = cudaMemcpyAsync(b->ptr, sdp_mem_data_const(m),
            bytes, cudaMemcpyHostToDevice, stream);
    if (err != cudaSuccess)
    {
        *status = SDP_ERR_MEM_COPY_FAILURE;
        SDP_LOG_CRITICAL("cudaMemcpy error: %s", cudaGetErrorString(err));
        (void) cudaGetLastError();
        return;
    }
    plan->stats.h2d_bytes += (int64_t) bytes;
}

static void stage_out(sdp_GridderUvwEsFft* plan, const EsBuffer* b,
        sdp_Mem* m, sdp_Error* status)
{
    const size_t bytes = mem_bytes(m);
    if (*status || bytes == 0) return;
    const cudaError_t err = cudaMemcpyAsync(sdp_mem_data(m), b->ptr, bytes,
            cudaMemcpyDeviceToHost, plan->stream);
    if (err != cudaSuccess)
    {
        *status = SDP_ERR_MEM_COPY_FAILURE;
        SDP_LOG_CRITICAL("cudaMemcpy error: %s", cudaGetErrorString(err));
        (void) cudaGetLastError();
        return;
    }
    plan->stats.d2h_bytes += (int64_t) bytes;
}

/* Device views of the five arrays; host arrays are copied in (extension over
 * the reference, which would fail with SDP_ERR_MEM_LOCATION). */
static void resolve_arrays(sdp_GridderUvwEsFft* plan, const sdp_Mem* uvw,
        const sdp_Mem* freq_hz, const sdp_Mem* vis, const sdp_Mem* weight,
        const sdp_Mem* dirty_image, int need_weight, EsArrays* a,
        sdp_Error* status)
{
    memset(a, 0, sizeof(*a));
    if (*status) return;
    if (sdp_mem_location(uvw) == SDP_MEM_GPU)
    {
        a->uvw = sdp_mem_data_const(uvw);
        a->freq = sdp_mem_data_const(freq_hz);
        a->vis = (void*) sdp_mem_data_const(vis);
        a->weight = sdp_mem_data_const(weight);
        a->dirty = (void*) sdp_mem_data_const(dirty_image);
        return;
    }
    a->staged = 1;
    /* The array that is only needed at the end of the call (gridding: the
     * image to accumulate into; degridding: the visibilities to add to) is
     * copied on a second stream, overlapping the presort and the w-planes;
     * the caller waits with late_copy_wait() before its first use. */
    cudaStream_t late = plan->stream;
    if (!plan->have_copy_stream &&
            cudaStreamCreateWithFlags(&plan->copy_stream,
                    cudaStreamNonBlocking) == cudaSuccess)
    {
        if (cudaEventCreateWithFlags(&plan->copy_ev,
                cudaEventDisableTiming) == cudaSuccess)
            plan->have_copy_stream = 1;
        else
            (void) cudaStreamDestroy(plan->copy_stream);
    }
    if (plan->have_copy_stream) late = plan->copy_stream;
    stage_in(plan, &plan->st_uvw, uvw, plan->stream, status);
    stage_in(plan, &plan->st_freq, freq_hz, plan->stream, status);
    stage_in(plan, &plan->st_vis, vis, need_weight ? plan->stream : late,
            status);
    if (need_weight)
        stage_in(plan, &plan->st_weight, weight, plan->stream, status);
    stage_in(plan, &plan->st_dirty, dirty_image,
            need_weight ? late : plan->stream, status);
    if (plan->have_copy_stream)
        (void) cudaEventRecord(plan->copy_ev, plan->copy_stream);
    a->uvw = plan->st_uvw.ptr;
    a->freq = plan->st_freq.ptr;
    a->vis = plan->st_vis.ptr;
    a->weight = plan->st_weight.ptr;
    a->dirty = plan->st_dirty.ptr;
}

/* Orders the main stream after the late staging copy of resolve_arrays. */
static void late_copy_wait(sdp_GridderUvwEsFft* plan, const EsArrays* a)
{
    if (a->staged && plan->have_copy_stream)
        (void) cudaStreamWaitEvent(plan->stream, plan->copy_ev, 0);
}

/* -------------------------------------------------------------- presort */

static int bits_for(uint64_t max_value)
{
    int bits = 1;
    while (bits < 64 && (max_value >> bits) != 0) bits++;
    return bits;
}

/* Expands the visibilities into (tile, plane)-sorted records for the owned
 * plane range. Leaves keys in keys_b, records in records; sets *num_entries. */
static void presort(sdp_GridderUvwEsFft* plan, const EsArrays* a,
        int for_gridding, int64_t* num_entries, sdp_Error* status)
{
    *num_entries = 0;
    if (*status) return;
    EsGeom* g = &plan->geom;
    cudaStream_t s = plan->stream;
    const int64_t nvis = (int64_t) g->num_rows * g->num_chan;
    const size_t ntiles = (size_t) g->ntile * g->ntile;
    buffer_reserve(&plan->tile_start, (ntiles + 2) * sizeof(uint32_t), status);
    buffer_reserve(&plan->counts, (size_t) nvis + 16, status);
    buffer_reserve(&plan->offsets, ((size_t) nvis + 1) * sizeof(uint32_t),
            status);
    if (*status) return;
    if (nvis == 0 || g->plane_begin >= g->plane_end)
    {
        ES_CUDA(es_launch_tile_bounds(g, 0, 0, (uint32_t*) plan->tile_start.ptr, s));
        return;
    }

    /* entries per visibility -> offsets (the extra zero count makes the
     * total come out as offsets[nvis]) */
    ES_CUDA(cudaMemsetAsync((uint8_t*) plan->counts.ptr + nvis, 0, 1, s));
    ES_CUDA(es_launch_count_entries(g, a->uvw, a->freq,
            (uint8_t*) plan->counts.ptr, s));
    size_t scan_bytes = 0;
    ES_CUDA(es_scan_counts((const uint8_t*) plan->counts.ptr,
            (uint32_t*) plan->offsets.ptr, nvis + 1, 0, &scan_bytes, s));
    buffer_reserve(&plan->sort_temp, scan_bytes, status);
    if (*status) return;
    scan_bytes = plan->sort_temp.bytes;
    ES_CUDA(es_scan_counts((const uint8_t*) plan->counts.ptr,
            (uint32_t*) plan->offsets.ptr, nvis + 1, plan->sort_temp.ptr,
            &scan_bytes, s));
    uint32_t total = 0;
    ES_CUDA(cudaMemcpyAsync(&total, (const uint32_t*) plan->offsets.ptr + nvis,
            sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    ES_CUDA(cudaStreamSynchronize(s));
    const int64_t m = (int64_t) total;
    plan->stats.kernel_launches += 1;
    if (m == 0)
    {
        ES_CUDA(es_launch_tile_bounds(g, 0, 0, (uint32_t*) plan->tile_start.ptr, s));
        return;
    }

    const size_t rec_bytes = es_record_bytes(g->is_double);
    buffer_reserve(&plan->keys_a, (size_t) m * sizeof(uint64_t), status);
    buffer_reserve(&plan->keys_b, (size_t) m * sizeof(uint64_t), status);
    buffer_reserve(&plan->vals_a, (size_t) m * sizeof(uint32_t), status);
    buffer_reserve(&plan->vals_b, (size_t) m * sizeof(uint32_t), status);
    buffer_reserve(&plan->records, (size_t) m * rec_bytes, status);
    /* fp64 gridding: the plane-independent ES weights of every record, once
     * per call (the fp64 scatter spent most of its time re-evaluating them
     * per plane); SKA_SDP_FUNC_ES_NO_KWEIGHTS=1: tests of the in-kernel path */
    const int use_kweights = for_gridding && g->is_double &&
            !getenv("SKA_SDP_FUNC_ES_NO_KWEIGHTS");
    plan->have_kweights = use_kweights;
    if (use_kweights)
        buffer_reserve(&plan->kweights, (size_t) m * es_kweight_count(g) *
                sizeof(double), status);
    if (*status) return;

    ES_CUDA(es_launch_emit_entries(g, a->uvw, a->freq,
            (const uint32_t*) plan->offsets.ptr, (uint64_t*) plan->keys_a.ptr,
            (uint32_t*) plan->vals_a.ptr, s));
    const uint64_t pk = (uint64_t) (g->plane_end - g->plane_begin);
    const int end_bit = bits_for((uint64_t) ntiles * pk);
    size_t sort_bytes = 0;
    ES_CUDA(es_sort_entries((const uint64_t*) plan->keys_a.ptr,
            (uint64_t*) plan->keys_b.ptr, (const uint32_t*) plan->vals_a.ptr,
            (uint32_t*) plan->vals_b.ptr, m, end_bit, 0, &sort_bytes, s));
    buffer_reserve(&plan->sort_temp, sort_bytes, status);
    if (*status) return;
    sort_bytes = plan->sort_temp.bytes;
    ES_CUDA(es_sort_entries((const uint64_t*) plan->keys_a.ptr,
            (uint64_t*) plan->keys_b.ptr, (const uint32_t*) plan->vals_a.ptr,
            (uint32_t*) plan->vals_b.ptr, m, end_bit, plan->sort_temp.ptr,
            &sort_bytes, s));
    ES_CUDA(es_launch_tile_bounds(g, (const uint64_t*) plan->keys_b.ptr, m,
            (uint32_t*) plan->tile_start.ptr, s));
    /* degridding: the sort input values are dead now; their buffer takes the
     * sorted position of every entry (entry_pos, used by the final
     * per-visibility sum) */
    ES_CUDA(es_launch_build_records(g, a->uvw, a->freq,
            for_gridding ? a->vis : 0, for_gridding ? a->weight : 0,
            (const uint64_t*) plan->keys_b.ptr,
            (const uint32_t*) plan->vals_b.ptr, m, plan->records.ptr,
            (const uint32_t*) plan->offsets.ptr,
            for_gridding ? 0 : (uint32_t*) plan->vals_a.ptr,
            use_kweights ? plan->kweights.ptr : 0, s));
    plan->stats.kernel_launches += 3;
    *num_entries = m;
}

/* ----------------------------------------------------------- work units */

#define ES_MAX_SPLIT_ITEMS 65536u
#define ES_PARTIALS_BYTES ((size_t) 512 << 20)
#define ES_MIN_SPLIT 8192u

typedef struct EsWorkLimits
{
    uint32_t split, max_units, max_partials;
} EsWorkLimits;

/* Sizes the unit list and the partial-tile scratch for batches of up to
 * `nb` planes over `m` records: a (tile, plane) item with more than `split`
 * records is cut into slices (es_scatter.cu). */
static void work_reserve(sdp_GridderUvwEsFft* plan, int nb, int64_t m,
        int with_partials, EsWorkLimits* lim, sdp_Error* status)
{
    const EsGeom* g = &plan->geom;
    const size_t tile_bytes = es_partial_tile_bytes(g->is_double);
    lim->max_partials = with_partials ?
            (uint32_t) (ES_PARTIALS_BYTES / tile_bytes) : 0u;
    uint32_t split = ES_MIN_SPLIT;
    const char* env = getenv("SKA_SDP_FUNC_ES_SPLIT");
    if (env && atoi(env) > 0) split = (uint32_t) atoi(env);
    /* worst case: every record touches every plane of the batch */
    const double visits = (double) nb * (double) m;
    if (with_partials && visits / split > 0.5 * lim->max_partials)
        split = (uint32_t) (visits / (0.5 * lim->max_partials)) + 1u;
    lim->split = split;
    const double units = (double) g->ntile * g->ntile * nb + visits / split + 64.0;
    lim->max_units = (uint32_t) units;
    buffer_reserve(&plan->units, (size_t) lim->max_units * es_unit_bytes(), status);
    buffer_reserve(&plan->unit_counters, 8 * sizeof(uint32_t), status);
    buffer_reserve(&plan->split_done, ES_MAX_SPLIT_ITEMS * sizeof(uint32_t), status);
    if (with_partials)
        buffer_reserve(&plan->partials, (size_t) lim->max_partials * tile_bytes, status);
}

/* Planes per batch for an owned range of `count` planes: as few batches as
 * the resident capacity allows, all of (nearly) the same size, so that a
 * range of e.g. 25 planes with 12 resident runs as 9 + 8 + 8 and not as
 * 12 + 12 + 1 (the line-FFT kernels fill the device per launch). */
static int even_batch(const sdp_GridderUvwEsFft* plan, int count)
{
    if (count <= plan->plane_batch) return count > 0 ? count : 1;
    const int nbatches = (count + plan->plane_batch - 1) / plan->plane_batch;
    return (count + nbatches - 1) / nbatches;
}

/* --------------------------------------------------------------- timing */

static void events_begin(sdp_GridderUvwEsFft* plan)
{
    memset(&plan->stats, 0, sizeof(plan->stats));
    plan->stats.plane_batch = plan->plane_batch;
    plan->stats.timing_enabled = plan->timing;
    plan->stats.pruned_fft = plan->pruned_fft ? 1 : 0;
    plan->stage_used = 0;
    plan->batch_used = 0;
    if (plan->occ_totals.ptr)
        (void) cudaMemsetAsync(plan->occ_totals.ptr, 0, 2 * sizeof(uint64_t),
                plan->stream);
    if (plan->timing && !plan->have_events)
    {
        int ok = 1;
        for (int k = 0; k < 8; ++k)
            if (cudaEventCreate(&plan->ev[k]) != cudaSuccess) ok = 0;
        plan->have_events = ok;
        if (!ok) (void) cudaGetLastError();
    }
    if (plan->timing && plan->have_events)
        (void) cudaEventRecord(plan->ev[0], plan->stream);
}

/* Per-stage timing: four events per plane batch (before stage 1, after
 * stages 1, 2, 3), summed after the call has drained. */
static void stage_mark(sdp_GridderUvwEsFft* plan)
{
    if (!plan->timing) return;
    if (plan->stage_used == plan->stage_cap)
    {
        const int cap = plan->stage_cap ? 2 * plan->stage_cap : 256;
        cudaEvent_t* grown = (cudaEvent_t*) realloc(plan->stage_ev,
                sizeof(cudaEvent_t) * (size_t) cap);
        if (!grown) return;
        plan->stage_ev = grown;
        for (int k = plan->stage_cap; k < cap; ++k)
        {
            if (cudaEventCreate(&plan->stage_ev[k]) != cudaSuccess)
            {
                (void) cudaGetLastError();
                plan->stage_cap = k;
                return;
            }
        }
        plan->stage_cap = cap;
    }
    (void) cudaEventRecord(plan->stage_ev[plan->stage_used++], plan->stream);
}

/* Remembers which planes the batch whose four stage events follow covers. */
static void batch_mark(sdp_GridderUvwEsFft* plan, int first, int count)
{
    if (!plan->timing) return;
    if (plan->batch_used == plan->batch_cap)
    {
        const int cap = plan->batch_cap ? 2 * plan->batch_cap : 64;
        int32_t* f = (int32_t*) realloc(plan->batch_first,
                sizeof(int32_t) * (size_t) cap);
        if (f) plan->batch_first = f;
        int32_t* c = (int32_t*) realloc(plan->batch_count,
                sizeof(int32_t) * (size_t) cap);
        if (c) plan->batch_count = c;
        if (!f || !c) return;
        plan->batch_cap = cap;
    }
    plan->batch_first[plan->batch_used] = first;
    plan->batch_count[plan->batch_used] = count;
    plan->batch_used++;
}

/* stage order: grid = scatter, fft, screen; degrid = screen, fft, gather */
static void stage_collect(sdp_GridderUvwEsFft* plan, int degrid)
{
    double sum[3] = {0.0, 0.0, 0.0};
    for (int k = 0; k + 3 < plan->stage_used; k += 4)
    {
        for (int j = 0; j < 3; ++j)
        {
            float ms = 0.0f;
            if (cudaEventElapsedTime(&ms, plan->stage_ev[k + j],
                    plan->stage_ev[k + j + 1]) != cudaSuccess)
                (void) cudaGetLastError();
            sum[j] += (double) ms;
        }
    }
    plan->stats.ms_scatter_gather = degrid ? sum[2] : sum[0];
    plan->stats.ms_fft = sum[1];
    plan->stats.ms_screen = degrid ? sum[0] : sum[2];
}

static void event_mark(sdp_GridderUvwEsFft* plan, int k)
{
    if (plan->timing && plan->have_events)
        (void) cudaEventRecord(plan->ev[k], plan->stream);
}

static double event_ms(sdp_GridderUvwEsFft* plan, int from, int to)
{
    float ms = 0.0f;
    if (cudaEventElapsedTime(&ms, plan->ev[from], plan->ev[to]) != cudaSuccess)
        (void) cudaGetLastError();
    return (double) ms;
}

/* ------------------------------------------------------------- gridding */

void sdp_grid_uvw_es_fft(sdp_GridderUvwEsFft* plan, const sdp_Mem* uvw,
        const sdp_Mem* freq_hz, const sdp_Mem* vis, const sdp_Mem* weight,
        sdp_Mem* dirty_image, sdp_Error* status)
{
    SDP_LOG_DEBUG("Executing sdp_GridderUvwEsFft...");
    if (*status || !plan) return;
    check_parameters(plan->pixsize_x_rad, plan->pixsize_y_rad, status);
    if (*status) return;
    check_buffers(uvw, freq_hz, vis, weight, dirty_image, 0, status);
    check_against_plan(plan, vis, dirty_image, status);
    if (*status) return;

    EsGeom* g = &plan->geom;
    g->plane_begin = plan->plane_begin;
    g->plane_end = plan->plane_end;
    cudaStream_t s = plan->stream;
    events_begin(plan);

    EsArrays a;
    resolve_arrays(plan, uvw, freq_hz, vis, weight, dirty_image, 1, &a, status);
    if (*status) return;

    int64_t m = 0;
    presort(plan, &a, 1, &m, status);
    if (*status) return;
    plan->stats.num_entries = m;
    if (!plan->pruned_fft) late_copy_wait(plan, &a); /* summed into per batch */
    if (plan->pruned_fft)
        ES_CUDA(cudaMemsetAsync(plan->image_t.ptr, 0, real_bytes(plan) *
                (size_t) g->image_size * (size_t) g->image_size, s));
    EsWorkLimits lim;
    work_reserve(plan, plan->plane_batch, m, 1, &lim, status);
    if (*status) return;
    event_mark(plan, 1);

    /* order of the scatter's work list: plane-major, or tile-major where the
     * per-record weight tables are re-read per plane (fp64);
     * SKA_SDP_FUNC_ES_UNIT_ORDER=1|2 overrides (experiments) */
    int unit_order = plan->have_kweights ? 2 : 1;
    if (getenv("SKA_SDP_FUNC_ES_UNIT_ORDER"))
        unit_order = atoi(getenv("SKA_SDP_FUNC_ES_UNIT_ORDER")) == 2 ? 2 : 1;
    const int batch = even_batch(plan, g->plane_end - g->plane_begin);
    for (int first = g->plane_begin; first < g->plane_end; first += batch)
    {
        int nb = g->plane_end - first;
        if (nb > batch) nb = batch;
        batch_mark(plan, first, nb);
        stage_mark(plan);
        /* pruned mode: only occupied tiles are written / transformed */
        if (plan->pruned_fft)
            ES_CUDA(es_launch_occupancy(g, (const uint64_t*) plan->keys_b.ptr,
                    (const uint32_t*) plan->tile_start.ptr, first, nb,
                    (uint32_t*) plan->occ.ptr,
                    (unsigned long long*) plan->occ_totals.ptr, s));
        ES_CUDA(es_launch_work_plan(g, (const uint64_t*) plan->keys_b.ptr,
                (const uint32_t*) plan->tile_start.ptr, first, nb,
                unit_order,
                lim.split, lim.max_units, ES_MAX_SPLIT_ITEMS, lim.max_partials,
                plan->pruned_fft ? 1 : 0, plan->units.ptr, (uint32_t*) plan->unit_counters.ptr,
                (uint32_t*) plan->split_done.ptr, s));
        ES_CUDA(es_launch_scatter2(g, plan->records.ptr, plan->units.ptr,
                (uint32_t*) plan->unit_counters.ptr, first, plan->planes.ptr,
                plan->partials.ptr, (uint32_t*) plan->split_done.ptr,
                plan->have_kweights ? plan->kweights.ptr : 0, s));
        plan->stats.kernel_launches += 1;
        stage_mark(plan);
        if (plan->pruned_fft)
        {
            /* inverse transform (es_fft.cpp:568) as two pruned line passes,
             * the second fused with the w-screen and the plane sum */
            ES_CUDA(es_fft_grid_rows(plan->pruned_fft, plan->planes.ptr, nb,
                    plan->strip.ptr, (const uint32_t*) plan->occ.ptr, s));
            stage_mark(plan);
            ES_CUDA(es_fft_grid_cols_screen(plan->pruned_fft, g,
                    plan->strip.ptr, first, nb, plan->image_t.ptr,
                    (const uint32_t*) plan->occ.ptr, s));
            plan->stats.kernel_launches += 1;
        }
        else
        {
            fft_exec(plan, nb, 0, status); /* CUFFT_INVERSE, es_fft.cpp:568 */
            if (*status) return;
            stage_mark(plan);
            ES_CUDA(es_launch_apply_screen(g, plan->planes.ptr, first, nb,
                    a.dirty, s));
        }
        stage_mark(plan);
        plan->stats.kernel_launches += 2;
        plan->stats.planes_processed += nb;
    }
    if (plan->pruned_fft)
    {
        /* dirty += transpose(accumulator) */
        late_copy_wait(plan, &a);
        ES_CUDA(es_fft_image_transpose(plan->pruned_fft, plan->image_t.ptr,
                a.dirty, 1, g->do_wstacking, s));
        plan->stats.kernel_launches += 1;
    }
    event_mark(plan, 2);
    if (plan->apply_correction)
    {
        ES_CUDA(es_launch_correct(g, a.dirty, plan->d_quad_kernel,
                plan->d_quad_nodes, plan->d_quad_weights, plan->d_conv_corr,
                s));
        plan->stats.kernel_launches += 1;
    }
    if (a.staged)
    {
        stage_out(plan, &plan->st_dirty, dirty_image, status);
        ES_CUDA(cudaStreamSynchronize(s));
    }
    event_mark(plan, 3);
    plan->stats_pending = 1; /* timings are collected by get_stats */
}

/* ----------------------------------------------------------- degridding */

void sdp_ifft_degrid_uvw_es(sdp_GridderUvwEsFft* plan, const sdp_Mem* uvw,
        const sdp_Mem* freq_hz, sdp_Mem* vis, const sdp_Mem* weight,
        sdp_Mem* dirty_image, sdp_Error* status)
{
    SDP_LOG_DEBUG("Executing sdp_GridderUvwEsFft...");
    if (*status || !plan) return;
    check_parameters(plan->pixsize_x_rad, plan->pixsize_y_rad, status);
    if (*status) return;
    check_buffers(uvw, freq_hz, vis, weight, dirty_image, 1, status);
    check_against_plan(plan, vis, dirty_image, status);
    if (*status) return;

    EsGeom* g = &plan->geom;
    g->plane_begin = plan->plane_begin;
    g->plane_end = plan->plane_end;
    cudaStream_t s = plan->stream;
    events_begin(plan);

    EsArrays a;
    resolve_arrays(plan, uvw, freq_hz, vis, weight, dirty_image, 0, &a, status);
    if (*status) return;

    int64_t m = 0;
    presort(plan, &a, 0, &m, status);
    if (*status) return;
    plan->stats.num_entries = m;
    const size_t cplx = 2 * real_bytes(plan);
    buffer_reserve(&plan->acc, (size_t) (m > 0 ? m : 1) * cplx, status);
    if (*status) return;
    ES_CUDA(cudaMemsetAsync(plan->acc.ptr, 0, (size_t) (m > 0 ? m : 1) * cplx, s));
    EsWorkLimits glim;
    work_reserve(plan, 1, m, 0, &glim, status);
    if (*status) return;
    event_mark(plan, 1);

    /* grid correction in place on the caller's image (es_fft.cpp:789-825) */
    if (plan->apply_correction)
    {
        ES_CUDA(es_launch_correct(g, a.dirty, plan->d_quad_kernel,
                plan->d_quad_nodes, plan->d_quad_weights, plan->d_conv_corr,
                s));
        plan->stats.kernel_launches += 1;
    }
    if (plan->pruned_fft)
    {
        /* the line FFTs read image columns: transposed copy, once */
        ES_CUDA(es_fft_image_transpose(plan->pruned_fft, a.dirty,
                plan->image_t.ptr, 0, g->do_wstacking, s));
        plan->stats.kernel_launches += 1;
    }
    const int batch = even_batch(plan, g->plane_end - g->plane_begin);
    for (int first = g->plane_begin; first < g->plane_end; first += batch)
    {
        int nb = g->plane_end - first;
        if (nb > batch) nb = batch;
        batch_mark(plan, first, nb);
        stage_mark(plan);
        if (plan->pruned_fft)
        {
            /* only the tiles the gather reads are produced */
            ES_CUDA(es_launch_occupancy(g, (const uint64_t*) plan->keys_b.ptr,
                    (const uint32_t*) plan->tile_start.ptr, first, nb,
                    (uint32_t*) plan->occ.ptr,
                    (unsigned long long*) plan->occ_totals.ptr, s));
            ES_CUDA(es_fft_degrid_cols(plan->pruned_fft, g,
                    plan->image_t.ptr, first, nb, plan->strip.ptr,
                    (const uint32_t*) plan->occ.ptr, s));
            stage_mark(plan);
            ES_CUDA(es_fft_degrid_rows(plan->pruned_fft, plan->strip.ptr, nb,
                    plan->planes.ptr, (const uint32_t*) plan->occ.ptr, s));
        }
        else
        {
            ES_CUDA(es_launch_reverse_screen(g, a.dirty, first, nb,
                    plan->planes.ptr, s));
            stage_mark(plan);
            fft_exec(plan, nb, 1, status); /* CUFFT_FORWARD, es_fft.cpp:782 */
            if (*status) return;
        }
        stage_mark(plan);
        /* (tile, record slice) units: heavy tiles are shared by many CTAs */
        ES_CUDA(es_launch_work_plan(g, (const uint64_t*) plan->keys_b.ptr,
                (const uint32_t*) plan->tile_start.ptr, first, nb, 0,
                glim.split, glim.max_units, 0, 0, 1, plan->units.ptr,
                (uint32_t*) plan->unit_counters.ptr, 0, s));
        ES_CUDA(es_launch_gather2(g, plan->records.ptr, plan->units.ptr,
                (uint32_t*) plan->unit_counters.ptr, first, nb,
                plan->planes.ptr, plan->acc.ptr, s));
        stage_mark(plan);
        plan->stats.kernel_launches += 3;
        plan->stats.planes_processed += nb;
    }
    event_mark(plan, 2);
    late_copy_wait(plan, &a);
    ES_CUDA(es_launch_finish_degrid(g, plan->records.ptr, plan->acc.ptr,
            (const uint32_t*) plan->offsets.ptr,
            (const uint32_t*) plan->vals_a.ptr, m, a.vis, s));
    plan->stats.kernel_launches += 1;
    if (a.staged)
    {
        stage_out(plan, &plan->st_vis, vis, status);
        stage_out(plan, &plan->st_dirty, dirty_image, status);
        ES_CUDA(cudaStreamSynchronize(s));
    }
    event_mark(plan, 3);
    plan->stats_pending = 2;
}

/* ----------------------------------------------------------- extensions */

void sdp_gridder_uvw_es_fft_correct_image(sdp_GridderUvwEsFft* plan,
        sdp_Mem* dirty_image, sdp_Error* status)
{
    if (*status || !plan) return;
    if (sdp_mem_shape_dim(dirty_image, 0) != plan->info.image_size ||
            sdp_mem_shape_dim(dirty_image, 1) != plan->info.image_size ||
            (sdp_mem_is_double(dirty_image) != 0) != (plan->info.is_double != 0)
            || sdp_mem_is_complex(dirty_image) ||
            !sdp_mem_is_c_contiguous(dirty_image))
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("Image does not match the plan.");
        return;
    }
    void* d = sdp_mem_data(dirty_image);
    const int staged = sdp_mem_location(dirty_image) != SDP_MEM_GPU;
    if (staged)
    {
        stage_in(plan, &plan->st_dirty, dirty_image, plan->stream, status);
        if (*status) return;
        d = plan->st_dirty.ptr;
    }
    ES_CUDA(es_launch_correct(&plan->geom, d, plan->d_quad_kernel,
            plan->d_quad_nodes, plan->d_quad_weights, plan->d_conv_corr,
            plan->stream));
    if (staged)
    {
        stage_out(plan, &plan->st_dirty, dirty_image, status);
        ES_CUDA(cudaStreamSynchronize(plan->stream));
    }
}

void sdp_gridder_uvw_es_fft_plane_histogram(sdp_GridderUvwEsFft* plan,
        const sdp_Mem* uvw, const sdp_Mem* freq_hz, int64_t* counts,
        sdp_Error* status)
{
    if (*status || !plan || !counts) return;
    if (sdp_mem_location(uvw) != sdp_mem_location(freq_hz))
    {
        *status = SDP_ERR_MEM_LOCATION;
        SDP_LOG_ERROR("Memory location mismatch.");
        return;
    }
    if (sdp_mem_shape_dim(uvw, 0) != plan->info.num_rows ||
            sdp_mem_shape_dim(uvw, 1) != 3 ||
            sdp_mem_shape_dim(freq_hz, 0) != plan->info.num_chan ||
            (sdp_mem_is_double(uvw) != 0) != (plan->info.is_double != 0) ||
            (sdp_mem_is_double(freq_hz) != 0) != (plan->info.is_double != 0))
    {
        *status = SDP_ERR_INVALID_ARGUMENT;
        SDP_LOG_ERROR("uvw / freq_hz do not match the plan.");
        return;
    }
    EsGeom g = plan->geom;
    g.plane_begin = 0;
    g.plane_end = g.num_planes;
    const void* d_uvw = sdp_mem_data_const(uvw);
    const void* d_freq = sdp_mem_data_const(freq_hz);
    if (sdp_mem_location(uvw) != SDP_MEM_GPU)
    {
        stage_in(plan, &plan->st_uvw, uvw, plan->stream, status);
        stage_in(plan, &plan->st_freq, freq_hz, plan->stream, status);
        if (*status) return;
        d_uvw = plan->st_uvw.ptr;
        d_freq = plan->st_freq.ptr;
    }
    const size_t bytes = (size_t) g.num_planes * sizeof(unsigned long long);
    buffer_reserve(&plan->sort_temp, bytes, status);
    if (*status) return;
    ES_CUDA(cudaMemsetAsync(plan->sort_temp.ptr, 0, bytes, plan->stream));
    ES_CUDA(es_launch_plane_histogram(&g, d_uvw, d_freq,
            (unsigned long long*) plan->sort_temp.ptr, plan->stream));
    ES_CUDA(cudaMemcpyAsync(counts, plan->sort_temp.ptr, bytes,
            cudaMemcpyDeviceToHost, plan->stream));
    ES_CUDA(cudaStreamSynchronize(plan->stream));
}
