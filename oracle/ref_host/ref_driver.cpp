/*
 * TEST INFRASTRUCTURE ONLY (oracle). Not part of the product library; nothing
 * under ska-sdp-func_b200/ links, loads or calls this.
 *
 * "oracle/_ref": the reference's OWN sources for the sdp_GridderUvwEsFft path,
 * compiled for the host CPU where they lie under /root/reference (never
 * copied), driven by a CPU restatement of the reference host orchestration:
 *
 *   - kernels: src/ska-sdp-func/grid_data/sdp_gridder_uvw_es_fft_kernels.cu
 *              (all of it, through cuda_host_shim.h; one call == one CUDA thread)
 *   - planning maths: src/ska-sdp-func/grid_data/sdp_gridder_uvw_es_fft_utils.cpp
 *              (KernelDB selection, Gauss-Legendre tables; included unmodified)
 *   - FFT: the reference's CPU backend, vendored PocketFFT
 *              src/ska-sdp-func/fourier_transforms/pocketfft_hdronly.h, called
 *              as in sdp_fft.cpp:632-670 (c2c, BACKWARD for gridding, FORWARD
 *              for degridding, scale 1)
 *   - this driver restates sdp_gridder_uvw_es_fft.cpp:301-424 (plan scalars),
 *     :578-741 (gridding loop) and :789-952 (degridding loop) over CPU buffers.
 *
 * The reference ships no CPU implementation of this path (plan buffers are
 * always SDP_MEM_GPU, es_fft.cpp:491-513), so this is the closest thing to
 * "the reference run on the host": the same per-thread arithmetic, the same
 * schedule (every visibility visited for every w-plane), OpenMP over threads.
 */
#include "cuda_host_shim.h"

#include <algorithm>
#include <chrono>
#include <complex>
#include <cstdio>
#include <cstring>
#include <vector>

#include <omp.h>

thread_local oracle_dim3 threadIdx = {0, 0, 0}, blockIdx = {0, 0, 0},
        blockDim = {1, 1, 1}, gridDim = {1, 1, 1};

// The reference sources, compiled in place.
#include "ska-sdp-func/grid_data/sdp_gridder_uvw_es_fft_kernels.cu"
#include "ska-sdp-func/grid_data/sdp_gridder_uvw_es_fft_utils.cpp"
#include "ska-sdp-func/fourier_transforms/pocketfft_hdronly.h"

namespace {

double now_s()
{
    using namespace std::chrono;
    return duration<double>(steady_clock::now().time_since_epoch()).count();
}

struct Plan
{
    int num_rows, num_chan, image_size, grid_size, support;
    int plan_prec, compute_prec, do_wstacking, num_total_w_grids;
    double epsilon, beta, pixel_size, uv_scale;
    double min_abs_w, max_abs_w, min_plane_w, max_plane_w;
    double w_scale, inv_w_scale, inv_w_range, conv_corr_norm_factor;
    std::vector<double> qk, qn, qw, cc; // tables in double
    std::vector<float> qk_f, qn_f, qw_f, cc_f; // cast as es_fft.cpp:466-483
};

template<typename T> struct Types;
template<> struct Types<float>
{
    typedef float2 T2;
    typedef float3 T3;
};
template<> struct Types<double>
{
    typedef double2 T2;
    typedef double3 T3;
};

template<typename T>
void fft2(std::complex<T>* plane, int g, bool forward, int nthreads)
{
    pocketfft::shape_t shape = {(size_t) g, (size_t) g};
    pocketfft::stride_t stride = {
        (ptrdiff_t) (g * sizeof(std::complex<T>)),
        (ptrdiff_t) sizeof(std::complex<T>)
    };
    pocketfft::shape_t axes = {0, 1};
    pocketfft::c2c<T>(shape, stride, stride, axes,
            forward ? pocketfft::FORWARD : pocketfft::BACKWARD,
            plane, plane, (T) 1, (size_t) nthreads
    );
}

// Conservative prefilter for the optional "fast" schedule: the set of planes a
// visibility may touch, with one plane of slack either side. The kernel's own
// exact early-out (kernels.cu:186-196) still decides.
template<typename T>
void plane_window(const Plan* p, const T* uvw, const T* freq, int row, int chan,
        int& lo, int& hi)
{
    if (!p->do_wstacking) { lo = 0; hi = 0; return; }
    const double w = std::fabs((double) uvw[3 * row + 2]) * (double) freq[chan] /
            299792458.0;
    const double pw = (w - p->min_plane_w) * p->w_scale;
    lo = (int) std::floor(pw - 0.5 * p->support) - 1;
    hi = (int) std::ceil(pw + 0.5 * p->support) + 1;
}

template<typename T>
void scatter_gather_plane(const Plan* p, const T* uvw, const T* freq,
        typename Types<T>::T2* vis, const T* weight,
        typename Types<T>::T2* plane, int grid_w, bool solving, int fast)
{
    typedef typename Types<T>::T2 T2;
    typedef typename Types<T>::T3 T3;
    const T beta = (T) p->beta;
    const T uv_scale = (T) p->uv_scale;
    const T w_scale = (T) p->w_scale;
    const T min_plane_w = (T) p->min_plane_w;
    const int R = p->num_rows, C = p->num_chan;
    #pragma omp parallel for schedule(dynamic, 16)
    for (int row = 0; row < R; ++row)
    {
        for (int chan = 0; chan < C; ++chan)
        {
            if (fast)
            {
                int lo, hi;
                plane_window<T>(p, uvw, freq, row, chan, lo, hi);
                if (grid_w < lo || grid_w > hi) continue;
            }
            blockDim.x = 1; blockDim.y = 1; threadIdx.x = 0; threadIdx.y = 0;
            blockIdx.x = (unsigned) chan; blockIdx.y = (unsigned) row;
            if (p->do_wstacking)
            {
                sdp_cuda_nifty_gridder_gridding_3d<T, T2, T, T2, T3>(
                        R, C, vis, weight, (const T3*) uvw, freq, plane,
                        p->grid_size, grid_w, 1u, (uint32_t) p->support,
                        beta, uv_scale, w_scale, min_plane_w, solving
                );
            }
            else
            {
                sdp_cuda_nifty_gridder_gridding_2d<T, T2, T, T2, T3>(
                        R, C, vis, weight, (const T3*) uvw, freq, plane,
                        p->grid_size, grid_w, 1u, (uint32_t) p->support,
                        beta, uv_scale, w_scale, min_plane_w, solving
                );
            }
        }
    }
}

template<typename T>
void conv_corr_image(const Plan* p, T* image, bool solving)
{
    const int half = p->image_size / 2;
    const T* qk = sizeof(T) == 8 ? (const T*) p->qk.data() : (const T*) p->qk_f.data();
    const T* qn = sizeof(T) == 8 ? (const T*) p->qn.data() : (const T*) p->qn_f.data();
    const T* qw = sizeof(T) == 8 ? (const T*) p->qw.data() : (const T*) p->qw_f.data();
    const T* cc = sizeof(T) == 8 ? (const T*) p->cc.data() : (const T*) p->cc_f.data();
    #pragma omp parallel for schedule(static)
    for (int j = 0; j <= half; ++j)
    {
        for (int i = 0; i <= half; ++i)
        {
            blockDim.x = 1; blockDim.y = 1; threadIdx.x = 0; threadIdx.y = 0;
            blockIdx.x = (unsigned) i; blockIdx.y = (unsigned) j;
            conv_corr_and_scaling<T>(image, (uint32_t) p->image_size,
                    (T) p->pixel_size, (uint32_t) p->support,
                    (T) p->conv_corr_norm_factor, cc, (T) p->inv_w_range,
                    (T) p->inv_w_scale, qk, qn, qw, solving,
                    (bool) p->do_wstacking
            );
        }
    }
}

template<typename T>
int run_grid(const Plan* p, const T* uvw, const T* freq, const void* vis_,
        const T* weight, T* dirty, int plane_begin, int plane_end, int fast,
        int nthreads, double* timings)
{
    typedef typename Types<T>::T2 T2;
    const int g = p->grid_size;
    const int half = p->image_size / 2;
    std::vector<std::complex<T> > plane((size_t) g * g);
    T2* plane2 = (T2*) plane.data();
    T2* vis = (T2*) const_cast<void*>(vis_);
    double t_scatter = 0, t_fft = 0, t_screen = 0, t_corr = 0;
    if (plane_end < 0 || plane_end > p->num_total_w_grids)
        plane_end = p->num_total_w_grids;
    for (int grid_w = plane_begin; grid_w < plane_end; ++grid_w)
    {
        double t0 = now_s();
        std::memset((void*) plane.data(), 0, plane.size() * sizeof(plane[0]));
        scatter_gather_plane<T>(p, uvw, freq, vis, weight, plane2, grid_w,
                true, fast
        );
        double t1 = now_s();
        fft2<T>(plane.data(), g, false, nthreads);
        double t2 = now_s();
        #pragma omp parallel for schedule(static)
        for (int j = 0; j <= half; ++j)
        {
            for (int i = 0; i <= half; ++i)
            {
                blockDim.x = 1; blockDim.y = 1;
                threadIdx.x = 0; threadIdx.y = 0;
                blockIdx.x = (unsigned) i; blockIdx.y = (unsigned) j;
                apply_w_screen_and_sum<T, T2>(dirty, (uint32_t) p->image_size,
                        (T) p->pixel_size, plane2, (uint32_t) g, grid_w, 1u,
                        (T) p->inv_w_scale, (T) p->min_plane_w, true,
                        (bool) p->do_wstacking
                );
            }
        }
        double t3 = now_s();
        t_scatter += t1 - t0; t_fft += t2 - t1; t_screen += t3 - t2;
    }
    double t0 = now_s();
    conv_corr_image<T>(p, dirty, true);
    t_corr = now_s() - t0;
    if (timings)
    {
        timings[0] = t_scatter; timings[1] = t_fft;
        timings[2] = t_screen; timings[3] = t_corr;
    }
    return 0;
}

template<typename T>
int run_degrid(const Plan* p, const T* uvw, const T* freq, void* vis_,
        const T* weight, T* dirty, int plane_begin, int plane_end, int fast,
        int nthreads, double* timings)
{
    typedef typename Types<T>::T2 T2;
    const int g = p->grid_size;
    const int half = p->image_size / 2;
    std::vector<std::complex<T> > plane((size_t) g * g);
    T2* plane2 = (T2*) plane.data();
    T2* vis = (T2*) vis_;
    double t_gather = 0, t_fft = 0, t_screen = 0, t_corr = 0;
    if (plane_end < 0 || plane_end > p->num_total_w_grids)
        plane_end = p->num_total_w_grids;
    double t0 = now_s();
    conv_corr_image<T>(p, dirty, false); // in place on the caller's image
    t_corr = now_s() - t0;
    for (int grid_w = plane_begin; grid_w < plane_end; ++grid_w)
    {
        double t1 = now_s();
        std::memset((void*) plane.data(), 0, plane.size() * sizeof(plane[0]));
        #pragma omp parallel for schedule(static)
        for (int j = 0; j <= half; ++j)
        {
            for (int i = 0; i <= half; ++i)
            {
                blockDim.x = 1; blockDim.y = 1;
                threadIdx.x = 0; threadIdx.y = 0;
                blockIdx.x = (unsigned) i; blockIdx.y = (unsigned) j;
                reverse_w_screen_to_stack<T, T2>(dirty,
                        (uint32_t) p->image_size, (T) p->pixel_size, plane2,
                        (uint32_t) g, grid_w, 1u, (T) p->inv_w_scale,
                        (T) p->min_plane_w, true, (bool) p->do_wstacking
                );
            }
        }
        double t2 = now_s();
        fft2<T>(plane.data(), g, true, nthreads);
        double t3 = now_s();
        scatter_gather_plane<T>(p, uvw, freq, vis, weight, plane2, grid_w,
                false, fast
        );
        double t4 = now_s();
        t_screen += t2 - t1; t_fft += t3 - t2; t_gather += t4 - t3;
    }
    if (timings)
    {
        timings[0] = t_gather; timings[1] = t_fft;
        timings[2] = t_screen; timings[3] = t_corr;
    }
    return 0;
}

} // namespace


extern "C" {

struct oref_PlanInfo
{
    int grid_size, support, num_total_w_grids, image_size;
    double beta, pixel_size, uv_scale, w_scale, inv_w_scale;
    double min_plane_w, max_plane_w, conv_corr_norm_factor;
    double min_abs_w, max_abs_w, inv_w_range;
};

// plan_prec selects the KernelDB branch (4: max support 8, 8: max support 16)
// exactly as vis_precision does in es_fft.cpp:321-328; compute_prec selects the
// arithmetic of the kernels (4: the literal float instantiations with all
// scalars and tables cast to float; 8: the double instantiations).
void* oref_create(int num_rows, int num_chan, int image_size, int plan_prec,
        int compute_prec, double pixel_size, double epsilon, double min_abs_w,
        double max_abs_w, int do_wstacking)
{
    Plan* p = new Plan();
    p->num_rows = num_rows; p->num_chan = num_chan;
    p->image_size = image_size;
    p->plan_prec = plan_prec; p->compute_prec = compute_prec;
    p->do_wstacking = do_wstacking; p->epsilon = epsilon;
    p->pixel_size = pixel_size;
    int grid_size = 0, support = 0;
    double beta = NAN;
    sdp_Error status = SDP_SUCCESS;
    sdp_calculate_params_from_epsilon(epsilon, image_size,
            plan_prec == 8 ? SDP_MEM_DOUBLE : SDP_MEM_FLOAT,
            grid_size, support, beta, &status
    );
    beta *= support; // es_fft.cpp:335
    p->grid_size = grid_size; p->support = support; p->beta = beta;
    p->uv_scale = grid_size * pixel_size; // es_fft.cpp:342
    if (do_wstacking) // es_fft.cpp:346-373
    {
        const double x0 = -0.5 * image_size * pixel_size;
        const double y0 = -0.5 * image_size * pixel_size;
        double nmin = sqrt(std::max(1.0 - x0 * x0 - y0 * y0, 0.0)) - 1.0;
        if (x0 * x0 + y0 * y0 > 1.0)
            nmin = -sqrt(fabs(1.0 - x0 * x0 - y0 * y0)) - 1.0;
        double w_scale = 0.25 / fabs(nmin);
        int num_total_w_grids = (max_abs_w - min_abs_w) / w_scale + 2;
        w_scale = 1.0 / ((1.0 + 1e-13) * (max_abs_w - min_abs_w) /
                (num_total_w_grids - 1));
        p->min_plane_w = min_abs_w - (0.5 * support - 1.0) / w_scale;
        p->max_plane_w = max_abs_w + (0.5 * support - 1.0) / w_scale;
        num_total_w_grids += support - 2;
        p->min_abs_w = min_abs_w; p->max_abs_w = max_abs_w;
        p->num_total_w_grids = num_total_w_grids;
        p->w_scale = w_scale;
        p->inv_w_range = p->max_plane_w - p->min_plane_w;
    }
    else // es_fft.cpp:374-383
    {
        p->min_abs_w = p->max_abs_w = 0.0;
        p->min_plane_w = p->max_plane_w = 0.0;
        p->num_total_w_grids = 1;
        p->inv_w_range = 1.0;
        p->w_scale = 1.0;
    }
    p->inv_w_scale = 1.0 / p->w_scale;
    if (compute_prec == 4)
    {
        // The float kernels receive float copies of every scalar
        // (es_fft.cpp:315,340,343,386-390,424); keep the rounded values.
        p->beta = (float) p->beta;
        p->pixel_size = (float) p->pixel_size;
        p->uv_scale = (float) p->uv_scale;
        p->w_scale = (float) p->w_scale;
        p->inv_w_scale = (float) p->inv_w_scale;
        p->min_plane_w = (float) p->min_plane_w;
        p->max_plane_w = (float) p->max_plane_w;
        p->inv_w_range = (float) p->inv_w_range;
    }
    // es_fft.cpp:399-424. NB tables are generated from the UNROUNDED beta.
    p->qk.assign(QUADRATURE_SUPPORT_BOUND, 0.0);
    p->qn.assign(QUADRATURE_SUPPORT_BOUND, 0.0);
    p->qw.assign(QUADRATURE_SUPPORT_BOUND, 0.0);
    p->cc.assign(image_size / 2 + 1, 0.0);
    sdp_generate_gauss_legendre_conv_kernel(image_size, grid_size, support,
            beta, p->qk.data(), p->qn.data(), p->qw.data(), p->cc.data()
    );
    const uint32_t np = (uint32_t) (int(1.5 * support + 2.0));
    double norm = 0.0;
    for (uint32_t i = 0; i < np; i++) norm += p->qk[i] * p->qw[i];
    norm *= (double) support;
    p->conv_corr_norm_factor = (compute_prec == 4) ? (double) (float) norm : norm;
    p->qk_f.assign(p->qk.begin(), p->qk.end());
    p->qn_f.assign(p->qn.begin(), p->qn.end());
    p->qw_f.assign(p->qw.begin(), p->qw.end());
    p->cc_f.assign(p->cc.begin(), p->cc.end());
    return p;
}

void oref_info(const void* plan, oref_PlanInfo* o)
{
    const Plan* p = (const Plan*) plan;
    o->grid_size = p->grid_size; o->support = p->support;
    o->num_total_w_grids = p->num_total_w_grids;
    o->image_size = p->image_size;
    o->beta = p->beta; o->pixel_size = p->pixel_size;
    o->uv_scale = p->uv_scale; o->w_scale = p->w_scale;
    o->inv_w_scale = p->inv_w_scale; o->min_plane_w = p->min_plane_w;
    o->max_plane_w = p->max_plane_w;
    o->conv_corr_norm_factor = p->conv_corr_norm_factor;
    o->min_abs_w = p->min_abs_w; o->max_abs_w = p->max_abs_w;
    o->inv_w_range = p->inv_w_range;
}

// qk/qn/qw: 32 doubles each; cc: image_size/2+1 doubles.
void oref_tables(const void* plan, double* qk, double* qn, double* qw,
        double* cc)
{
    const Plan* p = (const Plan*) plan;
    std::memcpy(qk, p->qk.data(), p->qk.size() * sizeof(double));
    std::memcpy(qn, p->qn.data(), p->qn.size() * sizeof(double));
    std::memcpy(qw, p->qw.data(), p->qw.size() * sizeof(double));
    std::memcpy(cc, p->cc.data(), p->cc.size() * sizeof(double));
}

// Arrays are in the plan's compute precision. dirty is accumulated into and
// then corrected, as in the reference. Planes [plane_begin, plane_end) only
// (plane_end < 0: all). fast != 0 skips visibilities that cannot touch the
// plane (identical results; the literal reference schedule is fast == 0).
// timings (may be NULL): seconds in {scatter, fft, screen, correction}.
int oref_grid(const void* plan, const void* uvw, const void* freq,
        const void* vis, const void* weight, void* dirty, int plane_begin,
        int plane_end, int fast, int nthreads, double* timings)
{
    const Plan* p = (const Plan*) plan;
    if (nthreads > 0) omp_set_num_threads(nthreads);
    else nthreads = omp_get_max_threads();
    if (p->compute_prec == 8)
        return run_grid<double>(p, (const double*) uvw, (const double*) freq,
                vis, (const double*) weight, (double*) dirty, plane_begin,
                plane_end, fast, nthreads, timings
        );
    return run_grid<float>(p, (const float*) uvw, (const float*) freq, vis,
            (const float*) weight, (float*) dirty, plane_begin, plane_end,
            fast, nthreads, timings
    );
}

// vis is accumulated into (+=); dirty is corrected in place (side effect of
// the reference, es_fft.cpp:789-825); weight is not read.
int oref_degrid(const void* plan, const void* uvw, const void* freq, void* vis,
        const void* weight, void* dirty, int plane_begin, int plane_end,
        int fast, int nthreads, double* timings)
{
    const Plan* p = (const Plan*) plan;
    if (nthreads > 0) omp_set_num_threads(nthreads);
    else nthreads = omp_get_max_threads();
    if (p->compute_prec == 8)
        return run_degrid<double>(p, (const double*) uvw, (const double*) freq,
                vis, (const double*) weight, (double*) dirty, plane_begin,
                plane_end, fast, nthreads, timings
        );
    return run_degrid<float>(p, (const float*) uvw, (const float*) freq, vis,
            (const float*) weight, (float*) dirty, plane_begin, plane_end,
            fast, nthreads, timings
    );
}

void oref_free(void* plan)
{
    delete (Plan*) plan;
}

int oref_max_threads(void)
{
    return omp_get_max_threads();
}

} // extern "C"
