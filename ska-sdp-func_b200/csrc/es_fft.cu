/*
 * Pruned w-plane FFT of the ES gridder for sm_100a: the 2-D transform of one
 * G x G w-plane, of which only the N x N image footprint is needed (gridding)
 * or non-zero (degridding), as two passes of whole-line FFTs held in shared
 * memory (es_fft_core.h), fused with what surrounds them:
 *
 *   gridding   pass 1  es_fft_grid_rows_kernel: rows of the plane (along v)
 *                      -> inverse FFT -> only the N needed outputs are stored,
 *                      transposed, into strip[v'][u]  (N x G)
 *              pass 2  es_fft_grid_cols_screen_kernel: rows of the strip
 *                      (along u) -> inverse FFT -> the N needed outputs are
 *                      multiplied by the w-screen and summed over all planes
 *                      of the resident batch in registers -> imageT[col][row]
 *   degridding pass 1  es_fft_degrid_cols_kernel: image column x w-screen,
 *                      zero-padded -> forward FFT (along u) -> strip[v'][u]
 *              pass 2  es_fft_degrid_rows_kernel: strip column (N values),
 *                      zero-padded -> forward FFT (along v) -> plane row
 *
 * Replaces, per plane, the reference's cufftExecC2C/Z2Z on the full plane
 * (sdp_fft.cpp:425-456) + apply_w_screen_and_sum / reverse_w_screen_to_stack
 * (kernels.cu:429-679) with 2 G^2 + 2 G N (+ image) element traversals
 * instead of ~10 G^2 + 2 N^2 (cuFFT's 2-D plan needs five passes on these
 * 11-smooth sizes, profiles/r01_launches_c2_grid_v1.csv).
 * The arithmetic is the same DFT; the w-screen maths is that of
 * es_kernels.cu (phase reduced in fp64).
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <cuda_runtime.h>

#include "es_device.h"
#include "es_fft_core.h"
#include "es_fft_plan.h"

#define ES_FFT_THREADS 512 /* upper bound of the CTA size */
#define ES_FFT_MAX_SMEM (227 * 1024)
#define ES_FFT_SMEM_PER_SM (228 * 1024)

/* CTA shape of one kernel family: several small CTAs per SM overlap the
 * load / butterfly / store phases of different lines. */
struct EsFftCfg
{
    int32_t threads, lines;
    size_t smem_bytes;
};

struct EsFft
{
    EsFftLine line;
    void* d_tw;       /* T2[num_twiddles] */
    int32_t* d_perm;  /* [n] */
    int32_t grid_size, image_size, is_double;
    int32_t pitch;
    EsFftCfg rows;    /* lines = adjacent plane rows (grid pass 1, degrid pass 2) */
    EsFftCfg cols;    /* lines = planes of the batch (degrid pass 1) */
    EsFftCfg screen;  /* lines = planes of the batch (grid pass 2) */
    int32_t max_pairs; /* pixel pairs per thread in the screen kernel */
};

namespace {

template<typename T> struct Cx;
template<> struct Cx<float> { typedef float2 T2; };
template<> struct Cx<double> { typedef double2 T2; };

template<int SGN, bool DIT, typename T, typename T2>
__device__ __forceinline__ void fft_lines(T2* x, const EsFftLine& pl,
        const T2* tw, int nlines, int pitch)
{
    for (int s = 0; s < pl.num_stages; ++s)
    {
        const int stage = DIT ? pl.num_stages - 1 - s : s;
        es_fft_pass<SGN, DIT, T>(x, pl, stage, tw, nlines, pitch,
                (int) threadIdx.x, (int) blockDim.x);
        __syncthreads();
    }
}

/* ---- TMA 1-D bulk copies global -> shared, completion on an mbarrier ---- */

__device__ __forceinline__ uint32_t smem_u32(const void* p)
{
    return (uint32_t) __cvta_generic_to_shared(p);
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;"
            :: "r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
            :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}

__device__ __forceinline__ void bulk_load(void* dst_smem, const void* src,
        uint32_t bytes, uint64_t* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes"
            " [%0], [%1], %2, [%3];"
            :: "r"(smem_u32(dst_smem)), "l"(src), "r"(bytes), "r"(smem_u32(bar))
            : "memory");
}

__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    asm volatile(
            "{\n"
            ".reg .pred P1;\n"
            "LAB_WAIT:\n"
            "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
            "@P1 bra DONE;\n"
            "bra LAB_WAIT;\n"
            "DONE:\n"
            "}"
            :: "r"(smem_u32(bar)), "r"(parity) : "memory");
}

/* Thread 0 queues one bulk copy per line (line_bytes each, a multiple of 16)
 * into x[l * pitch]; every thread then waits for all of them. The caller
 * guarantees that no thread still touches x (barrier) before the call. */
template<typename T2>
__device__ __forceinline__ void load_lines_tma(T2* x, int pitch,
        const T2* src, size_t src_pitch, int nl, uint32_t line_bytes,
        uint64_t* bar, uint32_t parity, bool reuse)
{
    if (threadIdx.x == 0)
    {
        if (reuse) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        mbar_expect_tx(bar, line_bytes * (uint32_t) nl);
        for (int l = 0; l < nl; ++l)
            bulk_load(x + (size_t) l * pitch, src + (size_t) l * src_pitch,
                    line_bytes, bar);
    }
    mbar_wait(bar, parity);
}

template<typename T2>
__device__ __forceinline__ void load_twiddles(T2* tw, const T2* tw_g, int count)
{
    for (int i = threadIdx.x; i < count; i += blockDim.x) tw[i] = tw_g[i];
}

/* n - 1 = -(l^2 + m^2) / (sqrt(1 - l^2 - m^2) + 1), kernels.cu:110-123.
 * fp32 data: reciprocal square root / reciprocal seeded in fp32 and refined
 * by one Newton step in fp64 (relative error ~2e-14, i.e. < 1e-10 turns of
 * phase at |w| = 1e5 wavelengths); fp64 data: exact sqrt and divide. */
template<typename T> __device__ __forceinline__ double es_nm1(double sos);

template<> __device__ __forceinline__ double es_nm1<float>(double sos)
{
    const double a = 1.0 - sos;
    double y = (double) rsqrtf((float) a);
    y = y * (1.5 - 0.5 * a * y * y);
    const double d = a * y + 1.0; /* sqrt(a) + 1 */
    double r = (double) __frcp_rn((float) d);
    r = r * (2.0 - d * r);
    return -sos * r;
}

template<> __device__ __forceinline__ double es_nm1<double>(double sos)
{
    return (-sos) / (sqrt(1.0 - sos) + 1.0);
}

/* (cos, sin)(2 pi sign w nm1): phase reduced to [-0.5, 0.5] turns in fp64. */
template<typename T>
__device__ __forceinline__ void es_screen_phasor(double w, double nm1,
        double sign, T& re, T& im);

template<> __device__ __forceinline__ void es_screen_phasor<float>(double w,
        double nm1, double sign, float& re, float& im)
{
    double t = w * nm1;
    t -= rint(t);
    sincospif((float) (2.0 * sign * t), &im, &re);
}

template<> __device__ __forceinline__ void es_screen_phasor<double>(double w,
        double nm1, double sign, double& re, double& im)
{
    double t = w * nm1;
    t -= rint(t);
    sincospi(2.0 * sign * t, &im, &re);
}

/* ------------------------------------------------------ gridding pass 1 */

template<typename T>
__global__ void __launch_bounds__(ES_FFT_THREADS, 1) es_fft_grid_rows_kernel(
        const EsFftLine pl, const typename Cx<T>::T2* __restrict__ tw_g,
        const int32_t* __restrict__ perm,
        const typename Cx<T>::T2* __restrict__ planes,
        typename Cx<T>::T2* __restrict__ strip, const int image_size,
        const int lines_per_cta, const int pitch)
{
    typedef typename Cx<T>::T2 T2;
    extern __shared__ __align__(16) unsigned char es_fft_smem[];
    T2* x = (T2*) es_fft_smem;
    T2* tw = x + (size_t) lines_per_cta * pitch;
    const int gsz = pl.n, tid = threadIdx.x, nthr = blockDim.x;
    const int u0 = blockIdx.x * lines_per_cta;
    const int nl = min(lines_per_cta, gsz - u0);
    const size_t plane_cells = (size_t) gsz * gsz;
    const T2* src = planes + (size_t) blockIdx.y * plane_cells +
            (size_t) u0 * gsz;
    __shared__ uint64_t bar;
    if (tid == 0) mbar_init(&bar, 1);
    load_twiddles(tw, tw_g, pl.num_twiddles);
    __syncthreads();

    /* rows are contiguous in the plane: one TMA bulk copy per row */
    constexpr int VEC = 16 / (int) sizeof(T2);
    const int vpl = gsz / VEC; /* grid_size is even */
    load_lines_tma(x, pitch, src, (size_t) gsz, nl,
            (uint32_t) ((size_t) gsz * sizeof(T2)), &bar, 0u, false);
    int any = 0;
    for (int l = 0; l < nl; ++l)
    {
        const uint4* d4 = reinterpret_cast<const uint4*>(x + (size_t) l * pitch);
#pragma unroll 4
        for (int c = tid; c < vpl; c += nthr)
        {
            const uint4 v = d4[c];
            any |= (int) ((v.x | v.y | v.z | v.w) != 0u);
        }
    }
    any = __syncthreads_or(any);

    const int n = image_size, k0 = gsz / 2 - n / 2;
    T2* dst = strip + (size_t) blockIdx.y * (size_t) n * gsz + u0;
    if (!any)
    {
        /* empty rows transform to zeros */
        T2 zero; zero.x = (T) 0; zero.y = (T) 0;
        for (int e = tid; e < n * nl; e += nthr)
        {
            const int idx = e / nl, l = e - idx * nl;
            dst[(size_t) idx * gsz + l] = zero;
        }
        return;
    }
    fft_lines<1, false, T>(x, pl, tw, nl, pitch);
    if (nl == 1)
    {
#pragma unroll 4
        for (int idx = tid; idx < n; idx += nthr)
            dst[(size_t) idx * gsz] = x[perm[k0 + idx]];
        return;
    }
#pragma unroll 4
    for (int e = tid; e < n * nl; e += nthr)
    {
        const int idx = e / nl, l = e - idx * nl;
        dst[(size_t) idx * gsz + l] = x[(size_t) l * pitch + perm[k0 + idx]];
    }
}

/* ------------------------------------------------------ gridding pass 2 */

/* One CTA per image column `col` (v' = k0 + col): for every plane of the
 * batch, FFT the strip row along u, pick the N outputs u' = k0 + row, apply
 * the w-screen (kernels.cu:444-515) and accumulate over the planes in
 * registers; one read-modify-write of imageT[col][row] at the end.
 * A thread owns the row offsets iu = tid + m * THREADS (m < MAXP), i.e. the
 * mirror pair of rows N/2 + iu and N/2 - iu, which share one phasor. */
template<typename T, int MAXP>
__global__ void __launch_bounds__(ES_FFT_THREADS, 1)
es_fft_grid_cols_screen_kernel(const EsFftLine pl,
        const typename Cx<T>::T2* __restrict__ tw_g,
        const int32_t* __restrict__ perm, const EsGeom g,
        const typename Cx<T>::T2* __restrict__ strip, const int first_plane,
        const int nplanes, T* __restrict__ image_t, const int lines_per_cta,
        const int pitch)
{
    typedef typename Cx<T>::T2 T2;
    extern __shared__ __align__(16) unsigned char es_fft_smem[];
    T2* x = (T2*) es_fft_smem;
    T2* tw = x + (size_t) lines_per_cta * pitch;
    const int gsz = pl.n, tid = threadIdx.x, nthr = blockDim.x;
    const int n = g.image_size, half = n / 2;
    const int col = blockIdx.x;
    const int jv = col - half;
    __shared__ uint64_t bar;
    if (tid == 0) mbar_init(&bar, 1);
    load_twiddles(tw, tw_g, pl.num_twiddles);

    T acc_p[MAXP], acc_m[MAXP];
#pragma unroll
    for (int m = 0; m < MAXP; ++m) { acc_p[m] = (T) 0; acc_m[m] = (T) 0; }
    const double mm = g.pixel_size * (double) jv;
    const double m2 = mm * mm;
    constexpr int VEC = 16 / (int) sizeof(T2);
    const int vpl = gsz / VEC;
    const size_t strip_plane = (size_t) n * gsz;

    for (int q0 = 0; q0 < nplanes; q0 += lines_per_cta)
    {
        const int nl = min(lines_per_cta, nplanes - q0);
        __syncthreads(); /* barrier initialised / previous group's reads done */
        load_lines_tma(x, pitch, strip + (size_t) q0 * strip_plane +
                (size_t) col * gsz, strip_plane, nl,
                (uint32_t) ((size_t) gsz * sizeof(T2)), &bar,
                (uint32_t) ((q0 / lines_per_cta) & 1), q0 > 0);
        fft_lines<1, false, T>(x, pl, tw, nl, pitch);
#pragma unroll
        for (int m = 0; m < MAXP; ++m)
        {
            const int iu = tid + m * nthr;
            if (iu <= half)
            {
                const int pos_p = perm[gsz / 2 + iu];
                const int pos_m = perm[gsz / 2 - iu];
                const double ll = g.pixel_size * (double) iu;
                const double nm1 = es_nm1<T>(ll * ll + m2);
                for (int l = 0; l < nl; ++l)
                {
                    T cr = (T) 1, ci = (T) 0;
                    if (g.do_wstacking)
                    {
                        const double w = (double) (first_plane + q0 + l) *
                                g.inv_w_scale + g.min_plane_w;
                        es_screen_phasor<T>(w, nm1, -1.0, cr, ci);
                    }
                    const T2 vp = x[(size_t) l * pitch + pos_p];
                    const T2 vm = x[(size_t) l * pitch + pos_m];
                    acc_p[m] += vp.x * cr - vp.y * ci;
                    acc_m[m] += vm.x * cr - vm.y * ci;
                }
            }
        }
    }
    T* out = image_t + (size_t) col * n + half;
#pragma unroll
    for (int m = 0; m < MAXP; ++m)
    {
        const int iu = tid + m * nthr;
        if (iu <= half)
        {
            T scale = (T) 1;
            if (g.do_wstacking)
            {
                const double ll = g.pixel_size * (double) iu;
                const double nm1 = es_nm1<T>(ll * ll + m2);
                scale = (T) (1.0 / (nm1 + 1.0));
            }
            if ((iu + jv) & 1) scale = -scale;
            if (iu < half) out[iu] += acc_p[m] * scale;
            if (iu > 0) out[-iu] += acc_m[m] * scale;
        }
    }
}

/* ---------------------------------------------------- degridding pass 1 */

/* One CTA per (image column, group of planes): line[u = G/2 + iu] =
 * imageT[col][N/2 + iu] * (-1)^(iu + jv) * w-screen(+) / n, zero elsewhere
 * (kernels.cu:580-676), forward FFT along u, all G outputs stored to
 * strip[plane][col][u]. */
template<typename T>
__global__ void __launch_bounds__(ES_FFT_THREADS, 1) es_fft_degrid_cols_kernel(
        const EsFftLine pl, const typename Cx<T>::T2* __restrict__ tw_g,
        const int32_t* __restrict__ perm, const EsGeom g,
        const T* __restrict__ image_t, const int first_plane,
        const int nplanes, typename Cx<T>::T2* __restrict__ strip,
        const int lines_per_cta, const int pitch)
{
    typedef typename Cx<T>::T2 T2;
    extern __shared__ __align__(16) unsigned char es_fft_smem[];
    T2* x = (T2*) es_fft_smem;
    T2* tw = x + (size_t) lines_per_cta * pitch;
    const int gsz = pl.n, tid = threadIdx.x, nthr = blockDim.x;
    const int n = g.image_size, half = n / 2;
    const int col = blockIdx.x;
    const int jv = col - half;
    const int q0 = blockIdx.y * lines_per_cta;
    const int nl = min(lines_per_cta, nplanes - q0);
    load_twiddles(tw, tw_g, pl.num_twiddles);

    constexpr int VEC = 16 / (int) sizeof(T2);
    const int vpl = gsz / VEC;
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int l = 0; l < nl; ++l)
    {
        float4* d4 = reinterpret_cast<float4*>(x + (size_t) l * pitch);
        for (int c = tid; c < vpl; c += nthr) d4[c] = zero4;
    }
    __syncthreads();
    const double mm = g.pixel_size * (double) jv;
    const double m2 = mm * mm;
    const T* img = image_t + (size_t) col * n + half;
    for (int iu = tid; iu <= half; iu += nthr)
    {
        T d_p = (iu < half) ? img[iu] : (T) 0;
        T d_m = (iu > 0) ? img[-iu] : (T) 0;
        const int pos_p = perm[gsz / 2 + iu];
        const int pos_m = perm[gsz / 2 - iu];
        const double ll = g.pixel_size * (double) iu;
        const double nm1 = es_nm1<T>(ll * ll + m2);
        T scale = (T) 1;
        if (g.do_wstacking) scale = (T) (1.0 / (nm1 + 1.0));
        if ((iu + jv) & 1) scale = -scale;
        d_p *= scale;
        d_m *= scale;
        for (int l = 0; l < nl; ++l)
        {
            T cr = (T) 1, ci = (T) 0;
            if (g.do_wstacking)
            {
                const double w = (double) (first_plane + q0 + l) *
                        g.inv_w_scale + g.min_plane_w;
                es_screen_phasor<T>(w, nm1, 1.0, cr, ci);
            }
            T2 o;
            if (iu < half)
            {
                o.x = cr * d_p; o.y = ci * d_p;
                x[(size_t) l * pitch + pos_p] = o;
            }
            if (iu > 0)
            {
                o.x = cr * d_m; o.y = ci * d_m;
                x[(size_t) l * pitch + pos_m] = o;
            }
        }
    }
    __syncthreads();
    fft_lines<-1, true, T>(x, pl, tw, nl, pitch);
    const size_t strip_plane = (size_t) n * gsz;
    for (int l = 0; l < nl; ++l)
    {
        float4* d4 = reinterpret_cast<float4*>(strip +
                (size_t) (q0 + l) * strip_plane + (size_t) col * gsz);
        const float4* s4 = reinterpret_cast<const float4*>(x + (size_t) l * pitch);
        for (int c = tid; c < vpl; c += nthr) d4[c] = s4[c];
    }
}

/* ---------------------------------------------------- degridding pass 2 */

/* One CTA per (group of adjacent plane rows u, plane): gather the N strip
 * values of each row (strip[col][u], col < N), zero-pad to G at v = k0 + col,
 * forward FFT along v, store the full plane row. */
template<typename T>
__global__ void __launch_bounds__(ES_FFT_THREADS, 1) es_fft_degrid_rows_kernel(
        const EsFftLine pl, const typename Cx<T>::T2* __restrict__ tw_g,
        const int32_t* __restrict__ perm,
        const typename Cx<T>::T2* __restrict__ strip,
        typename Cx<T>::T2* __restrict__ planes, const int image_size,
        const int lines_per_cta, const int pitch)
{
    typedef typename Cx<T>::T2 T2;
    extern __shared__ __align__(16) unsigned char es_fft_smem[];
    T2* x = (T2*) es_fft_smem;
    T2* tw = x + (size_t) lines_per_cta * pitch;
    const int gsz = pl.n, tid = threadIdx.x, nthr = blockDim.x;
    const int u0 = blockIdx.x * lines_per_cta;
    const int nl = min(lines_per_cta, gsz - u0);
    const int n = image_size, k0 = gsz / 2 - n / 2;
    load_twiddles(tw, tw_g, pl.num_twiddles);

    constexpr int VEC = 16 / (int) sizeof(T2);
    const int vpl = gsz / VEC;
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int l = 0; l < nl; ++l)
    {
        float4* d4 = reinterpret_cast<float4*>(x + (size_t) l * pitch);
        for (int c = tid; c < vpl; c += nthr) d4[c] = zero4;
    }
    __syncthreads();
    const T2* src = strip + (size_t) blockIdx.y * (size_t) n * gsz + u0;
    if (nl == 1)
    {
        for (int idx = tid; idx < n; idx += nthr)
            x[perm[k0 + idx]] = src[(size_t) idx * gsz];
    }
    else
    {
        for (int e = tid; e < n * nl; e += nthr)
        {
            const int idx = e / nl, l = e - idx * nl;
            x[(size_t) l * pitch + perm[k0 + idx]] = src[(size_t) idx * gsz + l];
        }
    }
    __syncthreads();
    fft_lines<-1, true, T>(x, pl, tw, nl, pitch);
    T2* dst = planes + (size_t) blockIdx.y * (size_t) gsz * gsz +
            (size_t) u0 * gsz;
    for (int l = 0; l < nl; ++l)
    {
        float4* d4 = reinterpret_cast<float4*>(dst + (size_t) l * gsz);
        const float4* s4 = reinterpret_cast<const float4*>(x + (size_t) l * pitch);
        for (int c = tid; c < vpl; c += nthr) d4[c] = s4[c];
    }
}

/* ------------------------------------------------------ image transposes */

/* dst[r][c] (+)= src[c][r], square n x n. */
template<typename T, bool ADD>
__global__ void __launch_bounds__(256) es_transpose_kernel(
        const T* __restrict__ src, T* __restrict__ dst, const int n)
{
    __shared__ T t[32][33];
    const int bx = blockIdx.x * 32, by = blockIdx.y * 32;
    for (int k = threadIdx.y; k < 32; k += 8)
    {
        const int r = by + k, c = bx + threadIdx.x;
        if (r < n && c < n) t[k][threadIdx.x] = src[(size_t) r * n + c];
    }
    __syncthreads();
    for (int k = threadIdx.y; k < 32; k += 8)
    {
        const int r = bx + k, c = by + threadIdx.x;
        if (r < n && c < n)
        {
            if (ADD) dst[(size_t) r * n + c] += t[threadIdx.x][k];
            else dst[(size_t) r * n + c] = t[threadIdx.x][k];
        }
    }
}

template<typename T>
int fft_set_smem(const void* kernel, size_t bytes)
{
    return (int) cudaFuncSetAttribute(kernel,
            cudaFuncAttributeMaxDynamicSharedMemorySize, (int) bytes);
}

/* CTA shape of a kernel family (measured on C2, profiles/README.md):
 *  - "rows" kernels store / gather `lines` adjacent elements per strip row, so
 *    more lines per CTA = fuller 32-byte sectors: take as many as fit (<= 64
 *    bytes), with up to 512 threads;
 *  - "plane" kernels (lines = planes of the batch) are coalesced either way:
 *    two lines and 256 threads leave room for two CTAs per SM, whose load,
 *    butterfly and store phases overlap. min_pixels > 0: the screen kernel
 *    keeps ceil(min_pixels / threads) <= 17 pixel pairs per thread. */
int choose_cfg(const EsFftLine* line, size_t line_bytes, size_t tw_bytes,
        int max_lines, int want_lines, int min_pixels, EsFftCfg* out)
{
    int fit = (int) ((ES_FFT_MAX_SMEM - tw_bytes - 64) / line_bytes);
    if (fit < 1) return 0;
    if (fit > max_lines) fit = max_lines;
    int lines = 1;
    while (2 * lines <= fit && 2 * lines <= want_lines) lines *= 2;
    /* at least two butterflies of the widest pass per thread */
    const int widest = line->stage[0].radix;
    int threads = ((lines * (line->n / widest) / 2 + 31) / 32) * 32;
    if (threads < 128) threads = 128;
    const int cap = want_lines >= 4 ? ES_FFT_THREADS : 256;
    if (threads > cap) threads = cap;
    while (min_pixels > 0 && (min_pixels + threads - 1) / threads > 17 &&
            threads < ES_FFT_THREADS)
        threads += 32;
    if (min_pixels > 0 && (min_pixels + threads - 1) / threads > 17) return 0;
    out->threads = threads;
    out->lines = lines;
    out->smem_bytes = (size_t) lines * line_bytes + tw_bytes;
    return 1;
}

} // namespace

/* ================================================================ C ABI */

extern "C" {

int es_fft_create(int grid_size, int image_size, int is_double, EsFft** out)
{
    *out = 0;
    if (getenv("SKA_SDP_FUNC_ES_USE_CUFFT")) return -1;
    EsFft* f = (EsFft*) calloc(1, sizeof(EsFft));
    if (!f) return (int) cudaErrorMemoryAllocation;
    if (es_fft_line_plan(grid_size, is_double ? 8 : 16, &f->line) ||
            (grid_size & 1) || (image_size & 1) || image_size > grid_size)
    {
        free(f);
        return -1; /* unsupported size: caller falls back to cuFFT */
    }
    f->grid_size = grid_size;
    f->image_size = image_size;
    f->is_double = is_double;
    const size_t cplx = is_double ? 16 : 8;
    /* pitch: lines 2 elements apart in bank space (16-byte aligned) */
    f->pitch = grid_size + 2;
    const size_t tw_bytes = (size_t) f->line.num_twiddles * cplx;
    const size_t line_bytes = (size_t) f->pitch * cplx;
    const int half1 = image_size / 2 + 1;
    const int cap = is_double ? 4 : 8; /* 64 bytes of adjacent lines */
    const int ok = choose_cfg(&f->line, line_bytes, tw_bytes, cap, cap, 0,
                    &f->rows) &&
            choose_cfg(&f->line, line_bytes, tw_bytes, cap, 2, 0, &f->cols) &&
            choose_cfg(&f->line, line_bytes, tw_bytes, cap, 2, half1,
                    &f->screen);
    if (!ok)
    {
        free(f);
        return -1;
    }
    /* experiments: SKA_SDP_FUNC_ES_FFT_CFG="rows_thr,rows_lines,screen_thr,screen_lines" */
    const char* env = getenv("SKA_SDP_FUNC_ES_FFT_CFG");
    int e[4];
    if (env && sscanf(env, "%d,%d,%d,%d", &e[0], &e[1], &e[2], &e[3]) == 4 &&
            e[0] >= 32 && e[0] <= ES_FFT_THREADS && e[2] >= 32 &&
            e[2] <= ES_FFT_THREADS && e[1] >= 1 && e[3] >= 1 &&
            (size_t) e[1] * line_bytes + tw_bytes <= ES_FFT_MAX_SMEM &&
            (size_t) e[3] * line_bytes + tw_bytes <= ES_FFT_MAX_SMEM &&
            (half1 + e[2] - 1) / e[2] <= 17)
    {
        f->rows.threads = f->cols.threads = e[0];
        f->rows.lines = f->cols.lines = e[1];
        f->rows.smem_bytes = f->cols.smem_bytes = (size_t) e[1] * line_bytes + tw_bytes;
        f->screen.threads = e[2];
        f->screen.lines = e[3];
        f->screen.smem_bytes = (size_t) e[3] * line_bytes + tw_bytes;
    }
    const int pairs = (half1 + f->screen.threads - 1) / f->screen.threads;
    f->max_pairs = pairs <= 4 ? 4 : (pairs <= 9 ? 9 : 17);
    const size_t sm = f->rows.smem_bytes > f->screen.smem_bytes ?
            f->rows.smem_bytes : f->screen.smem_bytes;
    if (getenv("SKA_SDP_FUNC_ES_DEBUG"))
    {
        fprintf(stderr, "es_fft: n=%d stages", grid_size);
        for (int k = 0; k < f->line.num_stages; ++k)
            fprintf(stderr, " %d", f->line.stage[k].radix);
        fprintf(stderr, "; rows %d thr x %d lines (%zu B), screen %d thr x %d "
                "lines (%zu B), pairs %d\n", f->rows.threads, f->rows.lines,
                f->rows.smem_bytes, f->screen.threads, f->screen.lines,
                f->screen.smem_bytes, f->max_pairs);
    }

    const int nt = f->line.num_twiddles;
    double* tw = (double*) malloc(sizeof(double) * 2 * (size_t) nt);
    int32_t* perm = (int32_t*) malloc(sizeof(int32_t) * (size_t) grid_size);
    void* tw_t = malloc(cplx * (size_t) nt);
    int err = 0;
    if (!tw || !perm || !tw_t) err = (int) cudaErrorMemoryAllocation;
    if (!err)
    {
        es_fft_line_tables(&f->line, tw, perm);
        for (int i = 0; i < 2 * nt; ++i)
        {
            if (is_double) ((double*) tw_t)[i] = tw[i];
            else ((float*) tw_t)[i] = (float) tw[i];
        }
        err = (int) cudaMalloc(&f->d_tw, cplx * (size_t) nt);
        if (!err) err = (int) cudaMalloc((void**) &f->d_perm,
                sizeof(int32_t) * (size_t) grid_size);
        if (!err) err = (int) cudaMemcpy(f->d_tw, tw_t, cplx * (size_t) nt,
                cudaMemcpyHostToDevice);
        if (!err) err = (int) cudaMemcpy(f->d_perm, perm,
                sizeof(int32_t) * (size_t) grid_size, cudaMemcpyHostToDevice);
    }
    free(tw); free(perm); free(tw_t);
    if (!err && is_double)
    {
        err = fft_set_smem<double>((const void*) es_fft_grid_rows_kernel<double>, sm);
        if (!err) err = fft_set_smem<double>((const void*) es_fft_grid_cols_screen_kernel<double, 4>, sm);
        if (!err) err = fft_set_smem<double>((const void*) es_fft_grid_cols_screen_kernel<double, 9>, sm);
        if (!err) err = fft_set_smem<double>((const void*) es_fft_grid_cols_screen_kernel<double, 17>, sm);
        if (!err) err = fft_set_smem<double>((const void*) es_fft_degrid_cols_kernel<double>, sm);
        if (!err) err = fft_set_smem<double>((const void*) es_fft_degrid_rows_kernel<double>, sm);
    }
    else if (!err)
    {
        err = fft_set_smem<float>((const void*) es_fft_grid_rows_kernel<float>, sm);
        if (!err) err = fft_set_smem<float>((const void*) es_fft_grid_cols_screen_kernel<float, 4>, sm);
        if (!err) err = fft_set_smem<float>((const void*) es_fft_grid_cols_screen_kernel<float, 9>, sm);
        if (!err) err = fft_set_smem<float>((const void*) es_fft_grid_cols_screen_kernel<float, 17>, sm);
        if (!err) err = fft_set_smem<float>((const void*) es_fft_degrid_cols_kernel<float>, sm);
        if (!err) err = fft_set_smem<float>((const void*) es_fft_degrid_rows_kernel<float>, sm);
    }
    if (err)
    {
        (void) cudaGetLastError();
        es_fft_destroy(f);
        return err;
    }
    *out = f;
    return 0;
}

void es_fft_destroy(EsFft* f)
{
    if (!f) return;
    if (f->d_tw) (void) cudaFree(f->d_tw);
    if (f->d_perm) (void) cudaFree(f->d_perm);
    free(f);
}

size_t es_fft_strip_bytes(const EsFft* f, int nplanes)
{
    return (size_t) nplanes * (size_t) f->image_size * (size_t) f->grid_size *
            (f->is_double ? 16 : 8);
}

int es_fft_grid_rows(const EsFft* f, const void* planes, int nplanes,
        void* strip, es_stream_t stream)
{
    cudaStream_t s = (cudaStream_t) stream;
    const EsFftCfg& c = f->rows;
    dim3 grid((unsigned) ((f->grid_size + c.lines - 1) / c.lines),
            (unsigned) nplanes);
    if (f->is_double)
        es_fft_grid_rows_kernel<double><<<grid, c.threads, c.smem_bytes,
                s>>>(f->line, (const double2*) f->d_tw, f->d_perm,
                (const double2*) planes, (double2*) strip, f->image_size,
                c.lines, f->pitch);
    else
        es_fft_grid_rows_kernel<float><<<grid, c.threads, c.smem_bytes,
                s>>>(f->line, (const float2*) f->d_tw, f->d_perm,
                (const float2*) planes, (float2*) strip, f->image_size,
                c.lines, f->pitch);
    return (int) cudaGetLastError();
}

#define ES_SCREEN_LAUNCH(T, T2, MP) \
    es_fft_grid_cols_screen_kernel<T, MP><<<grid, f->screen.threads, \
            f->screen.smem_bytes, s>>>(f->line, (const T2*) f->d_tw, \
            f->d_perm, *g, (const T2*) strip, first_plane, nplanes, \
            (T*) image_t, f->screen.lines, f->pitch)

int es_fft_grid_cols_screen(const EsFft* f, const EsGeom* g, const void* strip,
        int first_plane, int nplanes, void* image_t, es_stream_t stream)
{
    cudaStream_t s = (cudaStream_t) stream;
    dim3 grid((unsigned) f->image_size);
    if (f->is_double)
    {
        if (f->max_pairs == 4) ES_SCREEN_LAUNCH(double, double2, 4);
        else if (f->max_pairs == 9) ES_SCREEN_LAUNCH(double, double2, 9);
        else ES_SCREEN_LAUNCH(double, double2, 17);
    }
    else
    {
        if (f->max_pairs == 4) ES_SCREEN_LAUNCH(float, float2, 4);
        else if (f->max_pairs == 9) ES_SCREEN_LAUNCH(float, float2, 9);
        else ES_SCREEN_LAUNCH(float, float2, 17);
    }
    return (int) cudaGetLastError();
}

int es_fft_degrid_cols(const EsFft* f, const EsGeom* g, const void* image_t,
        int first_plane, int nplanes, void* strip, es_stream_t stream)
{
    cudaStream_t s = (cudaStream_t) stream;
    const EsFftCfg& c = f->cols;
    dim3 grid((unsigned) f->image_size, (unsigned) ((nplanes + c.lines - 1) /
            c.lines));
    if (f->is_double)
        es_fft_degrid_cols_kernel<double><<<grid, c.threads, c.smem_bytes,
                s>>>(f->line, (const double2*) f->d_tw, f->d_perm, *g,
                (const double*) image_t, first_plane, nplanes,
                (double2*) strip, c.lines, f->pitch);
    else
        es_fft_degrid_cols_kernel<float><<<grid, c.threads, c.smem_bytes,
                s>>>(f->line, (const float2*) f->d_tw, f->d_perm, *g,
                (const float*) image_t, first_plane, nplanes,
                (float2*) strip, c.lines, f->pitch);
    return (int) cudaGetLastError();
}

int es_fft_degrid_rows(const EsFft* f, const void* strip, int nplanes,
        void* planes, es_stream_t stream)
{
    cudaStream_t s = (cudaStream_t) stream;
    const EsFftCfg& c = f->rows;
    dim3 grid((unsigned) ((f->grid_size + c.lines - 1) / c.lines),
            (unsigned) nplanes);
    if (f->is_double)
        es_fft_degrid_rows_kernel<double><<<grid, c.threads, c.smem_bytes,
                s>>>(f->line, (const double2*) f->d_tw, f->d_perm,
                (const double2*) strip, (double2*) planes, f->image_size,
                c.lines, f->pitch);
    else
        es_fft_degrid_rows_kernel<float><<<grid, c.threads, c.smem_bytes,
                s>>>(f->line, (const float2*) f->d_tw, f->d_perm,
                (const float2*) strip, (float2*) planes, f->image_size,
                c.lines, f->pitch);
    return (int) cudaGetLastError();
}

int es_launch_transpose(const void* src, void* dst, int n, int is_double,
        int add, es_stream_t stream)
{
    cudaStream_t s = (cudaStream_t) stream;
    dim3 block(32, 8), grid((unsigned) ((n + 31) / 32), (unsigned) ((n + 31) / 32));
    if (is_double)
    {
        if (add) es_transpose_kernel<double, true><<<grid, block, 0, s>>>(
                (const double*) src, (double*) dst, n);
        else es_transpose_kernel<double, false><<<grid, block, 0, s>>>(
                (const double*) src, (double*) dst, n);
    }
    else
    {
        if (add) es_transpose_kernel<float, true><<<grid, block, 0, s>>>(
                (const float*) src, (float*) dst, n);
        else es_transpose_kernel<float, false><<<grid, block, 0, s>>>(
                (const float*) src, (float*) dst, n);
    }
    return (int) cudaGetLastError();
}

} // extern "C"
