/*
 * sm_100a kernels of the Hogbom CLEAN (sdp_hogbom_clean), behind a
 * plain-pointer C launcher (es_clean.h). Replaces the reference's GPU path
 * (src/ska-sdp-func/clean/sdp_hogbom_clean.cpp:281-730 + .cu:13-673: per
 * cycle a multi-stage maximum search, a one-thread model update and a PSF
 * subtraction as separate launches with the peak read back by the host, then
 * an FFT convolution of three padded complex images for the restore step).
 *
 * The minor cycle is a chain of `cycle_limit` dependent steps over an image
 * that fits on chip, so it is bound by LATENCY (launches, host round trips,
 * grid-wide agreement on the peak), not by bandwidth. Here ALL cycles run in
 * ONE cooperative launch:
 *  - every CTA owns a contiguous slice of the residual and keeps it in shared
 *    memory for the whole run (148 CTAs x ~200 KB: images up to ~1900^2 fp64 /
 *    ~2700^2 fp32; larger ones stay in global memory / L2);
 *  - the PSF subtraction of cycle k and the peak search of cycle k + 1 are one
 *    pass over the slice; the CTAs then publish one candidate each, meet at
 *    ONE grid barrier per cycle and each reduces the <= 148 candidates itself
 *    (double-buffered, so no second barrier);
 *  - the threshold test is made by every CTA on the same data: no host round
 *    trip anywhere in the loop.
 * Arithmetic is that of the reference's HOST loop (.cpp:183-231), operation
 * by operation: the peak is the FIRST largest element, the update is
 * r - (gain * peak) * psf with separate roundings (no fused multiply-add) and,
 * for fp32 images, in double with one rounding at the end, as the host's
 * `double highest_value` makes it. Residual and model are therefore
 * bit-identical to the reference's CPU path.
 * Restore step: the model has at most one non-zero pixel per cycle, so the
 * convolution with the CLEAN beam is evaluated directly over the list of
 * those pixels (kept in first-selection order by the loop) instead of through
 * FFTs of padded images; the reference's centring (FFT shift and crop put
 * beam sample (cbeam_dim / 2 - 1 - ...) under the component, sdp_fft_convolution
 * .cpp:127-244) is kept: see beam_offset().
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <cooperative_groups.h>
#include <cuda_runtime.h>

#include "es_clean.h"

namespace cg = cooperative_groups;

#define ES_CLEAN_THREADS 512
#define ES_CLEAN_MAX_CTAS 1024
#define ES_CLEAN_UNROLL 8 /* PSF loads in flight per thread */

namespace {

struct Cand
{
    double val;
    long long idx;
};

/* a beats b: larger value, or equal value at a lower index (the host loop
 * keeps the first of equal maxima, .cpp:189-196) */
__device__ __forceinline__ bool beats(const double av, const long long ai,
        const double bv, const long long bi)
{
    return ai >= 0 && (bi < 0 || av > bv || (av == bv && ai < bi));
}

/* Warp arg-max with the tie rule of beats(), all lanes get the result. The
 * value goes through an order-preserving 64-bit key (sign-magnitude -> biased)
 * so that the hardware warp reductions (redux.sync, 32 bits) can be used:
 * max of the high words, max of the low words among those, min of the indices
 * among those -- three instructions instead of five rounds of four shuffles
 * on the critical path of every cycle. Indices fit 32 bits (n <= 46340). */
__device__ __forceinline__ void warp_best(double& v, long long& i)
{
    const unsigned long long sign = 0x8000000000000000ULL;
    const unsigned long long bits =
            (unsigned long long) __double_as_longlong(v + 0.0); /* -0 -> +0 */
    const unsigned long long key = i < 0 ? 0ULL :
            (bits & sign) ? ~bits : (bits | sign);
    const unsigned hi = (unsigned) (key >> 32), lo = (unsigned) key;
    const unsigned mhi = __reduce_max_sync(0xffffffffu, hi);
    const unsigned mlo = __reduce_max_sync(0xffffffffu, hi == mhi ? lo : 0u);
    const bool winner = i >= 0 && hi == mhi && lo == mlo;
    const unsigned midx = __reduce_min_sync(0xffffffffu,
            winner ? (unsigned) i : 0xffffffffu);
    if (midx == 0xffffffffu)
    {
        v = 0.0;
        i = -1;
        return;
    }
    const unsigned long long best = ((unsigned long long) mhi << 32) | mlo;
    v = __longlong_as_double((long long) ((best & sign) ? (best & ~sign) : ~best));
    i = (long long) midx;
}

struct Header
{
    int cycles_done;
    int num_components;
};

template<typename T>
__device__ __forceinline__ T subtract(const T r, const T gain, const double peak,
        const T psf);

template<>
__device__ __forceinline__ double subtract(const double r, const double gain,
        const double peak, const double psf)
{
    /* residual -= (loop_gain * highest_value * psf), .cpp:222 */
    return __dsub_rn(r, __dmul_rn(__dmul_rn(gain, peak), psf));
}

template<>
__device__ __forceinline__ float subtract(const float r, const float gain,
        const double peak, const float psf)
{
    return (float) __dsub_rn((double) r,
            __dmul_rn(__dmul_rn((double) gain, peak), (double) psf));
}

/* All minor cycles. `slice` = elements per CTA. MODE:
 *  ES_CLEAN_GLOBAL  slices stay in global memory / L2 (images too large for
 *                   the shared memory of the grid), exchange through global
 *                   candidates + grid barrier;
 *  ES_CLEAN_SMEM    the CTA's slice lives in dynamic shared memory between the
 *                   first load and the final store, same exchange;
 *  ES_CLEAN_CLUSTER the grid is ONE thread-block cluster (8 or 16 CTAs; small
 *                   images): slices in shared memory, candidates in each
 *                   CTA's shared memory, read by the others through
 *                   distributed shared memory after the hardware cluster
 *                   barrier -- no global-memory round trip in the exchange. */
#define ES_CLEAN_GLOBAL 0
#define ES_CLEAN_SMEM 1
#define ES_CLEAN_CLUSTER 2
template<typename T, int MODE>
__global__ void __launch_bounds__(ES_CLEAN_THREADS) es_hogbom_cycles_kernel(
        const int64_t n, const int64_t slice, const T* __restrict__ psf,
        const T gain, const T threshold, const int cycle_limit,
        T* __restrict__ residual, T* __restrict__ clean_model,
        Cand* __restrict__ cand, Header* __restrict__ header,
        int* __restrict__ components)
{
    constexpr bool IN_SMEM = MODE != ES_CLEAN_GLOBAL;
    extern __shared__ __align__(16) unsigned char s_raw[];
    __shared__ Cand s_cand[2];
    __shared__ double s_val[ES_CLEAN_THREADS / 32];
    __shared__ long long s_idx[ES_CLEAN_THREADS / 32];
    __shared__ double s_peak;
    __shared__ long long s_peak_idx;

    const int64_t total = n * n;
    const int64_t lo = (int64_t) blockIdx.x * slice;
    const int64_t hi = lo + slice < total ? lo + slice : total;
    const int64_t mine = hi > lo ? hi - lo : 0;
    T* res = IN_SMEM ? reinterpret_cast<T*>(s_raw) : residual + lo;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

    /* Per-element work is kept to a handful of instructions (the kernel is
     * bound by instruction issue: ncu counted 715 warp instructions per cycle
     * and warp for 8 elements per thread before this form): 32-bit local
     * indices, the running peak in the image type with a strict `>` (a thread
     * visits its elements in rising index order, so the first of equal maxima
     * stays), (row, column) advanced by a fixed step. */
    const int ni = (int) n;
    const int mine32 = (int) mine;
    const T lowest = -(T) INFINITY;
    T bestv = lowest;
    int bestk = -1;
    /* first pass: load the slice, find its peak */
    for (int k = threadIdx.x; k < mine32; k += ES_CLEAN_THREADS)
    {
        const T r = residual[lo + k];
        if (IN_SMEM) res[k] = r;
        if (r > bestv) { bestv = r; bestk = k; }
    }
    double best = (double) bestv;
    long long best_idx = bestk < 0 ? -1 : lo + bestk;

    const int first_i = (int) ((lo + threadIdx.x) / n);
    const int first_j = (int) ((lo + threadIdx.x) - (int64_t) first_i * n);
    const int step_i = ES_CLEAN_THREADS / ni, step_j = ES_CLEAN_THREADS % ni;
    int cycle = 0, num_components = 0;
    for (;; ++cycle)
    {
        /* CTA candidate -> cand[cycle & 1][cta] */
        warp_best(best, best_idx);
        if (lane == 0) { s_val[warp] = best; s_idx[warp] = best_idx; }
        __syncthreads();
        if (warp == 0)
        {
            double v = lane < ES_CLEAN_THREADS / 32 ? s_val[lane] : 0.0;
            long long i = lane < ES_CLEAN_THREADS / 32 ? s_idx[lane] : -1;
            warp_best(v, i);
            if (lane == 0)
            {
                Cand c;
                c.val = v;
                c.idx = i;
                if (MODE == ES_CLEAN_CLUSTER)
                    s_cand[cycle & 1] = c;
                else
                    cand[(size_t) (cycle & 1) * gridDim.x + blockIdx.x] = c;
            }
        }
        /* all candidates of this cycle published: the one meeting point of a
         * cycle (cooperative_groups' grid barrier; a hand-rolled counter
         * barrier measured the same, profiles/r02_logs/clean_barrier_job31.txt) */
        if (MODE == ES_CLEAN_CLUSTER)
            cg::this_cluster().sync();
        else
            cg::this_grid().sync();
        /* every CTA reduces all candidates to the same peak */
        if (warp == 0)
        {
            double v = 0.0;
            long long i = -1;
            if (MODE == ES_CLEAN_CLUSTER)
            {
                if (lane < (int) gridDim.x)
                {
                    const Cand* theirs = cg::this_cluster().map_shared_rank(
                            &s_cand[cycle & 1], lane);
                    v = theirs->val;
                    i = theirs->idx;
                }
            }
            else
            {
                const Cand* all = cand + (size_t) (cycle & 1) * gridDim.x;
                for (unsigned c0 = lane; c0 < gridDim.x; c0 += 32 * 8)
                {
                    /* written by other SMs in this cycle: read past L1, eight
                     * loads in flight */
                    double ov[8];
                    long long oi[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u)
                    {
                        const unsigned c = c0 + 32 * u;
                        oi[u] = -1;
                        ov[u] = 0.0;
                        if (c < gridDim.x)
                        {
                            ov[u] = __ldcg(&all[c].val);
                            oi[u] = __ldcg(&all[c].idx);
                        }
                    }
#pragma unroll
                    for (int u = 0; u < 8; ++u)
                        if (beats(ov[u], oi[u], v, i)) { v = ov[u]; i = oi[u]; }
                }
            }
            warp_best(v, i);
            if (lane == 0) { s_peak = v; s_peak_idx = i; }
        }
        __syncthreads();
        const double peak = s_peak;
        const long long peak_idx = s_peak_idx;
        /* stop: cycle limit, or peak below the threshold (.cpp:199-202) */
        if (cycle >= cycle_limit || peak_idx < 0 || (T) peak < threshold) break;
        /* clean_comp[peak] += loop_gain * highest_value (.cpp:211-214), by
         * one thread of the grid: its load is issued here and used after the
         * subtraction pass, off the critical path of the cycle */
        const bool writer = blockIdx.x == 0 && threadIdx.x == 0;
        T before = (T) 0;
        if (writer) before = __ldcg(clean_model + peak_idx);
        /* subtract the shifted PSF from the slice, look for the next peak */
        const int mx = (int) ((unsigned) peak_idx / (unsigned) ni);
        const int my = (int) peak_idx - mx * ni;
        const T* window = psf + ((int64_t) (ni - mx) * 2 * ni + (ni - my));
        bestv = lowest;
        bestk = -1;
        int i = first_i, j = first_j;
        for (int k0 = threadIdx.x; k0 < mine32;
                k0 += ES_CLEAN_UNROLL * ES_CLEAN_THREADS)
        {
            /* all PSF loads of the group in flight before the first use (an
             * L2 round trip each: one at a time they are the whole cycle) */
            T p[ES_CLEAN_UNROLL];
#pragma unroll
            for (int u = 0; u < ES_CLEAN_UNROLL; ++u)
            {
                p[u] = (T) 0;
                if (k0 + u * ES_CLEAN_THREADS < mine32)
                    p[u] = window[(int64_t) i * (2 * ni) + j];
                i += step_i;
                j += step_j;
                if (j >= ni) { j -= ni; ++i; }
            }
#pragma unroll
            for (int u = 0; u < ES_CLEAN_UNROLL; ++u)
            {
                const int k = k0 + u * ES_CLEAN_THREADS;
                if (k >= mine32) break;
                const T r = subtract<T>(res[k], gain, peak, p[u]);
                res[k] = r;
                if (r > bestv) { bestv = r; bestk = k; }
            }
        }
        best = (double) bestv;
        best_idx = bestk < 0 ? -1 : lo + bestk;
        if (writer)
        {
            if (before == (T) 0) components[num_components++] = (int) peak_idx;
            __stcg(clean_model + peak_idx,
                    before + (T) __dmul_rn((double) gain, peak));
        }
        __syncthreads();
    }
    if (IN_SMEM)
        for (int k = threadIdx.x; k < mine32; k += ES_CLEAN_THREADS)
            residual[lo + k] = res[k];
    /* nobody may still be reading this CTA's candidates when it exits */
    if (MODE == ES_CLEAN_CLUSTER) cg::this_cluster().sync();
    if (blockIdx.x == 0 && threadIdx.x == 0)
    {
        header->cycles_done = cycle;
        header->num_components = num_components;
    }
}

/* cbeam[x, y] = exp(-(a dx^2 + 2 b dx dy + c dy^2)), dx = x - dim / 2
 * (sdp_create_cbeam, .cpp:33-80) */
template<typename T>
__global__ void es_clean_beam_kernel(const int64_t dim, const double a,
        const double b, const double c, T* __restrict__ cbeam)
{
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= dim * dim) return;
    const double dx = (double) (i / dim - dim / 2);
    const double dy = (double) (i % dim - dim / 2);
    cbeam[i] = (T) exp(-(a * dx * dx + 2.0 * b * dx * dy + c * dy * dy));
}

/* skymodel = sum over the model's pixels of value x beam + residual */
template<typename T>
__global__ void __launch_bounds__(256) es_clean_restore_kernel(const int64_t n,
        const int64_t dim, const int64_t offset, const T* __restrict__ cbeam,
        const T* __restrict__ clean_model, const T* __restrict__ residual,
        const Header* __restrict__ header, const int* __restrict__ components,
        T* __restrict__ skymodel)
{
    __shared__ int s_i[256], s_j[256];
    __shared__ double s_v[256];
    const int64_t f = (int64_t) blockIdx.x * 256 + threadIdx.x;
    const int64_t x = f / n, y = f - x * n;
    const int num = header->num_components;
    double acc = 0.0;
    for (int c0 = 0; c0 < num; c0 += 256)
    {
        __syncthreads();
        if (c0 + (int) threadIdx.x < num)
        {
            const int idx = components[c0 + threadIdx.x];
            s_i[threadIdx.x] = (int) (idx / n);
            s_j[threadIdx.x] = (int) (idx % n);
            s_v[threadIdx.x] = (double) clean_model[idx];
        }
        __syncthreads();
        const int m = min(256, num - c0);
        if (f < n * n)
            for (int c = 0; c < m; ++c)
            {
                const int64_t dx = x - s_i[c] + offset, dy = y - s_j[c] + offset;
                if (dx >= 0 && dx < dim && dy >= 0 && dy < dim)
                    acc += s_v[c] * (double) cbeam[dx * dim + dy];
            }
    }
    if (f < n * n) skymodel[f] = (T) acc + residual[f];
}

size_t align256(const size_t b) { return (b + 255) & ~(size_t) 255; }

/* entries of the component list: one per cycle, and never more than pixels */
size_t list_entries(const int64_t n, const int cycle_limit)
{
    const int64_t cycles = cycle_limit > 0 ? cycle_limit : 1;
    return (size_t) (cycles < n * n ? cycles : n * n);
}


/* Beam sample under a component: the reference pads both images to P (power
 * of two >= n + dim - 1) at (P - size) / 2, multiplies the spectra, shifts by
 * P / 2 and crops from (P - n) / 2 - 1 (sdp_fft_convolution.cpp:127-244):
 * out[x] = sum_i model[i] beam[x - i + P / 2 - (P - dim) / 2 - 1]. */
int64_t beam_offset(const int64_t n, const int64_t dim)
{
    int64_t p = n + dim - 1;
    while ((p & (p - 1)) != 0) ++p;
    return p / 2 - (p - dim) / 2 - 1;
}

template<typename T>
int run(const int64_t n, const T* psf, const double loop_gain,
        const double threshold, const int cycle_limit, const double a,
        const double b, const double c, const int64_t dim, T* clean_model,
        T* residual, T* skymodel, char* scratch, cudaStream_t st)
{
    /* scratch: header | candidates [2][ES_CLEAN_MAX_CTAS] | components | beam */
    Header* header = (Header*) scratch;
    Cand* cand = (Cand*) (scratch + 256);
    int* components = (int*) (scratch + 256 +
            align256(2 * ES_CLEAN_MAX_CTAS * sizeof(Cand)));
    T* cbeam = (T*) ((char*) components +
            align256(list_entries(n, cycle_limit) * 4));

    int dev = 0, sms = 0, coop = 0, smem_max = 0;
    cudaError_t err = cudaGetDevice(&dev);
    if (err != cudaSuccess) return (int) err;
    (void) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    (void) cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
    (void) cudaDeviceGetAttribute(&smem_max,
            cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    if (!coop || sms < 1) return (int) cudaErrorNotSupported;

    const int64_t total = n * n;
    const T gain = (T) loop_gain, thr = (T) threshold;
    int64_t slice = 0;
    void* args[] = {(void*) &n, (void*) &slice, (void*) &psf, (void*) &gain,
            (void*) &thr, (void*) &cycle_limit, (void*) &residual,
            (void*) &clean_model, (void*) &cand, (void*) &header,
            (void*) &components};
    bool launched = false;

    /* small images: ONE cluster of 16 (or 8) CTAs holds the whole residual in
     * its shared memory and exchanges the candidates through it */
    const char* no_cluster = getenv("SKA_SDP_FUNC_CLEAN_NO_CLUSTER");
    int max_groups = 1; /* groups of PSF loads per thread and cycle */
    if (const char* e = getenv("SKA_SDP_FUNC_CLEAN_CLUSTER_GROUPS"))
        if (atoi(e) > 0) max_groups = atoi(e);
    for (int csize = 16; csize >= 8 && !launched &&
            !(no_cluster && atoi(no_cluster) > 0); csize /= 2)
    {
        int64_t rows = (n + csize - 1) / csize;
        slice = rows * n;
        const size_t slice_bytes = (size_t) slice * sizeof(T);
        /* worth it while a thread has at most one group of PSF loads per
         * cycle (measured: 256^2 4.2 vs 5.0 us per cycle, 512^2 8.7 vs 5.0) */
        if (slice_bytes + 2048 > (size_t) smem_max ||
                total < (int64_t) csize * ES_CLEAN_THREADS ||
                slice > (int64_t) max_groups * ES_CLEAN_UNROLL * ES_CLEAN_THREADS)
            continue;
        const void* fn = (const void*) es_hogbom_cycles_kernel<T, ES_CLEAN_CLUSTER>;
        /* the opt-in belongs to the function and the device, not to a call:
         * always the device maximum, so that concurrent calls with other
         * slice sizes cannot lower it under each other */
        if (cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize,
                smem_max - 2048) != cudaSuccess ||
                (csize > 8 && cudaFuncSetAttribute(fn,
                cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess))
        {
            (void) cudaGetLastError();
            continue;
        }
        cudaLaunchConfig_t cfg;
        memset(&cfg, 0, sizeof(cfg));
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = (unsigned) csize;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.gridDim = dim3((unsigned) ((total + slice - 1) / slice));
        cfg.blockDim = dim3(ES_CLEAN_THREADS);
        cfg.dynamicSmemBytes = slice_bytes;
        cfg.stream = st;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        /* the cluster must be complete: rows that do not fill it fall back */
        if (cfg.gridDim.x != (unsigned) csize) continue;
        int clusters = 0;
        if (cudaOccupancyMaxActiveClusters(&clusters, fn, &cfg) != cudaSuccess ||
                clusters < 1)
        {
            (void) cudaGetLastError();
            continue;
        }
        err = cudaLaunchKernelExC(&cfg, fn, args);
        if (err != cudaSuccess) return (int) err;
        launched = true;
    }
    if (!launched)
    {
        /* few elements per thread and cycle; at most one CTA per SM */
        int per_thread = 4;
        if (const char* e = getenv("SKA_SDP_FUNC_CLEAN_PER_THREAD"))
            if (atoi(e) > 0) per_thread = atoi(e);
        int64_t ctas = (total + per_thread * ES_CLEAN_THREADS - 1) /
                (per_thread * ES_CLEAN_THREADS);
        int64_t cap = sms;
        if (const char* e = getenv("SKA_SDP_FUNC_CLEAN_MAX_CTAS"))
            if (atoi(e) > 0 && atoi(e) < sms) cap = atoi(e);
        if (ctas > cap) ctas = cap;
        if (ctas > ES_CLEAN_MAX_CTAS) ctas = ES_CLEAN_MAX_CTAS;
        if (ctas < 1) ctas = 1;
        /* slices of whole rows (aligned, coalesced PSF windows) where there
         * are enough rows, else of whole CTA strides */
        slice = (total + ctas - 1) / ctas;
        const int64_t unit = n >= 2 * ctas ? n : ES_CLEAN_THREADS;
        slice = (slice + unit - 1) / unit * unit;
        ctas = (total + slice - 1) / slice;
        const size_t slice_bytes = (size_t) slice * sizeof(T);
        const bool in_smem = slice_bytes + 2048 <= (size_t) smem_max;
        const void* fn = in_smem ?
                (const void*) es_hogbom_cycles_kernel<T, ES_CLEAN_SMEM> :
                (const void*) es_hogbom_cycles_kernel<T, ES_CLEAN_GLOBAL>;
        if (in_smem)
        {
            err = cudaFuncSetAttribute(fn,
                    cudaFuncAttributeMaxDynamicSharedMemorySize,
                    smem_max - 2048);
            if (err != cudaSuccess) return (int) err;
        }
        int per_sm = 0;
        err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn,
                ES_CLEAN_THREADS, in_smem ? slice_bytes : 0);
        if (err != cudaSuccess) return (int) err;
        if (per_sm < 1 || (int64_t) per_sm * sms < ctas)
            return (int) cudaErrorCooperativeLaunchTooLarge;
        err = cudaLaunchCooperativeKernel(fn, dim3((unsigned) ctas),
                dim3(ES_CLEAN_THREADS), args, in_smem ? slice_bytes : 0, st);
        if (err != cudaSuccess) return (int) err;
    }

    if (dim > 0)
    {
        es_clean_beam_kernel<T><<<(unsigned) ((dim * dim + 255) / 256), 256, 0,
                st>>>(dim, a, b, c, cbeam);
        if ((err = cudaGetLastError()) != cudaSuccess) return (int) err;
    }
    es_clean_restore_kernel<T><<<(unsigned) ((total + 255) / 256), 256, 0, st>>>(
            n, dim, beam_offset(n, dim), cbeam, clean_model, residual, header,
            components, skymodel);
    return (int) cudaGetLastError();
}

} // namespace

extern "C" {

size_t es_hogbom_scratch_bytes(int64_t n, int cycle_limit, int64_t cbeam_dim,
        int is_double)
{
    return 256 + align256(2 * ES_CLEAN_MAX_CTAS * sizeof(Cand)) +
            align256(list_entries(n, cycle_limit) * 4) +
            align256((size_t) (cbeam_dim > 0 ? cbeam_dim * cbeam_dim : 1) *
                    (is_double ? 8 : 4));
}

int es_hogbom_clean(int64_t n, const void* psf, double loop_gain,
        double threshold, int cycle_limit, double beam_a, double beam_b,
        double beam_c, int64_t cbeam_dim, void* clean_model, void* residual,
        void* skymodel, void* scratch, int is_double, void* stream)
{
    return is_double ?
            run<double>(n, (const double*) psf, loop_gain, threshold,
                    cycle_limit, beam_a, beam_b, beam_c, cbeam_dim,
                    (double*) clean_model, (double*) residual,
                    (double*) skymodel, (char*) scratch, (cudaStream_t) stream) :
            run<float>(n, (const float*) psf, loop_gain, threshold,
                    cycle_limit, beam_a, beam_b, beam_c, cbeam_dim,
                    (float*) clean_model, (float*) residual, (float*) skymodel,
                    (char*) scratch, (cudaStream_t) stream);
}

} // extern "C"
