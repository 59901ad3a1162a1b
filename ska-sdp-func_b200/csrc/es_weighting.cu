/*
 * sm_100a kernels of the uv-grid visibility weighting (uniform / Briggs),
 * behind plain-pointer C launchers (es_weighting.h). Replaces the reference's
 * grid_write_gpu / calc_sum_gpu / grid_briggs_read_gpu / grid_uniform_read_gpu
 * (src/ska-sdp-func/visibility/sdp_weighting.cu:19-313); the arithmetic that
 * decides the cell of a visibility is that of sdp_weighting.cpp:42-52, in
 * fp64, operation by operation, so that cell indices agree bit for bit.
 *
 * Layout of the work: one thread per (time, baseline, channel), channels
 * fastest, so that a warp reads consecutive weights (num_pols reals each) and
 * one or two uvw rows; the grid is updated / read with one atomic / load per
 * polarisation. HBM-bound on the weights stream: 2 x sizeof(W) x num_pols
 * bytes per visibility and pass, plus the scattered cell accesses (L2).
 */
#include <stdint.h>

#include <cuda_runtime.h>

#include "es_weighting.h"

#define ES_WT_C0 299792458.0 /* sdp_weighting.cpp:10 */
#define ES_WT_THREADS 256

namespace {

/* Cell of visibility (row = time * num_baselines + baseline, chan); false
 * when it falls outside the grid (the reference tests the upper bound only,
 * :53; a negative index would be out of bounds there and is skipped here). */
__device__ __forceinline__ bool wt_cell(const double* __restrict__ uvw,
        const double* __restrict__ freq_hz, const int64_t row, const int chan,
        const double max_abs_uv, const int64_t grid_size, int64_t& cell)
{
    const int64_t half = grid_size / 2;
    const double inv_wavelength = freq_hz[chan] / ES_WT_C0;
    const double grid_u = uvw[3 * row] * inv_wavelength;
    const double grid_v = uvw[3 * row + 1] * inv_wavelength;
    const int64_t iu = (int64_t) (floor(grid_u / max_abs_uv * (double) half) +
            (double) half);
    const int64_t iv = (int64_t) (floor(grid_v / max_abs_uv * (double) half) +
            (double) half);
    cell = iu * grid_size + iv;
    return iu >= 0 && iv >= 0 && iu < grid_size && iv < grid_size;
}

template<typename W>
__global__ void __launch_bounds__(ES_WT_THREADS) es_weight_grid_write_kernel(
        const int64_t num_rows, const int num_chan, const int num_pols,
        const int64_t grid_size, const double* __restrict__ uvw,
        const double* __restrict__ freq_hz, const double max_abs_uv,
        W* __restrict__ grid, const W* __restrict__ input)
{
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= num_rows * num_chan) return;
    const int64_t row = i / num_chan;
    const int chan = (int) (i - row * num_chan);
    int64_t cell;
    if (!wt_cell(uvw, freq_hz, row, chan, max_abs_uv, grid_size, cell)) return;
    for (int p = 0; p < num_pols; ++p)
        atomicAdd(grid + cell * num_pols + p, input[i * num_pols + p]);
}

/* sums[0] += sum grid[cell], sums[1] += sum grid[cell]^2 over the
 * visibilities inside the grid and their polarisations, in fp64. */
template<typename W>
__global__ void __launch_bounds__(ES_WT_THREADS) es_weight_sums_kernel(
        const int64_t num_rows, const int num_chan, const int num_pols,
        const int64_t grid_size, const double* __restrict__ uvw,
        const double* __restrict__ freq_hz, const double max_abs_uv,
        const W* __restrict__ grid, double* __restrict__ sums)
{
    __shared__ double s_part[2][ES_WT_THREADS / 32];
    double s1 = 0.0, s2 = 0.0;
    const int64_t total = num_rows * num_chan;
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
            i < total; i += (int64_t) gridDim.x * blockDim.x)
    {
        const int64_t row = i / num_chan;
        const int chan = (int) (i - row * num_chan);
        int64_t cell;
        if (!wt_cell(uvw, freq_hz, row, chan, max_abs_uv, grid_size, cell))
            continue;
        for (int p = 0; p < num_pols; ++p)
        {
            const double g = (double) grid[cell * num_pols + p];
            s1 += g;
            s2 += g * g;
        }
    }
    for (int o = 16; o > 0; o >>= 1)
    {
        s1 += __shfl_down_sync(0xffffffffu, s1, o);
        s2 += __shfl_down_sync(0xffffffffu, s2, o);
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) { s_part[0][warp] = s1; s_part[1][warp] = s2; }
    __syncthreads();
    if (threadIdx.x == 0)
    {
        double a = 0.0, b = 0.0;
        for (int w = 0; w < ES_WT_THREADS / 32; ++w)
        {
            a += s_part[0][w];
            b += s_part[1][w];
        }
        atomicAdd(sums, a);
        atomicAdd(sums + 1, b);
    }
}

/* BRIGGS: output = input / (1 + f2 grid[cell]) with f2 from the sums
 * (robustness_calc, sdp_weighting.cpp:136-144); else output = 1 / grid[cell]. */
template<typename W, bool BRIGGS>
__global__ void __launch_bounds__(ES_WT_THREADS) es_weight_read_kernel(
        const int64_t num_rows, const int num_chan, const int num_pols,
        const int64_t grid_size, const double* __restrict__ uvw,
        const double* __restrict__ freq_hz, const double max_abs_uv,
        const W* __restrict__ grid, const W* __restrict__ input,
        W* __restrict__ output, const double numerator,
        const double* __restrict__ sums)
{
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= num_rows * num_chan) return;
    const int64_t row = i / num_chan;
    const int chan = (int) (i - row * num_chan);
    int64_t cell;
    if (!wt_cell(uvw, freq_hz, row, chan, max_abs_uv, grid_size, cell)) return;
    double robustness = 0.0;
    if (BRIGGS) robustness = numerator / (sums[1] / sums[0]);
    for (int p = 0; p < num_pols; ++p)
    {
        const W g = grid[cell * num_pols + p];
        /* the reference evaluates these in the weights' precision
         * (WEIGHT_TYPE weight_val, :205,266; robustness is a double, so the
         * Briggs denominator is formed in fp64 and the quotient rounded) */
        if (BRIGGS)
            output[i * num_pols + p] = (W) ((double) input[i * num_pols + p] /
                    (1.0 + robustness * (double) g));
        else
            output[i * num_pols + p] = (W) 1.0 / g;
    }
}

} // namespace

extern "C" {

int es_weight_uniform(int64_t num_rows, int num_chan, int num_pols,
        int64_t grid_size, const double* uvw, const double* freq_hz,
        double max_abs_uv, void* grid, const void* input, void* output,
        int is_double, void* stream)
{
    cudaStream_t s = (cudaStream_t) stream;
    const int64_t total = num_rows * num_chan;
    if (total == 0) return 0;
    const unsigned blocks = (unsigned) ((total + ES_WT_THREADS - 1) / ES_WT_THREADS);
    if (is_double)
    {
        es_weight_grid_write_kernel<double><<<blocks, ES_WT_THREADS, 0, s>>>(
                num_rows, num_chan, num_pols, grid_size, uvw, freq_hz,
                max_abs_uv, (double*) grid, (const double*) input);
        es_weight_read_kernel<double, false><<<blocks, ES_WT_THREADS, 0, s>>>(
                num_rows, num_chan, num_pols, grid_size, uvw, freq_hz,
                max_abs_uv, (const double*) grid, (const double*) input,
                (double*) output, 0.0, 0);
    }
    else
    {
        es_weight_grid_write_kernel<float><<<blocks, ES_WT_THREADS, 0, s>>>(
                num_rows, num_chan, num_pols, grid_size, uvw, freq_hz,
                max_abs_uv, (float*) grid, (const float*) input);
        es_weight_read_kernel<float, false><<<blocks, ES_WT_THREADS, 0, s>>>(
                num_rows, num_chan, num_pols, grid_size, uvw, freq_hz,
                max_abs_uv, (const float*) grid, (const float*) input,
                (float*) output, 0.0, 0);
    }
    return (int) cudaGetLastError();
}

int es_weight_briggs(int64_t num_rows, int num_chan, int num_pols,
        int64_t grid_size, const double* uvw, const double* freq_hz,
        double max_abs_uv, double numerator, void* grid, const void* input,
        void* output, double* sums, int is_double, void* stream)
{
    cudaStream_t s = (cudaStream_t) stream;
    const int64_t total = num_rows * num_chan;
    if (total == 0) return 0;
    const unsigned blocks = (unsigned) ((total + ES_WT_THREADS - 1) / ES_WT_THREADS);
    /* persistent-style reduction: a few CTAs per SM */
    int dev = 0, sms = 0;
    cudaError_t err = cudaGetDevice(&dev);
    if (err == cudaSuccess) err = cudaDeviceGetAttribute(&sms,
            cudaDevAttrMultiProcessorCount, dev);
    if (err != cudaSuccess) return (int) err;
    unsigned rblocks = (unsigned) (sms * 8);
    if (rblocks > blocks) rblocks = blocks;
    err = cudaMemsetAsync(sums, 0, 2 * sizeof(double), s);
    if (err != cudaSuccess) return (int) err;
    if (is_double)
    {
        es_weight_grid_write_kernel<double><<<blocks, ES_WT_THREADS, 0, s>>>(
                num_rows, num_chan, num_pols, grid_size, uvw, freq_hz,
                max_abs_uv, (double*) grid, (const double*) input);
        es_weight_sums_kernel<double><<<rblocks, ES_WT_THREADS, 0, s>>>(
                num_rows, num_chan, num_pols, grid_size, uvw, freq_hz,
                max_abs_uv, (const double*) grid, sums);
        es_weight_read_kernel<double, true><<<blocks, ES_WT_THREADS, 0, s>>>(
                num_rows, num_chan, num_pols, grid_size, uvw, freq_hz,
                max_abs_uv, (const double*) grid, (const double*) input,
                (double*) output, numerator, sums);
    }
    else
    {
        es_weight_grid_write_kernel<float><<<blocks, ES_WT_THREADS, 0, s>>>(
                num_rows, num_chan, num_pols, grid_size, uvw, freq_hz,
                max_abs_uv, (float*) grid, (const float*) input);
        es_weight_sums_kernel<float><<<rblocks, ES_WT_THREADS, 0, s>>>(
                num_rows, num_chan, num_pols, grid_size, uvw, freq_hz,
                max_abs_uv, (const float*) grid, sums);
        es_weight_read_kernel<float, true><<<blocks, ES_WT_THREADS, 0, s>>>(
                num_rows, num_chan, num_pols, grid_size, uvw, freq_hz,
                max_abs_uv, (const float*) grid, (const float*) input,
                (float*) output, numerator, sums);
    }
    return (int) cudaGetLastError();
}

/* ---- Briggs weighting without any allocation ----
 * For callers that cannot call cudaMalloc (CUDA graph capture, serving
 * loops): the uv weight grid and the two Briggs sums live in this file-scope
 * device array instead of caller / cudaMalloc'ed memory. ES_WT_STATIC_CELLS
 * (grid_size^2 x num_pols) doubles + 2: 8 MiB + 16 B on every GPU that loads
 * the library, until exit. The array is shared by every call and keeps the
 * last call's contents: each call clears what it uses; calls that overlap on
 * the GPU (different streams or host threads) race on it.
 * In host code `es_wt_scratch` is a host-side stand-in, not a device address:
 * the address is asked for once per device (with that device current) by
 * cudaGetSymbolAddress and that pointer is what the kernels get. */
static __device__ __align__(256) double es_wt_scratch[ES_WT_STATIC_CELLS + 2];

#define ES_WT_MAX_DEVICES 64
static double* es_wt_scratch_ptr[ES_WT_MAX_DEVICES]; /* per device, 0 = not asked yet */

int64_t es_weight_static_max_cells(void)
{
    return (int64_t) ES_WT_STATIC_CELLS;
}

/* As es_weight_briggs with grid and sums in the static scratch. Returns a
 * cudaError_t as int; cudaErrorInvalidValue when the grid does not fit (the
 * caller checks the size first and reports it). Calls no cudaMalloc /
 * cudaFree: cudaGetSymbolAddress, cudaMemsetAsync and kernel launches only. */
int es_weight_briggs_static(int64_t num_rows, int num_chan, int num_pols,
        int64_t grid_size, const double* uvw, const double* freq_hz,
        double max_abs_uv, double numerator, const void* input, void* output,
        int is_double, void* stream)
{
    cudaStream_t s = (cudaStream_t) stream;
    const int64_t cells = grid_size * grid_size * (int64_t) num_pols;
    if (grid_size <= 0 || num_pols <= 0 || cells > (int64_t) ES_WT_STATIC_CELLS)
        return (int) cudaErrorInvalidValue;
    int dev = 0, sms = 0;
    cudaError_t err = cudaGetDevice(&dev);
    if (err != cudaSuccess) return (int) err;
    if (dev < 0 || dev >= ES_WT_MAX_DEVICES) return (int) cudaErrorInvalidDevice;
    if (!es_wt_scratch_ptr[dev])
    {
        /* set-up, once per device, that device being current */
        double* ptr = 0;
        err = cudaGetSymbolAddress((void**) &ptr, es_wt_scratch);
        if (err != cudaSuccess) return (int) err;
        es_wt_scratch_ptr[dev] = ptr;
    }
    double* scratch = es_wt_scratch_ptr[dev];
    void* grid = (void*) scratch;                /* W[cells], W = float / double */
    double* sums = scratch + ES_WT_STATIC_CELLS; /* [2] */
    const int64_t total = num_rows * num_chan;
    if (total == 0) return 0;
    /* the array keeps the last call's contents: clear the cells this call
     * accumulates into, and the sums */
    err = cudaMemsetAsync(grid, 0, (size_t) cells * (is_double ? 8 : 4), s);
    if (err == cudaSuccess) err = cudaMemsetAsync(sums, 0, 2 * sizeof(double), s);
    if (err == cudaSuccess) err = cudaDeviceGetAttribute(&sms,
            cudaDevAttrMultiProcessorCount, dev);
    if (err != cudaSuccess) return (int) err;
    const unsigned blocks = (unsigned) ((total + ES_WT_THREADS - 1) / ES_WT_THREADS);
    unsigned rblocks = (unsigned) (sms * 8);
    if (rblocks > blocks) rblocks = blocks;
    if (is_double)
    {
        es_weight_grid_write_kernel<double><<<blocks, ES_WT_THREADS, 0, s>>>(
                num_rows, num_chan, num_pols, grid_size, uvw, freq_hz,
                max_abs_uv, (double*) grid, (const double*) input);
        es_weight_sums_kernel<double><<<rblocks, ES_WT_THREADS, 0, s>>>(
                num_rows, num_chan, num_pols, grid_size, uvw, freq_hz,
                max_abs_uv, (const double*) grid, sums);
        es_weight_read_kernel<double, true><<<blocks, ES_WT_THREADS, 0, s>>>(
                num_rows, num_chan, num_pols, grid_size, uvw, freq_hz,
                max_abs_uv, (const double*) grid, (const double*) input,
                (double*) output, numerator, sums);
    }
    else
    {
        es_weight_grid_write_kernel<float><<<blocks, ES_WT_THREADS, 0, s>>>(
                num_rows, num_chan, num_pols, grid_size, uvw, freq_hz,
                max_abs_uv, (float*) grid, (const float*) input);
        es_weight_sums_kernel<float><<<rblocks, ES_WT_THREADS, 0, s>>>(
                num_rows, num_chan, num_pols, grid_size, uvw, freq_hz,
                max_abs_uv, (const float*) grid, sums);
        es_weight_read_kernel<float, true><<<blocks, ES_WT_THREADS, 0, s>>>(
                num_rows, num_chan, num_pols, grid_size, uvw, freq_hz,
                max_abs_uv, (const float*) grid, (const float*) input,
                (float*) output, numerator, sums);
    }
    return (int) cudaGetLastError();
}

} // extern "C"
