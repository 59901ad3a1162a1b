/*
 * sm_100a kernels of the tiled visibility functions -- tile count + prefix
 * sum, bucket sort, tiled indexing -- behind plain-pointer C launchers
 * (es_tiling.h). Replaces the reference's sdp_tile_count_simple_gpu /
 * sdp_prefix_sum_gpu / sdp_bucket_sort_simple_gpu / sdp_tiled_indexing_gpu
 * (src/ska-sdp-func/visibility/sdp_tiled_functions.cu:20-291). The arithmetic
 * that decides the tiles of a visibility is the reference's (.cu:96-114),
 * operation by operation and in its precisions, so that tiles, counts and
 * positions agree bit for bit.
 *
 * What differs is the schedule. The reference takes the slot of every
 * (visibility, tile) pair with an atomicAdd on the tile's offset, so the order
 * inside a tile changes from run to run, its prefix sum is one 512-thread CTA,
 * and its count issues one atomic per pair. Here:
 *   count : one thread per (row, channel), channels fastest; the lanes of a
 *           warp that hit the same tile are merged (match.any) into one atomic;
 *   scan  : device-wide scan (CUB decoupled look-back);
 *   sort  : the pairs are emitted in visibility order (per-CTA totals -> scan
 *           -> emit), sorted by tile with a STABLE radix sort over only the
 *           bits a tile index needs, and written through the caller's offsets:
 *           inside a tile the entries are in the order of the reference's host
 *           loops (time, baseline, channel; pv outer, pu inner), the same on
 *           every run. This is the presort the ES gridder uses (es_kernels.cu:
 *           count / emit / radix sort / bounds) on the reference's tile
 *           geometry.
 * HBM-bound: the outputs are 4 reals + 1 int per pair written once, the sort
 * moves 8 bytes per pair and pass; uvw and the per-channel 1 / wavelength
 * table are read through L1 / L2.
 * Tiles that do not exist: the centre of the grid is the middle of a tile, so
 * when grid_size is a multiple of the tile size the last half tile of each
 * axis lies in tile column / row num_tiles_u / num_tiles_v. The reference
 * forms the linear index pu + pv * num_tiles_u of such a pair all the same,
 * which lands in a neighbouring tile's counter and slots (and stores the
 * literal pv * 32768 + pu as its tile code). That is reproduced, so that
 * counts, offsets and sorted arrays equal the reference's; only a pair whose
 * linear index leaves [0, num_tiles) -- an out-of-bounds access there -- is
 * dropped.
 */
#include <stdint.h>

#include <mutex>

#include <cuda_runtime.h>
#include <cub/cub.cuh>

#include "es_tiling.h"

#define ES_TILE_C0 299792458.0 /* sdp_tiled_functions.cu:7 */
#define ES_TILE_THREADS 256
#define ES_TILE_VPT 4 /* visibilities per thread of the emit kernels */
#define ES_TILE_CHUNK (ES_TILE_THREADS * ES_TILE_VPT)
#define ES_TILE_MAX_DEVICES 64

namespace {

struct TileRange
{
    int u0, u1, v0, v1; /* [u0, u1) x [v0, v1), as the reference computes them */
};

/* The pu of tile row pv whose linear index pu + pv * num_tiles_u is inside
 * [0, num_tiles): [lo, hi), possibly empty. */
__device__ __forceinline__ void row_pairs(const EsTileGeom& g,
        const TileRange& r, const int pv, int& lo, int& hi)
{
    const int64_t first = (int64_t) pv * g.num_tiles_u;
    const int64_t num_tiles = g.num_tiles_u * g.num_tiles_v;
    lo = (int) max((int64_t) r.u0, -first);
    hi = (int) min((int64_t) r.u1, num_tiles - first);
    if (hi < lo) hi = lo;
}

__device__ __forceinline__ int count_pairs(const EsTileGeom& g,
        const TileRange& r)
{
    int n = 0;
    for (int pv = r.v0; pv < r.v1; ++pv)
    {
        int lo, hi;
        row_pairs(g, r, pv, lo, hi);
        n += hi - lo;
    }
    return n;
}

__device__ __forceinline__ float tl_round(float x) { return roundf(x); }
__device__ __forceinline__ double tl_round(double x) { return round(x); }

/* Position (in cells from the grid centre, in T) and tile range of
 * visibility (row, chan); false when its footprint leaves the grid
 * (.cu:96-114). */
template<typename T>
__device__ __forceinline__ bool tile_range(const EsTileGeom& g,
        const T* __restrict__ uvw, const T* __restrict__ inv_wl,
        const int64_t row, const int chan, T& pos_u, T& pos_v, TileRange& r)
{
    const T inv_wavelength = inv_wl[chan];
    const T lambda_u = uvw[3 * row] * inv_wavelength;
    const T lambda_v = uvw[3 * row + 1] * inv_wavelength;
    pos_u = (T) ((double) lambda_u * g.grid_scale);
    pos_v = (T) ((double) lambda_v * g.grid_scale);
    const int64_t centre = g.grid_size / 2;
    const int64_t grid_u = (int64_t) tl_round(pos_u) + centre;
    const int64_t grid_v = (int64_t) tl_round(pos_v) + centre;
    if (!((grid_u + g.support < g.grid_size) && (grid_u - g.support >= 0) &&
            (grid_v + g.support < g.grid_size) && (grid_v - g.support >= 0)))
        return false;
    const int rel_u = (int) (grid_u - g.top_left_u);
    const int rel_v = (int) (grid_v - g.top_left_v);
    const float u1 = __fmul_rn((float) (rel_u - g.support), g.inv_tile_size_u);
    const float u2 = __fmul_rn((float) (rel_u + g.support + 1), g.inv_tile_size_u);
    const float v1 = __fmul_rn((float) (rel_v - g.support), g.inv_tile_size_v);
    const float v2 = __fmul_rn((float) (rel_v + g.support + 1), g.inv_tile_size_v);
    r.u0 = (int) floorf(u1);
    r.u1 = (int) ceilf(u2);
    r.v0 = (int) floorf(v1);
    r.v1 = (int) ceilf(v2);
    return true;
}

/* inv_wl[chan] = freq / c in double, rounded to T (.cu:96): once per call
 * instead of one fp64 division per visibility and kernel. */
template<typename T>
__global__ void __launch_bounds__(ES_TILE_THREADS) es_tile_inv_wavelength_kernel(
        const int num_chan, const T* __restrict__ freqs, T* __restrict__ inv_wl)
{
    const int c = blockIdx.x * ES_TILE_THREADS + threadIdx.x;
    if (c < num_chan) inv_wl[c] = (T) ((double) freqs[c] / ES_TILE_C0);
}

/* (row, chan) of flat visibility i (< 2^31, checked by the host). */
__device__ __forceinline__ void row_chan(const EsTileGeom& g, const int64_t i,
        int64_t& row, int& chan)
{
    const uint32_t q = (uint32_t) i / (uint32_t) g.num_chan;
    row = (int64_t) q;
    chan = (int) ((uint32_t) i - q * (uint32_t) g.num_chan);
}

/* counts[tile] += pairs, *skipped += visibilities off the grid. */
template<typename T>
__global__ void __launch_bounds__(ES_TILE_THREADS) es_tile_count_kernel(
        const EsTileGeom g, const T* __restrict__ uvw,
        const T* __restrict__ inv_wl, int32_t* __restrict__ counts,
        int32_t* __restrict__ skipped)
{
    const int64_t total = g.num_rows * g.num_chan;
    const int lane = threadIdx.x & 31;
    for (int64_t base = (int64_t) blockIdx.x * ES_TILE_THREADS; base < total;
            base += (int64_t) gridDim.x * ES_TILE_THREADS)
    {
        const int64_t i = base + threadIdx.x;
        TileRange r = {0, 0, 0, 0};
        bool inside = false;
        if (i < total)
        {
            int64_t row;
            int chan;
            row_chan(g, i, row, chan);
            T pos_u, pos_v;
            inside = tile_range(g, uvw, inv_wl, row, chan, pos_u, pos_v, r);
        }
        const unsigned off_grid = __ballot_sync(0xffffffffu, i < total && !inside);
        if (lane == 0 && off_grid) atomicAdd(skipped, __popc(off_grid));
        const int nu = r.u1 - r.u0;
        const int n = inside ? nu * (r.v1 - r.v0) : 0;
        const int n_max = __reduce_max_sync(0xffffffffu, n);
        const int64_t num_tiles = g.num_tiles_u * g.num_tiles_v;
        for (int k = 0; k < n_max; ++k)
        {
            int tile = -1;
            if (k < n)
            {
                const int64_t at = (r.u0 + k % nu) +
                        (int64_t) (r.v0 + k / nu) * g.num_tiles_u;
                if (at >= 0 && at < num_tiles) tile = (int) at;
            }
            /* lanes on the same tile: one atomic for all of them */
            const unsigned peers = __match_any_sync(0xffffffffu, tile);
            if (tile >= 0 && lane == __ffs(peers) - 1)
                atomicAdd(counts + tile, __popc(peers));
        }
    }
}

/* The ES_TILE_VPT consecutive visibilities of a thread (blocked arrangement:
 * entries come out in visibility order) and the pairs each of them has. */
template<typename T>
__device__ __forceinline__ int thread_ranges(const EsTileGeom& g,
        const T* __restrict__ uvw, const T* __restrict__ inv_wl,
        const int64_t first, TileRange (&r)[ES_TILE_VPT], int (&n)[ES_TILE_VPT])
{
    const int64_t total = g.num_rows * g.num_chan;
    int sum = 0;
#pragma unroll
    for (int k = 0; k < ES_TILE_VPT; ++k)
    {
        const int64_t i = first + k;
        n[k] = 0;
        if (i < total)
        {
            int64_t row;
            int chan;
            row_chan(g, i, row, chan);
            T pos_u, pos_v;
            if (tile_range(g, uvw, inv_wl, row, chan, pos_u, pos_v, r[k]))
                n[k] = count_pairs(g, r[k]);
        }
        sum += n[k];
    }
    return sum;
}

/* totals[cta] = pairs of the ES_TILE_CHUNK visibilities of this CTA. */
template<typename T>
__global__ void __launch_bounds__(ES_TILE_THREADS) es_tile_totals_kernel(
        const EsTileGeom g, const T* __restrict__ uvw,
        const T* __restrict__ inv_wl, uint64_t* __restrict__ totals)
{
    typedef cub::BlockReduce<int, ES_TILE_THREADS> Reduce;
    __shared__ typename Reduce::TempStorage temp;
    TileRange r[ES_TILE_VPT];
    int n[ES_TILE_VPT];
    const int64_t first = ((int64_t) blockIdx.x * ES_TILE_THREADS + threadIdx.x)
            * ES_TILE_VPT;
    const int sum = Reduce(temp).Sum(thread_ranges(g, uvw, inv_wl, first, r, n));
    if (threadIdx.x == 0) totals[blockIdx.x] = (uint64_t) sum;
}

/* keys[e] = tile, vals[e] = i_vis for the pairs of this CTA, from bases[cta]
 * on, in visibility order; per visibility pv outer, pu inner. A CTA whose
 * pairs fit ES_TILE_STAGE slots of shared memory gathers them there and
 * writes them out coalesced. */
#define ES_TILE_STAGE (2 * ES_TILE_CHUNK)
template<typename T>
__global__ void __launch_bounds__(ES_TILE_THREADS) es_tile_emit_kernel(
        const EsTileGeom g, const T* __restrict__ uvw,
        const T* __restrict__ inv_wl, const uint64_t* __restrict__ bases,
        uint32_t* __restrict__ keys, uint32_t* __restrict__ vals)
{
    typedef cub::BlockScan<int, ES_TILE_THREADS> Scan;
    __shared__ typename Scan::TempStorage temp;
    __shared__ uint32_t s_keys[ES_TILE_STAGE];
    __shared__ uint32_t s_vals[ES_TILE_STAGE];
    TileRange r[ES_TILE_VPT];
    int n[ES_TILE_VPT];
    const int64_t first = ((int64_t) blockIdx.x * ES_TILE_THREADS + threadIdx.x)
            * ES_TILE_VPT;
    const int sum = thread_ranges(g, uvw, inv_wl, first, r, n);
    int offset, cta_total;
    Scan(temp).ExclusiveSum(sum, offset, cta_total);
    const uint64_t base = bases[blockIdx.x];
    const bool staged = cta_total <= ES_TILE_STAGE;
    uint32_t* out_keys = staged ? s_keys : keys + base;
    uint32_t* out_vals = staged ? s_vals : vals + base;
    int e = offset;
#pragma unroll
    for (int k = 0; k < ES_TILE_VPT; ++k)
    {
        if (n[k] == 0) continue;
        for (int pv = r[k].v0; pv < r[k].v1; ++pv)
        {
            int lo, hi;
            row_pairs(g, r[k], pv, lo, hi);
            for (int pu = lo; pu < hi; ++pu, ++e)
            {
                out_keys[e] = (uint32_t) (pu + (int64_t) pv * g.num_tiles_u);
                out_vals[e] = (uint32_t) (first + k);
            }
        }
    }
    if (!staged) return;
    __syncthreads();
    for (int i = threadIdx.x; i < cta_total; i += ES_TILE_THREADS)
    {
        keys[base + i] = s_keys[i];
        vals[base + i] = s_vals[i];
    }
}

/* start[t] = first sorted entry of tile t, for t in [0, num_tiles]
 * (start[num_tiles] = num_entries): thread j fills the tiles between the key
 * before it and its own. */
__global__ void __launch_bounds__(ES_TILE_THREADS) es_tile_bounds_kernel(
        const uint32_t* __restrict__ keys, const int64_t num_entries,
        const int64_t num_tiles, uint32_t* __restrict__ start)
{
    const int64_t j = (int64_t) blockIdx.x * ES_TILE_THREADS + threadIdx.x;
    if (j > num_entries) return;
    const int64_t prev = j > 0 ? (int64_t) keys[j - 1] : -1;
    const int64_t cur = j < num_entries ? (int64_t) keys[j] : num_tiles;
    for (int64_t t = prev + 1; t <= cur; ++t) start[t] = (uint32_t) j;
}

/* Sorted entry j of tile t goes to slot tile_offsets[t] + (j - start[t])
 * (.cu:200-205, with the rank in place of the atomic). */
template<typename T, bool VIS_COMPLEX>
__global__ void __launch_bounds__(ES_TILE_THREADS) es_tile_fill_kernel(
        const EsTileGeom g, const T* __restrict__ uvw,
        const T* __restrict__ inv_wl, const T* __restrict__ vis,
        const T* __restrict__ weights, const int num_pols,
        const uint32_t* __restrict__ keys, const uint32_t* __restrict__ vals,
        const uint32_t* __restrict__ start,
        const int32_t* __restrict__ tile_offsets, const int64_t num_entries,
        const int64_t num_out, T* __restrict__ sorted_uu,
        T* __restrict__ sorted_vv, T* __restrict__ sorted_vis,
        T* __restrict__ sorted_weight, int32_t* __restrict__ sorted_tile,
        int32_t* __restrict__ sorted_vis_index)
{
    const int64_t j = (int64_t) blockIdx.x * ES_TILE_THREADS + threadIdx.x;
    if (j >= num_entries) return;
    const uint32_t tile = keys[j];
    const int64_t i_vis = (int64_t) vals[j];
    const int64_t slot = (int64_t) tile_offsets[tile] + (j - (int64_t) start[tile]);
    if (slot < 0 || slot >= num_out) return;
    int64_t row;
    int chan;
    row_chan(g, i_vis, row, chan);
    T pos_u, pos_v;
    TileRange r;
    (void) tile_range(g, uvw, inv_wl, row, chan, pos_u, pos_v, r);
    sorted_uu[slot] = pos_u;
    sorted_vv[slot] = pos_v;
    /* The (pu, pv) this entry was emitted for: the pair of the visibility
     * with this linear index -- the dup-th of them in emission order if a
     * footprint as wide as the grid has several (the sort is stable). */
    int dup = 0;
    while (j - 1 - dup >= 0 && keys[j - 1 - dup] == tile &&
            vals[j - 1 - dup] == vals[j]) ++dup;
    int code = 0;
    for (int pv = r.v0; pv < r.v1; ++pv)
    {
        const int64_t pu = (int64_t) tile - (int64_t) pv * g.num_tiles_u;
        if (pu < r.u0 || pu >= r.u1) continue;
        code = pv * 32768 + (int) pu;
        if (dup-- == 0) break;
    }
    sorted_tile[slot] = code;
    if (sorted_vis_index) sorted_vis_index[slot] = (int32_t) i_vis;
    if (sorted_vis)
    {
        if (VIS_COMPLEX)
        {
            const int64_t at = 2 * i_vis * num_pols;
            sorted_vis[2 * slot] = vis[at];
            sorted_vis[2 * slot + 1] = vis[at + 1];
        }
        else
        {
            sorted_vis[slot] = vis[i_vis];
        }
    }
    if (sorted_weight) sorted_weight[slot] = weights[i_vis];
}

/* tile_offsets[t] += entries of tile t: what the reference's atomics leave. */
__global__ void __launch_bounds__(ES_TILE_THREADS) es_tile_advance_kernel(
        const uint32_t* __restrict__ start, const int64_t num_tiles,
        int32_t* __restrict__ tile_offsets)
{
    const int64_t t = (int64_t) blockIdx.x * ES_TILE_THREADS + threadIdx.x;
    if (t >= num_tiles) return;
    tile_offsets[t] += (int32_t) (start[t + 1] - start[t]);
}

/* Grow-only scratch per device, kept between calls. `small` holds what is
 * sized by tiles and CTAs, `big` what is sized by (visibility, tile) pairs. */
struct Scratch
{
    void* ptr[2];
    size_t bytes[2];
};
Scratch g_scratch[ES_TILE_MAX_DEVICES];
std::mutex g_mutex;

cudaError_t reserve(const int which, const size_t need, void** out)
{
    int dev = 0;
    cudaError_t err = cudaGetDevice(&dev);
    if (err != cudaSuccess) return err;
    if (dev < 0 || dev >= ES_TILE_MAX_DEVICES) return cudaErrorInvalidDevice;
    Scratch& s = g_scratch[dev];
    if (need > s.bytes[which])
    {
        if (s.ptr[which]) (void) cudaFree(s.ptr[which]); /* drains the device */
        s.ptr[which] = 0;
        s.bytes[which] = 0;
        const size_t want = need + need / 4 + 4096;
        err = cudaMalloc(&s.ptr[which], want);
        if (err != cudaSuccess) return err;
        s.bytes[which] = want;
    }
    *out = s.ptr[which];
    return cudaSuccess;
}

size_t align256(const size_t n) { return (n + 255) & ~(size_t) 255; }

int bits_for(const int64_t num_tiles)
{
    int bits = 1;
    while (((int64_t) 1 << bits) < num_tiles) ++bits;
    return bits;
}

template<typename T>
int count_t(const EsTileGeom* g, const T* uvw, const T* freqs, int32_t* counts,
        int32_t* offsets, int32_t* skipped, int* num_entries, cudaStream_t st)
{
    const int64_t total = g->num_rows * g->num_chan;
    const int64_t num_tiles = g->num_tiles_u * g->num_tiles_v;
    *num_entries = 0;
    if (num_tiles <= 0) return 0;
    cudaError_t err;
    std::lock_guard<std::mutex> lock(g_mutex);
    /* small scratch: inv_wl[num_chan] | scan temp */
    size_t temp_bytes = 0;
    err = cub::DeviceScan::InclusiveSum(0, temp_bytes, counts, offsets + 1,
            (int) num_tiles, st);
    if (err != cudaSuccess) return (int) err;
    const size_t o_temp = align256((size_t) g->num_chan * sizeof(T));
    char* small = 0;
    if ((err = reserve(0, o_temp + temp_bytes + 256, (void**) &small)) != cudaSuccess)
        return (int) err;
    T* inv_wl = (T*) small;
    if (total > 0)
    {
        es_tile_inv_wavelength_kernel<T><<<(unsigned) ((g->num_chan +
                ES_TILE_THREADS - 1) / ES_TILE_THREADS), ES_TILE_THREADS, 0,
                st>>>(g->num_chan, freqs, inv_wl);
        if ((err = cudaGetLastError()) != cudaSuccess) return (int) err;
        int sms = 148, dev = 0;
        (void) cudaGetDevice(&dev);
        (void) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        int64_t blocks = (total + ES_TILE_THREADS - 1) / ES_TILE_THREADS;
        if (blocks > (int64_t) sms * 8) blocks = (int64_t) sms * 8;
        es_tile_count_kernel<T><<<(unsigned) blocks, ES_TILE_THREADS, 0, st>>>(
                *g, uvw, inv_wl, counts, skipped);
        if ((err = cudaGetLastError()) != cudaSuccess) return (int) err;
    }
    /* offsets[0] = 0 (cleared by the caller), offsets[t + 1] = inclusive sum */
    err = cub::DeviceScan::InclusiveSum(small + o_temp, temp_bytes, counts,
            offsets + 1, (int) num_tiles, st);
    if (err != cudaSuccess) return (int) err;
    int32_t last = 0;
    err = cudaMemcpyAsync(&last, offsets + num_tiles, sizeof(int32_t),
            cudaMemcpyDeviceToHost, st);
    if (err == cudaSuccess) err = cudaStreamSynchronize(st);
    if (err != cudaSuccess) return (int) err;
    /* more than 2^31 - 1 pairs wrap the reference's int32 sums */
    if (last < 0) return ES_TILE_ERR_TOO_MANY;
    *num_entries = (int) last;
    return 0;
}

template<typename T>
int sort_t(const EsTileGeom* g, const T* uvw, const T* freqs, const T* vis,
        const T* weights, const int vis_complex, const int num_pols,
        int32_t* tile_offsets, T* sorted_uu, T* sorted_vv, T* sorted_vis,
        T* sorted_weight, int32_t* sorted_tile, int32_t* sorted_vis_index,
        const int64_t num_out, cudaStream_t st)
{
    const int64_t total = g->num_rows * g->num_chan;
    const int64_t num_tiles = g->num_tiles_u * g->num_tiles_v;
    if (total <= 0 || num_tiles <= 0) return 0;
    std::lock_guard<std::mutex> lock(g_mutex);
    cudaError_t err;
    const int64_t ctas = (total + ES_TILE_CHUNK - 1) / ES_TILE_CHUNK;

    /* small scratch: totals[ctas + 1] | bases[ctas + 1] | start[num_tiles + 1]
     * | inv_wl[num_chan] | scan temp */
    size_t scan_bytes = 0;
    err = cub::DeviceScan::ExclusiveSum(0, scan_bytes, (uint64_t*) 0,
            (uint64_t*) 0, (int) (ctas + 1), st);
    if (err != cudaSuccess) return (int) err;
    const size_t o_totals = 0;
    const size_t o_bases = o_totals + align256((size_t) (ctas + 1) * 8);
    const size_t o_start = o_bases + align256((size_t) (ctas + 1) * 8);
    const size_t o_invwl = o_start + align256((size_t) (num_tiles + 1) * 4);
    const size_t o_temp = o_invwl + align256((size_t) g->num_chan * sizeof(T));
    char* small = 0;
    if ((err = reserve(0, o_temp + scan_bytes + 256, (void**) &small)) != cudaSuccess)
        return (int) err;
    uint64_t* totals = (uint64_t*) (small + o_totals);
    uint64_t* bases = (uint64_t*) (small + o_bases);
    uint32_t* start = (uint32_t*) (small + o_start);
    T* inv_wl = (T*) (small + o_invwl);
    es_tile_inv_wavelength_kernel<T><<<(unsigned) ((g->num_chan +
            ES_TILE_THREADS - 1) / ES_TILE_THREADS), ES_TILE_THREADS, 0, st>>>(
            g->num_chan, freqs, inv_wl);
    if ((err = cudaGetLastError()) != cudaSuccess) return (int) err;

    /* pairs per CTA -> first pair of every CTA (and their number, read back) */
    err = cudaMemsetAsync(totals + ctas, 0, 8, st);
    if (err != cudaSuccess) return (int) err;
    es_tile_totals_kernel<T><<<(unsigned) ctas, ES_TILE_THREADS, 0, st>>>(*g,
            uvw, inv_wl, totals);
    if ((err = cudaGetLastError()) != cudaSuccess) return (int) err;
    err = cub::DeviceScan::ExclusiveSum(small + o_temp, scan_bytes, totals,
            bases, (int) (ctas + 1), st);
    if (err != cudaSuccess) return (int) err;
    uint64_t num_entries = 0;
    err = cudaMemcpyAsync(&num_entries, bases + ctas, 8, cudaMemcpyDeviceToHost,
            st);
    if (err == cudaSuccess) err = cudaStreamSynchronize(st);
    if (err != cudaSuccess) return (int) err;
    if (num_entries > (uint64_t) INT32_MAX) return ES_TILE_ERR_TOO_MANY;
    const int64_t m = (int64_t) num_entries;

    if (m > 0)
    {
        /* big scratch: keys x2 | vals x2 | sort temp */
        cub::DoubleBuffer<uint32_t> d_keys((uint32_t*) 0, (uint32_t*) 0);
        cub::DoubleBuffer<uint32_t> d_vals((uint32_t*) 0, (uint32_t*) 0);
        size_t sort_bytes = 0;
        const int end_bit = bits_for(num_tiles);
        err = cub::DeviceRadixSort::SortPairs(0, sort_bytes, d_keys, d_vals,
                (int) m, 0, end_bit, st);
        if (err != cudaSuccess) return (int) err;
        const size_t col = align256((size_t) m * 4);
        char* big = 0;
        if ((err = reserve(1, 4 * col + sort_bytes + 256, (void**) &big)) != cudaSuccess)
            return (int) err;
        d_keys = cub::DoubleBuffer<uint32_t>((uint32_t*) big,
                (uint32_t*) (big + col));
        d_vals = cub::DoubleBuffer<uint32_t>((uint32_t*) (big + 2 * col),
                (uint32_t*) (big + 3 * col));
        es_tile_emit_kernel<T><<<(unsigned) ctas, ES_TILE_THREADS, 0, st>>>(*g,
                uvw, inv_wl, bases, d_keys.Current(), d_vals.Current());
        if ((err = cudaGetLastError()) != cudaSuccess) return (int) err;
        err = cub::DeviceRadixSort::SortPairs(big + 4 * col, sort_bytes, d_keys,
                d_vals, (int) m, 0, end_bit, st);
        if (err != cudaSuccess) return (int) err;
        const unsigned nb = (unsigned) ((m + 1 + ES_TILE_THREADS - 1) / ES_TILE_THREADS);
        es_tile_bounds_kernel<<<nb, ES_TILE_THREADS, 0, st>>>(d_keys.Current(),
                m, num_tiles, start);
        if ((err = cudaGetLastError()) != cudaSuccess) return (int) err;
        if (vis_complex)
            es_tile_fill_kernel<T, true><<<nb, ES_TILE_THREADS, 0, st>>>(*g, uvw,
                    inv_wl, vis, weights, num_pols, d_keys.Current(),
                    d_vals.Current(), start, tile_offsets, m, num_out,
                    sorted_uu, sorted_vv, sorted_vis, sorted_weight,
                    sorted_tile, sorted_vis_index);
        else
            es_tile_fill_kernel<T, false><<<nb, ES_TILE_THREADS, 0, st>>>(*g, uvw,
                    inv_wl, vis, weights, num_pols, d_keys.Current(),
                    d_vals.Current(), start, tile_offsets, m, num_out,
                    sorted_uu, sorted_vv, sorted_vis, sorted_weight,
                    sorted_tile, sorted_vis_index);
        if ((err = cudaGetLastError()) != cudaSuccess) return (int) err;
        es_tile_advance_kernel<<<(unsigned) ((num_tiles + ES_TILE_THREADS - 1)
                / ES_TILE_THREADS), ES_TILE_THREADS, 0, st>>>(start, num_tiles,
                tile_offsets);
        if ((err = cudaGetLastError()) != cudaSuccess) return (int) err;
    }
    return 0;
}

} // namespace

extern "C" {

int es_tile_count(const EsTileGeom* g, const void* uvw, const void* freqs,
        int is_double, int32_t* counts, int32_t* offsets, int32_t* skipped,
        int* num_entries, void* stream)
{
    return is_double ?
            count_t<double>(g, (const double*) uvw, (const double*) freqs,
                    counts, offsets, skipped, num_entries, (cudaStream_t) stream) :
            count_t<float>(g, (const float*) uvw, (const float*) freqs,
                    counts, offsets, skipped, num_entries, (cudaStream_t) stream);
}

int es_tile_sort(const EsTileGeom* g, const void* uvw, const void* freqs,
        int is_double, const void* vis, const void* weights, int vis_complex,
        int num_pols, int32_t* tile_offsets, void* sorted_uu, void* sorted_vv,
        void* sorted_vis, void* sorted_weight, int32_t* sorted_tile,
        int32_t* sorted_vis_index, int64_t num_out, void* stream)
{
    return is_double ?
            sort_t<double>(g, (const double*) uvw, (const double*) freqs,
                    (const double*) vis, (const double*) weights, vis_complex,
                    num_pols, tile_offsets, (double*) sorted_uu,
                    (double*) sorted_vv, (double*) sorted_vis,
                    (double*) sorted_weight, sorted_tile, sorted_vis_index,
                    num_out, (cudaStream_t) stream) :
            sort_t<float>(g, (const float*) uvw, (const float*) freqs,
                    (const float*) vis, (const float*) weights, vis_complex,
                    num_pols, tile_offsets, (float*) sorted_uu,
                    (float*) sorted_vv, (float*) sorted_vis,
                    (float*) sorted_weight, sorted_tile, sorted_vis_index,
                    num_out, (cudaStream_t) stream);
}

void es_tile_release_scratch(void)
{
    std::lock_guard<std::mutex> lock(g_mutex);
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 ||
            dev >= ES_TILE_MAX_DEVICES) return;
    for (int k = 0; k < 2; ++k)
    {
        if (g_scratch[dev].ptr[k]) (void) cudaFree(g_scratch[dev].ptr[k]);
        g_scratch[dev].ptr[k] = 0;
        g_scratch[dev].bytes[k] = 0;
    }
}

} // extern "C"
