/*
 * Load-balanced scatter (gridding) of the sorted visibility records onto a
 * batch of resident w-planes, sm_100a.
 *
 * Replaces the reference's one-thread-per-visibility kernel with global
 * atomics (sdp_gridder_uvw_es_fft_kernels.cu:126-270, solving = 1) and its
 * per-plane cudaMemset (sdp_gridder_uvw_es_fft.cpp:586).
 *
 *   es_work_plan_kernel   one thread per (uv tile, w-plane) of the batch:
 *                         binary search of its record range, split of ranges
 *                         longer than `split` into slices, one work unit per
 *                         slice appended to a list (warp-aggregated atomics).
 *                         Real uv coverage is extremely uneven (SKA-Low: one
 *                         64x64 tile holds ~1e7 of 1.3e8 records on the first
 *                         planes); without the split one CTA does all of it.
 *   es_scatter2_kernel    persistent CTAs pull units from the list. A unit
 *                         accumulates its records into a 64x64 tile in shared
 *                         memory WITHOUT atomics: each warp owns a band of 16
 *                         tile rows. fp32: every warp walks the unit's records
 *                         on its own (no CTA barrier inside a unit), stages
 *                         the 2 (W+1) + 1 ES weights of the records touching
 *                         its band (one record per lane) and accumulates them,
 *                         its lanes covering (rows x W+1 columns) of a
 *                         footprint. fp64: the weights are staged once per
 *                         record CTA-wide, then each warp visits the records
 *                         of its band (ballot-compacted). Unsplit units
 *                         flush the tile with coalesced 16-byte stores (this
 *                         is also the zero fill of the plane); split units
 *                         write partial tiles to scratch and the last slice to
 *                         finish sums them in slice order, so the result does
 *                         not depend on scheduling (bitwise reproducible).
 */
#include <stdint.h>

#include <cuda_runtime.h>

#include "es_common.cuh"
#include "es_device.h"

using namespace es;

namespace {

/* per_plane != 0: one item per (tile, plane); else one item per tile covering
 * the whole plane range (gather side). */
__global__ void __launch_bounds__(256) es_work_plan_kernel(EsGeom g,
        const uint64_t* __restrict__ keys,
        const uint32_t* __restrict__ tile_start, int first_plane, int nplanes,
        int per_plane, uint32_t split, uint32_t max_units, uint32_t max_split,
        uint32_t max_partials, int skip_empty, EsUnit* __restrict__ units,
        uint32_t* __restrict__ counters)
{
    const int ntiles = g.ntile * g.ntile;
    const int nitems = per_plane ? ntiles * nplanes : ntiles;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    uint32_t a = 0, b = 0;
    int tile = 0, q = 0;
    uint32_t k = 0;
    if (i < nitems)
    {
        /* plane-major: the tiles of one plane are consecutive units, so the
         * stores of concurrently running CTAs fall into the same plane rows */
        q = per_plane ? i / ntiles : 0;
        tile = per_plane ? i - q * ntiles : i;
        if (per_plane == 2)
        {
            /* tile-major (fp64 gridding with per-record weight tables): the
             * planes of one tile are consecutive units, so the records and
             * weights of the tile are re-read from L2 instead of HBM */
            tile = i / nplanes;
            q = i - tile * nplanes;
        }
        const int p_lo = first_plane + q;
        const int p_hi = per_plane ? p_lo : first_plane + nplanes - 1;
        es_entry_range(g, keys, tile_start, tile, p_lo, p_hi, a, b);
        const uint32_t n = b - a;
        k = (n == 0 && skip_empty) ? 0u : 1u;
        if (n > split) k = (n + split - 1) / split;
    }
    int split_slot = -1;
    uint32_t partial_base = 0;
    if (k > 1 && per_plane)
    {
        partial_base = atomicAdd(&counters[CTR_PARTIALS], k);
        split_slot = (int) atomicAdd(&counters[CTR_SPLIT], 1u);
        if (partial_base + k > max_partials || (uint32_t) split_slot >= max_split)
        {
            k = 1; /* out of scratch: correct, only slower */
            split_slot = -1;
        }
    }
    /* warp-aggregated allocation of k list slots */
    uint32_t incl = k;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1)
    {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    const uint32_t warp_total = __shfl_sync(0xffffffffu, incl, 31);
    uint32_t base = 0;
    if (lane == 31 && warp_total > 0)
        base = atomicAdd(&counters[CTR_UNITS], warp_total);
    base = __shfl_sync(0xffffffffu, base, 31) + incl - k;
    if (k == 0 || base + k > max_units) return; /* host sizes the list for the worst case */
    const uint32_t n = b - a;
    const uint32_t per = (k > 1) ? (n + k - 1) / k : n;
    for (uint32_t s = 0; s < k; ++s)
    {
        EsUnit u;
        u.a = a + s * per;
        u.b = (s + 1 == k) ? b : a + (s + 1) * per;
        u.tile = tile;
        u.plane_q = q;
        u.num_slices = (int) k;
        u.slice = (int) s;
        u.split_slot = split_slot;
        u.partial_base = partial_base;
        units[base + s] = u;
    }
}

/* ----------------------------------------------------------- occupancy */

/* One thread per (uv tile, w-plane) of the batch: sets the bit of every pair
 * that has records (layout: es_device.h, es_occupancy_words). Everything
 * downstream moves only occupied tiles / tile rows: the scatter writes them,
 * the line FFTs read (gridding) or store (degridding) them, the gather reads
 * them. Real coverage is sparse per plane (SKA-Low snapshot: < 0.3 % of the
 * tiles, 0.5 .. 50 % of the tile rows). */
__global__ void __launch_bounds__(256) es_occupancy_kernel(EsGeom g,
        const uint64_t* __restrict__ keys,
        const uint32_t* __restrict__ tile_start, int first_plane, int nplanes,
        uint32_t* __restrict__ occ, unsigned long long* __restrict__ totals)
{
    const int ntiles = g.ntile * g.ntile;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= ntiles * nplanes) return;
    const int q = i / ntiles, tile = i - q * ntiles;
    uint32_t a, b;
    es_entry_range(g, keys, tile_start, tile, first_plane + q, first_plane + q,
            a, b);
    if (a == b) return;
    const int tu = tile / g.ntile, tv = tile - tu * g.ntile;
    const int nw = (g.ntile + 31) >> 5;
    uint32_t* o = occ + (size_t) q * (size_t) ((g.ntile + 1) * nw);
    atomicOr(&o[tu * nw + (tv >> 5)], 1u << (tv & 31));
    const uint32_t row_bit = 1u << (tu & 31);
    const uint32_t before = atomicOr(&o[g.ntile * nw + (tu >> 5)], row_bit);
    /* call totals for the statistics: occupied (tile, plane) and
     * (tile row, plane) pairs */
    if (totals)
    {
        atomicAdd(&totals[0], 1ULL);
        if (!(before & row_bit)) atomicAdd(&totals[1], 1ULL);
    }
}

/* ------------------------------------------------------------- scatter */

#define ES_SC_THREADS 128
#define ES_SC_BAND (ES_TILE / (ES_SC_THREADS / 32))

template<typename T> struct ScatterCfg;
template<> struct ScatterCfg<float> { static const int CHUNK = 128; };
template<> struct ScatterCfg<double> { static const int CHUNK = 32; };

/* NS: footprint slots per axis the staging arrays hold (support + 1 <= NS). */
template<typename T, int NS>
__global__ void __launch_bounds__(ES_SC_THREADS) es_scatter2_kernel(EsGeom g,
        const typename Prec<T>::Rec* __restrict__ records,
        const EsUnit* __restrict__ units, uint32_t* __restrict__ counters,
        int first_plane, typename Prec<T>::T2* __restrict__ planes,
        typename Prec<T>::T2* __restrict__ scratch,
        uint32_t* __restrict__ split_done, const T* __restrict__ kweights)
{
    typedef typename Prec<T>::T2 T2;
    typedef typename Prec<T>::Rec Rec;
    constexpr int CHUNK = ScatterCfg<T>::CHUNK;
    extern __shared__ __align__(16) unsigned char es_sc_smem[];
    T2* tile = (T2*) es_sc_smem;                       /* [ES_TILE][ES_TILE] */
    T* s_ku = (T*) (tile + ES_TILE * ES_TILE);         /* [CHUNK][NS] */
    T* s_kre = s_ku + CHUNK * NS;                      /* [CHUNK][NS] */
    T* s_kim = s_kre + CHUNK * NS;                     /* [CHUNK][NS] */
    int* s_pos = (int*) (s_kim + CHUNK * NS);          /* [CHUNK] */
    uint32_t* s_band = (uint32_t*) (s_pos + CHUNK);    /* [CHUNK] bands a record touches (fp64 path) */
    __shared__ uint32_t s_work;
    __shared__ int s_last;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int gsz = g.grid_size;
    const int ns = g.support + 1;          /* taps per axis */
    const int rps = 32 / ns;               /* footprint rows per warp step */
    const int c = lane % ns, rsub = lane / ns;
    const bool lane_active = rsub < rps;
    const T beta = (T) g.beta;
    const T inv_hs = (T) (2.0 / (double) g.support);
    const uint32_t num_units = counters[CTR_UNITS];
    constexpr int VEC = 16 / (int) sizeof(T2);
    constexpr int VPR = ES_TILE / VEC;
    const size_t plane_cells = (size_t) gsz * gsz;

    for (;;)
    {
        __syncthreads(); /* s_work / tile of the previous unit no longer read */
        if (tid == 0) s_work = atomicAdd(&counters[CTR_NEXT], 1u);
        __syncthreads();
        const uint32_t w = s_work;
        if (w >= num_units) break;
        const EsUnit u = units[w];
        const int tu = u.tile / g.ntile, tv = u.tile - tu * g.ntile;
        const int rows_valid = min(ES_TILE, gsz - tu * ES_TILE);
        const int cols_valid = min(ES_TILE, gsz - tv * ES_TILE);
        const int vec_valid = cols_valid / VEC; /* grid_size is even */
        T2* out = planes + (size_t) u.plane_q * plane_cells +
                (size_t) (tu * ES_TILE) * gsz + (size_t) (tv * ES_TILE);
        if (u.a == u.b)
        {
            /* nothing lands here on this plane: the fused "memset" */
            const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int v = tid; v < ES_TILE * VPR; v += ES_SC_THREADS)
            {
                const int r = v / VPR, cv = v % VPR;
                if (r < rows_valid && cv < vec_valid)
                    *reinterpret_cast<float4*>(out + (size_t) r * gsz + cv * VEC) = zero;
            }
            continue;
        }
        for (int v = tid; v < ES_TILE * VPR; v += ES_SC_THREADS)
            reinterpret_cast<float4*>(tile)[v] = make_float4(0.f, 0.f, 0.f, 0.f);

        const int plane = first_plane + u.plane_q;
        const int band_lo = warp * ES_SC_BAND;
        const int band_hi = min(band_lo + ES_SC_BAND, rows_valid) - 1;
        __syncthreads(); /* tile zeroed */
        /* Each warp walks the unit's records on its own (no CTA barrier
         * inside a unit): it picks the records whose footprint rows touch
         * its band, stages their ES weights into its private slice of the
         * staging arrays (one record per lane, kernels.cu:201-221 with the
         * signs of :233-235 folded in) and accumulates them in record order.
         * A record touching two bands is staged by both warps. */
        if constexpr (sizeof(T) == 4)
        {
            constexpr int WCH = CHUNK / (ES_SC_THREADS / 32); /* per warp */
            T* w_ku = s_ku + warp * WCH * NS;
            T* w_kre = s_kre + warp * WCH * NS;
            T* w_kim = s_kim + warp * WCH * NS;
            int* w_pos = s_pos + warp * WCH;
            for (uint32_t base = u.a; base < u.b; base += WCH)
            {
                const uint32_t idx = base + (uint32_t) lane;
                bool mine = false;
                Rec r;
                if (lane < WCH && idx < u.b)
                {
                    r = records[idx];
                    mine = max((int) r.lu0, band_lo) <=
                            min((int) r.lu0 + ns - 1, band_hi);
                }
                const unsigned m = __ballot_sync(0xffffffffu, mine);
                if (m == 0u) continue; /* warp-uniform */
                __syncwarp(); /* the previous group's weights are consumed */
                if (mine)
                {
                    const int e = __popc(m & ((1u << lane) - 1u));
                    const int dp = plane - r.p0;
                    const T kw = (dp >= 0 && dp <= g.support) ?
                            es_psi(beta, ((T) dp + r.dw) * inv_hs) : (T) 0;
                    const T vre = r.vre * kw, vim = r.vim * kw;
                    const int lu0 = r.lu0, lv0 = r.lv0;
                    w_pos[e] = (lu0 << 16) | (lv0 & 0xffff);
#pragma unroll
                    for (int k = 0; k < NS; ++k)
                    {
                        if (k < ns)
                        {
                            T ku = es_psi(beta, ((T) k + r.du) * inv_hs);
                            T kv = es_psi(beta, ((T) k + r.dv) * inv_hs);
                            if ((lu0 + k) & 1) ku = -ku;
                            if ((lv0 + k) & 1) kv = -kv;
                            w_ku[e * NS + k] = ku;
                            w_kre[e * NS + k] = vre * kv;
                            w_kim[e * NS + k] = vim * kv;
                        }
                    }
                }
                __syncwarp();
                const int cnt = __popc(m);
                for (int e = 0; e < cnt; ++e)
                {
                    const int pos = w_pos[e];
                    const int lu0 = pos >> 16;
                    const int r_lo = max(lu0, band_lo);
                    const int r_hi = min(lu0 + ns - 1, band_hi);
                    const int col = (int) (short) (pos & 0xffff) + c;
                    if (lane_active && col >= 0 && col < cols_valid)
                    {
                        const T pre = w_kre[e * NS + c], pim = w_kim[e * NS + c];
                        /* footprint rows of this lane, MS at a time: all
                         * loads, then all updates, then all stores (distinct
                         * cells): one shared-memory round trip per record */
                        constexpr int MS = (NS <= 9) ? 3 : 4;
                        for (int rb = r_lo + rsub; rb <= r_hi; rb += MS * rps)
                        {
                            T2 cell[MS];
                            T ku[MS];
#pragma unroll
                            for (int mm = 0; mm < MS; ++mm)
                            {
                                const int rr = rb + mm * rps;
                                if (rr <= r_hi)
                                {
                                    ku[mm] = w_ku[e * NS + (rr - lu0)];
                                    cell[mm] = tile[rr * ES_TILE + col];
                                }
                            }
#pragma unroll
                            for (int mm = 0; mm < MS; ++mm)
                            {
                                const int rr = rb + mm * rps;
                                if (rr <= r_hi)
                                {
                                    cell[mm].x += pre * ku[mm];
                                    cell[mm].y += pim * ku[mm];
                                    tile[rr * ES_TILE + col] = cell[mm];
                                }
                            }
                        }
                    }
                }
            }
        }
        else
        {
        if (kweights)
        {
            /* fp64 with the ES weights of every record already evaluated
             * (es_build_records_kernel: ku[ns], kv[ns] with the signs folded
             * in, kw[ns] -- they do not depend on the plane): the
             * warp-independent walk of the fp32 path, no CTA barrier inside a
             * unit. 32 records per step, one per lane (header, w weight);
             * those touching the warp's band are staged WCH at a time (lanes
             * k < ns copy ku[k], kv[k] x visibility) and accumulated in
             * record order. ncu on the CTA-wide form below (C3): 44 % of the
             * stall samples at its chunk barrier, 20 % issue-active. */
            constexpr int WCH = CHUNK / (ES_SC_THREADS / 32); /* per warp */
            T* w_ku = s_ku + warp * WCH * NS;
            T* w_kre = s_kre + warp * WCH * NS;
            T* w_kim = s_kim + warp * WCH * NS;
            int* w_pos = s_pos + warp * WCH;
            const int kstride = 3 * ns;
            for (uint32_t base = u.a; base < u.b; base += 32)
            {
                const uint32_t idx = base + (uint32_t) lane;
                bool mine = false;
                T vre = (T) 0, vim = (T) 0;
                int lu0 = 0, lv0 = 0;
                if (idx < u.b)
                {
                    const Rec& r = records[idx];
                    lu0 = r.lu0;
                    lv0 = r.lv0;
                    mine = max(lu0, band_lo) <= min(lu0 + ns - 1, band_hi);
                    if (mine)
                    {
                        const int dp = plane - r.p0;
                        const T kw = (dp >= 0 && dp <= g.support) ?
                                kweights[(size_t) idx * kstride + 2 * ns + dp] :
                                (T) 0;
                        vre = r.vre * kw;
                        vim = r.vim * kw;
                    }
                }
                unsigned m = __ballot_sync(0xffffffffu, mine);
                while (m) /* warp-uniform */
                {
                    __syncwarp(); /* the previous batch's weights are consumed */
                    /* all global loads of the batch first (the weights come
                     * from L2 / HBM: their latencies overlap), then the
                     * shared-memory stores */
                    int cnt = 0;
                    T bku[WCH], bkv[WCH], bvre[WCH], bvim[WCH];
                    int bpos[WCH];
#pragma unroll
                    for (int j = 0; j < WCH; ++j)
                    {
                        bku[j] = bkv[j] = bvre[j] = bvim[j] = (T) 0;
                        bpos[j] = 0;
                        if (m)
                        {
                            const int src = __ffs((int) m) - 1;
                            m &= m - 1u;
                            ++cnt;
                            bvre[j] = __shfl_sync(0xffffffffu, vre, src);
                            bvim[j] = __shfl_sync(0xffffffffu, vim, src);
                            const int slu = __shfl_sync(0xffffffffu, lu0, src);
                            const int slv = __shfl_sync(0xffffffffu, lv0, src);
                            bpos[j] = (slu << 16) | (slv & 0xffff);
                            if (lane < ns)
                            {
                                const T* kr = kweights +
                                        (size_t) (base + (uint32_t) src) * kstride;
                                bku[j] = kr[lane];
                                bkv[j] = kr[ns + lane];
                            }
                        }
                    }
#pragma unroll
                    for (int j = 0; j < WCH; ++j)
                    {
                        if (j < cnt)
                        {
                            if (lane < ns)
                            {
                                w_ku[j * NS + lane] = bku[j];
                                w_kre[j * NS + lane] = bvre[j] * bkv[j];
                                w_kim[j * NS + lane] = bvim[j] * bkv[j];
                            }
                            if (lane == 0) w_pos[j] = bpos[j];
                        }
                    }
                    __syncwarp();
                    for (int e = 0; e < cnt; ++e)
                    {
                        const int pos = w_pos[e];
                        const int elu0 = pos >> 16;
                        const int r_lo = max(elu0, band_lo);
                        const int r_hi = min(elu0 + ns - 1, band_hi);
                        const int col = (int) (short) (pos & 0xffff) + c;
                        if (lane_active && col >= 0 && col < cols_valid)
                        {
                            const T pre = w_kre[e * NS + c], pim = w_kim[e * NS + c];
                            constexpr int MS = (NS <= 9) ? 3 : 4;
                            for (int rb = r_lo + rsub; rb <= r_hi; rb += MS * rps)
                            {
                                T2 cell[MS];
                                T ku[MS];
#pragma unroll
                                for (int mm = 0; mm < MS; ++mm)
                                {
                                    const int rr = rb + mm * rps;
                                    if (rr <= r_hi)
                                    {
                                        ku[mm] = w_ku[e * NS + (rr - elu0)];
                                        cell[mm] = tile[rr * ES_TILE + col];
                                    }
                                }
#pragma unroll
                                for (int mm = 0; mm < MS; ++mm)
                                {
                                    const int rr = rb + mm * rps;
                                    if (rr <= r_hi)
                                    {
                                        cell[mm].x += pre * ku[mm];
                                        cell[mm].y += pim * ku[mm];
                                        tile[rr * ES_TILE + col] = cell[mm];
                                    }
                                }
                            }
                        }
                    }
                }
            }
        }
        else
        /* fp64 without precomputed weights: the ES weights are too expensive
         * to evaluate twice for records touching two bands: CTA-wide staging,
         * one record per thread, then the band walk */
        for (uint32_t chunk = u.a; chunk < u.b; chunk += CHUNK)
        {
            const int n = (int) min((uint32_t) CHUNK, u.b - chunk);
            __syncthreads(); /* staging arrays free, tile zeroed */
            /* stage: one record per thread, every ES weight it needs once
             * (kernels.cu:201-221, signs of :233-235 folded in) */
            for (int e = tid; e < n; e += ES_SC_THREADS)
            {
                const Rec r = records[chunk + e];
                const int dp = plane - r.p0;
                const T kw = (dp >= 0 && dp <= g.support) ?
                        es_psi(beta, ((T) dp + r.dw) * inv_hs) : (T) 0;
                const T vre = r.vre * kw, vim = r.vim * kw;
                const int lu0 = r.lu0, lv0 = r.lv0;
                s_pos[e] = (lu0 << 16) | (lv0 & 0xffff);
                {
                    /* bit w: the footprint rows overlap the band of warp w */
                    const int b_lo = max(lu0, 0) / ES_SC_BAND;
                    const int b_hi = min(lu0 + ns - 1, rows_valid - 1) / ES_SC_BAND;
                    s_band[e] = (b_hi >= b_lo) ?
                            ((2u << b_hi) - (1u << b_lo)) : 0u;
                }
#pragma unroll
                for (int k = 0; k < NS; ++k)
                {
                    if (k < ns)
                    {
                        T ku = es_psi(beta, ((T) k + r.du) * inv_hs);
                        T kv = es_psi(beta, ((T) k + r.dv) * inv_hs);
                        if ((lu0 + k) & 1) ku = -ku;
                        if ((lv0 + k) & 1) kv = -kv;
                        s_ku[e * NS + k] = ku;
                        s_kre[e * NS + k] = vre * kv;
                        s_kim[e * NS + k] = vim * kv;
                    }
                }
            }
            __syncthreads();
            /* accumulate: this warp owns tile rows band_lo .. band_hi */
            for (int base = 0; base < n; base += 32)
            {
            /* the records of this group of 32 that touch the warp's band */
            unsigned todo = __ballot_sync(0xffffffffu, base + lane < n &&
                    ((s_band[min(base + lane, n - 1)] >> warp) & 1u));
            while (todo)
            {
                const int e = base + __ffs((int) todo) - 1;
                todo &= todo - 1u;
                const int pos = s_pos[e];
                const int lu0 = pos >> 16;
                const int r_lo = max(lu0, band_lo);
                const int r_hi = min(lu0 + ns - 1, band_hi);
                const int col = (int) (short) (pos & 0xffff) + c;
                if (lane_active && col >= 0 && col < cols_valid)
                {
                    const T pre = s_kre[e * NS + c], pim = s_kim[e * NS + c];
                    /* footprint rows of this lane, MS at a time: all loads,
                     * then all updates, then all stores (distinct cells), so
                     * the shared-memory round trip is paid once per record */
                    constexpr int MS = (NS <= 9) ? 3 : 4;
                    for (int rb = r_lo + rsub; rb <= r_hi; rb += MS * rps)
                    {
                        T2 cell[MS];
                        T ku[MS];
#pragma unroll
                        for (int m = 0; m < MS; ++m)
                        {
                            const int r = rb + m * rps;
                            if (r <= r_hi)
                            {
                                ku[m] = s_ku[e * NS + (r - lu0)];
                                cell[m] = tile[r * ES_TILE + col];
                            }
                        }
#pragma unroll
                        for (int m = 0; m < MS; ++m)
                        {
                            const int r = rb + m * rps;
                            if (r <= r_hi)
                            {
                                cell[m].x += pre * ku[m];
                                cell[m].y += pim * ku[m];
                                tile[r * ES_TILE + col] = cell[m];
                            }
                        }
                    }
                }
            }
            }
        }
        }
        __syncthreads();
        if (u.num_slices == 1)
        {
            for (int v = tid; v < ES_TILE * VPR; v += ES_SC_THREADS)
            {
                const int r = v / VPR, cv = v % VPR;
                if (r < rows_valid && cv < vec_valid)
                    *reinterpret_cast<float4*>(out + (size_t) r * gsz + cv * VEC) =
                            reinterpret_cast<const float4*>(tile)[v];
            }
            continue;
        }
        /* split unit: publish the partial tile; the last slice to finish
         * sums all partials in slice order and stores the result */
        float4* part = reinterpret_cast<float4*>(scratch +
                (size_t) (u.partial_base + (uint32_t) u.slice) * (ES_TILE * ES_TILE));
        for (int v = tid; v < ES_TILE * VPR; v += ES_SC_THREADS)
            part[v] = reinterpret_cast<const float4*>(tile)[v];
        __threadfence();
        __syncthreads();
        if (tid == 0)
        {
            const uint32_t prev = atomicAdd(&split_done[u.split_slot], 1u);
            s_last = (prev + 1u == (uint32_t) u.num_slices);
        }
        __syncthreads();
        if (!s_last) continue;
        __threadfence();
        const T2* all = scratch + (size_t) u.partial_base * (ES_TILE * ES_TILE);
        for (int cell = tid; cell < ES_TILE * ES_TILE; cell += ES_SC_THREADS)
        {
            const int r = cell / ES_TILE, cc = cell % ES_TILE;
            if (r >= rows_valid || cc >= cols_valid) continue;
            T sx = (T) 0, sy = (T) 0;
            for (int s = 0; s < u.num_slices; ++s)
            {
                const T2 v = __ldcg(all + (size_t) s * (ES_TILE * ES_TILE) + cell);
                sx += v.x;
                sy += v.y;
            }
            T2 o; o.x = sx; o.y = sy;
            out[(size_t) r * gsz + cc] = o;
        }
    }
}

template<typename T, int NS>
size_t scatter2_smem()
{
    typedef typename Prec<T>::T2 T2;
    return (size_t) ES_TILE * ES_TILE * sizeof(T2) +
            (size_t) ScatterCfg<T>::CHUNK * (3 * NS * sizeof(T) + 2 * sizeof(int));
}

template<typename T, int NS>
int launch_scatter2_t(const EsGeom* g, const void* records, const void* units,
        uint32_t* counters, int first_plane, void* planes, void* scratch,
        uint32_t* split_done, const void* kweights, cudaStream_t s)
{
    typedef typename Prec<T>::T2 T2;
    typedef typename Prec<T>::Rec Rec;
    /* persistent grid and shared-memory opt-in, per device (the attribute,
     * the SM count and the occupancy all belong to the device) */
    static int grid_sizes[ES_MAX_DEVICES];
    const size_t smem = scatter2_smem<T, NS>();
    int dev = 0;
    cudaError_t err = cudaGetDevice(&dev);
    if (err != cudaSuccess) return (int) err;
    if (dev < 0 || dev >= ES_MAX_DEVICES) return (int) cudaErrorInvalidDevice;
    int grid_size = __atomic_load_n(&grid_sizes[dev], __ATOMIC_ACQUIRE);
    if (grid_size == 0)
    {
        err = cudaFuncSetAttribute(es_scatter2_kernel<T, NS>,
                cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
        if (err != cudaSuccess) return (int) err;
        int sms = 0, per_sm = 0;
        err = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount,
                dev);
        if (err == cudaSuccess) err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(
                &per_sm, es_scatter2_kernel<T, NS>, ES_SC_THREADS, smem);
        if (err != cudaSuccess) return (int) err;
        if (per_sm < 1) per_sm = 1;
        grid_size = sms * per_sm;
        __atomic_store_n(&grid_sizes[dev], grid_size, __ATOMIC_RELEASE);
    }
    es_scatter2_kernel<T, NS><<<grid_size, ES_SC_THREADS, smem, s>>>(*g,
            (const Rec*) records, (const EsUnit*) units, counters, first_plane,
            (T2*) planes, (T2*) scratch, split_done, (const T*) kweights);
    return (int) cudaGetLastError();
}

} // namespace

extern "C" {

size_t es_unit_bytes(void)
{
    return sizeof(EsUnit);
}

size_t es_partial_tile_bytes(int is_double)
{
    return (size_t) ES_TILE * ES_TILE * (is_double ? 16 : 8);
}

size_t es_occupancy_words(const EsGeom* g)
{
    return (size_t) (g->ntile + 1) * (size_t) ((g->ntile + 31) >> 5);
}

int es_launch_occupancy(const EsGeom* g, const uint64_t* keys,
        const uint32_t* tile_start, int first_plane, int nplanes,
        uint32_t* occ, unsigned long long* totals, es_stream_t stream)
{
    cudaStream_t s = (cudaStream_t) stream;
    cudaError_t err = cudaMemsetAsync(occ, 0, (size_t) nplanes *
            es_occupancy_words(g) * sizeof(uint32_t), s);
    if (err != cudaSuccess) return (int) err;
    const int nitems = g->ntile * g->ntile * nplanes;
    es_occupancy_kernel<<<(nitems + 255) / 256, 256, 0, s>>>(*g, keys,
            tile_start, first_plane, nplanes, occ, totals);
    return (int) cudaGetLastError();
}

int es_launch_work_plan(const EsGeom* g, const uint64_t* keys,
        const uint32_t* tile_start, int first_plane, int nplanes,
        int per_plane, uint32_t split, uint32_t max_units, uint32_t max_split,
        uint32_t max_partials, int skip_empty, void* units, uint32_t* counters,
        uint32_t* split_done, es_stream_t stream)
{
    cudaStream_t s = (cudaStream_t) stream;
    cudaError_t err = cudaMemsetAsync(counters, 0, 8 * sizeof(uint32_t), s);
    if (err == cudaSuccess && split_done && max_split > 0)
        err = cudaMemsetAsync(split_done, 0, (size_t) max_split * sizeof(uint32_t), s);
    if (err != cudaSuccess) return (int) err;
    const int ntiles = g->ntile * g->ntile;
    const int nitems = per_plane ? ntiles * nplanes : ntiles;
    es_work_plan_kernel<<<(nitems + 255) / 256, 256, 0, s>>>(*g, keys,
            tile_start, first_plane, nplanes, per_plane, split, max_units,
            max_split, max_partials, skip_empty, (EsUnit*) units, counters);
    return (int) cudaGetLastError();
}

int es_launch_scatter2(const EsGeom* g, const void* records, const void* units,
        uint32_t* counters, int first_plane, void* planes, void* scratch,
        uint32_t* split_done, const void* kweights, es_stream_t stream)
{
    cudaStream_t s = (cudaStream_t) stream;
    /* footprint slots per axis the kernel is unrolled for: 9, 13 or 17 */
    const int ns = g->support + 1;
    if (g->is_double)
    {
        if (ns <= 9) return launch_scatter2_t<double, 9>(g, records, units,
                counters, first_plane, planes, scratch, split_done, kweights, s);
        if (ns <= 13) return launch_scatter2_t<double, 13>(g, records, units,
                counters, first_plane, planes, scratch, split_done, kweights, s);
        return launch_scatter2_t<double, 17>(g, records, units, counters,
                first_plane, planes, scratch, split_done, kweights, s);
    }
    if (ns <= 9) return launch_scatter2_t<float, 9>(g, records, units,
            counters, first_plane, planes, scratch, split_done, kweights, s);
    if (ns <= 13) return launch_scatter2_t<float, 13>(g, records, units,
            counters, first_plane, planes, scratch, split_done, kweights, s);
    return launch_scatter2_t<float, 17>(g, records, units, counters,
            first_plane, planes, scratch, split_done, kweights, s);
}

} // extern "C"
