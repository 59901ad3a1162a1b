/*
 * sm_100a kernels of the B200-native w-stacking ES gridder / degridder and
 * their C-ABI launchers (see es_device.h).
 *
 * What the reference does with one thread per visibility, global atomics and
 * a full re-scan of every visibility for every w-plane
 * (src/ska-sdp-func/grid_data/sdp_gridder_uvw_es_fft_kernels.cu:126-270) is
 * done here as:
 *   presort   - every visibility is expanded into one record per 64x64 uv tile
 *               its footprint overlaps; records are radix sorted by
 *               (tile, first owned w-plane), so the records that touch plane p
 *               of tile t are ONE contiguous range found by binary search;
 *   scatter   - es_scatter.cu: load-balanced (tile, plane) work units, the
 *               tile accumulated in shared memory without atomics and
 *               flushed once, so the reference's per-plane cudaMemset
 *               (es_fft.cpp:586) disappears;
 *   gather    - the same ranges as (tile, record slice) units, one thread per
 *               record, plain loads;
 *   w-screen  - one thread per 4 mirror pixels as in kernels.cu:429-679, but
 *               looping over a BATCH of resident planes with register
 *               accumulators (one image read-modify-write per batch), and with
 *               the phase reduced in fp64 before an fp32 sincospi.
 * All coordinates (grid position, w-plane position, w-screen phase) are fp64
 * in both precisions; fp32 is used for kernel weights, accumulation and the
 * FFT only.
 */
#include <stdint.h>

#include <cuda_runtime.h>
#include <cub/cub.cuh>

#include "es_common.cuh"
#include "es_device.h"

#define ES_C0 299792458.0           /* kernels.cu:159 */
#define ES_PI 3.1415926535897931    /* kernels.cu:10 */
#define ES_VIS_BITS 30              /* sort values: visibility index | entry number << 30 */
#define ES_VIS_MASK ((1u << ES_VIS_BITS) - 1u)

using namespace es;

namespace {

/* Where one visibility lands. Restates kernels.cu:157-196 with fp64
 * coordinates; the footprint is described by its first slot (iu0, iv0, p0)
 * and the offsets du = iu0 - pos_u ...: slot k sits at (k + du) / (W/2) in
 * kernel units, slots 0..W (the last one has weight 0 unless du == -W/2). */
struct Foot
{
    int iu0, iv0, p0;
    int u_lo, u_hi, v_lo, v_hi, p_lo, p_hi;
    double du, dv, dw, flip;
    bool valid;
};

template<typename T>
__device__ __forceinline__ Foot es_footprint(const EsGeom& g,
        const T* __restrict__ uvw, const T* __restrict__ freq, int row, int chan)
{
    Foot f;
    const double u = (double) uvw[3 * (size_t) row];
    const double v = (double) uvw[3 * (size_t) row + 1];
    const double w = (double) uvw[3 * (size_t) row + 2];
    const double hs = 0.5 * (double) g.support;
    f.flip = (g.do_wstacking && w < 0.0) ? -1.0 : 1.0;
    const double inv_wavelength = f.flip * (double) freq[chan] / ES_C0;
    const double pos_u = u * inv_wavelength * g.uv_scale;
    const double pos_v = v * inv_wavelength * g.uv_scale;
    const double pos_w = g.do_wstacking ?
            (w * inv_wavelength - g.min_plane_w) * g.w_scale : 0.0;
    const double cu = ceil(pos_u - hs), cv = ceil(pos_v - hs);
    const double cw = ceil(pos_w - hs);
    const bool finite = isfinite(pos_u) && isfinite(pos_v) && isfinite(pos_w);
    f.iu0 = (int) cu; f.iv0 = (int) cv; f.p0 = (int) cw;
    f.du = cu - pos_u; f.dv = cv - pos_v; f.dw = cw - pos_w;
    const int grid_min = -(g.grid_size / 2);
    const int grid_max = (g.grid_size - 1) / 2;
    f.u_lo = max(f.iu0, grid_min);
    f.u_hi = min((int) floor(pos_u + hs), grid_max);
    f.v_lo = max(f.iv0, grid_min);
    f.v_hi = min((int) floor(pos_v + hs), grid_max);
    f.p_lo = max(f.p0, g.plane_begin);
    f.p_hi = min((int) floor(pos_w + hs), g.plane_end - 1);
    f.valid = finite && f.u_lo <= f.u_hi && f.v_lo <= f.v_hi &&
            f.p_lo <= f.p_hi;
    return f;
}

/* ------------------------------------------------------------- presort */

template<typename T>
__global__ void __launch_bounds__(256) es_count_entries_kernel(EsGeom g,
        const T* __restrict__ uvw, const T* __restrict__ freq,
        uint8_t* __restrict__ counts)
{
    const int64_t nvis = (int64_t) g.num_rows * g.num_chan;
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nvis) return;
    const int row = (int) (i / g.num_chan), chan = (int) (i % g.num_chan);
    const Foot f = es_footprint<T>(g, uvw, freq, row, chan);
    int n = 0;
    if (f.valid)
    {
        const int half = g.grid_size / 2;
        const int tu = (f.u_hi + half) / ES_TILE - (f.u_lo + half) / ES_TILE + 1;
        const int tv = (f.v_hi + half) / ES_TILE - (f.v_lo + half) / ES_TILE + 1;
        n = tu * tv;
    }
    counts[i] = (uint8_t) n;
}

template<typename T>
__global__ void __launch_bounds__(256) es_emit_entries_kernel(EsGeom g,
        const T* __restrict__ uvw, const T* __restrict__ freq,
        const uint32_t* __restrict__ offsets, uint64_t* __restrict__ keys,
        uint32_t* __restrict__ vals)
{
    const int64_t nvis = (int64_t) g.num_rows * g.num_chan;
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nvis) return;
    const int row = (int) (i / g.num_chan), chan = (int) (i % g.num_chan);
    const Foot f = es_footprint<T>(g, uvw, freq, row, chan);
    if (!f.valid) return;
    const int half = g.grid_size / 2;
    const uint64_t pk = (uint64_t) (g.plane_end - g.plane_begin);
    const uint64_t kp = (uint64_t) (f.p_lo - g.plane_begin);
    const uint32_t first_slot = offsets[i];
    uint32_t slot = first_slot;
    for (int tu = (f.u_lo + half) / ES_TILE; tu <= (f.u_hi + half) / ES_TILE; ++tu)
    {
        for (int tv = (f.v_lo + half) / ES_TILE; tv <= (f.v_hi + half) / ES_TILE; ++tv)
        {
            keys[slot] = (uint64_t) (tu * g.ntile + tv) * pk + kp;
            /* visibility index (< 2^30, checked at plan creation) and, in the
             * two top bits, the number of this entry within the visibility */
            vals[slot] = (uint32_t) i | ((slot - first_slot) << ES_VIS_BITS);
            ++slot;
        }
    }
}

struct CastU8
{
    __host__ __device__ __forceinline__ uint32_t operator()(const uint8_t& x) const
    {
        return (uint32_t) x;
    }
};

__global__ void __launch_bounds__(256) es_tile_bounds_kernel(EsGeom g,
        const uint64_t* __restrict__ keys, int64_t num_entries,
        uint32_t* __restrict__ tile_start)
{
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= num_entries) return;
    const uint64_t pk = (uint64_t) (g.plane_end - g.plane_begin);
    const int64_t t = (int64_t) (keys[i] / pk);
    const int64_t tprev = (i > 0) ? (int64_t) (keys[i - 1] / pk) : -1;
    for (int64_t x = tprev + 1; x <= t; ++x) tile_start[x] = (uint32_t) i;
    if (i == num_entries - 1)
    {
        const int64_t ntiles = (int64_t) g.ntile * g.ntile;
        for (int64_t x = t + 1; x <= ntiles; ++x)
            tile_start[x] = (uint32_t) num_entries;
    }
}

template<typename T>
__global__ void __launch_bounds__(256) es_build_records_kernel(EsGeom g,
        const T* __restrict__ uvw, const T* __restrict__ freq,
        const typename Prec<T>::T2* __restrict__ vis,
        const T* __restrict__ weight, const uint64_t* __restrict__ keys,
        const uint32_t* __restrict__ vals, int64_t num_entries,
        typename Prec<T>::Rec* __restrict__ records,
        const uint32_t* __restrict__ offsets, uint32_t* __restrict__ entry_pos,
        T* __restrict__ kweights)
{
    const int64_t e = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= num_entries) return;
    const uint32_t idx = vals[e] & ES_VIS_MASK;
    /* degridding: where the k-th entry of visibility idx went in the sort */
    if (entry_pos)
        entry_pos[offsets[idx] + (vals[e] >> ES_VIS_BITS)] = (uint32_t) e;
    const int row = (int) (idx / (uint32_t) g.num_chan);
    const int chan = (int) (idx % (uint32_t) g.num_chan);
    const Foot f = es_footprint<T>(g, uvw, freq, row, chan);
    const uint64_t pk = (uint64_t) (g.plane_end - g.plane_begin);
    const int tile_id = (int) (keys[e] / pk);
    const int tu = tile_id / g.ntile, tv = tile_id % g.ntile;
    const int half = g.grid_size / 2;
    typename Prec<T>::Rec r;
    r.lu0 = (short) (f.iu0 + half - tu * ES_TILE);
    r.lv0 = (short) (f.iv0 + half - tv * ES_TILE);
    r.p0 = f.p0;
    r.du = (T) f.du; r.dv = (T) f.dv; r.dw = (T) f.dw;
    r.idx = idx;
    if (vis)
    {
        /* kernels.cu:163-168 */
        const typename Prec<T>::T2 v = vis[idx];
        const T wt = weight[idx];
        r.vre = v.x * wt;
        r.vim = v.y * wt * (T) f.flip;
    }
    else
    {
        r.vre = (T) f.flip;
        r.vim = (T) 0;
    }
    records[e] = r;
    if (kweights)
    {
        /* exactly the expressions of the scatter's staging loop
         * (es_scatter.cu; kernels.cu:201-221, signs of :233-235) */
        const int ns = g.support + 1;
        const T beta = (T) g.beta;
        const T inv_hs = (T) (2.0 / (double) g.support);
        T* kr = kweights + (size_t) e * (size_t) (3 * ns);
        const int lu0 = r.lu0, lv0 = r.lv0;
        for (int k = 0; k < ns; ++k)
        {
            T ku = es_psi(beta, ((T) k + r.du) * inv_hs);
            T kv = es_psi(beta, ((T) k + r.dv) * inv_hs);
            if ((lu0 + k) & 1) ku = -ku;
            if ((lv0 + k) & 1) kv = -kv;
            kr[k] = ku;
            kr[ns + k] = kv;
            kr[2 * ns + k] = es_psi(beta, ((T) k + r.dw) * inv_hs);
        }
    }
}

template<typename T>
__global__ void __launch_bounds__(256) es_plane_histogram_kernel(EsGeom g,
        const T* __restrict__ uvw, const T* __restrict__ freq,
        unsigned long long* __restrict__ hist)
{
    const int64_t nvis = (int64_t) g.num_rows * g.num_chan;
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nvis) return;
    const int row = (int) (i / g.num_chan), chan = (int) (i % g.num_chan);
    const Foot f = es_footprint<T>(g, uvw, freq, row, chan);
    if (!f.valid) return;
    atomicAdd(&hist[f.p_lo], 1ULL);
}

/* -------------------------------------------------------------- gather */

#define ES_GATHER_THREADS 128

/* One CTA per uv tile, one thread per record, all planes of the resident
 * batch: a record is read once per batch, its u and v weights are evaluated
 * once, and its accumulator gets one read-modify-write (owned by exactly one
 * thread, so no atomics and a deterministic sum).
 * NS: footprint slots per axis handled (support + 1 <= NS). */
/* Records [a, b) of one tile: acc[e] += sum over the planes of the batch and
 * the footprint cells of plane * kernel (kernels.cu:201-268, solving = 0). */
template<typename T, int NS>
__device__ __forceinline__ void es_gather_range(const EsGeom& g,
        const typename Prec<T>::Rec* __restrict__ records, const int tile_id,
        const uint32_t a, const uint32_t b, const int first_plane,
        const int nplanes, const typename Prec<T>::T2* __restrict__ planes,
        typename Prec<T>::T2* __restrict__ acc)
{
    typedef typename Prec<T>::T2 T2;
    typedef typename Prec<T>::Rec Rec;
    const int tu = tile_id / g.ntile, tv = tile_id % g.ntile;
    const int gsz = g.grid_size;
    const size_t plane_cells = (size_t) gsz * gsz;
    const int rows_valid = min(ES_TILE, gsz - tu * ES_TILE);
    const int cols_valid = min(ES_TILE, gsz - tv * ES_TILE);
    const T2* in = planes + (size_t) (tu * ES_TILE) * gsz +
            (size_t) (tv * ES_TILE);
    const int w_sup = g.support;
    const T beta = (T) g.beta;
    const T inv_hs = (T) (2.0 / (double) w_sup);

    for (uint32_t e = a + threadIdx.x; e < b; e += ES_GATHER_THREADS)
    {
        const Rec r = records[e];
        /* planes of this batch the record touches: p0 .. p0 + W */
        const int q_lo = max(r.p0 - first_plane, 0);
        const int q_hi = min(r.p0 + w_sup - first_plane, nplanes - 1);
        if (q_lo > q_hi) continue;
        const int lu0 = r.lu0, lv0 = r.lv0;
        T kv[NS], ku[NS];
        #pragma unroll
        for (int c = 0; c < NS; ++c)
        {
            const int col = lv0 + c, row = lu0 + c;
            const bool col_ok = (c <= w_sup) && col >= 0 && col < cols_valid;
            const bool row_ok = (c <= w_sup) && row >= 0 && row < rows_valid;
            /* kernels.cu:201-213, with the (-1)^(gu+gv) of :233-235 folded in */
            const T kc = col_ok ? es_psi(beta, ((T) c + r.dv) * inv_hs) : (T) 0;
            const T kr = row_ok ? es_psi(beta, ((T) c + r.du) * inv_hs) : (T) 0;
            kv[c] = (col & 1) ? -kc : kc;
            ku[c] = (row & 1) ? -kr : kr;
        }
        T sum_re = (T) 0, sum_im = (T) 0;
        for (int q = q_lo; q <= q_hi; ++q)
        {
            const int dp = first_plane + q - r.p0;
            const T kw = es_psi(beta, ((T) dp + r.dw) * inv_hs);
            const T2* pl = in + (size_t) q * plane_cells + lv0;
            T pl_re = (T) 0, pl_im = (T) 0;
            #pragma unroll
            for (int s = 0; s < NS; ++s)
            {
                if (ku[s] != (T) 0)
                {
                    const T2* prow = pl + (size_t) (lu0 + s) * gsz;
                    T row_re = (T) 0, row_im = (T) 0;
                    #pragma unroll
                    for (int c = 0; c < NS; ++c)
                    {
                        if (kv[c] != (T) 0)
                        {
                            const T2 cell = prow[c];
                            row_re += cell.x * kv[c];
                            row_im += cell.y * kv[c];
                        }
                    }
                    pl_re += row_re * ku[s];
                    pl_im += row_im * ku[s];
                }
            }
            sum_re += pl_re * kw;
            sum_im += pl_im * kw;
        }
        T2 t = acc[e];
        t.x += sum_re;
        t.y += sum_im;
        acc[e] = t;
    }
}

/* Load-balanced form: persistent CTAs pull (tile, record slice) units from
 * the list of es_work_plan_kernel (per_plane = 0). Real uv coverage is very
 * uneven (a Mid- or Low-like core puts ~10 % of all records into one tile):
 * with one CTA per tile a single SM does most of the call. A record belongs
 * to exactly one unit, so acc[e] has one writer. */
template<typename T, int NS>
__global__ void __launch_bounds__(ES_GATHER_THREADS) es_gather2_kernel(
        EsGeom g, const typename Prec<T>::Rec* __restrict__ records,
        const EsUnit* __restrict__ units, uint32_t* __restrict__ counters,
        int first_plane, int nplanes,
        const typename Prec<T>::T2* __restrict__ planes,
        typename Prec<T>::T2* __restrict__ acc)
{
    __shared__ uint32_t s_work;
    const uint32_t num_units = counters[CTR_UNITS];
    for (;;)
    {
        __syncthreads();
        if (threadIdx.x == 0) s_work = atomicAdd(&counters[CTR_NEXT], 1u);
        __syncthreads();
        const uint32_t w = s_work;
        if (w >= num_units) break;
        const EsUnit u = units[w];
        es_gather_range<T, NS>(g, records, u.tile, u.a, u.b, first_plane,
                nplanes, planes, acc);
    }
}

/* vis[i] += sum of the partial sums of visibility i's (at most 4) tile
 * entries, taken in entry order: one thread per visibility, no atomics, so
 * degridded visibilities are bitwise reproducible (kernels.cu:267-268 adds
 * the one sum of its one thread the same way). */
template<typename T>
__global__ void __launch_bounds__(256) es_finish_degrid_kernel(
        const typename Prec<T>::Rec* __restrict__ records,
        const typename Prec<T>::T2* __restrict__ acc,
        const uint32_t* __restrict__ offsets,
        const uint32_t* __restrict__ entry_pos, int64_t num_vis,
        typename Prec<T>::T2* __restrict__ vis)
{
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= num_vis) return;
    const uint32_t lo = offsets[i], hi = offsets[i + 1];
    if (lo == hi) return;
    T sx = 0, sy = 0, flip = 1;
    for (uint32_t k = lo; k < hi; ++k)
    {
        const uint32_t e = entry_pos[k];
        const typename Prec<T>::T2 t = acc[e];
        if (k == lo) flip = records[e].vre;
        sx += t.x;
        sy += t.y;
    }
    typename Prec<T>::T2 v = vis[i];
    v.x += sx;
    v.y += sy * flip;
    vis[i] = v;
}

/* ------------------------------------------------------------ w-screen */

/* (cos, sin)(2 pi sign w (n-1)) / n with the phase reduced in fp64:
 * kernels.cu:110-123. nm1 = n - 1, xn = 1 / n. */
template<typename T>
__device__ __forceinline__ void es_phasor(double w, double nm1, double xn,
        double sign, T& re, T& im);

template<>
__device__ __forceinline__ void es_phasor<float>(double w, double nm1,
        double xn, double sign, float& re, float& im)
{
    double t = w * nm1;            /* turns */
    t -= rint(t);
    float s, c;
    sincospif((float) (2.0 * sign * t), &s, &c);
    re = c * (float) xn;
    im = s * (float) xn;
}

template<>
__device__ __forceinline__ void es_phasor<double>(double w, double nm1,
        double xn, double sign, double& re, double& im)
{
    double t = w * nm1;
    t -= rint(t);
    double s, c;
    sincospi(2.0 * sign * t, &s, &c);
    re = c * xn;
    im = s * xn;
}

#define ES_SCREEN_BX 32
#define ES_SCREEN_BY 8

template<typename T>
__global__ void __launch_bounds__(ES_SCREEN_BX * ES_SCREEN_BY)
es_apply_screen_kernel(EsGeom g,
        const typename Prec<T>::T2* __restrict__ planes, int first_plane,
        int nplanes, T* __restrict__ dirty)
{
    typedef typename Prec<T>::T2 T2;
    const int i = blockIdx.x * ES_SCREEN_BX + threadIdx.x;
    const int j = blockIdx.y * ES_SCREEN_BY + threadIdx.y;
    const int n = g.image_size, half = n / 2, gsz = g.grid_size;
    if (i > half || j > half) return;
    const double l = g.pixel_size * (double) i, m = g.pixel_size * (double) j;
    const double sos = l * l + m * m;
    const double nm1 = (-sos) / (sqrt(1.0 - sos) + 1.0);
    const double xn = 1.0 / (nm1 + 1.0);
    T s_pp = 0, s_pn = 0, s_np = 0, s_nn = 0;
    const size_t centre = (size_t) (gsz / 2) * gsz + (size_t) (gsz / 2);
    const ptrdiff_t dj = (ptrdiff_t) j * gsz;
    for (int q = 0; q < nplanes; ++q)
    {
        T cr = (T) 1, ci = (T) 0;
        if (g.do_wstacking)
        {
            const double w = (double) (first_plane + q) * g.inv_w_scale +
                    g.min_plane_w;
            es_phasor<T>(w, nm1, xn, -1.0, cr, ci);
        }
        const T2* p = planes + (size_t) q * gsz * gsz + centre;
        /* kernels.cu:487-502 */
        const T2 a = p[dj + i], b = p[-dj + i], c = p[dj - i], d = p[-dj - i];
        s_pp += a.x * cr - a.y * ci;
        s_pn += b.x * cr - b.y * ci;
        s_np += c.x * cr - c.y * ci;
        s_nn += d.x * cr - d.y * ci;
    }
    if ((i + j) & 1)
    {
        s_pp = -s_pp; s_pn = -s_pn; s_np = -s_np; s_nn = -s_nn;
    }
    /* kernels.cu:516-545: image offsets -N/2 .. N/2-1 */
    T* img = dirty + (size_t) half * n + half;
    const ptrdiff_t ij = (ptrdiff_t) j * n;
    if (i < half && j < half) img[ij + i] += s_pp;
    if (i > 0 && j < half) img[ij - i] += s_np;
    if (j > 0 && i < half) img[-ij + i] += s_pn;
    if (i > 0 && j > 0) img[-ij - i] += s_nn;
}

template<typename T>
__global__ void __launch_bounds__(ES_SCREEN_BX * ES_SCREEN_BY)
es_reverse_screen_kernel(EsGeom g, const T* __restrict__ dirty,
        int first_plane, int nplanes,
        typename Prec<T>::T2* __restrict__ planes)
{
    typedef typename Prec<T>::T2 T2;
    const int i = blockIdx.x * ES_SCREEN_BX + threadIdx.x;
    const int j = blockIdx.y * ES_SCREEN_BY + threadIdx.y;
    const int n = g.image_size, half = n / 2, gsz = g.grid_size;
    if (i > half || j > half) return;
    const T* img = dirty + (size_t) half * n + half;
    const ptrdiff_t ij = (ptrdiff_t) j * n;
    /* kernels.cu:580-613 */
    const bool w_pp = (i < half && j < half), w_np = (i > 0 && j < half);
    const bool w_pn = (j > 0 && i < half), w_nn = (i > 0 && j > 0);
    T d_pp = w_pp ? img[ij + i] : (T) 0;
    T d_np = w_np ? img[ij - i] : (T) 0;
    T d_pn = w_pn ? img[-ij + i] : (T) 0;
    T d_nn = w_nn ? img[-ij - i] : (T) 0;
    if ((i + j) & 1)
    {
        d_pp = -d_pp; d_np = -d_np; d_pn = -d_pn; d_nn = -d_nn;
    }
    const double l = g.pixel_size * (double) i, m = g.pixel_size * (double) j;
    const double sos = l * l + m * m;
    const double nm1 = (-sos) / (sqrt(1.0 - sos) + 1.0);
    const double xn = 1.0 / (nm1 + 1.0);
    const size_t centre = (size_t) (gsz / 2) * gsz + (size_t) (gsz / 2);
    const ptrdiff_t dj = (ptrdiff_t) j * gsz;
    for (int q = 0; q < nplanes; ++q)
    {
        T cr = (T) 1, ci = (T) 0;
        if (g.do_wstacking)
        {
            const double w = (double) (first_plane + q) * g.inv_w_scale +
                    g.min_plane_w;
            es_phasor<T>(w, nm1, xn, 1.0, cr, ci);
        }
        T2* p = planes + (size_t) q * gsz * gsz + centre;
        T2 o;
        /* kernels.cu:644-676 */
        if (w_pp) { o.x = cr * d_pp; o.y = ci * d_pp; p[dj + i] = o; }
        if (w_pn) { o.x = cr * d_pn; o.y = ci * d_pn; p[-dj + i] = o; }
        if (w_np) { o.x = cr * d_np; o.y = ci * d_np; p[dj - i] = o; }
        if (w_nn) { o.x = cr * d_nn; o.y = ci * d_nn; p[-dj - i] = o; }
    }
}

/* Zero every plane cell the reverse w-screen does not write (the padding
 * around the image footprint): replaces the per-plane cudaMemset of
 * es_fft.cpp:840 without touching the footprint twice. */
template<typename T>
__global__ void __launch_bounds__(256) es_zero_padding_kernel(EsGeom g,
        int nplanes, typename Prec<T>::T2* __restrict__ planes)
{
    typedef typename Prec<T>::T2 T2;
    const int gsz = g.grid_size, half = g.image_size / 2;
    const int row = blockIdx.x;
    const int col = blockIdx.y * blockDim.x + threadIdx.x;
    if (col >= gsz) return;
    const int a = row - gsz / 2, b = col - gsz / 2;
    const bool inside = (a >= -half && a < half && b >= -half && b < half);
    if (inside) return;
    T2 zero; zero.x = (T) 0; zero.y = (T) 0;
    for (int q = 0; q < nplanes; ++q)
        planes[(size_t) q * gsz * gsz + (size_t) row * gsz + col] = zero;
}

/* ---------------------------------------------------------- correction */

/* kernels.cu:690-769 (and conv_corr, :69-87; the quadrature is fp64 in both
 * precisions there too because PI is a double macro). */
template<typename T>
__global__ void __launch_bounds__(ES_SCREEN_BX * ES_SCREEN_BY)
es_correct_kernel(EsGeom g, T* __restrict__ dirty,
        const double* __restrict__ qk, const double* __restrict__ qn,
        const double* __restrict__ qw, const double* __restrict__ cc)
{
    const int i = blockIdx.x * ES_SCREEN_BX + threadIdx.x;
    const int j = blockIdx.y * ES_SCREEN_BY + threadIdx.y;
    const int n = g.image_size, half = n / 2;
    if (i > half || j > half) return;
    const double l = g.pixel_size * (double) i, m = g.pixel_size * (double) j;
    const double norm2 = g.conv_corr_norm_factor * g.conv_corr_norm_factor;
    double correction = cc[i] * cc[j];
    if (g.do_wstacking)
    {
        const double nn = sqrt(1.0 - l * l - m * m) - 1.0;
        const double k = nn * g.inv_w_scale;
        const double support = (double) g.support;
        const int p = (int) ceil(1.5 * support + 2.0);
        double sum = 0.0;
        for (int q = 0; q < p; ++q)
            sum += qk[q] * cos(ES_PI * k * support * qn[q]) * qw[q];
        correction *= (sum * support) * norm2;
    }
    else
    {
        correction *= norm2;
    }
    const T scale = (T) (1.0 / correction);
    T* img = dirty + (size_t) half * n + half;
    const ptrdiff_t ij = (ptrdiff_t) j * n;
    if (i < half && j < half) img[ij + i] *= scale;
    if (i > 0 && j < half) img[ij - i] *= scale;
    if (j > 0 && i < half) img[-ij + i] *= scale;
    if (i > 0 && j > 0) img[-ij - i] *= scale;
}

inline unsigned int blocks_for(int64_t n, int threads)
{
    return (unsigned int) ((n + threads - 1) / threads);
}

template<typename T, int NS>
static int launch_gather2_ts(const EsGeom* g, const void* records,
        const void* units, uint32_t* counters, int first_plane, int nplanes,
        const void* planes, void* acc, cudaStream_t s)
{
    typedef typename Prec<T>::T2 T2;
    typedef typename Prec<T>::Rec Rec;
    /* persistent grid, per device */
    static int grid_sizes[ES_MAX_DEVICES];
    int dev = 0;
    cudaError_t err = cudaGetDevice(&dev);
    if (err != cudaSuccess) return (int) err;
    if (dev < 0 || dev >= ES_MAX_DEVICES) return (int) cudaErrorInvalidDevice;
    int grid_size = __atomic_load_n(&grid_sizes[dev], __ATOMIC_ACQUIRE);
    if (grid_size == 0)
    {
        int sms = 0, per_sm = 0;
        err = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount,
                dev);
        if (err == cudaSuccess) err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(
                &per_sm, es_gather2_kernel<T, NS>, ES_GATHER_THREADS, 0);
        if (err != cudaSuccess) return (int) err;
        if (per_sm < 1) per_sm = 1;
        grid_size = sms * per_sm;
        __atomic_store_n(&grid_sizes[dev], grid_size, __ATOMIC_RELEASE);
    }
    es_gather2_kernel<T, NS><<<grid_size, ES_GATHER_THREADS, 0, s>>>(*g,
            (const Rec*) records, (const EsUnit*) units, counters, first_plane,
            nplanes, (const T2*) planes, (T2*) acc);
    return (int) cudaGetLastError();
}

} // namespace

/* ================================================================ C ABI */

extern "C" {

size_t es_kweight_count(const EsGeom* g)
{
    return (size_t) (3 * (g->support + 1));
}

size_t es_record_bytes(int is_double)
{
    return is_double ? sizeof(EsRecD) : sizeof(EsRecF);
}

int es_launch_count_entries(const EsGeom* g, const void* uvw, const void* freq,
        uint8_t* counts, es_stream_t stream)
{
    const int64_t nvis = (int64_t) g->num_rows * g->num_chan;
    if (nvis == 0) return 0;
    cudaStream_t s = (cudaStream_t) stream;
    if (g->is_double)
        es_count_entries_kernel<double><<<blocks_for(nvis, 256), 256, 0, s>>>(
                *g, (const double*) uvw, (const double*) freq, counts);
    else
        es_count_entries_kernel<float><<<blocks_for(nvis, 256), 256, 0, s>>>(
                *g, (const float*) uvw, (const float*) freq, counts);
    return (int) cudaGetLastError();
}

int es_scan_counts(const uint8_t* counts, uint32_t* offsets, int64_t n,
        void* temp, size_t* temp_bytes, es_stream_t stream)
{
    cub::TransformInputIterator<uint32_t, CastU8, const uint8_t*> in(counts,
            CastU8());
    return (int) cub::DeviceScan::ExclusiveSum(temp, *temp_bytes, in, offsets,
            n, (cudaStream_t) stream);
}

int es_launch_emit_entries(const EsGeom* g, const void* uvw, const void* freq,
        const uint32_t* offsets, uint64_t* keys, uint32_t* vals,
        es_stream_t stream)
{
    const int64_t nvis = (int64_t) g->num_rows * g->num_chan;
    if (nvis == 0) return 0;
    cudaStream_t s = (cudaStream_t) stream;
    if (g->is_double)
        es_emit_entries_kernel<double><<<blocks_for(nvis, 256), 256, 0, s>>>(
                *g, (const double*) uvw, (const double*) freq, offsets, keys,
                vals);
    else
        es_emit_entries_kernel<float><<<blocks_for(nvis, 256), 256, 0, s>>>(
                *g, (const float*) uvw, (const float*) freq, offsets, keys,
                vals);
    return (int) cudaGetLastError();
}

int es_sort_entries(const uint64_t* keys_in, uint64_t* keys_out,
        const uint32_t* vals_in, uint32_t* vals_out, int64_t num_entries,
        int end_bit, void* temp, size_t* temp_bytes, es_stream_t stream)
{
    return (int) cub::DeviceRadixSort::SortPairs(temp, *temp_bytes, keys_in,
            keys_out, vals_in, vals_out, num_entries, 0, end_bit,
            (cudaStream_t) stream);
}

int es_launch_tile_bounds(const EsGeom* g, const uint64_t* keys,
        int64_t num_entries, uint32_t* tile_start, es_stream_t stream)
{
    cudaStream_t s = (cudaStream_t) stream;
    const size_t ntiles = (size_t) g->ntile * g->ntile;
    if (num_entries == 0)
        return (int) cudaMemsetAsync(tile_start, 0,
                (ntiles + 1) * sizeof(uint32_t), s);
    es_tile_bounds_kernel<<<blocks_for(num_entries, 256), 256, 0, s>>>(*g,
            keys, num_entries, tile_start);
    return (int) cudaGetLastError();
}

int es_launch_build_records(const EsGeom* g, const void* uvw, const void* freq,
        const void* vis, const void* weight, const uint64_t* keys,
        const uint32_t* vals, int64_t num_entries, void* records,
        const uint32_t* offsets, uint32_t* entry_pos, void* kweights,
        es_stream_t stream)
{
    if (num_entries == 0) return 0;
    cudaStream_t s = (cudaStream_t) stream;
    if (g->is_double)
        es_build_records_kernel<double><<<blocks_for(num_entries, 256), 256, 0,
                s>>>(*g, (const double*) uvw, (const double*) freq,
                (const double2*) vis, (const double*) weight, keys, vals,
                num_entries, (EsRecD*) records, offsets, entry_pos,
                (double*) kweights);
    else
        es_build_records_kernel<float><<<blocks_for(num_entries, 256), 256, 0,
                s>>>(*g, (const float*) uvw, (const float*) freq,
                (const float2*) vis, (const float*) weight, keys, vals,
                num_entries, (EsRecF*) records, offsets, entry_pos,
                (float*) kweights);
    return (int) cudaGetLastError();
}

int es_launch_apply_screen(const EsGeom* g, const void* planes,
        int first_plane, int nplanes, void* dirty, es_stream_t stream)
{
    cudaStream_t s = (cudaStream_t) stream;
    const int half = g->image_size / 2;
    dim3 block(ES_SCREEN_BX, ES_SCREEN_BY);
    dim3 grid(blocks_for(half + 1, ES_SCREEN_BX), blocks_for(half + 1, ES_SCREEN_BY));
    if (g->is_double)
        es_apply_screen_kernel<double><<<grid, block, 0, s>>>(*g,
                (const double2*) planes, first_plane, nplanes, (double*) dirty);
    else
        es_apply_screen_kernel<float><<<grid, block, 0, s>>>(*g,
                (const float2*) planes, first_plane, nplanes, (float*) dirty);
    return (int) cudaGetLastError();
}

int es_launch_reverse_screen(const EsGeom* g, const void* dirty,
        int first_plane, int nplanes, void* planes, es_stream_t stream)
{
    cudaStream_t s = (cudaStream_t) stream;
    const int half = g->image_size / 2;
    dim3 zgrid((unsigned) g->grid_size, blocks_for(g->grid_size, 256));
    dim3 block(ES_SCREEN_BX, ES_SCREEN_BY);
    dim3 grid(blocks_for(half + 1, ES_SCREEN_BX), blocks_for(half + 1, ES_SCREEN_BY));
    if (g->is_double)
    {
        es_zero_padding_kernel<double><<<zgrid, 256, 0, s>>>(*g, nplanes,
                (double2*) planes);
        es_reverse_screen_kernel<double><<<grid, block, 0, s>>>(*g,
                (const double*) dirty, first_plane, nplanes, (double2*) planes);
    }
    else
    {
        es_zero_padding_kernel<float><<<zgrid, 256, 0, s>>>(*g, nplanes,
                (float2*) planes);
        es_reverse_screen_kernel<float><<<grid, block, 0, s>>>(*g,
                (const float*) dirty, first_plane, nplanes, (float2*) planes);
    }
    return (int) cudaGetLastError();
}

int es_launch_gather2(const EsGeom* g, const void* records, const void* units,
        uint32_t* counters, int first_plane, int nplanes, const void* planes,
        void* acc, es_stream_t stream)
{
    cudaStream_t s = (cudaStream_t) stream;
    /* footprint slots per axis the kernel is unrolled for: 9, 13 or 17 */
    const int ns = g->support + 1;
    if (g->is_double)
    {
        if (ns <= 9) return launch_gather2_ts<double, 9>(g, records, units,
                counters, first_plane, nplanes, planes, acc, s);
        if (ns <= 13) return launch_gather2_ts<double, 13>(g, records, units,
                counters, first_plane, nplanes, planes, acc, s);
        return launch_gather2_ts<double, 17>(g, records, units, counters,
                first_plane, nplanes, planes, acc, s);
    }
    if (ns <= 9) return launch_gather2_ts<float, 9>(g, records, units,
            counters, first_plane, nplanes, planes, acc, s);
    if (ns <= 13) return launch_gather2_ts<float, 13>(g, records, units,
            counters, first_plane, nplanes, planes, acc, s);
    return launch_gather2_ts<float, 17>(g, records, units, counters,
            first_plane, nplanes, planes, acc, s);
}

int es_launch_finish_degrid(const EsGeom* g, const void* records,
        const void* acc, const uint32_t* offsets, const uint32_t* entry_pos,
        int64_t num_entries, void* vis, es_stream_t stream)
{
    const int64_t nvis = (int64_t) g->num_rows * g->num_chan;
    if (num_entries == 0 || nvis == 0) return 0;
    cudaStream_t s = (cudaStream_t) stream;
    if (g->is_double)
        es_finish_degrid_kernel<double><<<blocks_for(nvis, 256), 256, 0,
                s>>>((const EsRecD*) records, (const double2*) acc, offsets,
                entry_pos, nvis, (double2*) vis);
    else
        es_finish_degrid_kernel<float><<<blocks_for(nvis, 256), 256, 0,
                s>>>((const EsRecF*) records, (const float2*) acc, offsets,
                entry_pos, nvis, (float2*) vis);
    return (int) cudaGetLastError();
}

int es_launch_correct(const EsGeom* g, void* dirty, const double* quad_kernel,
        const double* quad_nodes, const double* quad_weights,
        const double* conv_corr, es_stream_t stream)
{
    cudaStream_t s = (cudaStream_t) stream;
    const int half = g->image_size / 2;
    dim3 block(ES_SCREEN_BX, ES_SCREEN_BY);
    dim3 grid(blocks_for(half + 1, ES_SCREEN_BX), blocks_for(half + 1, ES_SCREEN_BY));
    if (g->is_double)
        es_correct_kernel<double><<<grid, block, 0, s>>>(*g, (double*) dirty,
                quad_kernel, quad_nodes, quad_weights, conv_corr);
    else
        es_correct_kernel<float><<<grid, block, 0, s>>>(*g, (float*) dirty,
                quad_kernel, quad_nodes, quad_weights, conv_corr);
    return (int) cudaGetLastError();
}

int es_launch_plane_histogram(const EsGeom* g, const void* uvw,
        const void* freq, unsigned long long* hist, es_stream_t stream)
{
    const int64_t nvis = (int64_t) g->num_rows * g->num_chan;
    if (nvis == 0) return 0;
    cudaStream_t s = (cudaStream_t) stream;
    if (g->is_double)
        es_plane_histogram_kernel<double><<<blocks_for(nvis, 256), 256, 0, s>>>(
                *g, (const double*) uvw, (const double*) freq, hist);
    else
        es_plane_histogram_kernel<float><<<blocks_for(nvis, 256), 256, 0, s>>>(
                *g, (const float*) uvw, (const float*) freq, hist);
    return (int) cudaGetLastError();
}

} // extern "C"
