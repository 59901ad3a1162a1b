/*
 * Kernels and launch functions of the pruned w-plane FFT (see es_fft.cu for
 * the description): included by es_fft.cu (generic kernel set: any planned
 * stage list, radix dispatch at run time) and by es_fft_seq.cu (one kernel set
 * per fixed stage list, EsFftSeq in es_fft_core.h).
 */
#ifndef ES_FFT_KERNELS_CUH_
#define ES_FFT_KERNELS_CUH_

#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <cuda_runtime.h>

#include "es_device.h"
#include "es_fft_core.h"
#include "es_fft_plan.h"
#include "es_screen.h"

#define ES_FFT_COL_GROUP 16 /* degridding column pass: bundles per CTA group */
#define ES_FFT_MAX_SMEM (227 * 1024 - 576) /* static: 512 B of occupancy flags + the work item */

/* CTA shape of one kernel family. */
struct EsFftCfg
{
    int32_t threads, lshift, pitch;
    int32_t tw_global; /* twiddle tables stay in global memory (L1): the line fills shared memory */
    size_t smem_bytes;
};

struct EsFftOps;

struct EsFft
{
    EsFftLine line;
    const EsFftOps* ops; /* kernel set: generic, or one fixed stage list */
    int32_t seq;         /* EsFftSeq id of that set (0: generic) */
    void* d_tw;       /* T2[num_twiddles] */
    int32_t* d_bpos;  /* [n / last radix] */
    int32_t grid_size, image_size, is_double;
    EsFftCfg rows;    /* one line per CTA */
    EsFftCfg cols;    /* 1 << lshift adjacent columns per CTA */
    int64_t plane_elems; /* elements of the intermediate per plane */
    /* row passes: persistent CTAs over the occupied tile rows of a batch */
    uint32_t* d_row_list; /* (plane << 16 | tile row), ascending */
    uint32_t* d_row_work; /* [0] work items of the batch, [1] next item */
    int32_t row_list_cap; /* entries */
    int32_t rows_grid;    /* resident CTAs of the row kernels on this device */
    /* w-screen (es_fft_set_screen): n - 1 per pixel of one image quadrant */
    double* d_nm1;        /* [image_size / 2 + 1][nm1_pitch] */
    int32_t nm1_pitch;
    int32_t screen_shift; /* 32 - fixed-point bits of the phase (es_screen.h) */
    double screen_magic;
};

namespace {

template<typename T> struct Cx;
template<> struct Cx<float> { typedef float2 T2; };
template<> struct Cx<double> { typedef double2 T2; };

/* Copies the tables to shared memory unless they are used in place. */
template<typename T2>
__device__ __forceinline__ void load_twiddles(const T2* tw, const T2* tw_g,
        int count)
{
    if (tw == tw_g) return;
    T2* dst = const_cast<T2*>(tw);
    for (int i = threadIdx.x; i < count; i += blockDim.x) dst[i] = tw_g[i];
}

/* ---- w-screen phasor exp(sign 2 pi i w (n - 1)), n = sqrt(1 - l^2 - m^2)
 * (kernels.cu:110-123). n - 1 comes from the plan's fp64 table (one entry per
 * pixel of an image quadrant), the 1 / n factor and the (-1)^(i+j) shift are
 * applied once per call by es_transpose_screen_kernel. The phase w (n - 1)
 * reaches 1e2 .. 1e4 turns: fp32 data reduce it with one fp64 fused
 * multiply-add into fixed point (es_screen.h), fp64 data with rint. */
struct ScreenTab
{
    const double* nm1;
    int32_t pitch, shift;
    double magic;
};

template<typename T> struct Phasor;

template<> struct Phasor<float>
{
    __device__ __forceinline__ static void eval(const double nm1,
            const double w_signed, const ScreenTab& tb, float& re, float& im)
    {
        es_phasor_f32(nm1, w_signed, tb.magic, tb.shift, re, im);
    }
};

template<> struct Phasor<double>
{
    __device__ __forceinline__ static void eval(const double nm1,
            const double w_signed, const ScreenTab&, double& re, double& im)
    {
        double t = w_signed * nm1;
        t -= rint(t);
        sincospi(2.0 * t, &im, &re);
    }
};

/* ------------------------------------------------ global-memory functors
 * (cursor protocol of es_line_pass: element k of a butterfly is index
 * idx0 + k * gs of line `line`). All of them address linearly,
 *   base + (idx - k0) * estride + line * lstride,
 * so a cursor is one pointer and one stride; WINDOW: indices outside
 * [k0, k0 + count) read as zero / are not stored (the pruning). */

/* MASK: index idx (k0 == 0) belongs to 64-wide block idx >> 6; blocks whose
 * flag in `bits` (shared memory, one byte per block: occ_fetch expands the
 * occupancy words, so the test is a shift, a byte load and a compare) is
 * clear read as zero / are not stored. */
#define ES_FFT_OCC_WORDS 16 /* blocks <= 512: lines up to 32768 elements */
#define ES_FFT_OCC_BYTES (ES_FFT_OCC_WORDS * 32)
static_assert(ES_TILE == 64, "occupancy blocks are addressed with idx >> 6");

__device__ __forceinline__ bool occ_bit(const uint8_t* bits, int idx)
{
    return bits[idx >> 6] != 0;
}

/* Non-zero iff any bit of the value is set: one or two 3-input ORs instead of
 * two floating-point compares per element. */
__device__ __forceinline__ unsigned int nonzero_bits(const float2 v)
{
    return __float_as_uint(v.x) | __float_as_uint(v.y);
}

__device__ __forceinline__ unsigned int nonzero_bits(const double2 v)
{
    return (unsigned int) (__double2hiint(v.x) | __double2loint(v.x) |
            __double2hiint(v.y) | __double2loint(v.y));
}

template<typename T2, bool WINDOW, bool ANY, bool MASK = false> struct LinearIn
{
    struct Cursor
    {
        const T2* p;
        ptrdiff_t step;
        int r0, gs;
    };
    const T2* base;          /* element k0 of line 0 */
    ptrdiff_t estride, lstride;
    int k0, count;
    const uint8_t* bits;     /* MASK */
    unsigned int any;        /* ANY: becomes non-zero when a loaded value is */
    unsigned int bf;         /* ANY: the same for the current butterfly */
    /* ANY: butterflies of the first pass with all-zero inputs are skipped */
    static constexpr bool skip_zero_butterflies = ANY;
    __device__ __forceinline__ void butterfly_begin()
    {
        if (ANY) { any |= bf; bf = 0u; }
    }
    __device__ __forceinline__ bool butterfly_nonzero()
    {
        return bf != 0u;
    }
    __device__ __forceinline__ Cursor cursor(int line, int idx0, int gs) const
    {
        Cursor c;
        c.r0 = idx0 - k0;
        c.gs = gs;
        c.p = base + (ptrdiff_t) c.r0 * estride + (ptrdiff_t) line * lstride;
        c.step = (ptrdiff_t) gs * estride;
        return c;
    }
    typedef T2 Raw;
    __device__ __forceinline__ unsigned int any_bits() const { return any | bf; }
    __device__ __forceinline__ void merge_any(unsigned int a) { any |= a; }
    __device__ __forceinline__ T2 finish(const Cursor&, int, const T2 v) const
    {
        return v;
    }
    __device__ __forceinline__ void finish2(const Cursor&, int, const T2 r0,
            const T2 r1, T2& v0, T2& v1) const
    {
        v0 = r0;
        v1 = r1;
    }
    __device__ __forceinline__ T2 fetch(const Cursor& c, int k)
    {
        T2 v;
        v.x = 0; v.y = 0;
        if ((!WINDOW || (unsigned) (c.r0 + k * c.gs) < (unsigned) count) &&
                (!MASK || occ_bit(bits, c.r0 + k * c.gs)))
            v = c.p[k * c.step];
        if (ANY) bf |= nonzero_bits(v); /* -0.0 counts: it only costs an FFT */
        return v;
    }
};

template<typename T2, bool WINDOW, bool MASK = false, bool SKIP = false>
struct LinearOut
{
    struct Cursor
    {
        T2* p;
        ptrdiff_t step;
        int r0, gs;
    };
    T2* base;
    ptrdiff_t estride, lstride;
    int k0, count;
    const uint8_t* bits;     /* MASK */
    __device__ __forceinline__ Cursor cursor(int line, int idx0, int gs) const
    {
        Cursor c;
        c.r0 = idx0 - k0;
        c.gs = gs;
        c.p = base + (ptrdiff_t) c.r0 * estride + (ptrdiff_t) line * lstride;
        c.step = (ptrdiff_t) gs * estride;
        return c;
    }
    __device__ __forceinline__ void store(const Cursor& c, int k, const T2 v) const
    {
        if ((!WINDOW || (unsigned) (c.r0 + k * c.gs) < (unsigned) count) &&
                (!MASK || occ_bit(bits, c.r0 + k * c.gs)))
            c.p[k * c.step] = v;
    }
    /* SKIP (with MASK): butterflies of the last pass none of whose outputs
     * is stored are skipped (row pass onto a few occupied tiles) */
    static constexpr bool skip_masked_butterflies = MASK && SKIP;
    int sparse;              /* SKIP: worth testing (few occupied blocks) */
    template<int R>
    __device__ __forceinline__ bool any_stored(const Cursor& c) const
    {
        if (!sparse) return true;
        bool on = false;
#pragma unroll
        for (int k = 0; k < R; ++k)
            on = on || ((!WINDOW || (unsigned) (c.r0 + k * c.gs) <
                    (unsigned) count) && occ_bit(bits, c.r0 + k * c.gs));
        return on;
    }
    /* nothing to prefetch */
    struct Pre {};
    __device__ __forceinline__ Pre prefetch(const Cursor&, int) const
    {
        return Pre();
    }
    __device__ __forceinline__ void store1(const Cursor& c, int k, const T2 v,
            const Pre) const
    {
        store(c, k, v);
    }
    __device__ __forceinline__ void store2(const Cursor& c, int k, const T2 v0,
            const T2 v1, const Pre, const Pre) const
    {
        store(c, k, v0);
        store(c, k + 1, v1);
    }
};

/* gridding pass 2 output: w-screen phasor and sum into imageT
 * (kernels.cu:444-515; 1 / n and the shift sign: es_transpose_screen_kernel) */
template<typename T, typename T2> struct ScreenOut
{
    struct Cursor
    {
        T* p;              /* imageT element of (line, iu0) */
        const double* tab; /* n - 1 of (|jv|, 0) */
        int iu0, gs;
    };
    T* image_t;      /* imageT + first column of the bundle * n + n / 2 */
    int n, half, gh; /* image size, n / 2, grid_size / 2 */
    int jv0;         /* first column - n / 2 */
    int wstack;
    double w;        /* sign * w of the plane */
    ScreenTab tb;
    __device__ __forceinline__ Cursor cursor(int line, int idx0, int gs) const
    {
        Cursor c;
        c.iu0 = idx0 - gh;
        c.gs = gs;
        c.tab = tb.nm1 + (size_t) abs(jv0 + line) * tb.pitch;
        c.p = image_t + (ptrdiff_t) line * n + c.iu0;
        return c;
    }
    static constexpr bool skip_masked_butterflies = false;
    /* n - 1 of output k, loaded before the butterfly arithmetic */
    typedef double Pre;
    __device__ __forceinline__ Pre prefetch(const Cursor& c, int k) const
    {
        const int iu = c.iu0 + k * c.gs;
        double nm1 = 0.0;
        if (wstack && (unsigned) (iu + half) < (unsigned) n)
            nm1 = __ldg(c.tab + abs(iu));
        return nm1;
    }
    __device__ __forceinline__ void store1(const Cursor& c, int k, const T2 v,
            const Pre nm1) const
    {
        const int iu = c.iu0 + k * c.gs;
        if ((unsigned) (iu + half) >= (unsigned) n) return;
        T val = v.x;
        if (wstack)
        {
            T cr, ci;
            Phasor<T>::eval(nm1, w, tb, cr, ci);
            val = v.x * cr - v.y * ci;
        }
        atomicAdd(c.p + k * c.gs, val);
    }
    /* outputs k and k + 1: when both fall into the image (the common case)
     * the two phasors are evaluated side by side without a branch between */
    __device__ __forceinline__ void store2(const Cursor& c, int k, const T2 v0,
            const T2 v1, const Pre nm0, const Pre nm1) const
    {
        const int iu0 = c.iu0 + k * c.gs, iu1 = iu0 + c.gs;
        const bool in0 = (unsigned) (iu0 + half) < (unsigned) n;
        const bool in1 = (unsigned) (iu1 + half) < (unsigned) n;
        if (!(in0 && in1) || !wstack)
        {
            if (in0) store1(c, k, v0, nm0);
            if (in1) store1(c, k + 1, v1, nm1);
            return;
        }
        T cr0, ci0, cr1, ci1;
        Phasor<T>::eval(nm0, w, tb, cr0, ci0);
        Phasor<T>::eval(nm1, w, tb, cr1, ci1);
        atomicAdd(c.p + k * c.gs, v0.x * cr0 - v0.y * ci0);
        atomicAdd(c.p + (k + 1) * c.gs, v1.x * cr1 - v1.y * ci1);
    }
};

/* degridding pass 1 input: imageT (already x shift sign / n) x conj w-screen
 * phasor, zero outside the image (kernels.cu:580-676) */
template<typename T, typename T2> struct ScreenIn
{
    struct Cursor
    {
        const T* p;
        const double* tab;
        int iu0, gs;
    };
    const T* image_t;
    int n, half, gh, jv0, wstack;
    double w;
    ScreenTab tb;
    __device__ __forceinline__ Cursor cursor(int line, int idx0, int gs) const
    {
        Cursor c;
        c.iu0 = idx0 - gh;
        c.gs = gs;
        c.tab = tb.nm1 + (size_t) abs(jv0 + line) * tb.pitch;
        c.p = image_t + (ptrdiff_t) line * n + c.iu0;
        return c;
    }
    struct Raw
    {
        T d;
        double nm1;
    };
    static constexpr bool skip_zero_butterflies = false;
    __device__ __forceinline__ void butterfly_begin() const {}
    __device__ __forceinline__ bool butterfly_nonzero() const { return true; }
    __device__ __forceinline__ unsigned int any_bits() const { return 0u; }
    __device__ __forceinline__ void merge_any(unsigned int) {}
    /* the image value (0 outside the image) and n - 1; no arithmetic yet */
    __device__ __forceinline__ Raw fetch(const Cursor& c, int k) const
    {
        const int iu = c.iu0 + k * c.gs;
        Raw r;
        r.d = 0;
        r.nm1 = 0.0;
        if ((unsigned) (iu + half) < (unsigned) n)
        {
            r.d = c.p[k * c.gs];
            if (wstack) r.nm1 = __ldg(c.tab + abs(iu));
        }
        return r;
    }
    __device__ __forceinline__ T2 finish(const Cursor& c, int k, const Raw r) const
    {
        T2 o;
        o.x = r.d; o.y = 0;
        const int iu = c.iu0 + k * c.gs;
        if (wstack && (unsigned) (iu + half) < (unsigned) n)
        {
            T cr, ci;
            Phasor<T>::eval(r.nm1, w, tb, cr, ci);
            o.x = cr * r.d;
            o.y = ci * r.d;
        }
        return o;
    }
    /* elements k and k + 1: the two phasors are evaluated side by side
     * without a branch between (outside the image d == 0 and n - 1 == 0:
     * the products are zeros) */
    __device__ __forceinline__ void finish2(const Cursor& c, int k, const Raw r0,
            const Raw r1, T2& o0, T2& o1) const
    {
        if (!wstack)
        {
            o0.x = r0.d; o0.y = 0;
            o1.x = r1.d; o1.y = 0;
            return;
        }
        T cr0, ci0, cr1, ci1;
        Phasor<T>::eval(r0.nm1, w, tb, cr0, ci0);
        Phasor<T>::eval(r1.nm1, w, tb, cr1, ci1);
        o0.x = cr0 * r0.d; o0.y = ci0 * r0.d;
        o1.x = cr1 * r1.d; o1.y = ci1 * r1.d;
    }
};

/* All passes of the lines held by this CTA; `after_first` runs once the
 * first pass is complete (all global loads done) and may end the CTA early by
 * returning false. */
template<int SEQ, int SGN, bool DIT, typename T, typename T2, typename In,
        typename Out>
__device__ __forceinline__ void fft_all_passes(T2* x, const EsFftLine& pl,
        const T2* tw, const int32_t* __restrict__ bpos, const int lshift,
        const int pitch, In& in, Out& out)
{
    for (int step = 0; step < pl.num_stages; ++step)
    {
        if (step > 0) __syncthreads();
        es_line_pass_seq<SEQ, SGN, DIT, T>(x, pl, step, tw, bpos, lshift, pitch,
                (int) threadIdx.x, (int) blockDim.x, in, out);
    }
}

/* CTA shapes and size of the intermediate, by value to the kernels. */
struct EsFftLayout
{
    int32_t lshift;    /* log2 of the columns per CTA of the column kernels */
    int32_t rshift;    /* log2 of the rows per CTA of the row kernels */
    int32_t rows_pitch, cols_pitch;
    int64_t plane_elems; /* elements of T per plane (grid_size x image_size) */
    int32_t ntile, occ_nw; /* occupancy words: es_device.h */
    int32_t rows_twg, cols_twg; /* twiddles read from global memory */
    ScreenTab screen;  /* w-screen table of the plan (es_fft_set_screen) */
};

/* Words of plane q: tile bits of tile row tu at + tu * nw, the tile-row bits
 * at + ntile * nw. */
__device__ __forceinline__ const uint32_t* occ_plane(const uint32_t* occ,
        const EsFftLayout& lay, int q)
{
    return occ + (size_t) q * (size_t) ((lay.ntile + 1) * lay.occ_nw);
}

/* Expands the nw occupancy words at src into one byte per block in shared
 * memory; true when any block is occupied (contains a barrier: every thread
 * of the CTA must call it). */
__device__ __forceinline__ bool occ_fetch(uint8_t* s_bits,
        const uint32_t* src, int nw)
{
    uint32_t any = 0;
    for (int t = threadIdx.x; t < nw * 32; t += blockDim.x)
    {
        const uint32_t b = (src[t >> 5] >> (t & 31)) & 1u;
        s_bits[t] = (uint8_t) b;
        any |= b;
    }
    return __syncthreads_or((int) any) != 0;
}

/* ----------------------------------------------- occupied tile-row list */

/* One CTA: the occupied (plane, tile row) pairs of the batch in ascending
 * order -> list, work[0] = pairs x row bundles per tile row, work[1] = 0.
 * The row kernels then run as persistent CTAs over exactly the rows that
 * hold data: on sparse coverage (SKA-Low: a quarter of the tile rows) most
 * CTAs of a (rows x planes) launch would only find out that they have
 * nothing to do, one shared-memory-filling CTA per SM at a time. */
__global__ void __launch_bounds__(1024) es_fft_row_list_kernel(
        const uint32_t* __restrict__ occ, const EsFftLayout lay,
        const int nplanes, uint32_t* __restrict__ list,
        uint32_t* __restrict__ work)
{
    __shared__ uint32_t s_base, s_warp[32];
    const int total = nplanes * lay.ntile;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) s_base = 0u;
    __syncthreads();
    for (int i0 = 0; i0 < total; i0 += (int) blockDim.x)
    {
        const int i = i0 + (int) threadIdx.x;
        bool on = false;
        int q = 0, tu = 0;
        if (i < total)
        {
            q = i / lay.ntile;
            tu = i - q * lay.ntile;
            const uint32_t* o = occ + (size_t) q *
                    (size_t) ((lay.ntile + 1) * lay.occ_nw) +
                    lay.ntile * lay.occ_nw;
            on = (o[tu >> 5] >> (tu & 31)) & 1u;
        }
        const uint32_t m = __ballot_sync(0xffffffffu, on);
        if (lane == 0) s_warp[warp] = (uint32_t) __popc(m);
        __syncthreads();
        uint32_t before = s_base;
        for (int w = 0; w < warp; ++w) before += s_warp[w];
        if (on)
            list[before + (uint32_t) __popc(m & ((1u << lane) - 1u))] =
                    ((uint32_t) q << 16) | (uint32_t) tu;
        __syncthreads();
        if (threadIdx.x == 0)
        {
            uint32_t sum = 0u;
            for (int w = 0; w < (int) (blockDim.x >> 5); ++w) sum += s_warp[w];
            s_base += sum;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0)
    {
        work[0] = s_base * (uint32_t) (ES_TILE >> lay.rshift);
        work[1] = 0u;
    }
}

/* Next work item of a persistent row CTA (contains barriers: the previous
 * item's shared-memory traffic is complete when it returns). Item ->
 * first row u of the bundle and plane q; false when the list is exhausted. */
__device__ __forceinline__ bool row_item(uint32_t* s_item,
        const uint32_t* __restrict__ list, uint32_t* __restrict__ work,
        const EsFftLayout& lay, int& q, int& u)
{
    __syncthreads();
    if (threadIdx.x == 0) *s_item = atomicAdd(&work[1], 1u);
    __syncthreads();
    const uint32_t item = *s_item;
    if (item >= work[0]) return false;
    const int per_shift = 6 - lay.rshift; /* row bundles per tile row */
    const uint32_t e = list[item >> per_shift];
    q = (int) (e >> 16);
    u = (int) ((e & 0xffffu) << 6) +
            (int) ((item & ((1u << per_shift) - 1u)) << lay.rshift);
    return true;
}

/* ------------------------------------------------------ gridding pass 1 */

template<typename T, int SEQ>
__global__ void __launch_bounds__(EsFftSeq<SEQ>::threads, 1) es_fft_grid_rows_kernel(
        const EsFftLine pl, const typename Cx<T>::T2* __restrict__ tw_g,
        const int32_t* __restrict__ bpos,
        const typename Cx<T>::T2* __restrict__ planes,
        typename Cx<T>::T2* __restrict__ strip, const int image_size,
        const uint32_t* __restrict__ occ, const EsFftLayout lay,
        const uint32_t* __restrict__ row_list, uint32_t* __restrict__ row_work)
{
    typedef typename Cx<T>::T2 T2;
    extern __shared__ __align__(16) unsigned char es_fft_smem[];
    __shared__ uint8_t s_bits[ES_FFT_OCC_BYTES];
    __shared__ uint32_t s_item;
    T2* x = (T2*) es_fft_smem;
    const T2* tw = lay.rows_twg ? tw_g :
            x + ((size_t) lay.rows_pitch << lay.rshift);
    const int gsz = pl.n, n = image_size;
    load_twiddles(tw, tw_g, pl.num_twiddles);
    /* persistent CTA: row bundles of tile rows that hold data, nothing else
     * (rows of the other tile rows are read as zeros by the column pass
     * without ever being stored) */
    int q, u;
    while (row_item(&s_item, row_list, row_work, lay, q, u))
    {
        if (u >= gsz) continue; /* the last tile row may be partial */
        occ_fetch(s_bits, occ_plane(occ, lay, q) + (u >> 6) * lay.occ_nw,
                lay.occ_nw);
        LinearIn<T2, false, true, true> in;
        in.bits = s_bits;
        in.base = planes + ((size_t) q * gsz + u) * gsz;
        in.estride = 1;
        in.lstride = gsz;
        in.k0 = 0;
        in.count = gsz;
        in.any = 0u;
        in.bf = 0u;
        LinearOut<T2, true> out;
        out.base = strip + (size_t) q * lay.plane_elems + (size_t) u * n;
        out.estride = 1;
        out.lstride = n;
        out.k0 = gsz / 2 - n / 2;
        out.count = n;
        for (int step = 0; step < pl.num_stages; ++step)
        {
            es_line_pass_seq<SEQ, 1, false, T>(x, pl,
                    step, tw, bpos, lay.rshift, lay.rows_pitch,
                    (int) threadIdx.x, (int) blockDim.x, in, out);
            if (step + 1 == pl.num_stages) break;
            if (step == 0)
            {
                /* rows nothing was gridded onto transform to zeros */
                if (!__syncthreads_or((int) (in.any != 0u)))
                {
                    T2 zero;
                    zero.x = 0; zero.y = 0;
                    for (int i = threadIdx.x; i < (n << lay.rshift);
                            i += blockDim.x)
                        out.base[i] = zero; /* the rows of a bundle are adjacent */
                    break;
                }
            }
            else
            {
                __syncthreads();
            }
        }
    }
}

/* ------------------------------------------------------ gridding pass 2 */

/* One CTA per bundle of adjacent columns v': the planes of the batch are
 * transformed one after the other and summed into the imageT slice of the
 * bundle (which stays in L2 meanwhile). A pixel always receives its
 * contributions from the same thread in plane order, so the sum does not
 * depend on scheduling (bitwise reproducible). */
template<typename T, int SEQ>
__global__ void __launch_bounds__(EsFftSeq<SEQ>::threads, 1)
es_fft_grid_cols_screen_kernel(const EsFftLine pl,
        const typename Cx<T>::T2* __restrict__ tw_g,
        const int32_t* __restrict__ bpos, const EsGeom g,
        const typename Cx<T>::T2* __restrict__ strip, const int first_plane,
        const int nplanes, T* __restrict__ image_t,
        const uint32_t* __restrict__ occ, const EsFftLayout lay)
{
    typedef typename Cx<T>::T2 T2;
    extern __shared__ __align__(16) unsigned char es_fft_smem[];
    __shared__ uint8_t s_bits[ES_FFT_OCC_BYTES];
    T2* x = (T2*) es_fft_smem;
    const T2* tw = lay.cols_twg ? tw_g :
            x + ((size_t) lay.cols_pitch << lay.lshift);
    const int gsz = pl.n, n = g.image_size;
    const int col0 = blockIdx.x << lay.lshift;
    LinearIn<T2, false, false, true> in;
    in.bits = s_bits;
    in.estride = n;
    in.lstride = 1;
    in.k0 = 0;
    in.count = gsz;
    in.any = 0u;
    in.bf = 0u;
    ScreenOut<T, T2> out;
    out.image_t = image_t + (size_t) col0 * n + n / 2;
    out.n = n;
    out.half = n / 2;
    out.gh = gsz / 2;
    out.jv0 = col0 - n / 2;
    out.wstack = g.do_wstacking;
    out.tb = lay.screen;
    load_twiddles(tw, tw_g, pl.num_twiddles);
    for (int q = 0; q < nplanes; ++q)
    {
        __syncthreads(); /* twiddles loaded / previous plane's last pass done */
        /* rows of unoccupied tile rows were not stored; an empty plane adds
         * nothing */
        if (!occ_fetch(s_bits, occ_plane(occ, lay, q) +
                lay.ntile * lay.occ_nw, lay.occ_nw)) continue;
        in.base = strip + (size_t) q * lay.plane_elems + col0;
        /* gridding: exp(-2 pi i w (n - 1)) */
        out.w = -((double) (first_plane + q) * g.inv_w_scale + g.min_plane_w);
        for (int step = 0; step < pl.num_stages; ++step)
        {
            if (step > 0) __syncthreads();
            es_line_pass_seq<SEQ, 1, false, T>(x, pl, step, tw, bpos, lay.lshift,
                    lay.cols_pitch, (int) threadIdx.x, (int) blockDim.x, in,
                    out);
        }
    }
}

/* ---------------------------------------------------- degridding pass 1 */

template<typename T, int SEQ>
__global__ void __launch_bounds__(EsFftSeq<SEQ>::threads, 1) es_fft_degrid_cols_kernel(
        const EsFftLine pl, const typename Cx<T>::T2* __restrict__ tw_g,
        const int32_t* __restrict__ bpos, const EsGeom g,
        const T* __restrict__ image_t, const int first_plane,
        const int nplanes, typename Cx<T>::T2* __restrict__ strip,
        const uint32_t* __restrict__ occ, const EsFftLayout lay)
{
    typedef typename Cx<T>::T2 T2;
    extern __shared__ __align__(16) unsigned char es_fft_smem[];
    __shared__ uint8_t s_bits[ES_FFT_OCC_BYTES];
    T2* x = (T2*) es_fft_smem;
    const T2* tw = lay.cols_twg ? tw_g :
            x + ((size_t) lay.cols_pitch << lay.lshift);
    const int gsz = pl.n, n = g.image_size;
    /* CTA order: groups of ES_FFT_COL_GROUP adjacent column bundles, all
     * planes of a group before the next group. Adjacent bundles of one plane
     * run together (their 16 / 32-byte pieces complete whole lines in L2
     * before eviction) and the group's image columns stay in L2 across the
     * planes of the batch. */
    const int nbundles = n >> lay.lshift;
    const int grp = blockIdx.x / (ES_FFT_COL_GROUP * nplanes);
    const int rem = blockIdx.x - grp * (ES_FFT_COL_GROUP * nplanes);
    const int q = rem / ES_FFT_COL_GROUP;
    const int bundle = grp * ES_FFT_COL_GROUP + (rem - q * ES_FFT_COL_GROUP);
    if (bundle >= nbundles) return;
    const int col0 = bundle << lay.lshift;
    /* only the tile rows the gather reads are stored; none: nothing to do */
    if (!occ_fetch(s_bits, occ_plane(occ, lay, q) + lay.ntile * lay.occ_nw,
            lay.occ_nw)) return;
    ScreenIn<T, T2> in;
    in.image_t = image_t + (size_t) col0 * n + n / 2;
    in.n = n;
    in.half = n / 2;
    in.gh = gsz / 2;
    in.jv0 = col0 - n / 2;
    in.wstack = g.do_wstacking;
    in.tb = lay.screen;
    in.w = (double) (first_plane + q) * g.inv_w_scale + g.min_plane_w;
    LinearOut<T2, false, true> out;
    out.bits = s_bits;
    out.base = strip + (size_t) q * lay.plane_elems + col0;
    out.estride = n;
    out.lstride = 1;
    out.k0 = 0;
    out.count = gsz;
    load_twiddles(tw, tw_g, pl.num_twiddles);
    __syncthreads();
    fft_all_passes<SEQ, -1, true, T>(x, pl, tw, bpos,
            lay.lshift, lay.cols_pitch, in, out);
}

/* ---------------------------------------------------- degridding pass 2 */

template<typename T, int SEQ>
__global__ void __launch_bounds__(EsFftSeq<SEQ>::threads, 1) es_fft_degrid_rows_kernel(
        const EsFftLine pl, const typename Cx<T>::T2* __restrict__ tw_g,
        const int32_t* __restrict__ bpos,
        const typename Cx<T>::T2* __restrict__ strip,
        typename Cx<T>::T2* __restrict__ planes, const int image_size,
        const uint32_t* __restrict__ occ, const EsFftLayout lay,
        const uint32_t* __restrict__ row_list, uint32_t* __restrict__ row_work)
{
    typedef typename Cx<T>::T2 T2;
    extern __shared__ __align__(16) unsigned char es_fft_smem[];
    __shared__ uint8_t s_bits[ES_FFT_OCC_BYTES];
    __shared__ uint32_t s_item;
    T2* x = (T2*) es_fft_smem;
    const T2* tw = lay.rows_twg ? tw_g :
            x + ((size_t) lay.rows_pitch << lay.rshift);
    const int gsz = pl.n, n = image_size;
    load_twiddles(tw, tw_g, pl.num_twiddles);
    /* persistent CTA over the tile rows the gather reads; only their
     * occupied tiles are stored */
    int q, u;
    while (row_item(&s_item, row_list, row_work, lay, q, u))
    {
        if (u >= gsz) continue;
        occ_fetch(s_bits, occ_plane(occ, lay, q) + (u >> 6) * lay.occ_nw,
                lay.occ_nw);
        /* occupied tiles of this tile row */
        int occupied = 0;
        for (int t0 = 0; t0 < lay.ntile; t0 += (int) blockDim.x)
        {
            const int t = t0 + (int) threadIdx.x;
            occupied += __syncthreads_count(t < lay.ntile && s_bits[t] != 0);
        }
        LinearIn<T2, true, false> in;
        in.base = strip + (size_t) q * lay.plane_elems + (size_t) u * n;
        in.estride = 1;
        in.lstride = n;
        in.k0 = gsz / 2 - n / 2;
        in.count = n;
        in.any = 0u;
        in.bf = 0u;
        LinearOut<T2, false, true, true> out;
        /* a butterfly of the last pass (stage 0, radix r) spreads its r
         * outputs over the whole line: it is skipped with probability
         * ~(1 - f)^r, f = occupied fraction; worth the r tests when f r is
         * small (C4: 0.07, C5: 1.8, C2: 4) */
        out.sparse = (occupied * ES_TILE * pl.stage[0].radix * 2 < gsz);
        out.bits = s_bits;
        out.base = planes + ((size_t) q * gsz + u) * gsz;
        out.estride = 1;
        out.lstride = gsz;
        out.k0 = 0;
        out.count = gsz;
        fft_all_passes<SEQ, -1, true, T>(x, pl, tw,
                bpos, lay.rshift, lay.rows_pitch, in, out);
    }
}

/* ------------------------------------------------------ image transposes */

/* dst[r][c] (+)= src[c][r] x (-1)^(r + c) / n(r, c), square n x n: the
 * transposes between the caller's image and the column-major accumulator /
 * model copy the line FFTs work on carry the two per-pixel factors of the
 * w-screen that do not depend on the plane (kernels.cu:119,499-507: xn and
 * the shift sign). tab == 0 (no w-stacking): the sign only. */
template<typename T, bool ADD>
__global__ void __launch_bounds__(256) es_transpose_screen_kernel(
        const T* __restrict__ src, T* __restrict__ dst, const int n,
        const double* __restrict__ tab, const int pitch)
{
    __shared__ T t[32][33];
    const int bx = blockIdx.x * 32, by = blockIdx.y * 32;
    for (int k = threadIdx.y; k < 32; k += 8)
    {
        const int r = by + k, c = bx + threadIdx.x;
        if (r < n && c < n) t[k][threadIdx.x] = src[(size_t) r * n + c];
    }
    __syncthreads();
    const int half = n / 2;
    for (int k = threadIdx.y; k < 32; k += 8)
    {
        const int r = bx + k, c = by + threadIdx.x;
        if (r < n && c < n)
        {
            T v = t[threadIdx.x][k];
            if (tab)
                v = (T) ((double) v / (tab[(size_t) abs(r - half) * pitch +
                        abs(c - half)] + 1.0));
            if ((r + c) & 1) v = -v;
            if (ADD) dst[(size_t) r * n + c] += v;
            else dst[(size_t) r * n + c] = v;
        }
    }
}

/* tab[j][i] = n - 1 at l = i pixel_size, m = j pixel_size (fp64). */
__global__ void __launch_bounds__(256) es_screen_table_kernel(
        double* __restrict__ tab, const int count, const int pitch,
        const double pix2)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y;
    if (i < count)
        tab[(size_t) j * pitch + i] = es_screen_nm1((uint32_t) (i * i + j * j),
                pix2);
}

} // namespace


namespace {

EsFftLayout fft_layout(const EsFft* f)
{
    EsFftLayout lay;
    lay.lshift = f->cols.lshift;
    lay.rshift = f->rows.lshift;
    lay.rows_twg = f->rows.tw_global;
    lay.cols_twg = f->cols.tw_global;
    lay.rows_pitch = f->rows.pitch;
    lay.cols_pitch = f->cols.pitch;
    lay.plane_elems = f->plane_elems;
    lay.ntile = (f->grid_size + ES_TILE - 1) / ES_TILE;
    lay.occ_nw = (lay.ntile + 31) >> 5;
    lay.screen.nm1 = f->d_nm1;
    lay.screen.pitch = f->nm1_pitch;
    lay.screen.shift = f->screen_shift;
    lay.screen.magic = f->screen_magic;
    return lay;
}

template<typename K>
int fft_set_smem(K kernel, size_t bytes)
{
    /* the opt-in limit of the device covers static + dynamic shared memory */
    cudaFuncAttributes attr;
    int dev = 0, optin = 0;
    if (cudaFuncGetAttributes(&attr, (const void*) kernel) == cudaSuccess &&
            cudaGetDevice(&dev) == cudaSuccess &&
            cudaDeviceGetAttribute(&optin,
            cudaDevAttrMaxSharedMemoryPerBlockOptin, dev) == cudaSuccess &&
            bytes + attr.sharedSizeBytes > (size_t) optin)
        bytes = (size_t) optin - attr.sharedSizeBytes;
    return (int) cudaFuncSetAttribute((const void*) kernel,
            cudaFuncAttributeMaxDynamicSharedMemorySize, (int) bytes);
}


/* (plane, tile row) list of a batch -> fft->d_row_list / d_row_work. */
int fft_row_list(EsFft* f, const uint32_t* occ, int nplanes,
        const EsFftLayout& lay, cudaStream_t s)
{
    const int need = nplanes * lay.ntile;
    if (need > f->row_list_cap)
    {
        /* grows only while the first batches of a plan are enqueued; the
         * stream is drained because kernels may still read the old list */
        cudaError_t err = cudaStreamSynchronize(s);
        if (err != cudaSuccess) return (int) err;
        if (f->d_row_list) (void) cudaFree(f->d_row_list);
        f->d_row_list = 0;
        f->row_list_cap = 0;
        err = cudaMalloc((void**) &f->d_row_list, sizeof(uint32_t) * (size_t) need);
        if (err != cudaSuccess) return (int) err;
        f->row_list_cap = need;
    }
    es_fft_row_list_kernel<<<1, 1024, 0, s>>>(occ, lay, nplanes, f->d_row_list,
            f->d_row_work);
    return (int) cudaGetLastError();
}

} // namespace

/* ------------------------------------------------------ launch functions
 * One set per (precision, stage list); es_fft.cu calls them through the
 * plan's EsFftOps table. */
struct EsFftOps
{
    int max_threads; /* CTA size limit of the set's kernels */
    int (*prepare)(size_t max_smem);
    int (*rows_per_sm)(int threads, size_t smem_bytes, int* per_sm);
    int (*grid_rows)(EsFft* f, const void* planes, int nplanes, void* strip,
            const uint32_t* occ, cudaStream_t s);
    int (*grid_cols_screen)(const EsFft* f, const EsGeom* g, const void* strip,
            int first_plane, int nplanes, void* image_t, const uint32_t* occ,
            cudaStream_t s);
    int (*degrid_cols)(const EsFft* f, const EsGeom* g, const void* image_t,
            int first_plane, int nplanes, void* strip, const uint32_t* occ,
            cudaStream_t s);
    int (*degrid_rows)(EsFft* f, const void* strip, int nplanes, void* planes,
            const uint32_t* occ, cudaStream_t s);
};

namespace {

template<typename T, int SEQ> struct EsFftLaunch
{
    typedef typename Cx<T>::T2 T2;

    /* The dynamic shared-memory limit belongs to the function and the
     * device, not to a plan: always opt in to the maximum any plan can ask
     * for, so that plans of different sizes can be alive together. */
    static int prepare(size_t max_smem)
    {
        int err = fft_set_smem(es_fft_grid_rows_kernel<T, SEQ>, max_smem);
        if (!err) err = fft_set_smem(es_fft_grid_cols_screen_kernel<T, SEQ>, max_smem);
        if (!err) err = fft_set_smem(es_fft_degrid_cols_kernel<T, SEQ>, max_smem);
        if (!err) err = fft_set_smem(es_fft_degrid_rows_kernel<T, SEQ>, max_smem);
        return err;
    }

    static int rows_per_sm(int threads, size_t smem_bytes, int* per_sm)
    {
        return (int) cudaOccupancyMaxActiveBlocksPerMultiprocessor(per_sm,
                es_fft_grid_rows_kernel<T, SEQ>, threads, smem_bytes);
    }

    static int grid_rows(EsFft* f, const void* planes, int nplanes,
            void* strip, const uint32_t* occ, cudaStream_t s)
    {
        const EsFftLayout lay = fft_layout(f);
        const EsFftCfg& c = f->rows;
        /* compacted list of the occupied tile rows, then persistent CTAs */
        const int rc = fft_row_list(f, occ, nplanes, lay, s);
        if (rc) return rc;
        es_fft_grid_rows_kernel<T, SEQ><<<(unsigned) f->rows_grid, c.threads,
                c.smem_bytes, s>>>(f->line, (const T2*) f->d_tw, f->d_bpos,
                (const T2*) planes, (T2*) strip, f->image_size, occ, lay,
                f->d_row_list, f->d_row_work);
        return (int) cudaGetLastError();
    }

    static int grid_cols_screen(const EsFft* f, const EsGeom* g,
            const void* strip, int first_plane, int nplanes, void* image_t,
            const uint32_t* occ, cudaStream_t s)
    {
        const EsFftLayout lay = fft_layout(f);
        const EsFftCfg& c = f->cols;
        es_fft_grid_cols_screen_kernel<T, SEQ><<<
                    (unsigned) (f->image_size >> c.lshift), c.threads,
                    c.smem_bytes, s>>>(f->line, (const T2*) f->d_tw, f->d_bpos,
                    *g, (const T2*) strip, first_plane, nplanes, (T*) image_t,
                    occ, lay);
        return (int) cudaGetLastError();
    }

    static int degrid_cols(const EsFft* f, const EsGeom* g,
            const void* image_t, int first_plane, int nplanes, void* strip,
            const uint32_t* occ, cudaStream_t s)
    {
        const EsFftLayout lay = fft_layout(f);
        const EsFftCfg& c = f->cols;
        const int nbundles = f->image_size >> c.lshift;
        const int groups = (nbundles + ES_FFT_COL_GROUP - 1) / ES_FFT_COL_GROUP;
        const unsigned ctas = (unsigned) (groups * ES_FFT_COL_GROUP * nplanes);
        es_fft_degrid_cols_kernel<T, SEQ><<<ctas, c.threads,
                    c.smem_bytes, s>>>(f->line, (const T2*) f->d_tw, f->d_bpos,
                    *g, (const T*) image_t, first_plane, nplanes, (T2*) strip,
                    occ, lay);
        return (int) cudaGetLastError();
    }

    static int degrid_rows(EsFft* f, const void* strip, int nplanes,
            void* planes, const uint32_t* occ, cudaStream_t s)
    {
        const EsFftLayout lay = fft_layout(f);
        const EsFftCfg& c = f->rows;
        const int rc = fft_row_list(f, occ, nplanes, lay, s);
        if (rc) return rc;
        es_fft_degrid_rows_kernel<T, SEQ><<<(unsigned) f->rows_grid, c.threads,
                c.smem_bytes, s>>>(f->line, (const T2*) f->d_tw, f->d_bpos,
                (const T2*) strip, (T2*) planes, f->image_size, occ, lay,
                f->d_row_list, f->d_row_work);
        return (int) cudaGetLastError();
    }

    static const EsFftOps* ops()
    {
        static const EsFftOps table = {EsFftSeq<SEQ>::threads, prepare,
                rows_per_sm, grid_rows,
                grid_cols_screen, degrid_cols, degrid_rows};
        return &table;
    }
};

} // namespace

#endif
