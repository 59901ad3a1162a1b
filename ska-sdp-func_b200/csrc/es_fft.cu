/*
 * Pruned w-plane FFT of the ES gridder for sm_100a: the 2-D transform of one
 * G x G w-plane, of which only the N x N image footprint is needed (gridding)
 * or non-zero (degridding), as two passes of whole-line FFTs held in shared
 * memory (es_fft_core.h), fused with what surrounds them. T = strip[q][u][v']
 * is the pruned intermediate (G rows of N elements per plane).
 *
 *   gridding   pass 1  es_fft_grid_rows_kernel: one plane row (along v) per
 *                      CTA -> inverse FFT -> the N needed outputs -> T[u][:]
 *              pass 2  es_fft_grid_cols_screen_kernel: a bundle of 1/2/4
 *                      adjacent columns of T (along u) per CTA -> inverse FFT
 *                      -> the N needed outputs x w-screen -> summed into
 *                      imageT[col][row] with fire-and-forget reductions
 *   degridding pass 1  es_fft_degrid_cols_kernel: a bundle of imageT columns
 *                      x conj w-screen, zero-padded -> forward FFT (along u)
 *                      -> T[:][cols]
 *              pass 2  es_fft_degrid_rows_kernel: T[u][:] zero-padded ->
 *                      forward FFT (along v) -> plane row
 *
 * The first radix pass of every line takes its operands straight from global
 * memory into registers and the last one stores straight from registers, so
 * a line crosses shared memory once per inner pass only; the row kernels are
 * fully coalesced, the column kernels move 8 x lines-byte pieces (a bundle of
 * 4 fp32 columns fills 32-byte sectors).
 *
 * Replaces, per plane, the reference's cufftExecC2C/Z2Z on the full plane
 * (sdp_fft.cpp:425-456) + apply_w_screen_and_sum / reverse_w_screen_to_stack
 * (kernels.cu:429-679) with 2 G^2 + 2 G N (+ image) element traversals
 * instead of ~10 G^2 + 2 N^2 (cuFFT's 2-D plan needs five passes on these
 * 11-smooth sizes, profiles/r01_launches_c2_grid_v1.csv).
 * The arithmetic is the same DFT; the w-screen maths is that of
 * kernels.cu:110-123 with the phase reduced in fp64.
 */
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
#define ES_FFT_THREADS 512 /* upper bound of the CTA size (128 registers each) */
#define ES_FFT_THREADS_WIDE 1024 /* variant for radices <= 16: 64 registers each,
                                  * twice the resident warps per line (fp32 only) */
#define ES_FFT_WIDE_MAX_RADIX 16
#define ES_FFT_MAX_SMEM (227 * 1024 - 576) /* static: 512 B of occupancy flags + the work item */

/* CTA shape of one kernel family. */
struct EsFftCfg
{
    int32_t threads, lshift, pitch;
    int32_t tw_global; /* twiddle tables stay in global memory (L1): the line fills shared memory */
    size_t smem_bytes;
};

struct EsFft
{
    EsFftLine line;
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
    __device__ __forceinline__ unsigned int any_bits() const { return any; }
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
        if (ANY) any |= nonzero_bits(v); /* -0.0 counts: it only costs an FFT */
        return v;
    }
};

template<typename T2, bool WINDOW, bool MASK = false> struct LinearOut
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
template<int SGN, bool DIT, typename T, int MAXR, typename T2, typename In,
        typename Out>
__device__ __forceinline__ void fft_all_passes(T2* x, const EsFftLine& pl,
        const T2* tw, const int32_t* __restrict__ bpos, const int lshift,
        const int pitch, In& in, Out& out)
{
    for (int step = 0; step < pl.num_stages; ++step)
    {
        if (step > 0) __syncthreads();
        es_line_pass<SGN, DIT, T, MAXR>(x, pl, step, tw, bpos, lshift, pitch,
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

template<typename T, int MAXT>
__global__ void __launch_bounds__(MAXT, 1) es_fft_grid_rows_kernel(
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
        LinearOut<T2, true> out;
        out.base = strip + (size_t) q * lay.plane_elems + (size_t) u * n;
        out.estride = 1;
        out.lstride = n;
        out.k0 = gsz / 2 - n / 2;
        out.count = n;
        for (int step = 0; step < pl.num_stages; ++step)
        {
            es_line_pass<1, false, T,
                    (MAXT > ES_FFT_THREADS ? ES_FFT_WIDE_MAX_RADIX : 32)>(x, pl,
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
template<typename T, int MAXT>
__global__ void __launch_bounds__(MAXT, 1)
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
            es_line_pass<1, false, T, (MAXT > ES_FFT_THREADS ? ES_FFT_WIDE_MAX_RADIX : 32)>(x, pl, step, tw, bpos, lay.lshift,
                    lay.cols_pitch, (int) threadIdx.x, (int) blockDim.x, in,
                    out);
        }
    }
}

/* ---------------------------------------------------- degridding pass 1 */

template<typename T, int MAXT>
__global__ void __launch_bounds__(MAXT, 1) es_fft_degrid_cols_kernel(
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
    fft_all_passes<-1, true, T,
            (MAXT > ES_FFT_THREADS ? ES_FFT_WIDE_MAX_RADIX : 32)>(x, pl, tw, bpos,
            lay.lshift, lay.cols_pitch, in, out);
}

/* ---------------------------------------------------- degridding pass 2 */

template<typename T, int MAXT>
__global__ void __launch_bounds__(MAXT, 1) es_fft_degrid_rows_kernel(
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
        LinearIn<T2, true, false> in;
        in.base = strip + (size_t) q * lay.plane_elems + (size_t) u * n;
        in.estride = 1;
        in.lstride = n;
        in.k0 = gsz / 2 - n / 2;
        in.count = n;
        in.any = 0u;
        LinearOut<T2, false, true> out;
        out.bits = s_bits;
        out.base = planes + ((size_t) q * gsz + u) * gsz;
        out.estride = 1;
        out.lstride = gsz;
        out.k0 = 0;
        out.count = gsz;
        fft_all_passes<-1, true, T,
                (MAXT > ES_FFT_THREADS ? ES_FFT_WIDE_MAX_RADIX : 32)>(x, pl, tw,
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

/* ================================================================ C ABI */

namespace {

/* Elements between the lines of a bundle: >= padded line, and
 * pitch mod 16 == 16 / lines so that the 16 threads of a shared-memory phase
 * (lines x consecutive elements) fall into distinct 8-byte banks. */
int bundle_pitch(const EsFftLine* line, int lshift)
{
    int pitch = line->padded_n;
    if (lshift == 0) return pitch + (pitch & 1); /* 16-byte aligned lines */
    const int want = 16 >> lshift;
    while (pitch % 16 != want) pitch++;
    return pitch;
}

void make_cfg(const EsFftLine* line, size_t cplx, int lshift, int threads,
        EsFftCfg* c)
{
    c->lshift = lshift;
    c->pitch = bundle_pitch(line, lshift);
    c->tw_global = 0;
    c->smem_bytes = ((size_t) c->pitch << lshift) * cplx +
            (size_t) (line->num_twiddles + 1) * cplx;
    /* tests: SKA_SDP_FUNC_ES_FFT_TW_GLOBAL=1 forces this mode at any size */
    const bool force = getenv("SKA_SDP_FUNC_ES_FFT_TW_GLOBAL") != 0;
    if (force || (c->smem_bytes > ES_FFT_MAX_SMEM && lshift == 0 &&
            (size_t) c->pitch * cplx <= ES_FFT_MAX_SMEM))
    {
        /* e.g. fp64 lines of 14400 elements (230 KB): everything but the
         * line itself leaves shared memory */
        c->tw_global = 1;
        c->smem_bytes = ((size_t) c->pitch << lshift) * cplx;
    }
    if (threads <= 0)
    {
        /* as many CTAs per SM as shared memory allows (at most 4), all of
         * them together 512 threads of 128 registers */
        int ctas = (int) ((size_t) ES_FFT_MAX_SMEM / (c->smem_bytes + 1024));
        if (ctas < 1) ctas = 1;
        if (ctas > 4) ctas = 4;
        threads = ES_FFT_THREADS / ctas;
        /* no more threads than butterflies of the narrowest pass */
        int min_bf = line->n;
        for (int i = 0; i < line->num_stages; ++i)
        {
            const int bf = (line->n / line->stage[i].radix) << lshift;
            if (bf < min_bf) min_bf = bf;
        }
        while (threads > 32 && threads / 2 >= min_bf) threads /= 2;
    }
    c->threads = threads;
}

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

} // namespace

extern "C" {

int es_fft_create(int grid_size, int image_size, int is_double, EsFft** out)
{
    *out = 0;
    if (getenv("SKA_SDP_FUNC_ES_USE_CUFFT")) return -1;
    EsFft* f = (EsFft*) calloc(1, sizeof(EsFft));
    if (!f) return (int) cudaErrorMemoryAllocation;
    /* experiments: SKA_SDP_FUNC_ES_FFT_RADICES="28,20,11" forces the passes */
    int32_t forced[ES_FFT_MAX_STAGES + 1];
    memset(forced, 0, sizeof(forced));
    const char* renv = getenv("SKA_SDP_FUNC_ES_FFT_RADICES");
    if (renv)
    {
        char buf[128];
        strncpy(buf, renv, sizeof(buf) - 1);
        buf[sizeof(buf) - 1] = 0;
        int c = 0;
        for (char* p = strtok(buf, ","); p && c < ES_FFT_MAX_STAGES;
                p = strtok(0, ","))
            forced[c++] = atoi(p);
    }
    const int max_radix = is_double ? ES_FFT_MAX_RADIX_DOUBLE : 32;
    int bad = es_fft_line_plan(grid_size, max_radix, forced[0] ? forced : 0,
            &f->line);
    if (bad && forced[0])
        bad = es_fft_line_plan(grid_size, max_radix, 0, &f->line);
    if (!bad && is_double)
        for (int i = 0; i < f->line.num_stages; ++i)
            if (f->line.stage[i].radix > ES_FFT_MAX_RADIX_DOUBLE) bad = 1;
    if (bad || (grid_size & 1) || (image_size & 1) || image_size > grid_size ||
            grid_size > ES_FFT_OCC_WORDS * 32 * ES_TILE)
    {
        free(f);
        return -1; /* unsupported size: caller falls back to cuFFT */
    }
    f->grid_size = grid_size;
    f->image_size = image_size;
    f->is_double = is_double;
    const size_t cplx = is_double ? 16 : 8;
    /* experiments: SKA_SDP_FUNC_ES_FFT_CFG="rows_threads,cols_threads,cols_lshift"
     * default: bundles of 4 (fp64: 2) adjacent columns / 2 (fp64: 1) rows per CTA where they fit
     * in shared memory (32-byte pieces of the row-major intermediate) */
    int e[4] = {0, 0, is_double ? 1 : 2, is_double ? 0 : 1};
    const char* env = getenv("SKA_SDP_FUNC_ES_FFT_CFG");
    int rows_forced = 0;
    if (env) rows_forced = sscanf(env, "%d,%d,%d,%d", &e[0], &e[1], &e[2],
            &e[3]) == 4;
    if (e[3] < 0 || e[3] > 2) e[3] = 0;
    /* more than 512 threads: fp32 lines whose radices all fit the wide kernel
     * variant (64 registers per thread) */
    int max_planned = 0;
    for (int i = 0; i < f->line.num_stages; ++i)
        if (f->line.stage[i].radix > max_planned)
            max_planned = f->line.stage[i].radix;
    const int thread_cap = (!is_double && max_planned <= ES_FFT_WIDE_MAX_RADIX) ?
            ES_FFT_THREADS_WIDE : ES_FFT_THREADS;
    for (int k = 0; k < 2; ++k)
        if (e[k] < 0 || e[k] > thread_cap || (e[k] & 31)) e[k] = 0;
    if (e[2] < 0 || e[2] > 2) e[2] = 0;
    /* rows of one CTA share a 64-row occupancy block and must tile the grid */
    int rshift = e[3];
    for (;; --rshift)
    {
        make_cfg(&f->line, cplx, rshift, e[0], &f->rows);
        /* default: a row bundle only where two CTAs still fit on an SM (C5:
         * 2 rows of 11520 leave one 512-thread CTA: rows 1365 vs 1314 ms);
         * a bundle forced by the environment only has to fit */
        const size_t need = rows_forced ? f->rows.smem_bytes :
                2 * (f->rows.smem_bytes + 1024);
        if ((need <= ES_FFT_MAX_SMEM && grid_size % (1 << rshift) == 0) ||
                rshift == 0) break;
    }
    int lshift = e[2];
    for (;; --lshift)
    {
        make_cfg(&f->line, cplx, lshift, e[1], &f->cols);
        if ((f->cols.smem_bytes <= ES_FFT_MAX_SMEM &&
                image_size % (1 << lshift) == 0) || lshift == 0) break;
    }
    f->plane_elems = (int64_t) grid_size * image_size;
    if (f->rows.smem_bytes > ES_FFT_MAX_SMEM ||
            f->cols.smem_bytes > ES_FFT_MAX_SMEM)
    {
        free(f);
        return -1;
    }
    if (getenv("SKA_SDP_FUNC_ES_DEBUG"))
    {
        fprintf(stderr, "es_fft: n=%d passes", grid_size);
        for (int k = 0; k < f->line.num_stages; ++k)
            fprintf(stderr, " %d", f->line.stage[k].radix);
        fprintf(stderr, "; pad %d; rows %d thr (%zu B), cols %d thr x %d lines "
                "(%zu B)\n", f->line.pad, f->rows.threads,
                f->rows.smem_bytes, f->cols.threads, 1 << f->cols.lshift,
                f->cols.smem_bytes);
    }

    const int nt = f->line.num_twiddles;
    const int nb = grid_size / f->line.stage[f->line.num_stages - 1].radix;
    double* tw = (double*) malloc(sizeof(double) * 2 * (size_t) (nt + 1));
    int32_t* bpos = (int32_t*) malloc(sizeof(int32_t) * (size_t) nb);
    void* tw_t = malloc(cplx * (size_t) (nt + 1));
    int err = 0;
    if (!tw || !bpos || !tw_t) err = (int) cudaErrorMemoryAllocation;
    if (!err)
    {
        memset(tw, 0, sizeof(double) * 2 * (size_t) (nt + 1));
        es_fft_line_tables(&f->line, tw, 0, bpos);
        for (int i = 0; i < 2 * (nt + 1); ++i)
        {
            if (is_double) ((double*) tw_t)[i] = tw[i];
            else ((float*) tw_t)[i] = (float) tw[i];
        }
        err = (int) cudaMalloc(&f->d_tw, cplx * (size_t) (nt + 1));
        if (!err) err = (int) cudaMalloc((void**) &f->d_bpos,
                sizeof(int32_t) * (size_t) nb);
        if (!err) err = (int) cudaMemcpy(f->d_tw, tw_t, cplx * (size_t) (nt + 1),
                cudaMemcpyHostToDevice);
        if (!err) err = (int) cudaMemcpy(f->d_bpos, bpos,
                sizeof(int32_t) * (size_t) nb, cudaMemcpyHostToDevice);
    }
    free(tw); free(bpos); free(tw_t);
    /* The dynamic shared-memory limit belongs to the function and the
     * device, not to this plan: always opt in to the maximum any plan can
     * ask for, so that plans of different sizes can be alive together. */
    const size_t sm = ES_FFT_MAX_SMEM;
    if (!err && is_double)
    {
        err = fft_set_smem(es_fft_grid_rows_kernel<double, ES_FFT_THREADS>, sm);
        if (!err) err = fft_set_smem(es_fft_grid_cols_screen_kernel<double, ES_FFT_THREADS>, sm);
        if (!err) err = fft_set_smem(es_fft_degrid_cols_kernel<double, ES_FFT_THREADS>, sm);
        if (!err) err = fft_set_smem(es_fft_degrid_rows_kernel<double, ES_FFT_THREADS>, sm);
    }
    else if (!err)
    {
        err = fft_set_smem(es_fft_grid_rows_kernel<float, ES_FFT_THREADS>, sm);
        if (!err) err = fft_set_smem(es_fft_grid_rows_kernel<float, ES_FFT_THREADS_WIDE>, sm);
        if (!err) err = fft_set_smem(es_fft_grid_cols_screen_kernel<float, ES_FFT_THREADS>, sm);
        if (!err) err = fft_set_smem(es_fft_grid_cols_screen_kernel<float, ES_FFT_THREADS_WIDE>, sm);
        if (!err) err = fft_set_smem(es_fft_degrid_cols_kernel<float, ES_FFT_THREADS>, sm);
        if (!err) err = fft_set_smem(es_fft_degrid_cols_kernel<float, ES_FFT_THREADS_WIDE>, sm);
        if (!err) err = fft_set_smem(es_fft_degrid_rows_kernel<float, ES_FFT_THREADS>, sm);
        if (!err) err = fft_set_smem(es_fft_degrid_rows_kernel<float, ES_FFT_THREADS_WIDE>, sm);
    }
    if (!err)
    {
        /* persistent row kernels: as many CTAs as the device holds at once */
        int dev = 0, sms = 0, per_sm = 0;
        err = (int) cudaGetDevice(&dev);
        if (!err) err = (int) cudaDeviceGetAttribute(&sms,
                cudaDevAttrMultiProcessorCount, dev);
        if (!err && is_double)
            err = (int) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm,
                    es_fft_grid_rows_kernel<double, ES_FFT_THREADS>,
                    f->rows.threads, f->rows.smem_bytes);
        else if (!err && f->rows.threads > ES_FFT_THREADS)
            err = (int) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm,
                    es_fft_grid_rows_kernel<float, ES_FFT_THREADS_WIDE>,
                    f->rows.threads, f->rows.smem_bytes);
        else if (!err)
            err = (int) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm,
                    es_fft_grid_rows_kernel<float, ES_FFT_THREADS>,
                    f->rows.threads, f->rows.smem_bytes);
        if (per_sm < 1) per_sm = 1;
        f->rows_grid = sms * per_sm;
        if (!err) err = (int) cudaMalloc((void**) &f->d_row_work,
                2 * sizeof(uint32_t));
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

/* (plane, tile row) list of a batch -> fft->d_row_list / d_row_work. */
static int fft_row_list(EsFft* f, const uint32_t* occ, int nplanes,
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

void es_fft_destroy(EsFft* f)
{
    if (!f) return;
    if (f->d_tw) (void) cudaFree(f->d_tw);
    if (f->d_bpos) (void) cudaFree(f->d_bpos);
    if (f->d_row_list) (void) cudaFree(f->d_row_list);
    if (f->d_row_work) (void) cudaFree(f->d_row_work);
    if (f->d_nm1) (void) cudaFree(f->d_nm1);
    free(f);
}

size_t es_fft_strip_bytes(const EsFft* f, int nplanes)
{
    return (size_t) nplanes * (size_t) f->plane_elems * (f->is_double ? 16 : 8);
}

int es_fft_grid_rows(const EsFft* f, const void* planes, int nplanes,
        void* strip, const uint32_t* occ, es_stream_t stream)
{
    cudaStream_t s = (cudaStream_t) stream;
    const EsFftLayout lay = fft_layout(f);
    const EsFftCfg& c = f->rows;
    /* compacted list of the occupied tile rows, then persistent CTAs */
    EsFft* fm = const_cast<EsFft*>(f);
    const int rc = fft_row_list(fm, occ, nplanes, lay, s);
    if (rc) return rc;
    dim3 grid((unsigned) f->rows_grid);
    if (f->is_double)
        es_fft_grid_rows_kernel<double, ES_FFT_THREADS><<<grid, c.threads, c.smem_bytes,
                s>>>(f->line, (const double2*) f->d_tw, f->d_bpos,
                (const double2*) planes, (double2*) strip, f->image_size,
                occ, lay, f->d_row_list, f->d_row_work);
    else if (c.threads > ES_FFT_THREADS)
        es_fft_grid_rows_kernel<float, ES_FFT_THREADS_WIDE><<<grid, c.threads, c.smem_bytes,
                s>>>(f->line, (const float2*) f->d_tw, f->d_bpos,
                (const float2*) planes, (float2*) strip, f->image_size,
                occ, lay, f->d_row_list, f->d_row_work);
    else
        es_fft_grid_rows_kernel<float, ES_FFT_THREADS><<<grid, c.threads, c.smem_bytes,
                s>>>(f->line, (const float2*) f->d_tw, f->d_bpos,
                (const float2*) planes, (float2*) strip, f->image_size,
                occ, lay, f->d_row_list, f->d_row_work);
    return (int) cudaGetLastError();
}

int es_fft_grid_cols_screen(const EsFft* f, const EsGeom* g, const void* strip,
        int first_plane, int nplanes, void* image_t, const uint32_t* occ,
        es_stream_t stream)
{
    cudaStream_t s = (cudaStream_t) stream;
    const EsFftLayout lay = fft_layout(f);
    const EsFftCfg& c = f->cols;
    dim3 grid((unsigned) (f->image_size >> c.lshift));
    if (f->is_double)
        es_fft_grid_cols_screen_kernel<double, ES_FFT_THREADS><<<grid, c.threads,
                c.smem_bytes, s>>>(f->line, (const double2*) f->d_tw,
                f->d_bpos, *g, (const double2*) strip, first_plane, nplanes,
                (double*) image_t, occ, lay);
    else if (c.threads > ES_FFT_THREADS)
        es_fft_grid_cols_screen_kernel<float, ES_FFT_THREADS_WIDE><<<grid, c.threads,
                c.smem_bytes, s>>>(f->line, (const float2*) f->d_tw,
                f->d_bpos, *g, (const float2*) strip, first_plane, nplanes,
                (float*) image_t, occ, lay);
    else
        es_fft_grid_cols_screen_kernel<float, ES_FFT_THREADS><<<grid, c.threads,
                c.smem_bytes, s>>>(f->line, (const float2*) f->d_tw,
                f->d_bpos, *g, (const float2*) strip, first_plane, nplanes,
                (float*) image_t, occ, lay);
    return (int) cudaGetLastError();
}

int es_fft_degrid_cols(const EsFft* f, const EsGeom* g, const void* image_t,
        int first_plane, int nplanes, void* strip, const uint32_t* occ,
        es_stream_t stream)
{
    cudaStream_t s = (cudaStream_t) stream;
    const EsFftLayout lay = fft_layout(f);
    const EsFftCfg& c = f->cols;
    const int nbundles = f->image_size >> c.lshift;
    const int groups = (nbundles + ES_FFT_COL_GROUP - 1) / ES_FFT_COL_GROUP;
    dim3 grid((unsigned) (groups * ES_FFT_COL_GROUP * nplanes));
    if (f->is_double)
        es_fft_degrid_cols_kernel<double, ES_FFT_THREADS><<<grid, c.threads, c.smem_bytes,
                s>>>(f->line, (const double2*) f->d_tw, f->d_bpos, *g,
                (const double*) image_t, first_plane, nplanes,
                (double2*) strip, occ, lay);
    else if (c.threads > ES_FFT_THREADS)
        es_fft_degrid_cols_kernel<float, ES_FFT_THREADS_WIDE><<<grid, c.threads, c.smem_bytes,
                s>>>(f->line, (const float2*) f->d_tw, f->d_bpos, *g,
                (const float*) image_t, first_plane, nplanes,
                (float2*) strip, occ, lay);
    else
        es_fft_degrid_cols_kernel<float, ES_FFT_THREADS><<<grid, c.threads, c.smem_bytes,
                s>>>(f->line, (const float2*) f->d_tw, f->d_bpos, *g,
                (const float*) image_t, first_plane, nplanes,
                (float2*) strip, occ, lay);
    return (int) cudaGetLastError();
}

int es_fft_degrid_rows(const EsFft* f, const void* strip, int nplanes,
        void* planes, const uint32_t* occ, es_stream_t stream)
{
    cudaStream_t s = (cudaStream_t) stream;
    const EsFftLayout lay = fft_layout(f);
    const EsFftCfg& c = f->rows;
    /* compacted list of the occupied tile rows, then persistent CTAs */
    EsFft* fm = const_cast<EsFft*>(f);
    const int rc = fft_row_list(fm, occ, nplanes, lay, s);
    if (rc) return rc;
    dim3 grid((unsigned) f->rows_grid);
    if (f->is_double)
        es_fft_degrid_rows_kernel<double, ES_FFT_THREADS><<<grid, c.threads, c.smem_bytes,
                s>>>(f->line, (const double2*) f->d_tw, f->d_bpos,
                (const double2*) strip, (double2*) planes, f->image_size,
                occ, lay, f->d_row_list, f->d_row_work);
    else if (c.threads > ES_FFT_THREADS)
        es_fft_degrid_rows_kernel<float, ES_FFT_THREADS_WIDE><<<grid, c.threads, c.smem_bytes,
                s>>>(f->line, (const float2*) f->d_tw, f->d_bpos,
                (const float2*) strip, (float2*) planes, f->image_size,
                occ, lay, f->d_row_list, f->d_row_work);
    else
        es_fft_degrid_rows_kernel<float, ES_FFT_THREADS><<<grid, c.threads, c.smem_bytes,
                s>>>(f->line, (const float2*) f->d_tw, f->d_bpos,
                (const float2*) strip, (float2*) planes, f->image_size,
                occ, lay, f->d_row_list, f->d_row_work);
    return (int) cudaGetLastError();
}

int es_fft_image_transpose(const EsFft* f, const void* src, void* dst,
        int add, int wstack, es_stream_t stream)
{
    cudaStream_t s = (cudaStream_t) stream;
    const int n = f->image_size;
    const double* tab = wstack ? f->d_nm1 : 0;
    if (wstack && !tab) return (int) cudaErrorInvalidValue;
    dim3 block(32, 8), grid((unsigned) ((n + 31) / 32), (unsigned) ((n + 31) / 32));
    if (f->is_double)
    {
        if (add) es_transpose_screen_kernel<double, true><<<grid, block, 0, s>>>(
                (const double*) src, (double*) dst, n, tab, f->nm1_pitch);
        else es_transpose_screen_kernel<double, false><<<grid, block, 0, s>>>(
                (const double*) src, (double*) dst, n, tab, f->nm1_pitch);
    }
    else
    {
        if (add) es_transpose_screen_kernel<float, true><<<grid, block, 0, s>>>(
                (const float*) src, (float*) dst, n, tab, f->nm1_pitch);
        else es_transpose_screen_kernel<float, false><<<grid, block, 0, s>>>(
                (const float*) src, (float*) dst, n, tab, f->nm1_pitch);
    }
    return (int) cudaGetLastError();
}

int es_fft_set_screen(EsFft* f, double pixel_size, double max_abs_w)
{
    const int count = f->image_size / 2 + 1;
    const int pitch = (count + 3) & ~3;
    if (!f->d_nm1)
    {
        cudaError_t err = cudaMalloc((void**) &f->d_nm1,
                sizeof(double) * (size_t) count * (size_t) pitch);
        if (err != cudaSuccess) return (int) err;
    }
    f->nm1_pitch = pitch;
    const double pix2 = pixel_size * pixel_size;
    /* largest |w (n - 1)| of the plan, in turns: the corner pixel */
    const double half = 0.5 * (double) f->image_size;
    const double sos = 2.0 * pix2 * half * half;
    const double nm1_max = sos / (sqrt(sos < 1.0 ? 1.0 - sos : 0.0) + 1.0);
    const int bits = es_screen_fixed_bits(fabs(max_abs_w) * nm1_max);
    f->screen_shift = 32 - bits;
    f->screen_magic = es_screen_magic(bits);
    dim3 grid((unsigned) ((count + 255) / 256), (unsigned) count);
    es_screen_table_kernel<<<grid, 256>>>(f->d_nm1, count, pitch, pix2);
    cudaError_t err = cudaGetLastError();
    if (err == cudaSuccess) err = cudaDeviceSynchronize();
    return (int) err;
}

} // extern "C"
