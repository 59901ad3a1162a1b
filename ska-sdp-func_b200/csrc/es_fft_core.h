/*
 * In-place mixed-radix line FFT for 11-smooth lengths, the arithmetic core of
 * the pruned w-plane FFT kernels (es_fft.cu). One line of length n lives in
 * (shared) memory; the transform is a sequence of radix-r passes, each a set
 * of n / r independent butterflies, so a pass maps onto "thread q does
 * butterfly q" with a barrier between passes.
 *
 * Two flow graphs over the same stage list (r_1 .. r_m, L_1 = n,
 * L_{i+1} = s_i = L_i / r_i):
 *   DIF (decimation in frequency): natural-order input, stages 1..m, twiddle
 *       AFTER the butterfly; X[k] ends up at position perm[k];
 *   DIT (decimation in time): input x[k] placed at perm[k], stages m..1,
 *       twiddle BEFORE the butterfly; natural-order output.
 * perm[k] = sum_i k_i s_i with k = k_1 + r_1 (k_2 + r_2 (k_3 + ...)).
 * The gridder uses DIF (dense input, pruned output), the degridder DIT
 * (zero-padded input, dense output).
 *
 * Twiddles w_L^j = exp(+2 pi i j / L) come from a two-level table per stage
 * (hi[j >> 6] * lo[j & 63], one complex product, ~1 ulp); the powers
 * w^(j k), k = 2 .. r-1, by the balanced recurrence w^2m = (w^m)^2,
 * w^(2m+1) = w^2m w. The forward transform conjugates the table values.
 *
 * Everything here is __host__ __device__ so that tests/test_fft_core.py can
 * run exactly this code on the CPU against numpy.
 */
#ifndef ES_FFT_CORE_H_
#define ES_FFT_CORE_H_

#include <stdint.h>

#if defined(__CUDACC__)
#define ES_HD __host__ __device__ __forceinline__
#else
#define ES_HD inline
#endif

#include "es_fft_core_types.h"

template<typename T> struct EsCplx { T x, y; };

template<typename T2> ES_HD T2 es_cadd(const T2 a, const T2 b)
{
    T2 r; r.x = a.x + b.x; r.y = a.y + b.y; return r;
}

template<typename T2> ES_HD T2 es_csub(const T2 a, const T2 b)
{
    T2 r; r.x = a.x - b.x; r.y = a.y - b.y; return r;
}

template<typename T2> ES_HD T2 es_cmul(const T2 a, const T2 b)
{
    T2 r;
    r.x = a.x * b.x - a.y * b.y;
    r.y = a.x * b.y + a.y * b.x;
    return r;
}

/* a * (SGN i) */
template<int SGN, typename T2> ES_HD T2 es_mul_i(const T2 a)
{
    T2 r;
    if (SGN > 0) { r.x = -a.y; r.y = a.x; }
    else { r.x = a.y; r.y = -a.x; }
    return r;
}

/* c + a * s with a real scalar s */
template<typename T2, typename T> ES_HD T2 es_caxpy(const T2 c, const T2 a,
        const T s)
{
    T2 r; r.x = c.x + a.x * s; r.y = c.y + a.y * s; return r;
}

#if defined(__CUDACC__) && !defined(ES_FFT_NO_PACKED)
/* fp32 on sm_100a: the fma pipe issues one instruction per two cycles and
 * scheduler, scalar or packed, so complex arithmetic goes through the packed
 * forms FADD2 / FMUL2 / FFMA2 on (re, im) register pairs. The operand swap,
 * the partial negation and the scalar broadcast below are operand modifiers
 * of those instructions (.LO_HI, .NP, .F32), not extra instructions: a
 * complex product is FMUL2 + FFMA2 instead of FMUL, FFMA, FMUL, FFMA.
 * Non-template overloads: chosen over the generic forms for float2. */
ES_HD float2 es_cadd(const float2 a, const float2 b)
{
#if defined(__CUDA_ARCH__)
    return __fadd2_rn(a, b);
#else
    return make_float2(a.x + b.x, a.y + b.y);
#endif
}

ES_HD float2 es_csub(const float2 a, const float2 b)
{
#if defined(__CUDA_ARCH__)
    return __fadd2_rn(a, make_float2(-b.x, -b.y));
#else
    return make_float2(a.x - b.x, a.y - b.y);
#endif
}

ES_HD float2 es_cmul(const float2 a, const float2 b)
{
#if defined(__CUDA_ARCH__)
    return __ffma2_rn(a, make_float2(b.x, b.x),
            __fmul2_rn(make_float2(a.y, a.x), make_float2(-b.y, b.y)));
#else
    return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
#endif
}

ES_HD float2 es_caxpy(const float2 c, const float2 a, const float s)
{
#if defined(__CUDA_ARCH__)
    return __ffma2_rn(a, make_float2(s, s), c);
#else
    return make_float2(c.x + a.x * s, c.y + a.y * s);
#endif
}
#endif

/* cos / sin (2 pi k / 16), k < 8 */
#define ES_COS16 {1.0, 0.92387953251128674, 0.70710678118654757, \
    0.38268343236508984, 0.0, -0.38268343236508973, -0.70710678118654746, \
    -0.92387953251128674}
#define ES_SIN16 {0.0, 0.38268343236508978, 0.70710678118654746, \
    0.92387953251128674, 1.0, 0.92387953251128674, 0.70710678118654757, \
    0.38268343236508989}

/* d * w_n^i with w_n = exp(SGN 2 pi i / n), n in {2, 4, 8, 16}; (i, n) are
 * compile-time constants after unrolling, so the branches fold away. */
template<int SGN, typename T, typename T2>
ES_HD T2 es_mul_w_pow2(const T2 d, const int i, const int n)
{
    const double c16[8] = ES_COS16, s16[8] = ES_SIN16;
    const int k = i * (16 / n); /* index into the n = 16 tables, k < 8 */
    if (k == 0) return d;
    if (k == 4) return es_mul_i<SGN>(d);
    T2 w;
    w.x = (T) c16[k];
    w.y = (T) (SGN > 0 ? s16[k] : -s16[k]);
    return es_cmul(d, w);
}

/* Radix-2^m DFT in registers (decimation in frequency); leaves output k at
 * v[bitrev(k)]. */
template<int R, int SGN, typename T, typename T2>
ES_HD void es_dft_pow2(T2 (&v)[R])
{
#pragma unroll
    for (int half = R / 2; half >= 1; half >>= 1)
    {
#pragma unroll
        for (int g = 0; g < R; g += 2 * half)
        {
#pragma unroll
            for (int i = 0; i < half; ++i)
            {
                const T2 a = v[g + i], b = v[g + i + half];
                v[g + i] = es_cadd(a, b);
                v[g + i + half] = es_mul_w_pow2<SGN, T>(es_csub(a, b), i,
                        2 * half);
            }
        }
    }
}

template<int R> ES_HD int es_bitrev(int k)
{
    int r = 0;
#pragma unroll
    for (int b = 1; b < R; b <<= 1)
    {
        r = (r << 1) | (k & 1);
        k >>= 1;
    }
    return r;
}

/* cos / sin (2 pi k / R) literals (generated, tools/gen_fft_trig.py) */
template<int R> struct EsTrigTab;
#include "es_fft_trig.inc"

/* Odd prime radix DFT in registers, natural-order output:
 * X[k], X[R-k] = (x0 + sum_m a_m cos(2 pi k m / R)) +- SGN i sum_m b_m sin(..)
 * with a_m = x[m] + x[R-m], b_m = x[m] - x[R-m]. */
template<int R, int SGN, typename T, typename T2>
ES_HD void es_dft_odd(T2 (&v)[R])
{
    const int H = (R - 1) / 2;
    T2 a[H], b[H];
#pragma unroll
    for (int m = 0; m < H; ++m)
    {
        a[m] = es_cadd(v[m + 1], v[R - 1 - m]);
        b[m] = es_csub(v[m + 1], v[R - 1 - m]);
    }
    const T2 x0 = v[0];
    T2 sum = x0;
#pragma unroll
    for (int m = 0; m < H; ++m) sum = es_cadd(sum, a[m]);
    v[0] = sum;
#pragma unroll
    for (int k = 1; k <= H; ++k)
    {
        T2 c = x0, d;
        d.x = (T) 0; d.y = (T) 0;
#pragma unroll
        for (int m = 1; m <= H; ++m)
        {
            const T cs = (T) EsTrigTab<R>::c((k * m) % R);
            const T sn = (T) EsTrigTab<R>::s((k * m) % R);
            c = es_caxpy(c, a[m - 1], cs);
            d = es_caxpy(d, b[m - 1], sn);
        }
        const T2 id = es_mul_i<SGN>(d);
        v[k] = es_cadd(c, id);
        v[R - k] = es_csub(c, id);
    }
}

/* ---- composite radices R = A * B (Cooley-Tukey inside the registers) ----
 * n = B n1 + n2, k = k1 + A k2:
 *   X[k1 + A k2] = sum_n2 w_B^(n2 k2) w_R^(n2 k1) sum_n1 w_A^(n1 k1) x[B n1 + n2]
 * i.e. B DFTs of length A over stride B, constant twiddles, A DFTs of length B
 * over contiguous registers. All indices are compile-time after unrolling. */
template<int R> struct EsSplit { static const int A = 0, B = 0; };
#define ES_FFT_SPLIT(R_, A_, B_) \
    template<> struct EsSplit<R_> { static const int A = A_, B = B_; };
ES_FFT_SPLIT(6, 2, 3) ES_FFT_SPLIT(9, 3, 3) ES_FFT_SPLIT(10, 2, 5)
ES_FFT_SPLIT(12, 4, 3) ES_FFT_SPLIT(14, 2, 7) ES_FFT_SPLIT(15, 3, 5)
ES_FFT_SPLIT(18, 2, 9) ES_FFT_SPLIT(20, 4, 5) ES_FFT_SPLIT(21, 3, 7)
ES_FFT_SPLIT(22, 2, 11) ES_FFT_SPLIT(24, 8, 3) ES_FFT_SPLIT(25, 5, 5)
ES_FFT_SPLIT(27, 3, 9) ES_FFT_SPLIT(28, 4, 7) ES_FFT_SPLIT(30, 2, 15)
ES_FFT_SPLIT(32, 4, 8)
#undef ES_FFT_SPLIT

template<int R> struct EsIsPow2
{
    static const bool value = (R == 2 || R == 4 || R == 8 || R == 16);
};

/* Radix-R DFT: output k is left in v[es_out_slot<R>(k)];
 * es_slot_index<R>(i) is the inverse map (which output sits in v[i]). */
template<int R> ES_HD int es_out_slot(int k)
{
    if constexpr (EsSplit<R>::A != 0)
    {
        const int A = EsSplit<R>::A, B = EsSplit<R>::B;
        return B * es_out_slot<A>(k % A) + es_out_slot<B>(k / A);
    }
    else if constexpr (EsIsPow2<R>::value) return es_bitrev<R>(k);
    else return k;
}

template<int R> ES_HD int es_slot_index(int i)
{
    if constexpr (EsSplit<R>::A != 0)
    {
        const int A = EsSplit<R>::A, B = EsSplit<R>::B;
        return es_slot_index<A>(i / B) + A * es_slot_index<B>(i % B);
    }
    else if constexpr (EsIsPow2<R>::value) return es_bitrev<R>(i);
    else return i;
}

/* d * exp(SGN 2 pi i num / den); (num, den) fold to constants. */
template<int DEN, int SGN, typename T, typename T2>
ES_HD T2 es_mul_w_const(const T2 d, const int num_)
{
    const int num = num_ % DEN;
    if (num == 0) return d;
    if (4 * num == DEN) return es_mul_i<SGN>(d);
    if (2 * num == DEN) { T2 r; r.x = -d.x; r.y = -d.y; return r; }
    if (4 * num == 3 * DEN) return es_mul_i<-SGN>(d);
    const T c = (T) EsTrigTab<DEN>::c(num);
    const T s = (T) (SGN > 0 ? EsTrigTab<DEN>::s(num) : -EsTrigTab<DEN>::s(num));
    T2 r;
    if (sizeof(T) == 8 && (8 * num == DEN || 8 * num == 5 * DEN))
    {
        /* c == s: (x - y) c, (x + y) c (fp32: the packed product below is
         * two instructions either way) */
        r.x = (d.x - d.y) * c;
        r.y = (d.x + d.y) * c;
        if (SGN < 0) { r.x = (d.x + d.y) * c; r.y = (d.y - d.x) * c; }
        return r;
    }
    T2 w;
    w.x = c;
    w.y = s;
    return es_cmul(d, w);
}

template<int R, int SGN, typename T, typename T2> ES_HD void es_dft(T2 (&v)[R]);

template<int A, int B, int SGN, typename T, typename T2>
ES_HD void es_dft_comp(T2 (&v)[A * B])
{
#pragma unroll
    for (int n2 = 0; n2 < B; ++n2)
    {
        T2 t[A];
#pragma unroll
        for (int n1 = 0; n1 < A; ++n1) t[n1] = v[B * n1 + n2];
        es_dft<A, SGN, T>(t);
#pragma unroll
        for (int i = 0; i < A; ++i)
            v[B * i + n2] = es_mul_w_const<A * B, SGN, T>(t[i],
                    n2 * es_slot_index<A>(i));
    }
#pragma unroll
    for (int i = 0; i < A; ++i)
    {
        T2 t[B];
#pragma unroll
        for (int n2 = 0; n2 < B; ++n2) t[n2] = v[B * i + n2];
        es_dft<B, SGN, T>(t);
#pragma unroll
        for (int n2 = 0; n2 < B; ++n2) v[B * i + n2] = t[n2];
    }
}

template<int R, int SGN, typename T, typename T2>
ES_HD void es_dft(T2 (&v)[R])
{
    if constexpr (EsSplit<R>::A != 0)
        es_dft_comp<EsSplit<R>::A, EsSplit<R>::B, SGN, T>(v);
    else if constexpr (EsIsPow2<R>::value)
        es_dft_pow2<R, SGN, T>(v);
    else
        es_dft_odd<R, SGN, T>(v);
}

/* w_L^j from the two-level table of a stage (conjugated for SGN < 0). */
template<int SGN, typename T2>
ES_HD T2 es_twiddle(const EsFftStage& st, const T2* tw, const int j)
{
    T2 w = tw[st.tw_lo + (j & (ES_FFT_TW_LO - 1))];
    if (st.stride > ES_FFT_TW_LO)
        w = es_cmul(tw[st.tw_hi + (j >> 6)], w);
    if (SGN < 0) w.y = -w.y;
    return w;
}

/* One radix-R butterfly in registers: v[k] = input k on entry; output k is
 * left in v[es_out_slot<R>(k)]. The stage twiddles w_L^(j k) are applied to
 * the inputs (DIT) or to the outputs (DIF); their powers come from the
 * balanced recurrence w^2m = (w^m)^2, w^(2m+1) = w^2m w. */
template<int R, int SGN, bool DIT, typename T, typename T2>
ES_HD void es_bf_core(T2 (&v)[R], const EsFftStage& st, const T2* tw,
        const int j)
{
    const bool twiddled = st.stride > 1;
    T2 w[R];
    if (twiddled)
    {
        w[1] = es_twiddle<SGN>(st, tw, j);
        if (DIT)
        {
#pragma unroll
            for (int k = 1; k < R; ++k)
            {
                if (k > 1)
                    w[k] = (k & 1) ? es_cmul(w[k - 1], w[1]) :
                            es_cmul(w[k / 2], w[k / 2]);
                v[k] = es_cmul(v[k], w[k]);
            }
        }
    }
    es_dft<R, SGN, T>(v);
    if (twiddled && !DIT)
    {
#pragma unroll
        for (int k = 1; k < R; ++k)
        {
            if (k > 1)
                w[k] = (k & 1) ? es_cmul(w[k - 1], w[1]) :
                        es_cmul(w[k / 2], w[k / 2]);
            v[es_out_slot<R>(k)] = es_cmul(v[es_out_slot<R>(k)], w[k]);
        }
    }
}

/* What one pass needs from the line plan, by value (the pass functions may
 * be real calls: no references to the kernel's parameter space). */
struct EsFftPassArgs
{
    EsFftStage st;
    int32_t n, pad;
    float inv_stride0; /* 1 / stage[0].stride */
};

/* Position of element `pos` of a line in (shared) memory: one pad of
 * pad elements after every block of stage[0].stride elements, so that
 * the digit-reversed last pass (consecutive threads stage[0].stride elements
 * apart) is free of bank conflicts. */
ES_HD int es_padded(const EsFftPassArgs& pa, const int pos)
{
    if (pa.pad == 0) return pos;
    return pos + pa.pad * (int) (((float) pos + 0.5f) * pa.inv_stride0);
}

/* Largest radix whose output-functor operands are loaded before the
 * butterfly arithmetic (es_line_pass_body). */
#ifndef ES_FFT_PREFETCH_MAX_RADIX
#define ES_FFT_PREFETCH_MAX_RADIX 11
#endif

/* Position of a stage in the stage list: it fixes the addressing pattern and,
 * with the flow graph, whether the pass reads from `in` / writes to `out`. */
#define ES_FFT_KIND_FIRST 0  /* stage 0 of several */
#define ES_FFT_KIND_MIDDLE 1
#define ES_FFT_KIND_LAST 2   /* last stage of several */
#define ES_FFT_KIND_ONLY 3   /* single-stage line */

/* In the generic kernels (any 11-smooth length, radix known at run time) the
 * pass functions are real device functions, one per (radix, kind), so that
 * each gets its own register allocation: inlined into one kernel the ~60 of
 * them share the worst case and spill (ptxas -v: 1 - 2.5 KB of spill stores
 * in the column kernels); -DES_FFT_PASS_INLINE restores that form. Kernels
 * instantiated for a fixed stage list (EsFftSeq below) inline their 3 - 4
 * passes. */
#if defined(__CUDACC__) && !defined(ES_FFT_PASS_INLINE)
#define ES_PASS_FN __device__ __noinline__
#else
#define ES_PASS_FN ES_HD
#endif

/* One pass of the line FFT over `1 << lshift` lines (pitch elements apart in
 * x) by nthreads workers, this one being tid.
 *  - the first processed pass takes its inputs from `in` instead of x, the
 *    last one hands its outputs to `out`; both are addressed through a cursor
 *    c = cursor(line, index0, index_stride), element k of the butterfly being
 *    index0 + k * index_stride (raw = fetch(c, k), value = finish(c, k, raw) /
 *    store(c, k, value)):
 *      natural-stride pass (stage 0):  index = j + k * stride0
 *      digit-reversed pass (last stage): index = t + k * (n / radix), where
 *      consecutive butterflies t own consecutive indices and bpos[t] is the
 *      padded position of their `radix` contiguous elements in x;
 *  - all other traffic is in place in x (no barrier inside a pass).
 * The caller synchronises between passes. Returns in.any_bits(). */
template<int R, int SGN, bool DIT, int KIND, typename T, typename T2,
        typename In, typename Out>
ES_HD unsigned int es_line_pass_body(T2* x, const EsFftPassArgs& pa,
        const T2* tw, const int32_t* bpos, const int lshift, const int pitch,
        const int tid, const int nthreads, In& in, Out& out)
{
    constexpr bool first_stage = (KIND == ES_FFT_KIND_FIRST ||
            KIND == ES_FFT_KIND_ONLY);
    constexpr bool last_stage = (KIND == ES_FFT_KIND_LAST ||
            KIND == ES_FFT_KIND_ONLY);
    /* processing order: DIF stage 0 first, DIT last stage first */
    constexpr bool from_in = DIT ? last_stage : first_stage;
    constexpr bool to_out = DIT ? first_stage : last_stage;
    const EsFftStage& st = pa.st;
    const int per_line = pa.n / R;
    const int total = per_line << lshift;
    const int lmask = (1 << lshift) - 1;
    for (int e = tid; e < total; e += nthreads)
    {
        const int line = e & lmask, ql = e >> lshift;
        T2* xl = x + (size_t) line * pitch;
        int j = 0, base, estride, gidx, gstride;
        if (last_stage)
        {
            /* stride 1: R contiguous elements, digit-reversed butterfly order */
            base = bpos[ql];
            estride = 1;
            gidx = ql;
            gstride = per_line;
        }
        else
        {
            int b;
            if (first_stage) { b = 0; j = ql; }
            else
            {
                b = (int) (((float) ql + 0.5f) * st.inv_stride);
                j = ql - b * st.stride;
            }
            const int pos = b * st.len + j;
            base = es_padded(pa, pos);
            /* stage 0 strides over the pads, later stages stay inside a block */
            estride = first_stage ? st.stride + pa.pad : st.stride;
            gidx = j; /* only used by stage 0 */
            gstride = st.stride;
        }
        if constexpr (to_out && Out::skip_masked_butterflies)
        {
            /* sparse output (row pass that stores a few occupied tiles):
             * nothing of this butterfly is kept */
            if (!out.template any_stored<R>(out.cursor(line, gidx, gstride)))
                continue;
        }
        T2 v[R];
        if constexpr (from_in)
        {
            /* two phases so that all R loads are in flight before the first
             * value is used */
            const typename In::Cursor c = in.cursor(line, gidx, gstride);
            typename In::Raw raw[R];
            in.butterfly_begin();
#pragma unroll
            for (int k = 0; k < R; ++k) raw[k] = in.fetch(c, k);
#pragma unroll
            for (int k = 0; k + 1 < R; k += 2)
                in.finish2(c, k, raw[k], raw[k + 1], v[k], v[k + 1]);
            if (R & 1) v[R - 1] = in.finish(c, R - 1, raw[R - 1]);
            if constexpr (In::skip_zero_butterflies && !to_out)
            {
                /* sparse lines (row pass over a few occupied tiles): a
                 * butterfly whose inputs are all zero stores zeros; runs of
                 * occupied cells are contiguous in the butterfly index, so
                 * warps mostly take one side */
                if (!in.butterfly_nonzero())
                {
                    T2 zero;
                    zero.x = 0; zero.y = 0;
#pragma unroll
                    for (int k = 0; k < R; ++k) xl[base + k * estride] = zero;
                    continue;
                }
            }
        }
        else
        {
#pragma unroll
            for (int k = 0; k < R; ++k) v[k] = xl[base + k * estride];
        }
        if constexpr (to_out)
        {
            /* the output functor may load per-output operands (the w-screen
             * table) before the butterfly arithmetic; large radices load
             * them afterwards to stay inside the register file */
            const typename Out::Cursor oc = out.cursor(line, gidx, gstride);
            constexpr bool early = (R <= ES_FFT_PREFETCH_MAX_RADIX);
            typename Out::Pre pre[early ? R : 1];
            if constexpr (early)
            {
#pragma unroll
                for (int k = 0; k < R; ++k) pre[k] = out.prefetch(oc, k);
            }
            es_bf_core<R, SGN, DIT, T>(v, st, tw, j);
            /* two outputs per call: the w-screen output functor evaluates
             * both phasors in one straight-line block (instruction-level
             * parallelism for its long dependent chains) */
#pragma unroll
            for (int k = 0; k + 1 < R; k += 2)
            {
                if constexpr (early)
                    out.store2(oc, k, v[es_out_slot<R>(k)],
                            v[es_out_slot<R>(k + 1)], pre[k], pre[k + 1]);
                else
                    out.store2(oc, k, v[es_out_slot<R>(k)],
                            v[es_out_slot<R>(k + 1)], out.prefetch(oc, k),
                            out.prefetch(oc, k + 1));
            }
            if (R & 1)
            {
                if constexpr (early)
                    out.store1(oc, R - 1, v[es_out_slot<R>(R - 1)], pre[R - 1]);
                else
                    out.store1(oc, R - 1, v[es_out_slot<R>(R - 1)],
                            out.prefetch(oc, R - 1));
            }
        }
        else
        {
            es_bf_core<R, SGN, DIT, T>(v, st, tw, j);
#pragma unroll
            for (int k = 0; k < R; ++k)
                xl[base + k * estride] = v[es_out_slot<R>(k)];
        }
    }
    return in.any_bits();
}

/* The same as a call (generic kernels): everything by value, x is known to
 * be shared memory. */
template<int R, int SGN, bool DIT, int KIND, typename T, typename T2,
        typename In, typename Out>
ES_PASS_FN unsigned int es_line_pass_call(T2* x, const EsFftPassArgs pa,
        const T2* tw, const int32_t* bpos, const int lshift, const int pitch,
        const int tid, const int nthreads, In in, Out out)
{
#if defined(__CUDA_ARCH__)
    __builtin_assume(__isShared(x));
#endif
    return es_line_pass_body<R, SGN, DIT, KIND, T>(x, pa, tw, bpos, lshift,
            pitch, tid, nthreads, in, out);
}

/* fp64 butterflies above radix 16 do not fit the register file and are
 * never planned (es_fft_line_plan): they are not instantiated. */
#define ES_FFT_MAX_RADIX_DOUBLE 16

template<int R, int SGN, bool DIT, int MAXR, typename T, typename T2,
        typename In, typename Out>
ES_HD unsigned int es_line_pass_r(T2* x, const EsFftPassArgs& pa,
        const int kind, const T2* tw, const int32_t* bpos, const int lshift,
        const int pitch, const int tid, const int nthreads, const In& in,
        const Out& out)
{
    if constexpr ((sizeof(T) == 4 || R <= ES_FFT_MAX_RADIX_DOUBLE) &&
            R <= MAXR)
    {
        switch (kind)
        {
        case ES_FFT_KIND_FIRST:
            return es_line_pass_call<R, SGN, DIT, ES_FFT_KIND_FIRST, T>(x, pa,
                    tw, bpos, lshift, pitch, tid, nthreads, in, out);
        case ES_FFT_KIND_MIDDLE:
            return es_line_pass_call<R, SGN, DIT, ES_FFT_KIND_MIDDLE, T>(x, pa,
                    tw, bpos, lshift, pitch, tid, nthreads, in, out);
        case ES_FFT_KIND_LAST:
            return es_line_pass_call<R, SGN, DIT, ES_FFT_KIND_LAST, T>(x, pa,
                    tw, bpos, lshift, pitch, tid, nthreads, in, out);
        default:
            return es_line_pass_call<R, SGN, DIT, ES_FFT_KIND_ONLY, T>(x, pa,
                    tw, bpos, lshift, pitch, tid, nthreads, in, out);
        }
    }
    return 0u;
}

/* Radices with a butterfly (ES_FFT_CASE is defined at the point of use). */
#ifndef ES_FFT_RADIX_CASES
#define ES_FFT_RADIX_CASES \
    ES_FFT_CASE(2) ES_FFT_CASE(3) ES_FFT_CASE(4) ES_FFT_CASE(5) \
    ES_FFT_CASE(6) ES_FFT_CASE(7) ES_FFT_CASE(8) ES_FFT_CASE(9) \
    ES_FFT_CASE(10) ES_FFT_CASE(11) ES_FFT_CASE(12) ES_FFT_CASE(14) \
    ES_FFT_CASE(15) ES_FFT_CASE(16) ES_FFT_CASE(20) ES_FFT_CASE(22) \
    ES_FFT_CASE(24) ES_FFT_CASE(28) ES_FFT_CASE(32)
#endif

/* step = 0 .. num_stages - 1 in processing order (DIF: stage = step, DIT:
 * stage = num_stages - 1 - step). MAXR: largest radix instantiated (kernel
 * variants with more threads and fewer registers per thread carry only the
 * small butterflies). */
template<int SGN, bool DIT, typename T, int MAXR = 32, typename T2, typename In,
        typename Out>
ES_HD void es_line_pass(T2* x, const EsFftLine& pl, const int step,
        const T2* tw, const int32_t* bpos, const int lshift, const int pitch,
        const int tid, const int nthreads, In& in, Out& out)
{
    const int stage = DIT ? pl.num_stages - 1 - step : step;
    const int kind = pl.num_stages == 1 ? ES_FFT_KIND_ONLY :
            stage == 0 ? ES_FFT_KIND_FIRST :
            stage == pl.num_stages - 1 ? ES_FFT_KIND_LAST : ES_FFT_KIND_MIDDLE;
    EsFftPassArgs pa;
    pa.st = pl.stage[stage];
    pa.n = pl.n;
    pa.pad = pl.pad;
    pa.inv_stride0 = pl.inv_stride0;
    unsigned int any = 0u;
#define ES_FFT_CASE(R) case R: any = es_line_pass_r<R, SGN, DIT, MAXR, T>(x, pa, \
        kind, tw, bpos, lshift, pitch, tid, nthreads, in, out); break;
    switch (pa.st.radix)
    {
    ES_FFT_RADIX_CASES
    default: break;
    }
#undef ES_FFT_CASE
    in.merge_any(any);
}

/* ---- fixed stage lists ----
 * A kernel instantiated for EsFftSeq<ID> (ID > 0) runs its passes with
 * compile-time radices and kinds: no radix switch, a handful of passes inlined
 * instead of ~60, registers allocated for exactly those (measured against the
 * generic kernels on a B200: -13 % on the 24640-point lines, -7 % on 6160).
 * The lists are the planner's (es_fft_line_plan) for the grid sizes of the
 * BASELINE configurations; es_fft.cu picks a specialised kernel set when the
 * plan's stage list equals one of them and the generic set otherwise. */
#define ES_FFT_THREADS 512 /* CTA size limit of the generic kernels (128 registers each) */
template<int ID> struct EsFftSeq
{
    static constexpr int n = 0;
    static constexpr int threads = ES_FFT_THREADS;
    static constexpr int radix(int) { return 0; }
};
/* THREADS_: CTA size limit of the set's kernels (ptxas keeps them within
 * 65536 / THREADS_ registers): lines that fill shared memory run one CTA per
 * SM, so every warp the register file can hold is one more to hide latency */
#define ES_FFT_SEQ(ID_, THREADS_, ...) \
    template<> struct EsFftSeq<ID_> \
    { \
        static constexpr int n = sizeof((int[]) {__VA_ARGS__}) / sizeof(int); \
        static constexpr int threads = THREADS_; \
        static constexpr int radix(int i) \
        { \
            constexpr int r[] = {__VA_ARGS__}; \
            return r[i]; \
        } \
    };
ES_FFT_SEQ(1, 512, 28, 20, 11)     /* 6160: C2 */
ES_FFT_SEQ(2, 640, 16, 14, 10, 11) /* 24640: C4, 96 registers */
ES_FFT_SEQ(3, 512, 32, 24, 15)     /* 11520: C5 */
ES_FFT_SEQ(4, 512, 15, 8, 8, 15)   /* 14400: C3 (fp64) */
/* Measured and not adopted (round 2, profiles/r02_logs/ab14_*): stage orders
 * that halve the shared-memory bank conflicts of the middle passes
 * (14.11.10.16, 24.15.32, 28.11.20: simulated wavefronts 1.17 / 1.08 / 1.00 x
 * ideal against 1.38 / 1.50 / 1.14 x for the lists above) speed the row passes
 * up by 3-9 % but put a large even radix last, where the gridding column pass
 * pays more for the w-screen operands of 16-32 outputs per butterfly than it
 * saves (C2 cols 9.8 -> 11.4 ms, C4 3.89 -> 4.05 s). */
#define ES_FFT_NUM_SEQ 4
/* (id, real type, is_double) of the kernel sets the Makefile builds (SEQ_IDS) */
#define ES_FFT_SEQ_TABLE(X) X(1, float, 0) X(2, float, 0) X(3, float, 0) \
    X(4, double, 1)

/* The line plan of a fixed stage list as compile-time constants: the same
 * arithmetic as es_fft_line_plan (es_fft_plan.c) -- lengths, strides, twiddle
 * table offsets, pad -- so that the passes of a specialised kernel address
 * shared memory with immediate offsets instead of per-element integer
 * arithmetic (profiles: LEA / IADD3 / IMAD / ISETP were ~30 % of the executed
 * instructions of the column kernels). es_fft.cu checks it against the
 * run-time plan before it selects a specialised kernel set. */
struct EsFftSeqPlan
{
    int32_t n, num_stages, pad, num_twiddles;
    int32_t radix[ES_FFT_MAX_STAGES], len[ES_FFT_MAX_STAGES];
    int32_t stride[ES_FFT_MAX_STAGES];
    int32_t tw_hi[ES_FFT_MAX_STAGES], tw_lo[ES_FFT_MAX_STAGES];
};

template<int ID> constexpr EsFftSeqPlan es_fft_seq_plan()
{
    EsFftSeqPlan p = {};
    p.num_stages = EsFftSeq<ID>::n;
    int32_t n = 1;
    for (int i = 0; i < EsFftSeq<ID>::n; ++i) n *= EsFftSeq<ID>::radix(i);
    p.n = n;
    int32_t len = n, tw = 0;
    for (int i = 0; i < EsFftSeq<ID>::n; ++i)
    {
        const int32_t r = EsFftSeq<ID>::radix(i);
        p.radix[i] = r;
        p.len[i] = len;
        p.stride[i] = len / r;
        p.tw_hi[i] = tw;
        if (p.stride[i] > ES_FFT_TW_LO)
            tw += (p.stride[i] + ES_FFT_TW_LO - 1) / ES_FFT_TW_LO;
        p.tw_lo[i] = tw;
        tw += p.stride[i] < ES_FFT_TW_LO ? p.stride[i] : ES_FFT_TW_LO;
        len = p.stride[i];
    }
    p.num_twiddles = tw;
    p.pad = (EsFftSeq<ID>::n > 1 && (p.stride[0] % 2 == 0)) ? 1 : 0;
    return p;
}

/* true when the run-time plan is the compile-time one */
template<int ID> inline bool es_fft_seq_matches(const EsFftLine& pl)
{
    constexpr EsFftSeqPlan p = es_fft_seq_plan<ID>();
    bool same = pl.n == p.n && pl.num_stages == p.num_stages &&
            pl.pad == p.pad && pl.num_twiddles == p.num_twiddles;
    for (int i = 0; same && i < p.num_stages; ++i)
        same = pl.stage[i].radix == p.radix[i] && pl.stage[i].len == p.len[i]
                && pl.stage[i].stride == p.stride[i] &&
                pl.stage[i].tw_hi == p.tw_hi[i] &&
                pl.stage[i].tw_lo == p.tw_lo[i];
    return same;
}

template<int SEQ, int STAGE, int SGN, bool DIT, typename T, typename T2,
        typename In, typename Out>
ES_HD void es_line_pass_fixed(T2* x, const T2* tw, const int32_t* bpos,
        const int lshift, const int pitch, const int tid, const int nthreads,
        In& in, Out& out)
{
    if constexpr (STAGE < EsFftSeq<SEQ>::n)
    {
        constexpr EsFftSeqPlan P = es_fft_seq_plan<SEQ>();
        constexpr int NS = P.num_stages;
        constexpr int R = P.radix[STAGE];
        constexpr int KIND = NS == 1 ? ES_FFT_KIND_ONLY :
                STAGE == 0 ? ES_FFT_KIND_FIRST :
                STAGE == NS - 1 ? ES_FFT_KIND_LAST : ES_FFT_KIND_MIDDLE;
        EsFftPassArgs pa;
        pa.st.radix = R;
        pa.st.len = P.len[STAGE];
        pa.st.stride = P.stride[STAGE];
        pa.st.tw_hi = P.tw_hi[STAGE];
        pa.st.tw_lo = P.tw_lo[STAGE];
        pa.st.inv_stride = 1.0f / (float) P.stride[STAGE];
        pa.n = P.n;
        pa.pad = P.pad;
        pa.inv_stride0 = 1.0f / (float) P.stride[0];
        in.merge_any(es_line_pass_body<R, SGN, DIT, KIND, T>(x, pa, tw, bpos,
                lshift, pitch, tid, nthreads, in, out));
    }
}

/* es_line_pass for a kernel instantiated with stage list SEQ (0: generic). */
template<int SEQ, int SGN, bool DIT, typename T, typename T2, typename In,
        typename Out>
ES_HD void es_line_pass_seq(T2* x, const EsFftLine& pl, const int step,
        const T2* tw, const int32_t* bpos, const int lshift, const int pitch,
        const int tid, const int nthreads, In& in, Out& out)
{
    if constexpr (SEQ == 0)
    {
        es_line_pass<SGN, DIT, T>(x, pl, step, tw, bpos, lshift, pitch, tid,
                nthreads, in, out);
    }
    else
    {
        const int stage = DIT ? EsFftSeq<SEQ>::n - 1 - step : step;
        switch (stage)
        {
        case 0: es_line_pass_fixed<SEQ, 0, SGN, DIT, T>(x, tw, bpos,
                lshift, pitch, tid, nthreads, in, out); break;
        case 1: es_line_pass_fixed<SEQ, 1, SGN, DIT, T>(x, tw, bpos,
                lshift, pitch, tid, nthreads, in, out); break;
        case 2: es_line_pass_fixed<SEQ, 2, SGN, DIT, T>(x, tw, bpos,
                lshift, pitch, tid, nthreads, in, out); break;
        case 3: es_line_pass_fixed<SEQ, 3, SGN, DIT, T>(x, tw, bpos,
                lshift, pitch, tid, nthreads, in, out); break;
        case 4: es_line_pass_fixed<SEQ, 4, SGN, DIT, T>(x, tw, bpos,
                lshift, pitch, tid, nthreads, in, out); break;
        default: break;
        }
    }
}

#endif
