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

template<int R> struct EsOddTab;
template<> struct EsOddTab<3>
{
    ES_HD static double c(int k)
    {
        const double t[3] = {1.0, -0.5, -0.5};
        return t[k];
    }
    ES_HD static double s(int k)
    {
        const double t[3] = {0.0, 0.86602540378443865, -0.86602540378443865};
        return t[k];
    }
};
template<> struct EsOddTab<5>
{
    ES_HD static double c(int k)
    {
        const double t[5] = {1.0, 0.30901699437494745, -0.80901699437494745,
            -0.80901699437494745, 0.30901699437494745};
        return t[k];
    }
    ES_HD static double s(int k)
    {
        const double t[5] = {0.0, 0.95105651629515353, 0.58778525229247314,
            -0.58778525229247314, -0.95105651629515353};
        return t[k];
    }
};
template<> struct EsOddTab<7>
{
    ES_HD static double c(int k)
    {
        const double t[7] = {1.0, 0.62348980185873348, -0.22252093395631439,
            -0.90096886790241915, -0.90096886790241915, -0.22252093395631439,
            0.62348980185873348};
        return t[k];
    }
    ES_HD static double s(int k)
    {
        const double t[7] = {0.0, 0.7818314824680298, 0.97492791218182362,
            0.43388373911755812, -0.43388373911755812, -0.97492791218182362,
            -0.7818314824680298};
        return t[k];
    }
};
template<> struct EsOddTab<11>
{
    ES_HD static double c(int k)
    {
        const double t[11] = {1.0, 0.84125353283118121, 0.41541501300188644,
            -0.14231483827328512, -0.65486073394528510, -0.95949297361449737,
            -0.95949297361449737, -0.65486073394528510, -0.14231483827328512,
            0.41541501300188644, 0.84125353283118121};
        return t[k];
    }
    ES_HD static double s(int k)
    {
        const double t[11] = {0.0, 0.54064081745559756, 0.90963199535451833,
            0.98982144188093268, 0.75574957435425827, 0.28173255684142967,
            -0.28173255684142967, -0.75574957435425827, -0.98982144188093268,
            -0.90963199535451833, -0.54064081745559756};
        return t[k];
    }
};

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
            const T cs = (T) EsOddTab<R>::c((k * m) % R);
            const T sn = (T) EsOddTab<R>::s((k * m) % R);
            c.x += a[m - 1].x * cs; c.y += a[m - 1].y * cs;
            d.x += b[m - 1].x * sn; d.y += b[m - 1].y * sn;
        }
        const T2 id = es_mul_i<SGN>(d);
        v[k] = es_cadd(c, id);
        v[R - k] = es_csub(c, id);
    }
}

/* Radix-R DFT, output k left in v[es_out_slot<R>(k)]. */
template<int R> ES_HD int es_out_slot(int k)
{
    return (R == 2 || R == 4 || R == 8 || R == 16) ? es_bitrev<R>(k) : k;
}

template<int R, int SGN, typename T, typename T2>
ES_HD void es_dft(T2 (&v)[R])
{
    if constexpr (R == 2 || R == 4 || R == 8 || R == 16)
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

/* Butterfly `ql` (< n / R) of one radix-R pass over the line at `x`. */
template<int R, int SGN, bool DIT, typename T, typename T2>
ES_HD void es_butterfly(T2* x, const EsFftStage& st, const T2* tw, const int ql)
{
    int b, j;
    if (st.stride == 1) { b = ql; j = 0; }
    else
    {
        b = (int) (((float) ql + 0.5f) * st.inv_stride);
        j = ql - b * st.stride;
    }
    T2* p = x + (size_t) b * st.len + j;
    const int s = st.stride;
    T2 v[R];
    T2 w[R]; /* w[k] = w_L^(j k), k >= 1; built in the order of use so that
              * only w[1] and the upper half of the computed powers stay live */
    if (s > 1) w[1] = es_twiddle<SGN>(st, tw, j);
    v[0] = p[0];
#pragma unroll
    for (int k = 1; k < R; ++k)
    {
        v[k] = p[k * s];
        if (DIT && s > 1)
        {
            if (k > 1)
                w[k] = (k & 1) ? es_cmul(w[k - 1], w[1]) :
                        es_cmul(w[k / 2], w[k / 2]);
            v[k] = es_cmul(v[k], w[k]);
        }
    }
    es_dft<R, SGN, T>(v);
    p[0] = v[es_out_slot<R>(0)];
#pragma unroll
    for (int k = 1; k < R; ++k)
    {
        T2 out = v[es_out_slot<R>(k)];
        if (!DIT && s > 1)
        {
            if (k > 1)
                w[k] = (k & 1) ? es_cmul(w[k - 1], w[1]) :
                        es_cmul(w[k / 2], w[k / 2]);
            out = es_cmul(out, w[k]);
        }
        p[k * s] = out;
    }
}

/* One radix-R pass over `nlines` lines (pitch elements apart) by `nthreads`
 * workers, this one being `tid`. Caller synchronises between passes. */
template<int R, int SGN, bool DIT, typename T, typename T2>
ES_HD void es_fft_pass_r(T2* x, const EsFftStage& st, const int n,
        const T2* tw, const int nlines, const int pitch, const int tid,
        const int nthreads)
{
    const int per_line = n / R;
    const int total = per_line * nlines;
    for (int q = tid; q < total; q += nthreads)
    {
        int line = 0, ql = q;
        while (ql >= per_line) { ql -= per_line; ++line; }
        es_butterfly<R, SGN, DIT, T>(x + (size_t) line * pitch, st, tw, ql);
    }
}

template<int SGN, bool DIT, typename T, typename T2>
ES_HD void es_fft_pass(T2* x, const EsFftLine& pl, const int stage,
        const T2* tw, const int nlines, const int pitch, const int tid,
        const int nthreads)
{
    const EsFftStage& st = pl.stage[stage];
#define ES_FFT_CASE(R) case R: es_fft_pass_r<R, SGN, DIT, T>(x, st, pl.n, tw, \
        nlines, pitch, tid, nthreads); break;
    switch (st.radix)
    {
    ES_FFT_CASE(16) ES_FFT_CASE(8) ES_FFT_CASE(4) ES_FFT_CASE(2)
    ES_FFT_CASE(3) ES_FFT_CASE(5) ES_FFT_CASE(7) ES_FFT_CASE(11)
    default: break;
    }
#undef ES_FFT_CASE
}

#endif
