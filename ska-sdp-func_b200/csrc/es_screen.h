/*
 * w-screen phasor of the w-stacking gridder,
 *     exp(sign 2 pi i w (n - 1)) / n,   n = sqrt(1 - l^2 - m^2),
 * (reference: phase_shift, sdp_gridder_uvw_es_fft_kernels.cu:110-123, used by
 * apply_w_screen_and_sum :429-547 and reverse_w_screen_to_stack :554-679),
 * evaluated once per (pixel, w-plane) inside the last / first pass of the
 * column FFT kernels (es_fft.cu).
 *
 * The phase w (n - 1) reaches 1e2 .. 1e4 turns, so n - 1 needs ~1e-13: it is
 * formed in fp64 in both precisions. For fp32 data everything that would go
 * through the 16-lane conversion / special-function unit (rint, sincospi's
 * float -> int) is done with integer and FMA-pipe instructions instead; the
 * int / fp32 <-> fp64 conversions exist in both forms (see es_screen_f32):
 *   1/n   = fp32 rsqrt seed + one Newton step in fp64       (rel. err ~2e-14)
 *   n - 1 = (1 - l^2 - m^2) / n - 1
 *   t     = w (n - 1) - round(w (n - 1))  (magic-constant rounding, fp64)
 *   cos, sin(2 pi t): device: special-function unit (abs. err ~4e-7);
 *                     host / -DES_SCREEN_POLY_SINCOS: quadrant by
 *                     magic-constant rounding in fp32, degree 8 / 9
 *                     polynomials on [-pi/4, pi/4]          (abs. err ~1e-7)
 * Host-callable so that tests/screen_host.cpp can check it against libm.
 */
#ifndef ES_SCREEN_H_
#define ES_SCREEN_H_

#include <math.h>
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define ES_SCR_HD __host__ __device__ __forceinline__
#else
#define ES_SCR_HD inline
#endif

ES_SCR_HD double es_hilo_to_double(const uint32_t hi, const uint32_t lo)
{
#if defined(__CUDA_ARCH__)
    return __hiloint2double((int) hi, (int) lo);
#else
    const uint64_t b = ((uint64_t) hi << 32) | lo;
    double d;
    memcpy(&d, &b, 8);
    return d;
#endif
}

ES_SCR_HD void es_double_to_hilo(const double d, uint32_t& hi, uint32_t& lo)
{
#if defined(__CUDA_ARCH__)
    hi = (uint32_t) __double2hiint(d);
    lo = (uint32_t) __double2loint(d);
#else
    uint64_t b;
    memcpy(&b, &d, 8);
    hi = (uint32_t) (b >> 32);
    lo = (uint32_t) b;
#endif
}

ES_SCR_HD uint32_t es_float_bits(const float f)
{
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    uint32_t b;
    memcpy(&b, &f, 4);
    return b;
#endif
}

ES_SCR_HD float es_bits_float(const uint32_t b)
{
#if defined(__CUDA_ARCH__)
    return __uint_as_float(b);
#else
    float f;
    memcpy(&f, &b, 4);
    return f;
#endif
}

/* exact, 0 <= s < 2^31 */
ES_SCR_HD double es_uint_to_double(const uint32_t s)
{
    return es_hilo_to_double(0x43300000u, s) - 4503599627370496.0;
}

/* exact, normal f */
ES_SCR_HD double es_float_to_double(const float f)
{
    const uint32_t b = es_float_bits(f);
    const uint32_t hi = (b & 0x80000000u) | (((b & 0x7fffffffu) >> 3) + 0x38000000u);
    return es_hilo_to_double(hi, b << 29);
}

/* truncating, |d| < 2^127; |d| < 2^-126 -> 0 */
ES_SCR_HD float es_double_to_float(const double d)
{
    uint32_t hi, lo;
    es_double_to_hilo(d, hi, lo);
    const uint32_t mag = hi & 0x7fffffffu;
    if (mag < 0x38100000u) return 0.0f;
    return es_bits_float((hi & 0x80000000u) | ((mag - 0x38000000u) << 3) | (lo >> 29));
}

/* cos, sin (2 pi t) for |t| <= 0.5 (slightly beyond is fine) */
ES_SCR_HD void es_sincos_turns(const float t, float& c, float& s)
{
    /* quarter turns: q = round(4 t), r = t - q / 4 in [-1/8, 1/8] */
    const float magic = 12582912.0f; /* 1.5 * 2^23 */
    const float qf = 4.0f * t + magic;
    const uint32_t q = es_float_bits(qf);
    const float r = t - 0.25f * (qf - magic);
    const float z = 6.28318530717958647692f * r;
    const float z2 = z * z;
    /* Taylor / near-minimax on |z| <= pi/4 */
    float sp = 2.7557313707e-6f;                 /* z^9 / 9! */
    sp = sp * z2 - 1.9841270114e-4f;
    sp = sp * z2 + 8.3333337680e-3f;
    sp = sp * z2 - 1.6666667163e-1f;
    const float sz = z + z * z2 * sp;
    float cp = 2.4801587642e-5f;                 /* z^8 / 8! */
    cp = cp * z2 - 1.3888889225e-3f;
    cp = cp * z2 + 4.1666667908e-2f;
    cp = cp * z2 - 0.5f;
    const float cz = 1.0f + z2 * cp;
    /* rotate by q quarter turns */
    const float a = (q & 1u) ? sz : cz;
    const float b = (q & 1u) ? cz : sz;
    c = ((q + 1u) & 2u) ? -a : a;
    s = (q & 2u) ? -b : b;
}

/* fp32 data: re + i im = exp(2 pi i w2 (n - 1) / 2) with w2 = 2 sign w folded
 * by the caller into `w_signed` = sign w; inv_n = 1 / n.
 * s = iu^2 + jv^2 (exact integer), pix2 = pixel_size^2. */
ES_SCR_HD void es_screen_f32(const uint32_t s, const double pix2,
        const float pix2f, const double w_signed, float& re, float& im,
        float& inv_n)
{
#if !defined(ES_SCREEN_INTEGER_CVT) && defined(__CUDA_ARCH__)
    /* device default: hardware conversions (quarter-rate unit, but one issue
     * slot each; with two phasors evaluated side by side they measured 2-3 %
     * faster in the column kernels than the integer forms below, which the
     * host build and -DES_SCREEN_INTEGER_CVT keep) */
    const double a = 1.0 - pix2 * (double) s;
    const float af = 1.0f - pix2f * (float) s;
    const float y0 = rsqrtf(af);
    const double yd = (double) y0;
    const double h = a * yd;
    const double e = 1.0 - h * yd;
    const double y = yd + (0.5 * yd) * e;
    const double nm1 = a * y - 1.0;
    double t = w_signed * nm1;
    const double big = 6755399441055744.0;
    t -= (t + big) - big;
#if !defined(ES_SCREEN_POLY_SINCOS)
    /* special-function unit: abs. error ~4e-7 on [-pi, pi], two MUFU + two
     * multiplies instead of ~22 FMA-pipe instructions; the fp32 parity errors
     * (2-5e-6, dominated by fp32 accumulation) move by < 2e-7
     * (profiles/parity_errors.py), the column kernels gain 5 % */
    const float z = 6.28318530717958647692f * __double2float_rz(t);
    re = __cosf(z);
    im = __sinf(z);
#else
    es_sincos_turns(__double2float_rz(t), re, im);
#endif
    inv_n = __double2float_rz(y);
#else
    const double a = 1.0 - pix2 * es_uint_to_double(s);
    const float af = 1.0f - pix2f * (float) s;
#if defined(__CUDA_ARCH__)
    const float y0 = rsqrtf(af);
#else
    const float y0 = 1.0f / sqrtf(af);
#endif
    const double yd = es_float_to_double(y0);
    const double h = a * yd;
    const double e = 1.0 - h * yd;           /* 1 - a y0^2 */
    const double y = yd + (0.5 * yd) * e;    /* 1 / n */
    const double nm1 = a * y - 1.0;          /* n - 1 */
    double t = w_signed * nm1;
    const double big = 6755399441055744.0;   /* 1.5 * 2^52 */
    t -= (t + big) - big;
    es_sincos_turns(es_double_to_float(t), re, im);
    inv_n = es_double_to_float(y);
#endif
}

/* ---- table form (es_fft.cu): n - 1 per image pixel comes from a table that
 * the plan fills once in full fp64 (it depends on |iu|, |jv| and the pixel
 * size only), 1 / n and the (-1)^(i+j) shift are applied once per call by the
 * image transposes, so that per (pixel, w-plane) only the phasor remains.
 *
 * Phase reduction without a rounding sequence or a float conversion: with
 * magic = 1.5 * 2^(52 - K) the sum r = w (n - 1) + magic (ONE fused
 * multiply-add, exact product) has ulp 2^-K, so the low word of r holds
 * frac(w (n - 1)) in K-bit fixed point, modulo 2^32 -- for K = 32 the whole
 * low word, valid while |w (n - 1)| < 2^19 turns; K = 24 reaches 2^27 turns
 * at 3e-8 turns resolution (es_screen_fixed_bits). `shift` = 32 - K moves the
 * fraction to the top of the word: a signed 32-bit fraction of a turn in
 * [-0.5, 0.5), error <= 2^-(K+1) turns. */
ES_SCR_HD int es_screen_fixed_bits(const double max_abs_turns)
{
    return max_abs_turns < 262144.0 ? 32 : 24;
}

ES_SCR_HD double es_screen_magic(const int fixed_bits)
{
    /* 1.5 * 2^(52 - K) */
    return fixed_bits == 32 ? 1572864.0 : 402653184.0;
}

/* cos, sin of 2 pi frac(w_signed nm1), fp32 data */
ES_SCR_HD void es_phasor_f32(const double nm1, const double w_signed,
        const double magic, const int shift, float& re, float& im)
{
#if defined(__CUDA_ARCH__)
    const double r = __fma_rn(w_signed, nm1, magic);
    const int f = __double2loint(r) << shift;
    /* 2 pi / 2^32; the special-function unit (abs. error ~4e-7) as before */
    const float z = (float) f * 1.46291807926715968e-9f;
    re = __cosf(z);
    im = __sinf(z);
#else
    const double r = fma(w_signed, nm1, magic);
    uint32_t hi, lo;
    es_double_to_hilo(r, hi, lo);
    const int32_t f = (int32_t) (lo << shift);
    const float z = (float) f * 1.46291807926715968e-9f;
    re = cosf(z);
    im = sinf(z);
#endif
}

/* n - 1 = -s / (sqrt(1 - s) + 1), s = l^2 + m^2 (kernels.cu:113-114), fp64 */
ES_SCR_HD double es_screen_nm1(const uint32_t s2, const double pix2)
{
    const double sos = pix2 * (double) s2;
    return (-sos) / (sqrt(1.0 - sos) + 1.0);
}

#endif
