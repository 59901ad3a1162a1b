// Host check of csrc/es_screen.h (the conversion-free fp32 w-screen phasor)
// against libm in double precision. Prints the maximum absolute error of
// (re, im) and the maximum relative error of 1/n over a sweep of fields of
// view, pixel offsets and w values.
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include "es_screen.h"
int main()
{
    double max_ph = 0.0, max_inv = 0.0, max_sc = 0.0;
    srand(7);
    const double pixs[] = {1e-6, 8.522e-6, 4.26e-5, 1.3635e-4, 3e-4};
    for (double pix : pixs)
    {
        for (int trial = 0; trial < 200000; ++trial)
        {
            const int half = 8192;
            const int iu = rand() % (2 * half) - half, jv = rand() % (2 * half) - half;
            const uint32_t s = (uint32_t) (iu * iu + jv * jv);
            const double pix2 = pix * pix;
            if (pix2 * s > 0.9) continue;
            const double w = (rand() / (double) RAND_MAX - 0.5) * 2e5;
            float re, im, inv_n;
            es_screen_f32(s, pix2, (float) pix2, w, re, im, inv_n);
            const double n = sqrt(1.0 - pix2 * s);
            const double nm1 = -(pix2 * s) / (n + 1.0);
            const double ph = 2.0 * M_PI * fmod(w * nm1, 1.0);
            const double er = fabs(re - cos(ph)), ei = fabs(im - sin(ph));
            if (er > max_ph) max_ph = er;
            if (ei > max_ph) max_ph = ei;
            const double ev = fabs(inv_n * n - 1.0);
            if (ev > max_inv) max_inv = ev;
        }
    }
    for (int i = -600000; i <= 600000; ++i)
    {
        const float t = (float) i * 1e-6f;
        float c, s;
        es_sincos_turns(t, c, s);
        const double e1 = fabs(c - cos(2.0 * M_PI * (double) t));
        const double e2 = fabs(s - sin(2.0 * M_PI * (double) t));
        if (e1 > max_sc) max_sc = e1;
        if (e2 > max_sc) max_sc = e2;
    }
    // table form: n - 1 from fp64, phase reduced in fixed point (K = 32 / 24)
    double max_fx[2] = {0.0, 0.0};
    for (int kb = 0; kb < 2; ++kb)
    {
        const int bits = kb == 0 ? 32 : 24;
        const double wmax = kb == 0 ? 2.5e5 : 6.0e7; // turns reach 2^18 / 2^26
        for (int trial = 0; trial < 400000; ++trial)
        {
            const int iu = rand() % 8193, jv = rand() % 8193;
            const double pix2 = 4.26e-5 * 4.26e-5;
            const double nm1 = es_screen_nm1((uint32_t) (iu * iu + jv * jv), pix2);
            const double w = (rand() / (double) RAND_MAX - 0.5) * 2.0 * wmax / 0.13;
            if (fabs(w * nm1) >= (kb == 0 ? 262144.0 : 67108864.0)) continue;
            if (es_screen_fixed_bits(fabs(w * nm1)) != bits && kb == 0) continue;
            float re, im;
            es_phasor_f32(nm1, w, es_screen_magic(bits), 32 - bits, re, im);
            const long double t = (long double) w * (long double) nm1;
            const long double fr = t - floorl(t);
            const double ph = 2.0 * M_PI * (double) fr;
            const double er = fabs(re - cos(ph)), ei = fabs(im - sin(ph));
            if (er > max_fx[kb]) max_fx[kb] = er;
            if (ei > max_fx[kb]) max_fx[kb] = ei;
        }
    }
    printf("%.3e %.3e %.3e %.3e %.3e\n", max_ph, max_inv, max_sc, max_fx[0],
            max_fx[1]);
    return 0;
}
