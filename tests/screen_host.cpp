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
    printf("%.3e %.3e %.3e\n", max_ph, max_inv, max_sc);
    return 0;
}
