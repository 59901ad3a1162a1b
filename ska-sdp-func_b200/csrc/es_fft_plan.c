/* See es_fft_plan.h. */
#include "es_fft_plan.h"

#include <math.h>
#include <string.h>

#define ES_TWO_PI 6.283185307179586476925286766559

int es_fft_line_plan(int32_t n, int32_t max_radix, EsFftLine* line)
{
    memset(line, 0, sizeof(*line));
    if (n < 2) return 1;
    int32_t radices[ES_FFT_MAX_STAGES + 8];
    int count = 0;
    int32_t m = n;
    int twos = 0;
    while (m % 2 == 0) { m /= 2; twos++; }
    /* powers of two: as few passes as possible, bits spread evenly */
    const int max_bits = (max_radix >= 16) ? 4 : 3;
    const int pow2_stages = (twos + max_bits - 1) / max_bits;
    for (int k = 0; k < pow2_stages; ++k)
    {
        const int bits = twos / pow2_stages + (k < twos % pow2_stages ? 1 : 0);
        radices[count++] = 1 << bits;
    }
    /* odd primes, smallest first: the last pass (stride 1) needs no twiddles
     * and odd strides keep the late passes free of bank conflicts */
    const int32_t odd[4] = {3, 5, 7, 11};
    for (int k = 0; k < 4; ++k)
    {
        while (m % odd[k] == 0)
        {
            if (count >= ES_FFT_MAX_STAGES) return 1;
            radices[count++] = odd[k];
            m /= odd[k];
        }
    }
    if (m != 1 || count > ES_FFT_MAX_STAGES) return 1;
    line->n = n;
    line->num_stages = count;
    int32_t len = n, tw = 0;
    for (int i = 0; i < count; ++i)
    {
        EsFftStage* st = &line->stage[i];
        st->radix = radices[i];
        st->len = len;
        st->stride = len / radices[i];
        st->inv_stride = 1.0f / (float) st->stride;
        st->tw_hi = tw;
        if (st->stride > ES_FFT_TW_LO)
            tw += (st->stride + ES_FFT_TW_LO - 1) / ES_FFT_TW_LO;
        st->tw_lo = tw;
        tw += st->stride < ES_FFT_TW_LO ? st->stride : ES_FFT_TW_LO;
        /* the kernels split a butterfly index with one float multiply */
        if (st->stride > 1)
        {
            for (int32_t q = 0; q < n / st->radix; ++q)
            {
                const int b = (int) (((float) q + 0.5f) * st->inv_stride);
                if (b != q / st->stride) return 1;
            }
        }
        len = st->stride;
    }
    line->num_twiddles = tw;
    return 0;
}

void es_fft_line_tables(const EsFftLine* line, double* tw, int32_t* perm)
{
    for (int i = 0; i < line->num_stages; ++i)
    {
        const EsFftStage* st = &line->stage[i];
        const double step = ES_TWO_PI / (double) st->len;
        if (st->stride > ES_FFT_TW_LO)
        {
            const int nhi = (st->stride + ES_FFT_TW_LO - 1) / ES_FFT_TW_LO;
            for (int h = 0; h < nhi; ++h)
            {
                tw[2 * (st->tw_hi + h)] = cos(step * (double) (h * ES_FFT_TW_LO));
                tw[2 * (st->tw_hi + h) + 1] = sin(step * (double) (h * ES_FFT_TW_LO));
            }
        }
        const int nlo = st->stride < ES_FFT_TW_LO ? st->stride : ES_FFT_TW_LO;
        for (int l = 0; l < nlo; ++l)
        {
            tw[2 * (st->tw_lo + l)] = cos(step * (double) l);
            tw[2 * (st->tw_lo + l) + 1] = sin(step * (double) l);
        }
    }
    if (!perm) return;
    for (int32_t k = 0; k < line->n; ++k)
    {
        int32_t rest = k, pos = 0;
        for (int i = 0; i < line->num_stages; ++i)
        {
            const EsFftStage* st = &line->stage[i];
            pos += (rest % st->radix) * st->stride;
            rest /= st->radix;
        }
        perm[k] = pos;
    }
}
