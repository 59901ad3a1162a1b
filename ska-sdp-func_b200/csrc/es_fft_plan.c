/* See es_fft_plan.h. */
#include "es_fft_plan.h"

#include <math.h>
#include <string.h>

#define ES_TWO_PI 6.283185307179586476925286766559

/* radices es_fft_core.h has butterflies for (ES_FFT_RADIX_CASES) */
static const int32_t es_radices[] = {
    32, 28, 24, 22, 20, 16, 15, 14, 12, 11, 10, 9, 8, 7, 6, 5, 4, 3, 2
};
#define ES_NUM_RADICES ((int) (sizeof(es_radices) / sizeof(es_radices[0])))

typedef struct EsSearch
{
    int32_t max_radix;
    int depth;                            /* stages wanted */
    int32_t cur[ES_FFT_MAX_STAGES];
    int32_t best[ES_FFT_MAX_STAGES];
    long best_score;
    int found;
} EsSearch;

/* Lower is better: an odd radix in the set (it becomes the last pass), then
 * the smallest largest radix (register pressure), then the largest smallest
 * radix (balance). */
static long score_of(const int32_t* r, int count)
{
    int32_t mx = 0, mn = 1 << 30;
    int has_odd = 0;
    for (int i = 0; i < count; ++i)
    {
        if (r[i] > mx) mx = r[i];
        if (r[i] < mn) mn = r[i];
        if (r[i] & 1) has_odd = 1;
    }
    return (has_odd ? 0 : 100000) + 100L * mx - mn;
}

/* non-increasing sequences of allowed radices with product n */
static void search(EsSearch* s, int32_t rest, int level, int first)
{
    if (level == s->depth)
    {
        if (rest != 1) return;
        const long sc = score_of(s->cur, s->depth);
        if (!s->found || sc < s->best_score)
        {
            s->found = 1;
            s->best_score = sc;
            memcpy(s->best, s->cur, sizeof(s->cur));
        }
        return;
    }
    for (int k = first; k < ES_NUM_RADICES; ++k)
    {
        const int32_t r = es_radices[k];
        if (r > s->max_radix || rest % r) continue;
        s->cur[level] = r;
        search(s, rest / r, level + 1, k);
    }
}

static int radix_supported(int32_t r)
{
    for (int k = 0; k < ES_NUM_RADICES; ++k)
        if (es_radices[k] == r) return 1;
    return 0;
}

int es_fft_line_plan(int32_t n, int32_t max_radix, const int32_t* forced,
        EsFftLine* line)
{
    memset(line, 0, sizeof(*line));
    if (n < 2) return 1;
    int32_t radices[ES_FFT_MAX_STAGES];
    int count = 0;
    if (forced && forced[0] > 0)
    {
        int32_t prod = 1;
        while (forced[count] > 0)
        {
            if (count >= ES_FFT_MAX_STAGES || !radix_supported(forced[count]))
                return 1;
            prod *= forced[count];
            radices[count] = forced[count];
            count++;
        }
        if (prod != n) return 1;
    }
    else
    {
        EsSearch s;
        memset(&s, 0, sizeof(s));
        s.max_radix = max_radix;
        for (s.depth = 1; s.depth <= ES_FFT_MAX_STAGES && !s.found; ++s.depth)
            search(&s, n, 0, 0);
        if (!s.found) return 1;
        count = s.depth - 1;
        /* largest radix first (the pass that talks to global memory keeps
         * the most loads in flight), the largest odd radix last */
        memcpy(radices, s.best, sizeof(int32_t) * (size_t) count);
        int odd_at = -1;
        for (int i = 0; i < count; ++i)
            if ((radices[i] & 1) && odd_at < 0) odd_at = i;
        if (odd_at >= 0 && count > 1)
        {
            const int32_t r = radices[odd_at];
            for (int i = odd_at; i + 1 < count; ++i) radices[i] = radices[i + 1];
            radices[count - 1] = r;
        }
    }
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
    /* one pad element per block of stride0 elements when that makes the
     * distance between the blocks odd */
    const int32_t s0 = line->stage[0].stride;
    line->pad = (count > 1 && (s0 % 2 == 0)) ? 1 : 0;
    line->inv_stride0 = 1.0f / (float) s0;
    line->padded_n = n + line->pad * line->stage[0].radix;
    if (line->pad)
    {
        for (int32_t pos = 0; pos < n; ++pos)
        {
            const int b = (int) (((float) pos + 0.5f) * line->inv_stride0);
            if (b != pos / s0) return 1;
        }
    }
    return 0;
}

void es_fft_line_tables(const EsFftLine* line, double* tw, int32_t* perm,
        int32_t* bpos)
{
    for (int i = 0; i < line->num_stages && tw; ++i)
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
    const int m = line->num_stages;
    if (perm)
    {
        for (int32_t k = 0; k < line->n; ++k)
        {
            int32_t rest = k, pos = 0;
            for (int i = 0; i < m; ++i)
            {
                const EsFftStage* st = &line->stage[i];
                pos += (rest % st->radix) * st->stride;
                rest /= st->radix;
            }
            perm[k] = pos;
        }
    }
    if (bpos)
    {
        const int32_t nb = line->n / line->stage[m - 1].radix;
        for (int32_t t = 0; t < nb; ++t)
        {
            /* t = k_0 + r_0 (k_1 + r_1 (...)) over the first m - 1 radices */
            int32_t rest = t, pos = 0;
            for (int i = 0; i + 1 < m; ++i)
            {
                const EsFftStage* st = &line->stage[i];
                pos += (rest % st->radix) * st->stride;
                rest /= st->radix;
            }
            bpos[t] = pos + line->pad * (pos / line->stage[0].stride);
        }
    }
}
