/* Plain-C descriptors of the line FFT (see es_fft_core.h). */
#ifndef ES_FFT_CORE_TYPES_H_
#define ES_FFT_CORE_TYPES_H_

#include <stdint.h>

#define ES_FFT_MAX_STAGES 12
#define ES_FFT_TW_LO 64 /* entries of the low twiddle table */

/* One radix pass. */
typedef struct EsFftStage
{
    int32_t radix;    /* r */
    int32_t len;      /* L: sub-transform length entering this stage (DIF order) */
    int32_t stride;   /* s = L / r */
    int32_t tw_hi;    /* offset of hi table (ceil(s / 64) entries) in the twiddle array */
    int32_t tw_lo;    /* offset of lo table (min(s, 64) entries) */
    float inv_stride; /* 1 / s for the butterfly-index split */
} EsFftStage;

typedef struct EsFftLine
{
    int32_t n;
    int32_t num_stages;
    int32_t num_twiddles; /* total entries of all tables */
    int32_t pad;          /* pad elements after every stage[0].stride elements */
    float inv_stride0;    /* 1 / stage[0].stride */
    int32_t padded_n;     /* n + pad * stage[0].radix */
    EsFftStage stage[ES_FFT_MAX_STAGES];
} EsFftLine;

#endif
