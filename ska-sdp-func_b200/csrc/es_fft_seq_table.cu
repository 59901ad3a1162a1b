/*
 * Registry of the kernel sets built for fixed stage lists (es_fft_seq.cu,
 * one object per entry of ES_FFT_SEQ_TABLE in es_fft_core.h).
 */
#include "es_fft_kernels.cuh"

#define X(ID, REAL, DBL) const EsFftOps* es_fft_seq_ops_##ID(void);
ES_FFT_SEQ_TABLE(X)
#undef X

/* 1 when kernel set `id` was built for this precision and exactly this line
 * plan (stage list, strides, twiddle offsets, pad). */
int es_fft_seq_match(int id, const EsFftLine* line, int is_double)
{
    switch (id)
    {
#define X(ID, REAL, DBL) case ID: \
        return (DBL == is_double) && es_fft_seq_matches<ID>(*line);
    ES_FFT_SEQ_TABLE(X)
#undef X
    default: return 0;
    }
}

const EsFftOps* es_fft_seq_ops(int id)
{
    switch (id)
    {
#define X(ID, REAL, DBL) case ID: return es_fft_seq_ops_##ID();
    ES_FFT_SEQ_TABLE(X)
#undef X
    default: return 0;
    }
}
