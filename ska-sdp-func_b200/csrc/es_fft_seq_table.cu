/*
 * Registry of the kernel sets built for fixed stage lists (es_fft_seq.cu,
 * one object per entry of ES_FFT_SEQ_TABLE in es_fft_core.h).
 */
#include "es_fft_kernels.cuh"

#define X(ID, REAL, DBL) const EsFftOps* es_fft_seq_ops_##ID(void);
ES_FFT_SEQ_TABLE(X)
#undef X

template<int ID> static int seq_radices(int32_t* radices)
{
    for (int i = 0; i < EsFftSeq<ID>::n; ++i) radices[i] = EsFftSeq<ID>::radix(i);
    return EsFftSeq<ID>::n;
}

int es_fft_seq_info(int id, int32_t* radices, int* is_double)
{
    *is_double = -1;
    switch (id)
    {
#define X(ID, REAL, DBL) case ID: *is_double = DBL; return seq_radices<ID>(radices);
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
