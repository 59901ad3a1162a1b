/*
 * Kernel set of the pruned w-plane FFT for ONE fixed stage list: compiled once
 * per list with -DES_SEQ_ID=<id> -DES_SEQ_REAL=<float|double> (Makefile), so
 * that the sets build in parallel. The kernels are those of
 * es_fft_kernels.cuh with the radix of every pass known at compile time
 * (EsFftSeq<id> in es_fft_core.h); es_fft.cu selects a set when a plan's
 * stage list equals it.
 */
#include "es_fft_kernels.cuh"

#ifndef ES_SEQ_ID
#error "compile with -DES_SEQ_ID=<id> -DES_SEQ_REAL=<float|double>"
#endif

#define ES_SEQ_CAT2(a, b) a##b
#define ES_SEQ_CAT(a, b) ES_SEQ_CAT2(a, b)

const EsFftOps* ES_SEQ_CAT(es_fft_seq_ops_, ES_SEQ_ID)(void)
{
    return EsFftLaunch<ES_SEQ_REAL, ES_SEQ_ID>::ops();
}
