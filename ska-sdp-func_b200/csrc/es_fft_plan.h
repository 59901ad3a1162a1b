/*
 * Host-side planning of the in-shared-memory line FFT (es_fft_core.h):
 * factorisation of the grid size into radix passes, twiddle tables, and the
 * digit-reversal map. Plain C interface; no device code.
 */
#ifndef ES_FFT_PLAN_H_
#define ES_FFT_PLAN_H_

#include "es_fft_core_types.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Factorise n (11-smooth) into passes of radix <= max_radix (8 or 16).
 * Returns 0 on success, 1 when n cannot be handled (other prime factors, too
 * many stages, or the float index split is not exact for this n). */
int es_fft_line_plan(int32_t n, int32_t max_radix, EsFftLine* line);

/* Tables of a planned line: tw = num_twiddles (re, im) pairs of doubles,
 * exp(+2 pi i j / L) per stage (hi then lo); perm[k] = position of index k. */
void es_fft_line_tables(const EsFftLine* line, double* tw, int32_t* perm);

#ifdef __cplusplus
}
#endif

#endif
