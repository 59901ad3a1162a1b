/*
 * Host-side planning of the in-shared-memory line FFT (es_fft_core.h):
 * factorisation of the grid size into radix passes, twiddle tables, and the
 * digit-reversal maps. Plain C interface; no device code.
 */
#ifndef ES_FFT_PLAN_H_
#define ES_FFT_PLAN_H_

#include "es_fft_core_types.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Factorise n (11-smooth) into as few passes as possible of radix <=
 * max_radix, each radix one es_fft_core.h has a butterfly for. The last pass
 * gets an odd radix where one exists (its stride-1 accesses are then free of
 * bank conflicts). `radices` (optional, 0-terminated) forces the stage list.
 * Returns 0 on success, 1 when n cannot be handled (other prime factors, too
 * many stages, or the float index splits are not exact for this n). */
int es_fft_line_plan(int32_t n, int32_t max_radix, const int32_t* radices,
        EsFftLine* line);

/* Tables of a planned line:
 *  tw    num_twiddles (re, im) pairs of doubles, exp(+2 pi i j / L) per stage
 *        (hi then lo);
 *  perm  (optional) perm[k] = unpadded position of index k, n entries;
 *  bpos  (optional) n / last radix entries: padded position of the elements of
 *        digit-reversed butterfly t of the last pass (the one owning indices
 *        t + k n / radix). */
void es_fft_line_tables(const EsFftLine* line, double* tw, int32_t* perm,
        int32_t* bpos);

#ifdef __cplusplus
}
#endif

#endif
