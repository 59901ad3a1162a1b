/*
 * C launchers of the Hogbom CLEAN kernels (es_clean.cu): plain device pointers
 * and sizes; work is enqueued on `stream`; the return value is a cudaError_t
 * as int.
 */
#ifndef ES_CLEAN_H_
#define ES_CLEAN_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Bytes of device scratch es_hogbom_clean needs for n x n images,
 * `cycle_limit` cycles and a cbeam_dim x cbeam_dim CLEAN beam (is_double:
 * element type of the images). */
size_t es_hogbom_scratch_bytes(int64_t n, int cycle_limit, int64_t cbeam_dim,
        int is_double);

/* The whole minor cycle and the restore step (sdp_hogbom_clean.cpp:113-277):
 * residual (n x n, holds the dirty image on entry) and clean_model (n x n,
 * zero on entry) are updated in place for at most cycle_limit cycles of
 * "find the peak, stop below threshold, add gain x peak to the model,
 * subtract gain x peak x shifted PSF (2n x 2n)"; then skymodel = model (*)
 * CLEAN beam (Gaussian a, b, c coefficients, cbeam_dim^2 samples) + residual.
 * One cooperative launch runs all cycles. *cycles_done (device int, inside
 * scratch) is left for the caller to read if wanted. */
int es_hogbom_clean(int64_t n, const void* psf, double loop_gain,
        double threshold, int cycle_limit, double beam_a, double beam_b,
        double beam_c, int64_t cbeam_dim, void* clean_model, void* residual,
        void* skymodel, void* scratch, int is_double, void* stream);

#ifdef __cplusplus
}
#endif

#endif
