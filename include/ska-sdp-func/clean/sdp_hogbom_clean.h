/*
 * Hogbom CLEAN: the image-plane minor cycle that consumes the gridder's dirty
 * image and PSF -- repeatedly find the peak of the residual, move loop_gain x
 * peak into the model and subtract loop_gain x peak x the PSF centred there --
 * and the restore step (model convolved with a Gaussian CLEAN beam, plus the
 * residual).
 * Drop-in for the reference's
 *     src/ska-sdp-func/clean/sdp_hogbom_clean.h:36-47
 * (same prototype, argument meaning, checks in the same order with the same
 * error codes, sdp_hogbom_clean.cpp:733-853; arithmetic of its host loop,
 * .cpp:113-277).
 *
 * Arrays, all of one data type (fp64 or fp32):
 *   dirty_img     [n, n]      input
 *   psf           [2n, 2n]    input, peak at [n, n]
 *   cbeam_details [4]         (sigma_x, sigma_y, theta in degrees, size of the
 *                             sampled beam) -- a HOST array, read by the host
 *                             as in the reference's GPU path
 *   clean_model   [n, n]      output: unconvolved components
 *   residual      [n, n]      output: what is left of the dirty image
 *   skymodel      [n, n]      output: model (*) CLEAN beam + residual
 * loop_gain > 0; the loop stops at cycle_limit cycles or when the peak falls
 * below threshold (a negative threshold runs exactly cycle_limit cycles).
 *
 * This implementation runs on the GPU only: SDP_MEM_GPU images are processed
 * on the legacy default stream (the call synchronises before it returns: its
 * scratch is freed); SDP_MEM_CPU images are staged through the current
 * device. Residual and model are bit-identical to the reference's CPU path;
 * the restore step sums the same products directly instead of through FFTs of
 * padded images (agrees to rounding).
 */
#ifndef SKA_SDP_PROC_FUNC_HOGBOM_CLEAN_H_
#define SKA_SDP_PROC_FUNC_HOGBOM_CLEAN_H_

#include "ska-sdp-func/utility/sdp_mem.h"

#ifdef __cplusplus
extern "C" {
#endif

void sdp_hogbom_clean(
        const sdp_Mem* dirty_img,
        const sdp_Mem* psf,
        const sdp_Mem* cbeam_details,
        const double loop_gain,
        const double threshold,
        const int cycle_limit,
        sdp_Mem* clean_model,
        sdp_Mem* residual,
        sdp_Mem* skymodel,
        sdp_Error* status
);

#ifdef __cplusplus
}
#endif

#endif /* include guard */
