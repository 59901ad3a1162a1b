/*
 * Point-source DFT predict: visibilities of a list of point components by
 * direct summation -- the exact (no gridding error) counterpart of the
 * degridder, used to synthesise and to check visibilities.
 * Drop-in for the reference's
 *     src/ska-sdp-func/visibility/sdp_dft.h:53-59   sdp_dft_point_v00
 *     src/ska-sdp-func/visibility/sdp_dft.h:88-96   sdp_dft_point_v01
 * (same prototypes, argument meaning, checks in the same order with the same
 * error codes, sdp_dft.cpp:102-147,289-344; kernels sdp_dft.cu:20-194).
 *
 *   vis[t, b, c, p] = sum_s flux[s, c, p] exp(-2 pi i (l_s u + m_s v + n_s w))
 * with (u, v, w) in wavelengths: given per channel (v00), or from metres and
 * the channel frequency channel_start_hz + c channel_step_hz (v01).
 *
 * Array shapes, slowest to fastest varying:
 *   source_directions [num_components, 3]                       fp64 (l, m, n)
 *   source_fluxes     [num_components, num_channels, num_pols]  complex fp64
 *   uvw_lambda        [num_times, num_baselines, num_channels, 3] fp64 (v00)
 *   uvw               [num_times, num_baselines, 3]             fp64, metres (v01)
 *   vis               [num_times, num_baselines, num_channels, num_pols]
 *                     complex fp64 or complex fp32, OVERWRITTEN; num_pols 1 or 4
 *
 * This implementation runs on the GPU only: SDP_MEM_GPU arrays are processed
 * in place on the legacy default stream without synchronising; SDP_MEM_CPU
 * arrays are staged through the current device (the reference has an OpenMP
 * loop for them). All arrays of a call must be in the same location.
 * The phase is evaluated in turns with an exact reduction (see
 * csrc/es_dft.cu); results agree with the reference's to its own rounding of
 * the phase (~1e-16 x the phase in radians).
 */
#ifndef SKA_SDP_PROC_FUNC_DFT_H_
#define SKA_SDP_PROC_FUNC_DFT_H_

#include "ska-sdp-func/utility/sdp_mem.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Deprecated in the reference in favour of v01; kept for its callers. */
void sdp_dft_point_v00(
        const sdp_Mem* source_directions,
        const sdp_Mem* source_fluxes,
        const sdp_Mem* uvw_lambda,
        sdp_Mem* vis,
        sdp_Error* status
);

void sdp_dft_point_v01(
        const sdp_Mem* source_directions,
        const sdp_Mem* source_fluxes,
        const sdp_Mem* uvw,
        const double channel_start_hz,
        const double channel_step_hz,
        sdp_Mem* vis,
        sdp_Error* status
);

#ifdef __cplusplus
}
#endif

#endif /* include guard */
