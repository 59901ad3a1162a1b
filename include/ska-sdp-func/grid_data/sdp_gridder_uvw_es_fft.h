/*
 * Wide-field w-stacking visibility gridder / degridder with the
 * "exponential of semicircle" (ES) kernel -- B200-native drop-in.
 *
 * The four entry points below are, symbol for symbol and argument for
 * argument, what the reference's FFI binds for this path:
 *   reference C API        src/ska-sdp-func/grid_data/sdp_gridder_uvw_es_fft.h:42-110
 *   reference ctypes view  src/ska_sdp_func/grid_data/gridder_uvw_es_fft.py:149-200
 * Semantics that are part of the contract and are kept:
 *  - status is in/out and last; a non-zero *status on entry makes every call a
 *    no-op; a NULL plan makes the exec calls no-ops (es_fft.cpp:291,543,756);
 *  - validation order and codes follow es_fft.cpp:72-262
 *    (location 6 -> real/complex kinds 3 -> shapes 2 -> square 2 ->
 *     precision 3 -> contiguity 2 -> writability 2; pix_x != pix_y -> 2);
 *  - gridding ACCUMULATES into dirty_image and then applies the grid
 *    correction to the whole image (pass zeros);
 *  - degridding applies the grid correction IN PLACE to the caller's
 *    dirty_image, ignores weight, and ADDS the prediction to vis;
 *  - min_abs_w / max_abs_w are in wavelengths and are supplied by the caller
 *    (the Python class derives them with get_w_range, py:90-106).
 * Extension over the reference (which only accepts device arrays and fails
 * with SDP_ERR_MEM_LOCATION otherwise): when all five arrays are
 * SDP_MEM_CPU they are staged to the current device and back inside the
 * call (synchronous). Device arrays are processed asynchronously on the
 * legacy default stream, like the reference.
 */
#ifndef SDP_B200_GRIDDER_UVW_ES_FFT_H_
#define SDP_B200_GRIDDER_UVW_ES_FFT_H_

#include "ska-sdp-func/utility/sdp_mem.h"

#ifdef __cplusplus
extern "C" {
#endif

struct sdp_GridderUvwEsFft;
typedef struct sdp_GridderUvwEsFft sdp_GridderUvwEsFft;

/* Replaces es_fft.h:42-55 / es_fft.cpp:276-529.
 * uvw [R,3] real (metres), freq_hz [C] real ascending, vis [R,C] complex,
 * weight [R,C] real, dirty_image [N,N] real; one precision, C-contiguous,
 * one location. Returns NULL and sets *status on failure. */
sdp_GridderUvwEsFft* sdp_gridder_uvw_es_fft_create_plan(
        const sdp_Mem* uvw,
        const sdp_Mem* freq_hz,
        const sdp_Mem* vis,
        const sdp_Mem* weight,
        const sdp_Mem* dirty_image,
        const double pixel_size_x_rad,
        const double pixel_size_y_rad,
        const double epsilon,
        const double min_abs_w,
        const double max_abs_w,
        const int do_w_stacking,
        sdp_Error* status
);

/* Replaces es_fft.h:71-79 / es_fft.cpp:532-742: visibilities -> dirty image. */
void sdp_grid_uvw_es_fft(
        sdp_GridderUvwEsFft* plan,
        const sdp_Mem* uvw,
        const sdp_Mem* freq_hz,
        const sdp_Mem* vis,
        const sdp_Mem* weight,
        sdp_Mem* dirty_image,
        sdp_Error* status
);

/* Replaces es_fft.h:95-103 / es_fft.cpp:745-956: image -> visibilities. */
void sdp_ifft_degrid_uvw_es(
        sdp_GridderUvwEsFft* plan,
        const sdp_Mem* uvw,
        const sdp_Mem* freq_hz,
        sdp_Mem* vis,
        const sdp_Mem* weight,
        sdp_Mem* dirty_image,
        sdp_Error* status
);

/* Replaces es_fft.h:110 / es_fft.cpp:60-69. NULL-safe. */
void sdp_gridder_uvw_es_fft_free_plan(sdp_GridderUvwEsFft* plan);

#ifdef __cplusplus
}
#endif

#endif
