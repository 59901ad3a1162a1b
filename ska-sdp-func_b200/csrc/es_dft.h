/*
 * C launchers of the point-source DFT kernels (es_dft.cu): plain device
 * pointers and sizes; work is enqueued on `stream`; the return value is a
 * cudaError_t as int. num_rows = num_times * num_baselines; num_pols is 1 or
 * 4; directions / fluxes / uvw are fp64, vis is complex fp64 (vis_double) or
 * complex fp32.
 */
#ifndef ES_DFT_H_
#define ES_DFT_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* vis[row, chan, pol] = sum_s flux[s, chan, pol] exp(-2 pi i (l_s u + m_s v +
 * n_s w)), uvw_lambda[row, chan, 3] in wavelengths (sdp_dft.cu:20-101) */
int es_dft_point_v00(int num_components, int num_pols, int num_chan,
        int64_t num_rows, const double* directions, const double* fluxes,
        const double* uvw_lambda, void* vis, int vis_double, void* stream);

/* the same with uvw[row, 3] in metres and the wavelength of channel c from
 * (channel_start_hz + c channel_step_hz) / c0 (sdp_dft.cu:107-194) */
int es_dft_point_v01(int num_components, int num_pols, int num_chan,
        int64_t num_rows, const double* directions, const double* fluxes,
        const double* uvw, double channel_start_hz, double channel_step_hz,
        void* vis, int vis_double, void* stream);

#ifdef __cplusplus
}
#endif

#endif
