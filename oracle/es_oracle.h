/*
 * TEST INFRASTRUCTURE ONLY (oracle). Not part of the product library: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may build, load or call this.
 *
 * CPU restatement (plain C, fp64 throughout, fp64 coordinates) of the
 * reference's w-stacking ES gridder / degridder:
 *   src/ska-sdp-func/grid_data/sdp_gridder_uvw_es_fft.cpp:578-741  (grid loop)
 *   src/ska-sdp-func/grid_data/sdp_gridder_uvw_es_fft.cpp:789-952  (degrid loop)
 *   src/ska-sdp-func/grid_data/sdp_gridder_uvw_es_fft_kernels.cu   (all kernels)
 * The plan scalars and tables are computed by oracle/es_oracle.py (restating
 * es_fft.cpp:301-424 and es_fft_utils.cpp) and passed in through eso_Params.
 *
 * Parity pin: checked in tests/test_oracle.py against oracle/_ref (the
 * reference's own kernel source compiled for the host), against the explicit
 * DFT (oracle/dft.py) and against the committed golden vectors generated from
 * oracle/_ref (tests/golden/).
 */
#ifndef ES_ORACLE_H_
#define ES_ORACLE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct eso_Params
{
    int32_t num_rows, num_chan, image_size, grid_size, support;
    int32_t num_total_w_grids, do_wstacking, num_quad; /* num_quad = ceil(1.5 W + 2) */
    double beta;            /* beta_row * support */
    double pixel_size, uv_scale, w_scale, inv_w_scale, min_plane_w;
    double conv_corr_norm_factor;
    const double* quad_kernel;   /* [32] */
    const double* quad_nodes;    /* [32] */
    const double* quad_weights;  /* [32] */
    const double* conv_corr;     /* [image_size / 2 + 1] */
} eso_Params;

/* dirty[N,N] is accumulated into, then corrected (reference semantics).
 * vis is complex128 [R,C] interleaved. Planes [plane_begin, plane_end) only;
 * plane_end < 0 means all. timings[4] (may be NULL): scatter, fft, screen,
 * correction seconds. Returns 0 on success. */
int eso_grid(const eso_Params* p, const double* uvw, const double* freq_hz,
        const double* vis, const double* weight, double* dirty,
        int plane_begin, int plane_end, int nthreads, double* timings);

/* vis += predicted; dirty is corrected in place; weight unused. */
int eso_degrid(const eso_Params* p, const double* uvw, const double* freq_hz,
        double* vis, double* dirty, int plane_begin, int plane_end,
        int nthreads, double* timings);

/* Unnormalised in-place 2-D complex FFT of an n x n array (interleaved
 * re,im), sign = -1 forward / +1 backward. Exposed so the tests can check it
 * against numpy.fft. */
int eso_fft2(double* data, int n, int sign, int nthreads);

int eso_max_threads(void);

#ifdef __cplusplus
}
#endif

#endif
