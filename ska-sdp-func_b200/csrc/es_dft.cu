/*
 * sm_100a kernels of the point-source DFT predict (sdp_dft_point_v00 / _v01),
 * behind plain-pointer C launchers (es_dft.h). Replaces the reference's
 * dft_point_v00 / dft_point_v01 (src/ska-sdp-func/visibility/sdp_dft.cu:
 * 20-194: one thread per visibility, per source an fp64 phase in radians and
 * a full-range fp64 sincos).
 *
 * Bound: the fp64 pipe (the phasor of every (visibility, source) pair), not
 * memory -- a visibility is written once and sums num_components terms. What
 * is done about it:
 *  - the phase is kept in TURNS, t = (l u + m v + n w) / lambda, reduced
 *    exactly (t - rint(t) is exact in fp64) and evaluated with sincospi: no
 *    multiplication by 2 pi before the reduction (the reference rounds the
 *    phase in radians at |phase| ~ 1e5, which costs it 1e-11 of absolute
 *    phase) and no large-argument path;
 *  - v01: the phase is linear in the channel index, t_c = t_0 + c dt, so a
 *    thread owns CH consecutive channels of one row and, per source, evaluates
 *    two phasors (first channel, channel step) and ROTATES through the other
 *    channels (one fp64 complex product each, error ~ 1e-16 per step) instead
 *    of CH sincos evaluations; long runs rotate on up to four interleaved
 *    chains (step r^4) so that no dependency chain is longer than 8 products;
 *  - lanes of a warp are consecutive rows of ONE channel group, so the source
 *    directions and the fluxes of the group are warp-uniform: they are staged
 *    in shared memory per chunk of sources (fluxes already cast to the
 *    accumulator's precision, as the reference's flux_cast) and read as
 *    broadcasts;
 *  - complex fp32 visibilities: phasors in fp64 as above, cast once, products
 *    and sums in fp32 (the reference accumulates in the visibility type too).
 * The output is overwritten, not accumulated into (sdp_dft.cu:92-100).
 */
#include <stdint.h>

#include <cuda_runtime.h>

#include "es_dft.h"

#define ES_DFT_C0 299792458.0 /* sdp_dft.cu:7 */
#define ES_DFT_THREADS 128
#define ES_DFT_CHUNK 32 /* sources staged per pass */

namespace {

template<typename A> struct Cplx { A x, y; };

/* exp(-2 pi i t) for t in turns: exact reduction, then sincospi. */
__device__ __forceinline__ void phasor_of_turns(const double t, double& c,
        double& s)
{
    const double r = t - rint(t);
    sincospi(-2.0 * r, &s, &c);
}

/* One thread: row `row`, channels [c0, c0 + CH) of it; A = accumulator type.
 * LINEAR: uvw in metres and per-channel 1 / wavelength = k0 + j dk (v01);
 * otherwise uvw_lambda[row, chan, 3] (v00). */
template<typename A, int P, int CH, bool LINEAR>
__global__ void __launch_bounds__(ES_DFT_THREADS) es_dft_kernel(
        const int num_components, const int num_chan, const int64_t num_rows,
        const double* __restrict__ directions,
        const double* __restrict__ fluxes, const double* __restrict__ uvw,
        const double channel_start_hz, const double channel_step_hz,
        Cplx<A>* __restrict__ vis)
{
    __shared__ double s_dir[ES_DFT_CHUNK][3];
    __shared__ Cplx<A> s_flux[ES_DFT_CHUNK][CH][P];

    const int64_t row = (int64_t) blockIdx.x * ES_DFT_THREADS + threadIdx.x;
    const int c0 = blockIdx.y * CH;
    const bool active = row < num_rows;
    const int nch = min(CH, num_chan - c0);

    double u[LINEAR ? 1 : CH], v[LINEAR ? 1 : CH], w[LINEAR ? 1 : CH];
    double k0 = 0.0, dk = 0.0;
    if (LINEAR)
    {
        u[0] = v[0] = w[0] = 0.0;
        if (active)
        {
            u[0] = uvw[3 * row];
            v[0] = uvw[3 * row + 1];
            w[0] = uvw[3 * row + 2];
        }
        k0 = (channel_start_hz + c0 * channel_step_hz) / ES_DFT_C0;
        dk = channel_step_hz / ES_DFT_C0;
    }
    else
    {
#pragma unroll
        for (int j = 0; j < CH; ++j)
        {
            u[j] = v[j] = w[j] = 0.0;
            if (active && j < nch)
            {
                const double* at = uvw + 3 * (row * num_chan + c0 + j);
                u[j] = at[0];
                v[j] = at[1];
                w[j] = at[2];
            }
        }
    }

    Cplx<A> acc[CH][P];
#pragma unroll
    for (int j = 0; j < CH; ++j)
#pragma unroll
        for (int p = 0; p < P; ++p) acc[j][p].x = acc[j][p].y = (A) 0;

    for (int s0 = 0; s0 < num_components; s0 += ES_DFT_CHUNK)
    {
        const int ns = min(ES_DFT_CHUNK, num_components - s0);
        __syncthreads();
        for (int i = threadIdx.x; i < ns * 3; i += ES_DFT_THREADS)
            s_dir[i / 3][i % 3] = directions[3 * (int64_t) s0 + i];
        for (int i = threadIdx.x; i < ns * CH * P; i += ES_DFT_THREADS)
        {
            const int s = i / (CH * P), jp = i % (CH * P);
            const int j = jp / P, p = jp % P;
            Cplx<A> f;
            f.x = f.y = (A) 0;
            if (j < nch)
            {
                const double* at = fluxes + 2 * (((int64_t) (s0 + s) * num_chan
                        + c0 + j) * P + p);
                f.x = (A) at[0];
                f.y = (A) at[1];
            }
            s_flux[s][j][p] = f;
        }
        __syncthreads();
        if (!active) continue;
        for (int s = 0; s < ns; ++s)
        {
            const double l = s_dir[s][0], m = s_dir[s][1], n = s_dir[s][2];
            /* NC interleaved rotation chains: channel j is on chain j % NC,
             * which starts at p0 r^(j % NC) and advances by r^NC -- NC times
             * shorter dependency chains for 5 more complex products */
            constexpr int NC = !LINEAR ? 1 : CH >= 16 ? 4 : CH >= 8 ? 2 : 1;
            double pc[NC], ps[NC], rc = 1.0, rs = 0.0;
            if (LINEAR)
            {
                const double d = l * u[0] + m * v[0] + n * w[0];
                phasor_of_turns(k0 * d, pc[0], ps[0]);
                if (CH > 1) phasor_of_turns(dk * d, rc, rs);
#pragma unroll
                for (int c = 1; c < NC; ++c)
                {
                    pc[c] = pc[c - 1] * rc - ps[c - 1] * rs;
                    ps[c] = pc[c - 1] * rs + ps[c - 1] * rc;
                }
#pragma unroll
                for (int k = 1; k < NC; k *= 2)
                {
                    const double sq = rc * rc - rs * rs;
                    rs = 2.0 * rc * rs;
                    rc = sq;
                }
            }
#pragma unroll
            for (int j = 0; j < CH; ++j)
            {
                const int c = j % NC;
                if (!LINEAR)
                    phasor_of_turns(l * u[j] + m * v[j] + n * w[j], pc[0], ps[0]);
                const A ac = (A) pc[c], as = (A) ps[c];
#pragma unroll
                for (int p = 0; p < P; ++p)
                {
                    const Cplx<A> f = s_flux[s][j][p];
                    acc[j][p].x += ac * f.x - as * f.y;
                    acc[j][p].y += ac * f.y + as * f.x;
                }
                if (LINEAR && j + NC < CH)
                {
                    const double nc = pc[c] * rc - ps[c] * rs;
                    ps[c] = pc[c] * rs + ps[c] * rc;
                    pc[c] = nc;
                }
            }
        }
    }
    if (!active) return;
#pragma unroll
    for (int j = 0; j < CH; ++j)
    {
        if (j >= nch) break;
#pragma unroll
        for (int p = 0; p < P; ++p)
            vis[(row * num_chan + c0 + j) * P + p] = acc[j][p];
    }
}

template<typename A, int P, int CH, bool LINEAR>
int launch(const int num_components, const int num_chan,
        const int64_t num_rows, const double* directions,
        const double* fluxes, const double* uvw, const double start_hz,
        const double step_hz, void* vis, cudaStream_t st)
{
    if (num_rows <= 0 || num_chan <= 0) return 0;
    if ((num_chan + CH - 1) / CH > 65535) return (int) cudaErrorInvalidValue;
    const dim3 grid((unsigned) ((num_rows + ES_DFT_THREADS - 1) / ES_DFT_THREADS),
            (unsigned) ((num_chan + CH - 1) / CH));
    es_dft_kernel<A, P, CH, LINEAR><<<grid, ES_DFT_THREADS, 0, st>>>(
            num_components, num_chan, num_rows, directions, fluxes, uvw,
            start_hz, step_hz, (Cplx<A>*) vis);
    return (int) cudaGetLastError();
}

/* Channels per thread (v01). Long runs amortise the two phasor evaluations
 * of a source (~80 fp64 operations) over CH rotations and CH x P complex
 * sums; the accumulators (CH x P complex) and the channels a last, partly
 * filled group wastes bound them. Cost model per row and source, in fp64
 * operation slots: groups x (80 + CH x (4 + 4 P w)), w = 1 for fp64 sums,
 * 1/2 for fp32 ones. */
int pick_channels(const int num_chan, const int num_pols, const int vis_double,
        const int max_ch)
{
    int best = 4;
    double best_cost = 0.0;
    for (int ch = 4; ch <= max_ch; ch *= 2)
    {
        const int groups = (num_chan + ch - 1) / ch;
        const double cost = groups * (80.0 + ch * (4.0 + 4.0 * num_pols *
                (vis_double ? 1.0 : 0.5)));
        if (ch == 4 || cost < best_cost)
        {
            best = ch;
            best_cost = cost;
        }
    }
    return best;
}

template<bool LINEAR>
int dispatch(const int num_components, const int num_pols, const int num_chan,
        const int64_t num_rows, const double* directions,
        const double* fluxes, const double* uvw, const double start_hz,
        const double step_hz, void* vis, const int vis_double, cudaStream_t st)
{
    if (num_pols != 1 && num_pols != 4) return (int) cudaErrorInvalidValue;
#define ES_DFT_GO(A, P, CH) launch<A, P, CH, LINEAR>(num_components, num_chan, \
        num_rows, directions, fluxes, uvw, start_hz, step_hz, vis, st)
    if constexpr (!LINEAR)
    {
        /* one phasor evaluation per (visibility, source): short groups */
        if (vis_double)
            return num_pols == 1 ? ES_DFT_GO(double, 1, 4) : ES_DFT_GO(double, 4, 2);
        return num_pols == 1 ? ES_DFT_GO(float, 1, 4) : ES_DFT_GO(float, 4, 2);
    }
    else
    {
    /* accumulators of at most 32 complex values per thread (measured: 16
     * channels x 4 polarisations in fp32 are slower than 8 x 4) */
    const int ch = pick_channels(num_chan, num_pols, vis_double, 32 / num_pols);
    if (vis_double && num_pols == 1)
        return ch == 32 ? ES_DFT_GO(double, 1, 32) : ch == 16 ?
                ES_DFT_GO(double, 1, 16) : ch == 8 ? ES_DFT_GO(double, 1, 8) :
                ES_DFT_GO(double, 1, 4);
    if (vis_double)
        return ch == 8 ? ES_DFT_GO(double, 4, 8) : ES_DFT_GO(double, 4, 4);
    if (num_pols == 1)
        return ch == 32 ? ES_DFT_GO(float, 1, 32) : ch == 16 ?
                ES_DFT_GO(float, 1, 16) : ch == 8 ? ES_DFT_GO(float, 1, 8) :
                ES_DFT_GO(float, 1, 4);
    return ch == 8 ? ES_DFT_GO(float, 4, 8) : ES_DFT_GO(float, 4, 4);
    }
#undef ES_DFT_GO
}

} // namespace

extern "C" {

int es_dft_point_v00(int num_components, int num_pols, int num_chan,
        int64_t num_rows, const double* directions, const double* fluxes,
        const double* uvw_lambda, void* vis, int vis_double, void* stream)
{
    return dispatch<false>(num_components, num_pols, num_chan, num_rows,
            directions, fluxes, uvw_lambda, 0.0, 0.0, vis, vis_double,
            (cudaStream_t) stream);
}

int es_dft_point_v01(int num_components, int num_pols, int num_chan,
        int64_t num_rows, const double* directions, const double* fluxes,
        const double* uvw, double channel_start_hz, double channel_step_hz,
        void* vis, int vis_double, void* stream)
{
    return dispatch<true>(num_components, num_pols, num_chan, num_rows,
            directions, fluxes, uvw, channel_start_hz, channel_step_hz, vis,
            vis_double, (cudaStream_t) stream);
}

} // extern "C"
