/*
 * TEST INFRASTRUCTURE ONLY (oracle). See es_oracle.h.
 *
 * Schedule differs from the reference (each visibility visits only the planes
 * it touches, rows of the grid are owned by threads so the sums are
 * deterministic); the arithmetic per visibility / pixel follows
 * sdp_gridder_uvw_es_fft_kernels.cu line by line, in fp64.
 */
#include "es_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include <omp.h>

#define C_0 299792458.0 /* kernels.cu:159 */
#define PI 3.1415926535897931 /* kernels.cu:10 */

typedef struct { double re, im; } cplx;

static double now_s(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double) ts.tv_sec + 1e-9 * (double) ts.tv_nsec;
}

/* ------------------------------------------------------------------ FFT */
/* Plain recursive mixed-radix decimation-in-time FFT for 11-smooth sizes
 * (the reference's grid sizes are 2 * good_size, es_fft_utils.cpp:181-218);
 * any other prime factor falls through to the generic O(r^2) butterfly too. */

typedef struct
{
    int n;
    int num_factors;
    int factors[64];
    cplx* twiddle; /* exp(sign * 2 pi i k / n), k < n */
} FftPlan;

static int fft_plan_init(FftPlan* fp, int n, int sign)
{
    fp->n = n;
    fp->num_factors = 0;
    int m = n;
    while (m % 4 == 0) { fp->factors[fp->num_factors++] = 4; m /= 4; }
    for (int f = 2; f * f <= m; ++f)
        while (m % f == 0) { fp->factors[fp->num_factors++] = f; m /= f; }
    if (m > 1) fp->factors[fp->num_factors++] = m;
    fp->twiddle = (cplx*) malloc(sizeof(cplx) * (size_t) n);
    if (!fp->twiddle) return 1;
    for (int k = 0; k < n; ++k)
    {
        const double a = (double) sign * 2.0 * PI * (double) k / (double) n;
        fp->twiddle[k].re = cos(a);
        fp->twiddle[k].im = sin(a);
    }
    return 0;
}

static void fft_plan_free(FftPlan* fp)
{
    free(fp->twiddle);
    fp->twiddle = 0;
}

/* out[0..n) = DFT of in[0], in[s], in[2 s], ...; n = product of
 * factors[level..). tw_step = N / n. */
static void fft_rec(const FftPlan* fp, int level, int n, int s, const cplx* in,
        cplx* out)
{
    if (n == 1) { out[0] = in[0]; return; }
    const int r = fp->factors[level];
    const int m = n / r;
    const int tw_step = fp->n / n;
    for (int k = 0; k < r; ++k)
        fft_rec(fp, level + 1, m, s * r, in + (size_t) k * s, out + (size_t) k * m);
    cplx t[64];
    const int r_step = fp->n / r;
    for (int j = 0; j < m; ++j)
    {
        for (int k = 0; k < r; ++k)
        {
            const cplx w = fp->twiddle[(size_t) j * k * tw_step];
            const cplx x = out[(size_t) k * m + j];
            t[k].re = x.re * w.re - x.im * w.im;
            t[k].im = x.re * w.im + x.im * w.re;
        }
        for (int q = 0; q < r; ++q)
        {
            double sr = t[0].re, si = t[0].im;
            for (int k = 1; k < r; ++k)
            {
                const cplx w = fp->twiddle[(size_t) ((q * k) % r) * r_step];
                sr += t[k].re * w.re - t[k].im * w.im;
                si += t[k].re * w.im + t[k].im * w.re;
            }
            out[(size_t) q * m + j].re = sr;
            out[(size_t) q * m + j].im = si;
        }
    }
}

int eso_fft2(double* data_, int n, int sign, int nthreads)
{
    cplx* data = (cplx*) data_;
    FftPlan fp;
    if (fft_plan_init(&fp, n, sign)) return 1;
    if (nthreads <= 0) nthreads = omp_get_max_threads();
    int err = 0;
    #pragma omp parallel num_threads(nthreads)
    {
        cplx* a = (cplx*) malloc(sizeof(cplx) * (size_t) n);
        cplx* b = (cplx*) malloc(sizeof(cplx) * (size_t) n);
        if (!a || !b)
        {
            #pragma omp atomic write
            err = 1;
        }
        #pragma omp barrier
        if (!err)
        {
            /* contiguous axis */
            #pragma omp for schedule(static)
            for (int row = 0; row < n; ++row)
            {
                cplx* x = data + (size_t) row * n;
                memcpy(a, x, sizeof(cplx) * (size_t) n);
                fft_rec(&fp, 0, n, 1, a, x);
            }
            /* strided axis */
            #pragma omp for schedule(static)
            for (int col = 0; col < n; ++col)
            {
                for (int row = 0; row < n; ++row) a[row] = data[(size_t) row * n + col];
                fft_rec(&fp, 0, n, 1, a, b);
                for (int row = 0; row < n; ++row) data[(size_t) row * n + col] = b[row];
            }
        }
        free(a);
        free(b);
    }
    fft_plan_free(&fp);
    return err;
}

/* --------------------------------------------------------- ES kernel bits */

/* kernels.cu:97-102 */
static inline double exp_semicircle(double beta, double x)
{
    const double xx = x * x;
    return (xx > 1.0) ? 0.0 : exp(beta * (sqrt(1.0 - xx) - 1.0));
}

/* kernels.cu:69-87 */
static double conv_corr(const eso_Params* p, double k)
{
    const double support = (double) p->support;
    double correction = 0.0;
    for (int i = 0; i < p->num_quad; ++i)
        correction += p->quad_kernel[i] *
                cos(PI * k * support * p->quad_nodes[i]) * p->quad_weights[i];
    return correction * support;
}

typedef struct
{
    double pos_u, pos_v, pos_w;
    double vre, vim; /* weighted, conjugated if flipped (scatter only) */
    double flip;
    int64_t index;   /* row * num_chan + chan */
} VisRec;

/* kernels.cu:157-181 (3-D) and :308-331 (2-D) */
static void make_records(const eso_Params* p, const double* uvw,
        const double* freq, const double* vis, const double* weight,
        VisRec* rec)
{
    const int R = p->num_rows, C = p->num_chan;
    #pragma omp parallel for schedule(static)
    for (int row = 0; row < R; ++row)
    {
        const double u = uvw[3 * row], v = uvw[3 * row + 1], w = uvw[3 * row + 2];
        for (int chan = 0; chan < C; ++chan)
        {
            VisRec* r = rec + (size_t) row * C + chan;
            const double flip = (p->do_wstacking && w < 0.0) ? -1.0 : 1.0;
            const double inv_wavelength = flip * freq[chan] / C_0;
            r->flip = flip;
            r->index = (int64_t) row * C + chan;
            r->pos_u = u * inv_wavelength * p->uv_scale;
            r->pos_v = v * inv_wavelength * p->uv_scale;
            r->pos_w = p->do_wstacking ?
                    (w * inv_wavelength - p->min_plane_w) * p->w_scale : 0.0;
            if (vis)
            {
                const double wt = weight[r->index];
                r->vre = vis[2 * r->index] * wt;
                r->vim = vis[2 * r->index + 1] * wt * flip;
            }
            else
            {
                r->vre = r->vim = 0.0;
            }
        }
    }
}

/* Bucket visibilities by their first plane ceil(pos_w - W/2): returns the
 * permutation and the bucket offsets [P + 1] (out-of-lattice first planes
 * are clamped into the end buckets; the exact per-plane test is re-done at
 * use, so this only has to be conservative). */
static int bucket_by_plane(const eso_Params* p, const VisRec* rec, int64_t nvis,
        int64_t** perm_out, int64_t** start_out)
{
    const int P = p->num_total_w_grids;
    const double hs = 0.5 * (double) p->support;
    int64_t* start = (int64_t*) calloc((size_t) P + 2, sizeof(int64_t));
    int64_t* perm = (int64_t*) malloc(sizeof(int64_t) * (size_t) (nvis > 0 ? nvis : 1));
    int32_t* first = (int32_t*) malloc(sizeof(int32_t) * (size_t) (nvis > 0 ? nvis : 1));
    if (!start || !perm || !first) { free(start); free(perm); free(first); return 1; }
    for (int64_t i = 0; i < nvis; ++i)
    {
        double f = ceil(rec[i].pos_w - hs);
        if (!(f > 0.0)) f = 0.0; /* also catches NaN */
        if (f > (double) (P - 1)) f = (double) (P - 1);
        first[i] = (int32_t) f;
        start[first[i] + 1]++;
    }
    for (int q = 0; q < P; ++q) start[q + 1] += start[q];
    int64_t* cursor = (int64_t*) malloc(sizeof(int64_t) * (size_t) (P + 1));
    if (!cursor) { free(start); free(perm); free(first); return 1; }
    memcpy(cursor, start, sizeof(int64_t) * (size_t) (P + 1));
    for (int64_t i = 0; i < nvis; ++i) perm[cursor[first[i]]++] = i;
    free(cursor);
    free(first);
    *perm_out = perm;
    *start_out = start;
    return 0;
}

/* Footprint of one visibility on plane grid_w: kernels.cu:174-221. Returns 0
 * when there is no overlap. */
typedef struct
{
    int u_min, u_max, v_min, v_max;
    double ku[33], kv[33], kw;
} Footprint;

static inline int footprint(const eso_Params* p, const VisRec* r, int grid_w,
        Footprint* f)
{
    const double half_support = (double) p->support / 2.0;
    const int g = p->grid_size;
    if (!isfinite(r->pos_u) || !isfinite(r->pos_v) || !isfinite(r->pos_w)) return 0;
    const int grid_min_uv = -g / 2;
    const int grid_max_uv = (g - 1) / 2;
    const int u_min = (int) fmax(ceil(r->pos_u - half_support), (double) grid_min_uv);
    const int u_max = (int) fmin(floor(r->pos_u + half_support), (double) grid_max_uv);
    const int v_min = (int) fmax(ceil(r->pos_v - half_support), (double) grid_min_uv);
    const int v_max = (int) fmin(floor(r->pos_v + half_support), (double) grid_max_uv);
    const double w_lo = ceil(r->pos_w - half_support);
    const double w_hi = floor(r->pos_w + half_support);
    if ((double) grid_w < w_lo || (double) grid_w > w_hi) return 0;
    if (u_min > u_max || v_min > v_max) return 0;
    if (u_max - u_min > 32 || v_max - v_min > 32) return 0; /* NaN/Inf guard */
    const double inv_half_support = 1.0 / half_support;
    f->u_min = u_min; f->u_max = u_max; f->v_min = v_min; f->v_max = v_max;
    for (int gu = u_min; gu <= u_max; ++gu)
        f->ku[gu - u_min] = exp_semicircle(p->beta,
                ((double) gu - r->pos_u) * inv_half_support);
    for (int gv = v_min; gv <= v_max; ++gv)
        f->kv[gv - v_min] = exp_semicircle(p->beta,
                ((double) gv - r->pos_v) * inv_half_support);
    f->kw = exp_semicircle(p->beta,
            ((double) grid_w - r->pos_w) * inv_half_support);
    return 1;
}

/* ------------------------------------------------------- image-side steps */

/* kernels.cu:110-123; signage -1 gridding, +1 degridding. */
static inline void phase_shift(double w, double l, double m, double signage,
        double* re, double* im)
{
    const double sos = l * l + m * m;
    const double nm1 = (-sos) / (sqrt(1.0 - sos) + 1.0);
    const double x = 2.0 * PI * w * nm1;
    const double xn = 1.0 / (nm1 + 1.0);
    *re = cos(signage * x) * xn;
    *im = sin(signage * x) * xn;
}

/* kernels.cu:690-769, applied to every pixel the reference's four guarded
 * mirror updates reach: offsets -N/2 .. N/2-1 on both axes. */
static void apply_correction(const eso_Params* p, double* image)
{
    const int n = p->image_size, half = n / 2;
    #pragma omp parallel for schedule(static)
    for (int a = -half; a < half; ++a)
    {
        for (int b = -half; b < half; ++b)
        {
            const int i = abs(b), j = abs(a);
            const double l = p->pixel_size * i, m = p->pixel_size * j;
            const double nn = sqrt(1.0 - l * l - m * m) - 1.0;
            const double l_conv = p->conv_corr[i], m_conv = p->conv_corr[j];
            double correction;
            if (p->do_wstacking)
            {
                double n_conv = conv_corr(p, nn * p->inv_w_scale);
                n_conv *= p->conv_corr_norm_factor * p->conv_corr_norm_factor;
                correction = l_conv * m_conv * n_conv;
            }
            else
            {
                correction = l_conv * m_conv * p->conv_corr_norm_factor *
                        p->conv_corr_norm_factor;
            }
            image[(size_t) (half + a) * n + (half + b)] *= 1.0 / correction;
        }
    }
}

/* ------------------------------------------------------------- pipelines */

int eso_grid(const eso_Params* p, const double* uvw, const double* freq,
        const double* vis, const double* weight, double* dirty,
        int plane_begin, int plane_end, int nthreads, double* timings)
{
    if (nthreads <= 0) nthreads = omp_get_max_threads();
    omp_set_num_threads(nthreads);
    const int g = p->grid_size, n = p->image_size, half = n / 2;
    const int P = p->num_total_w_grids, W = p->support;
    const int64_t nvis = (int64_t) p->num_rows * p->num_chan;
    if (plane_end < 0 || plane_end > P) plane_end = P;
    VisRec* rec = (VisRec*) malloc(sizeof(VisRec) * (size_t) (nvis > 0 ? nvis : 1));
    cplx* plane = (cplx*) malloc(sizeof(cplx) * (size_t) g * g);
    int64_t* perm = 0; int64_t* start = 0;
    if (!rec || !plane) { free(rec); free(plane); return 1; }
    make_records(p, uvw, freq, vis, weight, rec);
    if (bucket_by_plane(p, rec, nvis, &perm, &start)) { free(rec); free(plane); return 1; }
    double t_scatter = 0, t_fft = 0, t_screen = 0;
    const int origin = g / 2;
    for (int grid_w = plane_begin; grid_w < plane_end; ++grid_w)
    {
        double t0 = now_s();
        memset(plane, 0, sizeof(cplx) * (size_t) g * g);
        /* candidates: first plane in [grid_w - W - 1, grid_w]; the exact
         * per-plane test in footprint() decides. */
        int q_lo = grid_w - W - 1; if (q_lo < 0) q_lo = 0;
        const int64_t c_lo = start[q_lo];
        const int64_t c_hi = start[(grid_w + 1 > P) ? P : grid_w + 1];
        #pragma omp parallel
        {
            const int tid = omp_get_thread_num(), nt = omp_get_num_threads();
            const int row_lo = (int) ((int64_t) g * tid / nt);
            const int row_hi = (int) ((int64_t) g * (tid + 1) / nt);
            const double hs = 0.5 * W;
            Footprint f;
            for (int64_t c = c_lo; c < c_hi; ++c)
            {
                const VisRec* r = rec + perm[c];
                /* cheap row-ownership test first */
                if (r->pos_u + hs + origin < (double) row_lo - 1.0 ||
                        r->pos_u - hs + origin > (double) row_hi + 1.0)
                    continue;
                if (!footprint(p, r, grid_w, &f)) continue;
                for (int gu = f.u_min; gu <= f.u_max; ++gu)
                {
                    const int row = gu + origin;
                    if (row < row_lo || row >= row_hi) continue;
                    for (int gv = f.v_min; gv <= f.v_max; ++gv)
                    {
                        /* kernels.cu:231-252 */
                        double kv = f.ku[gu - f.u_min] * f.kv[gv - f.v_min] * f.kw;
                        if ((gu + gv) & 1) kv = -kv;
                        cplx* cell = plane + (size_t) row * g + (gv + origin);
                        cell->re += r->vre * kv;
                        cell->im += r->vim * kv;
                    }
                }
            }
        }
        double t1 = now_s();
        if (eso_fft2((double*) plane, g, +1, nthreads)) { free(rec); free(plane); free(perm); free(start); return 1; }
        double t2 = now_s();
        /* kernels.cu:429-547 over offsets -N/2 .. N/2-1 */
        const double w_plane = (double) grid_w * p->inv_w_scale + p->min_plane_w;
        #pragma omp parallel for schedule(static)
        for (int a = -half; a < half; ++a)
        {
            for (int b = -half; b < half; ++b)
            {
                const int i = abs(b), j = abs(a);
                double sre = 1.0, sim = 0.0;
                if (p->do_wstacking)
                    phase_shift(w_plane, p->pixel_size * i, p->pixel_size * j,
                            -1.0, &sre, &sim);
                const cplx v = plane[(size_t) (origin + a) * g + (origin + b)];
                double val = v.re * sre - v.im * sim;
                if ((i + j) & 1) val = -val;
                dirty[(size_t) (half + a) * n + (half + b)] += val;
            }
        }
        double t3 = now_s();
        t_scatter += t1 - t0; t_fft += t2 - t1; t_screen += t3 - t2;
    }
    double t0 = now_s();
    apply_correction(p, dirty);
    if (timings)
    {
        timings[0] = t_scatter; timings[1] = t_fft; timings[2] = t_screen;
        timings[3] = now_s() - t0;
    }
    free(rec); free(plane); free(perm); free(start);
    return 0;
}

int eso_degrid(const eso_Params* p, const double* uvw, const double* freq,
        double* vis, double* dirty, int plane_begin, int plane_end,
        int nthreads, double* timings)
{
    if (nthreads <= 0) nthreads = omp_get_max_threads();
    omp_set_num_threads(nthreads);
    const int g = p->grid_size, n = p->image_size, half = n / 2;
    const int P = p->num_total_w_grids, W = p->support;
    const int64_t nvis = (int64_t) p->num_rows * p->num_chan;
    if (plane_end < 0 || plane_end > P) plane_end = P;
    VisRec* rec = (VisRec*) malloc(sizeof(VisRec) * (size_t) (nvis > 0 ? nvis : 1));
    cplx* plane = (cplx*) malloc(sizeof(cplx) * (size_t) g * g);
    cplx* acc = (cplx*) calloc((size_t) (nvis > 0 ? nvis : 1), sizeof(cplx));
    int64_t* perm = 0; int64_t* start = 0;
    if (!rec || !plane || !acc) { free(rec); free(plane); free(acc); return 1; }
    make_records(p, uvw, freq, 0, 0, rec);
    if (bucket_by_plane(p, rec, nvis, &perm, &start)) { free(rec); free(plane); free(acc); return 1; }
    double t_gather = 0, t_fft = 0, t_screen = 0;
    const int origin = g / 2;
    double tc = now_s();
    apply_correction(p, dirty); /* in place on the caller's image, es_fft.cpp:789-825 */
    tc = now_s() - tc;
    for (int grid_w = plane_begin; grid_w < plane_end; ++grid_w)
    {
        double t0 = now_s();
        memset(plane, 0, sizeof(cplx) * (size_t) g * g);
        /* kernels.cu:554-679 */
        const double w_plane = (double) grid_w * p->inv_w_scale + p->min_plane_w;
        #pragma omp parallel for schedule(static)
        for (int a = -half; a < half; ++a)
        {
            for (int b = -half; b < half; ++b)
            {
                const int i = abs(b), j = abs(a);
                double sre = 1.0, sim = 0.0;
                if (p->do_wstacking)
                    phase_shift(w_plane, p->pixel_size * i, p->pixel_size * j,
                            1.0, &sre, &sim);
                double val = dirty[(size_t) (half + a) * n + (half + b)];
                if ((i + j) & 1) val = -val;
                cplx* cell = plane + (size_t) (origin + a) * g + (origin + b);
                cell->re = sre * val;
                cell->im = sim * val;
            }
        }
        double t1 = now_s();
        if (eso_fft2((double*) plane, g, -1, nthreads)) { free(rec); free(plane); free(acc); free(perm); free(start); return 1; }
        double t2 = now_s();
        int q_lo = grid_w - W - 1; if (q_lo < 0) q_lo = 0;
        const int64_t c_lo = start[q_lo];
        const int64_t c_hi = start[(grid_w + 1 > P) ? P : grid_w + 1];
        #pragma omp parallel for schedule(dynamic, 256)
        for (int64_t c = c_lo; c < c_hi; ++c)
        {
            const VisRec* r = rec + perm[c];
            Footprint f;
            if (!footprint(p, r, grid_w, &f)) continue;
            double sre = 0.0, sim = 0.0;
            for (int gu = f.u_min; gu <= f.u_max; ++gu)
            {
                for (int gv = f.v_min; gv <= f.v_max; ++gv)
                {
                    /* kernels.cu:231-235,254-260 */
                    double kv = f.ku[gu - f.u_min] * f.kv[gv - f.v_min] * f.kw;
                    if ((gu + gv) & 1) kv = -kv;
                    const cplx cell = plane[(size_t) (gu + origin) * g + (gv + origin)];
                    sre += cell.re * kv;
                    sim += cell.im * kv;
                }
            }
            acc[r->index].re += sre;
            acc[r->index].im += sim * r->flip; /* kernels.cu:267-268 */
        }
        double t3 = now_s();
        t_screen += t1 - t0; t_fft += t2 - t1; t_gather += t3 - t2;
    }
    for (int64_t i = 0; i < nvis; ++i)
    {
        vis[2 * i] += acc[i].re;
        vis[2 * i + 1] += acc[i].im;
    }
    if (timings)
    {
        timings[0] = t_gather; timings[1] = t_fft; timings[2] = t_screen;
        timings[3] = tc;
    }
    free(rec); free(plane); free(acc); free(perm); free(start);
    return 0;
}

int eso_max_threads(void)
{
    return omp_get_max_threads();
}
