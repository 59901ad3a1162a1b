/*
 * Host-side planning maths of the ES gridder. Pure C, no device code: these
 * numbers have to come out bit-identical to the reference's so that both
 * implementations work on the same grid, kernel and w-plane lattice
 * (SURVEY.md A.1/A.2). Checked against the host-compiled reference in
 * tests/test_plan.py over a sweep of (image size, epsilon, precision).
 */
#include <math.h>
#include <string.h>

#include "es_plan.h"

/* ES kernel parameter table (ducc0 "KernelDB" data as vendored by the
 * reference, es_fft_utils.cpp:241-486); generated file, one row per kernel. */
typedef struct
{
    int support;
    double sigma, epsilon, beta_over_support, corr_range;
} EsKernelRow;

static const EsKernelRow es_kernel_db[] = {
#define ES_ROW(w, s, e, b, c) {w, s, e, b, c},
#include "es_kernel_db.inc"
#undef ES_ROW
};
#define ES_NUM_KERNELS ((int) (sizeof(es_kernel_db) / sizeof(es_kernel_db[0])))

int es_good_size_complex(int n)
{
    /* es_fft_utils.cpp:181-218. Enumerate 5^a 7^b 11^c, then fill in the
     * powers of two and three. */
    if (n <= 12) return n;
    int best = 2 * n;
    for (int p11 = 1; p11 < best; p11 *= 11)
    {
        for (int p7 = p11; p7 < best; p7 *= 7)
        {
            for (int p5 = p7; p5 < best; p5 *= 5)
            {
                int x = p5;
                while (x < n) x *= 2;
                for (;;)
                {
                    if (x < n)
                    {
                        x *= 3;
                    }
                    else if (x > n)
                    {
                        if (x < best) best = x;
                        if (x & 1) break;
                        x >>= 1;
                    }
                    else
                    {
                        return n;
                    }
                }
            }
        }
    }
    return best;
}

static int grid_size_for(int image_size, double sigma)
{
    return 2 * es_good_size_complex((int) (image_size * sigma * 0.5) + 1);
}

void es_select_kernel(double epsilon, int image_size, int is_double,
        int* grid_size, int* support, double* beta_over_support)
{
    /* es_fft_utils.cpp:488-536 */
    enum { NUM_SLOTS = 20 };
    int row_of[NUM_SLOTS];
    double sigma_of[NUM_SLOTS];
    for (int s = 0; s < NUM_SLOTS; ++s)
    {
        row_of[s] = ES_NUM_KERNELS - 1;
        sigma_of[s] = 2.6;
    }
    const int max_support = is_double ? 16 : 8;
    for (int i = 0; i < ES_NUM_KERNELS; ++i)
    {
        const EsKernelRow* k = &es_kernel_db[i];
        if (k->support <= max_support && k->epsilon <= epsilon &&
                k->corr_range <= 10.0 && k->sigma <= sigma_of[k->support])
        {
            row_of[k->support] = i;
            sigma_of[k->support] = k->sigma;
        }
    }
    int best_nu = 80000, best_row = ES_NUM_KERNELS;
    for (int s = 0; s < NUM_SLOTS; ++s)
    {
        const int nu = grid_size_for(image_size, es_kernel_db[row_of[s]].sigma);
        if (nu <= best_nu)
        {
            best_nu = nu;
            best_row = row_of[s];
        }
    }
    if (best_row >= ES_NUM_KERNELS) best_row = ES_NUM_KERNELS - 1;
    const EsKernelRow* k = &es_kernel_db[best_row];
    *support = k->support;
    *beta_over_support = k->beta_over_support;
    *grid_size = grid_size_for(image_size, k->sigma);
}

/* P_n(x) and, for |x| < 1, P_n'(x), by the three-term recurrence
 * (es_fft_utils.cpp:13-57). */
static double legendre(double x, int n, double* derivative)
{
    if (n == 0) { *derivative = 0.0; return 1.0; }
    if (n == 1) { *derivative = 1.0; return x; }
    double older = 1.0, prev = x, cur = 0.0;
    for (int i = 2; i <= n; ++i)
    {
        cur = x * prev + (((double) i - 1.0) / (double) i) * (x * prev - older);
        older = prev;
        prev = cur;
    }
    if (x > -1.0 && x < 1.0)
        *derivative = (double) n / (1.0 - (x * x)) * (older - x * cur);
    return cur;
}

/* i-th largest root of P_n by Newton from Tricomi's estimate, and its
 * quadrature weight (es_fft_utils.cpp:60-110). The stopping rule is the
 * reference's: fdim(next, current) > 1e-16, at most 100 iterations. */
static double legendre_root(int i, int n, double* weight)
{
    const double shrink = 1.0 - (1.0 - 1.0 / (double) n) / (8.0 * n * n);
    double next = shrink * cos(M_PI * (4.0 * (double) i - 1.0) /
            (4.0 * n + 2.0));
    double cur, slope = 1.0;
    int iterations = 0;
    do
    {
        cur = next;
        const double value = legendre(cur, n, &slope);
        next = cur - value / slope;
        iterations++;
    }
    while (fdim(next, cur) > 1e-16 && iterations < 100);
    (void) legendre(next, n, &slope);
    *weight = 2.0 / ((1.0 - next * next) * slope * slope);
    return next;
}

double es_correction_tables(int image_size, int grid_size, int support,
        double beta, double* qk, double* qn, double* qw, double* conv_corr)
{
    /* es_fft_utils.cpp:122-175 */
    const int p = (int) (1.5 * support + 2.0);
    memset(qk, 0, sizeof(double) * ES_QUADRATURE_BOUND);
    memset(qn, 0, sizeof(double) * ES_QUADRATURE_BOUND);
    memset(qw, 0, sizeof(double) * ES_QUADRATURE_BOUND);
    for (int i = 1; i <= p; ++i)
    {
        double w = 0.0;
        const double x = legendre_root(i, 2 * p, &w);
        qn[i - 1] = x;
        qw[i - 1] = w;
        qk[i - 1] = exp(beta * (sqrt(1.0 - x * x) - 1.0));
    }
    double norm = 0.0;
    for (int i = 0; i < p; ++i) norm += qk[i] * qw[i];
    norm *= (double) support;
    if (conv_corr)
    {
        for (int m = 0; m <= image_size / 2; ++m)
        {
            const double l = ((double) m) * (1.0 / grid_size);
            double c = 0.0;
            for (int i = 0; i < p; ++i)
                c += qk[i] * qw[i] * cos(M_PI * l * support * qn[i]);
            conv_corr[m] = (double) (c * support) / norm;
        }
    }
    return norm;
}

void es_plan_scalars(int num_rows, int num_chan, int image_size, int is_double,
        double pixel_size, double epsilon, double min_abs_w, double max_abs_w,
        int do_wstacking, sdp_EsPlanInfo* info)
{
    memset(info, 0, sizeof(*info));
    info->num_rows = num_rows;
    info->num_chan = num_chan;
    info->image_size = image_size;
    info->is_double = is_double ? 1 : 0;
    info->do_wstacking = do_wstacking ? 1 : 0;
    info->epsilon = epsilon;
    info->pixel_size = pixel_size;
    int grid_size = 0, support = 0;
    double beta_row = NAN;
    es_select_kernel(epsilon, image_size, is_double, &grid_size, &support,
            &beta_row);
    info->grid_size = grid_size;
    info->support = support;
    info->beta = beta_row * support;           /* es_fft.cpp:335 */
    info->uv_scale = grid_size * pixel_size;   /* es_fft.cpp:342 */
    if (do_wstacking)
    {
        /* es_fft.cpp:346-373: plane spacing from the largest |n - 1| in the
         * image (its corner), then the lattice is stretched so the outermost
         * planes sit (W/2 - 1) spacings beyond [min_abs_w, max_abs_w]. */
        const double x0 = -0.5 * image_size * pixel_size;
        const double y0 = -0.5 * image_size * pixel_size;
        double nmin = sqrt(fmax(1.0 - x0 * x0 - y0 * y0, 0.0)) - 1.0;
        if (x0 * x0 + y0 * y0 > 1.0)
            nmin = -sqrt(fabs(1.0 - x0 * x0 - y0 * y0)) - 1.0;
        double w_scale = 0.25 / fabs(nmin);
        int planes = (int) ((max_abs_w - min_abs_w) / w_scale + 2);
        if (max_abs_w > min_abs_w)
        {
            w_scale = 1.0 / ((1.0 + 1e-13) * (max_abs_w - min_abs_w) /
                    (planes - 1));
        }
        else
        {
            /* A single |w| value: the reference divides by zero here
             * (es_fft.cpp:358). Keep the unstretched spacing instead. */
            planes = 2;
            w_scale = 1.0 / w_scale;
        }
        info->min_plane_w = min_abs_w - (0.5 * support - 1.0) / w_scale;
        info->max_plane_w = max_abs_w + (0.5 * support - 1.0) / w_scale;
        info->num_total_w_grids = planes + support - 2;
        info->min_abs_w = min_abs_w;
        info->max_abs_w = max_abs_w;
        info->w_scale = w_scale;
        info->inv_w_range = info->max_plane_w - info->min_plane_w;
    }
    else
    {
        /* es_fft.cpp:374-383 */
        info->num_total_w_grids = 1;
        info->inv_w_range = 1.0;
        info->w_scale = 1.0;
    }
    info->inv_w_scale = 1.0 / info->w_scale;
}
