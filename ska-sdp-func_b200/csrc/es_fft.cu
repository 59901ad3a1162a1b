/*
 * Pruned w-plane FFT of the ES gridder for sm_100a: the 2-D transform of one
 * G x G w-plane, of which only the N x N image footprint is needed (gridding)
 * or non-zero (degridding), as two passes of whole-line FFTs held in shared
 * memory (es_fft_core.h), fused with what surrounds them. T = strip[q][u][v']
 * is the pruned intermediate (G rows of N elements per plane).
 *
 *   gridding   pass 1  es_fft_grid_rows_kernel: one plane row (along v) per
 *                      CTA -> inverse FFT -> the N needed outputs -> T[u][:]
 *              pass 2  es_fft_grid_cols_screen_kernel: a bundle of 1/2/4
 *                      adjacent columns of T (along u) per CTA -> inverse FFT
 *                      -> the N needed outputs x w-screen -> summed into
 *                      imageT[col][row] with fire-and-forget reductions
 *   degridding pass 1  es_fft_degrid_cols_kernel: a bundle of imageT columns
 *                      x conj w-screen, zero-padded -> forward FFT (along u)
 *                      -> T[:][cols]
 *              pass 2  es_fft_degrid_rows_kernel: T[u][:] zero-padded ->
 *                      forward FFT (along v) -> plane row
 *
 * The first radix pass of every line takes its operands straight from global
 * memory into registers and the last one stores straight from registers, so
 * a line crosses shared memory once per inner pass only; the row kernels are
 * fully coalesced, the column kernels move 8 x lines-byte pieces (a bundle of
 * 4 fp32 columns fills 32-byte sectors).
 *
 * Replaces, per plane, the reference's cufftExecC2C/Z2Z on the full plane
 * (sdp_fft.cpp:425-456) + apply_w_screen_and_sum / reverse_w_screen_to_stack
 * (kernels.cu:429-679) with 2 G^2 + 2 G N (+ image) element traversals
 * instead of ~10 G^2 + 2 N^2 (cuFFT's 2-D plan needs five passes on these
 * 11-smooth sizes, profiles/r01_launches_c2_grid_v1.csv).
 * The arithmetic is the same DFT; the w-screen maths is that of
 * kernels.cu:110-123 with the phase reduced in fp64.
 */
#include "es_fft_kernels.cuh"

/* ================================================================ C ABI */

namespace {

/* Elements between the lines of a bundle: >= padded line, and
 * pitch mod 16 == 16 / lines so that the 16 threads of a shared-memory phase
 * (lines x consecutive elements) fall into distinct 8-byte banks. */
int bundle_pitch(const EsFftLine* line, int lshift)
{
    int pitch = line->padded_n;
    if (lshift == 0) return pitch + (pitch & 1); /* 16-byte aligned lines */
    const int want = 16 >> lshift;
    while (pitch % 16 != want) pitch++;
    return pitch;
}

void make_cfg(const EsFftLine* line, size_t cplx, int lshift, int threads,
        int max_threads, EsFftCfg* c)
{
    c->lshift = lshift;
    c->pitch = bundle_pitch(line, lshift);
    c->tw_global = 0;
    c->smem_bytes = ((size_t) c->pitch << lshift) * cplx +
            (size_t) (line->num_twiddles + 1) * cplx;
    /* tests: SKA_SDP_FUNC_ES_FFT_TW_GLOBAL=1 forces this mode at any size */
    const bool force = getenv("SKA_SDP_FUNC_ES_FFT_TW_GLOBAL") != 0;
    if (force || (c->smem_bytes > ES_FFT_MAX_SMEM && lshift == 0 &&
            (size_t) c->pitch * cplx <= ES_FFT_MAX_SMEM))
    {
        /* e.g. fp64 lines of 14400 elements (230 KB): everything but the
         * line itself leaves shared memory */
        c->tw_global = 1;
        c->smem_bytes = ((size_t) c->pitch << lshift) * cplx;
    }
    if (threads <= 0)
    {
        /* as many CTAs per SM as shared memory allows (at most 4), all of
         * them together max_threads (512 threads of 128 registers; more
         * where a specialised kernel set needs fewer registers) */
        int ctas = (int) ((size_t) ES_FFT_MAX_SMEM / (c->smem_bytes + 1024));
        if (ctas < 1) ctas = 1;
        if (ctas > 4) ctas = 4;
        threads = ((max_threads / ctas) / 32) * 32;
        /* no more threads than butterflies of the narrowest pass */
        int min_bf = line->n;
        for (int i = 0; i < line->num_stages; ++i)
        {
            const int bf = (line->n / line->stage[i].radix) << lshift;
            if (bf < min_bf) min_bf = bf;
        }
        while (threads > 32 && threads / 2 >= min_bf) threads /= 2;
    }
    c->threads = threads;
}

} // namespace

/* Kernel sets for fixed stage lists (es_fft_seq.cu, one object per list):
 * whether set `id` is for this precision and line plan, and its launch table
 * (0 when the set was not built into this library). */
int es_fft_seq_match(int id, const EsFftLine* line, int is_double);
const EsFftOps* es_fft_seq_ops(int id);

extern "C" {

int es_fft_create(int grid_size, int image_size, int is_double, EsFft** out)
{
    *out = 0;
    if (getenv("SKA_SDP_FUNC_ES_USE_CUFFT")) return -1;
    EsFft* f = (EsFft*) calloc(1, sizeof(EsFft));
    if (!f) return (int) cudaErrorMemoryAllocation;
    /* experiments: SKA_SDP_FUNC_ES_FFT_RADICES="28,20,11" forces the passes */
    int32_t forced[ES_FFT_MAX_STAGES + 1];
    memset(forced, 0, sizeof(forced));
    const char* renv = getenv("SKA_SDP_FUNC_ES_FFT_RADICES");
    if (renv)
    {
        char buf[128];
        strncpy(buf, renv, sizeof(buf) - 1);
        buf[sizeof(buf) - 1] = 0;
        int c = 0;
        for (char* p = strtok(buf, ","); p && c < ES_FFT_MAX_STAGES;
                p = strtok(0, ","))
            forced[c++] = atoi(p);
    }
    const int max_radix = is_double ? ES_FFT_MAX_RADIX_DOUBLE : 32;
    int bad = es_fft_line_plan(grid_size, max_radix, forced[0] ? forced : 0,
            &f->line);
    if (bad && forced[0])
        bad = es_fft_line_plan(grid_size, max_radix, 0, &f->line);
    if (!bad && is_double)
        for (int i = 0; i < f->line.num_stages; ++i)
            if (f->line.stage[i].radix > ES_FFT_MAX_RADIX_DOUBLE) bad = 1;
    if (bad || (grid_size & 1) || (image_size & 1) || image_size > grid_size ||
            grid_size > ES_FFT_OCC_WORDS * 32 * ES_TILE)
    {
        free(f);
        return -1; /* unsupported size: caller falls back to cuFFT */
    }
    f->grid_size = grid_size;
    f->image_size = image_size;
    f->is_double = is_double;
    /* kernel set: the one instantiated for exactly this stage list where
     * there is one (SKA_SDP_FUNC_ES_FFT_GENERIC=1: tests of the generic set) */
    f->ops = is_double ? EsFftLaunch<double, 0>::ops() :
            EsFftLaunch<float, 0>::ops();
    f->seq = 0;
    if (!getenv("SKA_SDP_FUNC_ES_FFT_GENERIC"))
    {
        for (int id = 1; id <= ES_FFT_NUM_SEQ; ++id)
        {
            const EsFftOps* ops = es_fft_seq_match(id, &f->line, is_double) ?
                    es_fft_seq_ops(id) : 0;
            if (ops)
            {
                f->ops = ops;
                f->seq = id;
                break;
            }
        }
    }
    const size_t cplx = is_double ? 16 : 8;
    /* experiments: SKA_SDP_FUNC_ES_FFT_CFG="rows_threads,cols_threads,cols_lshift"
     * default: bundles of 4 (fp64: 2) adjacent columns / 2 (fp64: 1) rows per CTA where they fit
     * in shared memory (32-byte pieces of the row-major intermediate) */
    int e[4] = {0, 0, is_double ? 1 : 2, is_double ? 0 : 1};
    const char* env = getenv("SKA_SDP_FUNC_ES_FFT_CFG");
    int rows_forced = 0;
    if (env) rows_forced = sscanf(env, "%d,%d,%d,%d", &e[0], &e[1], &e[2],
            &e[3]) == 4;
    if (e[3] < 0 || e[3] > 2) e[3] = 0;
    for (int k = 0; k < 2; ++k)
        if (e[k] < 0 || e[k] > f->ops->max_threads || (e[k] & 31)) e[k] = 0;
    if (e[2] < 0 || e[2] > 2) e[2] = 0;
    /* rows of one CTA share a 64-row occupancy block and must tile the grid */
    int rshift = e[3];
    for (;; --rshift)
    {
        make_cfg(&f->line, cplx, rshift, e[0], f->ops->max_threads, &f->rows);
        /* default: a row bundle only where two CTAs still fit on an SM (C5:
         * 2 rows of 11520 leave one 512-thread CTA: rows 1365 vs 1314 ms);
         * a bundle forced by the environment only has to fit */
        const size_t need = rows_forced ? f->rows.smem_bytes :
                2 * (f->rows.smem_bytes + 1024);
        if ((need <= ES_FFT_MAX_SMEM && grid_size % (1 << rshift) == 0) ||
                rshift == 0) break;
    }
    int lshift = e[2];
    for (;; --lshift)
    {
        make_cfg(&f->line, cplx, lshift, e[1], f->ops->max_threads, &f->cols);
        if ((f->cols.smem_bytes <= ES_FFT_MAX_SMEM &&
                image_size % (1 << lshift) == 0) || lshift == 0) break;
    }
    f->plane_elems = (int64_t) grid_size * image_size;
    if (f->rows.smem_bytes > ES_FFT_MAX_SMEM ||
            f->cols.smem_bytes > ES_FFT_MAX_SMEM)
    {
        free(f);
        return -1;
    }
    if (getenv("SKA_SDP_FUNC_ES_DEBUG"))
    {
        fprintf(stderr, "es_fft: kernel set %d (0: generic); n=%d passes",
                f->seq, grid_size);
        for (int k = 0; k < f->line.num_stages; ++k)
            fprintf(stderr, " %d", f->line.stage[k].radix);
        fprintf(stderr, "; pad %d; rows %d thr (%zu B), cols %d thr x %d lines "
                "(%zu B)\n", f->line.pad, f->rows.threads,
                f->rows.smem_bytes, f->cols.threads, 1 << f->cols.lshift,
                f->cols.smem_bytes);
    }

    const int nt = f->line.num_twiddles;
    const int nb = grid_size / f->line.stage[f->line.num_stages - 1].radix;
    double* tw = (double*) malloc(sizeof(double) * 2 * (size_t) (nt + 1));
    int32_t* bpos = (int32_t*) malloc(sizeof(int32_t) * (size_t) nb);
    void* tw_t = malloc(cplx * (size_t) (nt + 1));
    int err = 0;
    if (!tw || !bpos || !tw_t) err = (int) cudaErrorMemoryAllocation;
    if (!err)
    {
        memset(tw, 0, sizeof(double) * 2 * (size_t) (nt + 1));
        es_fft_line_tables(&f->line, tw, 0, bpos);
        for (int i = 0; i < 2 * (nt + 1); ++i)
        {
            if (is_double) ((double*) tw_t)[i] = tw[i];
            else ((float*) tw_t)[i] = (float) tw[i];
        }
        err = (int) cudaMalloc(&f->d_tw, cplx * (size_t) (nt + 1));
        if (!err) err = (int) cudaMalloc((void**) &f->d_bpos,
                sizeof(int32_t) * (size_t) nb);
        if (!err) err = (int) cudaMemcpy(f->d_tw, tw_t, cplx * (size_t) (nt + 1),
                cudaMemcpyHostToDevice);
        if (!err) err = (int) cudaMemcpy(f->d_bpos, bpos,
                sizeof(int32_t) * (size_t) nb, cudaMemcpyHostToDevice);
    }
    free(tw); free(bpos); free(tw_t);
    if (!err) err = f->ops->prepare(ES_FFT_MAX_SMEM);
    if (!err)
    {
        /* persistent row kernels: as many CTAs as the device holds at once */
        int dev = 0, sms = 0, per_sm = 0;
        err = (int) cudaGetDevice(&dev);
        if (!err) err = (int) cudaDeviceGetAttribute(&sms,
                cudaDevAttrMultiProcessorCount, dev);
        if (!err) err = f->ops->rows_per_sm(f->rows.threads,
                f->rows.smem_bytes, &per_sm);
        if (per_sm < 1) per_sm = 1;
        f->rows_grid = sms * per_sm;
        if (!err) err = (int) cudaMalloc((void**) &f->d_row_work,
                2 * sizeof(uint32_t));
    }
    if (err)
    {
        (void) cudaGetLastError();
        es_fft_destroy(f);
        return err;
    }
    *out = f;
    return 0;
}


void es_fft_destroy(EsFft* f)
{
    if (!f) return;
    if (f->d_tw) (void) cudaFree(f->d_tw);
    if (f->d_bpos) (void) cudaFree(f->d_bpos);
    if (f->d_row_list) (void) cudaFree(f->d_row_list);
    if (f->d_row_work) (void) cudaFree(f->d_row_work);
    if (f->d_nm1) (void) cudaFree(f->d_nm1);
    free(f);
}

size_t es_fft_strip_bytes(const EsFft* f, int nplanes)
{
    return (size_t) nplanes * (size_t) f->plane_elems * (f->is_double ? 16 : 8);
}

int es_fft_grid_rows(const EsFft* f, const void* planes, int nplanes,
        void* strip, const uint32_t* occ, es_stream_t stream)
{
    return f->ops->grid_rows(const_cast<EsFft*>(f), planes, nplanes, strip,
            occ, (cudaStream_t) stream);
}

int es_fft_grid_cols_screen(const EsFft* f, const EsGeom* g, const void* strip,
        int first_plane, int nplanes, void* image_t, const uint32_t* occ,
        es_stream_t stream)
{
    return f->ops->grid_cols_screen(f, g, strip, first_plane, nplanes, image_t,
            occ, (cudaStream_t) stream);
}

int es_fft_degrid_cols(const EsFft* f, const EsGeom* g, const void* image_t,
        int first_plane, int nplanes, void* strip, const uint32_t* occ,
        es_stream_t stream)
{
    return f->ops->degrid_cols(f, g, image_t, first_plane, nplanes, strip, occ,
            (cudaStream_t) stream);
}

int es_fft_degrid_rows(const EsFft* f, const void* strip, int nplanes,
        void* planes, const uint32_t* occ, es_stream_t stream)
{
    return f->ops->degrid_rows(const_cast<EsFft*>(f), strip, nplanes, planes,
            occ, (cudaStream_t) stream);
}

int es_fft_image_transpose(const EsFft* f, const void* src, void* dst,
        int add, int wstack, es_stream_t stream)
{
    cudaStream_t s = (cudaStream_t) stream;
    const int n = f->image_size;
    const double* tab = wstack ? f->d_nm1 : 0;
    if (wstack && !tab) return (int) cudaErrorInvalidValue;
    dim3 block(32, 8), grid((unsigned) ((n + 31) / 32), (unsigned) ((n + 31) / 32));
    if (f->is_double)
    {
        if (add) es_transpose_screen_kernel<double, true><<<grid, block, 0, s>>>(
                (const double*) src, (double*) dst, n, tab, f->nm1_pitch);
        else es_transpose_screen_kernel<double, false><<<grid, block, 0, s>>>(
                (const double*) src, (double*) dst, n, tab, f->nm1_pitch);
    }
    else
    {
        if (add) es_transpose_screen_kernel<float, true><<<grid, block, 0, s>>>(
                (const float*) src, (float*) dst, n, tab, f->nm1_pitch);
        else es_transpose_screen_kernel<float, false><<<grid, block, 0, s>>>(
                (const float*) src, (float*) dst, n, tab, f->nm1_pitch);
    }
    return (int) cudaGetLastError();
}

int es_fft_set_screen(EsFft* f, double pixel_size, double max_abs_w)
{
    const int count = f->image_size / 2 + 1;
    const int pitch = (count + 3) & ~3;
    if (!f->d_nm1)
    {
        cudaError_t err = cudaMalloc((void**) &f->d_nm1,
                sizeof(double) * (size_t) count * (size_t) pitch);
        if (err != cudaSuccess) return (int) err;
    }
    f->nm1_pitch = pitch;
    const double pix2 = pixel_size * pixel_size;
    /* largest |w (n - 1)| of the plan, in turns: the corner pixel */
    const double half = 0.5 * (double) f->image_size;
    const double sos = 2.0 * pix2 * half * half;
    const double nm1_max = sos / (sqrt(sos < 1.0 ? 1.0 - sos : 0.0) + 1.0);
    const int bits = es_screen_fixed_bits(fabs(max_abs_w) * nm1_max);
    f->screen_shift = 32 - bits;
    f->screen_magic = es_screen_magic(bits);
    dim3 grid((unsigned) ((count + 255) / 256), (unsigned) count);
    es_screen_table_kernel<<<grid, 256>>>(f->d_nm1, count, pitch, pix2);
    cudaError_t err = cudaGetLastError();
    if (err == cudaSuccess) err = cudaDeviceSynchronize();
    return (int) err;
}

} // extern "C"
