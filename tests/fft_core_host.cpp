// Host emulation of the in-shared-memory line FFT (csrc/es_fft_core.h + es_fft_plan.c):
// the same passes, butterflies, tables, padding and digit-reversal maps the
// kernels use, run one "thread" at a time over `lines` interleaved lines with
// the first pass reading from / the last pass writing to "global" arrays
// through functors, exactly like es_fft.cu. Writes input then output
// (complex128, line after line) to a file that tests/test_fft_core.py compares
// with numpy.fft.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cmath>
#include "es_fft_core.h"
#include "es_fft_plan.h"
template<typename T> struct C2 { T x, y; };
template<typename T> struct InF
{
    struct Cursor { const C2<T>* p; int gs; };
    const C2<T>* src; int n;
    Cursor cursor(int line, int idx0, int gs) const { return Cursor{src + (size_t) line * n + idx0, gs}; }
    typedef C2<T> Raw;
    C2<T> fetch(const Cursor& c, int k) const { return c.p[k * c.gs]; }
    C2<T> finish(const Cursor&, int, const C2<T> v) const { return v; }
    static constexpr bool skip_zero_butterflies = false;
    void butterfly_begin() const {}
    bool butterfly_nonzero() const { return true; }
    unsigned int any_bits() const { return 0u; }
    void merge_any(unsigned int) {}
    void finish2(const Cursor&, int, C2<T> r0, C2<T> r1, C2<T>& v0, C2<T>& v1) const
    {
        v0 = r0;
        v1 = r1;
    }
};
template<typename T> struct OutF
{
    struct Cursor { C2<T>* p; int gs; };
    C2<T>* dst; int n;
    Cursor cursor(int line, int idx0, int gs) const { return Cursor{dst + (size_t) line * n + idx0, gs}; }
    void store(const Cursor& c, int k, C2<T> v) const { c.p[k * c.gs] = v; }
    struct Pre {};
    static constexpr bool skip_masked_butterflies = false;
    Pre prefetch(const Cursor&, int) const { return Pre(); }
    void store1(const Cursor& c, int k, C2<T> v, Pre) const { store(c, k, v); }
    void store2(const Cursor& c, int k, C2<T> v0, C2<T> v1, Pre, Pre) const
    {
        store(c, k, v0);
        store(c, k + 1, v1);
    }
};
template<typename T, int SGN, bool DIT> int run(int n, int maxr, int lshift,
        const int32_t* forced, const char* out)
{
    EsFftLine pl;
    if (es_fft_line_plan(n, maxr, forced, &pl)) { printf("unsupported %d\n", n); return 1; }
    const int lines = 1 << lshift;
    std::vector<double> tw(2 * pl.num_twiddles + 2);
    std::vector<int32_t> bpos(n / pl.stage[pl.num_stages - 1].radix);
    es_fft_line_tables(&pl, tw.data(), 0, bpos.data());
    const int pitch = pl.padded_n + 3;
    std::vector<C2<T>> twt(pl.num_twiddles + 1), x((size_t) lines * pitch), in((size_t) lines * n),
            res((size_t) lines * n);
    for (int i = 0; i < pl.num_twiddles; ++i) { twt[i].x = (T) tw[2*i]; twt[i].y = (T) tw[2*i+1]; }
    srand(n);
    for (size_t i = 0; i < in.size(); ++i) { in[i].x = (T)(rand() / (double) RAND_MAX - 0.5); in[i].y = (T)(rand() / (double) RAND_MAX - 0.5); }
    // poison the work area: every element a pass reads must have been written
    for (size_t i = 0; i < x.size(); ++i) { x[i].x = (T) NAN; x[i].y = (T) NAN; }
    InF<T> fin{in.data(), n};
    OutF<T> fout{res.data(), n};
    const int nt = 96;
    for (int step = 0; step < pl.num_stages; ++step)
        for (int tid = 0; tid < nt; ++tid)
            es_line_pass<SGN, DIT, T>(x.data(), pl, step, twt.data(), bpos.data(), lshift,
                    pitch, tid, nt, fin, fout);
    FILE* f = fopen(out, "wb");
    for (size_t i = 0; i < in.size(); ++i) { double v[2] = {(double) in[i].x, (double) in[i].y}; fwrite(v, 8, 2, f); }
    for (size_t i = 0; i < res.size(); ++i) { double v[2] = {(double) res[i].x, (double) res[i].y}; fwrite(v, 8, 2, f); }
    fclose(f);
    printf("n=%d stages:", n); for (int s = 0; s < pl.num_stages; ++s) printf(" %d", pl.stage[s].radix);
    printf(" tw=%d pad=%d\n", pl.num_twiddles, pl.pad);
    return 0;
}
// fft_core_host n double sign dit lshift out [r0,r1,...]
int main(int argc, char** argv)
{
    int n = atoi(argv[1]); int dbl = atoi(argv[2]); int sgn = atoi(argv[3]); int dit = atoi(argv[4]);
    int lshift = atoi(argv[5]);
    const char* out = argv[6];
    int32_t forced[ES_FFT_MAX_STAGES + 1]; memset(forced, 0, sizeof(forced));
    if (argc > 7) { int c = 0; for (char* p = strtok(argv[7], ","); p && c < ES_FFT_MAX_STAGES; p = strtok(0, ",")) forced[c++] = atoi(p); }
    const int32_t* fr = forced[0] ? forced : 0;
    if (dbl) { if (sgn > 0) return dit ? run<double, 1, true>(n, 16, lshift, fr, out) : run<double, 1, false>(n, 16, lshift, fr, out);
               else return dit ? run<double, -1, true>(n, 16, lshift, fr, out) : run<double, -1, false>(n, 16, lshift, fr, out); }
    if (sgn > 0) return dit ? run<float, 1, true>(n, 32, lshift, fr, out) : run<float, 1, false>(n, 32, lshift, fr, out);
    return dit ? run<float, -1, true>(n, 32, lshift, fr, out) : run<float, -1, false>(n, 32, lshift, fr, out);
}
