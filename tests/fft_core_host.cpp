// Host emulation of the in-shared-memory line FFT (csrc/es_fft_core.h + es_fft_plan.c):
// the same butterflies, tables and digit-reversal map the kernels use, run one
// "thread" at a time. Writes input then output (complex128) to a file that
// tests/test_fft_core.py compares with numpy.fft.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cmath>
#include "es_fft_core.h"
#include "es_fft_plan.h"
template<typename T> struct C2 { T x, y; };
template<typename T, int SGN, bool DIT> int run(int n, int maxr, const char* out)
{
    EsFftLine pl;
    if (es_fft_line_plan(n, maxr, &pl)) { printf("unsupported %d\n", n); return 1; }
    std::vector<double> tw(2 * pl.num_twiddles); std::vector<int32_t> perm(n);
    es_fft_line_tables(&pl, tw.data(), perm.data());
    std::vector<C2<T>> twt(pl.num_twiddles), x(n), in(n);
    for (int i = 0; i < pl.num_twiddles; ++i) { twt[i].x = (T) tw[2*i]; twt[i].y = (T) tw[2*i+1]; }
    srand(n);
    for (int i = 0; i < n; ++i) { in[i].x = (T)(rand() / (double) RAND_MAX - 0.5); in[i].y = (T)(rand() / (double) RAND_MAX - 0.5); }
    if (DIT) for (int i = 0; i < n; ++i) x[perm[i]] = in[i]; else x = in;
    const int nt = 512;
    for (int s = 0; s < pl.num_stages; ++s) {
        const int stage = DIT ? pl.num_stages - 1 - s : s;
        for (int tid = 0; tid < nt; ++tid)
            es_fft_pass<SGN, DIT, T>(x.data(), pl, stage, twt.data(), 1, n, tid, nt);
    }
    FILE* f = fopen(out, "wb");
    for (int i = 0; i < n; ++i) { double v[2] = {(double) in[i].x, (double) in[i].y}; fwrite(v, 8, 2, f); }
    for (int i = 0; i < n; ++i) { C2<T> r = DIT ? x[i] : x[perm[i]]; double v[2] = {(double) r.x, (double) r.y}; fwrite(v, 8, 2, f); }
    fclose(f);
    printf("n=%d stages:", n); for (int s = 0; s < pl.num_stages; ++s) printf(" %d", pl.stage[s].radix); printf(" tw=%d\n", pl.num_twiddles);
    return 0;
}
int main(int argc, char** argv)
{
    int n = atoi(argv[1]); int dbl = atoi(argv[2]); int sgn = atoi(argv[3]); int dit = atoi(argv[4]);
    const char* out = argv[5];
    if (dbl) { if (sgn > 0) return dit ? run<double, 1, true>(n, 8, out) : run<double, 1, false>(n, 8, out);
               else return dit ? run<double, -1, true>(n, 8, out) : run<double, -1, false>(n, 8, out); }
    if (sgn > 0) return dit ? run<float, 1, true>(n, 16, out) : run<float, 1, false>(n, 16, out);
    return dit ? run<float, -1, true>(n, 16, out) : run<float, -1, false>(n, 16, out);
}
