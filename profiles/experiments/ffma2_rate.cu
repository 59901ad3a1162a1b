// Issue-rate microbenchmark on sm_100a: scalar FFMA / FADD vs packed
// FFMA2 / FADD2 (fma.rn.f32x2, add.rn.f32x2). Prints flop/clk/SM for each.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2_rate ffma2_rate.cu
#include <cstdio>
#include <cuda_runtime.h>

#define ITERS 4096
#define CHAINS 8

template<int MODE>
__global__ void __launch_bounds__(256) rate_kernel(float2* out, float2 a, float2 b)
{
    float2 v[CHAINS];
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) { v[c].x = threadIdx.x + c; v[c].y = c - 1.f; }
    for (int it = 0; it < ITERS; ++it)
    {
#pragma unroll
        for (int c = 0; c < CHAINS; ++c)
        {
            if (MODE == 0) { v[c].x = fmaf(v[c].x, a.x, b.x); v[c].y = fmaf(v[c].y, a.y, b.y); }
            if (MODE == 1) v[c] = __ffma2_rn(v[c], a, b);
            if (MODE == 2) { v[c].x = v[c].x + a.x; v[c].y = v[c].y + a.y; }
            if (MODE == 3) v[c] = __fadd2_rn(v[c], a);
            if (MODE == 4) { v[c].x = v[c].x * a.x; v[c].y = v[c].y * a.y; }
            if (MODE == 5) v[c] = __fmul2_rn(v[c], a);
            /* complex multiply by (a.x, a.y), interleaved (re, im) layout */
            if (MODE == 6) { const float2 t = v[c];
                v[c].x = fmaf(t.x, a.x, -t.y * a.y); v[c].y = fmaf(t.x, a.y, t.y * a.x); }
            if (MODE == 7) { const float2 t = v[c];
                v[c] = __ffma2_rn(t, make_float2(a.x, a.x),
                        __fmul2_rn(make_float2(t.y, t.x), make_float2(-a.y, a.y))); }
        }
    }
    float2 s = v[0];
#pragma unroll
    for (int c = 1; c < CHAINS; ++c) { s.x += v[c].x; s.y += v[c].y; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template<int MODE> void run(const char* name, float2* out, int sms, double clk_hz)
{
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int blocks = sms * 8;
    float2 a = {1.0001f, 0.9999f}, b = {1e-3f, -1e-3f};
    rate_kernel<MODE><<<blocks, 256>>>(out, a, b);
    cudaEventRecord(e0);
    rate_kernel<MODE><<<blocks, 256>>>(out, a, b);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double ops = (double) blocks * 256 * ITERS * CHAINS * 2; // scalar ops
    printf("%-8s %.3f ms  %.1f scalar-op/clk/SM (at %.0f MHz)\n", name, ms,
            ops / (ms * 1e-3) / clk_hz / sms, clk_hz / 1e6);
}

int main()
{
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    float2* out; cudaMalloc(&out, sizeof(float2) * p.multiProcessorCount * 8 * 256);
    const double hz = khz * 1e3;
    run<0>("FFMA", out, p.multiProcessorCount, hz);
    run<1>("FFMA2", out, p.multiProcessorCount, hz);
    run<2>("FADD", out, p.multiProcessorCount, hz);
    run<3>("FADD2", out, p.multiProcessorCount, hz);
    run<4>("FMUL", out, p.multiProcessorCount, hz);
    run<5>("FMUL2", out, p.multiProcessorCount, hz);
    // per chain step: 4 scalar instructions / 2 packed instructions (counted as 2 "scalar ops" like the rest)
    run<6>("CMUL", out, p.multiProcessorCount, hz);
    run<7>("CMUL2", out, p.multiProcessorCount, hz);
    return 0;
}
