// Experiment (not product code): how fast is cuFFT on one w-plane of size G for
// the layouts the gridder could use? Prints ms per plane and effective GB/s
// counted as 2 traversals (read + write) of the data each pass touches.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#include <cufft.h>

#define CK(x) do { auto e = (x); if (e != 0) { printf("fail %s = %d line %d\n", #x, (int) e, __LINE__); exit(1); } } while (0)

static float time_exec(cufftHandle h, cufftComplex* in, cufftComplex* out, int reps)
{
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    CK(cufftExecC2C(h, in, out, CUFFT_INVERSE));
    cudaDeviceSynchronize();
    cudaEventRecord(a);
    for (int i = 0; i < reps; ++i) CK(cufftExecC2C(h, in, out, CUFFT_INVERSE));
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    return ms / reps;
}

int main(int argc, char** argv)
{
    int G = argc > 1 ? atoi(argv[1]) : 6160;
    int N = argc > 2 ? atoi(argv[2]) : 4096;
    int B = argc > 3 ? atoi(argv[3]) : 8;
    int reps = 5;
    size_t cells = (size_t) G * G;
    cufftComplex *x, *y;
    CK(cudaMalloc(&x, cells * B * sizeof(cufftComplex)));
    CK(cudaMalloc(&y, cells * B * sizeof(cufftComplex)));
    cudaMemset(x, 0, cells * B * sizeof(cufftComplex));
    double plane_gb = cells * 8.0 / 1e9;
    printf("G=%d N=%d batch=%d plane=%.3f GB\n", G, N, B, plane_gb);
    {   // 1. 2-D in place, batch B
        cufftHandle h; int n[2] = {G, G}; size_t ws;
        CK(cufftCreate(&h));
        CK(cufftMakePlanMany(h, 2, n, 0, 1, 0, 0, 1, 0, CUFFT_C2C, B, &ws));
        float ms = time_exec(h, x, x, reps) / B;
        printf("2d in-place            : %.3f ms/plane  (%.0f GB/s at 4 traversals) ws=%.2f GB\n", ms, 4 * plane_gb / ms * 1e3, ws / 1e9);
        cufftDestroy(h);
    }
    {   // 2. rows only (contiguous), in place, batch G*B
        cufftHandle h; int n[1] = {G}; size_t ws;
        CK(cufftCreate(&h));
        CK(cufftMakePlanMany(h, 1, n, 0, 1, G, 0, 1, G, CUFFT_C2C, G * B, &ws));
        float ms = time_exec(h, x, x, reps) / B;
        printf("1d rows in-place       : %.3f ms/plane  (%.0f GB/s at 2 traversals) ws=%.2f GB\n", ms, 2 * plane_gb / ms * 1e3, ws / 1e9);
        cufftDestroy(h);
    }
    {   // 3. rows, out of place, transposed output (ostride=G, odist=1), per plane
        cufftHandle h; int n[1] = {G}; size_t ws;
        int inembed[1] = {G}, onembed[1] = {G};
        CK(cufftCreate(&h));
        CK(cufftMakePlanMany(h, 1, n, inembed, 1, G, onembed, G, 1, CUFFT_C2C, G, &ws));
        float ms = time_exec(h, x, y, reps);
        printf("1d rows -> transposed  : %.3f ms/plane  (%.0f GB/s at 2 traversals) ws=%.2f GB\n", ms, 2 * plane_gb / ms * 1e3, ws / 1e9);
        cufftDestroy(h);
    }
    {   // 4. columns strided in place, only N of them (istride=G, idist=1)
        cufftHandle h; int n[1] = {G}; size_t ws;
        int inembed[1] = {G};
        CK(cufftCreate(&h));
        CK(cufftMakePlanMany(h, 1, n, inembed, G, 1, inembed, G, 1, CUFFT_C2C, N, &ws));
        float ms = time_exec(h, x, x, reps);
        printf("1d cols strided (N)    : %.3f ms/plane  (%.0f GB/s at 2 traversals of N*G) ws=%.2f GB\n", ms, 2 * plane_gb * N / G / ms * 1e3, ws / 1e9);
        cufftDestroy(h);
    }
    {   // 5. rows of a pruned N x G array (what pass 2 would be after a transpose)
        cufftHandle h; int n[1] = {G}; size_t ws;
        CK(cufftCreate(&h));
        CK(cufftMakePlanMany(h, 1, n, 0, 1, G, 0, 1, G, CUFFT_C2C, N * B, &ws));
        float ms = time_exec(h, x, x, reps) / B;
        printf("1d rows pruned (N rows): %.3f ms/plane  (%.0f GB/s at 2 traversals of N*G)\n", ms, 2 * plane_gb * N / G / ms * 1e3);
        cufftDestroy(h);
    }
    {   // 6. plain device copy for reference
        cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
        cudaMemcpy(y, x, cells * B * 8, cudaMemcpyDeviceToDevice);
        cudaEventRecord(a);
        for (int i = 0; i < reps; ++i) cudaMemcpyAsync(y, x, cells * B * 8, cudaMemcpyDeviceToDevice);
        cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b); ms /= reps * B;
        printf("d2d copy               : %.3f ms/plane  (%.0f GB/s read+write)\n", ms, 2 * plane_gb / ms * 1e3);
    }
    return 0;
}
