// Experiment (not product code): how fast is cuFFT on one w-plane of size G for
// the layouts the gridder could use? Prints ms per plane and effective GB/s
// counted as 2 traversals (read + write) of the data each pass touches.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a cufft_layouts.cu -lcufft -o cufft_layouts
//   ./cufft_layouts G N batch [double]
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>
#include <cufft.h>

#define CK(x) do { auto e = (x); if (e != 0) { printf("fail %s = %d line %d\n", #x, (int) e, __LINE__); exit(1); } } while (0)

static bool g_double = false;
static size_t g_c = 8;

static float time_exec(cufftHandle h, void* in, void* out, int reps)
{
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    auto run = [&]() {
        if (g_double) CK(cufftExecZ2Z(h, (cufftDoubleComplex*) in, (cufftDoubleComplex*) out, CUFFT_INVERSE));
        else CK(cufftExecC2C(h, (cufftComplex*) in, (cufftComplex*) out, CUFFT_INVERSE));
    };
    run();
    CK(cudaDeviceSynchronize());
    cudaEventRecord(a);
    for (int i = 0; i < reps; ++i) run();
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    return ms / reps;
}

// out[c][r] = in[r][c0 + c] for r < rows, c < cols  (in pitch = ipitch, out pitch = opitch)
template<typename T2>
__global__ void transpose_kernel(const T2* __restrict__ in, T2* __restrict__ out,
        int rows, int cols, int c0, size_t ipitch, size_t opitch)
{
    __shared__ T2 t[32][33];
    const int bx = blockIdx.x * 32, by = blockIdx.y * 32;
    for (int k = threadIdx.y; k < 32; k += blockDim.y)
    {
        const int r = by + k, c = bx + threadIdx.x;
        if (r < rows && c < cols) t[k][threadIdx.x] = in[(size_t) r * ipitch + c0 + c];
    }
    __syncthreads();
    for (int k = threadIdx.y; k < 32; k += blockDim.y)
    {
        const int c = bx + k, r = by + threadIdx.x;
        if (r < rows && c < cols) out[(size_t) c * opitch + r] = t[threadIdx.x][k];
    }
}

int main(int argc, char** argv)
{
    int G = argc > 1 ? atoi(argv[1]) : 6160;
    int N = argc > 2 ? atoi(argv[2]) : 4096;
    int B = argc > 3 ? atoi(argv[3]) : 8;
    g_double = argc > 4 && !strcmp(argv[4], "double");
    g_c = g_double ? 16 : 8;
    cufftType ty = g_double ? CUFFT_Z2Z : CUFFT_C2C;
    int reps = 5;
    size_t cells = (size_t) G * G;
    void *x, *y;
    CK(cudaMalloc(&x, cells * B * g_c));
    CK(cudaMalloc(&y, cells * B * g_c));
    cudaMemset(x, 0, cells * B * g_c);
    cudaMemset(y, 0, cells * B * g_c);
    double plane_gb = cells * (double) g_c / 1e9;
    printf("G=%d N=%d batch=%d %s plane=%.3f GB\n", G, N, B, g_double ? "Z2Z" : "C2C", plane_gb);
    {   // 1. 2-D in place, batch B
        cufftHandle h; int n[2] = {G, G}; size_t ws;
        CK(cufftCreate(&h));
        CK(cufftMakePlanMany(h, 2, n, 0, 1, 0, 0, 1, 0, ty, B, &ws));
        float ms = time_exec(h, x, x, reps) / B;
        printf("2d in-place            : %.3f ms/plane  (%.0f GB/s at 4 traversals) ws=%.2f GB\n", ms, 4 * plane_gb / ms * 1e3, ws / 1e9);
        cufftDestroy(h);
    }
    {   // 2. rows only (contiguous), in place, batch G*B
        cufftHandle h; int n[1] = {G}; size_t ws;
        CK(cufftCreate(&h));
        CK(cufftMakePlanMany(h, 1, n, 0, 1, G, 0, 1, G, ty, G * B, &ws));
        float ms = time_exec(h, x, x, reps) / B;
        printf("1d rows in-place       : %.3f ms/plane  (%.0f GB/s at 2 traversals) ws=%.2f GB\n", ms, 2 * plane_gb / ms * 1e3, ws / 1e9);
        ms = time_exec(h, x, y, reps) / B;
        printf("1d rows out-of-place   : %.3f ms/plane  (%.0f GB/s at 2 traversals)\n", ms, 2 * plane_gb / ms * 1e3);
        cufftDestroy(h);
    }
    {   // 3. rows, out of place, transposed output (ostride=G, odist=1), per plane
        cufftHandle h; int n[1] = {G}; size_t ws;
        int inembed[1] = {G}, onembed[1] = {G};
        CK(cufftCreate(&h));
        CK(cufftMakePlanMany(h, 1, n, inembed, 1, G, onembed, G, 1, ty, G, &ws));
        float ms = time_exec(h, x, y, reps);
        printf("1d rows -> transposed  : %.3f ms/plane  (%.0f GB/s at 2 traversals) ws=%.2f GB\n", ms, 2 * plane_gb / ms * 1e3, ws / 1e9);
        cufftDestroy(h);
    }
    {   // 3b. strided in (columns) -> contiguous out
        cufftHandle h; int n[1] = {G}; size_t ws;
        int inembed[1] = {G}, onembed[1] = {G};
        CK(cufftCreate(&h));
        CK(cufftMakePlanMany(h, 1, n, inembed, G, 1, onembed, 1, G, ty, N, &ws));
        float ms = time_exec(h, x, y, reps);
        printf("1d cols(N) -> rows     : %.3f ms/plane  (%.0f GB/s at 2 traversals of N*G) ws=%.2f GB\n", ms, 2 * plane_gb * N / G / ms * 1e3, ws / 1e9);
        cufftDestroy(h);
    }
    {   // 4. columns strided in place, only N of them (istride=G, idist=1)
        cufftHandle h; int n[1] = {G}; size_t ws;
        int inembed[1] = {G};
        CK(cufftCreate(&h));
        CK(cufftMakePlanMany(h, 1, n, inembed, G, 1, inembed, G, 1, ty, N, &ws));
        float ms = time_exec(h, x, x, reps);
        printf("1d cols strided (N)    : %.3f ms/plane  (%.0f GB/s at 2 traversals of N*G) ws=%.2f GB\n", ms, 2 * plane_gb * N / G / ms * 1e3, ws / 1e9);
        cufftDestroy(h);
    }
    {   // 5. rows of a pruned N x G array (what pass 2 would be after a transpose)
        cufftHandle h; int n[1] = {G}; size_t ws;
        CK(cufftCreate(&h));
        CK(cufftMakePlanMany(h, 1, n, 0, 1, G, 0, 1, G, ty, N * B, &ws));
        float ms = time_exec(h, x, x, reps) / B;
        printf("1d rows pruned (N rows): %.3f ms/plane  (%.0f GB/s at 2 traversals of N*G)\n", ms, 2 * plane_gb * N / G / ms * 1e3);
        cufftDestroy(h);
    }
    {   // 6. plain device copy for reference
        cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
        cudaMemcpy(y, x, cells * B * g_c, cudaMemcpyDeviceToDevice);
        cudaEventRecord(a);
        for (int i = 0; i < reps; ++i) cudaMemcpyAsync(y, x, cells * B * g_c, cudaMemcpyDeviceToDevice);
        cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b); ms /= reps * B;
        printf("d2d copy               : %.3f ms/plane  (%.0f GB/s read+write)\n", ms, 2 * plane_gb / ms * 1e3);
    }
    {   // 7. own pruning transpose: G rows x N centre columns -> N x G
        cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
        dim3 blk(32, 8), grd((N + 31) / 32, (G + 31) / 32);
        auto run = [&]() {
            if (g_double) transpose_kernel<double2><<<grd, blk>>>((const double2*) x, (double2*) y, G, N, (G - N) / 2, G, G);
            else transpose_kernel<float2><<<grd, blk>>>((const float2*) x, (float2*) y, G, N, (G - N) / 2, G, G);
        };
        run(); CK(cudaDeviceSynchronize());
        cudaEventRecord(a);
        for (int i = 0; i < reps; ++i) run();
        cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b); ms /= reps;
        printf("prune-transpose GxN    : %.3f ms/plane  (%.0f GB/s read+write of N*G)\n", ms, 2 * plane_gb * N / G / ms * 1e3);
    }
    return 0;
}
