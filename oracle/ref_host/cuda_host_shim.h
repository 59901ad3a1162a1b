/*
 * TEST INFRASTRUCTURE ONLY (oracle). Not part of the product library.
 *
 * Minimal CUDA-on-host shim that lets the reference's kernel source
 *   /root/reference/src/ska-sdp-func/grid_data/sdp_gridder_uvw_es_fft_kernels.cu
 * be compiled, unmodified and in place, as plain host C++ by g++.
 * Each "kernel" becomes an ordinary function that processes ONE thread;
 * the driver (ref_driver.cpp) sets threadIdx/blockIdx before each call.
 *
 * The reference sources are #included from where they lie at build time and
 * are never copied into this repository.
 */
#ifndef ORACLE_CUDA_HOST_SHIM_H_
#define ORACLE_CUDA_HOST_SHIM_H_

#include <cmath>
#include <cstdint>
#include <cstddef>
#include <cstdlib>

/* Skip the reference's device-wrapper header (registry + CUDA launch glue). */
#define SKA_SDP_PROC_FUNC_DEVICE_WRAPPER_H_
#define SDP_CUDA_KERNEL(...)
struct sdp_CudaStream;
typedef struct sdp_CudaStream sdp_CudaStream;

#define __global__
#define __device__
#define __host__
#define __restrict__
#define __forceinline__ inline
#define __CUDA_ARCH__ 1000

struct float2 { float x, y; };
struct double2 { double x, y; };
struct float3 { float x, y, z; };
struct double3 { double x, y, z; };
struct oracle_dim3 { unsigned int x, y, z; };

extern thread_local oracle_dim3 threadIdx, blockIdx, blockDim, gridDim;

/* The kernels call unqualified math functions on FP = float or double; pick
 * the overload matching the argument type like nvcc does. */
using std::ceil;
using std::cos;
using std::exp;
using std::fabs;
using std::floor;
using std::sqrt;

static inline void sincos(float x, float* s, float* c) { ::sincosf(x, s, c); }

/* CUDA's integer min/max overload set: mixed int/unsigned compares unsigned
 * (reference kernels.cu:187-189 relies on this). */
static inline int min(int a, int b) { return a < b ? a : b; }
static inline int max(int a, int b) { return a > b ? a : b; }
static inline unsigned int min(unsigned int a, unsigned int b) { return a < b ? a : b; }
static inline unsigned int max(unsigned int a, unsigned int b) { return a > b ? a : b; }
static inline unsigned int min(int a, unsigned int b) { return min((unsigned int) a, b); }
static inline unsigned int min(unsigned int a, int b) { return min(a, (unsigned int) b); }
static inline unsigned int max(int a, unsigned int b) { return max((unsigned int) a, b); }
static inline unsigned int max(unsigned int a, int b) { return max(a, (unsigned int) b); }

static inline float atomicAdd(float* addr, float v)
{
    float old;
    #pragma omp atomic capture
    { old = *addr; *addr += v; }
    return old;
}

static inline double atomicAdd(double* addr, double v)
{
    double old;
    #pragma omp atomic capture
    { old = *addr; *addr += v; }
    return old;
}

#endif
