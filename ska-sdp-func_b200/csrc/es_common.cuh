/*
 * Definitions shared by the sm_100a kernels of the ES gridder (es_kernels.cu,
 * es_scatter.cu): sorted visibility records, the ES kernel function, and the
 * search for the records of one (tile, w-plane range).
 */
#ifndef ES_COMMON_CUH_
#define ES_COMMON_CUH_

#include <stdint.h>

#include <cuda_runtime.h>

#include "es_device.h"

namespace es {

/* ------------------------------------------------------------- records */

struct __align__(16) EsRecF
{
    short lu0, lv0;   /* first footprint slot, in tile-local cell indices */
    int p0;           /* first (unclipped) w-plane: ceil(pos_w - W/2) */
    float du, dv, dw; /* (first slot - position) in u, v, w: in [-W/2, -W/2 + 1) */
    float vre, vim;   /* scatter: weighted visibility; gather: vre = flip */
    unsigned int idx; /* row * num_chan + chan */
};

struct __align__(16) EsRecD
{
    short lu0, lv0;
    int p0;
    double du, dv, dw;
    double vre, vim;
    unsigned int idx;
    unsigned int pad0;
    double pad1;
};

static_assert(sizeof(EsRecF) == 32, "fp32 record must be 32 bytes");
static_assert(sizeof(EsRecD) == 64, "fp64 record must be 64 bytes");

template<typename T> struct Prec;
template<> struct Prec<float>
{
    typedef float2 T2;
    typedef EsRecF Rec;
};
template<> struct Prec<double>
{
    typedef double2 T2;
    typedef EsRecD Rec;
};

/* ------------------------------------------------------------ ES kernel */

/* psi(x) = x^2 > 1 ? 0 : exp(beta (sqrt(1 - x^2) - 1)), kernels.cu:97-102 */
__device__ __forceinline__ float es_psi(float beta, float x)
{
    const float t = 1.0f - x * x;
    return (t < 0.0f) ? 0.0f : __expf(beta * (sqrtf(t) - 1.0f));
}

__device__ __forceinline__ double es_psi(double beta, double x)
{
    const double t = 1.0 - x * x;
    return (t < 0.0) ? 0.0 : exp(beta * (sqrt(t) - 1.0));
}

/* Range of sorted entries of `tile_id` that can touch a plane in
 * [plane_first, plane_last]: first owned plane in [plane_first - W,
 * plane_last]. Every thread computes it (broadcast loads). */
__device__ __forceinline__ void es_entry_range(const EsGeom& g,
        const uint64_t* __restrict__ keys,
        const uint32_t* __restrict__ tile_start, int tile_id, int plane_first,
        int plane_last, uint32_t& a, uint32_t& b)
{
    uint32_t lo = tile_start[tile_id];
    const uint32_t hi = tile_start[tile_id + 1];
    a = b = lo;
    if (lo == hi) return;
    const uint64_t pk = (uint64_t) (g.plane_end - g.plane_begin);
    const uint64_t base = (uint64_t) tile_id * pk;
    const int kp_hi = plane_last - g.plane_begin;
    const int kp_lo = max(plane_first - g.plane_begin - g.support, 0);
    const uint64_t key_lo = base + (uint64_t) kp_lo;
    const uint64_t key_hi = base + (uint64_t) kp_hi; /* inclusive */
    uint32_t l = lo, h = hi;
    while (l < h) /* first key >= key_lo */
    {
        const uint32_t m = l + ((h - l) >> 1);
        if (keys[m] < key_lo) l = m + 1; else h = m;
    }
    a = l;
    h = hi;
    while (l < h) /* first key > key_hi */
    {
        const uint32_t m = l + ((h - l) >> 1);
        if (keys[m] <= key_hi) l = m + 1; else h = m;
    }
    b = l;
}

/* ------------------------------------------------------- work units */

/* One unit of the load-balanced work lists (es_scatter.cu: es_work_plan_kernel). */
struct __align__(16) EsUnit
{
    uint32_t a, b;        /* record range of this slice */
    int32_t tile;         /* tile index tu * ntile + tv */
    int32_t plane_q;      /* plane within the batch */
    int32_t num_slices;   /* K: 1 = unsplit */
    int32_t slice;        /* 0 .. K-1 */
    int32_t split_slot;   /* index of the completion counter (K > 1) */
    uint32_t partial_base; /* first partial tile in scratch (K > 1) */
};

static_assert(sizeof(EsUnit) == 32, "work unit must be 32 bytes");

enum { CTR_UNITS = 0, CTR_SPLIT = 1, CTR_PARTIALS = 2, CTR_NEXT = 3 };

} // namespace es

#endif
