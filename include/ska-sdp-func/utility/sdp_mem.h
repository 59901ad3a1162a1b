/*
 * sdp_Mem: opaque, reference-counted descriptor of a (possibly wrapped)
 * multi-dimensional array on the host or on the current CUDA device.
 *
 * This header re-exports, with identical names, argument order, enum values
 * and observable behaviour, the part of the reference's
 *   src/ska-sdp-func/utility/sdp_mem.h:59-577  (implementation sdp_mem.cpp:20-1000)
 * that the sdp_GridderUvwEsFft path, its Python bindings
 * (src/ska_sdp_func/utility/mem.py:45-146) and its tests
 * (tests/grid_data/test_gridder_uvw_es_fft.cpp) use. The struct is opaque in
 * the reference too (sdp_mem.h:59), so only function behaviour is ABI.
 *
 * Behaviour kept on purpose (callers rely on it):
 *  - strides are in BYTES (numpy compatible); a NULL stride means C order;
 *  - sdp_mem_create zero-fills host memory but not device memory;
 *  - queries on a handle whose data pointer is NULL answer "not contiguous,
 *    not complex, read-only, 0 elements" (sdp_mem.cpp:630-705);
 *  - sdp_mem_free only drops one reference; wrapped data is never freed;
 *  - sdp_mem_gpu_buffer returns the ADDRESS of the data pointer field
 *    (what cudaLaunchKernel wants) and flags CPU memory with
 *    SDP_ERR_MEM_LOCATION (sdp_mem.cpp:581-604).
 */
#ifndef SDP_B200_MEM_H_
#define SDP_B200_MEM_H_

#include <stdint.h>

#include "ska-sdp-func/utility/sdp_errors.h"
#include "ska-sdp-func/utility/sdp_logging.h"

#ifdef __cplusplus
extern "C" {
#endif

struct sdp_Mem;
typedef struct sdp_Mem sdp_Mem;

/* sdp_mem.h:73-99 */
enum sdp_MemType
{
    SDP_MEM_VOID = 0,
    SDP_MEM_CHAR = 1,
    SDP_MEM_INT = 2,
    SDP_MEM_FLOAT = 4,
    SDP_MEM_DOUBLE = 8,
    SDP_MEM_COMPLEX = 32,
    SDP_MEM_COMPLEX_FLOAT = SDP_MEM_FLOAT | SDP_MEM_COMPLEX,
    SDP_MEM_COMPLEX_DOUBLE = SDP_MEM_DOUBLE | SDP_MEM_COMPLEX
};
typedef enum sdp_MemType sdp_MemType;

/* sdp_mem.h:106-113 */
enum sdp_MemLocation
{
    SDP_MEM_CPU = 0,
    SDP_MEM_GPU = 1
};
typedef enum sdp_MemLocation sdp_MemLocation;

/* Allocation, wrapping, aliasing (sdp_mem.h:140-227). */
sdp_Mem* sdp_mem_create(sdp_MemType type, sdp_MemLocation location,
        int32_t num_dims, const int64_t* shape, sdp_Error* status);
sdp_Mem* sdp_mem_create_wrapper(void* data, sdp_MemType type,
        sdp_MemLocation location, int32_t num_dims, const int64_t* shape,
        const int64_t* stride_bytes, sdp_Error* status);
sdp_Mem* sdp_mem_create_alias(const sdp_Mem* src);
sdp_Mem* sdp_mem_create_copy(const sdp_Mem* src, sdp_MemLocation location,
        sdp_Error* status);

/* Contents (sdp_mem.h:229-305, 454, 486-505). */
void sdp_mem_clear_contents(sdp_Mem* mem, sdp_Error* status);
void sdp_mem_clear_portion(sdp_Mem* mem, int64_t start_index,
        int64_t num_elements, sdp_Error* status);
void sdp_mem_copy_contents(sdp_Mem* dst, const sdp_Mem* src,
        int64_t offset_dst, int64_t offset_src, int64_t num_elements,
        sdp_Error* status);
void sdp_mem_random_fill(sdp_Mem* mem, sdp_Error* status);

/* Raw access (sdp_mem.h:309-345). */
void* sdp_mem_data(sdp_Mem* mem);
const void* sdp_mem_data_const(const sdp_Mem* mem);
void* sdp_mem_gpu_buffer(sdp_Mem* mem, sdp_Error* status);
const void* sdp_mem_gpu_buffer_const(const sdp_Mem* mem, sdp_Error* status);

/* Lifetime (sdp_mem.h:357, 466-477). */
void sdp_mem_free(sdp_Mem* mem);
void sdp_mem_ref_dec(sdp_Mem* mem);
sdp_Mem* sdp_mem_ref_inc(sdp_Mem* mem);

/* Queries (sdp_mem.h:365-575). */
int32_t sdp_mem_is_c_contiguous(const sdp_Mem* mem);
int32_t sdp_mem_is_floating_point(const sdp_Mem* mem);
int32_t sdp_mem_is_complex(const sdp_Mem* mem);
int32_t sdp_mem_is_double(const sdp_Mem* mem);
int32_t sdp_mem_is_matching(const sdp_Mem* mem1, const sdp_Mem* mem2,
        int32_t check_location);
int32_t sdp_mem_is_read_only(const sdp_Mem* mem);
sdp_MemLocation sdp_mem_location(const sdp_Mem* mem);
int32_t sdp_mem_num_dims(const sdp_Mem* mem);
int64_t sdp_mem_num_elements(const sdp_Mem* mem);
void sdp_mem_set_read_only(sdp_Mem* mem, int32_t value);
int64_t sdp_mem_shape_dim(const sdp_Mem* mem, int32_t dim);
int64_t sdp_mem_stride_bytes_dim(const sdp_Mem* mem, int32_t dim);
int64_t sdp_mem_stride_elements_dim(const sdp_Mem* mem, int32_t dim);
sdp_MemType sdp_mem_type(const sdp_Mem* mem);
int64_t sdp_mem_type_size(sdp_MemType type);
const char* sdp_mem_location_name(sdp_MemLocation location);
const char* sdp_mem_type_name(sdp_MemType type);

#ifdef __cplusplus
}
#endif

#endif
