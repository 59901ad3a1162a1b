/*
 * sdp_Mem for the drop-in library: opaque ref-counted array descriptor.
 * Host C over the CUDA runtime API. Observable behaviour follows the
 * reference's src/ska-sdp-func/utility/sdp_mem.cpp (line references inline).
 */
#include <stdlib.h>
#include <string.h>

#include <cuda_runtime_api.h>

#include "ska-sdp-func/utility/sdp_mem.h"

struct sdp_Mem
{
    sdp_MemType type;
    sdp_MemLocation location;
    int32_t c_contiguous;
    int32_t owns_data;
    int32_t read_only;
    int32_t num_dims;
    int32_t ref_count;
    int64_t num_elements;
    int64_t* shape;        /* elements per dimension */
    int64_t* stride_bytes; /* numpy-style byte strides */
    void* data;
};

int64_t sdp_mem_type_size(sdp_MemType type)
{
    /* sdp_mem.cpp:1157-1176 */
    switch (type)
    {
    case SDP_MEM_CHAR:           return 1;
    case SDP_MEM_INT:            return (int64_t) sizeof(int);
    case SDP_MEM_FLOAT:          return 4;
    case SDP_MEM_DOUBLE:         return 8;
    case SDP_MEM_COMPLEX_FLOAT:  return 8;
    case SDP_MEM_COMPLEX_DOUBLE: return 16;
    default:                     return 0;
    }
}

sdp_Mem* sdp_mem_create_wrapper(void* data, sdp_MemType type,
        sdp_MemLocation location, int32_t num_dims, const int64_t* shape,
        const int64_t* stride_bytes, sdp_Error* status)
{
    /* sdp_mem.cpp:180-231: a handle is always returned, even on error. */
    sdp_Mem* mem = (sdp_Mem*) calloc(1, sizeof(sdp_Mem));
    if (!mem)
    {
        *status = SDP_ERR_MEM_ALLOC_FAILURE;
        return 0;
    }
    mem->data = data;
    mem->type = type;
    mem->location = location;
    mem->num_dims = num_dims;
    mem->ref_count = 1;
    if (type == SDP_MEM_VOID) return mem;
    const int64_t item = sdp_mem_type_size(type);
    if (item <= 0)
    {
        *status = SDP_ERR_DATA_TYPE;
        SDP_LOG_CRITICAL("Unsupported data type");
        return mem;
    }
    mem->num_elements = 1; /* a zero-dimensional array is a scalar */
    if (num_dims <= 0) return mem;
    mem->shape = (int64_t*) calloc((size_t) num_dims, sizeof(int64_t));
    mem->stride_bytes = (int64_t*) calloc((size_t) num_dims, sizeof(int64_t));
    if (!mem->shape || !mem->stride_bytes)
    {
        *status = SDP_ERR_MEM_ALLOC_FAILURE;
        return mem;
    }
    int64_t dense = item; /* byte stride a C-ordered array would have */
    mem->c_contiguous = 1;
    for (int32_t d = num_dims - 1; d >= 0; --d)
    {
        mem->shape[d] = shape[d];
        mem->stride_bytes[d] = stride_bytes ? stride_bytes[d] : dense;
        if (mem->stride_bytes[d] != dense) mem->c_contiguous = 0;
        dense *= shape[d];
        mem->num_elements *= shape[d];
    }
    return mem;
}

static void allocate_storage(sdp_Mem* mem, sdp_Error* status)
{
    /* sdp_mem.cpp:40-88 */
    if (!mem) return;
    mem->owns_data = 1;
    const size_t bytes =
            (size_t) (mem->num_elements * sdp_mem_type_size(mem->type));
    if (*status || bytes == 0) return;
    if (mem->location == SDP_MEM_CPU)
    {
        mem->data = calloc(bytes, 1);
        if (!mem->data)
        {
            *status = SDP_ERR_MEM_ALLOC_FAILURE;
            SDP_LOG_CRITICAL("Host memory allocation failure "
                    "(requested %zu bytes)", bytes);
        }
    }
    else if (mem->location == SDP_MEM_GPU)
    {
        const cudaError_t err = cudaMalloc(&mem->data, bytes);
        if (err != cudaSuccess || !mem->data)
        {
            *status = SDP_ERR_MEM_ALLOC_FAILURE;
            SDP_LOG_CRITICAL("GPU memory allocation failure: %s "
                    "(requested %zu bytes)", cudaGetErrorString(err), bytes);
            mem->data = 0;
            (void) cudaGetLastError();
        }
    }
    else
    {
        *status = SDP_ERR_MEM_LOCATION;
        SDP_LOG_CRITICAL("Unsupported memory location");
    }
}

sdp_Mem* sdp_mem_create(sdp_MemType type, sdp_MemLocation location,
        int32_t num_dims, const int64_t* shape, sdp_Error* status)
{
    sdp_Mem* mem = sdp_mem_create_wrapper(0, type, location, num_dims, shape,
            0, status);
    allocate_storage(mem, status);
    return mem;
}

sdp_Mem* sdp_mem_create_alias(const sdp_Mem* src)
{
    sdp_Error status = SDP_SUCCESS;
    if (!src) return 0;
    return sdp_mem_create_wrapper(src->data, src->type, src->location,
            src->num_dims, src->shape, src->stride_bytes, &status);
}

sdp_Mem* sdp_mem_create_copy(const sdp_Mem* src, sdp_MemLocation location,
        sdp_Error* status)
{
    /* sdp_mem.cpp:310-322 */
    if (!src) return 0;
    sdp_Mem* mem = sdp_mem_create_wrapper(0, src->type, location,
            src->num_dims, src->shape, src->stride_bytes, status);
    allocate_storage(mem, status);
    sdp_mem_copy_contents(mem, src, 0, 0, src->num_elements, status);
    return mem;
}

void sdp_mem_clear_portion(sdp_Mem* mem, int64_t start_index,
        int64_t num_elements, sdp_Error* status)
{
    /* sdp_mem.cpp:332-362 */
    if (*status || !mem || num_elements == 0 || !mem->data) return;
    const size_t item = (size_t) sdp_mem_type_size(mem->type);
    char* first = (char*) mem->data + (size_t) start_index * item;
    const size_t bytes = (size_t) num_elements * item;
    if (mem->location == SDP_MEM_CPU)
    {
        memset(first, 0, bytes);
    }
    else if (mem->location == SDP_MEM_GPU)
    {
        if (cudaMemset(first, 0, bytes) != cudaSuccess)
        {
            *status = SDP_ERR_RUNTIME;
            SDP_LOG_ERROR("cudaMemset failed");
            (void) cudaGetLastError();
        }
    }
    else
    {
        *status = SDP_ERR_MEM_LOCATION;
        SDP_LOG_CRITICAL("Unsupported memory location");
    }
}

void sdp_mem_clear_contents(sdp_Mem* mem, sdp_Error* status)
{
    if (*status || !mem || mem->num_elements == 0) return;
    sdp_mem_clear_portion(mem, 0, mem->num_elements, status);
}

void sdp_mem_copy_contents(sdp_Mem* dst, const sdp_Mem* src,
        int64_t offset_dst, int64_t offset_src, int64_t num_elements,
        sdp_Error* status)
{
    /* sdp_mem.cpp:444-499 */
    if (*status || !dst || !src || !dst->data || !src->data) return;
    if (src->num_elements == 0 || num_elements == 0) return;
    const int64_t item = sdp_mem_type_size(src->type);
    const void* from = (const char*) src->data + item * offset_src;
    void* to = (char*) dst->data + item * offset_dst;
    const size_t bytes = (size_t) (item * num_elements);
    const int src_gpu = src->location == SDP_MEM_GPU;
    const int dst_gpu = dst->location == SDP_MEM_GPU;
    if ((src->location != SDP_MEM_CPU && !src_gpu) ||
            (dst->location != SDP_MEM_CPU && !dst_gpu))
    {
        *status = SDP_ERR_MEM_LOCATION;
        SDP_LOG_CRITICAL("Unsupported memory location");
        return;
    }
    if (!src_gpu && !dst_gpu)
    {
        memcpy(to, from, bytes);
        return;
    }
    const enum cudaMemcpyKind kind = src_gpu ?
            (dst_gpu ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost) :
            cudaMemcpyHostToDevice;
    const cudaError_t err = cudaMemcpy(to, from, bytes, kind);
    if (err != cudaSuccess)
    {
        *status = SDP_ERR_MEM_COPY_FAILURE;
        SDP_LOG_CRITICAL("cudaMemcpy error: %s", cudaGetErrorString(err));
        (void) cudaGetLastError();
    }
}

void sdp_mem_random_fill(sdp_Mem* mem, sdp_Error* status)
{
    /* sdp_mem.cpp:708-738: rand()/RAND_MAX over every real component. */
    if (*status || !mem) return;
    if (mem->location != SDP_MEM_CPU)
    {
        *status = SDP_ERR_MEM_LOCATION;
        SDP_LOG_ERROR("Unsupported memory location");
        return;
    }
    int64_t count = mem->num_elements;
    if (sdp_mem_is_complex(mem)) count *= 2;
    const int precision = mem->type & 0x0F;
    if (precision == SDP_MEM_FLOAT)
    {
        float* v = (float*) mem->data;
        for (int64_t i = 0; i < count; ++i)
            v[i] = (float) rand() / (float) RAND_MAX;
    }
    else if (precision == SDP_MEM_DOUBLE)
    {
        double* v = (double*) mem->data;
        for (int64_t i = 0; i < count; ++i)
            v[i] = (double) rand() / (double) RAND_MAX;
    }
}

void* sdp_mem_data(sdp_Mem* mem)
{
    return mem ? mem->data : 0;
}

const void* sdp_mem_data_const(const sdp_Mem* mem)
{
    return mem ? mem->data : 0;
}

void* sdp_mem_gpu_buffer(sdp_Mem* mem, sdp_Error* status)
{
    /* sdp_mem.cpp:581-591 */
    if (*status || !mem) return 0;
    if (mem->location != SDP_MEM_GPU && mem->type != SDP_MEM_VOID)
    {
        *status = SDP_ERR_MEM_LOCATION;
        SDP_LOG_CRITICAL("Requested buffer is not in GPU memory");
        return 0;
    }
    return &mem->data;
}

const void* sdp_mem_gpu_buffer_const(const sdp_Mem* mem, sdp_Error* status)
{
    return sdp_mem_gpu_buffer((sdp_Mem*) mem, status);
}

void sdp_mem_free(sdp_Mem* mem)
{
    /* sdp_mem.cpp:607-627: drops one reference. */
    if (!mem) return;
    if (--mem->ref_count > 0) return;
    if (mem->owns_data && mem->data)
    {
        if (mem->location == SDP_MEM_CPU) free(mem->data);
        else if (mem->location == SDP_MEM_GPU) (void) cudaFree(mem->data);
    }
    free(mem->shape);
    free(mem->stride_bytes);
    free(mem);
}

void sdp_mem_ref_dec(sdp_Mem* mem)
{
    sdp_mem_free(mem);
}

sdp_Mem* sdp_mem_ref_inc(sdp_Mem* mem)
{
    if (mem) mem->ref_count++;
    return mem;
}

/* Queries: a handle without data answers like an empty read-only array
 * (sdp_mem.cpp:630-705). */
int32_t sdp_mem_is_c_contiguous(const sdp_Mem* mem)
{
    return (mem && mem->data) ? mem->c_contiguous : 0;
}

int32_t sdp_mem_is_floating_point(const sdp_Mem* mem)
{
    if (!mem || !mem->data) return 0;
    return mem->type == SDP_MEM_FLOAT || mem->type == SDP_MEM_DOUBLE;
}

int32_t sdp_mem_is_complex(const sdp_Mem* mem)
{
    if (!mem || !mem->data) return 0;
    return (mem->type & SDP_MEM_COMPLEX) == SDP_MEM_COMPLEX;
}

int32_t sdp_mem_is_double(const sdp_Mem* mem)
{
    if (!mem || !mem->data) return 0;
    return (mem->type & SDP_MEM_DOUBLE) == SDP_MEM_DOUBLE;
}

int32_t sdp_mem_is_matching(const sdp_Mem* mem1, const sdp_Mem* mem2,
        int32_t check_location)
{
    if (!mem1 || !mem2) return 0;
    if (mem1->type != mem2->type) return 0;
    if (check_location && mem1->location != mem2->location) return 0;
    if (mem1->num_dims != mem2->num_dims) return 0;
    for (int32_t d = 0; d < mem1->num_dims; ++d)
    {
        if (mem1->shape[d] != mem2->shape[d]) return 0;
        if (mem1->stride_bytes[d] != mem2->stride_bytes[d]) return 0;
    }
    return 1;
}

int32_t sdp_mem_is_read_only(const sdp_Mem* mem)
{
    return (mem && mem->data) ? mem->read_only : 1;
}

sdp_MemLocation sdp_mem_location(const sdp_Mem* mem)
{
    return mem ? mem->location : SDP_MEM_CPU;
}

int32_t sdp_mem_num_dims(const sdp_Mem* mem)
{
    return mem ? mem->num_dims : 0;
}

int64_t sdp_mem_num_elements(const sdp_Mem* mem)
{
    return (mem && mem->data) ? mem->num_elements : 0;
}

void sdp_mem_set_read_only(sdp_Mem* mem, int32_t value)
{
    if (mem) mem->read_only = value;
}

int64_t sdp_mem_shape_dim(const sdp_Mem* mem, int32_t dim)
{
    return (mem && dim >= 0 && dim < mem->num_dims) ? mem->shape[dim] : 0;
}

int64_t sdp_mem_stride_bytes_dim(const sdp_Mem* mem, int32_t dim)
{
    return (mem && dim >= 0 && dim < mem->num_dims) ?
            mem->stride_bytes[dim] : 0;
}

int64_t sdp_mem_stride_elements_dim(const sdp_Mem* mem, int32_t dim)
{
    if (!mem) return 0;
    const int64_t item = sdp_mem_type_size(mem->type);
    return item > 0 ? sdp_mem_stride_bytes_dim(mem, dim) / item : 0;
}

sdp_MemType sdp_mem_type(const sdp_Mem* mem)
{
    return mem ? mem->type : SDP_MEM_VOID;
}

const char* sdp_mem_location_name(sdp_MemLocation location)
{
    return location == SDP_MEM_CPU ? "CPU" :
           location == SDP_MEM_GPU ? "GPU" : "???";
}

const char* sdp_mem_type_name(sdp_MemType type)
{
    switch (type)
    {
    case SDP_MEM_VOID:           return "void";
    case SDP_MEM_CHAR:           return "char";
    case SDP_MEM_INT:            return "int";
    case SDP_MEM_FLOAT:          return "float";
    case SDP_MEM_DOUBLE:         return "double";
    case SDP_MEM_COMPLEX_FLOAT:  return "complex float";
    case SDP_MEM_COMPLEX_DOUBLE: return "complex double";
    default:                     return "???";
    }
}
