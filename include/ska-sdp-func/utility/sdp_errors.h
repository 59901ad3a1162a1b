/*
 * Status codes of the drop-in libska_sdp_func.so (B200-native build).
 *
 * ABI contract: identical enumerator values to the reference's
 *   src/ska-sdp-func/utility/sdp_errors.h:12-34
 * (ctypes sees a 4-byte int; the Python side maps 0..6 to messages in
 *  src/ska_sdp_func/utility/error_checking.py:11-19).
 */
#ifndef SDP_B200_ERRORS_H_
#define SDP_B200_ERRORS_H_

enum sdp_Error
{
    SDP_SUCCESS = 0,               /* no error */
    SDP_ERR_RUNTIME = 1,           /* generic runtime error (CUDA / cuFFT failure) */
    SDP_ERR_INVALID_ARGUMENT = 2,  /* bad shape, stride, flag or scalar */
    SDP_ERR_DATA_TYPE = 3,         /* unsupported or inconsistent element types */
    SDP_ERR_MEM_ALLOC_FAILURE = 4, /* host or device allocation failed */
    SDP_ERR_MEM_COPY_FAILURE = 5,  /* host<->device copy failed */
    SDP_ERR_MEM_LOCATION = 6       /* arrays are not where they must be */
};

typedef enum sdp_Error sdp_Error;

#endif
