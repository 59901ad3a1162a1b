/*
 * SKA-format log lines ("1|<UTC>|<LEVEL>||<func>|<file>#<line>|| <msg>").
 *
 * Replaces src/ska-sdp-func/utility/sdp_logging.h:27-95 / sdp_logging.c:13-113:
 * same symbol, same level enum, same environment switch
 * SKA_SDP_FUNC_LOG_LEVEL (debug|info|warn|err|crit, default info),
 * DEBUG/INFO to stdout, WARNING and above to stderr.
 */
#ifndef SDP_B200_LOGGING_H_
#define SDP_B200_LOGGING_H_

#include <stdio.h>

#ifdef __cplusplus
extern "C" {
#endif

enum sdp_LogLevel
{
    SDP_LOG_LEVEL_UNDEF,
    SDP_LOG_LEVEL_DEBUG,
    SDP_LOG_LEVEL_INFO,
    SDP_LOG_LEVEL_WARNING,
    SDP_LOG_LEVEL_ERROR,
    SDP_LOG_LEVEL_CRITICAL
};
typedef enum sdp_LogLevel sdp_LogLevel;

void sdp_log_message(sdp_LogLevel level, FILE* stream, const char* func,
        const char* file, int line, const char* message, ...);

#ifndef SOURCE_PATH_SIZE
#define SOURCE_PATH_SIZE 0
#endif
#define SDP_LOG_FILE_ ((__FILE__) + SOURCE_PATH_SIZE)

#define SDP_LOG_CRITICAL(...) sdp_log_message(SDP_LOG_LEVEL_CRITICAL, stderr, \
        __func__, SDP_LOG_FILE_, __LINE__, __VA_ARGS__)
#define SDP_LOG_ERROR(...) sdp_log_message(SDP_LOG_LEVEL_ERROR, stderr, \
        __func__, SDP_LOG_FILE_, __LINE__, __VA_ARGS__)
#define SDP_LOG_WARNING(...) sdp_log_message(SDP_LOG_LEVEL_WARNING, stderr, \
        __func__, SDP_LOG_FILE_, __LINE__, __VA_ARGS__)
#define SDP_LOG_INFO(...) sdp_log_message(SDP_LOG_LEVEL_INFO, stdout, \
        __func__, SDP_LOG_FILE_, __LINE__, __VA_ARGS__)
#define SDP_LOG_DEBUG(...) sdp_log_message(SDP_LOG_LEVEL_DEBUG, stdout, \
        __func__, SDP_LOG_FILE_, __LINE__, __VA_ARGS__)

#ifdef __cplusplus
}
#endif

#endif
