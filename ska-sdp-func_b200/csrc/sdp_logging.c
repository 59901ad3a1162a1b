/*
 * SKA-format logging for the drop-in library.
 * Behaviour follows the reference's src/ska-sdp-func/utility/sdp_logging.c:13-113
 * (line format, level names, SKA_SDP_FUNC_LOG_LEVEL filter, default INFO).
 */
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>
#include <strings.h>
#include <sys/time.h>
#include <time.h>

#include "ska-sdp-func/utility/sdp_logging.h"

static sdp_LogLevel g_threshold = SDP_LOG_LEVEL_UNDEF;

static sdp_LogLevel threshold_from_env(void)
{
    static const struct { const char* prefix; sdp_LogLevel level; } names[] = {
        {"debug", SDP_LOG_LEVEL_DEBUG}, {"info", SDP_LOG_LEVEL_INFO},
        {"warn", SDP_LOG_LEVEL_WARNING}, {"err", SDP_LOG_LEVEL_ERROR},
        {"crit", SDP_LOG_LEVEL_CRITICAL}
    };
    const char* env = getenv("SKA_SDP_FUNC_LOG_LEVEL");
    if (!env) return SDP_LOG_LEVEL_INFO;
    for (size_t i = 0; i < sizeof(names) / sizeof(names[0]); ++i)
    {
        /* the reference accepts all-lower or all-upper spellings */
        const size_t n = strlen(names[i].prefix);
        if (!strncasecmp(env, names[i].prefix, n)) return names[i].level;
    }
    /* Unrecognised value: the reference leaves the filter undefined, which
     * lets every message through (level < UNDEF is never true). */
    return SDP_LOG_LEVEL_UNDEF;
}

static const char* level_name(sdp_LogLevel level)
{
    switch (level)
    {
    case SDP_LOG_LEVEL_UNDEF:    return "UNDEF";
    case SDP_LOG_LEVEL_DEBUG:    return "DEBUG";
    case SDP_LOG_LEVEL_INFO:     return "INFO";
    case SDP_LOG_LEVEL_WARNING:  return "WARNING";
    case SDP_LOG_LEVEL_ERROR:    return "ERROR";
    case SDP_LOG_LEVEL_CRITICAL: return "CRITICAL";
    default:                     return "UNKNOWN";
    }
}

void sdp_log_message(sdp_LogLevel level, FILE* stream, const char* func,
        const char* file, int line, const char* message, ...)
{
    static int env_read = 0;
    if (!env_read)
    {
        g_threshold = threshold_from_env();
        env_read = 1;
    }
    if (level < g_threshold) return;

    struct timeval tv;
    gettimeofday(&tv, 0);
    struct tm utc;
    const time_t secs = tv.tv_sec;
    gmtime_r(&secs, &utc);
    char stamp[48];
    strftime(stamp, sizeof(stamp), "%Y-%m-%dT%H:%M:%S", &utc);

    va_list args;
    va_start(args, message);
    fprintf(stream, "1|%s.%03dZ|%s||%s|%s#%i|| ", stamp,
            (int) (tv.tv_usec / 1000), level_name(level), func, file, line);
    vfprintf(stream, message, args);
    fputc('\n', stream);
    va_end(args);
}
