/* Stub for <android/log.h>: logging is a no-op in the oracle build. Test infrastructure only. */
#pragma once
enum android_LogPriority { ANDROID_LOG_DEBUG = 3, ANDROID_LOG_INFO, ANDROID_LOG_WARN, ANDROID_LOG_ERROR };
#define __android_log_print(prio, tag, ...) ((void)0)
