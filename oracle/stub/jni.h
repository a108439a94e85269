/* Stub so that /root/reference/src/sdr-bridge-internal.h (which only needs the
 * JavaVM type name) compiles on a desktop toolchain. Test infrastructure only. */
#pragma once
typedef struct _JavaVM JavaVM;
