// Version / error plumbing of the C ABI.
#include <string.h>
#include "tdod_internal.h"

static thread_local char g_last_error[256] = "";

int tdod_fail(int code, const char* msg) {
    strncpy(g_last_error, msg ? msg : "", sizeof(g_last_error) - 1);
    g_last_error[sizeof(g_last_error) - 1] = 0;
    return code;
}

extern "C" int tdod_version(void) { return 200; }

extern "C" const char* tdod_last_error_string(void) { return g_last_error; }

extern "C" const char* tdod_strerror(int code) {
    if (code == TDOD_OK) return "ok";
    if (code == TDOD_EINVAL) return "invalid argument";
    if (code == TDOD_ERANGE) return "size outside the supported range";
    if (code > 0) return cudaGetErrorString((cudaError_t)code);
    return "unknown error";
}
