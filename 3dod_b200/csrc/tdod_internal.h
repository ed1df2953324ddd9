// Host-side helpers shared by the translation units of libtdod.
#pragma once
#include <cuda_runtime.h>
#include "../../include/tdod.h"

// records `msg` as the calling thread's last error and returns `code`
int tdod_fail(int code, const char* msg);

#define TDOD_CHECK_LAUNCH()                                \
    do {                                                   \
        cudaError_t e__ = cudaGetLastError();              \
        if (e__ != cudaSuccess) return (int)e__;           \
    } while (0)
