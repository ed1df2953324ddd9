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

// mode B of the fused entry (modeb.cu), launched by tdod_propose_score_select on the same stream
int tdod_modeb_launch(const tdod_pss_args& a, void* stream);

// Opt-in to more than 48 KB of dynamic shared memory for kernel `fn`, once per device and only ever upwards:
// cudaFuncSetAttribute is a driver call (not something to pay on every launch), and setting a smaller value later would
// lower the limit for a concurrent caller.  One cache object per kernel.
struct tdod_smem_cache { size_t seen[64]; };
inline cudaError_t tdod_smem_opt_in(const void* fn, size_t bytes, tdod_smem_cache* c) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    dev &= 63;
    if (bytes <= c->seen[dev]) return cudaSuccess;
    e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e == cudaSuccess) c->seen[dev] = bytes;
    return e;
}
