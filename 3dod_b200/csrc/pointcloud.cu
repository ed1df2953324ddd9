// score_point_cloud (ProposalNetwork/scoring/scorefunction.py:9-43): per proposal, the number of points strictly
// inside three intervals.  The reference reduces `verts[:, j].min(1)` over the xyz of VERTEX j (j = 0,1,2), not over
// the 8 vertices, so the intervals are x in (min v0, max v0), y in (min v1, max v1), z in (min v2, max v2);
// reproduced as is.  Points are staged through shared memory tile by tile; every thread owns one proposal.
#include "tdod_device.cuh"
#include "tdod_internal.h"

using namespace tdod;

namespace {

constexpr int kT = 128;       // proposals per CTA
constexpr int kTile = 1024;   // points per shared-memory tile

template <typename T>
__global__ void __launch_bounds__(kT) pc_kernel(const T* __restrict__ pts, int64_t P, const float* __restrict__ cubes,
                                                int64_t n, int64_t* __restrict__ out) {
    __shared__ T sx[kTile], sy[kTile], sz[kTile];
    const int64_t i = (int64_t)blockIdx.x * kT + threadIdx.x;
    T lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
    if (i < n) {
        float q[15];
#pragma unroll
        for (int k = 0; k < 15; ++k) q[k] = cubes[i * 15 + k];
        float X[8], Y[8], Z[8];
        cube_corners(q, X, Y, Z);
#pragma unroll
        for (int j = 0; j < 3; ++j) {   // comparisons promote the fp32 bounds to the dtype of the points
            lo[j] = (T)min_nan(min_nan(X[j], Y[j]), Z[j]);
            hi[j] = (T)max_nan(max_nan(X[j], Y[j]), Z[j]);
        }
    }
    // this grid.y slice of the point range
    const int64_t per = (P + gridDim.y - 1) / gridDim.y;
    const int64_t p0 = (int64_t)blockIdx.y * per, p1 = min(P, p0 + per);
    long long cnt = 0;
    for (int64_t base = p0; base < p1; base += kTile) {
        const int m = (int)min((int64_t)kTile, p1 - base);
        __syncthreads();
        for (int t = threadIdx.x; t < m; t += kT) {
            const T* p = pts + (base + t) * 3;
            sx[t] = p[0]; sy[t] = p[1]; sz[t] = p[2];
        }
        __syncthreads();
        int c = 0;
#pragma unroll 4
        for (int t = 0; t < m; ++t) {
            const T x = sx[t], y = sy[t], z = sz[t];
            c += (x > lo[0]) & (x < hi[0]) & (y > lo[1]) & (y < hi[1]) & (z > lo[2]) & (z < hi[2]);
        }
        cnt += c;
    }
    if (i < n && cnt) atomicAdd(reinterpret_cast<unsigned long long*>(out + i), (unsigned long long)cnt);
}

}  // namespace

extern "C" int tdod_score_point_cloud(const void* points, int points_f64, int64_t P, const float* cubes, int64_t n,
                                      int64_t* out, void* stream) {
    if (n == 0) return TDOD_OK;
    if (!cubes || !out || n < 0 || P < 0 || (P > 0 && !points))
        return tdod_fail(TDOD_EINVAL, "tdod_score_point_cloud: bad argument");
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t e = cudaMemsetAsync(out, 0, (size_t)n * sizeof(int64_t), st);
    if (e != cudaSuccess) return (int)e;
    if (P == 0) return TDOD_OK;
    const unsigned gx = (unsigned)((n + kT - 1) / kT);
    unsigned gy = 1;   // split the points when there are too few proposal tiles to fill 148 SMs
    while (gx * gy < 296 && (int64_t)gy * kTile * 2 <= P) gy *= 2;
    if (points_f64)
        pc_kernel<double><<<dim3(gx, gy), kT, 0, st>>>((const double*)points, P, cubes, n, out);
    else
        pc_kernel<float><<<dim3(gx, gy), kT, 0, st>>>((const float*)points, P, cubes, n, out);
    TDOD_CHECK_LAUNCH();
    return TDOD_OK;
}
