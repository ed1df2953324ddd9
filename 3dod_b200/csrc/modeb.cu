// Mode B of the fused entry: the three scores the MABO branch adds to IoU x dims x corners
// (cubercnn/modeling/roi_heads/roi_heads.py:535-538): score_segmentation and score_mod_segmentation
// (ProposalNetwork/scoring/scorefunction.py:88-126, mask_iou / mod_mask_iou utils.py:230-250) and score_point_cloud
// (scorefunction.py:9-43), for ALL objects of a batch in one launch.
//
// One CTA per object.  The object's proposals are read back as the 15 cube parameters K-C wrote (in this mode every
// cube is an OUTPUT of the call anyway -- the caller needs them for IoU3D, roi_heads.py:526); corners, projections,
// hulls and polygons exist only in registers / shared memory:
//   segmentation   the [::4, ::4] sampled mask becomes a row-prefix-sum table in shared memory ONCE per object (the
//                  per-function entry point rebuilds it for every 8 proposals); then one warp per proposal: cube ->
//                  8 projected corners -> strict hull -> closed-form fillPoly intervals per sampled row -> integer
//                  intersection / polygon sums; both scores come from the same pair of sums.
//   point cloud    the frame's non-ground points are first cut down to the union of the object's proposal intervals
//                  (a point outside the union is inside no proposal), compacted chunk by chunk into shared memory, and
//                  every thread counts them against the intervals of its proposals held in registers.
#include "raster_device.cuh"
#include "tdod_internal.h"

using namespace tdod;
using namespace tdod_raster;

namespace {

#ifndef TDOD_MODEB_MINB
#define TDOD_MODEB_MINB 3
#endif
constexpr int kMT = 256;                // threads per CTA
constexpr int kMW = kMT / 32;
constexpr int kPtChunk = 2048;          // points examined per round
constexpr int kPerThread = 4;           // proposals whose intervals a thread holds at a time

struct ModeBShared {
    int mask_total;
    int n_kept;
    float uni_lo[3], uni_hi[3];
    float red[kMW][6];
    Edge edges[kMW][8];
};

template <typename PT>
__device__ __forceinline__ void point_cloud_phase(const tdod_pss_args& a, int obj, int frame, unsigned char* dyn, ModeBShared& sh) {
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int n = a.n_prop;
    const float* cubes = a.all_cubes + (size_t)obj * n * 15;
    const PT* px = reinterpret_cast<const PT*>(a.points) + (size_t)frame * 3 * a.P_cap;
    const PT* py = px + a.P_cap;
    const PT* pz = py + a.P_cap;
    const int P = a.n_points[frame];
    PT* kx = reinterpret_cast<PT*>(dyn);            // compacted points of the current chunk
    PT* ky = kx + kPtChunk;
    PT* kz = ky + kPtChunk;
    // ---- union of the object's proposal intervals (scorefunction.py:15-24: interval j = min / max over the xyz of VERTEX j) ----
    float ulo[3] = {CUDART_INF_F, CUDART_INF_F, CUDART_INF_F}, uhi[3] = {-CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F};
    for (int i = tid; i < n; i += kMT) {
        float q[15];
#pragma unroll
        for (int k = 0; k < 15; ++k) q[k] = cubes[(size_t)i * 15 + k];
        float X[8], Y[8], Z[8];
        cube_corners(q, X, Y, Z);
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            const float lo = min_nan(min_nan(X[j], Y[j]), Z[j]), hi = max_nan(max_nan(X[j], Y[j]), Z[j]);
            // a NaN interval contains no point (every comparison is false): it does not widen the union
            if (lo == lo && hi == hi) { ulo[j] = fminf(ulo[j], lo); uhi[j] = fmaxf(uhi[j], hi); }
        }
    }
#pragma unroll
    for (int j = 0; j < 3; ++j) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            ulo[j] = fminf(ulo[j], __shfl_xor_sync(0xffffffffu, ulo[j], o));
            uhi[j] = fmaxf(uhi[j], __shfl_xor_sync(0xffffffffu, uhi[j], o));
        }
        if (lane == 0) { sh.red[wid][j] = ulo[j]; sh.red[wid][3 + j] = uhi[j]; }
    }
    __syncthreads();
    if (tid < 3) {
        float lo = CUDART_INF_F, hi = -CUDART_INF_F;
        for (int w = 0; w < kMW; ++w) { lo = fminf(lo, sh.red[w][tid]); hi = fmaxf(hi, sh.red[w][3 + tid]); }
        sh.uni_lo[tid] = lo; sh.uni_hi[tid] = hi;
    }
    __syncthreads();
    const PT U0 = (PT)sh.uni_lo[0], U1 = (PT)sh.uni_lo[1], U2 = (PT)sh.uni_lo[2];
    const PT V0 = (PT)sh.uni_hi[0], V1 = (PT)sh.uni_hi[1], V2 = (PT)sh.uni_hi[2];
    // ---- proposals in batches of kMT * kPerThread; for each batch the points stream through shared memory ----
    for (int i0 = 0; i0 < n; i0 += kMT * kPerThread) {
        PT lo[kPerThread][3], hi[kPerThread][3];
        int cnt[kPerThread];
#pragma unroll
        for (int u = 0; u < kPerThread; ++u) {
            const int i = i0 + u * kMT + tid;
            cnt[u] = 0;
#pragma unroll
            for (int j = 0; j < 3; ++j) { lo[u][j] = (PT)1; hi[u][j] = (PT)0; }     // empty interval
            if (i < n) {
                float q[15];
#pragma unroll
                for (int k = 0; k < 15; ++k) q[k] = cubes[(size_t)i * 15 + k];
                float X[8], Y[8], Z[8];
                cube_corners(q, X, Y, Z);
#pragma unroll
                for (int j = 0; j < 3; ++j) {   // comparisons promote the fp32 bounds to the dtype of the points
                    lo[u][j] = (PT)min_nan(min_nan(X[j], Y[j]), Z[j]);
                    hi[u][j] = (PT)max_nan(max_nan(X[j], Y[j]), Z[j]);
                }
            }
        }
        for (int p0 = (int)blockIdx.y * kPtChunk; p0 < P; p0 += (int)gridDim.y * kPtChunk) {   // chunks dealt round-robin to the splits
            __syncthreads();
            if (tid == 0) sh.n_kept = 0;
            __syncthreads();
            for (int t = tid; t < kPtChunk; t += kMT) {
                const int p = p0 + t;
                bool keep = false;
                PT x = 0, y = 0, z = 0;
                if (p < P) {
                    x = px[p]; y = py[p]; z = pz[p];
                    keep = (x > U0) & (x < V0) & (y > U1) & (y < V1) & (z > U2) & (z < V2);
                }
                const unsigned m = __ballot_sync(0xffffffffu, keep);
                int base = 0;
                if (lane == 0 && m) base = atomicAdd(&sh.n_kept, __popc(m));
                base = __shfl_sync(0xffffffffu, base, 0);
                if (keep) {
                    const int k = base + __popc(m & ((1u << lane) - 1u));
                    kx[k] = x; ky[k] = y; kz[k] = z;
                }
            }
            __syncthreads();
            const int m = sh.n_kept;
#pragma unroll 2
            for (int t = 0; t < m; ++t) {
                const PT x = kx[t], y = ky[t], z = kz[t];
#pragma unroll
                for (int u = 0; u < kPerThread; ++u)
                    cnt[u] += (x > lo[u][0]) & (x < hi[u][0]) & (y > lo[u][1]) & (y < hi[u][1]) & (z > lo[u][2]) & (z < hi[u][2]);
            }
        }
#pragma unroll
        for (int u = 0; u < kPerThread; ++u) {
            const int i = i0 + u * kMT + tid;
            if (i < n) {
                if (gridDim.y == 1) a.all_pc[(size_t)obj * n + i] = cnt[u];
                else if (cnt[u]) atomicAdd(&a.all_pc[(size_t)obj * n + i], cnt[u]);     // (zeroed by the launcher)
            }
        }
    }
}

__global__ void __launch_bounds__(kMT, TDOD_MODEB_MINB) modeb_kernel(const tdod_pss_args a) {
    extern __shared__ __align__(16) unsigned char dyn[];
    __shared__ ModeBShared sh;
    const int obj = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int frame = a.obj_frame ? a.obj_frame[obj] : 0;
    const int n = a.n_prop, H = a.H, W = a.W;
    const float* cubes = a.all_cubes + (size_t)obj * n * 15;
    if (a.mode_flags & TDOD_MODEB_SEG) {
        const int Hs = (H + 3) >> 2, Ws = (W + 3) >> 2, stride = Ws + 1;
        uint16_t* pre = reinterpret_cast<uint16_t*>(dyn);
        __shared__ float sK[9];
        if (tid == 0) sh.mask_total = 0;
        if (tid < 9) sK[tid] = a.K[(size_t)frame * 9 + tid];
        __syncthreads();
        build_sampled_prefix(a.masks + (size_t)obj * H * W, H, W, pre, &sh.mask_total, wid, lane, kMW);
        __syncthreads();
        const Clamp cl = make_clamp(W, H);
        const int mask_total = sh.mask_total;
        float* os = a.all_scores + (size_t)obj * a.score_channels * n;
        Edge* ed = sh.edges[wid];
        for (int i = wid + kMW * (int)blockIdx.y; i < n; i += kMW * (int)gridDim.y) {   // one warp per proposal
            float q[15];
#pragma unroll
            for (int k = 0; k < 15; ++k) q[k] = cubes[(size_t)i * 15 + k];
            float X[8], Y[8], Z[8], U[8], V[8];
            cube_corners(q, X, Y, Z);
#pragma unroll
            for (int k = 0; k < 8; ++k) project(sK, cl, X[k], Y[k], Z[k], U[k], V[k]);
            int hx[8], hy[8];
            const int nv = hull8(U, V, hx, hy);               // every lane, same result
            int ymin = 0x7fffffff, ymax = -0x7fffffff;
            for (int k = 0; k < nv; ++k) { ymin = min(ymin, hy[k]); ymax = max(ymax, hy[k]); }
            __syncwarp();                                     // the previous proposal's edges are no longer read
            if (lane < nv) {
                int px = hx[0], py = hy[0], qx = hx[0], qy = hy[0];
#pragma unroll
                for (int k = 0; k < 8; ++k) {                 // register arrays: select without dynamic indexing
                    if (k == lane) { qx = hx[k]; qy = hy[k]; }
                    if (k == (lane + nv - 1) % nv) { px = hx[k]; py = hy[k]; }
                }
                Edge e;
                make_edge(W, H, px, py, qx, qy, e);
                ed[lane] = e;
            }
            __syncwarp();
            i64 inter, poly;
            warp_raster_counts(ed, nv, ymin, ymax, pre, stride, Hs, W, lane, inter, poly);
            if (lane == 0) {
                os[4 * (size_t)n + i] = iou_from_counts(inter, poly, mask_total, false);
                os[5 * (size_t)n + i] = iou_from_counts(inter, poly, mask_total, true);
            }
        }
        __syncthreads();                                      // the table's memory is reused by the point phase
    }
    if (a.mode_flags & TDOD_MODEB_PC) {
        if (a.points_f64) point_cloud_phase<double>(a, obj, frame, dyn, sh);
        else point_cloud_phase<float>(a, obj, frame, dyn, sh);
    }
}

tdod_smem_cache g_sm_modeb;

}  // namespace

// called by tdod_propose_score_select after the mode-A kernel, same stream
int tdod_modeb_launch(const tdod_pss_args& a, void* stream) {
    if (!a.all_cubes) return tdod_fail(TDOD_EINVAL, "mode B reads the proposals back: all_cubes must be given");
    if ((a.mode_flags & TDOD_MODEB_SEG) && (!a.masks || !a.all_scores || a.score_channels < 6))
        return tdod_fail(TDOD_EINVAL, "mode B segmentation needs masks and all_scores with 6 channels");
    if ((a.mode_flags & TDOD_MODEB_PC) && (!a.points || !a.n_points || !a.all_pc || a.P_cap <= 0))
        return tdod_fail(TDOD_EINVAL, "mode B point cloud needs points [F,3,P_cap], n_points [F] and all_pc");
    size_t smem = 0;
    if (a.mode_flags & TDOD_MODEB_SEG) {
        const int Hs = (a.H + 3) >> 2, Ws = (a.W + 3) >> 2;
        smem = (size_t)Hs * (Ws + 1) * sizeof(uint16_t);
    }
    if (a.mode_flags & TDOD_MODEB_PC) {
        const size_t pc = (size_t)3 * kPtChunk * (a.points_f64 ? 8 : 4);
        smem = smem > pc ? smem : pc;
    }
    if (smem > 200 * 1024) return tdod_fail(TDOD_ERANGE, "mode B: sampled mask does not fit in shared memory");
    const cudaError_t e = tdod_smem_opt_in((const void*)modeb_kernel, smem, &g_sm_modeb);
    if (e != cudaSuccess) return (int)e;
    // few objects: an object's proposals (segmentation) and point chunks (point cloud) are dealt to `splits` CTAs, so that a
    // small batch still fills the device (one CTA per object takes ~3 ms at n = 1000 whatever the batch size)
    int splits = 1;
    while (splits < 8 && (int64_t)a.n_obj * splits < 4 * 148 && kMW * splits * 2 <= a.n_prop) splits *= 2;
    if (splits > 1 && (a.mode_flags & TDOD_MODEB_PC)) {
        const cudaError_t em = cudaMemsetAsync(a.all_pc, 0, (size_t)a.n_obj * a.n_prop * sizeof(int32_t), (cudaStream_t)stream);
        if (em != cudaSuccess) return (int)em;
    }
    modeb_kernel<<<dim3((unsigned)a.n_obj, (unsigned)splits), kMT, smem, (cudaStream_t)stream>>>(a);
    TDOD_CHECK_LAUNCH();
    return TDOD_OK;
}
