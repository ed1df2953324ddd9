// K-D: segmented descending order / top-k of per-object scores, and the MABO curves built on it.
//
//   tdod_segmented_order   np.argsort(scores)[::-1][:k] per object (cubercnn/modeling/roi_heads/roi_heads.py:278, and the
//                          top-10 refinement stage :585-586)
//   tdod_accumulate_scores ROIHeads_Boxer.accumulate_scores (roi_heads.py:277-281): IoU3D reordered by descending score,
//                          then its running maximum -- called 8 + 26 times per object by the MABO branch (:543-579)
//
// One CTA per (object, channel).  The n scores are spread over the threads (element i lives in thread i % T, register
// slot i / T); a bitonic network sorts them: compare-exchange partners at distance < 32 are reached with warp shuffles,
// partners in other warps of the same slot through shared memory, partners in another slot of the same thread in
// registers.  Order: value descending with NaN first (np.argsort puts NaN last, [::-1] reverses it); equal values by
// DESCENDING index (what a stable ascending sort followed by [::-1] gives; NumPy's default introsort leaves the order
// of ties unspecified, so parity with the reference is up to permutations of exactly tied scores).
#include "tdod_device.cuh"
#include "tdod_internal.h"

using namespace tdod;

namespace {

// ascending-orderable image of a double; every NaN maps to the top (np.argsort: NaN last)
__device__ __forceinline__ uint64_t d2key(double v) {
    if (v != v) return ~0ull;
    const uint64_t u = (uint64_t)__double_as_longlong(v);
    return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
}

struct Elem { uint64_t key; int idx; };
// true when a comes BEFORE b in the output (descending key, then descending index); padding has idx = -1 and key 0
__device__ __forceinline__ bool before(const Elem& a, const Elem& b) {
    return a.key > b.key || (a.key == b.key && a.idx > b.idx);
}

template <int E>
__device__ __forceinline__ void bitonic_sort(Elem (&v)[E], int T, uint64_t* sk, int* si) {
    const int tid = threadIdx.x;
    const int n_pad = E * T;
    for (int k = 2; k <= n_pad; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            if (j >= T) {                                      // partner in another register slot of this thread
                const int je = j / T;
#pragma unroll
                for (int e = 0; e < E; ++e) {
                    const int pe = e ^ je;
                    if (pe > e) {
                        const int i = e * T + tid;
                        const bool up = (i & k) == 0;          // this pair sorts "before-first" when up
                        const bool sw = up ? before(v[pe], v[e]) : before(v[e], v[pe]);
                        if (sw) { const Elem t = v[e]; v[e] = v[pe]; v[pe] = t; }
                    }
                }
            } else if (j >= 32) {                              // partner in another warp: through shared memory
#pragma unroll
                for (int e = 0; e < E; ++e) {
                    __syncthreads();
                    sk[tid] = v[e].key; si[tid] = v[e].idx;
                    __syncthreads();
                    Elem o; o.key = sk[tid ^ j]; o.idx = si[tid ^ j];
                    const int i = e * T + tid;
                    const bool up = (i & k) == 0, lower = (tid & j) == 0;
                    // the lower position of an "up" pair keeps whichever comes first
                    const bool take = (lower == up) ? before(o, v[e]) : before(v[e], o);
                    if (take) v[e] = o;
                }
            } else {                                           // partner in this warp: shuffles
#pragma unroll
                for (int e = 0; e < E; ++e) {
                    Elem o;
                    o.key = __shfl_xor_sync(0xffffffffu, v[e].key, j);
                    o.idx = __shfl_xor_sync(0xffffffffu, v[e].idx, j);
                    const int i = e * T + tid;
                    const bool up = (i & k) == 0, lower = (tid & j) == 0;
                    const bool take = (lower == up) ? before(o, v[e]) : before(v[e], o);
                    if (take) v[e] = o;
                }
            }
        }
    }
}

// kCurve == false: out_idx / out_val [S,k].  kCurve == true: scores are doubles [S*C, n], iou3d [S,n], curves [S*C, k].
template <int E, bool kCurve>
__global__ void __launch_bounds__(1024) order_kernel(const void* __restrict__ scores, const float* __restrict__ iou3d, int C,
                                                     int n, int k, int32_t* __restrict__ out_idx, float* __restrict__ out_val) {
    extern __shared__ __align__(16) unsigned char sm_raw[];
    const int T = blockDim.x, tid = threadIdx.x;
    uint64_t* sk = reinterpret_cast<uint64_t*>(sm_raw);
    int* si = reinterpret_cast<int*>(sm_raw + (size_t)T * 8);
    float* run = reinterpret_cast<float*>(sm_raw + (size_t)T * 12);   // [T/32] warp carries of the running maximum
    const int64_t seg = blockIdx.x;
    Elem v[E];
#pragma unroll
    for (int e = 0; e < E; ++e) {
        const int i = e * T + tid;
        v[e].key = 0ull; v[e].idx = -1;                        // padding sorts last
        if (i < n) {
            const double s = kCurve ? reinterpret_cast<const double*>(scores)[seg * n + i]
                                    : (double)reinterpret_cast<const float*>(scores)[seg * n + i];
            v[e].key = d2key(s); v[e].idx = i;
            // key 0 is reserved for the padding: the smallest real key is that of -inf... which is 0x000f.. > 0
        }
    }
    bitonic_sort<E>(v, T, sk, si);
    // position of v[e] in the output: e * T + tid
    if (!kCurve) {
#pragma unroll
        for (int e = 0; e < E; ++e) {
            const int pos = e * T + tid;
            if (pos < k && v[e].idx >= 0) {
                out_idx[seg * k + pos] = v[e].idx;
                if (out_val) out_val[seg * k + pos] = reinterpret_cast<const float*>(scores)[seg * n + v[e].idx];
            }
        }
        return;
    }
    // running maximum of IoU3D in sorted order (np.maximum.accumulate propagates NaN: max_nan)
    const float* iu = iou3d + (seg / C) * n;
    float carry = -CUDART_INF_F;
    bool carry_valid = false;
#pragma unroll
    for (int e = 0; e < E; ++e) {
        const int pos = e * T + tid;
        float x = (v[e].idx >= 0) ? iu[v[e].idx] : -CUDART_INF_F;
        // inclusive max-scan over the warp, then over the warps
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const float y = __shfl_up_sync(0xffffffffu, x, o);
            if ((tid & 31) >= o) x = max_nan(x, y);
        }
        __syncthreads();
        if ((tid & 31) == 31) run[tid >> 5] = x;
        __syncthreads();
        float pre = carry;
        bool have = carry_valid;
        for (int w = 0; w < (tid >> 5); ++w) { pre = have ? max_nan(pre, run[w]) : run[w]; have = true; }
        if (have) x = max_nan(x, pre);
        float tot = carry; bool th = carry_valid;
        for (int w = 0; w < (T >> 5); ++w) { tot = th ? max_nan(tot, run[w]) : run[w]; th = true; }
        carry = tot; carry_valid = true;
        if (pos < k && pos < n) out_val[seg * k + pos] = x;
    }
}

template <bool kCurve>
int launch_order(const void* scores, const float* iou3d, int64_t segs, int C, int n, int k, int32_t* out_idx, float* out_val,
                 cudaStream_t st) {
    int n_pad = 32;
    while (n_pad < n) n_pad <<= 1;
    const int T = n_pad < 1024 ? n_pad : 1024;
    const int E = n_pad / T;
    const size_t smem = (size_t)T * 12 + 32 * 4;
    if (segs > 0x7fffffff) return tdod_fail(TDOD_ERANGE, "too many segments for one launch");
#define TDOD_ORDER_CASE(EE)                                                                                          \
    case EE:                                                                                                         \
        order_kernel<EE, kCurve><<<(unsigned)segs, T, smem, st>>>(scores, iou3d, C, n, k, out_idx, out_val);         \
        break;
    switch (E) {
        TDOD_ORDER_CASE(1) TDOD_ORDER_CASE(2) TDOD_ORDER_CASE(4) TDOD_ORDER_CASE(8) TDOD_ORDER_CASE(16)
        default: return tdod_fail(TDOD_ERANGE, "segment longer than 16384 elements");
    }
#undef TDOD_ORDER_CASE
    TDOD_CHECK_LAUNCH();
    return TDOD_OK;
}

}  // namespace

extern "C" int tdod_segmented_order(const float* scores, int64_t S, int n, int k, int32_t* out_idx, float* out_val,
                                    void* stream) {
    if (S == 0) return TDOD_OK;
    if (S < 0 || n <= 0 || k <= 0 || k > n) return tdod_fail(TDOD_EINVAL, "tdod_segmented_order: need S >= 0, 0 < k <= n");
    if (!scores || !out_idx) return tdod_fail(TDOD_EINVAL, "tdod_segmented_order: a required pointer is NULL");
    return launch_order<false>(scores, nullptr, S, 1, n, k, out_idx, out_val, (cudaStream_t)stream);
}

extern "C" int tdod_accumulate_scores(const double* scores, const float* iou3d, int64_t S, int C, int n, int n_out,
                                      float* out_curves, void* stream) {
    if (S == 0 || C == 0) return TDOD_OK;
    if (S < 0 || C < 0 || n <= 0 || n_out <= 0 || n_out > n)
        return tdod_fail(TDOD_EINVAL, "tdod_accumulate_scores: need S, C >= 0 and 0 < n_out <= n");
    if (!scores || !iou3d || !out_curves) return tdod_fail(TDOD_EINVAL, "tdod_accumulate_scores: a required pointer is NULL");
    return launch_order<true>(scores, iou3d, S * C, C, n, n_out, nullptr, out_curves, (cudaStream_t)stream);
}
