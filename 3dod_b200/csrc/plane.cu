// K-A: depth -> strided point cloud, RANSAC ground plane, up-vector rectification.
//
// Restates cubercnn/modeling/roi_heads/roi_heads.py:345-428 (NumPy / pyransac3d flavour, fp64) and :1084-1136
// (torch flavour, fp32) with ProposalNetwork/utils/plane.py:79-134 (fit_parallel) / :154-209 (Plane_np.fit ==
// pyransac3d.Plane.fit).  The 1000 x P distance matrix of fit_parallel is never materialised: every CTA keeps a
// few points per thread in registers, streams the hypotheses from shared memory and counts inliers with
// warp ballots.
#include <cstdlib>
#include <cooperative_groups.h>
#include "tdod_device.cuh"
#include "tdod_internal.h"

using namespace tdod;

namespace {

// separately rounded arithmetic (each is its own NumPy / torch op in the reference)
template <typename T> struct Ar;
template <> struct Ar<float> {
    static __device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
    static __device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
    static __device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
    static __device__ __forceinline__ float div(float a, float b) { return __fdiv_rn(a, b); }
    static __device__ __forceinline__ float sqrt(float a) { return __fsqrt_rn(a); }
    static __device__ __forceinline__ float up(float a) { return nextafterf(a, CUDART_INF_F); }
    static __device__ __forceinline__ float down(float a) { return nextafterf(a, -CUDART_INF_F); }
};
template <> struct Ar<double> {
    static __device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
    static __device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
    static __device__ __forceinline__ double sub(double a, double b) { return __dsub_rn(a, b); }
    static __device__ __forceinline__ double div(double a, double b) { return __ddiv_rn(a, b); }
    static __device__ __forceinline__ double sqrt(double a) { return __dsqrt_rn(a); }
    static __device__ __forceinline__ double up(double a) { return nextafter(a, CUDART_INF); }
    static __device__ __forceinline__ double down(double a) { return nextafter(a, -CUDART_INF); }
};

// ------------------------------------------------------------------------------------------------
// back-projection with row-major ordered compaction: one cluster of kBpCluster CTAs per frame.
// CTA `rank` owns the contiguous slice [rank * per, (rank + 1) * per) of the strided grid.  Pass A reads only the
// ground samples and counts "on" / "off" points per (iteration, warp); the slice totals are exchanged through
// distributed shared memory, so every CTA knows where its slice starts in the two compacted outputs; pass B reads
// depth (once) and ground (again, now from L1/L2), back-projects and writes each point at its final position.
// ------------------------------------------------------------------------------------------------
#ifndef TDOD_BP_MINB
#define TDOD_BP_MINB 3
#endif
constexpr int kBpCluster = 8;
constexpr int kBpT = 512;
constexpr int kBpRegIt = 4;             // strided samples per thread the one-pass variant keeps in registers

template <typename T, int FLAVOUR>
__global__ void __cluster_dims__(kBpCluster, 1, 1) __launch_bounds__(kBpT, TDOD_BP_MINB)
    backproject_kernel(const float* __restrict__ depth, const uint8_t* __restrict__ ground, const float* __restrict__ Kall,
                       int H, int W, int stride, int iters, T* __restrict__ plane_pts, T* __restrict__ other_pts,
                       int32_t* __restrict__ counts) {
    extern __shared__ int bp_cnt[];            // [iters][kBpT/32] "on" counts, then the same for "off"; later their exclusive scans
    __shared__ int s_tot[2];                   // this CTA's totals, read by the whole cluster
    __shared__ int s_base[2];
    namespace cgx = cooperative_groups;
    cgx::cluster_group cl = cgx::this_cluster();
    constexpr int kW = kBpT / 32;
    const int rank = (int)cl.block_rank();
    const int f = blockIdx.x / kBpCluster, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int Hs = (H + stride - 1) / stride, Ws = (W + stride - 1) / stride;
    const int64_t P = (int64_t)Hs * Ws;
    const int64_t per = (int64_t)iters * kBpT;
    const int64_t lo = rank * per, hi = min(P, lo + per);
    const float* K = Kall + (size_t)f * 9;
    const float* dp = depth + (size_t)f * H * W;
    const uint8_t* gp = ground ? ground + (size_t)f * H * W : nullptr;
    int* c_on = bp_cnt;
    int* c_off = bp_cnt + iters * kW;
    if (iters <= kBpRegIt) {
        // ---- frames of up to kBpRegIt * 4096 samples (every BASELINE shape but 1920 x 1080): ONE pass.  All ground loads
        // are issued together, then all depth loads; the points wait in registers for the slice offsets, so a CTA pays two
        // load latencies instead of eight and the second read of the ground map disappears ----
        T fx, fy, cx, cy;
        if (FLAVOUR == TDOD_BACKPROJECT_NUMPY) {
            fx = (T)K[0]; fy = (T)K[4]; cx = (T)K[2]; cy = (T)K[5];
        } else {
            fx = (T)floorf(__fdiv_rn(K[0], (float)stride)); fy = (T)floorf(__fdiv_rn(K[4], (float)stride));
            cx = (T)((double)Ws / 2.0); cy = (T)((double)Hs / 2.0);
        }
        const bool want_off = other_pts != nullptr;
        bool valid[kBpRegIt], on[kBpRegIt];
        int uu[kBpRegIt], vv[kBpRegIt];
        uint8_t g[kBpRegIt];
        float zf[kBpRegIt];
#pragma unroll
        for (int it = 0; it < kBpRegIt; ++it) {
            const int64_t idx = lo + (int64_t)it * kBpT + tid;
            valid[it] = it < iters && idx < hi;
            const uint32_t i32 = valid[it] ? (uint32_t)idx : 0u;
            vv[it] = (int)(i32 / (uint32_t)Ws); uu[it] = (int)(i32 - (uint32_t)vv[it] * (uint32_t)Ws);
            g[it] = 1;
            if (valid[it] && gp) g[it] = __ldg(gp + (size_t)(vv[it] * stride) * W + (size_t)(uu[it] * stride));
        }
#pragma unroll
        for (int it = 0; it < kBpRegIt; ++it) {
            on[it] = valid[it] && g[it] > 0;
            zf[it] = 0.f;
            if (valid[it] && (on[it] || want_off)) zf[it] = __ldg(dp + (size_t)(vv[it] * stride) * W + (size_t)(uu[it] * stride));
        }
        unsigned m_on[kBpRegIt], m_off[kBpRegIt];
#pragma unroll
        for (int it = 0; it < kBpRegIt; ++it) {
            m_on[it] = __ballot_sync(0xffffffffu, on[it]); m_off[it] = __ballot_sync(0xffffffffu, valid[it] && !on[it]);
            if (lane == 0 && it < iters) { c_on[it * kW + wid] = __popc(m_on[it]); c_off[it * kW + wid] = __popc(m_off[it]); }
        }
        __syncthreads();
        if (wid < 2) {                               // iters * kW <= 64 entries per list: two rounds of a warp scan
            int* c = wid == 0 ? c_on : c_off;
            const int n = iters * kW;
            int carry = 0;
            for (int b0 = 0; b0 < n; b0 += 32) {
                const int v = (b0 + lane < n) ? c[b0 + lane] : 0;
                int inc = v;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const int t = __shfl_up_sync(0xffffffffu, inc, o);
                    if (lane >= o) inc += t;
                }
                if (b0 + lane < n) c[b0 + lane] = carry + inc - v;
                carry += __shfl_sync(0xffffffffu, inc, 31);
            }
            if (lane == 0) s_tot[wid] = carry;
        }
        cl.sync();
        if (tid < 2) {
            int base = 0;
            for (int r = 0; r < rank; ++r) base += *cl.map_shared_rank(&s_tot[tid], r);
            s_base[tid] = base;
            if (rank == kBpCluster - 1) counts[(size_t)tid * (gridDim.x / kBpCluster) + f] = base + s_tot[tid];   // [2,F]
        }
        cl.barrier_arrive();                         // (peers may leave once everybody has read their totals)
        __syncthreads();
        T* po = plane_pts + (size_t)f * 3 * P;
        T* oo = other_pts ? other_pts + (size_t)f * 3 * P : nullptr;
        const int base_on = s_base[0], base_off = s_base[1];
        const unsigned lt = (1u << lane) - 1u;
#pragma unroll
        for (int it = 0; it < kBpRegIt; ++it) {
            if (!valid[it] || !(on[it] || want_off)) continue;
            const T z = (T)zf[it];
            const T x = Ar<T>::div(Ar<T>::mul(Ar<T>::sub((T)uu[it], cx), z), fx);
            const T y = Ar<T>::div(Ar<T>::mul(Ar<T>::sub((T)vv[it], cy), z), fy);
            if (on[it]) {
                const int p = base_on + c_on[it * kW + wid] + __popc(m_on[it] & lt);
                po[p] = x; po[P + p] = y; po[2 * P + p] = z;
            } else {
                const int p = base_off + c_off[it * kW + wid] + __popc(m_off[it] & lt);
                oo[p] = x; oo[P + p] = y; oo[2 * P + p] = z;
            }
        }
        cl.barrier_wait();
        return;
    }
    // ---- pass A: counts ----
    for (int it = 0; it < iters; ++it) {
        const int64_t idx = lo + (int64_t)it * kBpT + tid;
        const bool valid = idx < hi;
        bool on = valid;
        if (valid && gp) {
            const int v = (int)((uint32_t)idx / (uint32_t)Ws), u = (int)((uint32_t)idx - (uint32_t)v * (uint32_t)Ws);   // idx < P < 2^31 (counts are int32)
            on = __ldg(gp + (size_t)(v * stride) * W + (size_t)(u * stride)) > 0;
        }
        const unsigned m_on = __ballot_sync(0xffffffffu, valid && on), m_off = __ballot_sync(0xffffffffu, valid && !on);
        if (lane == 0) { c_on[it * kW + wid] = __popc(m_on); c_off[it * kW + wid] = __popc(m_off); }
    }
    __syncthreads();
    // exclusive scan of the iters * kW entries of each list: warp 0 scans "on", warp 1 scans "off"
    if (wid < 2) {
        int* c = wid == 0 ? c_on : c_off;
        const int n = iters * kW;
        int carry = 0;
        for (int b0 = 0; b0 < n; b0 += 32) {
            const int v = (b0 + lane < n) ? c[b0 + lane] : 0;
            int inc = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += t;
            }
            if (b0 + lane < n) c[b0 + lane] = carry + inc - v;
            carry += __shfl_sync(0xffffffffu, inc, 31);
        }
        if (lane == 0) s_tot[wid] = carry;
    }
    cl.sync();
    if (tid < 2) {
        int base = 0;
        for (int r = 0; r < rank; ++r) base += *cl.map_shared_rank(&s_tot[tid], r);
        s_base[tid] = base;
        if (rank == kBpCluster - 1) counts[(size_t)tid * (gridDim.x / kBpCluster) + f] = base + s_tot[tid];   // [2,F]
    }
    __syncthreads();
    // ---- pass B: back-project and write in place ----
    T fx, fy, cx, cy;
    if (FLAVOUR == TDOD_BACKPROJECT_NUMPY) {   // roi_heads.py:347-352: full-resolution intrinsics
        fx = (T)K[0]; fy = (T)K[4]; cx = (T)K[2]; cy = (T)K[5];
    } else {                                   // roi_heads.py:1089-1092: K // stride, centre of the strided grid
        fx = (T)floorf(__fdiv_rn(K[0], (float)stride)); fy = (T)floorf(__fdiv_rn(K[4], (float)stride));
        cx = (T)((double)Ws / 2.0); cy = (T)((double)Hs / 2.0);
    }
    T* po = plane_pts + (size_t)f * 3 * P;
    T* oo = other_pts ? other_pts + (size_t)f * 3 * P : nullptr;
    const int base_on = s_base[0], base_off = s_base[1];
    const unsigned lt = (1u << lane) - 1u;
    for (int it = 0; it < iters; ++it) {
        const int64_t idx = lo + (int64_t)it * kBpT + tid;
        const bool valid = idx < hi;
        bool on = valid;
        T x = 0, y = 0, z = 0;
        if (valid) {
            const int v = (int)((uint32_t)idx / (uint32_t)Ws), u = (int)((uint32_t)idx - (uint32_t)v * (uint32_t)Ws);   // idx < P < 2^31 (counts are int32)
            const size_t pix = (size_t)(v * stride) * W + (size_t)(u * stride);
            if (gp) on = __ldg(gp + pix) > 0;
            if (on || oo) {
                z = (T)__ldg(dp + pix);
                x = Ar<T>::div(Ar<T>::mul(Ar<T>::sub((T)u, cx), z), fx);
                y = Ar<T>::div(Ar<T>::mul(Ar<T>::sub((T)v, cy), z), fy);
            }
        }
        const unsigned m_on = __ballot_sync(0xffffffffu, valid && on), m_off = __ballot_sync(0xffffffffu, valid && !on);
        if (valid && on) {
            const int p = base_on + c_on[it * kW + wid] + __popc(m_on & lt);
            po[p] = x; po[P + p] = y; po[2 * P + p] = z;
        } else if (valid && oo) {
            const int p = base_off + c_off[it * kW + wid] + __popc(m_off & lt);
            oo[p] = x; oo[P + p] = y; oo[2 * P + p] = z;
        }
    }
    cl.sync();   // nobody leaves while a peer may still read its totals
}

// ------------------------------------------------------------------------------------------------
// hypotheses: sample ids -> plane (a,b,c,k) and the largest |numerator| that is still an inlier
// ------------------------------------------------------------------------------------------------
template <typename T>
struct PlaneHyp {
    T a, b, c, k, bound;   // inlier  <=>  |a x + b y + c z + k| <= bound   (bound folds the division by ||n||)
};

template <typename T>
__global__ void hypotheses_kernel(const T* __restrict__ pts, const int32_t* __restrict__ n_pts, int64_t P_cap,
                                  const int32_t* __restrict__ ids, uint64_t seed, int64_t frame_id_base, int iters,
                                  T thresh, PlaneHyp<T>* __restrict__ hyps, int32_t* __restrict__ counts) {
    const int f = blockIdx.y;
    const int it = blockIdx.x * blockDim.x + threadIdx.x;
    if (it >= iters) return;
    const int P = n_pts[f];
    PlaneHyp<T> h;
    counts[(size_t)f * iters + it] = 0;
    if (P < 3) {   // random.sample would raise in the reference; report "no plane"
        h.a = h.b = h.c = h.k = 0; h.bound = (T)-1;
        hyps[(size_t)f * iters + it] = h;
        return;
    }
    int i0, i1, i2;
    if (ids) {
        const int32_t* s = ids + ((size_t)f * iters + it) * 3;
        i0 = s[0]; i1 = s[1]; i2 = s[2];
    } else {   // three distinct uniform indices
        const Philox ph(seed);
        const uint64_t gf = (uint64_t)(frame_id_base + f);   // global frame id: results do not depend on batching
        const uint4 r = ph((uint32_t)it, (uint32_t)gf, (uint32_t)(gf >> 32) ^ 0x52414Eu, 0u);
        i0 = (int)(r.x % (uint32_t)P);
        i1 = (int)(r.y % (uint32_t)(P - 1)); if (i1 >= i0) ++i1;
        i2 = (int)(r.z % (uint32_t)(P - 2));
        const int lo = min(i0, i1), hi = max(i0, i1);
        if (i2 >= lo) ++i2;
        if (i2 >= hi) ++i2;
    }
    const T* X = pts + (size_t)f * 3 * P_cap;
    const T* Y = X + P_cap;
    const T* Z = Y + P_cap;
    const T p0x = X[i0], p0y = Y[i0], p0z = Z[i0], p1x = X[i1], p1y = Y[i1], p1z = Z[i1], p2x = X[i2], p2y = Y[i2],
            p2z = Z[i2];
    typedef Ar<T> A;
    const T ax = A::sub(p1x, p0x), ay = A::sub(p1y, p0y), az = A::sub(p1z, p0z);       // plane.py:104 / :180
    const T bx = A::sub(p2x, p0x), by = A::sub(p2y, p0y), bz = A::sub(p2z, p0z);       // :105 / :181
    T cx = A::sub(A::mul(ay, bz), A::mul(az, by));                                     // cross, :108 / :184
    T cy = A::sub(A::mul(az, bx), A::mul(ax, bz));
    T cz = A::sub(A::mul(ax, by), A::mul(ay, bx));
    const T nrm = A::sqrt(A::add(A::add(A::mul(cx, cx), A::mul(cy, cy)), A::mul(cz, cz)));
    cx = A::div(cx, nrm); cy = A::div(cy, nrm); cz = A::div(cz, nrm);                   // :112 / :188
    const T k = -A::add(A::add(A::mul(cx, p1x), A::mul(cy, p1y)), A::mul(cz, p1z));    // :113 / :189
    const T den = A::sqrt(A::add(A::add(A::mul(cx, cx), A::mul(cy, cy)), A::mul(cz, cz)));   // :119 / :197
    // largest t with fl(t / den) <= thresh (IEEE division is monotone in the numerator)
    T t = A::mul(thresh, den);
    if (!(den == den) || !(t == t)) {
        t = (T)-1;   // degenerate (collinear) sample: NaN plane, no inliers (NaN <= thresh is false)
    } else {
        for (int g = 0; g < 8 && A::div(t, den) > thresh; ++g) t = A::down(t);
        for (int g = 0; g < 8 && A::div(A::up(t), den) <= thresh; ++g) t = A::up(t);
    }
    h.a = cx; h.b = cy; h.c = cz; h.k = k; h.bound = t;
    hyps[(size_t)f * iters + it] = h;
}

// ------------------------------------------------------------------------------------------------
// inlier counting: grid (point tiles, frames); PT points per thread in registers, hypotheses in smem
// ------------------------------------------------------------------------------------------------
template <typename T, int PT>
__global__ void __launch_bounds__(256) count_kernel(const T* __restrict__ pts, const int32_t* __restrict__ n_pts,
                                                    int64_t P_cap, const PlaneHyp<T>* __restrict__ hyps, int iters,
                                                    int32_t* __restrict__ counts) {
    extern __shared__ __align__(16) unsigned char smem[];
    PlaneHyp<T>* sh = reinterpret_cast<PlaneHyp<T>*>(smem);
    int* scount = reinterpret_cast<int*>(sh + iters);
    const int f = blockIdx.y, tid = threadIdx.x, lane = tid & 31;
    const int P = n_pts[f];
    const int tile0 = blockIdx.x * (256 * PT);
    if (tile0 >= P) return;
    for (int i = tid; i < iters; i += 256) { sh[i] = hyps[(size_t)f * iters + i]; scount[i] = 0; }
    const T* X = pts + (size_t)f * 3 * P_cap;
    const T* Y = X + P_cap;
    const T* Z = Y + P_cap;
    T px[PT], py[PT], pz[PT];
    bool ok[PT];
#pragma unroll
    for (int q = 0; q < PT; ++q) {
        const int i = tile0 + q * 256 + tid;
        ok[q] = i < P;
        px[q] = ok[q] ? X[i] : (T)0; py[q] = ok[q] ? Y[i] : (T)0; pz[q] = ok[q] ? Z[i] : (T)0;
    }
    __syncthreads();
    typedef Ar<T> A;
    for (int h = 0; h < iters; ++h) {
        const PlaneHyp<T> p = sh[h];
        int c = 0;
#pragma unroll
        for (int q = 0; q < PT; ++q) {
            // plane.py:116-118 / :194-196: ((a x + b y) + c z) + k, every product and sum rounded on its own
            const T s = A::add(A::add(A::add(A::mul(p.a, px[q]), A::mul(p.b, py[q])), A::mul(p.c, pz[q])), p.k);
            const bool in = ok[q] && (fabs(s) <= p.bound);
            c += __popc(__ballot_sync(0xffffffffu, in));
        }
        if (lane == 0 && c) atomicAdd(&scount[h], c);
    }
    __syncthreads();
    for (int i = tid; i < iters; i += 256)
        if (scount[i]) atomicAdd(&counts[(size_t)f * iters + i], scount[i]);
}

// ------------------------------------------------------------------------------------------------
// inlier counting, filtered: fp32 FMA evaluation with a rigorous error band, exact re-evaluation of the rare
// (hypothesis, tile) pairs that have a point inside the band.  Same counts as count_kernel, bit for bit.
//
// Layout: a CTA owns one tile of <= kTilePts points of one frame (fp32 copies in shared memory, each value
// duplicated so that it feeds both halves of a packed f32x2 FMA) and ALL hypotheses of the frame, four per thread in
// registers.  The inner loop broadcasts one point to the CTA and every thread advances its four hypotheses:
//   s'  = (a x + b y + c z + k) / 2e               (FFMA2: two hypotheses per instruction; the plane is pre-scaled)
//   t   = sat((bound + e) / 2e - |s'|)             one FADD.SAT: 1 = certainly in, 0 = certainly out, between = in the band
//   S  += t                                        (packed add)
//   mn  = min(mn, bits(t) - 1)  (unsigned)         one VIADDMNMX on the ALU pipe: below 0x3f7fffff iff some t was fractional
// The fma pipe issues one warp instruction every other cycle and was the limiter (three packed FMAs, a packed square, two
// FFMA.SAT, a packed add and a packed FMA per pair of hypotheses); the band test now runs on the otherwise idle ALU pipe
// and the ramp is linear in |s|, which takes the fma pipe from eight to six instructions per pair and point.
// The band: the fp32 value differs from the separately rounded fp64 / fp32 value of the reference by at most
// E = 5e-7 (max|x|+|y|+|z| + max|k| + bound) (inputs and the scaled plane rounded once, three fused operations,
// |a|,|b|,|c| <= 1, two roundings in the ramp's constant); with e = 2E, |s| <= bound - e gives t = 1 and |s| >= bound + e
// gives t = 0, both exactly (bit patterns 0x3f800000 and 0), so a tile without a fractional t has no undecided point and
// rint(S) is its count (sums of <= 1024 ones are exact).  Tiles with one are recounted exactly.
// ------------------------------------------------------------------------------------------------
constexpr int kTilePts = 1024;     // points per CTA tile
constexpr int kFiltT = 256;        // threads
constexpr int kHypT = 4;           // hypotheses per thread per pass

__device__ __forceinline__ float2 ffma2(const float2 a, const float2 b, const float2 c) {
    float2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;"
        : "=l"(reinterpret_cast<unsigned long long&>(d))
        : "l"(reinterpret_cast<const unsigned long long&>(a)), "l"(reinterpret_cast<const unsigned long long&>(b)),
          "l"(reinterpret_cast<const unsigned long long&>(c)));
    return d;
}
__device__ __forceinline__ float2 fmul2(const float2 a, const float2 b) {
    float2 d;
    asm("mul.rn.f32x2 %0, %1, %2;"
        : "=l"(reinterpret_cast<unsigned long long&>(d))
        : "l"(reinterpret_cast<const unsigned long long&>(a)), "l"(reinterpret_cast<const unsigned long long&>(b)));
    return d;
}
__device__ __forceinline__ float2 fadd2(const float2 a, const float2 b) {
    float2 d;
    asm("add.rn.f32x2 %0, %1, %2;"
        : "=l"(reinterpret_cast<unsigned long long&>(d))
        : "l"(reinterpret_cast<const unsigned long long&>(a)), "l"(reinterpret_cast<const unsigned long long&>(b)));
    return d;
}

template <typename T>
__global__ void __launch_bounds__(kFiltT) count_filter_kernel(const T* __restrict__ pts, const int32_t* __restrict__ n_pts,
                                                             int64_t P_cap, const PlaneHyp<T>* __restrict__ hyps,
                                                             int iters, int32_t* __restrict__ counts, int tile_pts) {
    __shared__ float4 s_xy[kTilePts];      // (x, x, y, y)
    __shared__ float2 s_z[kTilePts];       // (z, z)
    __shared__ int s_list[kFiltT * kHypT];
    __shared__ int s_nlist, s_mmax, s_kmax;
    __shared__ int s_cnt[kFiltT / 32];
    const int f = blockIdx.y, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int P = n_pts[f];
    const int tile0 = blockIdx.x * tile_pts;          // tile_pts <= kTilePts: small batches use smaller tiles, more CTAs
    if (tile0 >= P) return;
    const int np = min(tile_pts, P - tile0);
    const T* X = pts + (size_t)f * 3 * P_cap;
    const T* Y = X + P_cap;
    const T* Z = Y + P_cap;
    const PlaneHyp<T>* hf = hyps + (size_t)f * iters;
    if (tid == 0) { s_mmax = 0; s_kmax = 0; }
    __syncthreads();
    // ---- the tile's points, rounded once to fp32 ----
    float mmax = 0.f;
    for (int i = tid; i < np; i += kFiltT) {
        const float x = (float)X[tile0 + i], y = (float)Y[tile0 + i], z = (float)Z[tile0 + i];
        s_xy[i] = make_float4(x, x, y, y);
        s_z[i] = make_float2(z, z);
        float m = fabsf(x) + fabsf(y) + fabsf(z);
        if (!(m == m)) m = CUDART_INF_F;
        mmax = fmaxf(mmax, m);
    }
    // |k| + bound over ALL hypotheses of the frame (not only this pass): one band width per CTA
    float kmax = 0.f;
    for (int i = tid; i < iters; i += kFiltT) {
        const float k = fabsf((float)hf[i].k), b = (float)hf[i].bound;
        if (b >= 0.f) kmax = fmaxf(kmax, k + b);    // NaN / "no plane" hypotheses never count and do not widen the band
    }
    mmax = fmaxf(mmax, __shfl_xor_sync(0xffffffffu, mmax, 16)); kmax = fmaxf(kmax, __shfl_xor_sync(0xffffffffu, kmax, 16));
    mmax = fmaxf(mmax, __shfl_xor_sync(0xffffffffu, mmax, 8));  kmax = fmaxf(kmax, __shfl_xor_sync(0xffffffffu, kmax, 8));
    mmax = fmaxf(mmax, __shfl_xor_sync(0xffffffffu, mmax, 4));  kmax = fmaxf(kmax, __shfl_xor_sync(0xffffffffu, kmax, 4));
    mmax = fmaxf(mmax, __shfl_xor_sync(0xffffffffu, mmax, 2));  kmax = fmaxf(kmax, __shfl_xor_sync(0xffffffffu, kmax, 2));
    mmax = fmaxf(mmax, __shfl_xor_sync(0xffffffffu, mmax, 1));  kmax = fmaxf(kmax, __shfl_xor_sync(0xffffffffu, kmax, 1));
    if (lane == 0) { atomicMax(&s_mmax, __float_as_int(mmax)); atomicMax(&s_kmax, __float_as_int(kmax)); }   // non-negative floats
    __syncthreads();
    const float e = 1.0e-6f * (__int_as_float(s_mmax) + __int_as_float(s_kmax)) + 1e-30f;

    for (int h0 = 0; h0 < iters; h0 += kFiltT * kHypT) {
        if (tid == 0) s_nlist = 0;
        // ---- this thread's hypotheses: h0 + tid + j * kFiltT ----
        float2 a2[kHypT / 2], b2[kHypT / 2], c2[kHypT / 2], k2[kHypT / 2], S[kHypT / 2];
        float C[kHypT];
        uint32_t mn[kHypT];                                       // min over the points of bits(t) - 1 (unsigned)
        const float inv = 0.5f / e;
        const bool band_ok = (e < CUDART_INF_F) && (e > 1e-20f);
#pragma unroll
        for (int j = 0; j < kHypT; ++j) {
            const int h = h0 + tid + j * kFiltT;
            float a = 0.f, b = 0.f, c = 0.f, k = 0.f, bd = -1.f;
            if (h < iters) {
                const PlaneHyp<T> p = hf[h];
                a = (float)p.a; b = (float)p.b; c = (float)p.c; k = (float)p.k; bd = (float)p.bound;
            }
            // t = sat((bound + e - |s|) / 2e): the plane is scaled by 1 / 2e once, the ramp is one FADD.SAT per point
            if (bd < 0.f) { a = b = c = k = 0.f; C[j] = 0.f; }                          // no plane: t = 0 everywhere
            else if (!band_ok || !(bd == bd)) { a = b = c = k = 0.f; C[j] = 0.5f; }     // no usable band: t = 1/2, recount
            else { a *= inv; b *= inv; c *= inv; k *= inv; C[j] = (bd + e) * inv; }
            (j & 1 ? a2[j / 2].y : a2[j / 2].x) = a; (j & 1 ? b2[j / 2].y : b2[j / 2].x) = b;
            (j & 1 ? c2[j / 2].y : c2[j / 2].x) = c; (j & 1 ? k2[j / 2].y : k2[j / 2].x) = k;
            mn[j] = 0xffffffffu;
        }
#pragma unroll
        for (int j = 0; j < kHypT / 2; ++j) S[j] = make_float2(0.f, 0.f);
        __syncthreads();
        // ---- main loop: broadcast a point, advance four hypotheses.  Per pair of hypotheses the fma pipe sees three packed
        // FMAs, two FADD.SAT and one packed add; whether any t was fractional is tracked on the ALU pipe (VIADDMNMX) ----
#pragma unroll 4
        for (int i = 0; i < np; ++i) {
            const float4 q = s_xy[i];
            const float2 z = s_z[i];
            const float2 qx = make_float2(q.x, q.y), qy = make_float2(q.z, q.w);
#pragma unroll
            for (int j = 0; j < kHypT / 2; ++j) {
                float2 s = ffma2(c2[j], z, k2[j]);
                s = ffma2(b2[j], qy, s);
                s = ffma2(a2[j], qx, s);
                float2 t;
                t.x = __saturatef(C[2 * j] - fabsf(s.x));
                t.y = __saturatef(C[2 * j + 1] - fabsf(s.y));
                S[j] = fadd2(S[j], t);
                mn[2 * j] = __viaddmin_u32(__float_as_uint(t.x), 0xffffffffu, mn[2 * j]);
                mn[2 * j + 1] = __viaddmin_u32(__float_as_uint(t.y), 0xffffffffu, mn[2 * j + 1]);
            }
        }
        // ---- decided hypotheses go straight to the counters; the others are queued for the exact recount.  t is exactly
        // 0 or 1 for a decided point (bits 0 / 0x3f800000), so a minimum of bits(t) - 1 below 0x3f7fffff means some point
        // sat inside the band ----
#pragma unroll
        for (int j = 0; j < kHypT; ++j) {
            const int h = h0 + tid + j * kFiltT;
            const float s1 = (j & 1) ? S[j / 2].y : S[j / 2].x;
            if (h < iters) {
                if (mn[j] < 0x3f7fffffu) s_list[atomicAdd(&s_nlist, 1)] = h;
                else {
                    const int c = __float2int_rn(s1);
                    if (c) atomicAdd(&counts[(size_t)f * iters + h], c);
                }
            }
        }
        __syncthreads();
        const int nl = s_nlist;
        typedef Ar<T> A;
        for (int l = 0; l < nl; ++l) {
            const int h = s_list[l];
            const PlaneHyp<T> p = hf[h];
            int c = 0;
            for (int i = tid; i < np; i += kFiltT) {
                const T s = A::add(A::add(A::add(A::mul(p.a, X[tile0 + i]), A::mul(p.b, Y[tile0 + i])), A::mul(p.c, Z[tile0 + i])), p.k);
                c += (fabs(s) <= p.bound) ? 1 : 0;
            }
            c = __reduce_add_sync(0xffffffffu, c);
            if (lane == 0) s_cnt[wid] = c;
            __syncthreads();
            if (tid == 0) {
                int tot = 0;
                for (int w = 0; w < kFiltT / 32; ++w) tot += s_cnt[w];
                if (tot) atomicAdd(&counts[(size_t)f * iters + h], tot);
            }
            __syncthreads();
        }
        __syncthreads();                                // (more than 1024 iterations: s_nlist is reset by the next pass)
    }
}

// ------------------------------------------------------------------------------------------------
// best hypothesis (first maximum, plane.py:71 / :128 / :203) + rectification (roi_heads.py:411-428 / :1120-1136)
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) select_kernel(const PlaneHyp<T>* __restrict__ hyps,
                                                     const int32_t* __restrict__ counts, int iters, int negate,
                                                     T* __restrict__ out_eq, int32_t* __restrict__ out_best,
                                                     int32_t* __restrict__ out_count, T* __restrict__ out_normal) {
    __shared__ int sc[8], si[8];
    const int f = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    int bc = -1, bi = 0x7fffffff;
    for (int i = tid; i < iters; i += 256) {
        const int c = counts[(size_t)f * iters + i];
        if (c > bc) { bc = c; bi = i; }   // ascending i per thread: keeps the first maximum
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const int c = __shfl_xor_sync(0xffffffffu, bc, o), i = __shfl_xor_sync(0xffffffffu, bi, o);
        if (c > bc || (c == bc && i < bi)) { bc = c; bi = i; }
    }
    if (lane == 0) { sc[wid] = bc; si[wid] = bi; }
    __syncthreads();
    if (tid == 0) {
        for (int w = 1; w < 8; ++w)
            if (sc[w] > bc || (sc[w] == bc && si[w] < bi)) { bc = sc[w]; bi = si[w]; }
        const PlaneHyp<T> h = hyps[(size_t)f * iters + bi];
        if (out_eq) { out_eq[4 * f] = h.a; out_eq[4 * f + 1] = h.b; out_eq[4 * f + 2] = h.c; out_eq[4 * f + 3] = h.k; }
        if (out_best) out_best[f] = bi;
        if (out_count) out_count[f] = bc;
        if (out_normal) {
            T n0 = h.a, n1 = h.b, n2 = h.c;
            if (negate) { n0 = -n0; n1 = -n1; n2 = -n2; }   // fit_parallel returns -equation (plane.py:134)
            if (fabs(n2) > fabs(n1)) { const T t1 = n2, t2 = -n1; n1 = t1; n2 = t2; }              // back wall
            if (fabs(n0) > fabs(n1)) { const T t0 = -n2, t1 = n0, t2 = n1; n0 = t0; n1 = t1; n2 = t2; }   // side wall
            if (n1 < (T)0) { n0 = -n0; n1 = -n1; n2 = -n2; }
            out_normal[3 * f] = n0; out_normal[3 * f + 1] = n1; out_normal[3 * f + 2] = n2;
        }
    }
}

template <typename T>
__global__ void inlier_mask_kernel(const T* __restrict__ pts, const int32_t* __restrict__ n_pts, int64_t P_cap,
                                   const PlaneHyp<T>* __restrict__ hyps, const int32_t* __restrict__ best, int iters,
                                   uint8_t* __restrict__ mask) {
    const int f = blockIdx.y;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P_cap) return;
    uint8_t m = 0;
    if (i < n_pts[f]) {
        const PlaneHyp<T> p = hyps[(size_t)f * iters + best[f]];
        const T* X = pts + (size_t)f * 3 * P_cap;
        typedef Ar<T> A;
        const T s = A::add(A::add(A::add(A::mul(p.a, X[i]), A::mul(p.b, X[P_cap + i])), A::mul(p.c, X[2 * P_cap + i])), p.k);
        m = fabs(s) <= p.bound;
    }
    mask[(size_t)f * P_cap + i] = m;
}

template <typename T>
int ransac_impl(const T* pts, const int32_t* n_pts, int F, int64_t P_cap, const int32_t* ids, uint64_t seed,
                int64_t frame_id_base, int iters, double thresh, T* out_eq, int32_t* out_best, int32_t* out_count, T* out_normal,
                uint8_t* out_mask, void* ws, int64_t ws_bytes, cudaStream_t st, int negate) {
    const size_t hyp_bytes = ((size_t)F * iters * sizeof(PlaneHyp<T>) + 255) & ~size_t(255);
    const size_t need = hyp_bytes + (size_t)F * iters * 4 + (size_t)F * 4;
    if ((size_t)ws_bytes < need) return tdod_fail(TDOD_EINVAL, "tdod_ransac_plane: workspace too small");
    PlaneHyp<T>* hyps = reinterpret_cast<PlaneHyp<T>*>(ws);
    int32_t* counts = reinterpret_cast<int32_t*>((char*)ws + hyp_bytes);
    int32_t* best_tmp = counts + (size_t)F * iters;
    if (!out_best) out_best = best_tmp;
    hypotheses_kernel<T><<<dim3((iters + 127) / 128, F), 128, 0, st>>>(pts, n_pts, P_cap, ids, seed, frame_id_base, iters,
                                                                      (T)thresh, hyps, counts);
    TDOD_CHECK_LAUNCH();
    const bool exact_only = getenv("TDOD_RANSAC_EXACT") != nullptr;   // A/B switch (tests, profiling): unfiltered kernel
    if (exact_only) {
        constexpr int PT = 4;
        const size_t smem = (size_t)iters * (sizeof(PlaneHyp<T>) + sizeof(int));
        if (smem > 200 * 1024) return tdod_fail(TDOD_ERANGE, "tdod_ransac_plane: too many iterations for shared memory");
        static tdod_smem_cache sm_count;
        cudaError_t e = tdod_smem_opt_in((const void*)count_kernel<T, PT>, smem, &sm_count);
        if (e != cudaSuccess) return (int)e;
        const unsigned tiles = (unsigned)((P_cap + 256 * PT - 1) / (256 * PT));
        count_kernel<T, PT><<<dim3(tiles, F), 256, smem, st>>>(pts, n_pts, P_cap, hyps, iters, counts);
    } else {
        // few frames: smaller point tiles, so that the (tile, frame) grid still covers the device about twice over
        int tile_pts = kTilePts;
        while (tile_pts > 128 && (int64_t)F * ((P_cap + tile_pts - 1) / tile_pts) < 2 * 148 * 5) tile_pts >>= 1;
        const unsigned tiles = (unsigned)((P_cap + tile_pts - 1) / tile_pts);
        count_filter_kernel<T><<<dim3(tiles, F), kFiltT, 0, st>>>(pts, n_pts, P_cap, hyps, iters, counts, tile_pts);
    }
    TDOD_CHECK_LAUNCH();
    select_kernel<T><<<F, 256, 0, st>>>(hyps, counts, iters, negate, out_eq, out_best, out_count, out_normal);
    TDOD_CHECK_LAUNCH();
    if (out_mask) {
        inlier_mask_kernel<T><<<dim3((unsigned)((P_cap + 255) / 256), F), 256, 0, st>>>(pts, n_pts, P_cap, hyps, out_best,
                                                                                         iters, out_mask);
        TDOD_CHECK_LAUNCH();
    }
    return TDOD_OK;
}

}  // namespace

extern "C" int tdod_backproject(const float* depth, const uint8_t* ground, const float* K, int F, int H, int W,
                                int stride, int flavour, void* plane_pts, void* other_pts, int32_t* counts,
                                void* stream) {
    if (F == 0) return TDOD_OK;
    if (!depth || !K || !plane_pts || !counts || F < 0 || H <= 0 || W <= 0 || stride <= 0)
        return tdod_fail(TDOD_EINVAL, "tdod_backproject: bad argument");
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t P = (int64_t)((H + stride - 1) / stride) * ((W + stride - 1) / stride);
    const int iters = (int)((P + (int64_t)kBpCluster * kBpT - 1) / ((int64_t)kBpCluster * kBpT));
    const size_t smem = (size_t)2 * iters * (kBpT / 32) * sizeof(int);
    if (smem > 200 * 1024) return tdod_fail(TDOD_ERANGE, "tdod_backproject: frame too large for the per-slice count tables");
    const unsigned grid = (unsigned)F * kBpCluster;
    cudaError_t e;
    if (flavour == TDOD_BACKPROJECT_NUMPY) {
        static tdod_smem_cache sm_bp64;
        e = tdod_smem_opt_in((const void*)backproject_kernel<double, TDOD_BACKPROJECT_NUMPY>, smem, &sm_bp64);
        if (e != cudaSuccess) return (int)e;
        backproject_kernel<double, TDOD_BACKPROJECT_NUMPY><<<grid, kBpT, smem, st>>>(depth, ground, K, H, W, stride, iters,
                                                                                      (double*)plane_pts, (double*)other_pts, counts);
    } else if (flavour == TDOD_BACKPROJECT_TORCH) {
        static tdod_smem_cache sm_bp32;
        e = tdod_smem_opt_in((const void*)backproject_kernel<float, TDOD_BACKPROJECT_TORCH>, smem, &sm_bp32);
        if (e != cudaSuccess) return (int)e;
        backproject_kernel<float, TDOD_BACKPROJECT_TORCH><<<grid, kBpT, smem, st>>>(depth, ground, K, H, W, stride, iters,
                                                                                    (float*)plane_pts, (float*)other_pts, counts);
    } else {
        return tdod_fail(TDOD_EINVAL, "tdod_backproject: unknown flavour");
    }
    TDOD_CHECK_LAUNCH();
    return TDOD_OK;
}

extern "C" int64_t tdod_ransac_workspace_bytes(int F, int iters, int pts_f64) {
    const size_t hb = pts_f64 ? sizeof(PlaneHyp<double>) : sizeof(PlaneHyp<float>);
    return (int64_t)((((size_t)F * iters * hb + 255) & ~size_t(255)) + (size_t)F * iters * 4 + (size_t)F * 4);
}

extern "C" int tdod_ransac_plane(const void* pts, int pts_f64, const int32_t* n_pts, int F, int64_t P_cap,
                                 const int32_t* sample_ids, uint64_t seed, int64_t frame_id_base, int iters, double thresh,
                                 void* out_eq, int32_t* out_best_iter, int32_t* out_count, void* out_normal,
                                 uint8_t* out_inlier_mask, void* workspace, int64_t workspace_bytes, void* stream) {
    if (F == 0) return TDOD_OK;
    if (!pts || !n_pts || !workspace || F < 0 || P_cap <= 0 || iters <= 0)
        return tdod_fail(TDOD_EINVAL, "tdod_ransac_plane: bad argument");
    cudaStream_t st = (cudaStream_t)stream;
    if (pts_f64)
        return ransac_impl<double>((const double*)pts, n_pts, F, P_cap, sample_ids, seed, frame_id_base, iters, thresh, (double*)out_eq,
                                   out_best_iter, out_count, (double*)out_normal, out_inlier_mask, workspace,
                                   workspace_bytes, st, /*negate=*/0);
    return ransac_impl<float>((const float*)pts, n_pts, F, P_cap, sample_ids, seed, frame_id_base, iters, thresh, (float*)out_eq,
                              out_best_iter, out_count, (float*)out_normal, out_inlier_mask, workspace, workspace_bytes,
                              st, /*negate=*/1);
}

// ------------------------------------------------------------------------------------------------
// tdod_ransac_plane_static: tdod_ransac_plane without a caller-provided workspace and without any allocation.
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The scratch (hypotheses, inlier
// counts, best-iteration temporaries) lives in ONE file-scope device array of fixed size that every GPU loading the
// library holds until exit.  All calls on a device share it: calls that overlap on the GPU (different streams) race,
// and the array keeps the last call's contents -- the only part that must start at zero, the inlier counts, is cleared
// by hypotheses_kernel (counts[f, it] = 0) at the start of every call, on the call's stream.
// ------------------------------------------------------------------------------------------------
constexpr size_t kStaticHyps = TDOD_RANSAC_STATIC_MAX_HYPS;       // F * iters
constexpr size_t kStaticFrames = TDOD_RANSAC_STATIC_MAX_FRAMES;   // F
constexpr size_t kStaticBytes = ((kStaticHyps * sizeof(PlaneHyp<double>) + 255) & ~size_t(255)) + kStaticHyps * 4 + kStaticFrames * 4;
static __device__ __align__(256) unsigned char g_ransac_scratch[kStaticBytes];   // 2 887 680 B

// In host code `g_ransac_scratch` is a host-side stand-in for the symbol, NOT a device address: the address comes from
// cudaGetSymbolAddress, once per device with that device current, and is what the kernels receive.
static void* g_ransac_scratch_ptr[64] = {nullptr};

extern "C" int tdod_ransac_plane_static(const void* pts, int pts_f64, const int32_t* n_pts, int F, int64_t P_cap,
                                        const int32_t* sample_ids, uint64_t seed, int64_t frame_id_base, int iters,
                                        double thresh, void* out_eq, int32_t* out_best_iter, int32_t* out_count,
                                        void* out_normal, uint8_t* out_inlier_mask, void* stream) {
    if (F == 0) return TDOD_OK;
    if (!pts || !n_pts || F < 0 || P_cap <= 0 || iters <= 0)
        return tdod_fail(TDOD_EINVAL, "tdod_ransac_plane_static: bad argument");
    if ((size_t)F > kStaticFrames || (size_t)F * (size_t)iters > kStaticHyps)
        return tdod_fail(TDOD_ERANGE, "tdod_ransac_plane_static: F > 1024 or F * iters > 65536 exceeds the static scratch; "
                                      "use tdod_ransac_plane with a caller-provided workspace");
    if ((size_t)tdod_ransac_workspace_bytes(F, iters, pts_f64) > kStaticBytes)
        return tdod_fail(TDOD_ERANGE, "tdod_ransac_plane_static: workspace larger than the static scratch");
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= 64) return tdod_fail(TDOD_ERANGE, "tdod_ransac_plane_static: device index beyond 63");
    void* ws = g_ransac_scratch_ptr[dev];
    if (!ws) {                                          // set-up: once per device, that device being current
        e = cudaGetSymbolAddress(&ws, g_ransac_scratch);
        if (e != cudaSuccess) return (int)e;
        if (!ws) return tdod_fail(TDOD_EINVAL, "tdod_ransac_plane_static: cudaGetSymbolAddress returned NULL");
        g_ransac_scratch_ptr[dev] = ws;                 // (racing first calls store the same value)
    }
    cudaStream_t st = (cudaStream_t)stream;
    if (pts_f64)
        return ransac_impl<double>((const double*)pts, n_pts, F, P_cap, sample_ids, seed, frame_id_base, iters, thresh, (double*)out_eq,
                                   out_best_iter, out_count, (double*)out_normal, out_inlier_mask, ws, (int64_t)kStaticBytes, st,
                                   /*negate=*/0);
    return ransac_impl<float>((const float*)pts, n_pts, F, P_cap, sample_ids, seed, frame_id_base, iters, thresh, (float*)out_eq,
                              out_best_iter, out_count, (float*)out_normal, out_inlier_mask, ws, (int64_t)kStaticBytes, st,
                              /*negate=*/1);
}
