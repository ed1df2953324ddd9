// Device-side building blocks shared by every kernel of libtdod (sm_100a).
// Each function states the reference arithmetic it restates (file:line under /root/reference).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math_constants.h>

#define TDOD_NYAW 36
#define TDOD_MIN_PROP_S 0.05f

namespace tdod {

// ------------------------------------------------------------------------------------------
// warp / block reductions
// ------------------------------------------------------------------------------------------
// NaN-propagating min / max (torch.min / torch.max / torch.clamp semantics) in one instruction each: FMNMX.NAN
__device__ __forceinline__ float min_nan(float a, float b) {
    float r;
    asm("min.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ float max_nan(float a, float b) {
    float r;
    asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ float warp_max_nan(float v) {
    // torch.max semantics: NaN propagates
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = max_nan(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// (value, index) argmax with np.argmax semantics: the first NaN wins, else the first maximum
// (lowest index among equals).  `isn` = value is NaN.
struct ArgMax {
    float v;
    int i;
};
__device__ __forceinline__ bool argmax_better(float va, int ia, float vb, int ib) {
    // returns true when (vb, ib) should replace (va, ia)
    const bool na = va != va, nb = vb != vb;
    if (na || nb) {
        if (na && nb) return ib < ia;
        return nb;
    }
    if (vb > va) return true;
    if (vb == va) return ib < ia;
    return false;
}
__device__ __forceinline__ ArgMax warp_argmax(ArgMax a) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        float v = __shfl_xor_sync(0xffffffffu, a.v, o);
        int i = __shfl_xor_sync(0xffffffffu, a.i, o);
        if (argmax_better(a.v, a.i, v, i)) { a.v = v; a.i = i; }
    }
    return a;
}

// ------------------------------------------------------------------------------------------
// float <-> radix-orderable key (for the exact lower median, torch.median)
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t f2key(float f) {
    uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key2f(uint32_t k) {
    uint32_t u = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
    return __uint_as_float(u);
}

// ------------------------------------------------------------------------------------------
// sampling helpers
// ------------------------------------------------------------------------------------------
// ProposalNetwork/utils/utils.py:170-177 vectorized_linspace element i: arange(n)*step + lo with a
// separately rounded multiply and add (two torch ops), then .long() truncation toward zero
// (proposals.py:360-361).
__device__ __forceinline__ int linspace_px(float lo, float step, int i) {
    float v = __fadd_rn(__fmul_rn((float)i, step), lo);
    return (int)v;  // truncation toward zero == .long()
}

// proposals.py:366-375: pixel (gx, gy) with depth d -> metric seeds x, y, z_tmp.  Every product /
// sum is a separate torch op in the reference => explicit _rn intrinsics (no FMA contraction).
__device__ __forceinline__ void centre_seed(int gx, int gy, float d, float fx, float cx, float cy,
                                            float& x, float& y, float& zt) {
    const float ox = __fsub_rn((float)gx, cx);
    const float oy = __fsub_rn((float)gy, cy);
    const float ang_x = atan2f(ox, fx);
    const float r = __fsqrt_rn(__fadd_rn(__fmul_rn(ox, ox), __fmul_rn(fx, fx)));
    const float ang_d = atan2f(oy, r);
    y = __fmul_rn(d, sinf(ang_d));
    const float dx = __fsqrt_rn(__fsub_rn(__fmul_rn(d, d), __fmul_rn(y, y)));
    x = __fmul_rn(dx, sinf(ang_x));
    zt = __fsqrt_rn(__fsub_rn(__fmul_rn(dx, dx), __fmul_rn(x, x)));
}

// utils.py:45 torch.normal(mean, std) == z*std + mean, multiply and add rounded separately.
__device__ __forceinline__ float scale_normal(float z, float mean, float std) {
    return __fadd_rn(__fmul_rn(z, std), mean);
}

// Philox4x32-10 (counter-based; Salmon et al. 2011) -- the throughput-mode generator.
struct Philox {
    uint32_t k0, k1;
    __device__ __forceinline__ Philox(uint64_t seed) : k0((uint32_t)seed), k1((uint32_t)(seed >> 32)) {}
    __device__ __forceinline__ uint4 operator()(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3) const {
        uint32_t a = k0, b = k1;
#ifndef TDOD_PHILOX_UNROLL
#define TDOD_PHILOX_UNROLL 10
#endif
        constexpr int kUnroll = TDOD_PHILOX_UNROLL;      // build knob: rolled rounds shrink the fused kernel's code
#pragma unroll kUnroll
        for (int r = 0; r < 10; ++r) {
            const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
            const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
            const uint32_t n0 = hi1 ^ c1 ^ a, n2 = hi0 ^ c3 ^ b;
            c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
            a += 0x9E3779B9u; b += 0xBB67AE85u;
        }
        return make_uint4(c0, c1, c2, c3);
    }
};
// two uniforms in (0,1] -> two standard normals (Box-Muller)
__device__ __forceinline__ float2 box_muller(uint32_t a, uint32_t b) {
    const float u1 = ((float)a + 1.0f) * 2.3283064365386963e-10f;  // (0,1]
    const float u2 = (float)b * 2.3283064365386963e-10f;           // [0,1)
    float rad;                                                     // throughput-mode draws: approximate sqrt is enough
    asm("sqrt.approx.f32 %0, %1;" : "=f"(rad) : "f"(-2.0f * __logf(u1)));
    float s, c;
    __sincosf(6.283185307179586f * u2, &s, &c);
    return make_float2(rad * c, rad * s);
}

// ------------------------------------------------------------------------------------------
// geometry
// ------------------------------------------------------------------------------------------
// cubercnn/util/math_util.py:142-245 via ProposalNetwork/utils/spaces.py:192-204: corner j of
// (c, w,h,l, R row-major): local X = +-l/2 (neg for j in {0,3,4,7}), Y = +-h/2 (neg {0,1,4,5}),
// Z = +-w/2 (neg {0,1,2,3}); world = R @ local + c.
__device__ __forceinline__ void cube_corners(const float* __restrict__ q, float (&X)[8], float (&Y)[8],
                                             float (&Z)[8]) {
    const float hl = q[5] * 0.5f, hh = q[4] * 0.5f, hw = q[3] * 0.5f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const float sx = ((j & 3) == 0 || (j & 3) == 3) ? -hl : hl;   // j in {0,3,4,7} negative
        const float sy = ((j & 3) < 2) ? -hh : hh;                   // j in {0,1,4,5} negative
        const float sz = (j < 4) ? -hw : hw;                         // j in {0,1,2,3} negative
        X[j] = fmaf(q[8], sz, fmaf(q[7], sy, q[6] * sx)) + q[0];
        Y[j] = fmaf(q[11], sz, fmaf(q[10], sy, q[9] * sx)) + q[1];
        Z[j] = fmaf(q[14], sz, fmaf(q[13], sy, q[12] * sx)) + q[2];
    }
}

// ProposalNetwork/utils/spaces.py:224-245 get_bube_corners: p = K @ X; (px/pz, py/pz) with no z>0
// guard; clamp x to [int(-W/2+1), int(2W-1)], y to [int(-H/2+1), int(2H-1)].  torch.clamp
// propagates NaN (0/0), which fminf/fmaxf would drop, hence the explicit NaN test.
struct Clamp {
    float xlo, xhi, ylo, yhi;
};
__host__ __device__ __forceinline__ Clamp make_clamp(int W, int H) {
    Clamp c;
    c.xlo = (float)(int)(-(double)W / 2 + 1);
    c.xhi = (float)(W - 1 + W);
    c.ylo = (float)(int)(-(double)H / 2 + 1);
    c.yhi = (float)(H - 1 + H);
    return c;
}
__device__ __forceinline__ float clamp_nan(float v, float lo, float hi) {   // lo, hi are never NaN
    return min_nan(max_nan(v, lo), hi);
}
__device__ __forceinline__ void project(const float* __restrict__ K, const Clamp& cl, float X, float Y, float Z,
                                        float& u, float& v) {
    const float px = fmaf(K[2], Z, fmaf(K[1], Y, K[0] * X));
    const float py = fmaf(K[5], Z, fmaf(K[4], Y, K[3] * X));
    const float pz = fmaf(K[8], Z, fmaf(K[7], Y, K[6] * X));
    u = clamp_nan(__fdiv_rn(px, pz), cl.xlo, cl.xhi);
    v = clamp_nan(__fdiv_rn(py, pz), cl.ylo, cl.yhi);
}

// (torch.min / torch.max over the 8 corners propagate NaN, conversions.py:38-41: min_nan / max_nan above)

// detectron2 pairwise_iou (public formula; call site ProposalNetwork/utils/utils.py:186-192)
__device__ __forceinline__ float iou_xyxy(float ax0, float ay0, float ax1, float ay1, float bx0, float by0,
                                          float bx1, float by1) {
    const float a1 = __fmul_rn(ax1 - ax0, ay1 - ay0);
    const float a2 = __fmul_rn(bx1 - bx0, by1 - by0);
    float w = fminf(ax1, bx1) - fmaxf(ax0, bx0);
    float h = fminf(ay1, by1) - fmaxf(ay0, by0);
    // clamp_(min=0) keeps NaN; a NaN extent gives inter NaN -> (inter > 0) false -> 0
    w = (w != w) ? w : fmaxf(w, 0.f);
    h = (h != h) ? h : fmaxf(h, 0.f);
    const float inter = __fmul_rn(w, h);
    return (inter > 0.f) ? __fdiv_rn(inter, __fsub_rn(__fadd_rn(a1, a2), inter)) : 0.f;
}

// ProposalNetwork/scoring/scorefunction.py:151-152: mean_3(exp(-1/2 ((d-mu)/sigma)^2))
__device__ __forceinline__ float dims_gauss(float w, float h, float l, const float* mu, const float* sg) {
    const float a = __fdiv_rn(w - mu[0], sg[0]), b = __fdiv_rn(h - mu[1], sg[1]), c = __fdiv_rn(l - mu[2], sg[2]);
    const float ea = expf(-0.5f * __fmul_rn(a, a)), eb = expf(-0.5f * __fmul_rn(b, b)),
                ec = expf(-0.5f * __fmul_rn(c, c));
    return __fdiv_rn(__fadd_rn(__fadd_rn(ea, eb), ec), 3.0f);
}

// scorefunction.py:51-56 modified_chamfer_distance(rect, corners): mean over the 4 rectangle points of the
// distance to the nearest of the 8 projected corners, evaluated in fp64 like cKDTree, then stored to fp32
// (scores tensor is fp32, scorefunction.py:78-81).  The nearest corner is searched on fp32 squared distances
// (a near-tie can pick a corner whose true distance is larger by < 1e-7 relative -- far inside the 1e-5 bar);
// the four distances that enter the mean are then evaluated in fp64.
__device__ __forceinline__ float chamfer4x8(const float* __restrict__ rect /*[4][2]*/, const float (&U)[8],
                                            const float (&V)[8]) {
    bool any_nan = false;
#pragma unroll
    for (int j = 0; j < 8; ++j) any_nan |= (U[j] != U[j]) | (V[j] != V[j]);
    double acc = 0.0;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const float rx = rect[2 * r], ry = rect[2 * r + 1];
        float best = CUDART_INF_F, bu = U[0], bv = V[0];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float dx = rx - U[j], dy = ry - V[j];
            const float d2 = fmaf(dx, dx, dy * dy);
            const bool lt = d2 < best;
            best = lt ? d2 : best; bu = lt ? U[j] : bu; bv = lt ? V[j] : bv;
        }
        const double ex = (double)rx - (double)bu, ey = (double)ry - (double)bv;
        acc += sqrt(ex * ex + ey * ey);
    }
    // cKDTree rejects NaN input in the reference (it raises); we propagate NaN instead
    return any_nan ? CUDART_NAN_F : (float)(acc * 0.25);
}

}  // namespace tdod

// ------------------------------------------------------------------------------------------
// torch.std(x, dim=1) of a contiguous fp32 [N, n] tensor AS TORCH'S CUDA KERNEL ROUNDS IT
// ------------------------------------------------------------------------------------------
// On CUDA tensors -- how the reference runs in production (proposals.py:386-396 on reference_box.device) -- torch.std
// is a Welford reduction in fp32 (at::acc_type<float, true> is float) whose rounding depends on the launch
// configuration: which elements each thread folds in, in which order the partial results are merged.  To reproduce the
// proposal set of that flavour bit for bit the same tree is walked here.  Restated from the installed torch 2.11
// headers: ATen/native/cuda/Reduce.cuh (setReduceConfig :1033-1179, set_block_dimension :100-108,
// input_vectorized_thread_reduce_impl :500-557, thread_reduce_impl :559-622, block_x_reduce :624-663, block_y_reduce
// :665-682), ATen/native/SharedReduceOps.h WelfordOps :99-162, and std_var_kernel_impl (vt0 = 2, hence two-element
// vector loads and two accumulators per thread).  Checked against torch.std on the GPU (tests/test_gpu_parity.py).
namespace tstd {

struct W { float mean, m2, nf; };

__device__ __forceinline__ W reduce(W a, float x) {           // WelfordOps::reduce; nvcc contracts m2 + delta * new_delta
    const float nf = a.nf + 1.f;                              // counts stay far below 2^24: float(n + 1) == nf + 1
    const float delta = __fsub_rn(x, a.mean);
    const float mean = __fadd_rn(a.mean, __fdiv_rn(delta, nf));
    const float nd = __fsub_rn(x, mean);
    W r; r.mean = mean; r.m2 = __fmaf_rn(delta, nd, a.m2); r.nf = nf;
    return r;
}
__device__ __forceinline__ W combine(W a, W b) {              // WelfordOps::combine
    if (a.nf == 0.f) return b;
    if (b.nf == 0.f) return a;
    const float delta = __fsub_rn(b.mean, a.mean);
    const float cnt = __fadd_rn(a.nf, b.nf);
    const float r = __fdiv_rn(b.nf, cnt);
    W o;
    o.mean = __fmaf_rn(delta, r, a.mean);
    o.m2 = __fmaf_rn(__fmul_rn(__fmul_rn(delta, delta), a.nf), r, __fadd_rn(a.m2, b.m2));
    o.nf = cnt;
    return o;
}

struct Cfg { int vec, bw, Y, im1, step, ok; };
__host__ __device__ inline int last_pow2(int x) { int p = 1; while (2 * p <= x) p *= 2; return p; }
// launch shape torch picks for an [N, n] fp32 tensor reduced over its contiguous inner dimension
__host__ __device__ inline Cfg config(int n, int N) {
    Cfg c;
    c.vec = n >= 128;
    const int dim0 = c.vec ? n / 2 : n;
    const int d0 = dim0 < 512 ? last_pow2(dim0 < 1 ? 1 : dim0) : 512, d1 = N < 512 ? last_pow2(N < 1 ? 1 : N) : 512;
    int bw = d0 < 32 ? d0 : 32;
    const int bh = d1 < 512 / bw ? d1 : 512 / bw;
    bw = d0 < 512 / bh ? d0 : 512 / bh;
    c.bw = bw; c.step = bw;
    const int vpt = (n + bw - 1) / bw;
    const int thr = bh * 16 < 256 ? bh * 16 : 256;
    c.Y = 1; c.im1 = 0;
    if (vpt >= thr) { c.im1 = c.step; c.step *= bh; c.Y = bh; }
    // a reduction split over several CTAs (values_per_thread >= 256 after the warp split) is not restated
    c.ok = !(c.Y > 1 && (n + c.step - 1) / c.step >= 256);
    return c;
}

// All `nthreads` threads of the CTA call this.  `at(i)` returns element i of the row; `shift` is the row's misalignment
// in elements with respect to torch's 8-byte vector loads (row_index * n odd -> 1).  scratch: bw * Y entries.
template <typename F>
__device__ float block_std(F at, int n, int shift, const Cfg c, W* scratch, int tid, int nthreads) {
    const int nv = c.bw * c.Y;
    for (int v = tid; v < nv; v += nthreads) {
        const int x = v % c.bw, y = v / c.bw;
        const bool tail_ok = (c.Y == 1) || y == 0;             // should_reduce_tail()
        W a0 = {0.f, 0.f, 0.f}, a1 = {0.f, 0.f, 0.f};
        if (c.vec) {
            int end = n, base = 0;
            if (shift > 0) {                                   // unaligned head: torch steps back `shift` elements
                end += shift;
                if (x >= shift && x < 2 && tail_ok) a0 = reduce(a0, at(x - shift));
                end -= 2; base = 2 - shift;
            }
            int idx = x + y * c.im1;
            while (idx * 2 + 1 < end) {
                a0 = reduce(a0, at(base + 2 * idx)); a1 = reduce(a1, at(base + 2 * idx + 1));
                idx += c.step;
            }
            const int tail_start = end - end % 2;
            if (tail_ok) { const int i = tail_start + x; if (i < end) a0 = reduce(a0, at(base + i)); }
        } else {
            int idx = x + y * c.im1;
            while (idx + c.step < n) { a0 = reduce(a0, at(idx)); a1 = reduce(a1, at(idx + c.step)); idx += 2 * c.step; }
            if (idx < n) { a0 = reduce(a0, at(idx)); idx += c.step; if (idx < n) a1 = reduce(a1, at(idx)); }
        }
        scratch[v] = combine(a0, a1);
    }
    __syncthreads();
    for (int off = c.bw >> 1; off > 0; off >>= 1) {            // block_x_reduce: shared-memory levels, then the shuffle levels
        for (int v = tid; v < nv; v += nthreads) {
            const int x = v % c.bw;
            if (x < off) scratch[v] = combine(scratch[v], scratch[v + off]);
        }
        __syncthreads();
    }
    for (int off = c.Y >> 1; off > 0; off >>= 1) {             // block_y_reduce
        for (int y = tid; y < off; y += nthreads) scratch[y * c.bw] = combine(scratch[y * c.bw], scratch[(y + off) * c.bw]);
        __syncthreads();
    }
    const W r = scratch[0];
    const float div = r.nf > 1.f ? __fsub_rn(r.nf, 1.f) : 0.f; // correction = 1
    const float sd = __fsqrt_rn(__fdiv_rn(r.m2, div));
    __syncthreads();
    return sd;
}

}  // namespace tstd
