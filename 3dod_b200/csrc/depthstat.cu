// Depth order statistics of ROIHeads3DScore's pseudo-ground-truth losses (the losses that read the
// depth map instead of the cube geometry):
//   mode 0  pseudo_gt_z_box_loss   (cubercnn/modeling/roi_heads/roi_heads.py:1196-1231)  median of the depth crop under
//           the clipped 2D box; boxes with no area fall back to the depth at the clamped box centre
//   mode 1  pseudo_gt_z_point_loss (:1256-1277)  depth at the clamped predicted centre
//   mode 2  depth_range_loss       (:1279-1305)  quantile(0.9) - quantile(0.1) of the depth under the segmentation mask,
//           of the depth crop under the 2D box when the mask is empty
// The reference slices one crop per box in a Python loop and sorts it (torch.median / torch.quantile); here one CTA per
// box selects the wanted ranks with an 8-bit-per-pass radix select over the crop as it lies in the depth map (no copy, no
// sort).  Values are selected, never averaged, so the median is bit-exact; the quantile's interpolation follows
// at::lerp as the CUDA build of torch evaluates it.
#include <cstdint>

#include "tdod_device.cuh"
#include "tdod_internal.h"

using namespace tdod;

namespace {

constexpr int kSelT = 1024;

struct Region {                 // pixels (y0 + r, x0 + c), r < rows, c < cols, of one H x W frame; m == nullptr: all of them
    const float* d;
    const uint8_t* m;
    int x0, y0, cols, rows, W;
};

__device__ __forceinline__ unsigned key_of(float v) {          // order-preserving map float -> unsigned
    const unsigned b = __float_as_uint(v);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float val_of(unsigned k) {
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

struct SelShared {
    unsigned hist[514];         // two histograms of 257: bin 256 of the second swallows the lanes that have no pixel
    unsigned sel[4];
    unsigned cnt[5];            // pixels in the region, NaNs among them; later: values <= v, smallest key above v (x2)
    int box[4];                 // bounding box of the mask pixels
};

// Calls fn(valid, value, r, c) for every pixel of the region, warp-uniformly (all 32 lanes enter every call, so fn may
// use full-mask warp intrinsics): a warp per row, 128 columns per step with the four loads of a lane issued together.
template <bool kDepth = true, class Fn>
__device__ __forceinline__ void region_scan(const Region& R, Fn fn) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int r = warp; r < R.rows; r += kSelT / 32) {
        const size_t row = (size_t)(R.y0 + r) * R.W + R.x0;
        for (int c0 = 0; c0 < R.cols; c0 += 128) {
            float v[4];
            bool ok[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int c = c0 + u * 32 + lane;
                const bool in = c < R.cols;
                v[u] = (kDepth && in) ? R.d[row + c] : 0.0f;
                ok[u] = in && (!R.m || R.m[row + c]);
            }
#pragma unroll
            for (int u = 0; u < 4; u++) fn(ok[u], v[u], r, c0 + u * 32 + lane);
        }
    }
}

// pixels in the region and how many of them are NaN; with a mask only the mask is read (n_nan = 0, see region_next2) and
// the region shrinks to the bounding box of its pixels.  Every thread returns the same pair.
__device__ void region_count(Region& R, SelShared& s, unsigned& n, unsigned& n_nan) {
    if (threadIdx.x < 2) s.cnt[threadIdx.x] = 0;
    if (threadIdx.x == 0) { s.box[0] = s.box[1] = 0x7fffffff; s.box[2] = s.box[3] = -1; }
    __syncthreads();
    unsigned c = 0, cn = 0;
    int bx0 = 0x7fffffff, by0 = 0x7fffffff, bx1 = -1, by1 = -1;
    if (R.m) {                                      // the mask bytes alone: region_next2 reports the NaNs under the mask
        region_scan<false>(R, [&](bool ok, float, int r, int col) {
            if (ok) { c++; bx0 = min(bx0, col); bx1 = max(bx1, col); by0 = min(by0, r); by1 = max(by1, r); }
        });
    } else {
        region_scan(R, [&](bool ok, float v, int, int) {
            if (ok) { c++; cn += (v != v); }
        });
    }
    c = __reduce_add_sync(0xffffffffu, c);
    cn = __reduce_add_sync(0xffffffffu, cn);
    if (R.m) {
        bx0 = __reduce_min_sync(0xffffffffu, bx0); by0 = __reduce_min_sync(0xffffffffu, by0);
        bx1 = __reduce_max_sync(0xffffffffu, bx1); by1 = __reduce_max_sync(0xffffffffu, by1);
    }
    if ((threadIdx.x & 31) == 0 && c) {
        atomicAdd(&s.cnt[0], c); atomicAdd(&s.cnt[1], cn);
        if (R.m) { atomicMin(&s.box[0], bx0); atomicMin(&s.box[1], by0); atomicMax(&s.box[2], bx1); atomicMax(&s.box[3], by1); }
    }
    __syncthreads();
    n = s.cnt[0];
    n_nan = s.cnt[1];
    if (R.m && n) {
        R.x0 += s.box[0]; R.y0 += s.box[1];
        R.cols = s.box[2] - s.box[0] + 1; R.rows = s.box[3] - s.box[1] + 1;
    }
    __syncthreads();
}

// k-th smallest (0-based) of the region's values: four passes, most significant byte first
__device__ float region_select(const Region& R, unsigned k, SelShared& s) {
    unsigned prefix = 0, known = 0;
    for (int shift = 24; shift >= 0; shift -= 8) {
        for (int i = threadIdx.x; i < 257; i += kSelT) s.hist[i] = 0;
        __syncthreads();
        region_scan(R, [&](bool ok, float v, int, int) {
            const unsigned key = key_of(v);
            const unsigned bin = (ok && (key & known) == prefix) ? ((key >> shift) & 255u) : 256u;
            // neighbouring depths share their leading bytes: one shared-memory add per distinct bin of the warp
            const unsigned peers = __match_any_sync(0xffffffffu, bin);
            if (bin < 256 && (threadIdx.x & 31) == (unsigned)(__ffs(peers) - 1)) atomicAdd(&s.hist[bin], __popc(peers));
        });
        __syncthreads();
        if (threadIdx.x < 32) {                     // first bin whose running count passes k
            const int lane = threadIdx.x;
            unsigned mine = 0;
            for (int j = 0; j < 8; j++) mine += s.hist[lane * 8 + j];
            unsigned inc = mine;
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned a = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += a;
            }
            const unsigned hit = __ballot_sync(0xffffffffu, inc > k);
            const int owner = __ffs(hit) - 1;       // never -1: k < number of values with this prefix
            if (lane == owner) {
                unsigned acc = inc - mine;
                int b = lane * 8;
                while (acc + s.hist[b] <= k) acc += s.hist[b++];
                s.sel[0] = (unsigned)b;
                s.sel[1] = k - acc;
            }
        }
        __syncthreads();
        prefix |= s.sel[0] << shift;
        known |= 0xffu << shift;
        k = s.sel[1];
        __syncthreads();
    }
    return val_of(prefix);
}

// at::lerp (aten/src/ATen/native/Lerp.h) the way nvcc contracts it in torch's CUDA build
__device__ __forceinline__ float torch_lerp(float a, float b, float w) {
    const float diff = __fsub_rn(b, a);
    return (fabsf(w) < 0.5f) ? __fmaf_rn(w, diff, a) : __fmaf_rn(-diff, __fsub_rn(1.0f, w), b);
}

// one warp: the first bin of h[0..255] whose running count passes k -> sel[0], and k's rank inside that bin -> sel[1]
__device__ __forceinline__ void find_bin(const unsigned* h, unsigned k, unsigned* sel) {
    const int lane = threadIdx.x & 31;
    unsigned mine = 0;
    for (int j = 0; j < 8; j++) mine += h[lane * 8 + j];
    unsigned inc = mine;
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned a = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += a;
    }
    const int owner = __ffs(__ballot_sync(0xffffffffu, inc > k)) - 1;
    if (lane == owner) {
        unsigned acc = inc - mine;
        int b = lane * 8;
        while (acc + h[b] <= k) acc += h[b++];
        sel[0] = (unsigned)b;
        sel[1] = k - acc;
    }
}

// the order statistics of ranks ka and kb in the same four passes: a pixel counts in a's histogram, in b's, or, while the
// two prefixes still coincide, in the one they share
__device__ void region_select2(const Region& R, unsigned ka, unsigned kb, SelShared& s, float& va, float& vb) {
    unsigned pa = 0, pb = 0, known = 0;
    for (int shift = 24; shift >= 0; shift -= 8) {
        for (int i = threadIdx.x; i < 514; i += kSelT) s.hist[i] = 0;
        __syncthreads();
        region_scan(R, [&](bool ok, float v, int, int) {
            const unsigned key = key_of(v), top = key & known, dig = (key >> shift) & 255u;
            const unsigned bin = !ok ? 513u : (top == pa) ? dig : (top == pb) ? 257u + dig : 513u;
            const unsigned peers = __match_any_sync(0xffffffffu, bin);
            if (bin < 513 && (threadIdx.x & 31) == (unsigned)(__ffs(peers) - 1)) atomicAdd(&s.hist[bin], __popc(peers));
        });
        __syncthreads();
        if (threadIdx.x < 32) find_bin(s.hist, ka, s.sel);
        else if (threadIdx.x < 64) find_bin(pa == pb ? s.hist : s.hist + 257, kb, s.sel + 2);
        __syncthreads();
        pa |= s.sel[0] << shift; ka = s.sel[1];
        pb |= s.sel[2] << shift; kb = s.sel[3];
        known |= 0xffu << shift;
        __syncthreads();
    }
    va = val_of(pa);
    vb = val_of(pb);
}

// the order statistics of ranks ka + 1 and kb + 1, given va and vb of ranks ka and kb, in one pass: v again when more than
// k + 1 values are <= v, else the smallest value above v
__device__ unsigned region_next2(const Region& R, unsigned ka, float va, unsigned kb, float vb, SelShared& s, float& na, float& nb) {
    if (threadIdx.x < 5) s.cnt[threadIdx.x] = (threadIdx.x & 1) ? 0xffffffffu : 0u;
    __syncthreads();
    const unsigned kva = key_of(va), kvb = key_of(vb);
    unsigned lea = 0, upa = 0xffffffffu, leb = 0, upb = 0xffffffffu, cn = 0;
    region_scan(R, [&](bool ok, float x, int, int) {
        const unsigned key = key_of(x);
        if (ok) {
            cn += (x != x);
            if (key <= kva) lea++; else upa = min(upa, key);
            if (key <= kvb) leb++; else upb = min(upb, key);
        }
    });
    lea = __reduce_add_sync(0xffffffffu, lea); upa = __reduce_min_sync(0xffffffffu, upa);
    leb = __reduce_add_sync(0xffffffffu, leb); upb = __reduce_min_sync(0xffffffffu, upb);
    cn = __reduce_add_sync(0xffffffffu, cn);
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&s.cnt[0], lea); atomicMin(&s.cnt[1], upa);
        atomicAdd(&s.cnt[2], leb); atomicMin(&s.cnt[3], upb);
        if (cn) atomicAdd(&s.cnt[4], cn);
    }
    __syncthreads();
    na = (s.cnt[0] > ka + 1) ? va : val_of(s.cnt[1]);
    nb = (s.cnt[2] > kb + 1) ? vb : val_of(s.cnt[3]);
    const unsigned n_nan = s.cnt[4];
    __syncthreads();
    return n_nan;
}

// torch.quantile(x, q) with linear interpolation (aten/src/ATen/native/Sorting.cpp, quantile_compute): rank = q * (n - 1)
// in fp32, the two neighbouring order statistics, lerp by the fractional rank.  Returns quantile(qb) - quantile(qa), qa < qb.
__device__ float region_quantile_range(const Region& R, unsigned n, float qa, float qb, SelShared& s) {
    const float ra = __fmul_rn(qa, (float)(n - 1)), rb = __fmul_rn(qb, (float)(n - 1));
    const float la = floorf(ra), lb = floorf(rb);
    float va, vb, na, nb;
    region_select2(R, (unsigned)la, (unsigned)lb, s, va, vb);
    // always run: the pass also reports the NaNs of the region, which make torch.quantile NaN
    if (region_next2(R, (unsigned)la, va, (unsigned)lb, vb, s, na, nb)) return __uint_as_float(0x7fc00000u);
    if (ceilf(ra) == la) na = va;          // an integral rank interpolates to itself (its weight is 0; keeps NaN-free)
    if (ceilf(rb) == lb) nb = vb;
    return __fsub_rn(torch_lerp(vb, nb, __fsub_rn(rb, lb)), torch_lerp(va, na, __fsub_rn(ra, la)));
}

__device__ __forceinline__ float clampf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }

// depth at (clamp(x, 10, W-11), clamp(y, 10, H-11)), truncated: roi_heads.py:1224-1227 and :1272-1274
__device__ __forceinline__ float point_depth(const float* d, float x, float y, int H, int W) {
    const int xi = (int)clampf(x, 10.0f, (float)(W - 11)), yi = (int)clampf(y, 10.0f, (float)(H - 11));
    return d[(size_t)min(max(yi, 0), H - 1) * W + min(max(xi, 0), W - 1)];
}

__global__ void __launch_bounds__(kSelT)
depth_select_kernel(const float* __restrict__ depth, const int32_t* __restrict__ frame, const float* __restrict__ boxes,
                    const uint8_t* __restrict__ masks, const int32_t* __restrict__ mask_idx, int mode, int H, int W,
                    float* __restrict__ out, int32_t* __restrict__ flag) {
    __shared__ SelShared s;
    const int item = blockIdx.x;
    const float* d = depth + (size_t)frame[item] * H * W;
    const float bx1 = boxes[item * 4 + 0], by1 = boxes[item * 4 + 1], bx2 = boxes[item * 4 + 2], by2 = boxes[item * 4 + 3];
    const float nan = __uint_as_float(0x7fc00000u);
    float res;
    int fl = 0;
    if (mode == 1) {
        res = point_depth(d, bx1, by1, H, W);
    } else if (mode == 0) {
        // Boxes.clip((h, w)) then area() > 0 (:1209-1212)
        const float x1 = clampf(bx1, 0.0f, (float)W), y1 = clampf(by1, 0.0f, (float)H);
        const float x2 = clampf(bx2, 0.0f, (float)W), y2 = clampf(by2, 0.0f, (float)H);
        if (__fmul_rn(__fsub_rn(x2, x1), __fsub_rn(y2, y1)) > 0.0f) {
            fl = 1;
            Region R{d, nullptr, (int)x1, (int)y1, (int)x2 - (int)x1, (int)y2 - (int)y1, W};
            unsigned n, n_nan;
            region_count(R, s, n, n_nan);
            // an empty slice raises in the reference; a NaN anywhere makes torch.median NaN
            res = (n == 0 || n_nan) ? nan : region_select(R, (n - 1) >> 1, s);
        } else {
            res = point_depth(d, __fmul_rn(__fadd_rn(x1, x2), 0.5f), __fmul_rn(__fadd_rn(y1, y2), 0.5f), H, W);
        }
    } else {
        Region R{d, masks + (size_t)mask_idx[item] * H * W, 0, 0, W, H, W};
        unsigned n, n_nan;
        region_count(R, s, n, n_nan);
        fl = 1;
        if (n == 0) {                               // no mask pixel: the slice under the 2D box (:1296-1297)
            fl = 0;
            const int x1 = min(max((int)bx1, 0), W), x2 = min(max((int)bx2, 0), W);
            const int y1 = min(max((int)by1, 0), H), y2 = min(max((int)by2, 0), H);
            R = Region{d, nullptr, x1, y1, max(x2 - x1, 0), max(y2 - y1, 0), W};
            region_count(R, s, n, n_nan);
        }
        if (n == 0 || n_nan) res = nan;
        else res = region_quantile_range(R, n, 0.1f, 0.9f, s);
    }
    if (threadIdx.x == 0) {
        out[item] = res;
        if (flag) flag[item] = fl;
    }
}

}  // namespace

extern "C" int tdod_depth_select(const float* depth, int64_t F, int H, int W, const int32_t* frame, const float* boxes,
                                 const uint8_t* masks, const int32_t* mask_idx, int mode, int64_t n, float* out,
                                 int32_t* out_flag, void* stream) {
    if (n == 0) return TDOD_OK;
    if (!depth || !frame || !boxes || !out || n < 0 || F <= 0 || H <= 0 || W <= 0 || mode < 0 || mode > 2 ||
        (int64_t)H * W >= (1ll << 31))
        return tdod_fail(TDOD_EINVAL, "tdod_depth_select: bad argument");
    if (mode == 2 && (!masks || !mask_idx)) return tdod_fail(TDOD_EINVAL, "tdod_depth_select: mode 2 needs masks and mask_idx");
    if (n > 0x7fffffff) return tdod_fail(TDOD_EINVAL, "tdod_depth_select: too many boxes");
    depth_select_kernel<<<(unsigned)n, kSelT, 0, (cudaStream_t)stream>>>(depth, frame, boxes, masks, mask_idx, mode, H, W,
                                                                         out, out_flag);
    TDOD_CHECK_LAUNCH();
    return TDOD_OK;
}
