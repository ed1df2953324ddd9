// Device code of the batched segmentation scores (modeb.cu: all objects of a batch; raster.cu keeps its own
// self-contained copy of the same rules for the one-object entry point):
// cv2.convexHull on 8 points, and cv2.fillPoly on the truncated hull as closed-form per-row intervals.  The interval
// rules restate OpenCV's clipLine / Line / FillEdgeCollection (unpinned dependency of the reference, requirements.txt:13)
// and were validated bit-for-bit against cv2 4.13 (oracle/cv_spec.py, tests/test_cv_spec.py).
#pragma once
#include "tdod_device.cuh"

namespace tdod_raster {

typedef long long i64;

// cv2.clipLine on the image rectangle [0,W) x [0,H)  (Cohen-Sutherland, y first, truncation toward zero)
__device__ inline bool clip_line(int W, int H, i64& x1, i64& y1, i64& x2, i64& y2) {
    const i64 r = W - 1, b = H - 1;
    int c1 = (x1 < 0) + (x1 > r) * 2 + (y1 < 0) * 4 + (y1 > b) * 8;
    int c2 = (x2 < 0) + (x2 > r) * 2 + (y2 < 0) * 4 + (y2 > b) * 8;
    if ((c1 & c2) == 0 && (c1 | c2) != 0) {
        if (c1 & 12) {
            const i64 a = c1 < 8 ? 0 : b;
            x1 += (i64)((double)(a - y1) * (double)(x2 - x1) / (double)(y2 - y1));
            y1 = a;
            c1 = (x1 < 0) + (x1 > r) * 2;
        }
        if (c2 & 12) {
            const i64 a = c2 < 8 ? 0 : b;
            x2 += (i64)((double)(a - y2) * (double)(x2 - x1) / (double)(y2 - y1));
            y2 = a;
            c2 = (x2 < 0) + (x2 > r) * 2;
        }
        if ((c1 & c2) == 0 && (c1 | c2) != 0) {
            if (c1) {
                const i64 a = c1 == 1 ? 0 : r;
                y1 += (i64)((double)(a - x1) * (double)(y2 - y1) / (double)(x2 - x1));
                x1 = a; c1 = 0;
            }
            if (c2) {
                const i64 a = c2 == 1 ? 0 : r;
                y2 += (i64)((double)(a - x2) * (double)(y2 - y1) / (double)(x2 - x1));
                x2 = a; c2 = 0;
            }
        }
    }
    return (c1 | c2) == 0;
}

struct Edge {
    // Bresenham part (cv2.line, LINE_8) after clipping, x increasing
    int lx0, ly0, lsy, lmajor, lminor;
    bool lvalid, lsteep;
    // scanline part: active for ya <= y < yb, x(y) = xs + (y - ya) * dx  (16.16 fixed point)
    int ya, yb;
    i64 xs, dx;
    bool svalid;
};

__device__ inline void make_edge(int W, int H, int ax, int ay, int bx, int by, Edge& e) {
    // ---- line ----
    i64 x0 = ax, y0 = ay, x1 = bx, y1 = by;
    const bool in0 = ax >= 0 && ax < W && ay >= 0 && ay < H, in1 = bx >= 0 && bx < W && by >= 0 && by < H;
    e.lvalid = true;
    if (!(in0 && in1)) e.lvalid = clip_line(W, H, x0, y0, x1, y1);
    if (e.lvalid) {
        if (x1 < x0) { i64 t = x0; x0 = x1; x1 = t; t = y0; y0 = y1; y1 = t; }
        const int dx = (int)(x1 - x0);
        int dy = (int)(y1 - y0);
        e.lsy = dy >= 0 ? 1 : -1;
        dy = dy < 0 ? -dy : dy;
        e.lsteep = dy > dx;
        e.lmajor = e.lsteep ? dy : dx;
        e.lminor = e.lsteep ? dx : dy;
        e.lx0 = (int)x0; e.ly0 = (int)y0;
    }
    // ---- scan ----
    e.svalid = ay != by;
    if (e.svalid) {
        i64 cx0 = ax, cy0 = ay, cx1 = bx, cy1 = by;
        if (!(in0 && in1)) {
            i64 tx0 = ax, ty0 = ay, tx1 = bx, ty1 = by;
            clip_line(W, H, tx0, ty0, tx1, ty1);
            cx0 = tx0; cx1 = tx1;                       // x from the clipped copy, always
            if (ty0 != ty1) { cy0 = ty0; cy1 = ty1; }   // y only if the clipped copy is not flat
        }
        const i64 X0 = cx0 * 65536, X1 = cx1 * 65536;   // no +1/2 offset
        const i64 d = (X1 - X0) / (cy1 - cy0);          // C division: toward zero
        e.dx = d;
        if (ay < by) { e.ya = ay; e.yb = by; e.xs = X0 + ((i64)ay - cy0) * d; }
        else         { e.ya = by; e.yb = ay; e.xs = X1 + ((i64)by - cy1) * d; }
    }
}

// extent [xa, xb] of the edge's Bresenham pixels in row y (closed form of the err recurrence)
__device__ __forceinline__ bool line_row(const Edge& e, int y, int& xa, int& xb) {
    if (!e.lvalid) return false;
    const int J = (y - e.ly0) * e.lsy;
    // extents below 2^14 pixels (every image the fused path accepts: clamped coordinates span at most 2.5 W): all products
    // stay below 2^30 and the divisions are 32-bit (a 64-bit division is ~6 x the instructions on this machine)
    const bool small = e.lmajor < 16384;
    if (e.lsteep) {
        if (J < 0 || J > e.lmajor) return false;
        const int j = small ? (int)((2u * (unsigned)e.lminor * (unsigned)J + (unsigned)e.lmajor - 1u) / (2u * (unsigned)e.lmajor))
                            : (int)((2ll * e.lminor * J + e.lmajor - 1) / (2ll * e.lmajor));
        xa = xb = e.lx0 + j;
        return true;
    }
    if (J < 0 || J > e.lminor) return false;
    if (e.lminor == 0) { xa = e.lx0; xb = e.lx0 + e.lmajor; return true; }
    if (small) {
        const int num32 = 2 * e.lmajor * J - e.lmajor + 1;
        const unsigned den = 2u * (unsigned)e.lminor;
        const int kmin32 = num32 <= 0 ? 0 : (int)(((unsigned)num32 + den - 1u) / den);
        int kmax32 = (int)((unsigned)(2 * e.lmajor * (J + 1) - e.lmajor) / den);
        kmax32 = min(kmax32, e.lmajor);
        if (kmin32 > kmax32) return false;
        xa = e.lx0 + kmin32; xb = e.lx0 + kmax32;
        return true;
    }
    const i64 num = 2ll * e.lmajor * J - e.lmajor + 1;
    const int kmin = num <= 0 ? 0 : (int)((num + 2ll * e.lminor - 1) / (2ll * e.lminor));
    int kmax = (int)((2ll * e.lmajor * (J + 1) - e.lmajor) / (2ll * e.lminor));
    kmax = min(kmax, e.lmajor);
    if (kmin > kmax) return false;
    xa = e.lx0 + kmin; xb = e.lx0 + kmax;
    return true;
}

// strict monotone-chain hull of 8 fp32 points (== cv2.convexHull's vertex set), truncated to int32
__device__ inline int hull8(const float* U, const float* V, int* hx, int* hy) {
    float px[8], py[8];
    int n = 8;
#pragma unroll
    for (int i = 0; i < 8; ++i) { px[i] = U[i]; py[i] = V[i]; }
    // insertion sort by (x, y)
    for (int i = 1; i < 8; ++i) {
        const float x = px[i], y = py[i];
        int j = i - 1;
        while (j >= 0 && (px[j] > x || (px[j] == x && py[j] > y))) { px[j + 1] = px[j]; py[j + 1] = py[j]; --j; }
        px[j + 1] = x; py[j + 1] = y;
    }
    // drop exact duplicates
    int m = 1;
    for (int i = 1; i < n; ++i)
        if (px[i] != px[m - 1] || py[i] != py[m - 1]) { px[m] = px[i]; py[m] = py[i]; ++m; }
    n = m;
    if (n <= 2) {
        for (int i = 0; i < n; ++i) { hx[i] = (int)px[i]; hy[i] = (int)py[i]; }
        return n;
    }
    float sx[16], sy[16];
    int k = 0;
    auto cross = [](float ox, float oy, float ax, float ay, float bx, float by) {
        const float dax = __fsub_rn(ax, ox), day = __fsub_rn(ay, oy), dbx = __fsub_rn(bx, ox), dby = __fsub_rn(by, oy);
        return __dsub_rn(__dmul_rn((double)dax, (double)dby), __dmul_rn((double)day, (double)dbx));
    };
    for (int i = 0; i < n; ++i) {   // lower chain
        while (k >= 2 && cross(sx[k - 2], sy[k - 2], sx[k - 1], sy[k - 1], px[i], py[i]) <= 0) --k;
        sx[k] = px[i]; sy[k] = py[i]; ++k;
    }
    const int lo = k + 1;
    for (int i = n - 2; i >= 0; --i) {   // upper chain
        while (k >= lo && cross(sx[k - 2], sy[k - 2], sx[k - 1], sy[k - 1], px[i], py[i]) <= 0) --k;
        sx[k] = px[i]; sy[k] = py[i]; ++k;
    }
    --k;   // last point == first point
    for (int i = 0; i < k; ++i) { hx[i] = (int)sx[i]; hy[i] = (int)sy[i]; }   // np.int32 cast truncates toward zero
    return k;
}


// Mask overlap of ONE proposal's filled hull on the [::4, ::4] grid, rows split over the 32 lanes of a warp:
// ed[0..nv) the hull's edges (shared memory), [ymin, ymax] its row range, pre [Hs][stride] exclusive row prefix sums of
// the sampled mask.  Returns (inter, poly) reduced over the warp (integer sums: order-free, exact).
__device__ __forceinline__ void warp_raster_counts(const Edge* ed, int nv, int ymin, int ymax, const uint16_t* pre, int stride,
                                                   int Hs, int W, int lane, i64& inter_out, i64& poly_out) {
    i64 inter = 0, poly = 0;
    const int r0 = max(0, (ymin + 3) >> 2), r1 = min(Hs - 1, ymax >> 2);
    for (int r = r0 + lane; r <= r1 && ymax >= 0; r += 32) {
        const int y = 4 * r;
        // intervals in sample-index space [clo, chi]
        int clo[12], chi[12];
        int m = 0;
        // scanline spans: active edges sorted by x, paired even-odd
        i64 xs[8];
        int na = 0;
        for (int k = 0; k < nv; ++k) {
            const Edge& e = ed[k];
            if (!e.svalid || y < e.ya || y >= e.yb) continue;
            const i64 x = e.xs + (i64)(y - e.ya) * e.dx;
            int j = na - 1;
            while (j >= 0 && xs[j] > x) { xs[j + 1] = xs[j]; --j; }
            xs[j + 1] = x; ++na;
        }
        for (int k = 0; k + 1 < na; k += 2) {
            const i64 xa = (xs[k] + 65535) >> 16, xb = xs[k + 1] >> 16;    // ceil / floor of 16.16
            if (xa < W && xb >= 0) {
                const int a = (int)max((i64)0, xa), b = (int)min((i64)(W - 1), xb);
                if (a <= b && ((a + 3) >> 2) <= (b >> 2)) { clo[m] = (a + 3) >> 2; chi[m] = b >> 2; ++m; }
            }
        }
        for (int k = 0; k < nv; ++k) {
            int a, b;
            if (line_row(ed[k], y, a, b)) {
                a = max(a, 0); b = min(b, W - 1);
                if (a <= b && ((a + 3) >> 2) <= (b >> 2)) { clo[m] = (a + 3) >> 2; chi[m] = b >> 2; ++m; }
            }
        }
        if (m == 0) continue;
        // union of the intervals: insertion sort by start, then merge
        for (int a = 1; a < m; ++a) {
            const int l = clo[a], h = chi[a];
            int j = a - 1;
            while (j >= 0 && clo[j] > l) { clo[j + 1] = clo[j]; chi[j + 1] = chi[j]; --j; }
            clo[j + 1] = l; chi[j + 1] = h;
        }
        int cl = clo[0], ch = chi[0];
        const uint16_t* pr = pre + r * stride;
        for (int a = 1; a <= m; ++a) {
            if (a < m && clo[a] <= ch + 1) { ch = max(ch, chi[a]); continue; }
            poly += ch - cl + 1;
            inter += (int)pr[ch + 1] - (int)pr[cl];
            if (a < m) { cl = clo[a]; ch = chi[a]; }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        inter += __shfl_xor_sync(0xffffffffu, inter, o);
        poly += __shfl_xor_sync(0xffffffffu, poly, o);
    }
    inter_out = inter; poly_out = poly;
}

// mask_iou (utils.py:230-239) / mod_mask_iou (:241-250, int64 inter**5 wraparound as torch computes it) from the counts
__device__ __forceinline__ float iou_from_counts(i64 inter, i64 poly, int mask_total, bool power5) {
    if (inter == 0) return 0.f;
    const i64 uni = (i64)mask_total + poly - inter;
    i64 num = inter;
    if (power5) {
        const unsigned long long u = (unsigned long long)inter;
        num = (i64)(u * u * u * u * u);
    }
    return __fdiv_rn(__ll2float_rn(num), __ll2float_rn(uni));   // int64 / int64 -> fp32 true division
}

// Exclusive row prefix sums of mask[4r][4c] (one warp per row, ballot + popc) and the number of set samples.
// pre [Hs][Ws + 1]; `total` must be zeroed and a barrier must follow before use.  Called by all `nwarps` warps.
__device__ __forceinline__ void build_sampled_prefix(const uint8_t* __restrict__ mask, int H, int W, uint16_t* pre, int* total,
                                                     int wid, int lane, int nwarps) {
    const int Hs = (H + 3) >> 2, Ws = (W + 3) >> 2, stride = Ws + 1;
    for (int r = wid; r < Hs; r += nwarps) {
        const uint8_t* row = mask + (size_t)(4 * r) * W;
        int base = 0;
        if (lane == 0) pre[r * stride] = 0;
        for (int c0 = 0; c0 < Ws; c0 += 32) {
            const int c = c0 + lane;
            const bool bit = c < Ws && row[4 * c] != 0;
            const unsigned m = __ballot_sync(0xffffffffu, bit);
            if (c < Ws) pre[r * stride + c + 1] = (uint16_t)(base + __popc(m & (0xffffffffu >> (31 - lane))));
            base += __popc(m);
        }
        if (lane == 0 && base) atomicAdd(total, base);
    }
}

}  // namespace tdod_raster
