// K-B: object mask -> minimum-area rectangle of its largest external contour
// (first half of score_corners, ProposalNetwork/scoring/scorefunction.py:59-68:
//  cv2.findContours(RETR_EXTERNAL, CHAIN_APPROX_SIMPLE) -> max(contourArea) -> minAreaRect -> boxPoints).
//
// OpenCV is an unpinned binary dependency of the reference (requirements.txt:13); its behaviour is restated
// here from its published algorithms and was validated bit-for-bit against cv2 4.13 in the build container
// (oracle/cv_spec.py, tests/test_cv_spec.py):
//   * an external contour is the outer border of an 8-connected component that is not nested in a hole;
//     contourArea of that border polygon (through pixel centres) equals the sum over 2x2 pixel cells of
//     {1 if 4 corners, 1/2 if 3 corners} in the HOLE-FILLED component;
//   * python max() keeps the first maximum of cv2's list, which is ordered last-raster-first: ties go to
//     the component whose first pixel comes LAST in raster order;
//   * minAreaRect = Sklansky hull (strict, start at (max x, max y), image-clockwise) -> fp32 rotating calipers.
//
// One CTA per mask.  The mask is streamed once from HBM with 16-byte loads and bit-packed into shared
// memory; everything after that (run-length encoding, union-find over runs, areas, hull, calipers) runs out
// of shared memory.  Pass 1 decides every mask whose winner is certain from bounds (area without hole filling
// <= true area <= bounding box); pass 2 (flood.cu path below) fills holes exactly for the ambiguous rest.
#include "tdod_device.cuh"
#include "tdod_internal.h"

using namespace tdod;

namespace {

constexpr int kT = 512;
constexpr int kRunCap = 2048;     // runs per mask held in shared memory (more -> the global-memory overflow pass)

struct RunSet {                   // shared-memory views
    uint16_t* xs; uint16_t* xe; uint16_t* row;
    int* parent; int* area2;      // area2: twice the cell-formula area, accumulated at the root
    int* row_off;                 // [H+1] first run of each row
};

__device__ __forceinline__ uint32_t nz_bits16(const uint4 v) {
    // 16 mask bytes -> 16 bits (bit i = byte i != 0)
    uint32_t out = 0;
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const uint32_t t = w[k];
        const uint32_t hi = (t | ((t & 0x7f7f7f7fu) + 0x7f7f7f7fu)) & 0x80808080u;
        out |= ((((hi >> 7) * 0x00204081u) >> 21) & 0xFu) << (4 * k);
    }
    return out;
}

// OR `nbits` bits (LSB first) into the row-packed bitmap at flat pixel index p (may straddle rows)
__device__ __forceinline__ void scatter_bits(uint32_t* bm, int W, int Wp, int64_t p, uint32_t bits, int nbits) {
    while (nbits > 0 && bits) {
        const int r = (int)(p / W), c = (int)(p - (int64_t)r * W);
        const int take = min(nbits, W - c);
        const uint32_t seg = bits & ((take >= 32) ? 0xffffffffu : ((1u << take) - 1u));
        if (seg) {
            const int wi = c >> 5, sh = c & 31;
            atomicOr(&bm[r * Wp + wi], seg << sh);
            if (sh && (sh + take > 32)) atomicOr(&bm[r * Wp + wi + 1], seg >> (32 - sh));
        }
        bits = (take >= 32) ? 0u : (bits >> take);
        nbits -= take; p += take;
    }
}

// stream one mask (H*W bytes, any alignment) into the bitmap
__device__ void load_pack(const uint8_t* __restrict__ m, int H, int W, int Wp, uint32_t* bm) {
    const int64_t total = (int64_t)H * W;
    const uintptr_t addr = reinterpret_cast<uintptr_t>(m);
    const int64_t head_raw = (int64_t)((16 - (addr & 15)) & 15);
    const int64_t head = head_raw < total ? head_raw : total;
    for (int64_t i = threadIdx.x; i < head; i += kT)
        if (m[i]) scatter_bits(bm, W, Wp, i, 1u, 1);
    const int64_t nchunks = (total - head) / 16;
    const uint4* v = reinterpret_cast<const uint4*>(m + head);
    int64_t c = threadIdx.x;
    for (; c + 3 * kT < nchunks; c += 4 * kT) {   // four independent 16-byte loads in flight per thread
        const uint4 a0 = __ldcs(v + c), a1 = __ldcs(v + c + kT), a2 = __ldcs(v + c + 2 * kT), a3 = __ldcs(v + c + 3 * kT);
        const uint32_t b0 = nz_bits16(a0), b1 = nz_bits16(a1), b2 = nz_bits16(a2), b3 = nz_bits16(a3);
        if (b0) scatter_bits(bm, W, Wp, head + 16 * c, b0, 16);
        if (b1) scatter_bits(bm, W, Wp, head + 16 * (c + kT), b1, 16);
        if (b2) scatter_bits(bm, W, Wp, head + 16 * (c + 2 * kT), b2, 16);
        if (b3) scatter_bits(bm, W, Wp, head + 16 * (c + 3 * kT), b3, 16);
    }
    for (; c < nchunks; c += kT) {
        const uint32_t b = nz_bits16(__ldcs(v + c));
        if (b) scatter_bits(bm, W, Wp, head + 16 * c, b, 16);
    }
    for (int64_t i = head + nchunks * 16 + threadIdx.x; i < total; i += kT)
        if (m[i]) scatter_bits(bm, W, Wp, i, 1u, 1);
}

// number of runs in row r / emit them starting at index `base`
__device__ int row_runs(const uint32_t* bm, int Wp, int W, int r, RunSet* rs, int base, int cap) {
    int n = 0, start = 0;
    uint32_t prev = 0;
    for (int w = 0; w < Wp; ++w) {
        const uint32_t m = bm[r * Wp + w];
        uint32_t t = m ^ ((m << 1) | prev);
        prev = m >> 31;
        while (t) {
            const int b = __ffs(t) - 1;
            t &= t - 1;
            const int pos = 32 * w + b;
            if ((m >> b) & 1u) {
                start = pos;
            } else {
                if (rs && base + n < cap) { rs->xs[base + n] = (uint16_t)start; rs->xe[base + n] = (uint16_t)(pos - 1); rs->row[base + n] = (uint16_t)r; }
                ++n;
            }
        }
    }
    if (prev) {
        if (rs && base + n < cap) { rs->xs[base + n] = (uint16_t)start; rs->xe[base + n] = (uint16_t)(W - 1); rs->row[base + n] = (uint16_t)r; }
        ++n;
    }
    return n;
}

__device__ __forceinline__ int uf_find(const int* parent, int x) {
    int p = parent[x];
    while (p != x) { x = p; p = parent[x]; }
    return x;
}
__device__ __forceinline__ void uf_union(int* parent, int a, int b) {
    while (true) {
        a = uf_find(parent, a); b = uf_find(parent, b);
        if (a == b) return;
        if (a < b) { const int t = a; a = b; b = t; }
        const int old = atomicMin(&parent[a], b);   // a > b: hang the larger root under the smaller
        if (old == a) return;
        a = old;
    }
}

// block-wide exclusive scan of one int per row (H rows, any H) into row_off[0..H]; returns the total
__device__ int scan_rows(int* row_off, int H, int* scratch /*[kT/32 + 1]*/) {
    __shared__ int carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int base = 0; base < H; base += kT) {
        const int r = base + threadIdx.x;
        const int v = (r < H) ? row_off[r] : 0;
        int inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) scratch[wid] = inc;
        __syncthreads();
        int woff = 0;
        for (int w = 0; w < wid; ++w) woff += scratch[w];
        const int c0 = carry;
        __syncthreads();
        if (r < H) row_off[r] = c0 + woff + inc - v;
        if (threadIdx.x == kT - 1) carry = c0 + woff + inc;
        __syncthreads();
    }
    if (threadIdx.x == 0) row_off[H] = carry;
    __syncthreads();
    return carry;
}

// label the runs of `bm` (8-connectivity), accumulate twice the cell-formula area at each root.
// returns the number of runs, or -1 if they exceed the capacity.
__device__ int label_runs(const uint32_t* bm, int H, int W, int Wp, RunSet rs, int* scratch, int cap) {
    for (int r = threadIdx.x; r < H; r += kT) rs.row_off[r] = row_runs(bm, Wp, W, r, nullptr, 0, 0);
    __syncthreads();
    const int R = scan_rows(rs.row_off, H, scratch);
    if (R > cap) return -1;
    for (int r = threadIdx.x; r < H; r += kT) row_runs(bm, Wp, W, r, &rs, rs.row_off[r], cap);
    for (int i = threadIdx.x; i < R; i += kT) { rs.parent[i] = i; rs.area2[i] = 0; }
    __syncthreads();
    for (int i = threadIdx.x; i < R; i += kT) {
        const int r = rs.row[i];
        if (r == 0) continue;
        const int s = rs.xs[i], e = rs.xe[i];
        for (int j = rs.row_off[r - 1]; j < rs.row_off[r]; ++j) {
            if ((int)rs.xs[j] > e + 1) break;
            if ((int)rs.xe[j] >= s - 1) uf_union(rs.parent, i, j);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < R; i += kT) rs.parent[i] = uf_find(rs.parent, i);
    __syncthreads();
    // cell formula on pairs of overlapping runs in adjacent rows: 2*full + (#3-corner end cells)
    for (int i = threadIdx.x; i < R; i += kT) {
        const int r = rs.row[i];
        if (r == 0) continue;
        const int s = rs.xs[i], e = rs.xe[i];
        int acc = 0;
        for (int j = rs.row_off[r - 1]; j < rs.row_off[r]; ++j) {
            const int s2 = rs.xs[j], e2 = rs.xe[j];
            if (s2 > e) break;
            if (e2 < s) continue;
            acc += 2 * (min(e, e2) - max(s, s2)) + (s != s2) + (e != e2);
        }
        if (acc) atomicAdd(&rs.area2[rs.parent[i]], acc);
    }
    __syncthreads();
    return R;
}

// ---- hull + rotating calipers of the component `root` ----------------------------------------------------------
// Only two things are inherently sequential: the monotone-chain stacks (one thread per chain, top of stack in
// registers) and the calipers' walk around the hull (one thread, everything it reads precomputed in parallel).
constexpr int kChainCap = 512;    // vertices of one strictly convex chain (a lattice polygon in a <=1024^2 grid has < 360)
constexpr int kHullCap = 2 * kChainCap;
struct HullBuf { int16_t* rx; int16_t* ry; int16_t* lx; int16_t* ly; float* hx; float* hy; float* vx; float* vy; float* inv; };

// OpenCV rotatingCalipers (CALIPERS_MINAREARECT), fp32 with every product/sum rounded on its own (the x86
// build has no FMA contraction), then minAreaRect's centre/size/angle and RotatedRect::points (boxPoints).
// Called by all threads; `ext` = {left, right, top, bottom} start indices (first extreme in hull order).
__device__ void calipers_box(const HullBuf& hb, int n, int* ext /*[4] smem*/, float* out8) {
    const float* px = hb.hx; const float* py = hb.hy;
    const int tid = threadIdx.x;
    if (n >= 3) {
        // edge vectors and inverse lengths (double sqrt + divide each): one edge per thread
        for (int i = tid; i < n; i += kT) {
            const int j = (i + 1 < n) ? i + 1 : 0;
            const double dx = (double)px[j] - (double)px[i], dy = (double)py[j] - (double)py[i];
            hb.vx[i] = (float)dx; hb.vy[i] = (float)dy;
            hb.inv[i] = (float)(1.0 / sqrt(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy))));
        }
        // extremes with the serial scan's semantics (strict compares from index 0: the FIRST extreme wins)
        if (tid < 128) {
            const int which = tid >> 5, lane = tid & 31;     // 0 left (min x), 1 right (max x), 2 top (max y), 3 bottom (min y)
            const float* src = (which < 2) ? px : py;
            const bool want_min = (which == 0 || which == 3);
            float bv = src[0]; int bi = 0;
            for (int i = lane; i < n; i += 32) {
                const float v = src[i];
                if (want_min ? (v < bv || (v == bv && i < bi)) : (v > bv || (v == bv && i < bi))) { bv = v; bi = i; }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float v = __shfl_xor_sync(0xffffffffu, bv, o);
                const int i2 = __shfl_xor_sync(0xffffffffu, bi, o);
                if (want_min ? (v < bv || (v == bv && i2 < bi)) : (v > bv || (v == bv && i2 < bi))) { bv = v; bi = i2; }
            }
            if (lane == 0) ext[which] = bi;
        }
    }
    __syncthreads();
    if (tid != 0) return;
    if (n == 1) {
        for (int k = 0; k < 4; ++k) { out8[2 * k] = px[0]; out8[2 * k + 1] = py[0]; }
        return;
    }
    float cx, cy, bw, bh;
    double ang_deg;
    if (n == 2) {
        cx = __fmul_rn(__fadd_rn(px[0], px[1]), 0.5f); cy = __fmul_rn(__fadd_rn(py[0], py[1]), 0.5f);
        const double dx = (double)__fsub_rn(px[1], px[0]), dy = (double)__fsub_rn(py[1], py[0]);
        bw = (float)sqrt(dx * dx + dy * dy); bh = 0.f;
        ang_deg = (double)(float)((double)(float)atan2(dy, dx) * 180.0 / 3.14159265358979323846);
    } else {
        const int left = ext[0], right = ext[1], top = ext[2], bottom = ext[3];
        float orientation = 0.f;
        {
            double ax = hb.vx[n - 1], ay = hb.vy[n - 1];
            for (int i = 0; i < n; ++i) {
                const double bx = hb.vx[i], by = hb.vy[i];
                const double conv = __dsub_rn(__dmul_rn(ax, by), __dmul_rn(ay, bx));
                if (conv != 0) { orientation = (conv > 0) ? 1.f : -1.f; break; }
                ax = bx; ay = by;
            }
        }
        float base_a = orientation, base_b = 0.f;
        int seq[4] = {bottom, right, top, left};
        // the four calipers' current edge (vx, vy, inv) and point (px, py) live in registers; only the caliper
        // that advances reloads from shared memory
        float evx[4], evy[4], einv[4], ppx[4], ppy[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            evx[k] = hb.vx[seq[k]]; evy[k] = hb.vy[seq[k]]; einv[k] = hb.inv[seq[k]];
            ppx[k] = px[seq[k]]; ppy[k] = py[seq[k]];
        }
        float minarea = 3.402823466e+38f;
        int b_left = 0, b_bottom = 0; float b_a = 0, b_w = 0, b_b = 0, b_h = 0;
        for (int k = 0; k < n; ++k) {
            const float dp0 = __fadd_rn(__fmul_rn(base_a, evx[0]), __fmul_rn(base_b, evy[0]));
            const float dp1 = __fadd_rn(__fmul_rn(-base_b, evx[1]), __fmul_rn(base_a, evy[1]));
            const float dp2 = __fsub_rn(__fmul_rn(-base_a, evx[2]), __fmul_rn(base_b, evy[2]));
            const float dp3 = __fsub_rn(__fmul_rn(base_b, evx[3]), __fmul_rn(base_a, evy[3]));
            float maxcos = __fmul_rn(dp0, einv[0]);
            int main_el = 0;
            const float c1 = __fmul_rn(dp1, einv[1]); if (c1 > maxcos) { main_el = 1; maxcos = c1; }
            const float c2 = __fmul_rn(dp2, einv[2]); if (c2 > maxcos) { main_el = 2; maxcos = c2; }
            const float c3 = __fmul_rn(dp3, einv[3]); if (c3 > maxcos) { main_el = 3; maxcos = c3; }
            float lvx = evx[0], lvy = evy[0], linv = einv[0];
            if (main_el == 1) { lvx = evx[1]; lvy = evy[1]; linv = einv[1]; }
            else if (main_el == 2) { lvx = evx[2]; lvy = evy[2]; linv = einv[2]; }
            else if (main_el == 3) { lvx = evx[3]; lvy = evy[3]; linv = einv[3]; }
            const float lead_x = __fmul_rn(lvx, linv), lead_y = __fmul_rn(lvy, linv);
            if (main_el == 0) { base_a = lead_x; base_b = lead_y; }
            else if (main_el == 1) { base_a = lead_y; base_b = -lead_x; }
            else if (main_el == 2) { base_a = -lead_x; base_b = -lead_y; }
            else { base_a = -lead_y; base_b = lead_x; }
            // advance that caliper and reload its edge / point
            int nx;
#pragma unroll
            for (int m = 0; m < 4; ++m)
                if (m == main_el) {
                    nx = (seq[m] + 1 == n) ? 0 : seq[m] + 1;
                    seq[m] = nx;
                    evx[m] = hb.vx[nx]; evy[m] = hb.vy[nx]; einv[m] = hb.inv[nx]; ppx[m] = px[nx]; ppy[m] = py[nx];
                }
            float dx = __fsub_rn(ppx[1], ppx[3]), dy = __fsub_rn(ppy[1], ppy[3]);
            const float width = __fadd_rn(__fmul_rn(dx, base_a), __fmul_rn(dy, base_b));
            dx = __fsub_rn(ppx[2], ppx[0]); dy = __fsub_rn(ppy[2], ppy[0]);
            const float height = __fadd_rn(__fmul_rn(-dx, base_b), __fmul_rn(dy, base_a));
            const float area = __fmul_rn(width, height);
            if (area <= minarea) {
                minarea = area; b_left = seq[3]; b_a = base_a; b_w = width; b_b = base_b; b_h = height; b_bottom = seq[0];
            }
        }
        const float A1 = b_a, B1 = b_b, A2 = -b_b, B2 = b_a;
        const float C1 = __fadd_rn(__fmul_rn(A1, px[b_left]), __fmul_rn(py[b_left], B1));
        const float C2 = __fadd_rn(__fmul_rn(A2, px[b_bottom]), __fmul_rn(py[b_bottom], B2));
        const float idet = __fdiv_rn(1.f, __fsub_rn(__fmul_rn(A1, B2), __fmul_rn(A2, B1)));
        const float ox = __fmul_rn(__fsub_rn(__fmul_rn(C1, B2), __fmul_rn(C2, B1)), idet);
        const float oy = __fmul_rn(__fsub_rn(__fmul_rn(A1, C2), __fmul_rn(A2, C1)), idet);
        const float o2 = __fmul_rn(A1, b_w), o3 = __fmul_rn(B1, b_w), o4 = __fmul_rn(A2, b_h), o5 = __fmul_rn(B2, b_h);
        cx = __fadd_rn(ox, __fmul_rn(__fadd_rn(o2, o4), 0.5f));
        cy = __fadd_rn(oy, __fmul_rn(__fadd_rn(o3, o5), 0.5f));
        bw = (float)sqrt(__dadd_rn(__dmul_rn((double)o2, (double)o2), __dmul_rn((double)o3, (double)o3)));
        bh = (float)sqrt(__dadd_rn(__dmul_rn((double)o4, (double)o4), __dmul_rn((double)o5, (double)o5)));
        // cv2 4.13 reports the angle in [-90, 0): (w,h,a) -> (h,w,a-90) for a in [0,90), (w,h,a-180) for a == 90
        const double ad = atan2((double)o3, (double)o2) * 180.0 / 3.14159265358979323846;
        if (ad >= 90.0) { ang_deg = (double)(float)(ad - 180.0); }
        else { const float t = bw; bw = bh; bh = t; ang_deg = (double)(float)(ad - 90.0); }
    }
    // RotatedRect::points
    const double ar = ang_deg * 3.14159265358979323846 / 180.0;
    const float b = __fmul_rn((float)cos(ar), 0.5f), a = __fmul_rn((float)sin(ar), 0.5f);
    const float p0x = __fsub_rn(__fsub_rn(cx, __fmul_rn(a, bh)), __fmul_rn(b, bw));
    const float p0y = __fsub_rn(__fadd_rn(cy, __fmul_rn(b, bh)), __fmul_rn(a, bw));
    const float p1x = __fsub_rn(__fadd_rn(cx, __fmul_rn(a, bh)), __fmul_rn(b, bw));
    const float p1y = __fsub_rn(__fsub_rn(cy, __fmul_rn(b, bh)), __fmul_rn(a, bw));
    out8[0] = p0x; out8[1] = p0y; out8[2] = p1x; out8[3] = p1y;
    out8[4] = __fsub_rn(__fmul_rn(2.f, cx), p0x); out8[5] = __fsub_rn(__fmul_rn(2.f, cy), p0y);
    out8[6] = __fsub_rn(__fmul_rn(2.f, cx), p1x); out8[7] = __fsub_rn(__fmul_rn(2.f, cy), p1y);
}

// Strict convex hull of the component `root` from its per-row extremes, in cv2.convexHull's output order
// (start at (max x, max y), then image-clockwise), followed by the calipers.  Called by all threads.
// counts: int[8] shared scratch.  Returns false if a chain outgrew kChainCap (cannot happen below 1024^2 pixels).
__device__ bool component_rect(const RunSet& rs, int R, int root, int H, int* rowL, int* rowR, const HullBuf& hb,
                               int* counts, float* out8) {
    const int tid = threadIdx.x;
    if (tid == 0) { counts[0] = 0; counts[1] = 0; counts[2] = 0x7fffffff; counts[3] = -1; counts[7] = 0; }
    for (int r = tid; r < H; r += kT) { rowL[r] = 0x7fffffff; rowR[r] = -1; }
    __syncthreads();
    for (int i = tid; i < R; i += kT)
        if (rs.parent[i] == root) {
            const int r = rs.row[i];
            atomicMin(&rowL[r], (int)rs.xs[i]); atomicMax(&rowR[r], (int)rs.xe[i]);
            atomicMin(&counts[2], r); atomicMax(&counts[3], r);
        }
    __syncthreads();
    const int rmin = counts[2], rmax = counts[3];      // a connected component occupies every row in between
    if (tid == 0 || tid == 32) {
        // tid 0: right chain, tid 32: left chain, both top -> bottom; pop while the turn is not strictly convex.
        // (ax,ay) / (bx,by) mirror the two topmost stack entries so that a push costs no shared-memory read.
        const bool right = tid == 0;
        int16_t* X = right ? hb.rx : hb.lx; int16_t* Y = right ? hb.ry : hb.ly;
        const int* src = right ? rowR : rowL;
        int n = 0, ax = 0, ay = 0, bx = 0, by = 0;
        bool ok = true;
        for (int r = rmin; r <= rmax; ++r) {
            const int cx = src[r], cy = r;
            while (n >= 2) {
                const int cr = (bx - ax) * (cy - by) - (by - ay) * (cx - bx);
                if (right ? (cr <= 0) : (cr >= 0)) {
                    --n; bx = ax; by = ay;
                    if (n >= 2) { ax = X[n - 2]; ay = Y[n - 2]; }
                } else break;
            }
            if (n >= kChainCap) { ok = false; break; }
            X[n] = (int16_t)cx; Y[n] = (int16_t)cy; ++n;
            ax = bx; ay = by; bx = cx; by = cy;
        }
        counts[right ? 0 : 1] = n;
        if (!ok) counts[7] = 1;
    }
    __syncthreads();
    if (counts[7]) return false;
    const int nr = counts[0], nl = counts[1];
    // cyclic order: right chain downward, then left chain upward (dropping duplicated end points)
    const bool dup_bottom = hb.lx[nl - 1] == hb.rx[nr - 1] && hb.ly[nl - 1] == hb.ry[nr - 1];
    const bool dup_top = hb.lx[0] == hb.rx[0] && hb.ly[0] == hb.ry[0];
    // left-chain entries kept: indices [lo, hi] walked downward from hi
    const int hi = nl - 1 - (dup_bottom ? 1 : 0), lo = dup_top ? 1 : 0;
    int n = nr + max(0, hi - lo + 1);
    for (int i = tid; i < n; i += kT) {
        if (i < nr) { hb.hx[i] = (float)hb.rx[i]; hb.hy[i] = (float)hb.ry[i]; }
        else { const int k = hi - (i - nr); hb.hx[i] = (float)hb.lx[k]; hb.hy[i] = (float)hb.ly[k]; }
    }
    __syncthreads();
    // collinear triples can only survive across the chain joins of degenerate (1- or 2-row, needle-like) components:
    // detect in parallel, repair serially only then
    int bad = 0;
    if (n > 2)
        for (int i = tid; i < n; i += kT) {
            const int a = (i + n - 1) % n, c = (i + 1) % n;
            const float cr = (hb.hx[i] - hb.hx[a]) * (hb.hy[c] - hb.hy[i]) - (hb.hy[i] - hb.hy[a]) * (hb.hx[c] - hb.hx[i]);
            bad |= (cr == 0.f);
        }
    if (__syncthreads_or(bad)) {
        if (tid == 0) {
            bool changed = n > 2;
            while (changed) {
                changed = false;
                for (int i = 0; i < n && n > 2; ++i) {
                    const int a = (i + n - 1) % n, c = (i + 1) % n;
                    const float cr = (hb.hx[i] - hb.hx[a]) * (hb.hy[c] - hb.hy[i]) - (hb.hy[i] - hb.hy[a]) * (hb.hx[c] - hb.hx[i]);
                    if (cr == 0.f) {
                        for (int k = i; k + 1 < n; ++k) { hb.hx[k] = hb.hx[k + 1]; hb.hy[k] = hb.hy[k + 1]; }
                        --n; changed = true; break;
                    }
                }
            }
            counts[4] = n;
        }
        __syncthreads();
        n = counts[4];
    }
    // rotate so that the start is (max x, max y among those): first such index in a serial scan == the unique argmax
    if (tid < 32) {
        float bxv = hb.hx[0], byv = hb.hy[0]; int bi = 0;
        for (int i = tid; i < n; i += 32) {
            const float x = hb.hx[i], y = hb.hy[i];
            if (x > bxv || (x == bxv && y > byv) || (x == bxv && y == byv && i < bi)) { bxv = x; byv = y; bi = i; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float x = __shfl_xor_sync(0xffffffffu, bxv, o), y = __shfl_xor_sync(0xffffffffu, byv, o);
            const int i2 = __shfl_xor_sync(0xffffffffu, bi, o);
            if (x > bxv || (x == bxv && y > byv) || (x == bxv && y == byv && i2 < bi)) { bxv = x; byv = y; bi = i2; }
        }
        if (tid == 0) counts[5] = bi;
    }
    __syncthreads();
    const int s0 = counts[5];
    if (s0) {   // rotate left by s0 through the (still unused) vx/vy scratch
        for (int i = tid; i < n; i += kT) { hb.vx[i] = hb.hx[(i + s0) % n]; hb.vy[i] = hb.hy[(i + s0) % n]; }
        __syncthreads();
        for (int i = tid; i < n; i += kT) { hb.hx[i] = hb.vx[i]; hb.hy[i] = hb.vy[i]; }
        __syncthreads();
    }
    if (n == 2 && hb.hx[0] == hb.hx[1] && hb.hy[0] == hb.hy[1]) n = 1;
    calipers_box(hb, n, counts, out8);
    __syncthreads();
    return true;
}

struct Smem {
    uint32_t* bmA; uint32_t* bmB;
    RunSet rs;
    int* rowL; int* rowR;
    HullBuf hb;
    int* scratch;     // [kT/32 + 24]
    float* out8;
};

__host__ __device__ inline size_t align16(size_t x) { return (x + 15) & ~size_t(15); }

__host__ __device__ inline size_t smem_layout(int H, int Wp, bool two_bitmaps, Smem* s, unsigned char* base) {
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align16(off + bytes); return o; };
    const size_t o_bmA = take((size_t)H * Wp * 4);
    const size_t o_bmB = two_bitmaps ? take((size_t)H * Wp * 4) : 0;
    const size_t o_xs = take(kRunCap * 2), o_xe = take(kRunCap * 2), o_row = take(kRunCap * 2);
    const size_t o_par = take(kRunCap * 4), o_area = take(kRunCap * 4), o_roff = take((size_t)(H + 1) * 4);
    const size_t o_rowL = take((size_t)H * 4), o_rowR = take((size_t)H * 4);
    const size_t o_rx = take(kChainCap * 2), o_ry = take(kChainCap * 2), o_lx = take(kChainCap * 2), o_ly = take(kChainCap * 2);
    const size_t hn = (size_t)kHullCap;
    const size_t o_hx = take(hn * 4), o_hy = take(hn * 4), o_vx = take(hn * 4), o_vy = take(hn * 4), o_inv = take(hn * 4);
    const size_t o_scr = take((kT / 32 + 32) * 4), o_out = take(8 * 4);
    if (s) {
        s->bmA = (uint32_t*)(base + o_bmA); s->bmB = two_bitmaps ? (uint32_t*)(base + o_bmB) : nullptr;
        s->rs.xs = (uint16_t*)(base + o_xs); s->rs.xe = (uint16_t*)(base + o_xe); s->rs.row = (uint16_t*)(base + o_row);
        s->rs.parent = (int*)(base + o_par); s->rs.area2 = (int*)(base + o_area); s->rs.row_off = (int*)(base + o_roff);
        s->rowL = (int*)(base + o_rowL); s->rowR = (int*)(base + o_rowR);
        s->hb.rx = (int16_t*)(base + o_rx); s->hb.ry = (int16_t*)(base + o_ry);
        s->hb.lx = (int16_t*)(base + o_lx); s->hb.ly = (int16_t*)(base + o_ly);
        s->hb.hx = (float*)(base + o_hx); s->hb.hy = (float*)(base + o_hy);
        s->hb.vx = (float*)(base + o_vx); s->hb.vy = (float*)(base + o_vy); s->hb.inv = (float*)(base + o_inv);
        s->scratch = (int*)(base + o_scr); s->out8 = (float*)(base + o_out);
    }
    return off;
}

// pick the component with the largest area2 (ties: the larger root index = later first pixel in raster order).
// In bounds mode also reports whether any OTHER component's bounding-box area could reach the winner's
// hole-free area (then hole filling could change the answer and pass 2 must decide).
__device__ int select_component(const RunSet& rs, int R, bool bounds_mode, int* scratch, bool* ambiguous) {
    // scratch[0] = best key hi (area2), scratch[1] = best root, via a two-step atomicMax on a packed 64-bit key
    unsigned long long* key = reinterpret_cast<unsigned long long*>(scratch + 20);   // 8-byte aligned slot
    if (threadIdx.x == 0) { *key = 0ull; scratch[24] = 0; }
    __syncthreads();
    for (int i = threadIdx.x; i < R; i += kT)
        if (rs.parent[i] == i) atomicMax(key, ((unsigned long long)(unsigned)rs.area2[i] << 32) | (unsigned)i);
    __syncthreads();
    const int best = (int)(*key & 0xffffffffu);
    const int best_a2 = (int)(*key >> 32);
    if (bounds_mode) {
        // bounding boxes per root: reuse area2 of non-root runs?  no -- compute on the fly per root with atomics
        // on four small arrays would cost smem; instead every run contributes to its root's extents held in
        // rowL/rowR-free scratch is not available here, so scan: a root's bbox is reduced by the threads that
        // own its runs through atomics on packed words stored in area2 of its FIRST non-root run ... too
        // intricate; simple exact alternative: for every non-best root, one thread walks all runs once.
        for (int i = threadIdx.x; i < R; i += kT) {
            if (rs.parent[i] != i || i == best) continue;
            int x0 = 0x7fffffff, x1 = -1, y0 = 0x7fffffff, y1 = -1;
            for (int j = i; j < R; ++j)
                if (rs.parent[j] == i) {
                    x0 = min(x0, (int)rs.xs[j]); x1 = max(x1, (int)rs.xe[j]);
                    y0 = min(y0, (int)rs.row[j]); y1 = max(y1, (int)rs.row[j]);
                }
            const long long ub2 = 2ll * (x1 - x0) * (y1 - y0);
            if (ub2 >= (long long)best_a2) atomicOr(&scratch[24], 1);
        }
        __syncthreads();
        *ambiguous = scratch[24] != 0;
    } else {
        *ambiguous = false;
    }
    __syncthreads();
    return best;
}

// ---- exact hole filling: outside = 4-connected flood of the background from the (virtual) image border ----
// Bit-parallel, iterated to convergence.  `out` starts from everything visible from the border along a row or a
// column (exact already for hole-free convex-ish blobs); every sweep lets the frontier run through a whole word
// horizontally and one row vertically.
__device__ __forceinline__ uint32_t fill_word(uint32_t seed, uint32_t free_) {
    // propagate seed bits through runs of free bits, both directions, inside one word
    uint32_t s = seed & free_;
    // towards higher bits: carry trick  (free + s) flips the run above each seed
    const uint32_t up = (free_ ^ (free_ + s)) & free_;   // f + s ripples a carry through the run above each seed
    // towards lower bits: same on the bit-reversed word
    const uint32_t rf = __brev(free_), rsd = __brev(s);
    const uint32_t dn = __brev((rf ^ (rf + rsd)) & rf);
    return s | up | dn;
}

__device__ void flood_outside(const uint32_t* mask, uint32_t* out, int H, int W, int Wp, int* flag) {
    const uint32_t tail = (W & 31) ? ((1u << (W & 31)) - 1u) : 0xffffffffu;   // valid bits of the last word
    // seed: per row, everything left of the first / right of the last foreground pixel
    for (int r = threadIdx.x; r < H; r += kT) {
        int first = -1, last = -1;
        for (int w = 0; w < Wp; ++w) {
            const uint32_t m = mask[r * Wp + w];
            if (m) { if (first < 0) first = 32 * w + __ffs(m) - 1; last = 32 * w + 31 - __clz(m); }
        }
        for (int w = 0; w < Wp; ++w) {
            uint32_t o;
            if (first < 0) o = 0xffffffffu;
            else {
                const int lo = 32 * w, hi = lo + 31;
                uint32_t a = (first > hi) ? 0xffffffffu : (first <= lo ? 0u : ((1u << (first - lo)) - 1u));
                uint32_t b = (last < lo) ? 0xffffffffu : (last >= hi ? 0u : ~((2u << (last - lo)) - 1u));
                o = a | b;
            }
            if (w == Wp - 1) o &= tail;
            out[r * Wp + w] = o;
        }
    }
    __syncthreads();
    // seed: per word-column, everything above the first / below the last foreground pixel of each bit column
    for (int w = threadIdx.x; w < Wp; w += kT) {
        uint32_t seen = 0;
        for (int r = 0; r < H; ++r) { seen |= mask[r * Wp + w]; out[r * Wp + w] |= ~seen & ((w == Wp - 1) ? tail : 0xffffffffu); }
        seen = 0;
        for (int r = H - 1; r >= 0; --r) { seen |= mask[r * Wp + w]; out[r * Wp + w] |= ~seen & ((w == Wp - 1) ? tail : 0xffffffffu); }
    }
    __syncthreads();
    for (int iter = 0; iter < 4 * (H + 32 * Wp); ++iter) {
        if (threadIdx.x == 0) *flag = 0;
        __syncthreads();
        bool changed = false;
        for (int i = threadIdx.x; i < H * Wp; i += kT) {
            const int r = i / Wp, w = i - r * Wp;
            const uint32_t free_ = ~mask[i] & ((w == Wp - 1) ? tail : 0xffffffffu);
            const uint32_t cur = out[i];
            uint32_t seed = cur;
            if (r > 0) seed |= out[i - Wp];
            if (r + 1 < H) seed |= out[i + Wp];
            if (w > 0) seed |= out[i - 1] >> 31;
            if (w + 1 < Wp) seed |= out[i + 1] << 31;
            const uint32_t nw = fill_word(seed, free_);
            if (nw != cur) { out[i] = nw | cur; changed = true; }
        }
        if (changed) *flag = 1;
        __syncthreads();
        const int any = *flag;
        __syncthreads();
        if (!any) break;
    }
}

// worst-case number of runs of an H x W mask, and the bytes of one global-memory run set of that capacity
__host__ __device__ inline int64_t run_cap_global(int H, int W) { return (int64_t)H * ((W + 1) / 2); }
__host__ __device__ inline size_t slot_bytes(int H, int W) { return align16((size_t)run_cap_global(H, W) * 14 + 64); }
constexpr int kSlots = 32;        // CTAs (and workspace slots) of the overflow pass

// kExact: fill holes before labelling.  kGlobal: the run arrays live in a workspace slot (masks with more runs than
// shared memory holds -- speckle noise); the CTA then walks every mask whose state is 3.
template <bool kExact, bool kGlobal>
__global__ void __launch_bounds__(kT, kExact ? 1 : 2) mask_rect_kernel(const uint8_t* __restrict__ masks, int64_t n_masks, int H, int W,
                                                          float* __restrict__ out_rects, uint8_t* __restrict__ out_has,
                                                          unsigned char* __restrict__ ws) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int Wp = (W + 31) >> 5;
    Smem s;
    smem_layout(H, Wp, kExact, &s, smem_raw);
    int cap = kRunCap;
    if (kGlobal) {
        const int64_t c = run_cap_global(H, W);
        unsigned char* b = ws + (size_t)blockIdx.x * slot_bytes(H, W);
        s.rs.parent = (int*)b; s.rs.area2 = (int*)(b + 4 * c);
        s.rs.xs = (uint16_t*)(b + 8 * c); s.rs.xe = (uint16_t*)(b + 10 * c); s.rs.row = (uint16_t*)(b + 12 * c);
        cap = (int)c;
    }
    for (int64_t mi = blockIdx.x; mi < n_masks; mi += gridDim.x) {
        __syncthreads();                             // previous mask's shared state is dead
        const int state = out_has[mi];
        if (kGlobal && state != 3) continue;         // overflow pass: only masks the shared-memory passes gave up on
        if (!kGlobal && kExact && state != 2) continue;   // pass 2 only revisits the masks pass 1 could not decide
        for (int i = threadIdx.x; i < H * Wp; i += kT) s.bmA[i] = 0u;
        __syncthreads();
        load_pack(masks + mi * (int64_t)H * W, H, W, Wp, s.bmA);
        __syncthreads();
        const uint32_t* comp_bm = s.bmA;
        if (kExact) {
            flood_outside(s.bmA, s.bmB, H, W, Wp, s.scratch + 30);
            const uint32_t tail = (W & 31) ? ((1u << (W & 31)) - 1u) : 0xffffffffu;
            for (int i = threadIdx.x; i < H * Wp; i += kT) {
                const int w = i % Wp;
                s.bmB[i] = ~s.bmB[i] & ((w == Wp - 1) ? tail : 0xffffffffu);    // M = not outside (foreground + holes)
            }
            __syncthreads();
            comp_bm = s.bmB;
        }
        const int R = label_runs(comp_bm, H, W, Wp, s.rs, s.scratch, cap);
        if (R == 0) {                                   // empty mask: scorefunction.py:69-75 fallback applies
            if (threadIdx.x == 0) out_has[mi] = 0;
            continue;
        }
        if (R < 0) {                                    // more runs than shared memory holds -> overflow pass
            if (threadIdx.x == 0) out_has[mi] = 3;
            continue;
        }
        bool ambiguous = false;
        const int root = select_component(s.rs, R, !kExact, s.scratch, &ambiguous);
        if (ambiguous) {
            if (threadIdx.x == 0) out_has[mi] = 2;
            continue;
        }
        const bool ok = component_rect(s.rs, R, root, H, s.rowL, s.rowR, s.hb, s.scratch + 32, s.out8);
        if (threadIdx.x < 8 && ok) out_rects[mi * 8 + threadIdx.x] = s.out8[threadIdx.x];
        if (threadIdx.x == 0) out_has[mi] = ok ? 1 : 4;
    }
}

}  // namespace

extern "C" int64_t tdod_mask_rect_workspace_bytes(int64_t n_masks, int H, int W) {
    if (n_masks <= 0 || H <= 0 || W <= 0) return 16;
    const int64_t slots = n_masks < kSlots ? n_masks : kSlots;
    return (int64_t)(slots * slot_bytes(H, W));
}

extern "C" int tdod_mask_rect(const uint8_t* masks, int64_t n_masks, int H, int W, float* out_rects,
                              uint8_t* out_has_rect, void* workspace, int64_t workspace_bytes, void* stream) {
    if (n_masks == 0) return TDOD_OK;
    if (!masks || !out_rects || !out_has_rect || n_masks < 0 || H <= 0 || W < 16)
        return tdod_fail(TDOD_EINVAL, "tdod_mask_rect: bad argument (W must be >= 16)");
    if (H > 32767 || W > 32767) return tdod_fail(TDOD_ERANGE, "tdod_mask_rect: image larger than 32767");
    if (!workspace || workspace_bytes < tdod_mask_rect_workspace_bytes(n_masks, H, W))
        return tdod_fail(TDOD_EINVAL, "tdod_mask_rect: workspace smaller than tdod_mask_rect_workspace_bytes()");
    const int Wp = (W + 31) >> 5;
    const size_t smem1 = smem_layout(H, Wp, false, nullptr, nullptr);
    const size_t smem2 = smem_layout(H, Wp, true, nullptr, nullptr);
    if (smem2 > 227 * 1024)
        return tdod_fail(TDOD_ERANGE, "tdod_mask_rect: mask does not fit in shared memory (H*ceil(W/32)*8 bytes + 60 KB <= 227 KB)");
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t e = cudaFuncSetAttribute(mask_rect_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem1);
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(mask_rect_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2);
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(mask_rect_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2);
    if (e != cudaSuccess) return (int)e;
    // the state byte doubles as the work list of the later passes: every mask starts undecided
    e = cudaMemsetAsync(out_has_rect, 0xff, (size_t)n_masks, st);
    if (e != cudaSuccess) return (int)e;
    mask_rect_kernel<false, false><<<(unsigned)n_masks, kT, smem1, st>>>(masks, n_masks, H, W, out_rects, out_has_rect, nullptr);
    TDOD_CHECK_LAUNCH();
    // passes 2 and 3 are persistent: a handful of CTAs walk the state bytes and almost always find nothing to do
    mask_rect_kernel<true, false><<<(unsigned)(n_masks < 148 ? n_masks : 148), kT, smem2, st>>>(masks, n_masks, H, W, out_rects,
                                                                                          out_has_rect, nullptr);
    TDOD_CHECK_LAUNCH();
    const unsigned slots = (unsigned)(n_masks < kSlots ? n_masks : kSlots);
    mask_rect_kernel<true, true><<<slots, kT, smem2, st>>>(masks, n_masks, H, W, out_rects, out_has_rect,
                                                           (unsigned char*)workspace);
    TDOD_CHECK_LAUNCH();
    return TDOD_OK;
}
