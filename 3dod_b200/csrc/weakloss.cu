// Weak-loss geometry of ROIHeads3DScore: the differentiable consumers of Cubes / cubes_to_box.
//   cube_box_fwd / _bwd : Cubes -> 8 corners -> K projection -> clamp -> box (+ GIoU loss against the 2D box), forward
//                         and the hand-derived backward; replaces the per-object Python loops of
//                         cubercnn/modeling/roi_heads/roi_heads.py:1571-1583,1599 and their autograd graphs
//   z_sweep             : z_loss's 50-step depth sweep (roi_heads.py:1151-1194), one warp per cube
//   hull / hull_fill    : segment_loss's jarvis_march + fill_polygon (ProposalNetwork/utils/utils.py:424-507,
//                         roi_heads.py:1030-1052) as per-object power sums of the soft mask, split by the
//                         ground-truth mask value, and the backward of those sums to the 8 corners; the H x W
//                         per-edge tensors of the reference (6 per edge per object, saved for autograd) never exist
#include <cstdint>

#include "tdod_device.cuh"
#include "tdod_internal.h"

using namespace tdod;

namespace {

constexpr int kT = 128;
constexpr int kMom = 12;       // per object: [x=0|1][sum t^0..t^4] then [x=0|1] sum sigmoid(t)
constexpr int kFillT = 256;

__host__ __device__ inline int64_t cdiv64(int64_t a, int64_t b) { return (a + b - 1) / b; }

// index of torch.min / torch.max along a dim: NaN wins, first occurrence on ties
__device__ __forceinline__ int arg_first(const float (&v)[8], bool want_max) {
    int idx = 0;
    float best = v[0];
#pragma unroll
    for (int j = 1; j < 8; ++j) {
        const float c = v[j];
        if (best != best) break;
        if (c != c || (want_max ? c > best : c < best)) { best = c; idx = j; }
    }
    return idx;
}

// weight with which torch.maximum(a, b) / torch.minimum(a, b) route the gradient to `a` (derivatives.yaml: ties halve)
__device__ __forceinline__ float w_max(float a, float b) { return a > b ? 1.f : (a == b ? 0.5f : 0.f); }
__device__ __forceinline__ float w_min(float a, float b) { return a < b ? 1.f : (a == b ? 0.5f : 0.f); }

// torchvision.ops.generalized_box_iou_loss(gt, box) and its derivative with respect to `box`
__device__ inline float giou_loss(const float* g, const float* b, float* d) {
    const float eps = 1e-7f;
    const float x1 = g[0], y1 = g[1], x2 = g[2], y2 = g[3], x1g = b[0], y1g = b[1], x2g = b[2], y2g = b[3];
    const float xk1 = max_nan(x1, x1g), yk1 = max_nan(y1, y1g), xk2 = min_nan(x2, x2g), yk2 = min_nan(y2, y2g);
    const bool m = (yk2 > yk1) && (xk2 > xk1);
    const float iw = xk2 - xk1, ih = yk2 - yk1;
    const float I = m ? __fmul_rn(iw, ih) : 0.f;
    const float gw = x2g - x1g, gh = y2g - y1g;
    const float U = __fsub_rn(__fadd_rn(__fmul_rn(x2 - x1, y2 - y1), __fmul_rn(gw, gh)), I);
    const float Ue = U + eps;
    const float iou = __fdiv_rn(I, Ue);
    const float xc1 = min_nan(x1, x1g), yc1 = min_nan(y1, y1g), xc2 = max_nan(x2, x2g), yc2 = max_nan(y2, y2g);
    const float cw = xc2 - xc1, ch = yc2 - yc1;
    const float C = __fmul_rn(cw, ch);
    const float Ce = C + eps;
    const float loss = 1.f - (iou - __fdiv_rn(C - U, Ce));
    if (d) {
        // dI, dA (area of `box`), dC per coordinate (x1g, y1g, x2g, y2g)
        const float dI[4] = {m ? -ih * w_max(x1g, x1) : 0.f, m ? -iw * w_max(y1g, y1) : 0.f,
                             m ? ih * w_min(x2g, x2) : 0.f, m ? iw * w_min(y2g, y2) : 0.f};
        const float dA[4] = {-gh, -gw, gh, gw};
        const float dC[4] = {-ch * w_min(x1g, x1), -cw * w_min(y1g, y1), ch * w_max(x2g, x2), cw * w_max(y2g, y2)};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float dU = dA[k] - dI[k];
            d[k] = -dI[k] / Ue + I * dU / (Ue * Ue) + (dC[k] * Ue - dU * Ce) / (Ce * Ce);
        }
    }
    return loss;
}

__global__ void __launch_bounds__(kT) cube_box_fwd_kernel(const float* __restrict__ cubes, const float* __restrict__ K,
                                                          const int32_t* __restrict__ im, const float* __restrict__ gt,
                                                          int64_t N, float* __restrict__ corners, float* __restrict__ seg,
                                                          float* __restrict__ boxes, int32_t* __restrict__ arg,
                                                          float* __restrict__ giou, float* __restrict__ dgiou) {
    const int64_t i = (int64_t)blockIdx.x * kT + threadIdx.x;
    if (i >= N) return;
    float q[15], k[9];
#pragma unroll
    for (int t = 0; t < 15; ++t) q[t] = cubes[i * 15 + t];
#pragma unroll
    for (int t = 0; t < 9; ++t) k[t] = K[i * 9 + t];
    const int W = im[i * 2], H = im[i * 2 + 1];
    const Clamp cl = make_clamp(W, H);
    float X[8], Y[8], Z[8], u[8], v[8];
    cube_corners(q, X, Y, Z);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        project(k, cl, X[j], Y[j], Z[j], u[j], v[j]);
        corners[i * 16 + 2 * j] = u[j];
        corners[i * 16 + 2 * j + 1] = v[j];
        if (seg) {                                                   // roi_heads.py:1574-1575
            seg[i * 16 + 2 * j] = clamp_nan(u[j], 0.f, (float)(W - 1));
            seg[i * 16 + 2 * j + 1] = clamp_nan(v[j], 0.f, (float)(H - 1));
        }
    }
    const int a0 = arg_first(u, false), a1 = arg_first(v, false), a2 = arg_first(u, true), a3 = arg_first(v, true);
    const float b[4] = {u[a0], v[a1], u[a2], v[a3]};
#pragma unroll
    for (int t = 0; t < 4; ++t) boxes[i * 4 + t] = b[t];
    arg[i * 4] = a0; arg[i * 4 + 1] = a1; arg[i * 4 + 2] = a2; arg[i * 4 + 3] = a3;
    if (gt) {
        float d[4];
        giou[i] = giou_loss(gt + i * 4, b, d);
#pragma unroll
        for (int t = 0; t < 4; ++t) dgiou[i * 4 + t] = d[t];
    }
}

// backward of the chain above for one object per thread: everything is recomputed from the 15 cube parameters
__global__ void __launch_bounds__(kT) cube_box_bwd_kernel(const float* __restrict__ cubes, const float* __restrict__ K,
                                                          const int32_t* __restrict__ im, const int32_t* __restrict__ arg,
                                                          int64_t N, const float* __restrict__ g_corners,
                                                          const float* __restrict__ g_seg, const float* __restrict__ g_boxes,
                                                          const float* __restrict__ g_giou, const float* __restrict__ dgiou,
                                                          float* __restrict__ g_cubes) {
    const int64_t i = (int64_t)blockIdx.x * kT + threadIdx.x;
    if (i >= N) return;
    float q[15], k[9];
#pragma unroll
    for (int t = 0; t < 15; ++t) q[t] = cubes[i * 15 + t];
#pragma unroll
    for (int t = 0; t < 9; ++t) k[t] = K[i * 9 + t];
    const int W = im[i * 2], H = im[i * 2 + 1];
    const Clamp cl = make_clamp(W, H);
    float gb[4] = {0.f, 0.f, 0.f, 0.f};
    if (g_boxes)
#pragma unroll
        for (int t = 0; t < 4; ++t) gb[t] = g_boxes[i * 4 + t];
    if (g_giou)
#pragma unroll
        for (int t = 0; t < 4; ++t) gb[t] = fmaf(g_giou[i], dgiou[i * 4 + t], gb[t]);
    const int a0 = arg[i * 4], a1 = arg[i * 4 + 1], a2 = arg[i * 4 + 2], a3 = arg[i * 4 + 3];
    float X[8], Y[8], Z[8];
    cube_corners(q, X, Y, Z);
    float gq[15];
#pragma unroll
    for (int t = 0; t < 15; ++t) gq[t] = 0.f;
    const float hl = q[5] * 0.5f, hh = q[4] * 0.5f, hw = q[3] * 0.5f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const float px = fmaf(k[2], Z[j], fmaf(k[1], Y[j], k[0] * X[j]));
        const float py = fmaf(k[5], Z[j], fmaf(k[4], Y[j], k[3] * X[j]));
        const float pz = fmaf(k[8], Z[j], fmaf(k[7], Y[j], k[6] * X[j]));
        const float ur = __fdiv_rn(px, pz), vr = __fdiv_rn(py, pz);
        const float uc = clamp_nan(ur, cl.xlo, cl.xhi), vc = clamp_nan(vr, cl.ylo, cl.yhi);
        float gu = (j == a0 ? gb[0] : 0.f) + (j == a2 ? gb[2] : 0.f);
        float gv = (j == a1 ? gb[1] : 0.f) + (j == a3 ? gb[3] : 0.f);
        if (g_corners) { gu += g_corners[i * 16 + 2 * j]; gv += g_corners[i * 16 + 2 * j + 1]; }
        if (g_seg) {                                         // second clamp: passes where 0 <= uc <= W-1 (inclusive)
            if (uc >= 0.f && uc <= (float)(W - 1)) gu += g_seg[i * 16 + 2 * j];
            if (vc >= 0.f && vc <= (float)(H - 1)) gv += g_seg[i * 16 + 2 * j + 1];
        }
        if (!(ur >= cl.xlo && ur <= cl.xhi)) gu = 0.f;       // torch.clamp backward mask (NaN -> 0)
        if (!(vr >= cl.ylo && vr <= cl.yhi)) gv = 0.f;
        if (gu == 0.f && gv == 0.f) continue;
        // u = px / pz: d/dpx = 1/pz, d/dpz = -(px/pz)/pz
        const float gpx = gu / pz, gpy = gv / pz, gpz = -(gu * ur + gv * vr) / pz;
        const float gX = k[0] * gpx + k[3] * gpy + k[6] * gpz;
        const float gY = k[1] * gpx + k[4] * gpy + k[7] * gpz;
        const float gZ = k[2] * gpx + k[5] * gpy + k[8] * gpz;
        const float sx = ((j & 3) == 0 || (j & 3) == 3) ? -hl : hl;
        const float sy = ((j & 3) < 2) ? -hh : hh;
        const float sz = (j < 4) ? -hw : hw;
        gq[0] += gX; gq[1] += gY; gq[2] += gZ;
        gq[6] += gX * sx; gq[7] += gX * sy; gq[8] += gX * sz;
        gq[9] += gY * sx; gq[10] += gY * sy; gq[11] += gY * sz;
        gq[12] += gZ * sx; gq[13] += gZ * sy; gq[14] += gZ * sz;
        // local = (+-l/2, +-h/2, +-w/2): d local / d (w,h,l) = sign / 2
        const float gsx = q[6] * gX + q[9] * gY + q[12] * gZ;
        const float gsy = q[7] * gX + q[10] * gY + q[13] * gZ;
        const float gsz = q[8] * gX + q[11] * gY + q[14] * gZ;
        gq[5] += (((j & 3) == 0 || (j & 3) == 3) ? -0.5f : 0.5f) * gsx;
        gq[4] += (((j & 3) < 2) ? -0.5f : 0.5f) * gsy;
        gq[3] += ((j < 4) ? -0.5f : 0.5f) * gsz;
    }
#pragma unroll
    for (int t = 0; t < 15; ++t) g_cubes[i * 15 + t] = gq[t];
}

// z_loss (roi_heads.py:1151-1194): one warp per cube, the depth steps spread over the lanes
__global__ void __launch_bounds__(kT) z_sweep_kernel(const float* __restrict__ gt, const float* __restrict__ cubes,
                                                     const float* __restrict__ K, const int32_t* __restrict__ im,
                                                     const float* __restrict__ proj, const float* __restrict__ values,
                                                     int steps, int64_t N, float* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * (kT / 32) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (i >= N) return;
    const float* g = gt + i * 4;
    const float* p = proj + i * 4;
    const float gt_area = __fmul_rn(g[2] - g[0], g[3] - g[1]);
    const float pred_area = __fmul_rn(p[2] - p[0], p[3] - p[1]);
    const float cx = __fdiv_rn(p[0] + p[2], 2.f), cy = __fdiv_rn(p[1] + p[3], 2.f);
    const float mc = (float)steps;
    // :1166-1167 as written: (lo <= c) is a bool, compared as 0/1 with the upper bound
    const bool within = ((g[0] - mc <= cx ? 1.f : 0.f) <= g[2] + mc) && ((g[1] - mc <= cy ? 1.f : 0.f) <= g[3] + mc);
    if (!within) {
        if (lane == 0) out[i] = __fdiv_rn(0.1f * mc, 2.f);
        return;
    }
    float q[15], k[9];
#pragma unroll
    for (int t = 0; t < 15; ++t) q[t] = cubes[i * 15 + t];
#pragma unroll
    for (int t = 0; t < 9; ++t) k[t] = K[i * 9 + t];
    const Clamp cl = make_clamp(im[i * 2], im[i * 2 + 1]);
    const bool smaller = gt_area < pred_area;
    const float z = q[2];
    float best = 0.f, bestz = 0.f;
    int bidx = 0x7fffffff;
    bool bnan = false;
    for (int s = lane; s < steps; s += 32) {
        const float zm = smaller ? z + values[s] : z - values[s];
        q[2] = zm;
        float X[8], Y[8], Z[8], u[8], v[8];
        cube_corners(q, X, Y, Z);
#pragma unroll
        for (int j = 0; j < 8; ++j) project(k, cl, X[j], Y[j], Z[j], u[j], v[j]);
        const float x0 = u[arg_first(u, false)], y0 = v[arg_first(v, false)], x1 = u[arg_first(u, true)],
                    y1 = v[arg_first(v, true)];
        float a = __fmul_rn(x1 - x0, y1 - y0);
        if (a == 0.f) a += 10000000.f;
        const float d = fabsf(gt_area - a);
        const bool dn = d != d;
        // torch.argmin: NaN is the smallest, first index on ties; this lane's steps come in ascending order
        const bool better = (bidx == 0x7fffffff) || (!bnan && (dn || d < best));
        if (better) { best = d; bestz = zm; bidx = s; bnan = dn; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ob = __shfl_xor_sync(0xffffffffu, best, o), oz = __shfl_xor_sync(0xffffffffu, bestz, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bidx, o);
        const bool on = __shfl_xor_sync(0xffffffffu, (int)bnan, o) != 0;
        if (oi == 0x7fffffff) continue;
        bool take;
        if (bidx == 0x7fffffff) take = true;
        else if (on != bnan) take = on;
        else if (on) take = oi < bidx;
        else take = (ob < best) || (ob == best && oi < bidx);
        if (take) { best = ob; bestz = oz; bidx = oi; bnan = on; }
    }
    if (lane == 0) out[i] = __fdiv_rn(fabsf(z - bestz), 2.f);
}

// ---- convex hull of the 8 projected corners: jarvis_march as the reference wrote it (utils.py:424-468) -------------
// meta[i] = {m, corner index of hull vertex 0..7}; hull [n,8,2] holds the (shifted) vertices in the reference's order
__device__ __forceinline__ float turn(float ax, float ay, float bx, float by, float cx, float cy) {
    return __fsub_rn(__fmul_rn(bx - ax, cy - ay), __fmul_rn(by - ay, cx - ax));
}
__device__ __forceinline__ float dist2(float ax, float ay, float bx, float by) {
    const float dx = bx - ax, dy = by - ay;
    return __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
}

__global__ void __launch_bounds__(kT) hull_kernel(const float* __restrict__ pts, int64_t n, float* __restrict__ hull,
                                                  int32_t* __restrict__ meta) {
    const int64_t i = (int64_t)blockIdx.x * kT + threadIdx.x;
    if (i >= n) return;
    float px[8], py[8];
    for (int j = 0; j < 8; ++j) { px[j] = pts[i * 16 + 2 * j]; py[j] = pts[i * 16 + 2 * j + 1]; }
    // findDuplicates (:397-422): every index that has an equal point later on is shifted by its rank + 1 in x and y
    unsigned dup = 0;
    for (int a = 0; a < 7; ++a)
        for (int b = a + 1; b < 8; ++b)
            if (px[a] == px[b] && py[a] == py[b]) dup |= 1u << a;
    for (int a = 0, r = 0; a < 8; ++a)
        if (dup >> a & 1) { ++r; px[a] += (float)r; py[a] += (float)r; }
    // lowest x; among several the highest y (first of equal maxima)
    float mx = px[0];
    for (int j = 1; j < 8; ++j) mx = min_nan(mx, px[j]);
    int index = -1, ncand = 0;
    float besty = 0.f;
    for (int j = 0; j < 8; ++j)
        if (px[j] == mx) {
            if (ncand == 0 || py[j] > besty) { index = j; besty = py[j]; }
            ++ncand;
        }
    if (index < 0) index = 0;                         // NaN coordinates: no candidate (the reference raises)
    int order[8], cnt = 0, l = index;
    order[cnt++] = index;
    while (true) {
        int q = (l + 1) & 7;
        for (int j = 0; j < 8; ++j) {
            if (j == l) continue;
            const float d = turn(px[l], py[l], px[j], py[j], px[q], py[q]);
            if (d > 0.f || (d == 0.f && dist2(px[j], py[j], px[l], py[l]) > dist2(px[q], py[q], px[l], py[l]))) q = j;
        }
        l = q;
        if (l == index || cnt >= 8) break;            // cnt >= 8 cannot happen for a closed march (the reference would not end)
        order[cnt++] = q;
    }
    meta[i * 9] = cnt;
    for (int t = 0; t < 8; ++t) {                     // torch.flip(..., [0])
        const int c = t < cnt ? order[cnt - 1 - t] : 0;
        meta[i * 9 + 1 + t] = t < cnt ? c : -1;
        hull[i * 16 + 2 * t] = t < cnt ? px[c] : 0.f;
        hull[i * 16 + 2 * t + 1] = t < cnt ? py[c] : 0.f;
    }
}

struct Edges {
    float v1x[8], v1y[8], ex[8], ey[8];
    float del[8];              // bound on |fp32 edge function - exact edge function| anywhere in the image
    int m;
    int bx0, bx1, by0, by1;    // pixels outside this box have a zero soft mask (see load_edges); empty when bx1 < bx0
};

// The soft mask is the product of clamp(e_i, 0, 1); a pixel where it (or any product leaving one edge out times the
// indicator 0 <= e_i <= 1) is non-zero lies in the closed intersection of the left half-planes of a closed chain of
// edges, hence inside the convex hull of the vertices, hence inside their bounding box: everything else is skipped.
__device__ __forceinline__ void load_edges(const float* __restrict__ hull, const int32_t* __restrict__ meta, int64_t i,
                                           int H, int W, Edges& E) {
    E.m = meta[i * 9];
    float lox = CUDART_INF_F, loy = CUDART_INF_F, hix = -CUDART_INF_F, hiy = -CUDART_INF_F;
    bool bad = false;
    for (int t = 0; t < 8; ++t) {
        const int t2 = (t + 1 < E.m) ? t + 1 : 0;
        const float ax = hull[i * 16 + 2 * t], ay = hull[i * 16 + 2 * t + 1];
        E.v1x[t] = ax; E.v1y[t] = ay;
        E.ex[t] = hull[i * 16 + 2 * t2] - ax;
        E.ey[t] = hull[i * 16 + 2 * t2 + 1] - ay;
        // three roundings of relative size 2^-24 on products no larger than these: 1.8e-7 * B; 1e-6 * B is used
        E.del[t] = 1e-6f * (((float)W + fabsf(ax)) * fabsf(E.ey[t]) + ((float)H + fabsf(ay)) * fabsf(E.ex[t])) + 1e-30f;
        if (t < E.m) {
            bad |= !(fabsf(ax) < 1e9f) || !(fabsf(ay) < 1e9f);
            lox = fminf(lox, ax); hix = fmaxf(hix, ax); loy = fminf(loy, ay); hiy = fmaxf(hiy, ay);
        }
    }
    if (bad) {                 // NaN / huge vertices: no shortcut, every pixel is evaluated
        E.bx0 = 0; E.bx1 = W - 1; E.by0 = 0; E.by1 = H - 1;
        for (int t = 0; t < 8; ++t) E.del[t] = CUDART_INF_F;
    } else {
        E.bx0 = max(0, (int)floorf(lox)); E.bx1 = min(W - 1, (int)ceilf(hix));
        E.by0 = max(0, (int)floorf(loy)); E.by1 = min(H - 1, (int)ceilf(hiy));
        if (E.m < 3) E.bx1 = E.bx0 - 1;     // one or two vertices: e and -e (or e = 0) -> the product is zero everywhere
    }
}

// fill_polygon (:470-502) at one pixel: edge functions with every product and difference rounded as torch does
__device__ __forceinline__ float edge_fn(const Edges& E, int t, float X, float Y) {
    return __fsub_rn(__fmul_rn(X - E.v1x[t], E.ey[t]), __fmul_rn(Y - E.v1y[t], E.ex[t]));
}
__device__ __forceinline__ float clamp01(float e) { return min_nan(max_nan(e, 0.f), 1.f); }

// Hierarchical evaluation: the bounding box is cut into 32 x 8 pixel tiles, one warp per tile.  An edge function is
// affine, so over a tile it lies between its values at the four corners; lane i < m evaluates edge i there.
//   some edge below -2 del at all corners  -> its fp32 value is negative at every pixel: the product is exactly 0
//   every edge above 1 + 2 del everywhere  -> every factor clamps to exactly 1: the product is exactly 1 (no gradient)
// and only the remaining tiles (those the hull's boundary passes through) are evaluated pixel by pixel.
constexpr int kTileW = 32, kTileH = 8;
enum { TILE_OUT = 0, TILE_IN = 1, TILE_MIXED = 2 };
__device__ __forceinline__ int classify_tile(const Edges& E, int m, int lane, int x0, int y0, int x1, int y1) {
    bool out_e = false, in_e = false;
    if (lane < m) {
        const float a = edge_fn(E, lane, (float)x0, (float)y0), b = edge_fn(E, lane, (float)x1, (float)y0);
        const float c = edge_fn(E, lane, (float)x0, (float)y1), d = edge_fn(E, lane, (float)x1, (float)y1);
        const float lo = fminf(fminf(a, b), fminf(c, d)), hi = fmaxf(fmaxf(a, b), fmaxf(c, d));
        const float dl = E.del[lane];
        out_e = hi < -2.f * dl;                      // comparisons with NaN / inf are false: such tiles stay MIXED
        in_e = lo > 1.f + 2.f * dl;                  // strictly above 1: also no tie of min(e, 1) in such a tile
    }
    if (__ballot_sync(0xffffffffu, out_e)) return TILE_OUT;
    return __ballot_sync(0xffffffffu, in_e) == ((1u << m) - 1u) ? TILE_IN : TILE_MIXED;
}

// number of non-zero bytes in [p, p + n): word loads where the address allows
__device__ __forceinline__ int count_nonzero(const uint8_t* __restrict__ p, int64_t n, int tid, int nthreads) {
    int c = 0;
    const int64_t head = min(n, (int64_t)((4 - ((uintptr_t)p & 3)) & 3));
    for (int64_t k = tid; k < head; k += nthreads) c += p[k] != 0;
    const uint32_t* w = (const uint32_t*)(p + head);
    const int64_t nw = (n - head) >> 2;
    for (int64_t k = tid; k < nw; k += nthreads) {
        const uint32_t t = w[k];
        c += __popc((t | ((t & 0x7f7f7f7fu) + 0x7f7f7f7fu)) & 0x80808080u);
    }
    for (int64_t k = head + (nw << 2) + tid; k < n; k += nthreads) c += p[k] != 0;
    return c;
}

// forward: per object the sums over pixels of t^k (k = 0..4) and sigmoid(t), split by mask value x.  gridDim.y CTAs
// per object: each counts the set pixels of one slice of the whole mask (the t = 0 terms: t^0 and sigmoid(0) = 1/2)
// and evaluates the soft mask on one band of rows of the hull's bounding box.
template <bool SIG>
__global__ void __launch_bounds__(kFillT) hull_moments_kernel(const uint8_t* __restrict__ masks,
                                                              const int32_t* __restrict__ mask_idx,
                                                              const float* __restrict__ hull,
                                                              const int32_t* __restrict__ meta, int H, int W,
                                                              double* __restrict__ mom) {
    __shared__ Edges E;
    __shared__ double s_red[kFillT / 32][kMom];
    const int64_t obj = blockIdx.x;
    if (threadIdx.x == 0) load_edges(hull, meta, obj, H, W, E);
    __syncthreads();
    const uint8_t* mk = masks + (int64_t)mask_idx[obj] * H * W;
    const int64_t HW = (int64_t)H * W;
    const int64_t slice = cdiv64(HW, gridDim.y);
    const int64_t s0 = min(HW, slice * blockIdx.y), s1 = min(HW, s0 + slice);
    const int cnt_one = count_nonzero(mk + s0, s1 - s0, threadIdx.x, kFillT);
    // sums over the pixels with t > 0: all of them, and those with mask == 1 (registers only)
    double all[5] = {0.0, 0.0, 0.0, 0.0, 0.0}, one[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
    int pos_all = 0, pos_one = 0, in_all = 0, in_one = 0;       // in_*: pixels of tiles where t is exactly 1
    const int m = E.m;
    const int bw = E.bx1 - E.bx0 + 1, bh = E.by1 - E.by0 + 1;
    const int lane = threadIdx.x & 31;
    const int ntx = (bw + kTileW - 1) / kTileW, nty = (bh + kTileH - 1) / kTileH;
    const int ntiles = (bw > 0 && bh > 0) ? ntx * nty : 0;
    for (int tile = blockIdx.y * (kFillT / 32) + (threadIdx.x >> 5); tile < ntiles; tile += gridDim.y * (kFillT / 32)) {
        // columns rotated by the tile row: with a fixed stride a warp would own whole tile columns, and the warps owning
        // the columns of the hull's left / right boundary would do all the per-pixel work of the CTA
        const int ty = tile / ntx, tx = (tile - ty * ntx + 5 * ty) % ntx;
        const int x0 = E.bx0 + tx * kTileW, y0 = E.by0 + ty * kTileH;
        const int x1 = min(x0 + kTileW - 1, E.bx1), y1 = min(y0 + kTileH - 1, E.by1);
        const int cls = classify_tile(E, m, lane, x0, y0, x1, y1);
        if (cls == TILE_OUT) continue;
        const int x = x0 + lane;
        if (x > x1) continue;
        const float X = (float)x;
        for (int y = y0; y <= y1; ++y) {
            if (cls == TILE_IN) {
                ++in_all;
                in_one += mk[(int64_t)y * W + x] != 0;
                continue;
            }
            const float Y = (float)y;
            float t = 1.f;
            for (int e = 0; e < m; ++e) {
                t = __fmul_rn(t, clamp01(edge_fn(E, e, X, Y)));
                if (t == 0.f) break;
            }
            if (t == 0.f) continue;
            const int b = mk[(int64_t)y * W + x] != 0;
            ++pos_all;
            pos_one += b;
            const double td = (double)t, t2 = td * td, bd = (double)b;
            const double p[5] = {td, t2, t2 * td, t2 * t2, SIG ? (double)__fdiv_rn(1.f, 1.f + expf(-t)) : 0.0};
#pragma unroll
            for (int q = 0; q < 5; ++q) { all[q] += p[q]; one[q] += bd * p[q]; }
        }
    }
    {   // t == 1 pixels: every power is 1
        const double sg1 = SIG ? (double)__fdiv_rn(1.f, 1.f + expf(-1.f)) : 0.0;
        const double na = (double)in_all, no = (double)in_one;
#pragma unroll
        for (int q = 0; q < 4; ++q) { all[q] += na; one[q] += no; }
        all[4] += sg1 * na; one[4] += sg1 * no;
        pos_all += in_all; pos_one += in_one;
    }
    // pixels with t == 0 contribute t^0 = 1 and sigmoid(0) = 1/2: they come from the slice counts
    const double c1 = (double)cnt_one, slice_px = threadIdx.x == 0 ? (double)(s1 - s0) : 0.0;
    double v[kMom];
    v[0] = slice_px - c1;
    v[5] = c1;
#pragma unroll
    for (int k = 1; k < 5; ++k) { v[k] = all[k - 1] - one[k - 1]; v[5 + k] = one[k - 1]; }
    v[10] = (all[4] - one[4]) + 0.5 * (slice_px - c1) - 0.5 * (double)(pos_all - pos_one);
    v[11] = one[4] + 0.5 * c1 - 0.5 * (double)pos_one;
#pragma unroll
    for (int k = 0; k < kMom; ++k) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
        if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5][k] = v[k];
    }
    __syncthreads();
    if (threadIdx.x < kMom) {
        double sum = 0.0;
        for (int w = 0; w < kFillT / 32; ++w) sum += s_red[w][threadIdx.x];
        atomicAdd(mom + obj * kMom + threadIdx.x, sum);
    }
}

// convex_hull(mask, coords) itself (utils.py:504-507): the dense soft mask, for callers that want the image
__global__ void __launch_bounds__(kFillT) hull_fill_kernel(const float* __restrict__ hull, const int32_t* __restrict__ meta,
                                                           int H, int W, int rows, float* __restrict__ out) {
    __shared__ Edges E;
    const int64_t obj = blockIdx.x;
    if (threadIdx.x == 0) load_edges(hull, meta, obj, H, W, E);
    __syncthreads();
    const int r0 = blockIdx.y * rows, r1 = min(H, r0 + rows);
    const int m = E.m;
    for (int y = r0; y < r1; ++y) {
        const float Y = (float)y;
        const bool row_in = y >= E.by0 && y <= E.by1;
        for (int x = threadIdx.x; x < W; x += kFillT) {
            float t = (m == 0) ? 1.f : 0.f;
            if (row_in && x >= E.bx0 && x <= E.bx1) {
                t = 1.f;
                for (int e = 0; e < m; ++e) t = __fmul_rn(t, clamp01(edge_fn(E, e, (float)x, Y)));
            }
            out[(obj * H + y) * W + x] = t;
        }
    }
}

// backward: d(sum_k G[x][k] S[x][k]) / d(edge function) -- or, DENSE, of sum_pixels g[pixel] t[pixel] -- accumulated
// per edge as (sum g, sum g X, sum g Y)
template <bool DENSE>
__global__ void __launch_bounds__(kFillT) hull_grad_kernel(const uint8_t* __restrict__ masks,
                                                           const int32_t* __restrict__ mask_idx,
                                                           const float* __restrict__ hull, const int32_t* __restrict__ meta,
                                                           int H, int W, const double* __restrict__ gmom,
                                                           const float* __restrict__ gdense, double* __restrict__ edge_acc) {
    __shared__ Edges E;
    __shared__ double s_acc[kFillT / 32][8][3];     // one set of accumulators per warp: no contention between warps
    __shared__ double s_g[kMom];
    const int64_t obj = blockIdx.x;
    if (threadIdx.x == 0) load_edges(hull, meta, obj, H, W, E);
    for (int k = threadIdx.x; k < (kFillT / 32) * 24; k += kFillT) (&s_acc[0][0][0])[k] = 0.0;
    if (!DENSE && threadIdx.x >= 32 && threadIdx.x < 32 + kMom) s_g[threadIdx.x - 32] = gmom[obj * kMom + threadIdx.x - 32];
    __syncthreads();
    const uint8_t* mk = DENSE ? nullptr : masks + (int64_t)mask_idx[obj] * H * W;
    const int m = E.m;
    const int bw = E.bx1 - E.bx0 + 1, bh = E.by1 - E.by0 + 1;
    const int lane = threadIdx.x & 31;
    const int ntx = (bw + kTileW - 1) / kTileW, nty = (bh + kTileH - 1) / kTileH;
    const int ntiles = (bw > 0 && bh > 0) ? ntx * nty : 0;
    for (int tile = blockIdx.y * (kFillT / 32) + (threadIdx.x >> 5); tile < ntiles; tile += gridDim.y * (kFillT / 32)) {
        // columns rotated by the tile row: with a fixed stride a warp would own whole tile columns, and the warps owning
        // the columns of the hull's left / right boundary would do all the per-pixel work of the CTA
        const int ty = tile / ntx, tx = (tile - ty * ntx + 5 * ty) % ntx;
        const int x0 = E.bx0 + tx * kTileW, y0 = E.by0 + ty * kTileH;
        const int x1 = min(x0 + kTileW - 1, E.bx1), y1 = min(y0 + kTileH - 1, E.by1);
        // a tile with t == 0 or with every edge function above 1 has no pixel with 0 <= e_i <= 1: no gradient
        if (classify_tile(E, m, lane, x0, y0, x1, y1) != TILE_MIXED) continue;
        const int x = x0 + lane;
        const bool live = x <= x1;
        const float X = (float)x;
        for (int y = y0; y <= y1; ++y) {               // warp-uniform: contributions are reduced over the warp
            const float Y = (float)y;
            float c[8], w[8];
            bool ok = live;
            if (ok) {
                int zeros = 0;
                bool inband = false;
                for (int e = 0; e < m; ++e) {
                    const float f = edge_fn(E, e, X, Y);
                    // a strictly negative edge function zeroes its own weight and every product leaving another edge out
                    if (f < 0.f) { zeros = 2; break; }
                    c[e] = clamp01(f);
                    // max(f, 0) then min(., 1): ties route half of the gradient (derivatives.yaml maximum / minimum)
                    w[e] = (f > 0.f && f < 1.f) ? 1.f : ((f == 0.f || f == 1.f) ? 0.5f : 0.f);
                    zeros += c[e] == 0.f;
                    inband |= w[e] != 0.f;
                    if (zeros >= 2) break;
                }
                ok = zeros < 2 && inband;
            }
            double dLdt = 0.0;
            if (ok) {
                if (DENSE) {
                    dLdt = (double)gdense[(obj * H + y) * W + x];
                } else {
                    float t = 1.f;
                    for (int e = 0; e < m; ++e) t = __fmul_rn(t, c[e]);
                    const int b = mk[(int64_t)y * W + x] != 0;
                    const double td = (double)t;
                    const double sg = s_g[10 + b] != 0.0 ? (double)__fdiv_rn(1.f, 1.f + expf(-t)) : 0.0;
                    const double* G = s_g + b * 5;
                    dLdt = G[1] + td * (2.0 * G[2] + td * (3.0 * G[3] + td * 4.0 * G[4])) + s_g[10 + b] * sg * (1.0 - sg);
                }
                ok = dLdt != 0.0;
            }
            if (!__any_sync(0xffffffffu, ok)) continue;
            for (int e = 0; e < m; ++e) {
                double g = 0.0;
                if (ok && w[e] != 0.f) {
                    float P = 1.f;
                    for (int o = 0; o < m; ++o)
                        if (o != e) P *= c[o];
                    g = dLdt * (double)P * (double)w[e];
                }
                if (!__any_sync(0xffffffffu, g != 0.0)) continue;
                double g0 = g, g1 = g * (double)X, g2 = g * (double)Y;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    g0 += __shfl_xor_sync(0xffffffffu, g0, o);
                    g1 += __shfl_xor_sync(0xffffffffu, g1, o);
                    g2 += __shfl_xor_sync(0xffffffffu, g2, o);
                }
                if (lane == 0) {                       // this warp's own accumulators
                    double* acc = s_acc[threadIdx.x >> 5][e];
                    acc[0] += g0; acc[1] += g1; acc[2] += g2;
                }
            }
        }
    }
    __syncthreads();
    if (threadIdx.x < 24) {
        double v = 0.0;
        for (int w = 0; w < kFillT / 32; ++w) v += (&s_acc[w][0][0])[threadIdx.x];
        if (v != 0.0) atomicAdd(edge_acc + obj * 24 + threadIdx.x, v);
    }
}

// edge sums -> hull vertices -> the 8 corners (the shift added to duplicated corners is a constant)
__global__ void __launch_bounds__(kT) hull_grad_finish_kernel(const float* __restrict__ hull, const int32_t* __restrict__ meta,
                                                              const double* __restrict__ edge_acc, int64_t n,
                                                              float* __restrict__ g_pts) {
    const int64_t i = (int64_t)blockIdx.x * kT + threadIdx.x;
    if (i >= n) return;
    const int m = meta[i * 9];
    double gx[8], gy[8];
    for (int t = 0; t < 8; ++t) gx[t] = gy[t] = 0.0;
    for (int t = 0; t < m; ++t) {
        const int t2 = (t + 1 < m) ? t + 1 : 0;
        const double A = edge_acc[i * 24 + t * 3], B = edge_acc[i * 24 + t * 3 + 1], C = edge_acc[i * 24 + t * 3 + 2];
        const double v1x = hull[i * 16 + 2 * t], v1y = hull[i * 16 + 2 * t + 1];
        const double v2x = hull[i * 16 + 2 * t2], v2y = hull[i * 16 + 2 * t2 + 1];
        gx[t] += -(v2y - v1y) * A + C - v1y * A;
        gy[t] += -B + v2x * A;
        gx[t2] += -C + v1y * A;
        gy[t2] += B - v1x * A;
    }
    for (int t = 0; t < 16; ++t) g_pts[i * 16 + t] = 0.f;
    for (int t = 0; t < m; ++t) {
        const int c = meta[i * 9 + 1 + t];
        g_pts[i * 16 + 2 * c] = (float)gx[t];
        g_pts[i * 16 + 2 * c + 1] = (float)gy[t];
    }
}

inline unsigned blocks_for(int64_t n, int t) { return (unsigned)cdiv64(n, t); }

// rows of the image per CTA of the fill kernels: about 4 CTAs per SM over the whole launch, at least 8 rows each
inline int fill_rows(int64_t n, int H) {
    const int64_t want = cdiv64((int64_t)148 * 4, n > 0 ? n : 1);
    int64_t rows = cdiv64(H, want);
    if (rows < 8) rows = 8;
    if (rows > H) rows = H;
    return (int)rows;
}

}  // namespace

extern "C" int tdod_cube_box_fwd(const float* cubes, const float* K, const int32_t* im_wh, const float* gt_boxes,
                                 int64_t N, float* out_corners, float* out_seg_corners, float* out_boxes,
                                 int32_t* out_arg, float* out_giou, float* out_dgiou, void* stream) {
    if (N == 0) return TDOD_OK;
    if (!cubes || !K || !im_wh || !out_corners || !out_boxes || !out_arg || N < 0)
        return tdod_fail(TDOD_EINVAL, "tdod_cube_box_fwd: bad argument");
    if (gt_boxes && (!out_giou || !out_dgiou))
        return tdod_fail(TDOD_EINVAL, "tdod_cube_box_fwd: gt_boxes given without out_giou / out_dgiou");
    cube_box_fwd_kernel<<<blocks_for(N, kT), kT, 0, (cudaStream_t)stream>>>(cubes, K, im_wh, gt_boxes, N, out_corners,
                                                                            out_seg_corners, out_boxes, out_arg,
                                                                            out_giou, out_dgiou);
    TDOD_CHECK_LAUNCH();
    return TDOD_OK;
}

extern "C" int tdod_cube_box_bwd(const float* cubes, const float* K, const int32_t* im_wh, const int32_t* arg, int64_t N,
                                 const float* g_corners, const float* g_seg_corners, const float* g_boxes,
                                 const float* g_giou, const float* dgiou, float* out_g_cubes, void* stream) {
    if (N == 0) return TDOD_OK;
    if (!cubes || !K || !im_wh || !arg || !out_g_cubes || N < 0)
        return tdod_fail(TDOD_EINVAL, "tdod_cube_box_bwd: bad argument");
    if (g_giou && !dgiou) return tdod_fail(TDOD_EINVAL, "tdod_cube_box_bwd: g_giou given without dgiou");
    cube_box_bwd_kernel<<<blocks_for(N, kT), kT, 0, (cudaStream_t)stream>>>(cubes, K, im_wh, arg, N, g_corners,
                                                                            g_seg_corners, g_boxes, g_giou, dgiou,
                                                                            out_g_cubes);
    TDOD_CHECK_LAUNCH();
    return TDOD_OK;
}

extern "C" int tdod_z_sweep(const float* gt_boxes, const float* cubes, const float* K, const int32_t* im_wh,
                            const float* proj_boxes, const float* values, int steps, int64_t N, float* out, void* stream) {
    if (N == 0) return TDOD_OK;
    if (!gt_boxes || !cubes || !K || !im_wh || !proj_boxes || !values || !out || N < 0 || steps <= 0)
        return tdod_fail(TDOD_EINVAL, "tdod_z_sweep: bad argument");
    z_sweep_kernel<<<blocks_for(N, kT / 32), kT, 0, (cudaStream_t)stream>>>(gt_boxes, cubes, K, im_wh, proj_boxes, values,
                                                                            steps, N, out);
    TDOD_CHECK_LAUNCH();
    return TDOD_OK;
}

extern "C" int tdod_hull(const float* corners2d, int64_t n, float* out_hull, int32_t* out_meta, void* stream) {
    if (n == 0) return TDOD_OK;
    if (!corners2d || !out_hull || !out_meta || n < 0) return tdod_fail(TDOD_EINVAL, "tdod_hull: bad argument");
    hull_kernel<<<blocks_for(n, kT), kT, 0, (cudaStream_t)stream>>>(corners2d, n, out_hull, out_meta);
    TDOD_CHECK_LAUNCH();
    return TDOD_OK;
}

extern "C" int tdod_hull_fill_moments(const uint8_t* masks, const int32_t* mask_idx, const float* hull,
                                      const int32_t* meta, int64_t n, int H, int W, int with_sigmoid,
                                      double* out_moments, void* stream) {
    if (n == 0) return TDOD_OK;
    if (!masks || !mask_idx || !hull || !meta || !out_moments || n < 0 || H <= 0 || W <= 0)
        return tdod_fail(TDOD_EINVAL, "tdod_hull_fill_moments: bad argument");
    cudaError_t e = cudaMemsetAsync(out_moments, 0, sizeof(double) * kMom * n, (cudaStream_t)stream);
    if (e != cudaSuccess) return (int)e;
    const int rows = fill_rows(n, H);
    dim3 grid((unsigned)n, (unsigned)cdiv64(H, rows));     // objects along x: no 65535 limit
    if (with_sigmoid)
        hull_moments_kernel<true><<<grid, kFillT, 0, (cudaStream_t)stream>>>(masks, mask_idx, hull, meta, H, W, out_moments);
    else
        hull_moments_kernel<false><<<grid, kFillT, 0, (cudaStream_t)stream>>>(masks, mask_idx, hull, meta, H, W, out_moments);
    TDOD_CHECK_LAUNCH();
    return TDOD_OK;
}

extern "C" int tdod_hull_fill_grad(const uint8_t* masks, const int32_t* mask_idx, const float* hull, const int32_t* meta,
                                   int64_t n, int H, int W, const double* g_moments, double* workspace,
                                   float* out_g_corners, void* stream) {
    if (n == 0) return TDOD_OK;
    if (!masks || !mask_idx || !hull || !meta || !g_moments || !workspace || !out_g_corners || n < 0 || H <= 0 ||
        W <= 0)
        return tdod_fail(TDOD_EINVAL, "tdod_hull_fill_grad: bad argument");
    cudaError_t e = cudaMemsetAsync(workspace, 0, sizeof(double) * 24 * n, (cudaStream_t)stream);
    if (e != cudaSuccess) return (int)e;
    const int rows = fill_rows(n, H);
    dim3 grid((unsigned)n, (unsigned)cdiv64(H, rows));     // objects along x: no 65535 limit
    hull_grad_kernel<false><<<grid, kFillT, 0, (cudaStream_t)stream>>>(masks, mask_idx, hull, meta, H, W, g_moments,
                                                                      nullptr, workspace);
    TDOD_CHECK_LAUNCH();
    hull_grad_finish_kernel<<<blocks_for(n, kT), kT, 0, (cudaStream_t)stream>>>(hull, meta, workspace, n, out_g_corners);
    TDOD_CHECK_LAUNCH();
    return TDOD_OK;
}

extern "C" int tdod_hull_fill(const float* hull, const int32_t* meta, int64_t n, int H, int W, float* out_mask,
                              void* stream) {
    if (n == 0) return TDOD_OK;
    if (!hull || !meta || !out_mask || n < 0 || H <= 0 || W <= 0)
        return tdod_fail(TDOD_EINVAL, "tdod_hull_fill: bad argument");
    const int rows = fill_rows(n, H);
    dim3 grid((unsigned)n, (unsigned)cdiv64(H, rows));     // objects along x: no 65535 limit
    hull_fill_kernel<<<grid, kFillT, 0, (cudaStream_t)stream>>>(hull, meta, H, W, rows, out_mask);
    TDOD_CHECK_LAUNCH();
    return TDOD_OK;
}

extern "C" int tdod_hull_fill_dense_grad(const float* hull, const int32_t* meta, int64_t n, int H, int W,
                                         const float* g_mask, double* workspace, float* out_g_corners, void* stream) {
    if (n == 0) return TDOD_OK;
    if (!hull || !meta || !g_mask || !workspace || !out_g_corners || n < 0 || H <= 0 || W <= 0)
        return tdod_fail(TDOD_EINVAL, "tdod_hull_fill_dense_grad: bad argument");
    cudaError_t e = cudaMemsetAsync(workspace, 0, sizeof(double) * 24 * n, (cudaStream_t)stream);
    if (e != cudaSuccess) return (int)e;
    const int rows = fill_rows(n, H);
    dim3 grid((unsigned)n, (unsigned)cdiv64(H, rows));     // objects along x: no 65535 limit
    hull_grad_kernel<true><<<grid, kFillT, 0, (cudaStream_t)stream>>>(nullptr, nullptr, hull, meta, H, W, nullptr,
                                                                     g_mask, workspace);
    TDOD_CHECK_LAUNCH();
    hull_grad_finish_kernel<<<blocks_for(n, kT), kT, 0, (cudaStream_t)stream>>>(hull, meta, workspace, n, out_g_corners);
    TDOD_CHECK_LAUNCH();
    return TDOD_OK;
}
