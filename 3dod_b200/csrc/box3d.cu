// IoU3D of oriented boxes: the `iou_3d` of ProposalNetwork/utils/utils.py:194-210, which hands two corner sets
// (vertex order of cubercnn/util/math_util.py:178-207) to pytorch3d.ops.box3d_overlap.  pytorch3d is a third-party
// dependency that is not under /root/reference (requirements: unpinned git install); its published contract is the
// EXACT intersection volume of two convex boxes and iou = vol / (vol1 + vol2 - vol).  The reference's own known
// answer (ProposalNetwork/utils/tests/test_iou.py:3-27 -> 0.9944) pins it.
//
// Method (not pytorch3d's triangle clipping): the intersection of two boxes is the convex polytope bounded by their
// 12 face planes.  Each of its faces is a polygon in one of those planes: a quad spanning the box face generously,
// clipped (Sutherland-Hodgman) by the four neighbouring half-spaces of its own box and the six of the OTHER box --
// so the result depends on the planes alone and stays consistent when the 8 corners are not an exact cuboid (the
// reference's known-answer vector has corners rounded to 1e-4) -- and V = 1/3 * sum_f (d_f - n_f . c) * area_f for any
// reference point c.  One lane per (pair, face): 16 lanes per
// pair (12 active), reduced with shuffles.  Coplanar faces: faces of box 1 are clipped inclusively, faces of box 2
// strictly, so a shared face is counted once; c is the mean of the clipped polygons' vertices, which lies in the
// intersection, so a flat (touching) intersection gives exactly 0.
#include "tdod_device.cuh"
#include "tdod_internal.h"

namespace {

struct V3 { float x, y, z; };
__device__ __forceinline__ V3 sub(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ float dot(V3 a, V3 b) { return fmaf(a.x, b.x, fmaf(a.y, b.y, a.z * b.z)); }
__device__ __forceinline__ V3 cross(V3 a, V3 b) {
    return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}

// quads of the six faces in the vertex order of math_util.py:178-207
__constant__ int kFace[6][4] = {{0, 1, 2, 3}, {4, 5, 6, 7}, {0, 1, 5, 4}, {1, 2, 6, 5}, {2, 3, 7, 6}, {3, 0, 4, 7}};

struct Planes { V3 n[6]; float d[6]; };   // outward unit normals, n . x <= d inside

__device__ __forceinline__ void box_planes(const V3* v, Planes& P, float& vol) {
    V3 c = {0.f, 0.f, 0.f};
#pragma unroll
    for (int k = 0; k < 8; ++k) { c.x += v[k].x; c.y += v[k].y; c.z += v[k].z; }
    c.x *= 0.125f; c.y *= 0.125f; c.z *= 0.125f;
#pragma unroll
    for (int f = 0; f < 6; ++f) {
        const V3 p0 = v[kFace[f][0]];
        V3 n = cross(sub(v[kFace[f][1]], p0), sub(v[kFace[f][3]], p0));
        const float len = sqrtf(dot(n, n));
        const float inv = len > 0.f ? 1.f / len : 0.f;
        n.x *= inv; n.y *= inv; n.z *= inv;
        float d = dot(n, p0);
        if (dot(n, c) > d) { n.x = -n.x; n.y = -n.y; n.z = -n.z; d = -d; }
        P.n[f] = n; P.d[f] = d;
    }
    const V3 e1 = sub(v[1], v[0]), e2 = sub(v[3], v[0]), e3 = sub(v[4], v[0]);
    vol = fabsf(dot(e1, cross(e2, e3)));
}

constexpr int kMaxPoly = 16;
__constant__ int kOpp[6] = {1, 0, 4, 5, 2, 3};   // the face opposite to face f

// one pair per 16 lanes: gt box g against proposal i
__global__ void __launch_bounds__(256) box3d_iou_kernel(const float* __restrict__ gt, const float* __restrict__ pr,
                                                        int64_t G, int64_t n, float* __restrict__ out_iou,
                                                        float* __restrict__ out_vol) {
    const int64_t pair = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 4;
    const int sub16 = threadIdx.x & 15;
    const bool live = pair < G * n;
    const int64_t pc = live ? pair : 0;
    const int64_t g = pc / n;
    V3 a[8], b[8];
    // translate both boxes so that box 1's first corner region is near the origin (fp32 cancellation in d - n.c)
    V3 o = {0.f, 0.f, 0.f};
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        a[k] = {gt[g * 24 + 3 * k], gt[g * 24 + 3 * k + 1], gt[g * 24 + 3 * k + 2]};
        o.x += a[k].x; o.y += a[k].y; o.z += a[k].z;
    }
    o.x *= 0.125f; o.y *= 0.125f; o.z *= 0.125f;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        a[k] = sub(a[k], o);
        b[k] = sub(V3{pr[pc * 24 + 3 * k], pr[pc * 24 + 3 * k + 1], pr[pc * 24 + 3 * k + 2]}, o);
    }
    Planes PA, PB;
    float volA, volB;
    box_planes(a, PA, volA);
    box_planes(b, PB, volB);
    const float scale = fmaxf(cbrtf(fmaxf(volA, volB)), 1e-6f);
    const float eps = 1e-5f * scale;   // coplanarity tolerance (pytorch3d uses an absolute 1e-4)

    // lane f < 6: face f of box 1 clipped inclusively by box 2; 6 <= f < 12: face f-6 of box 2 clipped strictly by box 1
    float s0 = 0.f, snx = 0.f, sny = 0.f, snz = 0.f, cx = 0.f, cy = 0.f, cz = 0.f, cn = 0.f;
    if (sub16 < 12) {
        const bool first = sub16 < 6;
        const int f = first ? sub16 : sub16 - 6;
        const V3* own = first ? a : b;
        const Planes& CP = first ? PB : PA;
        const Planes& OP = first ? PA : PB;
        const float tol2 = first ? eps : -eps;
        V3 poly[kMaxPoly], tmp[kMaxPoly];
        int m = 4;
        {   // the face's corners dropped onto its plane and pushed out to twice their distance from their mean
            const V3 nf = OP.n[f];
            V3 cen = {0.f, 0.f, 0.f};
            for (int k = 0; k < 4; ++k) {
                const V3 p = own[kFace[f][k]];
                const float h = dot(nf, p) - OP.d[f];
                poly[k] = {p.x - h * nf.x, p.y - h * nf.y, p.z - h * nf.z};
                cen.x += 0.25f * poly[k].x; cen.y += 0.25f * poly[k].y; cen.z += 0.25f * poly[k].z;
            }
            for (int k = 0; k < 4; ++k)
                poly[k] = {2.f * poly[k].x - cen.x, 2.f * poly[k].y - cen.y, 2.f * poly[k].z - cen.z};
        }
        // clips 0..5: the own box's planes (the face's own and the opposite one are skipped), 6..11: the other box's
        for (int c = 0; c < 12 && m > 0; ++c) {
            if (c < 6 && (c == f || c == kOpp[f])) continue;
            const V3 nn = c < 6 ? OP.n[c] : CP.n[c - 6];
            const float dd = c < 6 ? OP.d[c] : CP.d[c - 6];
            const float tol = c < 6 ? eps : tol2;
            int mo = 0;
            V3 prev = poly[m - 1];
            float dp = dot(nn, prev) - dd;
            for (int k = 0; k < m; ++k) {
                const V3 cur = poly[k];
                const float dc = dot(nn, cur) - dd;
                const bool in_p = dp <= tol, in_c = dc <= tol;
                if (in_p != in_c && mo < kMaxPoly) {
                    const float t = dp / (dp - dc);
                    tmp[mo++] = {fmaf(t, cur.x - prev.x, prev.x), fmaf(t, cur.y - prev.y, prev.y), fmaf(t, cur.z - prev.z, prev.z)};
                }
                if (in_c && mo < kMaxPoly) tmp[mo++] = cur;
                prev = cur; dp = dc;
            }
            m = mo;
            for (int k = 0; k < m; ++k) poly[k] = tmp[k];
        }
        if (m >= 3) {
            V3 ar = {0.f, 0.f, 0.f};
            for (int k = 1; k + 1 < m; ++k) {
                const V3 cr = cross(sub(poly[k], poly[0]), sub(poly[k + 1], poly[0]));
                ar.x += cr.x; ar.y += cr.y; ar.z += cr.z;
            }
            const float area = 0.5f * fabsf(dot(ar, OP.n[f]));
            s0 = OP.d[f] * area;
            snx = OP.n[f].x * area; sny = OP.n[f].y * area; snz = OP.n[f].z * area;
            for (int k = 0; k < m; ++k) { cx += poly[k].x; cy += poly[k].y; cz += poly[k].z; }
            cn = (float)m;
        }
    }
#pragma unroll
    for (int off = 8; off > 0; off >>= 1) {
        s0 += __shfl_xor_sync(0xffffffffu, s0, off);
        snx += __shfl_xor_sync(0xffffffffu, snx, off);
        sny += __shfl_xor_sync(0xffffffffu, sny, off);
        snz += __shfl_xor_sync(0xffffffffu, snz, off);
        cx += __shfl_xor_sync(0xffffffffu, cx, off);
        cy += __shfl_xor_sync(0xffffffffu, cy, off);
        cz += __shfl_xor_sync(0xffffffffu, cz, off);
        cn += __shfl_xor_sync(0xffffffffu, cn, off);
    }
    if (live && sub16 == 0) {
        float vol = 0.f;
        if (cn > 0.f) {
            const float ic = 1.f / cn;
            vol = (s0 - (cx * ic * snx + cy * ic * sny + cz * ic * snz)) * (1.f / 3.f);
            vol = fminf(fmaxf(vol, 0.f), fminf(volA, volB));
        }
        const float uni = volA + volB - vol;
        out_iou[pair] = uni > 0.f ? vol / uni : 0.f;
        if (out_vol) out_vol[pair] = vol;
    }
}

}  // namespace

extern "C" int tdod_box3d_iou(const float* gt_corners, const float* corners, int64_t G, int64_t n, float* out_iou,
                              float* out_vol, void* stream) {
    if (G == 0 || n == 0) return TDOD_OK;
    if (G < 0 || n < 0) return tdod_fail(TDOD_EINVAL, "G and n must be >= 0");
    if (!gt_corners || !corners || !out_iou) return tdod_fail(TDOD_EINVAL, "a required pointer is NULL");
    const int64_t lanes = G * n * 16;
    const int64_t blocks = (lanes + 255) / 256;
    if (blocks > 0x7fffffff) return tdod_fail(TDOD_ERANGE, "too many pairs for one launch");
    box3d_iou_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(gt_corners, corners, G, n, out_iou, out_vol);
    TDOD_CHECK_LAUNCH();
    return TDOD_OK;
}
