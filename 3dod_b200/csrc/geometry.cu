// Per-function entry points behind the reference's individual Python functions (drop-in granularity).
// The hot loop of a real run goes through fused.cu; these exist so every reference call site keeps working
// one call at a time, and so each stage can be checked against the oracle in isolation.
#include "tdod_device.cuh"
#include "tdod_internal.h"

using namespace tdod;

namespace {

constexpr int kT = 256;

__global__ void corners3d_kernel(const float* __restrict__ cubes, int64_t M, float* __restrict__ out) {
    const int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= M) return;
    float q[15];
#pragma unroll
    for (int k = 0; k < 15; ++k) q[k] = cubes[m * 15 + k];
    float X[8], Y[8], Z[8];
    cube_corners(q, X, Y, Z);
    float* o = out + m * 24;
#pragma unroll
    for (int j = 0; j < 8; ++j) { o[3 * j] = X[j]; o[3 * j + 1] = Y[j]; o[3 * j + 2] = Z[j]; }
}

__global__ void project_kernel(const float* __restrict__ cubes, int64_t M, const float* __restrict__ Kp, Clamp cl,
                               float* __restrict__ c2d, float* __restrict__ boxes) {
    __shared__ float K[9];
    if (threadIdx.x < 9) K[threadIdx.x] = Kp[threadIdx.x];
    __syncthreads();
    const int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= M) return;
    float q[15];
#pragma unroll
    for (int k = 0; k < 15; ++k) q[k] = cubes[m * 15 + k];
    float X[8], Y[8], Z[8];
    cube_corners(q, X, Y, Z);
    float x0 = CUDART_INF_F, y0 = CUDART_INF_F, x1 = -CUDART_INF_F, y1 = -CUDART_INF_F;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        float u, v;
        project(K, cl, X[j], Y[j], Z[j], u, v);
        if (c2d) { c2d[m * 16 + 2 * j] = u; c2d[m * 16 + 2 * j + 1] = v; }
        x0 = min_nan(x0, u); x1 = max_nan(x1, u); y0 = min_nan(y0, v); y1 = max_nan(y1, v);
    }
    if (boxes) *reinterpret_cast<float4*>(boxes + m * 4) = make_float4(x0, y0, x1, y1);
}

// one K per cube, picked through cube_frame (the MABO branch projects every object's GT cube with its frame's intrinsics)
__global__ void project_indexed_kernel(const float* __restrict__ cubes, int64_t M, const float* __restrict__ Kall,
                                       const int32_t* __restrict__ cube_frame, Clamp cl, float* __restrict__ c2d,
                                       float* __restrict__ boxes) {
    const int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= M) return;
    float K[9];
    const float* Kp = Kall + (size_t)cube_frame[m] * 9;
#pragma unroll
    for (int k = 0; k < 9; ++k) K[k] = __ldg(Kp + k);
    float q[15];
#pragma unroll
    for (int k = 0; k < 15; ++k) q[k] = cubes[m * 15 + k];
    float X[8], Y[8], Z[8];
    cube_corners(q, X, Y, Z);
    float x0 = CUDART_INF_F, y0 = CUDART_INF_F, x1 = -CUDART_INF_F, y1 = -CUDART_INF_F;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        float u, v;
        project(K, cl, X[j], Y[j], Z[j], u, v);
        if (c2d) { c2d[m * 16 + 2 * j] = u; c2d[m * 16 + 2 * j + 1] = v; }
        x0 = min_nan(x0, u); x1 = max_nan(x1, u); y0 = min_nan(y0, v); y1 = max_nan(y1, v);
    }
    if (boxes) *reinterpret_cast<float4*>(boxes + m * 4) = make_float4(x0, y0, x1, y1);
}

__global__ void iou_kernel(const float* __restrict__ ref, const float* __restrict__ boxes, int64_t n,
                           float* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 b = *reinterpret_cast<const float4*>(boxes + i * 4);
    out[i] = iou_xyxy(ref[0], ref[1], ref[2], ref[3], b.x, b.y, b.z, b.w);
}

// block-wide NaN-propagating max (torch.max)
__device__ float block_max_nan(float v, float* red) {
    v = warp_max_nan(v);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    float m = red[0];
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w) {
        const float t = red[w];
        m = (m != m || t != t) ? CUDART_NAN_F : fmaxf(m, t);
    }
    __syncthreads();
    return m;
}

// one CTA: pass 1 writes |r_ref - r_i| into out and reduces the max, pass 2 rescales
__global__ void __launch_bounds__(1024) dims_kernel(const float* __restrict__ mu, const float* __restrict__ sg,
                                                    const float* __restrict__ dims, int64_t stride,
                                                    const float* __restrict__ ref, const float* __restrict__ boxes,
                                                    int64_t n, float* __restrict__ out) {
    __shared__ float red[32];
    const float r_ref = __fdiv_rn(__fsub_rn(ref[2], ref[0]), __fsub_rn(ref[3], ref[1]));
    float lmax = -CUDART_INF_F;
    for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
        const float4 b = *reinterpret_cast<const float4*>(boxes + i * 4);
        const float r = __fdiv_rn(__fsub_rn(b.z, b.x), __fsub_rn(b.w, b.y));
        const float diff = fabsf(__fsub_rn(r_ref, r));
        out[i] = diff;
        lmax = (lmax != lmax || diff != diff) ? CUDART_NAN_F : fmaxf(lmax, diff);
    }
    const float mx = block_max_nan(lmax, red);
    const float m3[3] = {mu[0], mu[1], mu[2]}, s3[3] = {sg[0], sg[1], sg[2]};
    for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
        const float* d = dims + i * stride;
        const float g = dims_gauss(d[0], d[1], d[2], m3, s3);
        out[i] = __fmul_rn(__fsub_rn(1.f, __fdiv_rn(out[i], mx)), g);
    }
}

__global__ void __launch_bounds__(1024) corners_kernel(const float* __restrict__ rect_in, int has_rect,
                                                       const float* __restrict__ c2d, int64_t n,
                                                       float* __restrict__ out) {
    __shared__ float red[32];
    __shared__ float redf[32][8];
    __shared__ float rect[8];
    if (has_rect) {
        if (threadIdx.x < 8) rect[threadIdx.x] = rect_in[threadIdx.x];
    } else {
        // scorefunction.py:69-75 as the callers' [1,n,8,2] corners make it behave: x0/x1 = mean of (min / max over
        // PROPOSALS of vertex 0's x, of vertex 0's y); y0/y1 likewise from vertex 1 (see oracle fallback_rect)
        float mn[4] = {CUDART_INF_F, CUDART_INF_F, CUDART_INF_F, CUDART_INF_F};
        float mx4[4] = {-CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F};
        for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float v = c2d[i * 16 + k];          // v0x, v0y, v1x, v1y
                mn[k] = min_nan(mn[k], v); mx4[k] = max_nan(mx4[k], v);
            }
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) { mn[k] = -warp_max_nan(-mn[k]); mx4[k] = warp_max_nan(mx4[k]); }
        if ((threadIdx.x & 31) == 0)
            for (int k = 0; k < 4; ++k) { redf[threadIdx.x >> 5][k] = mn[k]; redf[threadIdx.x >> 5][4 + k] = mx4[k]; }
        __syncthreads();
        if (threadIdx.x == 0) {
            float t[8];
            for (int k = 0; k < 8; ++k) {
                float m = (k < 4) ? CUDART_INF_F : -CUDART_INF_F;
                for (int w = 0; w < (int)(blockDim.x >> 5); ++w) m = (k < 4) ? min_nan(m, redf[w][k]) : max_nan(m, redf[w][k]);
                t[k] = m;
            }
            const float x0 = __fdiv_rn(__fadd_rn(t[0], t[1]), 2.f), x1 = __fdiv_rn(__fadd_rn(t[4], t[5]), 2.f);
            const float y0 = __fdiv_rn(__fadd_rn(t[2], t[3]), 2.f), y1 = __fdiv_rn(__fadd_rn(t[6], t[7]), 2.f);
            rect[0] = x0; rect[1] = y0; rect[2] = x1; rect[3] = y0;
            rect[4] = x1; rect[5] = y1; rect[6] = x0; rect[7] = y1;
        }
    }
    __syncthreads();
    float lmax = -CUDART_INF_F;
    for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
        float U[8], V[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) { U[j] = c2d[i * 16 + 2 * j]; V[j] = c2d[i * 16 + 2 * j + 1]; }
        const float d = chamfer4x8(rect, U, V);
        out[i] = d;
        lmax = (lmax != lmax || d != d) ? CUDART_NAN_F : fmaxf(lmax, d);
    }
    const float mx = block_max_nan(lmax, red);
    for (int64_t i = threadIdx.x; i < n; i += blockDim.x) out[i] = __fsub_rn(1.f, __fdiv_rn(out[i], mx));
}

// ProposalNetwork/utils/utils.py:112-146 in the dtype of the normal (fp64 when it comes from the NumPy plane
// fit); the yaw angles torch.linspace(0, pi, 36) and their cos/sin are fp32 and are promoted.
// cos / sin / (1 - cos) of the 36 angles as torch-CPU 2.11 evaluates them in fp32 (utils.py:138-139): the reference's
// yaw table is built on the host, so its trigonometry is the CPU's, not libdevice's; pinning the 36 values makes
// the table bit-comparable with the reference.
__constant__ uint32_t kYawCosBits[TDOD_NYAW] = {
    0x3f800000, 0x3f7ef82c, 0x3f7be2cf, 0x3f76c645, 0x3f6fad17, 0x3f66a5e5, 0x3f5bc34d, 0x3f4f1bbd, 0x3f40c94b,
    0x3f30e97c, 0x3f1f9d07, 0x3f0d0796, 0x3ef29eeb, 0x3ec93a94, 0x3e9e3779, 0x3e63dc8a, 0x3e097473, 0x3d37c428,
    0xbd37c440, 0xbe097479, 0xbe63dc90, 0xbe9e377b, 0xbec93a97, 0xbef29eee, 0xbf0d0798, 0xbf1f9d0a, 0xbf30e97b,
    0xbf40c94b, 0xbf4f1bbd, 0xbf5bc34e, 0xbf66a5e6, 0xbf6fad18, 0xbf76c645, 0xbf7be2cf, 0xbf7ef82c, 0xbf800000};
__constant__ uint32_t kYawSinBits[TDOD_NYAW] = {
    0x00000000, 0x3db794d0, 0x3e36d79e, 0x3e8835fb, 0x3eb3e767, 0x3ede2603, 0x3f034d5e, 0x3f167918, 0x3f286eac,
    0x3f390915, 0x3f48261c, 0x3f55a698, 0x3f616eb5, 0x3f6b662c, 0x3f737871, 0x3f7994e0, 0x3f7daee2, 0x3f7fbe02,
    0x3f7fbe02, 0x3f7daee2, 0x3f7994e0, 0x3f737870, 0x3f6b662c, 0x3f616eb4, 0x3f55a696, 0x3f48261a, 0x3f390916,
    0x3f286eac, 0x3f167917, 0x3f034d5c, 0x3ede25fe, 0x3eb3e761, 0x3e8835fc, 0x3e36d79d, 0x3db794ca, 0xb3bbbd2e};
__constant__ uint32_t kYawOmcBits[TDOD_NYAW] = {
    0x00000000, 0x3b83ea00, 0x3c83a620, 0x3d139bb0, 0x3d829748, 0x3dcad0d8, 0x3e10f2cc, 0x3e43910c, 0x3e7cdad4,
    0x3e9e2d08, 0x3ec0c5f2, 0x3ee5f0d4, 0x3f06b08a, 0x3f1b62b6, 0x3f30e444, 0x3f4708de, 0x3f5da2e3, 0x3f7483be,
    0x3f85be22, 0x3f912e8f, 0x3f9c7b92, 0x3fa78ddf, 0x3fb24ea6, 0x3fbca7bc, 0x3fc683cc, 0x3fcfce85, 0x3fd874be,
    0x3fe064a6, 0x3fe78dde, 0x3fede1a7, 0x3ff352f3, 0x3ff7d68c, 0x3ffb6322, 0x3ffdf168, 0x3fff7c16, 0x40000000};

// separately rounded products / sums (each is its own torch op in utils.py:134-146)
__device__ __forceinline__ float mulr(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ double mulr(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ float addr(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ double addr(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ float subr(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ double subr(double a, double b) { return __dsub_rn(a, b); }

template <typename T>
__global__ void yaw_table_kernel(const T* __restrict__ normals, int F, float* __restrict__ out) {
    const int f = blockIdx.x, t = threadIdx.x;
    if (f >= F || t >= TDOD_NYAW) return;
    const T a = normals[f * 3], b = normals[f * 3 + 1], c = normals[f * 3 + 2];
    T v0, v1, v2;
    if (a == (T)0) {                 // vec_perp_t: exact compare, not re-normalised
        v0 = 0; v1 = c; v2 = -b;
    } else {
        // normalize_vector: the magnitude passes through an fp32 tensor (torch.max with tensor([1e-8]))
        const T mag = sqrt(addr(mulr(b, b), mulr(a, a)));
        const float magf = fmaxf((float)mag, 1e-8f);
        v0 = b / (T)magf; v1 = -a / (T)magf; v2 = 0;
    }
    const T cs = (T)__uint_as_float(kYawCosBits[t]), sn = (T)__uint_as_float(kYawSinBits[t]);
    const T omc = (T)__uint_as_float(kYawOmcBits[t]);
    // cross(normal, v), normal . v
    const T k0 = subr(mulr(b, v2), mulr(c, v1)), k1 = subr(mulr(c, v0), mulr(a, v2)), k2 = subr(mulr(a, v1), mulr(b, v0));
    const T kdv = addr(addr(mulr(a, v0), mulr(b, v1)), mulr(c, v2));
    // Rodrigues (utils.py:141-146): (v cos + (k x v) sin) + k (k.v) (1 - cos)
    const T x0 = addr(addr(mulr(v0, cs), mulr(k0, sn)), mulr(mulr(a, kdv), omc));
    const T x1 = addr(addr(mulr(v1, cs), mulr(k1, sn)), mulr(mulr(b, kdv), omc));
    const T x2 = addr(addr(mulr(v2, cs), mulr(k2, sn)), mulr(mulr(c, kdv), omc));
    // cross(normal, x)
    const T y0 = subr(mulr(b, x2), mulr(c, x1)), y1 = subr(mulr(c, x0), mulr(a, x2)), y2 = subr(mulr(a, x1), mulr(b, x0));
    float* o = out + ((size_t)f * TDOD_NYAW + t) * 9;   // rows of [x | n | y] columns
    o[0] = (float)x0; o[1] = (float)a; o[2] = (float)y0;
    o[3] = (float)x1; o[4] = (float)b; o[5] = (float)y1;
    o[6] = (float)x2; o[7] = (float)c; o[8] = (float)y2;
}

inline unsigned blocks_for(int64_t n, int t) { return (unsigned)((n + t - 1) / t); }

}  // namespace

extern "C" int tdod_cube_corners(const float* cubes, int64_t M, float* out, void* stream) {
    if (M == 0) return TDOD_OK;
    if (!cubes || !out || M < 0) return tdod_fail(TDOD_EINVAL, "tdod_cube_corners: bad argument");
    corners3d_kernel<<<blocks_for(M, kT), kT, 0, (cudaStream_t)stream>>>(cubes, M, out);
    TDOD_CHECK_LAUNCH();
    return TDOD_OK;
}

extern "C" int tdod_project_cubes(const float* cubes, int64_t M, const float* K, int clamp_w, int clamp_h,
                                  float* out_corners2d, float* out_boxes, void* stream) {
    if (M == 0) return TDOD_OK;
    if (!cubes || !K || M < 0 || (!out_corners2d && !out_boxes))
        return tdod_fail(TDOD_EINVAL, "tdod_project_cubes: bad argument");
    if (clamp_w <= 0 || clamp_h <= 0)
        return tdod_fail(TDOD_EINVAL, "tdod_project_cubes: clamp (image width, height) is mandatory (spaces.py:240-243)");
    project_kernel<<<blocks_for(M, kT), kT, 0, (cudaStream_t)stream>>>(cubes, M, K, make_clamp(clamp_w, clamp_h),
                                                                      out_corners2d, out_boxes);
    TDOD_CHECK_LAUNCH();
    return TDOD_OK;
}

extern "C" int tdod_project_cubes_indexed(const float* cubes, int64_t M, const float* K, const int32_t* cube_frame, int clamp_w,
                                          int clamp_h, float* out_corners2d, float* out_boxes, void* stream) {
    if (M == 0) return TDOD_OK;
    if (!cubes || !K || !cube_frame || M < 0 || (!out_corners2d && !out_boxes))
        return tdod_fail(TDOD_EINVAL, "tdod_project_cubes_indexed: bad argument");
    if (clamp_w <= 0 || clamp_h <= 0)
        return tdod_fail(TDOD_EINVAL, "tdod_project_cubes_indexed: clamp (image width, height) is mandatory (spaces.py:240-243)");
    project_indexed_kernel<<<blocks_for(M, kT), kT, 0, (cudaStream_t)stream>>>(cubes, M, K, cube_frame, make_clamp(clamp_w, clamp_h),
                                                                              out_corners2d, out_boxes);
    TDOD_CHECK_LAUNCH();
    return TDOD_OK;
}

extern "C" int tdod_score_iou(const float* ref_box, const float* boxes, int64_t n, float* out, void* stream) {
    if (n == 0) return TDOD_OK;
    if (!ref_box || !boxes || !out || n < 0) return tdod_fail(TDOD_EINVAL, "tdod_score_iou: bad argument");
    iou_kernel<<<blocks_for(n, kT), kT, 0, (cudaStream_t)stream>>>(ref_box, boxes, n, out);
    TDOD_CHECK_LAUNCH();
    return TDOD_OK;
}

extern "C" int tdod_score_dimensions(const float* prior_mean, const float* prior_std, const float* dims,
                                     int64_t dims_stride, const float* ref_box, const float* boxes, int64_t n,
                                     float* out, void* stream) {
    if (n == 0) return TDOD_OK;
    if (!prior_mean || !prior_std || !dims || !ref_box || !boxes || !out || n < 0 || dims_stride < 3)
        return tdod_fail(TDOD_EINVAL, "tdod_score_dimensions: bad argument");
    dims_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(prior_mean, prior_std, dims, dims_stride, ref_box, boxes, n, out);
    TDOD_CHECK_LAUNCH();
    return TDOD_OK;
}

extern "C" int tdod_score_corners(const float* rect, int has_rect, const float* corners2d, int64_t n, float* out,
                                  void* stream) {
    if (n == 0) return TDOD_OK;
    if ((has_rect && !rect) || !corners2d || !out || n < 0)
        return tdod_fail(TDOD_EINVAL, "tdod_score_corners: bad argument");
    corners_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(rect, has_rect, corners2d, n, out);
    TDOD_CHECK_LAUNCH();
    return TDOD_OK;
}

extern "C" int tdod_yaw_table(const void* normals, int normals_f64, int F, float* out_table, void* stream) {
    if (F == 0) return TDOD_OK;
    if (!normals || !out_table || F < 0) return tdod_fail(TDOD_EINVAL, "tdod_yaw_table: bad argument");
    if (normals_f64)
        yaw_table_kernel<double><<<F, 64, 0, (cudaStream_t)stream>>>((const double*)normals, F, out_table);
    else
        yaw_table_kernel<float><<<F, 64, 0, (cudaStream_t)stream>>>((const float*)normals, F, out_table);
    TDOD_CHECK_LAUNCH();
    return TDOD_OK;
}

// ---- depth / ground map ingest: PIL NEAREST resize on the device (cubercnn/data/dataset_mapper.py:125-147) --------------
// Pillow's nearest resize steps through the source with a double-precision accumulator (xx = 0.5*s; xx += s); the caller
// builds the two index tables that way on the host (tdod/ingest.py) and the kernel is a pure gather: every output
// element written once, coalesced; the source rows it touches are read through L1/L2.
namespace {
template <typename E>
__global__ void __launch_bounds__(256) resize_nearest_kernel(const E* __restrict__ src, int h, int w, const int32_t* __restrict__ xmap,
                                                             const int32_t* __restrict__ ymap, E* __restrict__ dst, int H, int W) {
    const int f = blockIdx.z, y = blockIdx.y;
    const E* s = src + ((size_t)f * h + ymap[y]) * w;
    E* d = dst + ((size_t)f * H + y) * W;
    for (int x = blockIdx.x * blockDim.x + threadIdx.x; x < W; x += gridDim.x * blockDim.x) d[x] = __ldg(s + xmap[x]);
}
}  // namespace

extern "C" int tdod_resize_nearest(const void* src, int elem_bytes, int F, int h, int w, const int32_t* xmap, const int32_t* ymap,
                                   void* dst, int H, int W, void* stream) {
    if (F == 0) return TDOD_OK;
    if (!src || !dst || !xmap || !ymap || F < 0 || h <= 0 || w <= 0 || H <= 0 || W <= 0)
        return tdod_fail(TDOD_EINVAL, "tdod_resize_nearest: bad argument");
    if (H > 65535 || F > 65535) return tdod_fail(TDOD_ERANGE, "tdod_resize_nearest: H and F are grid dimensions (<= 65535)");
    const dim3 grid((unsigned)((W + 255) / 256), (unsigned)H, (unsigned)F);
    cudaStream_t st = (cudaStream_t)stream;
    if (elem_bytes == 4)
        resize_nearest_kernel<uint32_t><<<grid, 256, 0, st>>>((const uint32_t*)src, h, w, xmap, ymap, (uint32_t*)dst, H, W);
    else if (elem_bytes == 1)
        resize_nearest_kernel<uint8_t><<<grid, 256, 0, st>>>((const uint8_t*)src, h, w, xmap, ymap, (uint8_t*)dst, H, W);
    else
        return tdod_fail(TDOD_EINVAL, "tdod_resize_nearest: elem_bytes must be 1 or 4");
    TDOD_CHECK_LAUNCH();
    return TDOD_OK;
}

// ---- torch.std with torch's CUDA rounding (tdod_device.cuh, namespace tstd), one CTA per row ----------------------
namespace {
__global__ void __launch_bounds__(256) torch_std_kernel(const float* __restrict__ x, int N, int n, float* __restrict__ out) {
    __shared__ tstd::W scr[512];
    const int row = blockIdx.x;
    const tstd::Cfg cfg = tstd::config(n, N);
    const float* p = x + (size_t)row * n;
    const int shift = (int)(((long long)row * n) & 1);
    const float sd = tstd::block_std([&](int i) { return p[i]; }, n, shift, cfg, scr, threadIdx.x, 256);
    if (threadIdx.x == 0) out[row] = sd;
}
}  // namespace

extern "C" int tdod_torch_std(const float* x, int64_t N, int n, float* out, void* stream) {
    if (N == 0) return TDOD_OK;
    if (N < 0 || N > 0x7fffffff || n <= 0) return tdod_fail(TDOD_EINVAL, "tdod_torch_std: bad shape");
    if (!x || !out) return tdod_fail(TDOD_EINVAL, "tdod_torch_std: a pointer is NULL");
    if (!tstd::config(n, (int)N).ok) return tdod_fail(TDOD_ERANGE, "tdod_torch_std: torch splits this reduction over several CTAs");
    if ((reinterpret_cast<uintptr_t>(x) & 7) != 0) return tdod_fail(TDOD_EINVAL, "tdod_torch_std: x must be 8-byte aligned (torch tensors are)");
    torch_std_kernel<<<(unsigned)N, 256, 0, (cudaStream_t)stream>>>(x, (int)N, n, out);
    TDOD_CHECK_LAUNCH();
    return TDOD_OK;
}
