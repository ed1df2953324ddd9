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
    __shared__ double redd[32][4];
    __shared__ float rect[8];
    if (has_rect) {
        if (threadIdx.x < 8) rect[threadIdx.x] = rect_in[threadIdx.x];
    } else {
        // scorefunction.py:69-75: mean over proposals of the per-proposal corner extents
        double s[4] = {0, 0, 0, 0};
        for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
            float x0 = CUDART_INF_F, y0 = CUDART_INF_F, x1 = -CUDART_INF_F, y1 = -CUDART_INF_F;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float u = c2d[i * 16 + 2 * j], v = c2d[i * 16 + 2 * j + 1];
                x0 = min_nan(x0, u); x1 = max_nan(x1, u); y0 = min_nan(y0, v); y1 = max_nan(y1, v);
            }
            s[0] += x0; s[1] += y0; s[2] += x1; s[3] += y1;
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) s[k] = warp_sum_d(s[k]);
        if ((threadIdx.x & 31) == 0)
            for (int k = 0; k < 4; ++k) redd[threadIdx.x >> 5][k] = s[k];
        __syncthreads();
        if (threadIdx.x == 0) {
            float t[4];
            for (int k = 0; k < 4; ++k) {
                double a = 0;
                for (int w = 0; w < (int)(blockDim.x >> 5); ++w) a += redd[w][k];
                t[k] = (float)(a / (double)n);
            }
            rect[0] = t[0]; rect[1] = t[1]; rect[2] = t[2]; rect[3] = t[1];
            rect[4] = t[2]; rect[5] = t[3]; rect[6] = t[0]; rect[7] = t[3];
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
// fit); the yaw angles and their cos/sin are fp32 (torch.linspace(0, pi, 36) is fp32) and are promoted.
__constant__ uint32_t kYawAngleBits[TDOD_NYAW] = {
    0x00000000, 0x3db7d3fb, 0x3e37d3fb, 0x3e89defc, 0x3eb7d3fb, 0x3ee5c8fa, 0x3f09defc, 0x3f20d97c, 0x3f37d3fb,
    0x3f4ece7a, 0x3f65c8fa, 0x3f7cc379, 0x3f89defc, 0x3f955c3c, 0x3fa0d97c, 0x3fac56bb, 0x3fb7d3fb, 0x3fc3513b,
    0x3fcece7b, 0x3fda4bbb, 0x3fe5c8fb, 0x3ff1463a, 0x3ffcc37a, 0x4004205d, 0x4009defd, 0x400f9d9d, 0x40155c3c,
    0x401b1adc, 0x4020d97c, 0x4026981c, 0x402c56bc, 0x4032155c, 0x4037d3fb, 0x403d929b, 0x4043513b, 0x40490fdb};

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
        const T mag = sqrt(b * b + a * a);
        const float magf = fmaxf((float)mag, 1e-8f);
        v0 = b / (T)magf; v1 = -a / (T)magf; v2 = 0;
    }
    // torch.linspace(0, pi, 36) fp32 as torch-CPU 2.11 produces it (its vectorised kernel rounds two of the 36
    // entries differently from start + i*step); the angles are constants, so the bit patterns are pinned here.
    const float ang = __uint_as_float(kYawAngleBits[t]);
    const T cs = (T)cosf(ang), sn = (T)sinf(ang);
    const T omc = (T)(1.0f - cosf(ang));
    const T k0 = b * v2 - c * v1, k1 = c * v0 - a * v2, k2 = a * v1 - b * v0;   // cross(normal, v)
    const T kdv = a * v0 + b * v1 + c * v2;
    const T x0 = v0 * cs + k0 * sn + (a * kdv) * omc;
    const T x1 = v1 * cs + k1 * sn + (b * kdv) * omc;
    const T x2 = v2 * cs + k2 * sn + (c * kdv) * omc;
    const T y0 = b * x2 - c * x1, y1 = c * x0 - a * x2, y2 = a * x1 - b * x0;   // cross(normal, x)
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
