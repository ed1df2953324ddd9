// K-C: fused propose + project + score + select.  One thread-block cluster (1..8 CTAs) per object.
//
// Restates, per object, ProposalNetwork/proposals/proposals.py:338-424 (propose),
// utils/spaces.py:192-245 (corners + projection), utils/conversions.py:25-48 (box),
// scoring/scorefunction.py:47-49 (IoU), :144-160 (dims x aspect), :69-85 + :51-56 (chamfer to the mask
// rectangle), cubercnn/modeling/roi_heads/roi_heads.py:499-502 (product + np.argmax).
//
// Nothing per-proposal goes to HBM in select mode: the n proposals of an object live in shared memory /
// registers of its cluster; the three per-object reductions the reference's scores need (median+std of the
// centre seeds, max aspect difference, max chamfer distance) and the argmax run over DSMEM.
#include <cooperative_groups.h>
#include "tdod_device.cuh"
#include "tdod_internal.h"

namespace cg = cooperative_groups;
using namespace tdod;

namespace {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kMaxCluster = 8;
#ifndef TDOD_PSS_MINBLOCKS
#define TDOD_PSS_MINBLOCKS 3
#endif

struct Mail {               // per-CTA mailbox read by every CTA of the cluster through DSMEM
    int hist[4][3][256];    // radix-select histograms, one buffer per pass (no WAR hazard)
    double sum[3];          // sum of seeds
    double meanv[3];        // one CTA per object: the mean, formed once by the thread that owns the sum
    double m2[3];           // sum of squared deviations
    float fb[8];            // fallback rectangle: min/max over proposals of vertex 0 and vertex 1 (x, y)
    float maxdiff, maxd;
    float sdv[3];           // std_mode TORCH_CUDA: the three standard deviations, published by rank 0
    int nanflag[3];
    int nfail;              // Philox dimensions: proposals whose first draw fell outside the truncation bounds
    float am_v;
    int am_i;
};

struct Shared {
    Mail mail;
    // block-level scratch
    double red_d[kWarps][4];
    float red_f[kWarps][2];
    float red_fb[kWarps][8];
    int red_i[kWarps];
    uint32_t prefix[3];
    int kth[3];
    float med[3];
    float sd_loc[3];        // clusters, exact std: formed by this CTA's owner threads from all ranks' partial sums
    float rect[8];
    float K[9];
    float yaw[TDOD_NYAW * 9];
};

struct ObjCtx {
    int obj, frame;
    float bx0, by0, bx1, by1;      // sampling / aspect box
    float lox, stepx, loy, stepy;  // vectorized_linspace parameters
    float mu[3], sg[3];            // priors (w,h,l)
    float sgh;                     // std_h * 1.1
    float hi[3];                   // upper truncation per dimension
    float nx0, nsx, ny0, nsy;      // PATCH centre modes: vectorized_linspace in normalised image space
    float rh, rl;                  // DIMS_ASPECT ratios
};

// proposals.py:12-14 rescale_interval(u, MIN_PROP_S, 2) = (0.05 - 2) * u + 2, product and sum rounded separately
__device__ __forceinline__ float rescale_dim(float u) { return __fadd_rn(__fmul_rn(u, -1.95f), 2.f); }
// U[0,1) from 24 random bits
__device__ __forceinline__ float u01(uint32_t r) { return (float)(r & 0xffffffu) * 5.9604644775390625e-8f; }

// Philox mode: the standard normal behind the x centre of proposal i.  With prior dimensions it is the fourth normal of
// the first (w,h,l) try; the uniform-dimension modes spend that block on w,h,l and take it from the next block.
__device__ __forceinline__ float philox_xn(const tdod_pss_args& a, const ObjCtx& c, int i) {
    const uint64_t gid = (uint64_t)(a.obj_id_base + c.obj);
    const uint4 r = Philox(a.seed)((uint32_t)i, (uint32_t)gid, (uint32_t)(gid >> 32) ^ 0x57484Cu,
                                   a.dims_mode == TDOD_DIMS_PRIOR ? 0u : 1u);
    return box_muller(r.z, r.w).y;
}

// accepted standard normals for (w,h,l) of proposal i
// `xn`: Philox mode only -- the standard normal of the x draw (fourth normal of the first try's Philox block)
// `zn`: the three numbers behind (w,h,l) as HOST mode would be fed them (accepted standard normals, or the U[0,1) draws of
// the uniform-dimension modes) -- what out_noise records so that a PHILOX run can be replayed in HOST mode
// one round of the rejection loop of draw_whl (Philox stream, prior dimensions): redraws the components that are not ok yet
__device__ __forceinline__ void draw_whl_try(const tdod_pss_args& a, const ObjCtx& c, int i, uint32_t t, float& w, float& h, float& l,
                                             bool& okw, bool& okh, bool& okl, float& xn, float (&zn)[3]) {
    const Philox ph(a.seed);
    const uint64_t gid = (uint64_t)(a.obj_id_base + c.obj);
    const uint4 r = ph((uint32_t)i, (uint32_t)gid, (uint32_t)(gid >> 32) ^ 0x57484Cu, t);
    const float2 n01 = box_muller(r.x, r.y), n23 = box_muller(r.z, r.w);
    if (t == 0) xn = n23.y;
    if (!okw) { zn[0] = n01.x; w = scale_normal(n01.x, c.mu[0], c.sg[0]); okw = !(w < TDOD_MIN_PROP_S || w > c.hi[0]); }
    if (!okh) { zn[1] = n01.y; h = scale_normal(n01.y, c.mu[1], c.sgh);   okh = !(h < TDOD_MIN_PROP_S || h > c.hi[1]); }
    if (!okl) { zn[2] = n23.x; l = scale_normal(n23.x, c.mu[2], c.sg[2]); okl = !(l < TDOD_MIN_PROP_S || l > c.hi[2]); }
}

__device__ __forceinline__ void draw_whl(const tdod_pss_args& a, const ObjCtx& c, int i, float& w, float& h, float& l, float& xn,
                                         float (&zn)[3]) {
    xn = 0.f;
    zn[0] = zn[1] = zn[2] = 0.f;
    if (a.rng_mode == TDOD_RNG_HOST) {
        const float* nz = a.noise + (size_t)c.obj * 6 * a.n_prop;
        zn[0] = nz[i]; zn[1] = nz[a.n_prop + i]; zn[2] = nz[2 * a.n_prop + i];
        if (a.dims_mode == TDOD_DIMS_PRIOR) {
            w = scale_normal(nz[i], c.mu[0], c.sg[0]);
            h = scale_normal(nz[a.n_prop + i], c.mu[1], c.sgh);
            l = scale_normal(nz[2 * a.n_prop + i], c.mu[2], c.sg[2]);
        } else if (a.dims_mode == TDOD_DIMS_UNIFORM) {                       // proposals.py:28-30
            w = rescale_dim(nz[i]); h = rescale_dim(nz[a.n_prop + i]); l = rescale_dim(nz[2 * a.n_prop + i]);
        } else {                                                            // proposals.py:230-239
            w = rescale_dim(nz[i]); h = __fmul_rn(w, c.rh); l = __fmul_rn(w, c.rl);
        }
        return;
    }
    if (a.dims_mode != TDOD_DIMS_PRIOR) {
        const Philox ph(a.seed);
        const uint64_t gid = (uint64_t)(a.obj_id_base + c.obj);
        const uint4 r = ph((uint32_t)i, (uint32_t)gid, (uint32_t)(gid >> 32) ^ 0x57484Cu, 0u);
        w = rescale_dim(u01(r.x));
        zn[0] = u01(r.x);
        if (a.dims_mode == TDOD_DIMS_UNIFORM) { h = rescale_dim(u01(r.y)); l = rescale_dim(u01(r.z)); zn[1] = u01(r.y); zn[2] = u01(r.z); }
        else { h = __fmul_rn(w, c.rh); l = __fmul_rn(w, c.rl); }
        xn = philox_xn(a, c, i);
        return;
    }
    // utils.py:49-58: invalid entries are replaced by fresh draws until all are in [0.05, mean + k std];
    // per element that is an independent rejection loop (same distribution, Philox stream).
    bool okw = false, okh = false, okl = false;
    w = h = l = 0.f;
    for (uint32_t t = 0; t < 10001u && !(okw && okh && okl); ++t) draw_whl_try(a, c, i, t, w, h, l, okw, okh, okl, xn, zn);
}

// standard normals for x,y,z and the rotation of proposal i -> q[0..2], q[6..14]
// `rec`: out_noise / out_yaw_idx are written when true (the winner's recomputation passes false)
__device__ __forceinline__ void draw_xyz_rot(const tdod_pss_args& a, const ObjCtx& c, const float* yaw, int i,
                                             const float (&mean)[3], const float (&sd)[3], float xn, float* q, bool rec) {
    float* on = (rec && a.out_noise) ? a.out_noise + (size_t)c.obj * 6 * a.n_prop : nullptr;
    if (a.rng_mode == TDOD_RNG_HOST) {
        if (on) {
            const float* nz = a.noise + (size_t)c.obj * 6 * a.n_prop;
            on[3 * a.n_prop + i] = nz[3 * a.n_prop + i]; on[4 * a.n_prop + i] = nz[4 * a.n_prop + i];
            on[5 * a.n_prop + i] = nz[5 * a.n_prop + i];
        }
        const float* nz = a.noise + (size_t)c.obj * 6 * a.n_prop;
        if (a.centre_mode == TDOD_CENTRE_DEPTH) {
            q[0] = scale_normal(nz[3 * a.n_prop + i], mean[0], sd[0]);
            q[1] = scale_normal(nz[4 * a.n_prop + i], mean[1], sd[1]);
            q[2] = scale_normal(nz[5 * a.n_prop + i], mean[2], sd[2]);
        } else if (a.centre_mode == TDOD_CENTRE_UNIFORM) {                   // proposals.py:23-25
            q[0] = __fsub_rn(__fmul_rn(nz[3 * a.n_prop + i], 4.f), 2.f);
            q[1] = __fsub_rn(__fmul_rn(nz[4 * a.n_prop + i], 2.f), 1.f);
            q[2] = __fadd_rn(__fmul_rn(nz[5 * a.n_prop + i], 4.f), 1.f);
        } else {                                                            // proposals.py:59-62 / :106-115
            q[0] = __fadd_rn(__fmul_rn((float)i, c.nsx), c.nx0);
            q[1] = __fadd_rn(__fmul_rn((float)i, c.nsy), c.ny0);
            const float zc = nz[5 * a.n_prop + i];
            q[2] = (a.centre_mode == TDOD_CENTRE_PATCH) ? __fadd_rn(__fmul_rn(zc, 4.f), 1.f) : zc;
        }
        if (a.yaw_table) {
            const int yi = a.yaw_idx[(size_t)c.obj * a.n_prop + i];
            if (rec && a.out_yaw_idx) a.out_yaw_idx[(size_t)c.obj * a.n_prop + i] = (int16_t)yi;
#pragma unroll
            for (int k = 0; k < 9; ++k) q[6 + k] = yaw[yi * 9 + k];
        } else {
            const float* r = a.rot + ((size_t)c.obj * a.n_prop + i) * 9;
#pragma unroll
            for (int k = 0; k < 9; ++k) q[6 + k] = r[k];
        }
        return;
    }
    const Philox ph(a.seed);
    const uint64_t gid = (uint64_t)(a.obj_id_base + c.obj);
    // one Philox block: two uniforms -> (y, z) normals, one -> yaw bin; the x normal came with the (w,h,l) block
    const uint4 r = ph((uint32_t)i, (uint32_t)gid, (uint32_t)(gid >> 32) ^ 0x58595Au, 0u);
    if (a.centre_mode == TDOD_CENTRE_DEPTH) {
        const float2 n01 = box_muller(r.x, r.y);
        q[0] = scale_normal(xn, mean[0], sd[0]);
        q[1] = scale_normal(n01.x, mean[1], sd[1]);
        q[2] = scale_normal(n01.y, mean[2], sd[2]);
        if (on) { on[3 * a.n_prop + i] = xn; on[4 * a.n_prop + i] = n01.x; on[5 * a.n_prop + i] = n01.y; }
    } else if (a.centre_mode == TDOD_CENTRE_UNIFORM) {
        q[0] = __fsub_rn(__fmul_rn(u01(r.x), 4.f), 2.f);
        q[1] = __fsub_rn(__fmul_rn(u01(r.y), 2.f), 1.f);
        q[2] = __fadd_rn(__fmul_rn(u01(r.w), 4.f), 1.f);
        if (on) { on[3 * a.n_prop + i] = u01(r.x); on[4 * a.n_prop + i] = u01(r.y); on[5 * a.n_prop + i] = u01(r.w); }
    } else {                                                                // PATCH (PATCH_Z needs the caller's ramp)
        q[0] = __fadd_rn(__fmul_rn((float)i, c.nsx), c.nx0);
        q[1] = __fadd_rn(__fmul_rn((float)i, c.nsy), c.ny0);
        q[2] = __fadd_rn(__fmul_rn(u01(r.w), 4.f), 1.f);
        if (on) { on[3 * a.n_prop + i] = 0.f; on[4 * a.n_prop + i] = 0.f; on[5 * a.n_prop + i] = u01(r.w); }
    }
    if (a.yaw_table) {
        const int yi = (int)(r.z % TDOD_NYAW);  // torch.randint(36, ...): u32 % 36
        if (rec && a.out_yaw_idx) a.out_yaw_idx[(size_t)c.obj * a.n_prop + i] = (int16_t)yi;
#pragma unroll
        for (int k = 0; k < 9; ++k) q[6 + k] = yaw[yi * 9 + k];
    } else {
        // utils.py:62-69 randn_orthobasis_torch: rows normalised, row0 = r1 x r2, row1 = r2 x row0
        float m[9];
#pragma unroll
        for (int t = 0; t < 3; ++t) {
            const uint4 s = ph((uint32_t)i, (uint32_t)gid, (uint32_t)(gid >> 32) ^ 0x524F54u, 1u + t);
            const float2 a01 = box_muller(s.x, s.y), a23 = box_muller(s.z, s.w);
            const float inv = rsqrtf(a01.x * a01.x + a01.y * a01.y + a23.x * a23.x);
            m[3 * t] = a01.x * inv; m[3 * t + 1] = a01.y * inv; m[3 * t + 2] = a23.x * inv;
        }
        float c0 = m[4] * m[8] - m[5] * m[7], c1 = m[5] * m[6] - m[3] * m[8], c2 = m[3] * m[7] - m[4] * m[6];
        float inv = rsqrtf(c0 * c0 + c1 * c1 + c2 * c2);
        m[0] = c0 * inv; m[1] = c1 * inv; m[2] = c2 * inv;
        c0 = m[7] * m[2] - m[8] * m[1]; c1 = m[8] * m[0] - m[6] * m[2]; c2 = m[6] * m[1] - m[7] * m[0];
        inv = rsqrtf(c0 * c0 + c1 * c1 + c2 * c2);
        m[3] = c0 * inv; m[4] = c1 * inv; m[5] = c2 * inv;
#pragma unroll
        for (int k = 0; k < 9; ++k) q[6 + k] = m[k];
    }
}

// The per-object reductions run over the CTAs of a cluster through DSMEM.  With one CTA per object (the common
// case: n <= 2048 proposals) the same code runs on plain CTA barriers and local pointers.
template <bool kCluster> struct ClusterCtx;
template <> struct ClusterCtx<true> {
    cg::cluster_group cl = cg::this_cluster();
    __device__ __forceinline__ int size() { return (int)cl.num_blocks(); }
    __device__ __forceinline__ int rank() { return (int)cl.block_rank(); }
    __device__ __forceinline__ void sync() { cl.sync(); }
    template <typename T> __device__ __forceinline__ T* remote(T* p, int r) { return cl.map_shared_rank(p, r); }
};
template <> struct ClusterCtx<false> {
    __device__ __forceinline__ int size() { return 1; }
    __device__ __forceinline__ int rank() { return 0; }
    __device__ __forceinline__ void sync() { __syncthreads(); }
    template <typename T> __device__ __forceinline__ T* remote(T* p, int) { return p; }
};

// kHot: the configuration of the throughput path -- in-kernel Philox, depth-seeded centres, prior dimensions, a yaw table,
// winners only.  Every mode switch of the argument block becomes a compile-time constant there, which removes the code of
// the other samplers, of the audit / return-all outputs and of the fallback sweep's bookkeeping from the instruction
// stream (ncu: the general kernel is 107 KB of SASS and stalls on instruction fetch for 1.2 warp-cycles per issue).
template <bool kCluster, bool kHot>
__global__ void __launch_bounds__(kThreads, TDOD_PSS_MINBLOCKS) pss_kernel(const tdod_pss_args a_in, const int n_loc) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    tdod_pss_args a = a_in;
    if (kHot) {
        a.rng_mode = TDOD_RNG_PHILOX; a.centre_mode = TDOD_CENTRE_DEPTH; a.dims_mode = TDOD_DIMS_PRIOR;
        a.propose_only = 0; a.std_mode = TDOD_STD_EXACT;
        a.noise = nullptr; a.yaw_idx = nullptr; a.rot = nullptr; a.aux = nullptr;
        a.all_cubes = nullptr; a.all_boxes = nullptr; a.all_scores = nullptr; a.stats = nullptr;
        a.out_noise = nullptr; a.out_yaw_idx = nullptr; a.out_seeds = nullptr;
        __builtin_assume(a.yaw_table != nullptr);
        __builtin_assume(a.obj_frame != nullptr);
    }
    ClusterCtx<kCluster> cl;
    const int S = cl.size();
    const int rank = cl.rank();
    const int obj = (int)(blockIdx.x / S);
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;

    Shared& sh = *reinterpret_cast<Shared*>(smem_raw);
    float* arr = reinterpret_cast<float*>(smem_raw + ((sizeof(Shared) + 15) & ~size_t(15)));
    // seven per-proposal arrays of n_loc floats:
    float* s_x = arr;                  // seed x        -> later: iou
    float* s_y = arr + n_loc;          // seed y        -> later: dims gauss
    float* s_z = arr + 2 * n_loc;      // seed z_tmp+l/2-> later: aspect difference
    float* s_w = arr + 3 * n_loc;
    float* s_h = arr + 4 * n_loc;
    float* s_l = arr + 5 * n_loc;
    float* s_d = arr + 6 * n_loc;      // chamfer distance

    const int n = a.n_prop;
    const int i0 = rank * n_loc;
    const int cnt = max(0, min(n - i0, n_loc));

    // ---------------- phase 0: object context ----------------
    ObjCtx c;
    c.obj = obj;
    c.frame = a.obj_frame ? a.obj_frame[obj] : 0;
    {
        const float* b = a.boxes + (size_t)obj * 4;
        c.bx0 = b[0]; c.by0 = b[1]; c.bx1 = b[2]; c.by1 = b[3];
        const float bw = __fsub_rn(c.bx1, c.bx0), bh = __fsub_rn(c.by1, c.by0);
        const float qx = __fdiv_rn(bw, 4.f), qy = __fdiv_rn(bh, 4.f);
        c.lox = __fadd_rn(c.bx0, qx);
        c.loy = __fadd_rn(c.by0, qy);
        const float hix = __fsub_rn(c.bx1, qx), hiy = __fsub_rn(c.by1, qy);
        c.stepx = __fdiv_rn(__fsub_rn(hix, c.lox), (float)(n - 1));
        c.stepy = __fdiv_rn(__fsub_rn(hiy, c.loy), (float)(n - 1));
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            c.mu[k] = a.prior_mean[(size_t)obj * 3 + k];
            c.sg[k] = a.prior_std[(size_t)obj * 3 + k];
        }
        c.sgh = __fmul_rn(c.sg[1], 1.1f);                                  // proposals.py:382
        c.hi[0] = __fadd_rn(c.mu[0], __fmul_rn(2.f, c.sg[0]));             // :381
        c.hi[1] = __fadd_rn(c.mu[1], __fmul_rn(2.2f, c.sg[1]));            // :382
        c.hi[2] = __fadd_rn(c.mu[2], __fmul_rn(2.f, c.sg[2]));             // :383
        c.nx0 = c.nsx = c.ny0 = c.nsy = 0.f;
        if (a.centre_mode >= TDOD_CENTRE_PATCH) {
            // conversions.py:50-67 pixel_to_normalised_space: (p - 0.5*old) * (new/old), python-float constants cast
            // to fp32; x uses (im_w, 3), y uses (im_h, 2) (proposals.py:56-57)
            const int iw = a.im_w > 0 ? a.im_w : a.W, ih = a.im_h > 0 ? a.im_h : a.H;
            const float hx = (float)(0.5 * (double)iw), sx = (float)(3.0 / (double)iw);
            const float hy = (float)(0.5 * (double)ih), sy = (float)(2.0 / (double)ih);
            const float x0n = __fmul_rn(__fsub_rn(c.lox, hx), sx), x1n = __fmul_rn(__fsub_rn(hix, hx), sx);
            const float y0n = __fmul_rn(__fsub_rn(c.loy, hy), sy), y1n = __fmul_rn(__fsub_rn(hiy, hy), sy);
            c.nx0 = x0n; c.nsx = __fdiv_rn(__fsub_rn(x1n, x0n), (float)(n - 1));
            c.ny0 = y0n; c.nsy = __fdiv_rn(__fsub_rn(y1n, y0n), (float)(n - 1));
        }
        c.rh = c.rl = 1.f;
        if (a.dims_mode == TDOD_DIMS_ASPECT) { c.rh = a.aux[(size_t)obj * 2]; c.rl = a.aux[(size_t)obj * 2 + 1]; }
    }
    if (tid < 9) sh.K[tid] = a.K[(size_t)c.frame * 9 + tid];
    if (a.yaw_table)
        for (int t = tid; t < TDOD_NYAW * 9; t += kThreads) sh.yaw[t] = a.yaw_table[(size_t)c.frame * TDOD_NYAW * 9 + t];
    // only code 1 is a rectangle; 0 (empty mask) and the error codes 3 / 4 of tdod_mask_rect take the fallback of :69-75
    const bool has_rect = a.propose_only ? true : (a.has_rect ? (a.has_rect[obj] == 1) : true);
    if (tid < 8) sh.rect[tid] = (has_rect && a.rects) ? a.rects[(size_t)obj * 8 + tid] : 0.f;
    for (int t = tid; t < 4 * 3 * 256; t += kThreads) (&sh.mail.hist[0][0][0])[t] = 0;
    if (tid < 3) { sh.prefix[tid] = 0u; sh.kth[tid] = (n - 1) / 2; sh.mail.nanflag[tid] = 0; }
    if (tid == 0) sh.mail.nfail = 0;
    __syncthreads();
    if (S > 1) cl.sync();   // (0) every CTA's histograms are zero before a peer may add to rank 0's
    // Clusters: each CTA counts into its own histograms and then ADDS its non-zero bins to rank 0's (DSMEM atomics), so
    // that the warps that resolve a pass read one merged histogram -- two 16-byte loads per lane -- instead of walking all
    // S mailboxes (ncu on C3, S = 8: 57 % of the kernel's stall samples sat in those 64 dependent remote loads per lane
    // and in the peers waiting for them at the cluster barrier).
    auto merge_hist = [&](int pass) {
        if (S > 1 && rank != 0) {
            const int* loc = &sh.mail.hist[pass][0][0];
            int* dst = cl.remote(&sh.mail.hist[pass][0][0], 0);
            for (int t = tid; t < 3 * 256; t += kThreads) {
                const int v = loc[t];
                if (v) atomicAdd(dst + t, v);
            }
        }
    };

    // ---------------- phase 1: centre seeds + dimensions ----------------
    const float* depth = a.depth + (size_t)c.frame * a.H * a.W;
    const float fx = sh.K[0], cx = sh.K[2], cy = sh.K[5];
    // Philox dimensions (utils.py:49-58 as an independent rejection loop per element): 13 % of the proposals need a second
    // draw, and a warp that loops until its last lane is done spends 2.2 rounds per proposal.  Round 0 runs for everybody,
    // the proposals that are not done are queued and their redraws run densely packed (one queued proposal per thread).
    const bool split_draw = a.rng_mode == TDOD_RNG_PHILOX && a.dims_mode == TDOD_DIMS_PRIOR;
    if (split_draw) {
        constexpr int kFailCap = 3 * 256;
        int* fail = &sh.mail.hist[3][0][0];            // free until the last radix pass; every used entry is zeroed again
        float* on = a.out_noise ? a.out_noise + (size_t)obj * 6 * n : nullptr;
        for (int j = tid; j < cnt; j += kThreads) {
            const int i = i0 + j;
            float w = 0.f, h = 0.f, l = 0.f, xn = 0.f, zn[3];
            bool okw = false, okh = false, okl = false;
            draw_whl_try(a, c, i, 0u, w, h, l, okw, okh, okl, xn, zn);
            if (!(okw && okh && okl)) {
                const int slot = atomicAdd(&sh.mail.nfail, 1);
                if (slot < kFailCap) fail[slot] = j | (okw ? 0x10000 : 0) | (okh ? 0x20000 : 0) | (okl ? 0x40000 : 0);
                else for (uint32_t t = 1; t < 10001u && !(okw && okh && okl); ++t) draw_whl_try(a, c, i, t, w, h, l, okw, okh, okl, xn, zn);
            }
            s_w[j] = w; s_h[j] = h; s_l[j] = l; s_d[j] = xn;
            if (on) { on[i] = zn[0]; on[n + i] = zn[1]; on[2 * n + i] = zn[2]; }
        }
        __syncthreads();
        const int nf = min(sh.mail.nfail, kFailCap);
        for (int f = tid; f < nf; f += kThreads) {
            const int e = fail[f];
            fail[f] = 0;
            const int j = e & 0xffff, i = i0 + j;
            float w = s_w[j], h = s_h[j], l = s_l[j], xn = 0.f, zn[3] = {0.f, 0.f, 0.f};
            bool okw = (e & 0x10000) != 0, okh = (e & 0x20000) != 0, okl = (e & 0x40000) != 0;
            const bool hw = !okw, hh = !okh, hl = !okl;
            for (uint32_t t = 1; t < 10001u && !(okw && okh && okl); ++t) draw_whl_try(a, c, i, t, w, h, l, okw, okh, okl, xn, zn);
            s_w[j] = w; s_h[j] = h; s_l[j] = l;
            if (on) { if (hw) on[i] = zn[0]; if (hh) on[n + i] = zn[1]; if (hl) on[2 * n + i] = zn[2]; }
        }
        __syncthreads();
    }
    double lsum[3] = {0.0, 0.0, 0.0};
    int lnan[3] = {0, 0, 0};
    for (int j = tid; j < cnt; j += kThreads) {
        const int i = i0 + j;
        int gx = linspace_px(c.lox, c.stepx, i);
        int gy = linspace_px(c.loy, c.stepy, i);
        // torch indexing: negative indices wrap once; anything else out of range would raise in the
        // reference -- we clamp instead of faulting.
        int ix = gx < 0 ? gx + a.W : gx, iy = gy < 0 ? gy + a.H : gy;
        ix = min(max(ix, 0), a.W - 1); iy = min(max(iy, 0), a.H - 1);
        float x = 0.f, y = 0.f, zt = 0.f, w, h, l, xn;
        if (a.centre_mode == TDOD_CENTRE_DEPTH) {
            const float d = __ldg(depth + (size_t)iy * a.W + ix);
            centre_seed(gx, gy, d, fx, cx, cy, x, y, zt);
        }
        if (split_draw) { w = s_w[j]; h = s_h[j]; l = s_l[j]; }
        else {
            float zn[3];
            draw_whl(a, c, i, w, h, l, xn, zn);
            s_d[j] = xn;                                                   // parked until phase 2 reads it back
            if (a.out_noise) {
                float* on = a.out_noise + (size_t)obj * 6 * n;
                on[i] = zn[0]; on[n + i] = zn[1]; on[2 * n + i] = zn[2];
            }
        }
        const float z = __fadd_rn(zt, __fdiv_rn(l, 2.f));                  // proposals.py:394
        if (a.out_seeds) {
            float* os = a.out_seeds + (size_t)obj * 3 * n;
            os[i] = x; os[n + i] = y; os[2 * n + i] = z;
        }
        s_x[j] = x; s_y[j] = y; s_z[j] = z; s_w[j] = w; s_h[j] = h; s_l[j] = l;
        lsum[0] += (double)x; lsum[1] += (double)y; lsum[2] += (double)z;
        lnan[0] |= (x != x); lnan[1] |= (y != y); lnan[2] |= (z != z);
        atomicAdd(&sh.mail.hist[0][0][f2key(x) >> 24], 1);
        atomicAdd(&sh.mail.hist[0][1][f2key(y) >> 24], 1);
        atomicAdd(&sh.mail.hist[0][2][f2key(z) >> 24], 1);
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        lsum[k] = warp_sum_d(lsum[k]);
        if (__any_sync(0xffffffffu, lnan[k]) && lane == 0) atomicOr(&sh.mail.nanflag[k], 1);
    }
    if (lane == 0) { sh.red_d[wid][0] = lsum[0]; sh.red_d[wid][1] = lsum[1]; sh.red_d[wid][2] = lsum[2]; }
    __syncthreads();
    if (tid < 3) {
        double s = 0.0;
        for (int w2 = 0; w2 < kWarps; ++w2) s += sh.red_d[w2][tid];
        sh.mail.sum[tid] = s;
        if (S == 1) sh.mail.meanv[tid] = s / (double)n;
    }
    merge_hist(0);           // (the barrier above closed this CTA's pass-0 counting)
    cl.sync();   // (1) pass-0 histograms, partial sums and NaN flags visible cluster-wide

    // ---------------- exact lower median (torch.median) by 4-pass radix select over the cluster ----------------
    for (int pass = 0; pass < 4; ++pass) {
        if (pass > 0) {
            const int shift = 24 - 8 * pass;
            const uint32_t himask = 0xffffffffu << (shift + 8);
            const uint32_t p0 = sh.prefix[0], p1 = sh.prefix[1], p2 = sh.prefix[2];
            for (int j = tid; j < cnt; j += kThreads) {
                const uint32_t kx = f2key(s_x[j]), ky = f2key(s_y[j]), kz = f2key(s_z[j]);
                if ((kx & himask) == p0) atomicAdd(&sh.mail.hist[pass][0][(kx >> shift) & 255u], 1);
                if ((ky & himask) == p1) atomicAdd(&sh.mail.hist[pass][1][(ky >> shift) & 255u], 1);
                if ((kz & himask) == p2) atomicAdd(&sh.mail.hist[pass][2][(kz >> shift) & 255u], 1);
            }
            if (S > 1) { __syncthreads(); merge_hist(pass); }
            cl.sync();   // (2..4)
        }
        if (wid < 3) {   // warp `wid` resolves array `wid`: lane owns bins [8*lane, 8*lane+8)
            int bins[8];
            {
                const int4* hr = reinterpret_cast<const int4*>(cl.remote(&sh.mail.hist[pass][wid][0], 0));   // rank 0 holds the sum
                const int4 lo4 = hr[2 * lane], hi4 = hr[2 * lane + 1];
                bins[0] = lo4.x; bins[1] = lo4.y; bins[2] = lo4.z; bins[3] = lo4.w;
                bins[4] = hi4.x; bins[5] = hi4.y; bins[6] = hi4.z; bins[7] = hi4.w;
            }
            int tot = 0;
#pragma unroll
            for (int b = 0; b < 8; ++b) tot += bins[b];
            int incl = tot;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += v;
            }
            const int excl = incl - tot;
            const int k = sh.kth[wid];
            __syncwarp();
            if (k >= excl && k < incl) {   // exactly one lane
                int run = excl, b = 0;
#pragma unroll
                for (b = 0; b < 8; ++b) {
                    if (k < run + bins[b]) break;
                    run += bins[b];
                }
                sh.prefix[wid] |= (uint32_t)(8 * lane + b) << (24 - 8 * pass);
                sh.kth[wid] = k - run;
            }
        }
        __syncthreads();
    }
    if (tid < 3) {
        int anynan = 0;
        for (int r = 0; r < S; ++r) anynan |= *cl.remote(&sh.mail.nanflag[tid], r);
        sh.med[tid] = anynan ? CUDART_NAN_F : key2f(sh.prefix[tid]);
        if (S > 1) {                                    // clusters: the three owners gather the partial sums (in rank order)
            double s = 0.0;
            for (int r = 0; r < S; ++r) s += *cl.remote(&sh.mail.sum[tid], r);
            sh.mail.meanv[tid] = s / (double)n;
        }
    }
    if (S > 1) __syncthreads();
    // mean over the cluster, then squared deviations (two-pass, fp64: torch CPU accumulates std in double)
    double mean[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        mean[k] = sh.mail.meanv[k];                     // formed once by the thread that owns the sum (same operations, same bits)
    }
    double lm2[3] = {0.0, 0.0, 0.0};
    for (int j = tid; j < cnt; j += kThreads) {
        const double dx = (double)s_x[j] - mean[0], dy = (double)s_y[j] - mean[1], dz = (double)s_z[j] - mean[2];
        lm2[0] += dx * dx; lm2[1] += dy * dy; lm2[2] += dz * dz;
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) lm2[k] = warp_sum_d(lm2[k]);
    if (lane == 0) { sh.red_d[wid][0] = lm2[0]; sh.red_d[wid][1] = lm2[1]; sh.red_d[wid][2] = lm2[2]; }
    __syncthreads();
    if (tid < 3) {
        double s = 0.0;
        for (int w2 = 0; w2 < kWarps; ++w2) s += sh.red_d[w2][tid];
        sh.mail.m2[tid] = s;
        // one CTA per object: the unbiased std is formed here, once, instead of by all 256 threads after the barrier
        if (S == 1 && a.std_mode != TDOD_STD_TORCH_CUDA) sh.mail.sdv[tid] = (float)sqrt(s / (double)(n - 1));
    }
    if (a.std_mode == TDOD_STD_TORCH_CUDA && rank == 0) {
        // torch.std as torch's CUDA kernel rounds it (tdod_device.cuh, namespace tstd): the launch shape depends on how
        // many objects the frame's `propose` call holds and the vector-load alignment on the object's row in it
        if (tid == 0) {
            int before = 0, after = 0;
            if (a.obj_frame) {
                while (obj - before - 1 >= 0 && a.obj_frame[obj - before - 1] == c.frame) ++before;
                while (obj + after + 1 < a.n_obj && a.obj_frame[obj + after + 1] == c.frame) ++after;
            } else { before = obj; after = a.n_obj - 1 - obj; }
            sh.red_i[0] = before + after + 1; sh.red_i[1] = before;
        }
        __syncthreads();
        const tstd::Cfg cfg = tstd::config(n, sh.red_i[0]);
        const int shift = (int)(((long long)sh.red_i[1] * n) & 1);
        tstd::W* scr = reinterpret_cast<tstd::W*>(&sh.mail.hist[0][0][0]);   // passes 0-2 are dead cluster-wide by now
        for (int k = 0; k < 3; ++k) {
            float* loc = k == 0 ? s_x : (k == 1 ? s_y : s_z);
            const float sd = tstd::block_std([&](int i) { return *cl.remote(loc + (i % n_loc), i / n_loc); }, n, shift, cfg,
                                             scr, tid, kThreads);
            if (tid == 0) sh.mail.sdv[k] = sd;
        }
    }
    cl.sync();   // (5)
    if (S > 1 && a.std_mode != TDOD_STD_TORCH_CUDA) {   // clusters: the three owners gather the squared deviations (rank order)
        if (tid < 3) {
            double s = 0.0;
            for (int r = 0; r < S; ++r) s += *cl.remote(&sh.mail.m2[tid], r);
            sh.sd_loc[tid] = (float)sqrt(s / (double)(n - 1));             // torch.std: unbiased (fp64 == torch on CPU)
        }
        __syncthreads();
    }
    float smean[3], ssd[3];
    {
        const float coef_a[3] = {1.15f, 1.1f, 0.85f}, coef_b[3] = {0.f, 0.f, 0.35f}, coef_s[3] = {1.2f, 0.8f, 1.2f};
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            float sd;
            if (S == 1 || a.std_mode == TDOD_STD_TORCH_CUDA) sd = *cl.remote(&sh.mail.sdv[k], 0);
            else sd = sh.sd_loc[k];
            smean[k] = __fadd_rn(__fmul_rn(coef_a[k], sh.med[k]), coef_b[k]);   // lin_fun, proposals.py:16-18
            ssd[k] = __fmul_rn(sd, coef_s[k]);
        }
    }
    if (a.stats && rank == 0 && tid == 0) {
        float* st = a.stats + (size_t)obj * 6;
        st[0] = smean[0]; st[1] = ssd[0]; st[2] = smean[1]; st[3] = ssd[1]; st[4] = smean[2]; st[5] = ssd[2];
    }

    // ---------------- phase 2: cube -> corners -> projection -> box -> scores ----------------
    const Clamp clampr = make_clamp(a.W, a.H);
    const float* ib = a.iou_boxes ? a.iou_boxes + (size_t)obj * 4 : a.boxes + (size_t)obj * 4;
    const float ib0 = ib[0], ib1 = ib[1], ib2 = ib[2], ib3 = ib[3];
    const float r_ref = __fdiv_rn(__fsub_rn(c.bx1, c.bx0), __fsub_rn(c.by1, c.by0));   // scorefunction.py:154

    for (int sweep = (has_rect || a.propose_only) ? 1 : 0; sweep < 2; ++sweep) {
        // sweep 0 (only without a mask rectangle): fallback rectangle of scorefunction.py:69-75.  With the callers'
        // [1,n,8,2] corners, `bube_corners[:,:,0].min(1)[0].mean()` is the mean of (min over PROPOSALS of vertex 0's x,
        // min over proposals of vertex 0's y); the 'y' values use vertex 1 the same way.
        float fmn[4] = {CUDART_INF_F, CUDART_INF_F, CUDART_INF_F, CUDART_INF_F};      // v0x v0y v1x v1y
        float fmx[4] = {-CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F};
        float lmaxdiff = -CUDART_INF_F, lmaxd = -CUDART_INF_F;
        for (int j = tid; j < cnt; j += kThreads) {
            const int i = i0 + j;
            float q[15];
            draw_xyz_rot(a, c, sh.yaw, i, smean, ssd, s_d[j], q, sweep == 1);
            q[3] = s_w[j]; q[4] = s_h[j]; q[5] = s_l[j];
            float X[8], Y[8], Z[8], U[8], V[8];
            cube_corners(q, X, Y, Z);
            float x0 = CUDART_INF_F, y0 = CUDART_INF_F, x1 = -CUDART_INF_F, y1 = -CUDART_INF_F;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                project(sh.K, clampr, X[k], Y[k], Z[k], U[k], V[k]);
                x0 = min_nan(x0, U[k]); x1 = max_nan(x1, U[k]);
                y0 = min_nan(y0, V[k]); y1 = max_nan(y1, V[k]);
            }
            if (sweep == 0) {
                fmn[0] = min_nan(fmn[0], U[0]); fmn[1] = min_nan(fmn[1], V[0]);
                fmn[2] = min_nan(fmn[2], U[1]); fmn[3] = min_nan(fmn[3], V[1]);
                fmx[0] = max_nan(fmx[0], U[0]); fmx[1] = max_nan(fmx[1], V[0]);
                fmx[2] = max_nan(fmx[2], U[1]); fmx[3] = max_nan(fmx[3], V[1]);
                continue;
            }
            if (a.propose_only) goto emit;   // uniform over the grid: proposals.propose / cubes_to_box only
            {
            const float iou = iou_xyxy(ib0, ib1, ib2, ib3, x0, y0, x1, y1);
            const float g = dims_gauss(q[3], q[4], q[5], c.mu, c.sg);
            const float ratio = __fdiv_rn(__fsub_rn(x1, x0), __fsub_rn(y1, y0));
            const float diff = fabsf(__fsub_rn(r_ref, ratio));
            const float d = chamfer4x8(sh.rect, U, V);
            s_x[j] = iou; s_y[j] = g; s_z[j] = diff; s_d[j] = d;
            lmaxdiff = max_nan(lmaxdiff, diff);
            lmaxd = max_nan(lmaxd, d);
            }
        emit:
            if (a.all_cubes) {
                float* oc = a.all_cubes + ((size_t)obj * n + i) * 15;
#pragma unroll
                for (int k = 0; k < 15; ++k) oc[k] = q[k];
            }
            if (a.all_boxes) {
                float4* ob = reinterpret_cast<float4*>(a.all_boxes + ((size_t)obj * n + i) * 4);
                *ob = make_float4(x0, y0, x1, y1);
            }
        }
        if (sweep == 0) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                fmn[k] = -warp_max_nan(-fmn[k]);
                fmx[k] = warp_max_nan(fmx[k]);
            }
            if (lane == 0) {
                for (int k = 0; k < 4; ++k) { sh.red_fb[wid][k] = fmn[k]; sh.red_fb[wid][4 + k] = fmx[k]; }
            }
            __syncthreads();
            if (tid < 8) {
                const int k = tid & 3;
                float m = (tid < 4) ? CUDART_INF_F : -CUDART_INF_F;
                for (int w2 = 0; w2 < kWarps; ++w2)
                    m = (tid < 4) ? min_nan(m, sh.red_fb[w2][k]) : max_nan(m, sh.red_fb[w2][4 + k]);
                // layout: fb[0..1] = min v0 (x,y), fb[2..3] = max v0, fb[4..5] = min v1, fb[6..7] = max v1
                const int slot = (k < 2) ? ((tid < 4) ? k : 2 + k) : ((tid < 4) ? 4 + (k - 2) : 6 + (k - 2));
                sh.mail.fb[slot] = m;
            }
            cl.sync();   // (5b)
            if (tid == 0) {
                float t8[8];
                for (int k = 0; k < 8; ++k) {
                    float m = (k & 2) ? -CUDART_INF_F : CUDART_INF_F;
                    for (int r = 0; r < S; ++r) {
                        const float v = *cl.remote(&sh.mail.fb[k], r);
                        m = (k & 2) ? max_nan(m, v) : min_nan(m, v);
                    }
                    t8[k] = m;
                }
                const float fx0 = __fdiv_rn(__fadd_rn(t8[0], t8[1]), 2.f), fx1 = __fdiv_rn(__fadd_rn(t8[2], t8[3]), 2.f);
                const float fy0 = __fdiv_rn(__fadd_rn(t8[4], t8[5]), 2.f), fy1 = __fdiv_rn(__fadd_rn(t8[6], t8[7]), 2.f);
                // [[x0,y0],[x1,y0],[x1,y1],[x0,y1]]
                sh.rect[0] = fx0; sh.rect[1] = fy0; sh.rect[2] = fx1; sh.rect[3] = fy0;
                sh.rect[4] = fx1; sh.rect[5] = fy1; sh.rect[6] = fx0; sh.rect[7] = fy1;
            }
            __syncthreads();
            continue;
        }
        lmaxdiff = warp_max_nan(lmaxdiff);
        lmaxd = warp_max_nan(lmaxd);
        if (lane == 0) { sh.red_f[wid][0] = lmaxdiff; sh.red_f[wid][1] = lmaxd; }
        __syncthreads();
        if (tid < 2) {
            float m = -CUDART_INF_F;
            for (int w2 = 0; w2 < kWarps; ++w2) {
                const float v = sh.red_f[w2][tid];
                m = max_nan(m, v);
            }
            if (tid == 0) sh.mail.maxdiff = m; else sh.mail.maxd = m;
        }
    }
    cl.sync();   // (6)
    if (a.propose_only) return;
    float maxdiff = -CUDART_INF_F, maxd = -CUDART_INF_F;
    for (int r = 0; r < S; ++r) {
        const float v0 = *cl.remote(&sh.mail.maxdiff, r), v1 = *cl.remote(&sh.mail.maxd, r);
        maxdiff = max_nan(maxdiff, v0);
        maxd = max_nan(maxd, v1);
    }

    // ---------------- phase 3: final scores, argmax (np.argmax: first NaN, else first maximum) ----------------
    ArgMax best;
    best.v = -CUDART_INF_F; best.i = 0x7fffffff;
    for (int j = tid; j < cnt; j += kThreads) {
        const int i = i0 + j;
        const float iou = s_x[j], g = s_y[j], diff = s_z[j], d = s_d[j];
        const float dim = __fmul_rn(__fsub_rn(1.f, __fdiv_rn(diff, maxdiff)), g);     // scorefunction.py:160
        const float cor = __fsub_rn(1.f, __fdiv_rn(d, maxd));                        // :85
        const float comb = __fmul_rn(__fmul_rn(iou, dim), cor);                      // roi_heads.py:499
        if (a.all_scores) {
            float* os = a.all_scores + (size_t)obj * (a.score_channels ? a.score_channels : 4) * n;
            os[i] = iou; os[n + i] = dim; os[2 * n + i] = cor; os[3 * n + i] = comb;
        }
        if (argmax_better(best.v, best.i, comb, i)) { best.v = comb; best.i = i; }
    }
    best = warp_argmax(best);
    if (lane == 0) { sh.red_f[wid][0] = best.v; sh.red_i[wid] = best.i; }
    __syncthreads();
    if (tid == 0) {
        ArgMax b; b.v = sh.red_f[0][0]; b.i = sh.red_i[0];
        for (int w2 = 1; w2 < kWarps; ++w2)
            if (argmax_better(b.v, b.i, sh.red_f[w2][0], sh.red_i[w2])) { b.v = sh.red_f[w2][0]; b.i = sh.red_i[w2]; }
        sh.mail.am_v = b.v; sh.mail.am_i = b.i;
    }
    cl.sync();   // (7)
    ArgMax win; win.v = -CUDART_INF_F; win.i = 0x7fffffff;
    for (int r = 0; r < S; ++r) {
        const float v = *cl.remote(&sh.mail.am_v, r);
        const int ii = *cl.remote(&sh.mail.am_i, r);
        if (argmax_better(win.v, win.i, v, ii)) { win.v = v; win.i = ii; }
    }
    // the CTA that owns the winner recomputes its cube and box and writes the 84-byte record
    if (win.i >= i0 && win.i < i0 + cnt && tid == 0) {
        const int j = win.i - i0;
        float q[15];
        float xn = 0.f;
        if (a.rng_mode == TDOD_RNG_PHILOX) xn = philox_xn(a, c, win.i);    // s_d[j] now holds the chamfer distance
        draw_xyz_rot(a, c, sh.yaw, win.i, smean, ssd, xn, q, false);
        q[3] = s_w[j]; q[4] = s_h[j]; q[5] = s_l[j];
        float X[8], Y[8], Z[8];
        cube_corners(q, X, Y, Z);
        float x0 = CUDART_INF_F, y0 = CUDART_INF_F, x1 = -CUDART_INF_F, y1 = -CUDART_INF_F;
        for (int k = 0; k < 8; ++k) {
            float u, v;
            project(sh.K, clampr, X[k], Y[k], Z[k], u, v);
            x0 = min_nan(x0, u); x1 = max_nan(x1, u); y0 = min_nan(y0, v); y1 = max_nan(y1, v);
        }
        a.best_idx[obj] = win.i;
        a.best_score[obj] = win.v;
        for (int k = 0; k < 15; ++k) a.best_cube[(size_t)obj * 15 + k] = q[k];
        float* bb = a.best_box + (size_t)obj * 4;
        bb[0] = x0; bb[1] = y0; bb[2] = x1; bb[3] = y1;
    }
    cl.sync();   // (8) nobody leaves while a peer may still read its mailbox
}

}  // namespace

extern "C" int tdod_propose_score_select(const tdod_pss_args* ap, void* stream) {
    if (!ap) return tdod_fail(TDOD_EINVAL, "args is NULL");
    const tdod_pss_args& a = *ap;
    if (a.n_obj == 0) return TDOD_OK;
    if (a.n_obj < 0 || a.n_prop < 2) return tdod_fail(TDOD_EINVAL, "n_obj must be >= 0 and n_prop >= 2");
    if (!a.depth || !a.K || !a.boxes || !a.prior_mean || !a.prior_std)
        return tdod_fail(TDOD_EINVAL, "a required pointer is NULL");
    if (a.propose_only) {
        if (!a.all_cubes && !a.all_boxes) return tdod_fail(TDOD_EINVAL, "propose_only needs all_cubes or all_boxes");
    } else {
        if (!a.best_idx || !a.best_score || !a.best_cube || !a.best_box)
            return tdod_fail(TDOD_EINVAL, "a winner output pointer is NULL");
        if (!a.rects && !a.has_rect) return tdod_fail(TDOD_EINVAL, "rects is NULL and has_rect does not say so");
    }
    if (a.H <= 0 || a.W <= 0 || a.F <= 0) return tdod_fail(TDOD_EINVAL, "bad frame shape");
    if (a.rng_mode == TDOD_RNG_HOST) {
        if (!a.noise) return tdod_fail(TDOD_EINVAL, "rng_mode HOST needs noise");
        if (a.yaw_table ? !a.yaw_idx : !a.rot) return tdod_fail(TDOD_EINVAL, "rng_mode HOST needs yaw_idx (with yaw_table) or rot");
    } else if (a.rng_mode != TDOD_RNG_PHILOX) {
        return tdod_fail(TDOD_EINVAL, "unknown rng_mode");
    }
    if (a.centre_mode < TDOD_CENTRE_DEPTH || a.centre_mode > TDOD_CENTRE_PATCH_Z) return tdod_fail(TDOD_EINVAL, "unknown centre_mode");
    if (a.dims_mode < TDOD_DIMS_PRIOR || a.dims_mode > TDOD_DIMS_ASPECT) return tdod_fail(TDOD_EINVAL, "unknown dims_mode");
    if (a.dims_mode == TDOD_DIMS_ASPECT && !a.aux) return tdod_fail(TDOD_EINVAL, "dims_mode ASPECT needs aux [n_obj,2]");
    if (a.centre_mode == TDOD_CENTRE_PATCH_Z && a.rng_mode != TDOD_RNG_HOST)
        return tdod_fail(TDOD_EINVAL, "centre_mode PATCH_Z takes its z ramp from the caller (rng_mode HOST)");
    if (a.std_mode != TDOD_STD_EXACT && a.std_mode != TDOD_STD_TORCH_CUDA) return tdod_fail(TDOD_EINVAL, "unknown std_mode");
    if (a.score_channels != 0 && a.score_channels != 4 && a.score_channels != 6) return tdod_fail(TDOD_EINVAL, "score_channels must be 0, 4 or 6");
    if (a.mode_flags & ~(TDOD_MODEB_SEG | TDOD_MODEB_PC)) return tdod_fail(TDOD_EINVAL, "unknown mode_flags");
    if (a.mode_flags && a.propose_only) return tdod_fail(TDOD_EINVAL, "mode B scores need the scoring pass (propose_only is set)");
    if (a.std_mode == TDOD_STD_TORCH_CUDA && !tstd::config(a.n_prop, 1).ok)
        return tdod_fail(TDOD_ERANGE, "std_mode TORCH_CUDA: n_prop so large that torch splits the reduction over several CTAs");
    int S = a.cluster_size;
    if (S == 0) {
        // one CTA handles up to 2048 proposals comfortably; larger objects are spread over a cluster.  (Round 1 also
        // clustered small batches of short objects: measured on C1 -- 5 objects x 1000 -- one CTA per object is faster,
        // 0.089 against 0.097 / 0.099 / 0.122 ms per step at cluster sizes 2 / 4 / 8: the cluster barriers of the radix
        // passes cost more than the halved per-proposal work saves.)
        S = 1;
        while (S < kMaxCluster && (a.n_prop + S - 1) / S > 2048) S *= 2;
    }
    if (S != 1 && S != 2 && S != 4 && S != 8) return tdod_fail(TDOD_EINVAL, "cluster_size must be 0,1,2,4,8");
    const int n_loc = (a.n_prop + S - 1) / S;
    const size_t smem = ((sizeof(Shared) + 15) & ~size_t(15)) + (size_t)7 * n_loc * sizeof(float);
    if (smem > 227 * 1024) return tdod_fail(TDOD_ERANGE, "n_prop too large for shared memory at this cluster size");
    cudaError_t e;
    static tdod_smem_cache sm_cta, sm_cluster, sm_hot;
    const bool hot = a.rng_mode == TDOD_RNG_PHILOX && a.centre_mode == TDOD_CENTRE_DEPTH && a.dims_mode == TDOD_DIMS_PRIOR &&
                     a.yaw_table && a.obj_frame && !a.propose_only && a.std_mode == TDOD_STD_EXACT && !a.all_cubes && !a.all_boxes &&
                     !a.all_scores && !a.stats && !a.out_noise && !a.out_yaw_idx && !a.out_seeds && !a.mode_flags;
    if (S == 1) {
        if (hot) {
            e = tdod_smem_opt_in((const void*)pss_kernel<false, true>, smem, &sm_hot);
            if (e != cudaSuccess) return (int)e;
            pss_kernel<false, true><<<(unsigned)a.n_obj, kThreads, smem, (cudaStream_t)stream>>>(a, n_loc);
            return (int)cudaGetLastError();
        }
        e = tdod_smem_opt_in((const void*)pss_kernel<false, false>, smem, &sm_cta);
        if (e != cudaSuccess) return (int)e;
        pss_kernel<false, false><<<(unsigned)a.n_obj, kThreads, smem, (cudaStream_t)stream>>>(a, n_loc);
        e = cudaGetLastError();
        if (e != cudaSuccess) return (int)e;
        return a.mode_flags ? tdod_modeb_launch(a, stream) : TDOD_OK;
    }
    e = tdod_smem_opt_in((const void*)pss_kernel<true, false>, smem, &sm_cluster);
    if (e != cudaSuccess) return (int)e;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)((int64_t)a.n_obj * S), 1, 1);
    cfg.blockDim = dim3(kThreads, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = S; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    e = cudaLaunchKernelEx(&cfg, pss_kernel<true, false>, a, n_loc);
    if (e != cudaSuccess) return (int)e;
    return a.mode_flags ? tdod_modeb_launch(a, stream) : TDOD_OK;
}
