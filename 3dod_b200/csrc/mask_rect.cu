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
// Two kernels.  `pack_kernel` is the only one that touches the masks: it streams them once from HBM with 16-byte
// loads and writes a 1-bit-per-pixel copy (rows padded to whole words) plus the occupied row range of every mask;
// it is bandwidth-bound and sized to keep ~100 KB of loads in flight per SM.  `rect_kernel` then works per mask on
// the bitmap (L2-resident, 1/8 of the bytes): run-length encoding, union-find over runs, cell-formula areas, strict
// hull, rotating calipers.  Those steps are short dependent chains, so the kernel is built small (256 threads,
// ~36 KB of shared memory) to keep 4+ masks per SM in flight and overlap their serial tails.  Pass 1 decides every
// mask whose winner is certain from bounds (area without hole filling <= true area <= bounding box); pass 2 fills
// holes exactly for the ambiguous rest; pass 3 handles speckle masks whose runs outgrow shared memory.
#include <cstdlib>
#include "tdod_device.cuh"
#include "tdod_internal.h"

using namespace tdod;

namespace {

#ifndef TDOD_KB_T
#define TDOD_KB_T 256
#endif
#ifndef TDOD_KB_MINB
#define TDOD_KB_MINB 5
#endif
#ifndef TDOD_KB_RUNCAP
#define TDOD_KB_RUNCAP 2048
#endif
constexpr int kT = TDOD_KB_T;    // threads of the rectangle kernels
constexpr int kPackT = 256;      // threads of the pack kernel
constexpr int kPackRows = 32;    // mask rows per pack tile

// phase timers for tools/kb_phases.cu (compiled out of the library)
#ifdef TDOD_KB_TIMING
__device__ long long g_kb_t[32 * 8192];
#define KB_T(k) do { if (threadIdx.x == 0 && blockIdx.x < 8192) g_kb_t[blockIdx.x * 32 + (k)] = clock64(); } while (0)
#else
#define KB_T(k) do { } while (0)
#endif
constexpr int kRunCap = TDOD_KB_RUNCAP;     // runs per mask held in shared memory (more -> the global-memory overflow pass)

struct RunSet {                   // shared-memory views
    uint16_t* xs; uint16_t* xe; uint16_t* row;
    int* parent; int* area2;      // area2: twice the cell-formula area, accumulated at the root
    int* row_off;                 // [H+1] first run of each row
};

__device__ __forceinline__ uint32_t nz_bits16(const uint4 v) {
    // 16 mask bytes -> 16 bits (bit i = byte i != 0).  Per word: bit 7 of every byte := (byte != 0) (three logic / add
    // operations), then one multiply gathers the four flags into the top nibble: flag j sits at bit 8j+7 and the
    // multiplier 2^21 + 2^14 + 2^7 + 1 moves it to bit 28+j; all partial products land on distinct bits, so no carry
    // reaches the nibble.
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    uint32_t p[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const uint32_t t = w[k];
        const uint32_t hi = (t | ((t & 0x7f7f7f7fu) + 0x7f7f7f7fu)) & 0x80808080u;
        p[k] = hi * 0x00204081u;
    }
    return (p[0] >> 28) | ((p[1] >> 24) & 0xF0u) | ((p[2] >> 20) & 0xF00u) | ((p[3] >> 16) & 0xF000u);
}

// ------------------------------------------------------------------------------------------------------------
// pack: masks [n,H,W] bytes -> bitmap [n,H,Wp] words (bit c&31 of word c>>5 of a row) + occupied row range
// ------------------------------------------------------------------------------------------------------------
// A tile is kPackRows consecutive rows of one mask.  Its bytes are fetched as aligned 16-byte chunks of the flat
// mask array (a row is W bytes, so tiles start at any alignment; the chunks that straddle the tile ends are shared
// with the neighbouring tile), reduced to 16 bits each and laid down as a contiguous bit stream in shared memory
// with plain half-word stores; every output word is then a 32-bit window of that stream (funnel shift).
__global__ void __launch_bounds__(kPackT, 8) pack_kernel(const uint8_t* __restrict__ masks, int64_t n_masks, int H, int W, int Wp,
                                                        uint32_t* __restrict__ bitmap, int* __restrict__ rowrange,
                                                        int* __restrict__ lists, int tiles_per_mask) {
    extern __shared__ __align__(16) unsigned char pk_raw[];
    uint16_t* sb16 = reinterpret_cast<uint16_t*>(pk_raw);
    const uint32_t* sb32 = reinterpret_cast<const uint32_t*>(pk_raw);
    __shared__ int s_lo, s_hi;
    const int tid = threadIdx.x;
    const uintptr_t base = reinterpret_cast<uintptr_t>(masks);
    const int64_t mis = (int64_t)(base & 15);
    const uint4* v = reinterpret_cast<const uint4*>(base - mis);
    const int64_t HW = (int64_t)H * W;
    const int64_t n_tiles = n_masks * tiles_per_mask;
    if (blockIdx.x == 0 && tid < 4) lists[tid] = 0;       // work-list counters of the later passes
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int64_t m = tile / tiles_per_mask;
        const int r0 = (int)(tile - m * tiles_per_mask) * kPackRows, r1 = min(H, r0 + kPackRows);
        const int64_t b0 = mis + m * HW + (int64_t)r0 * W, b1 = mis + m * HW + (int64_t)r1 * W;
        const int64_t c0 = b0 >> 4;
        const int nch = (int)(((b1 + 15) >> 4) - c0);
        if (tid == 0) { s_lo = 0x7fffffff; s_hi = -1; }
        // bytes outside [b0, b1) that ride along in the first / last chunk are never selected below
        int c = tid;
        for (; c + 3 * kPackT < nch; c += 4 * kPackT) {
            const uint4 a0 = __ldcs(v + c0 + c), a1 = __ldcs(v + c0 + c + kPackT), a2 = __ldcs(v + c0 + c + 2 * kPackT),
                        a3 = __ldcs(v + c0 + c + 3 * kPackT);
            sb16[c] = (uint16_t)nz_bits16(a0); sb16[c + kPackT] = (uint16_t)nz_bits16(a1);
            sb16[c + 2 * kPackT] = (uint16_t)nz_bits16(a2); sb16[c + 3 * kPackT] = (uint16_t)nz_bits16(a3);
        }
        for (; c < nch; c += kPackT) sb16[c] = (uint16_t)nz_bits16(__ldcs(v + c0 + c));
        if (tid < 4) sb16[nch + tid] = 0;                       // the funnel reads one word past the last chunk
        __syncthreads();
        const int nwords = (r1 - r0) * Wp;
        int lo = 0x7fffffff, hi = -1;
        for (int k = tid; k < nwords; k += kPackT) {
            const int rr = k / Wp, w = k - rr * Wp, r = r0 + rr;
            const int o = (int)(b0 - (c0 << 4)) + rr * W + 32 * w;          // bit offset of the window in the stream
            uint32_t word = __funnelshift_r(sb32[o >> 5], sb32[(o >> 5) + 1], o & 31);
            const int nb = W - 32 * w;
            if (nb < 32) word &= (1u << nb) - 1u;
            bitmap[((size_t)m * H + r) * Wp + w] = word;
            if (word) { lo = min(lo, r); hi = max(hi, r); }
        }
        lo = __reduce_min_sync(0xffffffffu, lo); hi = __reduce_max_sync(0xffffffffu, hi);
        if ((tid & 31) == 0 && hi >= 0) { atomicMin(&s_lo, lo); atomicMax(&s_hi, hi); }
        __syncthreads();
        if (tid == 0 && s_hi >= 0) {                            // rowrange[m] = {first row, H-1-last row}, both via atomicMin
            atomicMin(&rowrange[2 * m], s_lo);
            atomicMin(&rowrange[2 * m + 1], H - 1 - s_hi);
        }
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------------------------
// pack, TMA-staged: the same tiles, but the raw bytes of tile k+1 are brought into shared memory by one bulk
// asynchronous copy (cp.async.bulk, completion on an mbarrier) while the CTA turns tile k into bits.  Loads are no
// longer tied to registers or to the instruction stream of the packing code, so HBM stays busy across the barriers
// of the packing phases.  Two raw buffers per CTA; used whenever they fit in shared memory.
// ------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_load(void* dst, const void* src, uint32_t bytes, uint64_t* b) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(b))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(b)),
        "r"(parity)
        : "memory");
}

__global__ void __launch_bounds__(kPackT) pack_tma_kernel(const uint8_t* __restrict__ masks, int64_t n_masks, int H, int W, int Wp,
                                                         uint32_t* __restrict__ bitmap, int* __restrict__ rowrange,
                                                         int* __restrict__ lists, int tiles_per_mask, int raw_cap /*bytes, x16*/) {
    extern __shared__ __align__(128) unsigned char pk_raw[];
    unsigned char* raw0 = pk_raw;
    unsigned char* raw1 = pk_raw + raw_cap;
    uint16_t* sb16 = reinterpret_cast<uint16_t*>(pk_raw + 2 * (size_t)raw_cap);
    const uint32_t* sb32 = reinterpret_cast<const uint32_t*>(sb16);
    __shared__ __align__(8) uint64_t mbar[2];
    __shared__ int s_lo, s_hi;
    const int tid = threadIdx.x;
    const uintptr_t base = reinterpret_cast<uintptr_t>(masks);
    const int64_t mis = (int64_t)(base & 15);
    const unsigned char* vbase = reinterpret_cast<const unsigned char*>(base - mis);
    const int64_t HW = (int64_t)H * W;
    const int64_t n_tiles = n_masks * tiles_per_mask;
    if (blockIdx.x == 0 && tid < 4) lists[tid] = 0;       // work-list counters of the later passes

    // every thread needs the tile geometry of every tile: 32-bit division (the entry point bounds n_tiles), and the
    // word -> (row, column) split of the inner loop by a multiply-high with a precomputed reciprocal of Wp
    const uint32_t wp_magic = (uint32_t)((0x100000000ull + (uint32_t)Wp - 1u) / (uint32_t)Wp);   // exact for q * Wp < 2^32
    auto geometry = [&](int64_t tile, int64_t& m, int& r0, int& r1, int64_t& b0, int64_t& c0, int& nch) {
        const uint32_t mm = (uint32_t)tile / (uint32_t)tiles_per_mask;
        m = mm;
        r0 = (int)((uint32_t)tile - mm * (uint32_t)tiles_per_mask) * kPackRows; r1 = min(H, r0 + kPackRows);
        b0 = mis + m * HW + (int64_t)r0 * W;
        const int64_t b1 = mis + m * HW + (int64_t)r1 * W;
        c0 = b0 >> 4;
        nch = (int)(((b1 + 15) >> 4) - c0);
    };
    if (tid == 0) { mbar_init(&mbar[0], 1); mbar_init(&mbar[1], 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();
    if (tid == 0 && (int64_t)blockIdx.x < n_tiles) {
        int64_t m, b0, c0; int r0, r1, nch;
        geometry(blockIdx.x, m, r0, r1, b0, c0, nch);
        mbar_expect_tx(&mbar[0], (uint32_t)nch * 16u);
        bulk_load(raw0, vbase + (c0 << 4), (uint32_t)nch * 16u, &mbar[0]);
    }
    int k = 0;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++k) {
        const int cur = k & 1;
        int64_t m, b0, c0; int r0, r1, nch;
        geometry(tile, m, r0, r1, b0, c0, nch);
        if (tid == 0) {
            s_lo = 0x7fffffff; s_hi = -1;
            const int64_t nxt = tile + gridDim.x;
            if (nxt < n_tiles) {   // buffer cur^1 was last read two barriers ago
                int64_t m2, b02, c02; int r02, r12, nch2;
                geometry(nxt, m2, r02, r12, b02, c02, nch2);
                mbar_expect_tx(&mbar[cur ^ 1], (uint32_t)nch2 * 16u);
                bulk_load(cur ? raw0 : raw1, vbase + (c02 << 4), (uint32_t)nch2 * 16u, &mbar[cur ^ 1]);
            }
        }
        mbar_wait(&mbar[cur], (uint32_t)(k >> 1) & 1u);
        const uint4* rv = reinterpret_cast<const uint4*>(cur ? raw1 : raw0);
        for (int c = tid; c < nch; c += kPackT) sb16[c] = (uint16_t)nz_bits16(rv[c]);
        if (tid < 4) sb16[nch + tid] = 0;                       // the funnel reads one word past the last chunk
        __syncthreads();
        const int nwords = (r1 - r0) * Wp;
        int lo = 0x7fffffff, hi = -1;
        for (int q = tid; q < nwords; q += kPackT) {
            const int rr = Wp == 1 ? q : (int)__umulhi((uint32_t)q, wp_magic), w = q - rr * Wp, r = r0 + rr;   // (the magic is 2^32 for Wp = 1)
            const int o = (int)(b0 - (c0 << 4)) + rr * W + 32 * w;          // bit offset of the window in the stream
            uint32_t word = __funnelshift_r(sb32[o >> 5], sb32[(o >> 5) + 1], o & 31);
            const int nb = W - 32 * w;
            if (nb < 32) word &= (1u << nb) - 1u;
            bitmap[((size_t)m * H + r) * Wp + w] = word;
            if (word) { lo = min(lo, r); hi = max(hi, r); }
        }
        lo = __reduce_min_sync(0xffffffffu, lo); hi = __reduce_max_sync(0xffffffffu, hi);
        if ((tid & 31) == 0 && hi >= 0) { atomicMin(&s_lo, lo); atomicMax(&s_hi, hi); }
        __syncthreads();
        if (tid == 0 && s_hi >= 0) {
            atomicMin(&rowrange[2 * m], s_lo);
            atomicMin(&rowrange[2 * m + 1], H - 1 - s_hi);
        }
        __syncthreads();
    }
}

// Where the labelling reads a mask's bits from.  GlobalBits: the row-padded 1 bit/px bitmap the pack kernels write
// (word w of row r at r * Wp + w).  StreamBits: the fused front kernel's shared-memory bit stream -- bit i of the
// stream is byte i of the mask's 16-byte-aligned chunk range, rows are NOT padded, so a row word is a 32-bit window at
// bit offset off0 + r * W + 32 w (one funnel shift over two shared-memory words).
struct GlobalBits {
    const uint32_t* bm; int Wp;
    __device__ __forceinline__ uint32_t word(int r, int w) const { return bm[(size_t)r * Wp + w]; }
};
struct StreamBits {
    const uint32_t* s32; int off0, W;
    __device__ __forceinline__ uint32_t word(int r, int w) const {
        const int o = off0 + r * W + 32 * w;
        uint32_t v = __funnelshift_r(s32[o >> 5], s32[(o >> 5) + 1], o & 31);
        const int nb = W - 32 * w;
        if (nb < 32) v &= (1u << nb) - 1u;
        return v;
    }
};

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

// Runs of one bitmap row, one warp per row: lane l owns word l (+32, +64, ... for rows wider than 1024 px).  A run
// starts where a set bit follows a clear one and ends where a set bit precedes a clear one; the k-th start of a row
// pairs with its k-th end, so lanes can write starts and ends independently at prefix-summed positions.
// emit == false: returns the row's run count (valid in every lane).
template <bool kEmit, typename Bits>
__device__ __forceinline__ int row_runs_warp(const Bits& bm, int Wp, int r, const RunSet& rs, int base, int cap) {
    const int lane = threadIdx.x & 31;
    int S = 0, E = 0;
    uint32_t carry = 0;
    for (int w0 = 0; w0 < Wp; w0 += 32) {
        const int w = w0 + lane;
        const uint32_t m = (w < Wp) ? bm.word(r, w) : 0u;
        uint32_t prev = __shfl_up_sync(0xffffffffu, m >> 31, 1);
        if (lane == 0) prev = carry;
        carry = __shfl_sync(0xffffffffu, m >> 31, 31);
        const uint32_t starts = m & ~((m << 1) | prev);
        const int ns = __popc(starts);
        if (!kEmit) { S += ns; continue; }
        uint32_t next = __shfl_down_sync(0xffffffffu, m & 1u, 1);
        if (lane == 31) next = (w0 + 32 < Wp) ? (bm.word(r, w0 + 32) & 1u) : 0u;
        const uint32_t ends = m & ~((m >> 1) | (next << 31));
        const int ne = __popc(ends);
        int ps = ns, pe = ne;                                   // inclusive warp prefix sums
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int a = __shfl_up_sync(0xffffffffu, ps, o), b = __shfl_up_sync(0xffffffffu, pe, o);
            if (lane >= o) { ps += a; pe += b; }
        }
        int ks = base + S + ps - ns, ke = base + E + pe - ne;
        for (uint32_t t = starts; t; t &= t - 1, ++ks)
            if (ks < cap) { rs.xs[ks] = (uint16_t)(32 * w + __ffs(t) - 1); rs.row[ks] = (uint16_t)r; }
        for (uint32_t t = ends; t; t &= t - 1, ++ke)
            if (ke < cap) rs.xe[ke] = (uint16_t)(32 * w + __ffs(t) - 1);
        S += __shfl_sync(0xffffffffu, ps, 31);
        E += __shfl_sync(0xffffffffu, pe, 31);
    }
    if (!kEmit) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) S += __shfl_xor_sync(0xffffffffu, S, o);
    }
    return S;
}

// label the runs of rows [r0, r1] of `bm` (8-connectivity; the rows outside are empty), accumulate twice the
// cell-formula area at each root.  returns the number of runs, or -1 if they exceed the capacity.
// A group of threads that synchronises among itself: the whole CTA, or a warp-aligned subset on a named barrier (the
// extract / label half of a front CTA, while the other half keeps turning incoming bytes into bits).
struct CtaGrp {
    static __device__ __forceinline__ int tid() { return (int)threadIdx.x; }
    static __device__ __forceinline__ int size() { return (int)blockDim.x; }
    static __device__ __forceinline__ void sync() { __syncthreads(); }
    static __device__ __forceinline__ int sync_or(int p) { return __syncthreads_or(p); }
};
template <int ID, int N, int T0> struct SubGrp {           // threads T0 .. T0 + N - 1 on named barrier ID
    static __device__ __forceinline__ int tid() { return (int)threadIdx.x - T0; }
    static __device__ __forceinline__ int size() { return N; }
    static __device__ __forceinline__ void sync() { asm volatile("bar.sync %0, %1;" ::"n"(ID), "n"(N) : "memory"); }
    static __device__ __forceinline__ int sync_or(int p) {
        int r;
        asm volatile("{\n.reg .pred pi, po;\nsetp.ne.s32 pi, %1, 0;\nbar.red.or.pred po, %2, %3, pi;\nselp.s32 %0, 1, 0, po;\n}"
                     : "=r"(r) : "r"(p), "n"(ID), "n"(N) : "memory");
        return r;
    }
};

template <typename TG> __device__ int label_from_runs(RunSet rs, int R);

template <typename Bits>
__device__ int label_runs(const Bits bm, int H, int W, int Wp, RunSet rs, int* scratch, int cap, int r0, int r1) {
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int r = threadIdx.x; r <= H; r += kT) rs.row_off[r] = 0;
    __syncthreads();
    constexpr int kWarps = kT / 32, kU = 8;                  // kU rows in flight per warp: the bitmap sits in L2 / HBM
    if (Wp <= 32) {
        for (int rb = r0 + wid * kU; rb <= r1; rb += kWarps * kU) {
            uint32_t m[kU];
#pragma unroll
            for (int u = 0; u < kU; ++u) m[u] = (rb + u <= r1 && lane < Wp) ? bm.word(rb + u, lane) : 0u;
#pragma unroll
            for (int u = 0; u < kU; ++u) {
                uint32_t prev = __shfl_up_sync(0xffffffffu, m[u] >> 31, 1);
                if (lane == 0) prev = 0u;
                const int c = __reduce_add_sync(0xffffffffu, __popc(m[u] & ~((m[u] << 1) | prev)));
                if (lane == 0 && rb + u <= r1) rs.row_off[rb + u] = c;
            }
        }
    } else {
        for (int r = r0 + wid; r <= r1; r += kWarps) {
            const int c = row_runs_warp<false, Bits>(bm, Wp, r, rs, 0, 0);
            if (lane == 0) rs.row_off[r] = c;
        }
    }
    __syncthreads();
    KB_T(16);
    const int R = scan_rows(rs.row_off + r0, r1 - r0 + 1, scratch);
    KB_T(17);     // rows outside [r0, r1] hold no runs
    if (R > cap) return -1;
    if (Wp <= 32) {
        for (int rb = r0 + wid * kU; rb <= r1; rb += kWarps * kU) {
            uint32_t m[kU];
#pragma unroll
            for (int u = 0; u < kU; ++u) m[u] = (rb + u <= r1 && lane < Wp) ? bm.word(rb + u, lane) : 0u;
#pragma unroll
            for (int u = 0; u < kU; ++u) {
                const int r = rb + u;
                if (r > r1) break;
                uint32_t prev = __shfl_up_sync(0xffffffffu, m[u] >> 31, 1);
                if (lane == 0) prev = 0u;
                uint32_t next = __shfl_down_sync(0xffffffffu, m[u] & 1u, 1);
                if (lane == 31) next = 0u;
                const uint32_t starts = m[u] & ~((m[u] << 1) | prev), ends = m[u] & ~((m[u] >> 1) | (next << 31));
                if (!__any_sync(0xffffffffu, starts | ends)) continue;
                const int ns = __popc(starts), ne = __popc(ends);
                int ps, pe;                                       // inclusive prefix sums over the lanes
                if (!__any_sync(0xffffffffu, (ns | ne) > 1)) {
                    // the usual row: no word holds more than one run start or end -> two ballots instead of a scan
                    const unsigned le = (2u << lane) - 1u;        // lanes 0..lane
                    ps = __popc(__ballot_sync(0xffffffffu, ns != 0) & le);
                    pe = __popc(__ballot_sync(0xffffffffu, ne != 0) & le);
                } else {
                    int pk = ns | (ne << 16);                     // a word has at most 16 runs, a row at most 512
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        const int a = __shfl_up_sync(0xffffffffu, pk, o);
                        if (lane >= o) pk += a;
                    }
                    ps = pk & 0xffff; pe = pk >> 16;
                }
                const int base = rs.row_off[r];
                int ks = base + ps - ns, ke = base + pe - ne;
                for (uint32_t t = starts; t; t &= t - 1, ++ks)
                    if (ks < cap) { rs.xs[ks] = (uint16_t)(32 * lane + __ffs(t) - 1); rs.row[ks] = (uint16_t)r; }
                for (uint32_t t = ends; t; t &= t - 1, ++ke)
                    if (ke < cap) rs.xe[ke] = (uint16_t)(32 * lane + __ffs(t) - 1);
            }
        }
    } else {
        for (int r = r0 + wid; r <= r1; r += kWarps) row_runs_warp<true, Bits>(bm, Wp, r, rs, rs.row_off[r], cap);
    }
    __syncthreads();                                    // the runs were written by other warps
    return label_from_runs<CtaGrp>(rs, R);
}

// Labels the R runs in `rs` (row_off filled for every row) and accumulates twice the cell-formula area at each root.
// Called by all threads; returns R.
template <typename TG>
__device__ int label_from_runs(RunSet rs, int R) {
    // Labelling (8-connectivity).  A blob is a chain of one run per row, so hanging run i under the run above it
    // through union-find makes every later find walk the whole chain (quadratic in the rows).  Instead:
    //   1. every run points straight at the FIRST run of the previous row it touches (no find, no atomics): a forest
    //      whose trees are pieces of components and whose roots are the smallest index of their tree;
    //   2. pointer jumping flattens the forest (each round walks up to 8 links: 3-4 rounds for 500 rows);
    //   3. the remaining adjacencies -- a run touching several runs of the previous row -- are unions of roots;
    //   4. one short walk per run gives the final root = smallest run index of the component, as before.
    for (int i = TG::tid(); i < R; i += TG::size()) {
        const int r = rs.row[i];
        int first = i;
        if (r != 0) {
            const int s = rs.xs[i], e = rs.xe[i];
            for (int j = rs.row_off[r - 1]; j < rs.row_off[r]; ++j) {
                if ((int)rs.xs[j] > e + 1) break;
                if ((int)rs.xe[j] >= s - 1) { first = j; break; }
            }
        }
        rs.parent[i] = first;
        rs.area2[i] = 0;
    }
    TG::sync();
    KB_T(18);
    while (true) {
        int changed = 0;
        for (int i = TG::tid(); i < R; i += TG::size()) {
            const int p0 = rs.parent[i];
            int p = p0;
#pragma unroll
            for (int k = 0; k < 8; ++k) p = rs.parent[p];          // any value read is an ancestor: races are benign
            if (p != p0) { rs.parent[i] = p; changed = 1; }
        }
        if (!TG::sync_or(changed)) break;
    }
    for (int i = TG::tid(); i < R; i += TG::size()) {
        const int r = rs.row[i];
        if (r == 0) continue;
        const int s = rs.xs[i], e = rs.xe[i];
        int first = -1;
        for (int j = rs.row_off[r - 1]; j < rs.row_off[r]; ++j) {
            if ((int)rs.xs[j] > e + 1) break;
            if ((int)rs.xe[j] >= s - 1) {
                if (first < 0) first = j;
                else uf_union(rs.parent, first, j);
            }
        }
    }
    TG::sync();
    KB_T(19);
    for (int i = TG::tid(); i < R; i += TG::size()) rs.parent[i] = uf_find(rs.parent, i);
    TG::sync();
    KB_T(20);
    // cell formula on pairs of overlapping runs in adjacent rows: 2*full + (#3-corner end cells)
    // (the runs of a blob sit next to each other in the list and share their root: lanes with the same root are summed with
    // match.any + a masked reduction, one shared-memory add per root and warp instead of thirty-two on one address)
    for (int i0 = 0; i0 < R; i0 += TG::size()) {
        const int i = i0 + TG::tid();
        int acc = 0, root = -1;
        if (i < R) {
            const int r = rs.row[i];
            if (r != 0) {
                const int s = rs.xs[i], e = rs.xe[i];
                for (int j = rs.row_off[r - 1]; j < rs.row_off[r]; ++j) {
                    const int s2 = rs.xs[j], e2 = rs.xe[j];
                    if (s2 > e) break;
                    if (e2 < s) continue;
                    acc += 2 * (min(e, e2) - max(s, s2)) + (s != s2) + (e != e2);
                }
            }
            if (acc) root = rs.parent[i];
        }
        const unsigned peers = __match_any_sync(0xffffffffu, root);
        const int sum = __reduce_add_sync(peers, acc);
        if (root >= 0 && (int)(__ffs(peers) - 1) == (int)(threadIdx.x & 31)) atomicAdd(&rs.area2[root], sum);
    }
    TG::sync();
    return R;
}

// The hull / calipers stage runs either on a whole CTA (the exact passes) or on ONE WARP per mask (the tail kernel of the
// fused path, where many masks' serial chains overlap on an SM instead of stalling a CTA at barriers).
template <int G> struct Grp {
    static __device__ __forceinline__ int tid() { return G == 32 ? (int)(threadIdx.x & 31) : (int)threadIdx.x; }
    static __device__ __forceinline__ void sync() { if (G == 32) __syncwarp(); else __syncthreads(); }
    static __device__ __forceinline__ int sync_or(int p) { return G == 32 ? __any_sync(0xffffffffu, p) : __syncthreads_or(p); }
};
// exclusive scan of n ints in place (+ total at [n]) by one warp
__device__ __forceinline__ int scan_rows_warp(int* a, int n) {
    const int lane = threadIdx.x & 31;
    int carry = 0;
    for (int base = 0; base < n; base += 32) {
        const int i = base + lane;
        const int v = (i < n) ? a[i] : 0;
        int inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (i < n) a[i] = carry + inc - v;
        carry += __shfl_sync(0xffffffffu, inc, 31);
    }
    if (lane == 0) a[n] = carry;
    __syncwarp();
    return carry;
}

// ---- hull + rotating calipers of the component `root` ----------------------------------------------------------
// Only two things are inherently sequential: the monotone-chain stacks (one thread per chain, top of stack in
// registers) and the calipers' walk around the hull (one thread, everything it reads precomputed in parallel).
#ifndef TDOD_KB_CHAINCAP
#define TDOD_KB_CHAINCAP 512
#endif
constexpr int kChainCap = TDOD_KB_CHAINCAP;    // vertices of one strictly convex chain (a lattice polygon in a <=1024^2 grid has < 360)
constexpr int kHullCap = 2 * kChainCap;
struct HullBuf { int16_t* rx; int16_t* ry; int16_t* lx; int16_t* ly; float* hx; float* hy; float* vx; float* vy; float* invf; double* inv; };

// OpenCV rotatingCalipers (CALIPERS_MINAREARECT) as the installed 4.13 computes it: edge vectors, inverse lengths
// and the cosines that decide which caliper advances in double, everything else fp32 with every product / sum
// rounded on its own (no FMA contraction), then minAreaRect's centre/size/angle and RotatedRect::points (boxPoints).
// Called by all threads; `ext` = {left, right, top, bottom} start indices (first extreme in hull order).
template <int G>
__device__ void calipers_box(const HullBuf& hb, int n, int* ext /*[4] smem*/, float* out8) {
    const float* px = hb.hx; const float* py = hb.hy;
    const int tid = Grp<G>::tid();
    if (n >= 3) {
        // edge vectors and inverse lengths (double sqrt + divide each): one edge per thread
        for (int i = tid; i < n; i += G) {
            const int j = (i + 1 < n) ? i + 1 : 0;
            const double dx = (double)px[j] - (double)px[i], dy = (double)py[j] - (double)py[i];
            hb.vx[i] = (float)dx; hb.vy[i] = (float)dy;
            const double iv = 1.0 / sqrt(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
            hb.inv[i] = iv; hb.invf[i] = (float)iv;
        }
        // extremes with the serial scan's semantics (strict compares from index 0: the FIRST extreme wins)
        for (int which = (G == 32 ? 0 : (tid >> 5)); which < 4; which += (G == 32 ? 1 : 4)) {
            if (G != 32 && tid >= 128) break;
            const int lane = tid & 31;                       // 0 left (min x), 1 right (max x), 2 top (max y), 3 bottom (min y)
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
    Grp<G>::sync();
    KB_T(8);
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
        // the four calipers (0 bottom, 1 right, 2 top, 3 left): current vertex index, its edge (vx, vy, 1/len) and its
        // coordinates, all in scalar registers; only the caliper that advances reloads from shared memory
        int s0 = bottom, s1 = right, s2 = top, s3 = left;
        // cv2 4.13 keeps the edge vectors and inverse lengths in double and forms the four cosines in double; the base
        // vector, width, height and area stay float (established against the installed library, oracle/cv_spec.py)
        double id0 = hb.inv[s0], id1 = hb.inv[s1], id2 = hb.inv[s2], id3 = hb.inv[s3];
        float vx0 = hb.vx[s0], vy0 = hb.vy[s0], iv0 = hb.invf[s0], px0 = px[s0], py0 = py[s0];
        float vx1 = hb.vx[s1], vy1 = hb.vy[s1], iv1 = hb.invf[s1], px1 = px[s1], py1 = py[s1];
        float vx2 = hb.vx[s2], vy2 = hb.vy[s2], iv2 = hb.invf[s2], px2 = px[s2], py2 = py[s2];
        float vx3 = hb.vx[s3], vy3 = hb.vy[s3], iv3 = hb.invf[s3], px3 = px[s3], py3 = py[s3];
        float minarea = 3.402823466e+38f;
        int b_left = 0, b_bottom = 0; float b_a = 0, b_w = 0, b_b = 0, b_h = 0;
        for (int k = 0; k < n; ++k) {
            // which caliper advances is decided by double-precision cosines (centrally symmetric hulls make two of them
            // mathematically equal at almost every step, so an fp32 pre-test would rarely settle it)
            const double da = (double)base_a, db = (double)base_b;       // float x float products are exact in double
            const double dp0 = __dadd_rn(__dmul_rn(da, (double)vx0), __dmul_rn(db, (double)vy0));
            const double dp1 = __dadd_rn(__dmul_rn(-db, (double)vx1), __dmul_rn(da, (double)vy1));
            const double dp2 = __dsub_rn(__dmul_rn(-da, (double)vx2), __dmul_rn(db, (double)vy2));
            const double dp3 = __dsub_rn(__dmul_rn(db, (double)vx3), __dmul_rn(da, (double)vy3));
            double maxcos = __dmul_rn(dp0, id0);
            int main_el = 0;
            const double c1 = __dmul_rn(dp1, id1); if (c1 > maxcos) { main_el = 1; maxcos = c1; }
            const double c2 = __dmul_rn(dp2, id2); if (c2 > maxcos) { main_el = 2; maxcos = c2; }
            const double c3 = __dmul_rn(dp3, id3); if (c3 > maxcos) { main_el = 3; maxcos = c3; }
            int nx;
            if (main_el == 0) {
                const float lx = __fmul_rn(vx0, iv0), ly = __fmul_rn(vy0, iv0);
                base_a = lx; base_b = ly;
                nx = (s0 + 1 == n) ? 0 : s0 + 1; s0 = nx;
                vx0 = hb.vx[nx]; vy0 = hb.vy[nx]; iv0 = hb.invf[nx]; id0 = hb.inv[nx]; px0 = px[nx]; py0 = py[nx];
            } else if (main_el == 1) {
                const float lx = __fmul_rn(vx1, iv1), ly = __fmul_rn(vy1, iv1);
                base_a = ly; base_b = -lx;
                nx = (s1 + 1 == n) ? 0 : s1 + 1; s1 = nx;
                vx1 = hb.vx[nx]; vy1 = hb.vy[nx]; iv1 = hb.invf[nx]; id1 = hb.inv[nx]; px1 = px[nx]; py1 = py[nx];
            } else if (main_el == 2) {
                const float lx = __fmul_rn(vx2, iv2), ly = __fmul_rn(vy2, iv2);
                base_a = -lx; base_b = -ly;
                nx = (s2 + 1 == n) ? 0 : s2 + 1; s2 = nx;
                vx2 = hb.vx[nx]; vy2 = hb.vy[nx]; iv2 = hb.invf[nx]; id2 = hb.inv[nx]; px2 = px[nx]; py2 = py[nx];
            } else {
                const float lx = __fmul_rn(vx3, iv3), ly = __fmul_rn(vy3, iv3);
                base_a = -ly; base_b = lx;
                nx = (s3 + 1 == n) ? 0 : s3 + 1; s3 = nx;
                vx3 = hb.vx[nx]; vy3 = hb.vy[nx]; iv3 = hb.invf[nx]; id3 = hb.inv[nx]; px3 = px[nx]; py3 = py[nx];
            }
            float dx = __fsub_rn(px1, px3), dy = __fsub_rn(py1, py3);
            const float width = __fadd_rn(__fmul_rn(dx, base_a), __fmul_rn(dy, base_b));
            dx = __fsub_rn(px2, px0); dy = __fsub_rn(py2, py0);
            const float height = __fadd_rn(__fmul_rn(-dx, base_b), __fmul_rn(dy, base_a));
            const float area = __fmul_rn(width, height);
            if (area <= minarea) {
                minarea = area; b_left = s3; b_a = base_a; b_w = width; b_b = base_b; b_h = height; b_bottom = s0;
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

// Per-row extremes of the component `root` (its runs' leftmost start / rightmost end per row) and its row range ->
// rowL / rowR [H], counts[2] = first row, counts[3] = last row.  Called by all threads of the CTA.
template <typename TG>
__device__ void component_rows(const RunSet& rs, int R, int root, int H, int* rowL, int* rowR, int* counts) {
    const int tid = TG::tid();
    if (tid == 0) { counts[0] = 0; counts[1] = 0; counts[2] = 0x7fffffff; counts[3] = -1; }
    for (int r = tid; r < H; r += TG::size()) { rowL[r] = 0x7fffffff; rowR[r] = -1; }
    TG::sync();
    for (int i0 = 0; i0 < R; i0 += TG::size()) {
        const int i = i0 + tid;
        int rlo = 0x7fffffff, rhi = -1;
        if (i < R && rs.parent[i] == root) {
            const int r = rs.row[i];
            atomicMin(&rowL[r], (int)rs.xs[i]); atomicMax(&rowR[r], (int)rs.xe[i]);
            rlo = rhi = r;
        }
        rlo = __reduce_min_sync(0xffffffffu, rlo); rhi = __reduce_max_sync(0xffffffffu, rhi);   // one pair of atomics per warp
        if ((threadIdx.x & 31) == 0 && rhi >= 0) { atomicMin(&counts[2], rlo); atomicMax(&counts[3], rhi); }
    }
    TG::sync();
}

// Strict convex hull of a component from its per-row extremes (rows rmin..rmax, every row occupied: a connected
// component has no empty row in between), in cv2.convexHull's output order (start at (max x, max y), then
// image-clockwise), followed by the calipers.  Called by all G threads of the group (a CTA of kT threads, or one warp).
// keep: int[nrows + 1] scratch; counts: int[>= 8 + kT/32 + 2] scratch.  Returns false if the hull outgrew `hull_cap`
// vertices (cannot happen below 1024^2 pixels with the CTA capacities).
template <int G, typename RT>
__device__ bool hull_rect(int rmin, int rmax, const RT* rowL, const RT* rowR, int* keep, const HullBuf& hb, int hull_cap,
                          int chain_cap, int* counts, float* out8) {
    const int tid = Grp<G>::tid();
    KB_T(4);
    const int nrows = rmax - rmin + 1;
    // Candidates of the two chains are the per-row extremes.  A candidate that lies on or inside ANY chord between two
    // other candidates of its chain is not a strict hull vertex, so a cheap parallel sieve (chords to the rows
    // +-1, 2, 4, ... 32 away) removes most of them exactly; the survivors are compacted in row order and the
    // monotone-chain stack -- the only sequential step -- runs over those, in place, one thread per chain.
    for (int r = tid; r < nrows + 1; r += G) keep[r] = 0;    // low half: right chain, high half: left chain
    if (tid == 0) { counts[0] = 0; counts[1] = 0; }
    Grp<G>::sync();
    for (int it = tid; it < 2 * nrows; it += G) {
        const int side = it >= nrows;                  // 0 right chain, 1 left chain
        const int r = rmin + (side ? it - nrows : it);
        const RT* src = side ? rowL : rowR;
        const int xi = src[r];
        bool kp = true;
        if (r != rmin && r != rmax) {
#pragma unroll
            for (int d = 1; d <= 32; d <<= 1) {
                const int a = max(r - d, rmin), c = min(r + d, rmax);
                const int cr = (xi - (int)src[a]) * (c - r) - (r - a) * ((int)src[c] - xi);
                kp &= side ? (cr < 0) : (cr > 0);      // same sign test as the stack's pop condition below
            }
        }
        if (kp) atomicAdd(&keep[r - rmin], side ? 0x10000 : 1);
    }
    Grp<G>::sync();
    KB_T(5);
    // exclusive scan of both survivor counts at once
    const int packed = (G == 32) ? scan_rows_warp(keep, nrows) : scan_rows(keep, nrows, counts + 8);
    const int sr_ = packed & 0xffff, sl_ = packed >> 16;
    if (sr_ > chain_cap || sl_ > chain_cap) return false;    // more sieve survivors than the group's chain buffers hold
    for (int k = tid; k < nrows; k += G) {
        const int before = keep[k], after = keep[k + 1];
        const int r = rmin + k;
        if ((after & 0xffff) != (before & 0xffff)) { hb.rx[before & 0xffff] = (int16_t)rowR[r]; hb.ry[before & 0xffff] = (int16_t)r; }
        if ((after >> 16) != (before >> 16)) { hb.lx[before >> 16] = (int16_t)rowL[r]; hb.ly[before >> 16] = (int16_t)r; }
    }
    Grp<G>::sync();
    KB_T(6);
    // the two chains walk the same code with different data: as two lanes of one warp they stay converged (the group's
    // cost is the longer chain, not the sum); on a CTA they sit in two different warps
    const int chain_b = (G == 32) ? 1 : 32;
    if (tid == 0 || tid == chain_b) {
        // right chain / left chain, top -> bottom; pop while the turn is not strictly convex.  The stack
        // is written over the survivor list itself (it never gets ahead of the read position); (ax,ay) / (bx,by)
        // mirror its two topmost entries so that a push costs no shared-memory read.
        const bool right = tid == 0;
        int16_t* X = right ? hb.rx : hb.lx; int16_t* Y = right ? hb.ry : hb.ly;
        const int ns = right ? sr_ : sl_;
        int n = 0, ax = 0, ay = 0, bx = 0, by = 0;
        for (int k = 0; k < ns; ++k) {
            const int cx = X[k], cy = Y[k];
            while (n >= 2) {
                const int cr = (bx - ax) * (cy - by) - (by - ay) * (cx - bx);
                if (right ? (cr <= 0) : (cr >= 0)) {
                    --n; bx = ax; by = ay;
                    if (n >= 2) { ax = X[n - 2]; ay = Y[n - 2]; }
                } else break;
            }
            X[n] = (int16_t)cx; Y[n] = (int16_t)cy; ++n;
            ax = bx; ay = by; bx = cx; by = cy;
        }
        counts[right ? 0 : 1] = n;
    }
    Grp<G>::sync();
    if (counts[0] + counts[1] > hull_cap) return false;
    const int nr = counts[0], nl = counts[1];
    // cyclic order: right chain downward, then left chain upward (dropping duplicated end points)
    const bool dup_bottom = hb.lx[nl - 1] == hb.rx[nr - 1] && hb.ly[nl - 1] == hb.ry[nr - 1];
    const bool dup_top = hb.lx[0] == hb.rx[0] && hb.ly[0] == hb.ry[0];
    // left-chain entries kept: indices [lo, hi] walked downward from hi
    const int hi = nl - 1 - (dup_bottom ? 1 : 0), lo = dup_top ? 1 : 0;
    int n = nr + max(0, hi - lo + 1);
    Grp<G>::sync();        // (the float hull may alias the sieve's scratch: everybody is done with `keep`)
    for (int i = tid; i < n; i += G) {
        if (i < nr) { hb.hx[i] = (float)hb.rx[i]; hb.hy[i] = (float)hb.ry[i]; }
        else { const int k = hi - (i - nr); hb.hx[i] = (float)hb.lx[k]; hb.hy[i] = (float)hb.ly[k]; }
    }
    Grp<G>::sync();
    // collinear triples can only survive across the chain joins of degenerate (1- or 2-row, needle-like) components:
    // detect in parallel, repair serially only then
    int bad = 0;
    if (n > 2)
        for (int i = tid; i < n; i += G) {
            const int a = (i + n - 1) % n, c = (i + 1) % n;
            const float cr = (hb.hx[i] - hb.hx[a]) * (hb.hy[c] - hb.hy[i]) - (hb.hy[i] - hb.hy[a]) * (hb.hx[c] - hb.hx[i]);
            bad |= (cr == 0.f);
        }
    if (Grp<G>::sync_or(bad)) {
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
        Grp<G>::sync();
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
    Grp<G>::sync();
    const int s0 = counts[5];
    if (s0) {   // rotate left by s0 through the (still unused) vx/vy scratch
        for (int i = tid; i < n; i += G) { hb.vx[i] = hb.hx[(i + s0) % n]; hb.vy[i] = hb.hy[(i + s0) % n]; }
        Grp<G>::sync();
        for (int i = tid; i < n; i += G) { hb.hx[i] = hb.vx[i]; hb.hy[i] = hb.vy[i]; }
        Grp<G>::sync();
    }
    if (n == 2 && hb.hx[0] == hb.hx[1] && hb.hy[0] == hb.hy[1]) n = 1;
    KB_T(7);
    calipers_box<G>(hb, n, counts, out8);
    Grp<G>::sync();
    KB_T(9);
#ifdef TDOD_KB_TIMING
    if (threadIdx.x == 0 && blockIdx.x < 8192) { g_kb_t[blockIdx.x * 32 + 10] = n; g_kb_t[blockIdx.x * 32 + 11] = nrows; }
#endif
    return true;
}

// the CTA form used by the rectangle kernels: rows of the component `root`, then its hull and calipers
__device__ bool component_rect(const RunSet& rs, int R, int root, int H, int* rowL, int* rowR, const HullBuf& hb,
                               int* counts, float* out8) {
    component_rows<CtaGrp>(rs, R, root, H, rowL, rowR, counts);
    const int rmin = counts[2], rmax = counts[3];
    __syncthreads();
    // rs.row_off [H+1] is free from here on: the sieve's scratch
    return hull_rect<kT, int>(rmin, rmax, rowL, rowR, rs.row_off, hb, kHullCap, 0x7fff, counts, out8);
}

struct Smem {
    uint32_t* bmA; uint32_t* bmB;     // exact passes only: the mask's bitmap and the flood-fill result
    RunSet rs;
    int* rowL; int* rowR;
    HullBuf hb;
    int* scratch;                     // [2*kT/32 + 48]
    float* out8;
};

__host__ __device__ inline size_t align16(size_t x) { return (x + 15) & ~size_t(15); }

// Shared-memory carve-up.  The hull buffers alias the run arrays: once the selected component's per-row extremes are
// in rowL/rowR the runs are dead (when the runs live in a global workspace slot the region simply holds the hull).
__host__ __device__ inline size_t smem_layout(int H, int Wp, bool two_bitmaps, Smem* s, unsigned char* base) {
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align16(off + bytes); return o; };
    const size_t o_bmA = two_bitmaps ? take((size_t)H * Wp * 4) : 0;
    const size_t o_bmB = two_bitmaps ? take((size_t)H * Wp * 4) : 0;
    const size_t o_roff = take((size_t)(H + 1) * 4);
    const size_t o_rowL = take((size_t)H * 4), o_rowR = take((size_t)H * 4);
    const size_t o_scr = take((2 * (kT / 32) + 48) * 4), o_out = take(8 * 4);
    const size_t region = off;
    // view 1: runs
    const size_t o_xs = take(kRunCap * 2), o_xe = take(kRunCap * 2), o_row = take(kRunCap * 2);
    const size_t o_par = take(kRunCap * 4), o_area = take(kRunCap * 4);
    const size_t end_runs = off;
    // view 2: hull (chain candidates are at most one per row and side; the hull itself at most kHullCap vertices)
    off = region;
    const size_t o_rx = take((size_t)H * 2), o_ry = take((size_t)H * 2), o_lx = take((size_t)H * 2), o_ly = take((size_t)H * 2);
    const size_t hn = (size_t)kHullCap;
    const size_t o_hx = take(hn * 4), o_hy = take(hn * 4), o_vx = take(hn * 4), o_vy = take(hn * 4), o_invf = take(hn * 4), o_inv = take(hn * 8);
    const size_t end_hull = off;
    off = end_runs > end_hull ? end_runs : end_hull;
    if (s) {
        s->bmA = two_bitmaps ? (uint32_t*)(base + o_bmA) : nullptr; s->bmB = two_bitmaps ? (uint32_t*)(base + o_bmB) : nullptr;
        s->rs.xs = (uint16_t*)(base + o_xs); s->rs.xe = (uint16_t*)(base + o_xe); s->rs.row = (uint16_t*)(base + o_row);
        s->rs.parent = (int*)(base + o_par); s->rs.area2 = (int*)(base + o_area); s->rs.row_off = (int*)(base + o_roff);
        s->rowL = (int*)(base + o_rowL); s->rowR = (int*)(base + o_rowR);
        s->hb.rx = (int16_t*)(base + o_rx); s->hb.ry = (int16_t*)(base + o_ry);
        s->hb.lx = (int16_t*)(base + o_lx); s->hb.ly = (int16_t*)(base + o_ly);
        s->hb.hx = (float*)(base + o_hx); s->hb.hy = (float*)(base + o_hy);
        s->hb.vx = (float*)(base + o_vx); s->hb.vy = (float*)(base + o_vy); s->hb.inv = (double*)(base + o_inv); s->hb.invf = (float*)(base + o_invf);
        s->scratch = (int*)(base + o_scr); s->out8 = (float*)(base + o_out);
    }
    return off;
}

// pick the component with the largest area2 (ties: the larger root index = later first pixel in raster order).
// In bounds mode also reports whether any OTHER component's hole-filled area could reach the winner's hole-free
// area (then hole filling could change the answer and pass 2 must decide).
template <typename TG>
__device__ int select_component(const RunSet& rs, int R, bool bounds_mode, int* scratch, bool* ambiguous) {
    // scratch[0] = best key hi (area2), scratch[1] = best root, via a two-step atomicMax on a packed 64-bit key
    unsigned long long* key = reinterpret_cast<unsigned long long*>(scratch + 20);   // 8-byte aligned slot
    if (TG::tid() == 0) { *key = 0ull; scratch[24] = 0; }
    TG::sync();
    for (int i = TG::tid(); i < R; i += TG::size())
        if (rs.parent[i] == i) atomicMax(key, ((unsigned long long)(unsigned)rs.area2[i] << 32) | (unsigned)i);
    TG::sync();
    const int best = (int)(*key & 0xffffffffu);
    const int best_a2 = (int)(*key >> 32);
    if (bounds_mode) {
        // bounding boxes per root: reuse area2 of non-root runs?  no -- compute on the fly per root with atomics
        // on four small arrays would cost smem; instead every run contributes to its root's extents held in
        // rowL/rowR-free scratch is not available here, so scan: a root's bbox is reduced by the threads that
        // own its runs through atomics on packed words stored in area2 of its FIRST non-root run ... too
        // intricate; simple exact alternative: for every non-best root, one thread walks all runs once.
        // Upper bound on a rival's hole-filled area: filling holes never leaves the per-row extent [L_r, R_r] of the
        // component, and the cell formula is monotone in the pixel set, so the area of the row-convexified component
        // bounds it.  One thread per rival walks the runs (raster order) once.
        for (int i = TG::tid(); i < R; i += TG::size()) {
            if (rs.parent[i] != i || i == best) continue;
            int pl = 0, pr = -1, cl = rs.xs[i], cr = rs.xe[i], crow = rs.row[i];
            bool have_prev = false;
            long long ub2 = 0;
            for (int j = i + 1; j <= R; ++j) {
                bool mine = j < R && rs.parent[j] == i;
                if (j < R && !mine) {
                    // runs are in raster order and a component has no empty row inside: a foreign run two rows below the
                    // component's last one means it is complete -- flush as at the end of the list
                    if ((int)rs.row[j] > crow + 1) j = R; else continue;
                }
                if (mine && (int)rs.row[j] == crow) { cl = min(cl, (int)rs.xs[j]); cr = max(cr, (int)rs.xe[j]); continue; }
                // row `crow` is complete: pair it with the previous one
                if (have_prev && pl <= cr && pr >= cl) ub2 += 2 * (min(cr, pr) - max(cl, pl)) + (cl != pl) + (cr != pr);
                pl = cl; pr = cr; have_prev = true;
                if (mine) { cl = rs.xs[j]; cr = rs.xe[j]; crow = rs.row[j]; }
            }
            // a rival that can at most TIE at area 0 cannot win either: ties go to the later component, which `best` is
            if (ub2 > (long long)best_a2 || (ub2 == (long long)best_a2 && ub2 > 0)) atomicOr(&scratch[24], 1);
        }
        TG::sync();
        *ambiguous = scratch[24] != 0;
    } else {
        *ambiguous = false;
    }
    TG::sync();
    return best;
}

// Front kernel: select_component (bounds mode) and component_rows in ONE sweep over the runs and two barriers.  The caller has
// already set *key = 0, scratch[24] = 0, counts[2] = INT_MAX, counts[3] = -1 and rowL / rowR to their sentinels, with a
// barrier of the group since (the labelling's).  Every run of the winner updates the per-row extremes; every other root
// walks its component once for the hole-filled upper bound (see select_component).
template <typename TG>
__device__ int pick_component_rows(const RunSet& rs, int R, int* rowL, int* rowR, int* scratch, bool* ambiguous) {
    unsigned long long* key = reinterpret_cast<unsigned long long*>(scratch + 20);
    int* counts = scratch + 32;
    for (int i = TG::tid(); i < R; i += TG::size())
        if (rs.parent[i] == i) atomicMax(key, ((unsigned long long)(unsigned)rs.area2[i] << 32) | (unsigned)i);
    TG::sync();
    const int best = (int)(*key & 0xffffffffu);
    const int best_a2 = (int)(*key >> 32);
    for (int i0 = 0; i0 < R; i0 += TG::size()) {
        const int i = i0 + TG::tid();
        int rlo = 0x7fffffff, rhi = -1;
        if (i < R) {
            const int par = rs.parent[i];
            if (par == best) {
                const int r = rs.row[i];
                atomicMin(&rowL[r], (int)rs.xs[i]); atomicMax(&rowR[r], (int)rs.xe[i]);
                rlo = rhi = r;
            } else if (par == i) {
                int pl = 0, pr = -1, cl = rs.xs[i], cr = rs.xe[i], crow = rs.row[i];
                bool have_prev = false;
                long long ub2 = 0;
                for (int j = i + 1; j <= R; ++j) {
                    const bool mine = j < R && rs.parent[j] == i;
                    if (j < R && !mine) {
                        if ((int)rs.row[j] > crow + 1) j = R; else continue;
                    }
                    if (mine && (int)rs.row[j] == crow) { cl = min(cl, (int)rs.xs[j]); cr = max(cr, (int)rs.xe[j]); continue; }
                    if (have_prev && pl <= cr && pr >= cl) ub2 += 2 * (min(cr, pr) - max(cl, pl)) + (cl != pl) + (cr != pr);
                    pl = cl; pr = cr; have_prev = true;
                    if (mine) { cl = rs.xs[j]; cr = rs.xe[j]; crow = rs.row[j]; }
                }
                if (ub2 > (long long)best_a2 || (ub2 == (long long)best_a2 && ub2 > 0)) atomicOr(&scratch[24], 1);
            }
        }
        rlo = __reduce_min_sync(0xffffffffu, rlo); rhi = __reduce_max_sync(0xffffffffu, rhi);   // one pair of atomics per warp
        if ((threadIdx.x & 31) == 0 && rhi >= 0) { atomicMin(&counts[2], rlo); atomicMax(&counts[3], rhi); }
    }
    TG::sync();
    *ambiguous = scratch[24] != 0;
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

// Workspace: [bitmap n*H*Wp words][rowrange n*2 ints][kSlots run-set slots for the overflow pass]
__host__ __device__ inline size_t ws_bitmap_bytes(int64_t n, int H, int W) { return align16((size_t)n * H * ((W + 31) >> 5) * 4); }
__host__ __device__ inline size_t ws_range_bytes(int64_t n) { return align16((size_t)n * 8); }
// work lists of passes 2 and 3: [count2, count3, pad, pad][list2 n ints][list3 n ints]
__host__ __device__ inline size_t ws_list_bytes(int64_t n) { return align16(16 + (size_t)n * 8); }

// kExact: fill holes before labelling (bitmap staged in shared memory).  kGlobal: the run arrays live in a workspace
// slot (masks with more runs than shared memory holds -- speckle noise).  Pass 1 (neither) launches one CTA per mask;
// the other two are persistent and walk the state bytes: 2 = ambiguous without hole filling, 3 = run overflow.
template <bool kExact, bool kGlobal>
__global__ void __launch_bounds__(kT, kExact ? 1 : TDOD_KB_MINB) rect_kernel(const uint32_t* __restrict__ bitmap, const int* __restrict__ rowrange,
                                                                  int64_t n_masks, int H, int W, float* __restrict__ out_rects,
                                                                  uint8_t* __restrict__ out_has, unsigned char* __restrict__ slots,
                                                                  int* __restrict__ lists, uint32_t* __restrict__ fill_slots) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int Wp = (W + 31) >> 5;
    Smem s;
    // masks too large for two shared-memory bitmaps (above ~0.75 Mpx): the flood reads the packed mask where it lies and
    // keeps its state in a per-CTA workspace slot (L2-resident); only the rare exact passes pay for it
    const bool bm_global = kExact && fill_slots != nullptr;
    smem_layout(H, Wp, kExact && !bm_global, &s, smem_raw);
    int cap = kRunCap;
    if (kGlobal) {
        const int64_t c = run_cap_global(H, W);
        unsigned char* b = slots + (size_t)blockIdx.x * slot_bytes(H, W);
        s.rs.parent = (int*)b; s.rs.area2 = (int*)(b + 4 * c);
        s.rs.xs = (uint16_t*)(b + 8 * c); s.rs.xe = (uint16_t*)(b + 10 * c); s.rs.row = (uint16_t*)(b + 12 * c);
        cap = (int)c;
    }
    int* list2 = lists + 4;
    int* list3 = lists + 4 + n_masks;
    const int64_t n_work = (!kExact && !kGlobal) ? n_masks : (int64_t)lists[kGlobal ? 1 : 0];
    for (int64_t wk = blockIdx.x; wk < n_work; wk += gridDim.x) {
        __syncthreads();                             // previous mask's shared state is dead
        const int64_t mi = (!kExact && !kGlobal) ? wk : (int64_t)(kGlobal ? list3 : list2)[wk];
        KB_T(0);
        const int r0 = rowrange[2 * mi], r1 = H - 1 - rowrange[2 * mi + 1];
        if (r0 > r1) {                               // empty mask: scorefunction.py:69-75 fallback applies
            if (threadIdx.x == 0) out_has[mi] = 0;
            continue;
        }
        const uint32_t* comp_bm = bitmap + (size_t)mi * H * Wp;
        if (kExact) {
            if (bm_global) {
                s.bmA = const_cast<uint32_t*>(comp_bm);
                s.bmB = fill_slots + (size_t)blockIdx.x * (align16((size_t)H * Wp * 4) / 4);
            } else {
                for (int i = threadIdx.x; i < H * Wp; i += kT) s.bmA[i] = comp_bm[i];
            }
            __syncthreads();
            flood_outside(s.bmA, s.bmB, H, W, Wp, s.scratch + 30);
            const uint32_t tail = (W & 31) ? ((1u << (W & 31)) - 1u) : 0xffffffffu;
            for (int i = threadIdx.x; i < H * Wp; i += kT) {
                const int w = i % Wp;
                s.bmB[i] = ~s.bmB[i] & ((w == Wp - 1) ? tail : 0xffffffffu);    // M = not outside (foreground + holes)
            }
            __syncthreads();
            comp_bm = s.bmB;
        }
        KB_T(1);
        const int R = label_runs(GlobalBits{comp_bm, Wp}, H, W, Wp, s.rs, s.scratch, cap, r0, r1);
        KB_T(2);
        if (R < 0) {                                    // more runs than shared memory holds -> overflow pass
            if (threadIdx.x == 0) { out_has[mi] = 3; list3[atomicAdd(&lists[1], 1)] = (int)mi; }
            continue;
        }
        bool ambiguous = false;
        const int root = select_component<CtaGrp>(s.rs, R, !kExact, s.scratch, &ambiguous);
        KB_T(3);
        if (ambiguous) {
            if (threadIdx.x == 0) { out_has[mi] = 2; list2[atomicAdd(&lists[0], 1)] = (int)mi; }
            continue;
        }
        const bool ok = component_rect(s.rs, R, root, H, s.rowL, s.rowR, s.hb, s.scratch + 32, s.out8);
        if (threadIdx.x < 8 && ok) out_rects[mi * 8 + threadIdx.x] = s.out8[threadIdx.x];
        if (threadIdx.x == 0) out_has[mi] = ok ? 1 : 4;
    }
}

// ------------------------------------------------------------------------------------------------------------
// Fused path (masks whose bit stream and run arrays fit in shared memory): the 1 bit/px copy never leaves the SM.
//
//   front kernel  persistent CTAs, two per SM.  Per mask: the raw bytes arrive in tiles of whole rows through a kRing-slot
//                 TMA ring (cp.async.bulk + mbarrier; the loads of the next tiles -- also the next mask's first -- in flight
//                 while this one is turned into bits), 16 bytes -> 16 bits -> one half-word store into the ROW SLOT of a
//                 bit stream in shared memory (every row owns a 16-byte-aligned slot, see front_slot_hw); then run
//                 extraction, labelling, cell-formula areas and the choice of the component straight from those slots.
//                 What leaves the SM is the winner's per-row extremes: two int16 per occupied row (~1 KB per mask
//                 instead of a 48 KB bitmap written and read back).
//   tail kernel   ONE WARP per mask: sieve, monotone-chain stacks, rotating calipers from those extremes.  The serial
//                 chains of 12-16 masks overlap on an SM; nothing waits at a CTA barrier.
// Masks the front kernel cannot decide (hole filling could change the winner; more runs than shared memory holds) go
// to the work lists; a list-driven pack writes only THEIR bitmaps and the exact passes (rect_kernel<true, *>) finish them.
// ------------------------------------------------------------------------------------------------------------
#ifndef TDOD_KB_RING
#define TDOD_KB_RING 3
#endif
#ifndef TDOD_KB_STREAMS
#define TDOD_KB_STREAMS 3
#endif
constexpr int kRing = TDOD_KB_RING;     // TMA ring slots of a front CTA (loads of kRing - 1 tiles in flight behind the one being packed)
constexpr int kSB = TDOD_KB_STREAMS;    // bit-stream buffers (row groups the pack half may run ahead of the extract half)
constexpr int kTailHullCap = 192;       // hull vertices one tail warp holds (beyond: the overflow pass)
constexpr int kTailChainCap = 256;      // sieve survivors per chain it holds (beyond: the overflow pass)
constexpr int kTailWarps = 4;
constexpr uint8_t kPending = 0xF0;      // state byte between the front and the tail kernel

__host__ __device__ inline size_t cand_stride(int H) { return align16(8 + (size_t)H * 4); }   // {rmin, rmax} + int16 L/R per row

// Front CTA = two halves that never meet at a CTA barrier.  PACK half (threads 0..127, thread 0 also drives the TMA ring):
// bytes -> bits, group of rows after group of rows, into one of kSB bit-stream buffers.  EXTRACT half (threads 128..383):
// runs of a finished group (two threads per row), and after a mask's last group its labels, areas, winner and the record.
// The halves hand over through mbarriers (stream full / stream free), so while the extract half works on mask k the pack
// half keeps HBM streaming for mask k + 1: the kernel's loads never pause for the labelling.
#ifdef TDOD_KB_PROF
__device__ unsigned long long g_kb_prof[16];
#define KBP_DECL long long kbp_t = clock64(); long long kbp_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#define KBP(k) do { const long long t__ = clock64(); kbp_acc[k] += t__ - kbp_t; kbp_t = t__; } while (0)
#define KBP_FLUSH(base, leader) do { if (leader) for (int q__ = 0; q__ < 8; ++q__) atomicAdd(&g_kb_prof[(base) + q__], (unsigned long long)kbp_acc[q__]); } while (0)
#else
#define KBP_DECL
#define KBP(k) do { } while (0)
#define KBP_FLUSH(base, leader) do { } while (0)
#endif
constexpr int kHalf = 128;              // pack threads
constexpr int kExT = 256;               // extract / label threads
constexpr int kGroupRows = kExT / 2;    // rows per bit-stream buffer: two extract threads per row
constexpr int kTileChunks = 768;        // 16-byte chunks per TMA tile (12 KB): six per pack thread
using ExGrp = SubGrp<1, kExT, kHalf>;

// Row slots of the bit stream: row r of a mask starts at byte off_m + r W of the 16-byte-aligned base, i.e. at bit
// dl = (off_m + r W) & 15 of the 16-byte chunk that holds its first pixel.  The pack half converts, for every row, the
// chunks that cover it (first to last, the straddling chunk of two neighbouring rows twice) into that row's OWN slot of
// 16 QH half-words, so the row's pixels sit at bits [dl, dl + W) of a 16-byte-aligned slot: the extract half reads whole
// quads with LDS.128 and never funnel-shifts; the bits outside the row are cleared in registers.
__host__ __device__ inline int front_slot_hw(int W) { return (((W + 30) >> 4) + 15) & ~15; }
__host__ __device__ inline int front_tile_rows(int W) {             // rows per TMA tile: 768 (row, chunk) pairs, 12 KB of bytes
    int tr = kTileChunks / front_slot_hw(W);
    const int by_bytes = (kTileChunks * 16 - 32) / W;
    if (by_bytes < tr) tr = by_bytes;
    return tr < kGroupRows ? tr : kGroupRows;
}

struct FrontSmem {
    unsigned char* raw0; size_t raw_stride;            // ring slot k at raw0 + k * raw_stride
    unsigned char* sb0; size_t sb_stride;               // kSB bit-stream buffers (a group of rows each): sb0 + b * sb_stride
    RunSet rs; int* rowL; int* rowR; int* scratch;
};
__host__ __device__ inline size_t front_layout(int H, int W, FrontSmem* f, unsigned char* base) {
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = (off + bytes + 127) & ~size_t(127); return o; };
    const size_t o_raw = take((size_t)kTileChunks * 16);
    const size_t raw_stride = off - o_raw;
    for (int k = 1; k < kRing; ++k) take((size_t)kTileChunks * 16);
    const size_t sbytes = (size_t)kGroupRows * front_slot_hw(W) * 2;        // one 16-byte-aligned slot per row
    const size_t o_sb0 = take(sbytes), o_sb1 = take(sbytes);
    for (int k = 2; k < kSB; ++k) take(sbytes);
    const size_t o_xs = take(kRunCap * 2), o_xe = take(kRunCap * 2), o_row = take(kRunCap * 2);
    const size_t o_par = take(kRunCap * 4), o_area = take(kRunCap * 4);
    const size_t o_roff = take((size_t)(H + 1) * 4), o_rl = take((size_t)H * 4), o_rr = take((size_t)H * 4);
    const size_t o_scr = take((2 * (kT / 32) + 48) * 4);
    if (f) {
        f->raw0 = base + o_raw; f->raw_stride = raw_stride;
        f->sb0 = base + o_sb0; f->sb_stride = o_sb1 - o_sb0;
        f->rs.xs = (uint16_t*)(base + o_xs); f->rs.xe = (uint16_t*)(base + o_xe); f->rs.row = (uint16_t*)(base + o_row);
        f->rs.parent = (int*)(base + o_par); f->rs.area2 = (int*)(base + o_area); f->rs.row_off = (int*)(base + o_roff);
        f->rowL = (int*)(base + o_rl); f->rowR = (int*)(base + o_rr); f->scratch = (int*)(base + o_scr);
    }
    return off;
}

// 16 mask bytes that are all 0 or 1 (torch.bool storage) -> 16 bits.  Two words are merged first (word b's flags moved
// to bit 4 of each byte: a + 16 b), then ONE multiply gathers the eight flags: a flag at bit 8k (+4) goes to bit 24 + k
// (28 + k) through the multiplier term 2^(24 - 7k); every other partial product lands below bit 24 or above bit 31, on
// pairwise distinct bits, so nothing carries into the byte.  `other` collects every bit that is NOT a bit 0 of a byte:
// non-zero means the mask holds bytes above 1 and must take the general (normalising) path.
__device__ __forceinline__ uint32_t bits16_of_01(const uint4 v, uint32_t& other) {
    other |= (v.x | v.y) | (v.z | v.w);
    const uint32_t p01 = (v.y * 16u + v.x) * 0x01020408u, p23 = (v.w * 16u + v.z) * 0x01020408u;
    return (p01 >> 24) | ((p23 >> 16) & 0xFF00u);
}
__device__ __forceinline__ void mbar_arrive(uint64_t* b) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}

// W <= 1024.  All positions inside a mask are 32-bit byte offsets from its 16-byte-aligned base.  QH = quads (16 bytes of
// bit stream) per half row: a row slot is 2 QH quads = 16 QH half-words = front_slot_hw(W).
template <int QH>
__global__ void __launch_bounds__(kHalf + kExT, 2) kb_front_kernel(const uint8_t* __restrict__ masks, int64_t n_masks, int H, int W,
                                                               unsigned char* __restrict__ cand, uint8_t* __restrict__ out_has,
                                                               int* __restrict__ lists) {
    constexpr int CR = 16 * QH;                         // half-words (= chunks) per row slot
    constexpr int WH = 4 * QH;                          // words per half row
    extern __shared__ __align__(128) unsigned char fs_raw[];
    __shared__ __align__(8) uint64_t mb_full[kRing];   // the tile's bytes have landed (TMA transaction count)
    __shared__ __align__(8) uint64_t mb_free[kRing];   // all pack threads have turned the slot's bytes into bits
    __shared__ __align__(8) uint64_t sb_full[kSB];       // the group's bit stream is complete (pack threads arrive)
    __shared__ __align__(8) uint64_t sb_free[kSB];       // its runs have been extracted (extract threads arrive)
    __shared__ int s_warp[kExT / 32];
    __shared__ uint32_t s_other[8];                     // per mask (index m' & 7): some byte of the mask is above 1
    FrontSmem f;
    front_layout(H, W, &f, fs_raw);
    const int tid = threadIdx.x;
    const uintptr_t base = reinterpret_cast<uintptr_t>(masks);
    const int64_t HW = (int64_t)H * W;
    const int TR = front_tile_rows(W);                  // rows per tile
    if (tid == 0) {
        for (int k = 0; k < kRing; ++k) { mbar_init(&mb_full[k], 1); mbar_init(&mb_free[k], kHalf); }
        for (int k = 0; k < kSB; ++k) { mbar_init(&sb_full[k], kHalf); mbar_init(&sb_free[k], kExT); }
        for (int k = 0; k < 8; ++k) s_other[k] = 0u;
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();

    if (tid < kHalf) {
        // =============================== PACK half ===============================
        // producer state (thread 0): the next tile to request, walking (mask, rows) in the order the tiles are consumed.
        // A tile = up to TR rows of one row group = ONE contiguous run of chunks (neighbouring tiles share a chunk).
        int64_t pm = blockIdx.x;
        int pr0 = 0, issued = 0;
        auto issue_next = [&]() {                       // thread 0 only
            if (pm >= n_masks) return;
            const uintptr_t a = base + (uintptr_t)(pm * HW);
            const int off_m = (int)(a & 15);
            const int r1 = min(min(H, (pr0 / kGroupRows + 1) * kGroupRows), pr0 + TR);
            const int c0 = (off_m + pr0 * W) >> 4, c1 = (off_m + r1 * W + 15) >> 4;
            const int slot = issued % kRing;
            // refilling a slot: its previous tile (number issued - kRing) must have been packed by every pack thread
            if (issued >= kRing) mbar_wait(&mb_free[slot], (uint32_t)(issued / kRing - 1) & 1u);
            mbar_expect_tx(&mb_full[slot], (uint32_t)(c1 - c0) * 16u);
            bulk_load(f.raw0 + slot * f.raw_stride, reinterpret_cast<const unsigned char*>(a - off_m) + ((size_t)c0 << 4),
                      (uint32_t)(c1 - c0) * 16u, &mb_full[slot]);
            ++issued;
            pr0 = r1;
            if (pr0 >= H) { pr0 = 0; pm += gridDim.x; }
        };
        if (tid == 0)
            for (int k = 0; k < kRing - 1; ++k) issue_next();
        // this thread's six (row of the tile, chunk of the row) pairs: the same in every tile
        constexpr int kPairs = kTileChunks / kHalf;
        int rw_u[kPairs], k_u[kPairs];
#pragma unroll
        for (int u = 0; u < kPairs; ++u) {
            const int pp = tid + u * kHalf, rr = pp / CR;
            k_u[u] = pp - rr * CR; rw_u[u] = rr * W;
        }
        int g = 0, gs = 0;                              // tiles / groups handled so far by this CTA
        int mi = 0;                                     // masks handled so far
        KBP_DECL
        for (int64_t m = blockIdx.x; m < n_masks; m += gridDim.x, ++mi) {
            uint32_t other = 0u;
            const int off_m = (int)((base + (uintptr_t)(m * HW)) & 15);
            for (int gr0 = 0; gr0 < H; gr0 += kGroupRows, ++gs) {
                const int gr1 = min(H, gr0 + kGroupRows);
                const int b = gs % kSB;
                // the buffer's previous group (kSB back) must have been read by the extract half
                KBP(0);
                if (gs >= kSB) mbar_wait(&sb_free[b], (uint32_t)(gs / kSB - 1) & 1u);
                KBP(1);                                 // 1: waiting for the extract half to free a stream buffer
                uint16_t* stream = reinterpret_cast<uint16_t*>(f.sb0 + b * f.sb_stride);
                for (int tr0 = gr0; tr0 < gr1; tr0 += TR, ++g) {
                    const int tr1 = min(gr1, tr0 + TR);
                    const int slot = g % kRing;
                    if (tid == 0) issue_next();         // tile g + kRing - 1
                    const int t0 = off_m + tr0 * W;
                    const int nch = ((off_m + tr1 * W + 15) >> 4) - (t0 >> 4), off_t = t0 & 15;
                    const int np = (tr1 - tr0) * CR;    // pairs of this tile
                    KBP(0);
                    mbar_wait(&mb_full[slot], (uint32_t)(g / kRing) & 1u);
                    KBP(2);                             // 2: waiting for the tile's bytes
                    const uint4* rv = reinterpret_cast<const uint4*>(f.raw0 + slot * f.raw_stride);
                    uint16_t* dst = stream + (tr0 - gr0) * CR;
                    // all six loads first (index clamped, so always inside the tile): six LDS.128 in flight per thread
                    uint4 v[kPairs];
#pragma unroll
                    for (int u = 0; u < kPairs; ++u) v[u] = rv[min(((off_t + rw_u[u]) >> 4) + k_u[u], nch - 1)];
#pragma unroll
                    for (int u = 0; u < kPairs; ++u) {
                        const int pp = tid + u * kHalf;
                        const uint32_t bits = bits16_of_01(v[u], other);     // (chunks past the row's end: cleared by the reader)
                        if (pp < np) dst[pp] = (uint16_t)bits;
                    }
                    mbar_arrive(&mb_free[slot]);
                }
                if (gr1 == H) {                         // last group of the mask: publish whether any byte was above 1
                    other &= 0xFEFEFEFEu;
                    if (__any_sync(0xffffffffu, other != 0u) && (tid & 31) == 0) atomicOr(&s_other[mi & 7], 1u);
                }
                mbar_arrive(&sb_full[b]);               // (release: the stream and the flag are visible to whoever waits)
            }
        }
        KBP(0);                                         // 0: packing (and issuing)
        KBP_FLUSH(0, tid == 32);
        return;
    }

    // =============================== EXTRACT half ===============================
    const int et = tid - kHalf, lane = et & 31, wid = et >> 5;
    int* list2 = lists + 4;
    int* list3 = lists + 4 + n_masks;
    int gs = 0, mi = 0;
    KBP_DECL
    for (int64_t m = blockIdx.x; m < n_masks; m += gridDim.x, ++mi) {
        int Rbase = 0;                                  // runs emitted so far (uniform over the half)
        const int off_m = (int)((base + (uintptr_t)(m * HW)) & 15);
        for (int gr0 = 0; gr0 < H; gr0 += kGroupRows, ++gs) {
            const int gr1 = min(H, gr0 + kGroupRows);
            const int b = gs % kSB;
            KBP(3);
            mbar_wait(&sb_full[b], (uint32_t)(gs / kSB) & 1u);
            KBP(4);                                     // 4: waiting for the pack half
            const uint32_t* sb32 = reinterpret_cast<const uint32_t*>(f.sb0 + b * f.sb_stride);
            // ---- runs of the group's rows, TWO THREADS PER ROW, each holding its half of the row's slot (QH quads) in
            // registers.  Thread order = raster order of the halves, so one scan of the packed (starts, ends) counts places
            // every half in the run arrays: the k-th start of a row pairs with its k-th end wherever the halves meet ----
            const int rloc = et >> 1, half = et & 1;
            const int r = gr0 + rloc;
            const bool live = r < gr1;
            const int dl = (off_m + r * W) & 15;        // the row's pixels are bits [dl, dl + W) of its slot
            const int lim = dl + W;
            const uint32_t* slot32 = sb32 + rloc * (2 * WH);
            uint32_t mw[WH];
            uint32_t acc = 0u;
            if (live) {
                const uint4* sq = reinterpret_cast<const uint4*>(slot32) + half * QH;
#pragma unroll
                for (int q = 0; q < QH; ++q) {
                    const uint4 v4 = sq[q];
                    mw[4 * q] = v4.x; mw[4 * q + 1] = v4.y; mw[4 * q + 2] = v4.z; mw[4 * q + 3] = v4.w;
                    acc |= (v4.x | v4.y) | (v4.z | v4.w);
                }
            }
            // most halves are empty (rows above / below the object, words beside it); bits of the neighbouring rows in the
            // edge words can only make a half look occupied, which costs the walk and nothing else
            int ns = 0, ne = 0;
            uint32_t carry_in = 0u, after = 0u;
            if (acc != 0u) {
                // clear what is not this row: bits below dl in the slot's first word, bits from lim on (a real branch per
                // quad: only the last quad(s) of the second half can reach lim)
                if (half == 0) mw[0] &= 0xffffffffu << dl;
#pragma unroll
                for (int q = 0; q < QH; ++q) {
                    const int w0 = half * WH + 4 * q;   // slot word of the quad's first word
                    if (32 * w0 + 128 > lim) {
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const int rem = lim - 32 * (w0 + k);
                            if (rem < 32) mw[4 * q + k] = rem <= 0 ? 0u : (mw[4 * q + k] & ((1u << rem) - 1u));
                        }
                    }
                }
                if (half) carry_in = (32 * WH - 1 < lim) ? slot32[WH - 1] >> 31 : 0u;     // (bit 32 WH - 1 >= 127 > dl)
                uint32_t carry = carry_in;
#pragma unroll
                for (int k = 0; k < WH; ++k) {
                    ns += __popc(mw[k] & ~((mw[k] << 1) | carry));
                    carry = mw[k] >> 31;
                }
                // ends inside the half = starts inside it + a run open at its left edge - a run open at its right edge
                if (carry && half == 0 && 32 * WH < lim) after = slot32[WH] & 1u;
                ne = ns + (int)(carry_in & mw[0] & 1u) - (int)(carry & after);
            }
            const int pk = ns | (ne << 16);
            int inc = pk;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t2 = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += t2;
            }
            if (lane == 31) s_warp[wid] = inc;
            ExGrp::sync();
            int before = 0, total = 0;
#pragma unroll
            for (int w2 = 0; w2 < kExT / 32; ++w2) { if (w2 < wid) before += s_warp[w2]; total += s_warp[w2]; }
            total &= 0xffff;                            // starts == ends over whole rows
            const int excl = before + inc - pk;
            if (live) {
                if (half == 0) f.rs.row_off[r] = Rbase + (excl & 0xffff);
                if (pk) {
                    int ks = Rbase + (excl & 0xffff), ke = Rbase + (excl >> 16);
                    uint32_t carry = carry_in;
                    const int xb = 32 * half * WH - dl - 1;                  // x of bit i of word k = xb + 32 k + ffs
#pragma unroll
                    for (int k = 0; k < WH; ++k) {
                        const uint32_t w = mw[k], nx = k + 1 < WH ? mw[k + 1 < WH ? k + 1 : k] : after;
                        uint32_t st = w & ~((w << 1) | carry), en = w & ~((w >> 1) | (nx << 31));
                        for (; st; st &= st - 1, ++ks)
                            if (ks < kRunCap) { f.rs.xs[ks] = (uint16_t)(xb + 32 * k + __ffs(st)); f.rs.row[ks] = (uint16_t)r; }
                        for (; en; en &= en - 1, ++ke)
                            if (ke < kRunCap) f.rs.xe[ke] = (uint16_t)(xb + 32 * k + __ffs(en));
                        carry = w >> 31;
                    }
                }
            }
            Rbase += total;
            mbar_arrive(&sb_free[b]);                   // this thread is done with the stream buffer
            ExGrp::sync();                              // s_warp is free again; the runs are visible to the half
            KBP(5);                                     // 5: run extraction
        }
        const int R = Rbase;
        const bool not01 = s_other[mi & 7] != 0u;       // (visible: ordered by the wait on the last group's sb_full)
        if (et == 0) f.rs.row_off[H] = R;
        ExGrp::sync();
        if (et == 0) s_other[mi & 7] = 0u;              // reused eight masks on; the pack half is at most kSB groups ahead
        if (not01 || R == 0 || R > kRunCap) {
            if (et == 0) {
                if (not01 || R > kRunCap) { out_has[m] = 3; list3[atomicAdd(&lists[1], 1)] = (int)m; }   // general / overflow pass
                else out_has[m] = 0;                                                        // empty mask: scorefunction.py:69-75
            }
            ExGrp::sync();
            continue;
        }
        KBP(3);
        // state of the winner search: set here, published by the labelling's barriers; the previous mask's readers of
        // rowL / rowR / scratch are at least one extraction barrier behind
        if (et == 0) {
            *reinterpret_cast<unsigned long long*>(f.scratch + 20) = 0ull;
            f.scratch[24] = 0; f.scratch[32 + 2] = 0x7fffffff; f.scratch[32 + 3] = -1;
        }
        for (int rr = et; rr < H; rr += kExT) { f.rowL[rr] = 0x7fffffff; f.rowR[rr] = -1; }
        label_from_runs<ExGrp>(f.rs, R);
        KBP(6);                                         // 6: labelling
        bool ambiguous = false;
        pick_component_rows<ExGrp>(f.rs, R, f.rowL, f.rowR, f.scratch, &ambiguous);
        KBP(7);                                         // 7: winner, rival bounds, per-row extremes
        if (ambiguous) {                                // hole filling could change the winner: the exact pass decides
            if (et == 0) { out_has[m] = 2; list2[atomicAdd(&lists[0], 1)] = (int)m; }
            continue;
        }
        const int rmin = f.scratch[32 + 2], rmax = f.scratch[32 + 3];
        unsigned char* rec = cand + (size_t)m * cand_stride(H);
        int16_t* lr = reinterpret_cast<int16_t*>(rec + 8);
        const int nrows = rmax - rmin + 1;
        for (int rr = et; rr < nrows; rr += kExT) {
            lr[2 * rr] = (int16_t)f.rowL[rmin + rr]; lr[2 * rr + 1] = (int16_t)f.rowR[rmin + rr];
        }
        if (et == 0) { reinterpret_cast<int*>(rec)[0] = rmin; reinterpret_cast<int*>(rec)[1] = rmax; out_has[m] = kPending; }
        KBP(3);                                         // 3: the record
    }
    KBP_FLUSH(8, et == 0);
}

struct TailSmem { int16_t* L; int16_t* R; int* keep; HullBuf hb; int* counts; float* out8; };
__host__ __device__ inline size_t tail_layout(int H, TailSmem* t, unsigned char* base) {
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align16(off + bytes); return o; };
    const size_t o_cnt = take(24 * 4), o_out = take(8 * 4);
    const size_t cc = (size_t)(H < kTailChainCap ? H : kTailChainCap);
    const size_t o_rx = take(cc * 2), o_ry = take(cc * 2), o_lx = take(cc * 2), o_ly = take(cc * 2);
    const size_t region = off;
    const size_t o_L = take((size_t)H * 2), o_R = take((size_t)H * 2), o_keep = take((size_t)(H + 1) * 4);
    const size_t end_a = off;
    off = region;                                      // the float hull aliases the sieve's inputs (dead once compacted)
    const size_t hn = (size_t)kTailHullCap;
    const size_t o_inv = take(hn * 8), o_hx = take(hn * 4), o_hy = take(hn * 4), o_vx = take(hn * 4), o_vy = take(hn * 4), o_invf = take(hn * 4);
    off = off > end_a ? off : end_a;
    if (t) {
        t->counts = (int*)(base + o_cnt); t->out8 = (float*)(base + o_out);
        t->hb.rx = (int16_t*)(base + o_rx); t->hb.ry = (int16_t*)(base + o_ry); t->hb.lx = (int16_t*)(base + o_lx); t->hb.ly = (int16_t*)(base + o_ly);
        t->L = (int16_t*)(base + o_L); t->R = (int16_t*)(base + o_R); t->keep = (int*)(base + o_keep);
        t->hb.inv = (double*)(base + o_inv); t->hb.hx = (float*)(base + o_hx); t->hb.hy = (float*)(base + o_hy);
        t->hb.vx = (float*)(base + o_vx); t->hb.vy = (float*)(base + o_vy); t->hb.invf = (float*)(base + o_invf);
    }
    return off;
}

__global__ void __launch_bounds__(32 * kTailWarps) kb_tail_kernel(const unsigned char* __restrict__ cand, int64_t n_masks, int H,
                                                                float* __restrict__ out_rects, uint8_t* __restrict__ out_has,
                                                                int* __restrict__ lists, size_t per_warp) {
    extern __shared__ __align__(16) unsigned char ts_raw[];
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t m = (int64_t)blockIdx.x * kTailWarps + wid;
    if (m >= n_masks) return;
    if (out_has[m] != kPending) return;                // empty, or handed to the exact passes
    TailSmem t;
    tail_layout(H, &t, ts_raw + (size_t)wid * per_warp);
    const unsigned char* rec = cand + (size_t)m * cand_stride(H);
    const int rmin = reinterpret_cast<const int*>(rec)[0], rmax = reinterpret_cast<const int*>(rec)[1];
    const int nrows = rmax - rmin + 1;
    const uint32_t* lr = reinterpret_cast<const uint32_t*>(rec + 8);
    for (int r = lane; r < nrows; r += 32) {
        const uint32_t v = lr[r];
        t.L[r] = (int16_t)(v & 0xffffu); t.R[r] = (int16_t)(v >> 16);
    }
    __syncwarp();
    // the hull code indexes the extremes by absolute row
    const bool ok = hull_rect<32, int16_t>(rmin, rmax, t.L - rmin, t.R - rmin, t.keep, t.hb, kTailHullCap, kTailChainCap, t.counts,
                                           t.out8);
    if (ok) {
        if (lane < 8) out_rects[m * 8 + lane] = t.out8[lane];
        if (lane == 0) out_has[m] = 1;
    } else if (lane == 0) {                             // a hull beyond the warp's capacity: the overflow pass has room
        int* list3 = lists + 4 + n_masks;
        out_has[m] = 3; list3[atomicAdd(&lists[1], 1)] = (int)m;
    }
}

// pack only the masks on the two work lists (what the exact passes will read); grid-stride over (listed mask, tile)
__global__ void __launch_bounds__(kPackT, 8) pack_list_kernel(const uint8_t* __restrict__ masks, int64_t n_masks, int H, int W, int Wp,
                                                             uint32_t* __restrict__ bitmap, int* __restrict__ rowrange,
                                                             const int* __restrict__ lists, int tiles_per_mask) {
    extern __shared__ __align__(16) unsigned char pk_raw[];
    uint16_t* sb16 = reinterpret_cast<uint16_t*>(pk_raw);
    const uint32_t* sb32 = reinterpret_cast<const uint32_t*>(pk_raw);
    __shared__ int s_lo, s_hi;
    const int tid = threadIdx.x;
    const uintptr_t base = reinterpret_cast<uintptr_t>(masks);
    const int64_t mis = (int64_t)(base & 15);
    const uint4* v = reinterpret_cast<const uint4*>(base - mis);
    const int64_t HW = (int64_t)H * W;
    const int n2 = lists[0], n3 = lists[1];
    const int* list2 = lists + 4;
    const int* list3 = lists + 4 + n_masks;
    const int64_t n_tiles = (int64_t)(n2 + n3) * tiles_per_mask;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int li = (int)(tile / tiles_per_mask);
        const int64_t m = li < n2 ? list2[li] : list3[li - n2];
        const int r0 = (int)(tile - (int64_t)li * tiles_per_mask) * kPackRows, r1 = min(H, r0 + kPackRows);
        const int64_t b0 = mis + m * HW + (int64_t)r0 * W, b1 = mis + m * HW + (int64_t)r1 * W;
        const int64_t c0 = b0 >> 4;
        const int nch = (int)(((b1 + 15) >> 4) - c0);
        if (tid == 0) { s_lo = 0x7fffffff; s_hi = -1; }
        for (int c = tid; c < nch; c += kPackT) sb16[c] = (uint16_t)nz_bits16(__ldcs(v + c0 + c));
        if (tid < 4) sb16[nch + tid] = 0;
        __syncthreads();
        const int nwords = (r1 - r0) * Wp;
        int lo = 0x7fffffff, hi = -1;
        for (int q = tid; q < nwords; q += kPackT) {
            const int rr = q / Wp, w = q - rr * Wp, r = r0 + rr;
            const int o = (int)(b0 - (c0 << 4)) + rr * W + 32 * w;
            uint32_t word = __funnelshift_r(sb32[o >> 5], sb32[(o >> 5) + 1], o & 31);
            const int nb = W - 32 * w;
            if (nb < 32) word &= (1u << nb) - 1u;
            bitmap[((size_t)m * H + r) * Wp + w] = word;
            if (word) { lo = min(lo, r); hi = max(hi, r); }
        }
        lo = __reduce_min_sync(0xffffffffu, lo); hi = __reduce_max_sync(0xffffffffu, hi);
        if ((tid & 31) == 0 && hi >= 0) { atomicMin(&s_lo, lo); atomicMax(&s_hi, hi); }
        __syncthreads();
        if (tid == 0 && s_hi >= 0) {
            atomicMin(&rowrange[2 * m], s_lo);
            atomicMin(&rowrange[2 * m + 1], H - 1 - s_hi);
        }
        __syncthreads();
    }
}

// occupied row range of every mask of a caller-supplied bitmap (what the pack kernels record on the fly)
__global__ void __launch_bounds__(256) rowrange_kernel(const uint32_t* __restrict__ bitmap, int H, int Wp, int* __restrict__ rowrange) {
    const int64_t m = blockIdx.x;
    const uint32_t* bm = bitmap + (size_t)m * H * Wp;
    int lo = 0x7fffffff, hi = -1;
    for (int i = threadIdx.x; i < H * Wp; i += 256)
        if (bm[i]) { const int r = i / Wp; lo = min(lo, r); hi = max(hi, r); }
    lo = __reduce_min_sync(0xffffffffu, lo); hi = __reduce_max_sync(0xffffffffu, hi);
    __shared__ int s_lo, s_hi;
    if (threadIdx.x == 0) { s_lo = 0x7fffffff; s_hi = -1; }
    __syncthreads();
    if ((threadIdx.x & 31) == 0 && hi >= 0) { atomicMin(&s_lo, lo); atomicMax(&s_hi, hi); }
    __syncthreads();
    if (threadIdx.x == 0) {
        rowrange[2 * m] = s_hi >= 0 ? s_lo : 0x7f7f7f7f;
        rowrange[2 * m + 1] = s_hi >= 0 ? H - 1 - s_hi : 0x7f7f7f7f;
    }
}

}  // namespace

constexpr int kExactCtas = 148;   // persistent CTAs of the exact pass

// true when the exact passes cannot hold two H x Wp bitmaps in shared memory
static bool exact_needs_global(int H, int W) { return smem_layout(H, (W + 31) >> 5, true, nullptr, nullptr) > 227 * 1024; }
static size_t ws_fill_bytes(int H, int W) {
    return exact_needs_global(H, W) ? (size_t)kExactCtas * align16((size_t)H * ((W + 31) >> 5) * 4) : 0;
}

// workspace up to (not including) the fused path's per-mask extreme records
static size_t ws_base_bytes(int64_t n_masks, int H, int W) {
    const int64_t slots = n_masks < kSlots ? n_masks : kSlots;
    return ws_bitmap_bytes(n_masks, H, W) + ws_range_bytes(n_masks) + ws_list_bytes(n_masks) + slots * slot_bytes(H, W) +
           ws_fill_bytes(H, W);
}

extern "C" int64_t tdod_mask_rect_workspace_bytes(int64_t n_masks, int H, int W) {
    if (n_masks <= 0 || H <= 0 || W <= 0) return 16;
    return (int64_t)(ws_base_bytes(n_masks, H, W) + (size_t)n_masks * cand_stride(H));
}

// shape of the fused path for an H x W mask: shared memory of a front CTA and CTAs per SM (0: does not fit -- wider than
// 1024 px, or the row tables + run arrays + ring exceed an SM)
static int fused_plan(int H, int W, size_t* smem) {
    if (getenv("TDOD_KB_NO_FUSED") != nullptr || W > 1024 || W < 32) return 0;
    if ((int64_t)H * W + 64 > 0x7fffffff) return 0;       // positions inside a mask are 32-bit
    *smem = front_layout(H, W, nullptr, nullptr);
    if (*smem <= 113u * 1024) return 2;
    if (*smem <= 227u * 1024) return 1;
    return 0;
}

namespace {
struct KbPlan {
    int Wp; size_t smem1, smem2; bool big;
    uint32_t* bitmap; int* rowrange; int* lists; unsigned char* slots; uint32_t* fill_slots;
};

int kb_plan(int64_t n_masks, int H, int W, void* workspace, int64_t workspace_bytes, KbPlan* p, const char* who) {
    if (n_masks < 0 || H <= 0 || W < 16) return tdod_fail(TDOD_EINVAL, "tdod_mask_*: bad argument (W must be >= 16)");
    if (H > 32767 || W > 32767) return tdod_fail(TDOD_ERANGE, "tdod_mask_*: image larger than 32767");
    if (!workspace || workspace_bytes < tdod_mask_rect_workspace_bytes(n_masks, H, W))
        return tdod_fail(TDOD_EINVAL, "tdod_mask_*: workspace smaller than tdod_mask_rect_workspace_bytes()");
    if ((reinterpret_cast<uintptr_t>(workspace) & 15) != 0)
        return tdod_fail(TDOD_EINVAL, "tdod_mask_*: workspace must be 16-byte aligned");
    (void)who;
    p->Wp = (W + 31) >> 5;
    p->smem1 = smem_layout(H, p->Wp, false, nullptr, nullptr);
    p->big = exact_needs_global(H, W);
    p->smem2 = p->big ? p->smem1 : smem_layout(H, p->Wp, true, nullptr, nullptr);
    if (p->smem1 > 227 * 1024) return tdod_fail(TDOD_ERANGE, "tdod_mask_*: image too tall for the per-row tables in shared memory");
    unsigned char* wsb = (unsigned char*)workspace;
    p->bitmap = (uint32_t*)wsb;
    p->rowrange = (int*)(wsb + ws_bitmap_bytes(n_masks, H, W));
    p->lists = (int*)(wsb + ws_bitmap_bytes(n_masks, H, W) + ws_range_bytes(n_masks));
    p->slots = wsb + ws_bitmap_bytes(n_masks, H, W) + ws_range_bytes(n_masks) + ws_list_bytes(n_masks);
    const int64_t n_run_slots = n_masks < kSlots ? n_masks : kSlots;
    p->fill_slots = p->big ? (uint32_t*)(p->slots + n_run_slots * slot_bytes(H, W)) : nullptr;
    return TDOD_OK;
}
}  // namespace

static tdod_smem_cache g_sm_front[5], g_sm_tail, g_sm_r0, g_sm_r1, g_sm_r2, g_sm_pack;

extern "C" int tdod_mask_pack(const uint8_t* masks, int64_t n_masks, int H, int W, void* workspace, int64_t workspace_bytes,
                              void* stream) {
    if (n_masks == 0) return TDOD_OK;
    if (!masks) return tdod_fail(TDOD_EINVAL, "tdod_mask_pack: masks is NULL");
    KbPlan p;
    const int rc = kb_plan(n_masks, H, W, workspace, workspace_bytes, &p, "pack");
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    // rowrange = {0x7f7f7f7f, 0x7f7f7f7f} is "no row seen" for the two atomicMin's of the pack kernel
    cudaError_t e = cudaMemsetAsync(p.rowrange, 0x7f, (size_t)n_masks * 8, st);
    if (e != cudaSuccess) return (int)e;
    const int tiles_per_mask = (H + kPackRows - 1) / kPackRows;
    const int64_t n_tiles = n_masks * tiles_per_mask;
    const size_t pk_smem = (size_t)(((int64_t)kPackRows * W + 30) / 16 + 8) * 2;
    // TMA-staged variant: two raw tile buffers (+ the bit stream) per CTA, as many CTAs per SM as shared memory allows
    const int max_chunks = (int)(((int64_t)kPackRows * W + 30) / 16 + 1);
    const int raw_cap = ((max_chunks * 16 + 127) / 128) * 128;
    const size_t tma_smem = 2 * (size_t)raw_cap + pk_smem + 16;
    const bool use_tma = tma_smem <= 100 * 1024 && n_tiles + 148 * 8 < 0x7fffffffll &&    // tile ids are divided in 32 bits
                         getenv("TDOD_PACK_NO_TMA") == nullptr;
    if (use_tma) {
        e = tdod_smem_opt_in((const void*)pack_tma_kernel, tma_smem, &g_sm_pack);
        if (e != cudaSuccess) return (int)e;
        int per_sm = (int)((220 * 1024) / (tma_smem + 1024));
        per_sm = per_sm < 1 ? 1 : (per_sm > 8 ? 8 : per_sm);
        const int64_t cap = (int64_t)148 * per_sm;
        const unsigned grid = (unsigned)(n_tiles < cap ? n_tiles : cap);
        pack_tma_kernel<<<grid, kPackT, tma_smem, st>>>(masks, n_masks, H, W, p.Wp, p.bitmap, p.rowrange, p.lists, tiles_per_mask,
                                                       raw_cap);
    } else {
        const unsigned pk_grid = (unsigned)(n_tiles < 148 * 8 * 4 ? n_tiles : 148 * 8 * 4);
        pack_kernel<<<pk_grid, kPackT, pk_smem, st>>>(masks, n_masks, H, W, p.Wp, p.bitmap, p.rowrange, p.lists, tiles_per_mask);
    }
    TDOD_CHECK_LAUNCH();
    return TDOD_OK;
}

// the three rectangle passes over a row-padded bitmap (the workspace's, or a caller's) whose row ranges are in p.rowrange
static int rect_passes(const KbPlan& p, const uint32_t* bitmap, int64_t n_masks, int H, int W, float* out_rects, uint8_t* out_has_rect,
                       cudaStream_t st) {
    cudaError_t e = tdod_smem_opt_in((const void*)rect_kernel<false, false>, p.smem1, &g_sm_r0);
    if (e != cudaSuccess) return (int)e;
    e = tdod_smem_opt_in((const void*)rect_kernel<true, false>, p.smem2, &g_sm_r1);
    if (e != cudaSuccess) return (int)e;
    e = tdod_smem_opt_in((const void*)rect_kernel<true, true>, p.smem2, &g_sm_r2);
    if (e != cudaSuccess) return (int)e;
    // the state byte doubles as the work list of the later passes: every mask starts undecided
    e = cudaMemsetAsync(out_has_rect, 0xff, (size_t)n_masks, st);
    if (e != cudaSuccess) return (int)e;
    rect_kernel<false, false><<<(unsigned)n_masks, kT, p.smem1, st>>>(bitmap, p.rowrange, n_masks, H, W, out_rects, out_has_rect,
                                                                   nullptr, p.lists, nullptr);
    TDOD_CHECK_LAUNCH();
    rect_kernel<true, false><<<(unsigned)(n_masks < kExactCtas ? n_masks : kExactCtas), kT, p.smem2, st>>>(
        bitmap, p.rowrange, n_masks, H, W, out_rects, out_has_rect, nullptr, p.lists, p.fill_slots);
    TDOD_CHECK_LAUNCH();
    const unsigned nslots = (unsigned)(n_masks < kSlots ? n_masks : kSlots);
    rect_kernel<true, true><<<nslots, kT, p.smem2, st>>>(bitmap, p.rowrange, n_masks, H, W, out_rects, out_has_rect, p.slots,
                                                        p.lists, p.fill_slots);
    TDOD_CHECK_LAUNCH();
    return TDOD_OK;
}

extern "C" int tdod_mask_rect_packed(int64_t n_masks, int H, int W, float* out_rects, uint8_t* out_has_rect, void* workspace,
                                     int64_t workspace_bytes, void* stream) {
    if (n_masks == 0) return TDOD_OK;
    if (!out_rects || !out_has_rect) return tdod_fail(TDOD_EINVAL, "tdod_mask_rect_packed: an output pointer is NULL");
    KbPlan p;
    const int rc = kb_plan(n_masks, H, W, workspace, workspace_bytes, &p, "rect");
    if (rc) return rc;
    return rect_passes(p, p.bitmap, n_masks, H, W, out_rects, out_has_rect, (cudaStream_t)stream);
}

#ifdef TDOD_KB_PROF
extern "C" int tdod_debug_kb_prof(unsigned long long* out16_host, int reset) {
    unsigned long long z[16] = {0};
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out16_host, g_kb_prof, sizeof(z));
    if (reset) cudaMemcpyToSymbol(g_kb_prof, z, sizeof(z));
    return 0;
}
#endif

extern "C" int tdod_mask_rect_bitmap(const uint32_t* bitmap, int64_t n_masks, int H, int W, float* out_rects,
                                     uint8_t* out_has_rect, void* workspace, int64_t workspace_bytes, void* stream) {
    if (n_masks == 0) return TDOD_OK;
    if (!bitmap || !out_rects || !out_has_rect) return tdod_fail(TDOD_EINVAL, "tdod_mask_rect_bitmap: a pointer is NULL");
    if (n_masks > 0x7fffffff) return tdod_fail(TDOD_ERANGE, "tdod_mask_rect_bitmap: too many masks for one call");
    KbPlan p;
    const int rc = kb_plan(n_masks, H, W, workspace, workspace_bytes, &p, "bitmap");
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const cudaError_t e = cudaMemsetAsync(p.lists, 0, 16, st);
    if (e != cudaSuccess) return (int)e;
    rowrange_kernel<<<(unsigned)n_masks, 256, 0, st>>>(bitmap, H, p.Wp, p.rowrange);
    TDOD_CHECK_LAUNCH();
    return rect_passes(p, bitmap, n_masks, H, W, out_rects, out_has_rect, st);
}

extern "C" int tdod_mask_rect(const uint8_t* masks, int64_t n_masks, int H, int W, float* out_rects,
                              uint8_t* out_has_rect, void* workspace, int64_t workspace_bytes, void* stream) {
    if (n_masks == 0) return TDOD_OK;
    if (!masks || !out_rects || !out_has_rect) return tdod_fail(TDOD_EINVAL, "tdod_mask_rect: a pointer is NULL");
    size_t fsmem;
    const int per_sm = fused_plan(H, W, &fsmem);
    if (per_sm == 0) {                                  // masks too large for the shared-memory bit stream: two-kernel path
        const int rc = tdod_mask_pack(masks, n_masks, H, W, workspace, workspace_bytes, stream);
        if (rc) return rc;
        return tdod_mask_rect_packed(n_masks, H, W, out_rects, out_has_rect, workspace, workspace_bytes, stream);
    }
    KbPlan p;
    const int rc = kb_plan(n_masks, H, W, workspace, workspace_bytes, &p, "fused");
    if (rc) return rc;
    if (n_masks > 0x7fffffff / 4) return tdod_fail(TDOD_ERANGE, "tdod_mask_rect: too many masks for one call");
    cudaStream_t st = (cudaStream_t)stream;
    unsigned char* cand = (unsigned char*)workspace + ws_base_bytes(n_masks, H, W);
    const size_t per_warp = tail_layout(H, nullptr, nullptr);
    const size_t tsmem = per_warp * kTailWarps;
    if (tsmem > 227 * 1024) return tdod_fail(TDOD_ERANGE, "tdod_mask_rect: image too tall for the tail kernel");
    cudaError_t e;
    typedef void (*front_fn)(const uint8_t*, int64_t, int, int, unsigned char*, uint8_t*, int*);
    static const front_fn fronts[5] = {kb_front_kernel<1>, kb_front_kernel<2>, kb_front_kernel<3>, kb_front_kernel<4>, kb_front_kernel<5>};
    const int qh = front_slot_hw(W) / 16;               // 1..5 for 32 <= W <= 1024
    if ((e = tdod_smem_opt_in((const void*)fronts[qh - 1], fsmem, &g_sm_front[qh - 1])) != cudaSuccess) return (int)e;
    if ((e = tdod_smem_opt_in((const void*)kb_tail_kernel, tsmem, &g_sm_tail)) != cudaSuccess) return (int)e;
    if ((e = tdod_smem_opt_in((const void*)rect_kernel<true, false>, p.smem2, &g_sm_r1)) != cudaSuccess) return (int)e;
    if ((e = tdod_smem_opt_in((const void*)rect_kernel<true, true>, p.smem2, &g_sm_r2)) != cudaSuccess) return (int)e;
    // rowrange = "no row seen" for the list-driven pack; work-list counters = 0
    if ((e = cudaMemsetAsync(p.rowrange, 0x7f, (size_t)n_masks * 8, st)) != cudaSuccess) return (int)e;
    if ((e = cudaMemsetAsync(p.lists, 0, 16, st)) != cudaSuccess) return (int)e;
    const int64_t cap = (int64_t)148 * per_sm;
    fronts[qh - 1]<<<(unsigned)(n_masks < cap ? n_masks : cap), kHalf + kExT, fsmem, st>>>(masks, n_masks, H, W, cand, out_has_rect,
                                                                                p.lists);
    TDOD_CHECK_LAUNCH();
    kb_tail_kernel<<<(unsigned)((n_masks + kTailWarps - 1) / kTailWarps), 32 * kTailWarps, tsmem, st>>>(cand, n_masks, H, out_rects,
                                                                                                      out_has_rect, p.lists, per_warp);
    TDOD_CHECK_LAUNCH();
    // the undecided rest (usually nobody): their bitmaps only, then the exact passes
    const int tiles_per_mask = (H + kPackRows - 1) / kPackRows;
    const size_t pk_smem = (size_t)(((int64_t)kPackRows * W + 30) / 16 + 8) * 2;
    pack_list_kernel<<<148 * 2, kPackT, pk_smem, st>>>(masks, n_masks, H, W, p.Wp, p.bitmap, p.rowrange, p.lists, tiles_per_mask);
    TDOD_CHECK_LAUNCH();
    rect_kernel<true, false><<<(unsigned)(n_masks < kExactCtas ? n_masks : kExactCtas), kT, p.smem2, st>>>(
        p.bitmap, p.rowrange, n_masks, H, W, out_rects, out_has_rect, nullptr, p.lists, p.fill_slots);
    TDOD_CHECK_LAUNCH();
    const unsigned nslots = (unsigned)(n_masks < kSlots ? n_masks : kSlots);
    rect_kernel<true, true><<<nslots, kT, p.smem2, st>>>(p.bitmap, p.rowrange, n_masks, H, W, out_rects, out_has_rect, p.slots,
                                                        p.lists, p.fill_slots);
    TDOD_CHECK_LAUNCH();
    return TDOD_OK;
}
