// Host-side transfer codec for object masks: 1 byte/px (torch.bool storage, what the reference's callers hold, e.g.
// the SAM masks of cubercnn/modeling/roi_heads/roi_heads.py:432) -> 1 bit/px in the layout tdod_mask_rect_bitmap reads
// ([n_masks, H, Wp] 32-bit words, bit (c & 31) of word (c >> 5) of a row = pixel c).  The masks are 75 % of the bytes a
// step moves over PCIe; packing them on the host cores while the previous chunk's copy is in flight cuts the step's
// host->device traffic from 1.98 GB to 0.68 GB.  This is a data-format step on the host side of the boundary (like the
// JSON writer), not a CPU implementation of any kernel: every score and rectangle is still computed on the GPU.
#include <immintrin.h>

#include <condition_variable>
#include <cstring>
#include <mutex>
#include <thread>

#include "tdod_internal.h"

namespace {

// one row of W bytes -> Wp words; any byte != 0 sets its bit
inline void pack_row(const uint8_t* src, int W, uint32_t* dst, int Wp) {
    const __m128i zero = _mm_setzero_si128();
    int x = 0, w = 0;
    for (; x + 32 <= W; x += 32, ++w) {
        const __m128i a = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src + x));
        const __m128i b = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src + x + 16));
        const uint32_t za = (uint32_t)_mm_movemask_epi8(_mm_cmpeq_epi8(a, zero));
        const uint32_t zb = (uint32_t)_mm_movemask_epi8(_mm_cmpeq_epi8(b, zero));
        dst[w] = ~(za | (zb << 16));
    }
    if (x < W) {                                        // the row's last, partial word: never read past the row
        uint8_t tmp[32];
        std::memset(tmp, 0, sizeof(tmp));
        std::memcpy(tmp, src + x, (size_t)(W - x));
        const __m128i a = _mm_loadu_si128(reinterpret_cast<const __m128i*>(tmp));
        const __m128i b = _mm_loadu_si128(reinterpret_cast<const __m128i*>(tmp + 16));
        const uint32_t za = (uint32_t)_mm_movemask_epi8(_mm_cmpeq_epi8(a, zero));
        const uint32_t zb = (uint32_t)_mm_movemask_epi8(_mm_cmpeq_epi8(b, zero));
        dst[w++] = ~(za | (zb << 16));
    }
    for (; w < Wp; ++w) dst[w] = 0u;
}

// the same with 32-byte vectors where the host has AVX2 (checked once at run time): half the instructions per byte
// staging words are written once and read by the DMA engine only: non-temporal stores skip the read-for-ownership
#define NT(p, v) _mm_stream_si32(reinterpret_cast<int*>(p), (int)(v))
__attribute__((target("avx2"))) inline void pack_row_avx2(const uint8_t* src, int W, uint32_t* dst, int Wp) {
    const __m256i zero = _mm256_setzero_si256();
    int x = 0, w = 0;
    for (; x + 64 <= W; x += 64, w += 2) {
        const __m256i a = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(src + x));
        const __m256i b = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(src + x + 32));
        NT(dst + w, ~(uint32_t)_mm256_movemask_epi8(_mm256_cmpeq_epi8(a, zero)));
        NT(dst + w + 1, ~(uint32_t)_mm256_movemask_epi8(_mm256_cmpeq_epi8(b, zero)));
    }
    for (; x + 32 <= W; x += 32, ++w) {
        const __m256i a = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(src + x));
        NT(dst + w, ~(uint32_t)_mm256_movemask_epi8(_mm256_cmpeq_epi8(a, zero)));
    }
    if (x < W) {                                        // the row's last, partial word: never read past the row
        uint8_t tmp[32];
        std::memset(tmp, 0, sizeof(tmp));
        std::memcpy(tmp, src + x, (size_t)(W - x));
        const __m256i a = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(tmp));
        NT(dst + w++, ~(uint32_t)_mm256_movemask_epi8(_mm256_cmpeq_epi8(a, zero)));
    }
    for (; w < Wp; ++w) NT(dst + w, 0u);
}

#undef NT

__attribute__((target("avx2"))) void pack_rows_avx2(const uint8_t* masks, int W, uint32_t* out, int Wp, int64_t r0, int64_t r1) {
    for (int64_t r = r0; r < r1; ++r) pack_row_avx2(masks + r * W, W, out + r * Wp, Wp);
}

void pack_rows_sse2(const uint8_t* masks, int W, uint32_t* out, int Wp, int64_t r0, int64_t r1) {
    for (int64_t r = r0; r < r1; ++r) pack_row(masks + r * W, W, out + r * Wp, Wp);
}

void pack_rows(const uint8_t* masks, int W, uint32_t* out, int Wp, int64_t r0, int64_t r1) {
    static const bool avx2 = __builtin_cpu_supports("avx2");
    if (avx2) pack_rows_avx2(masks, W, out, Wp, r0, r1);
    else pack_rows_sse2(masks, W, out, Wp, r0, r1);
}

// Persistent workers (created on first use, never destroyed: the process's exit ends them).  Spawning 16 threads per call
// costs ~0.5 ms, a tenth of what packing one chunk takes; parked workers cost a condition-variable wake-up.
struct PackPool {
    std::mutex mu;
    std::condition_variable work, done;
    int n_workers = 0, generation = 0, pending = 0;
    // the current job
    const uint8_t* masks = nullptr; uint32_t* out = nullptr;
    int W = 0, Wp = 0, parts = 0;
    int64_t rows = 0, per = 0;

    void worker(int id) {
        int seen = 0;
        std::unique_lock<std::mutex> lk(mu);
        for (;;) {
            work.wait(lk, [&] { return generation != seen; });
            seen = generation;
            const int part = id + 1;                    // part 0 is the caller's
            if (part < parts) {
                const uint8_t* m = masks; uint32_t* o = out;
                const int w = W, wp = Wp;
                const int64_t r0 = part * per, r1 = r0 + per < rows ? r0 + per : rows;
                lk.unlock();
                if (r0 < r1) pack_rows(m, w, o, wp, r0, r1);
                lk.lock();
            }
            if (--pending == 0) done.notify_one();
        }
    }

    void run(const uint8_t* m, int w, uint32_t* o, int wp, int64_t n_rows, int n_parts) {
        {
            std::unique_lock<std::mutex> lk(mu);
            while (n_workers < n_parts - 1) {
                std::thread(&PackPool::worker, this, n_workers).detach();
                ++n_workers;
            }
            masks = m; out = o; W = w; Wp = wp; rows = n_rows; parts = n_parts;
            per = (n_rows + n_parts - 1) / n_parts;
            pending = n_workers;
            ++generation;
        }
        work.notify_all();
        pack_rows(m, w, o, wp, 0, per < n_rows ? per : n_rows);
        std::unique_lock<std::mutex> lk(mu);
        done.wait(lk, [&] { return pending == 0; });
    }
};

std::mutex g_pack_call;                 // one packing job at a time
PackPool* g_pack_pool = nullptr;

}  // namespace

extern "C" int tdod_host_pack_masks(const uint8_t* masks, int64_t n_masks, int H, int W, uint32_t* out_bitmap, int n_threads) {
    if (n_masks == 0) return TDOD_OK;
    if (!masks || !out_bitmap || n_masks < 0 || H <= 0 || W <= 0) return tdod_fail(TDOD_EINVAL, "tdod_host_pack_masks: bad argument");
    const int Wp = (W + 31) >> 5;
    const int64_t rows = n_masks * (int64_t)H;
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 256) n_threads = 256;
    if ((int64_t)n_threads > rows) n_threads = (int)rows;
    if (n_threads == 1) {
        pack_rows(masks, W, out_bitmap, Wp, 0, rows);
        _mm_sfence();
        return TDOD_OK;
    }
    std::lock_guard<std::mutex> call(g_pack_call);
    if (!g_pack_pool) g_pack_pool = new PackPool();
    g_pack_pool->run(masks, W, out_bitmap, Wp, rows, n_threads);
    _mm_sfence();
    return TDOD_OK;
}
