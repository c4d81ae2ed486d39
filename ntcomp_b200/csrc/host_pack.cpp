// host_pack.cpp -- ASCII <-> 2-bit conversion on host threads for the host-buffer entry points: reads cross PCIe as
// 0.25 B per base instead of 1 (pack2bit_kernel's output made on the host), decoded bases come back the same way and are
// expanded into the caller's buffer.  Same bit layout as pack2bit_kernel / bitnuc::as_2bit (src/encode.rs:154): base j of the
// stream at bits 2(j % 32) of word j / 32, A0 C1 G2 T3; a byte that is not A, C, G or T is reported, as the kernel does.
// A persistent worker pool (spawning threads per 48 M-base chunk cost more than the packing).
#include "host_pack.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>

#if defined(__x86_64__)
#include <immintrin.h>
#endif

namespace ntc {

// ---- worker pool: run(n_items, f) calls f(item) for every item on the pool's threads + the caller ------------------------
struct HostPool::Impl {
    std::vector<std::thread> th;
    std::mutex m;
    std::condition_variable cv, cv_done;
    const std::function<void(uint64_t)>* fn = nullptr;
    uint64_t n_items = 0;
    std::atomic<uint64_t> generation{0};
    std::atomic<uint64_t> next{0};
    std::atomic<int> active{0};
    bool stop = false;
    std::atomic<int64_t> last_large{0};   // steady-clock ns of the last large job (a chunk of a big batch): the next is likely ~0.5 ms away, worth polling for
    static int64_t now_ns() { return std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
    bool streaming() const { return now_ns() - last_large.load(std::memory_order_relaxed) < 5000000; }
    void work() {
        for (;;) {
            const uint64_t i = next.fetch_add(1);
            if (i >= n_items) break;
            (*fn)(i);
        }
    }
    static void pause() {
#if defined(__x86_64__)
        _mm_pause();
#endif
    }
    // A worker polls for the next job for ~2 ms before it sleeps, but only while a LARGE job (16 tasks or more) ran in the last 5 ms; otherwise it sleeps at once: the host-buffer encode hands over a
    // chunk every ~0.5 ms, and waking 15 sleeping vCPUs cost more than packing the chunk (measured on the B200 box: +0.6 ms per chunk).
    void loop() {
        uint64_t seen = 0;
        for (;;) {
            bool got = false;
            const auto t0 = std::chrono::steady_clock::now();
            for (int spin = 0; !got && (spin & 1023 ? true : streaming()); spin++) {
                if (generation.load(std::memory_order_acquire) != seen) { got = true; break; }
                pause();
                if ((spin & 1023) == 1023 && std::chrono::steady_clock::now() - t0 > std::chrono::milliseconds(2)) break;
            }
            if (!got) {
                std::unique_lock<std::mutex> lk(m);
                cv.wait(lk, [&] { return stop || generation.load() != seen; });
                if (stop) return;
            }
            seen = generation.load(std::memory_order_acquire);
            work();
            if (active.fetch_sub(1) == 1) { std::lock_guard<std::mutex> lk(m); cv_done.notify_all(); }
        }
    }
};

HostPool::HostPool(int n_threads) : impl_(new Impl()) {
    n_ = std::max(1, n_threads);
    for (int i = 1; i < n_; i++) impl_->th.emplace_back([this] { impl_->loop(); });
}
HostPool::~HostPool() {
    { std::lock_guard<std::mutex> lk(impl_->m); impl_->stop = true; impl_->cv.notify_all(); }
    for (auto& t : impl_->th) t.join();
    delete impl_;
}
void HostPool::run(uint64_t n_items, const std::function<void(uint64_t)>& f) {
    if (n_items == 0) return;
    if (n_items == 1 || n_ == 1) { for (uint64_t i = 0; i < n_items; i++) f(i); return; }   // nothing to share: nobody is woken
    Impl& p = *impl_;
    {
        std::lock_guard<std::mutex> lk(p.m);
        p.fn = &f;
        p.n_items = n_items;
        p.next.store(0);
        p.active.store((int)p.th.size());
        if (n_items >= 16) p.last_large.store(Impl::now_ns(), std::memory_order_relaxed);
        p.generation.fetch_add(1, std::memory_order_release);
        p.cv.notify_all();
    }
    p.work();
    // the caller polls too: the last worker is at most one task behind
    for (int spin = 0; p.active.load(std::memory_order_acquire) != 0; spin++) {
        Impl::pause();
        if (spin > (1 << 16)) {
            std::unique_lock<std::mutex> lk(p.m);
            p.cv_done.wait(lk, [&] { return p.active.load() == 0; });
            break;
        }
    }
}

// ---- ASCII -> 2 bit ----------------------------------------------------------------------------------------------------------
namespace {

struct CodeLut {                                                      // byte -> 2-bit code, 0xFF for anything but A, C, G, T
    uint8_t t[256];
    CodeLut() { memset(t, 0xFF, sizeof t); t['A'] = 0; t['C'] = 1; t['G'] = 2; t['T'] = 3; }
};
const CodeLut kLut;

#if defined(__x86_64__)
// 32 bases -> one u64, AVX2
__attribute__((target("avx2"))) inline uint64_t pack32_avx2(const uint8_t* s, __m256i* bad_acc) {
    const __m256i v = _mm256_loadu_si256((const __m256i*)s);
    const __m256i x = _mm256_and_si256(_mm256_srli_epi16(v, 1), _mm256_set1_epi8(3));
    const __m256i c = _mm256_xor_si256(x, _mm256_and_si256(_mm256_srli_epi16(x, 1), _mm256_set1_epi8(1)));
    const __m256i letters = _mm256_setr_epi8('A', 'C', 'G', 'T', 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 'A', 'C', 'G', 'T', 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0);
    *bad_acc = _mm256_or_si256(*bad_acc, _mm256_xor_si256(_mm256_shuffle_epi8(letters, c), v));
    // c0 + 4 c1 per 16-bit lane, then (..) + 16 (..) per 32-bit lane: one byte per four bases
    const __m256i p16 = _mm256_maddubs_epi16(c, _mm256_set1_epi16(0x0401));
    const __m256i p32 = _mm256_madd_epi16(p16, _mm256_set1_epi32(0x00100001));
    // the low byte of each of the 8 dwords, in order
    const __m256i sh = _mm256_shuffle_epi8(p32, _mm256_setr_epi8(0, 4, 8, 12, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1,
                                                                 0, 4, 8, 12, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1));
    const uint32_t lo = (uint32_t)_mm256_extract_epi32(sh, 0), hi = (uint32_t)_mm256_extract_epi32(sh, 4);
    return (uint64_t)lo | ((uint64_t)hi << 32);
}
__attribute__((target("avx2"))) bool pack_range_avx2(const uint8_t* src, uint64_t n_bases, uint64_t w0, uint64_t w1, uint64_t* dst) {
    __m256i bad = _mm256_setzero_si256();
    uint64_t w = w0;
    for (; w < w1 && (w + 1) * 32 <= n_bases; w++) dst[w] = pack32_avx2(src + w * 32, &bad);
    bool any_bad = !_mm256_testz_si256(bad, bad);
    for (; w < w1; w++) {                                             // the stream's last, partial word
        uint64_t out = 0;
        for (uint64_t j = w * 32; j < n_bases && j < (w + 1) * 32; j++) {
            const uint32_t c = kLut.t[src[j]];
            any_bad |= c > 3;
            out |= (uint64_t)(c & 3u) << (2 * (j & 31));
        }
        dst[w] = out;
    }
    return any_bad;
}
// 32 bases from one u64 -> ASCII
__attribute__((target("avx2"))) void unpack_range_avx2(const uint64_t* src, uint64_t first_base, uint64_t n, uint8_t* dst) {
    // expand byte b (4 bases) -> 4 letters with two shuffles: nibble tables of letter pairs are too wide, so go through codes
    const __m256i letters = _mm256_setr_epi8('A', 'C', 'G', 'T', 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 'A', 'C', 'G', 'T', 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0);
    uint64_t j = 0;
    if ((first_base & 31) == 0) {
        for (; j + 32 <= n; j += 32) {
            const uint64_t w = src[(first_base + j) >> 5];
            // byte i of the word -> dword i: broadcast then shift each byte lane by 0/2/4/6
            const __m256i bytes = _mm256_cvtepu8_epi32(_mm_cvtsi64_si128((long long)w));                 // 8 dwords, one packed byte each
            const __m256i spread = _mm256_or_si256(_mm256_or_si256(bytes, _mm256_slli_epi32(bytes, 6)),      // byte1 = b >> 2
                                                   _mm256_or_si256(_mm256_slli_epi32(bytes, 12), _mm256_slli_epi32(bytes, 18)));
            const __m256i codes = _mm256_and_si256(spread, _mm256_set1_epi8(3));
            _mm256_storeu_si256((__m256i*)(dst + j), _mm256_shuffle_epi8(letters, codes));
        }
    }
    for (; j < n; j++) {
        const uint64_t p = first_base + j;
        dst[j] = (uint8_t)"ACGT"[(src[p >> 5] >> (2 * (p & 31))) & 3];
    }
}
#endif

bool pack_range_scalar(const uint8_t* src, uint64_t n_bases, uint64_t w0, uint64_t w1, uint64_t* dst) {
    uint32_t bad = 0;
    for (uint64_t w = w0; w < w1; w++) {
        uint64_t out = 0;
        const uint64_t hi = std::min(n_bases, (w + 1) * 32);
        for (uint64_t j = w * 32; j < hi; j++) {
            const uint32_t c = kLut.t[src[j]];
            bad |= c;
            out |= (uint64_t)(c & 3u) << (2 * (j & 31));
        }
        dst[w] = out;
    }
    return (bad & 0xFCu) != 0;
}
void unpack_range_scalar(const uint64_t* src, uint64_t first_base, uint64_t n, uint8_t* dst) {
    for (uint64_t j = 0; j < n; j++) {
        const uint64_t p = first_base + j;
        dst[j] = (uint8_t)"ACGT"[(src[p >> 5] >> (2 * (p & 31))) & 3];
    }
}

bool have_avx2() {
#if defined(__x86_64__)
    static const bool v = __builtin_cpu_supports("avx2") && !getenv("NTC_NO_AVX2");
    return v;
#else
    return false;
#endif
}

}  // namespace

bool host_pack2bit(HostPool& pool, const uint8_t* src, uint64_t n_bases, uint64_t* dst) {
    const uint64_t n_words = (n_bases + 31) / 32;
    const uint64_t per = 1 << 15;                                     // 1 M bases per task
    std::atomic<int> bad(0);
    pool.run((n_words + per - 1) / per, [&](uint64_t t) {
        const uint64_t w0 = t * per, w1 = std::min(n_words, w0 + per);
#if defined(__x86_64__)
        const bool b = have_avx2() ? pack_range_avx2(src, n_bases, w0, w1, dst) : pack_range_scalar(src, n_bases, w0, w1, dst);
#else
        const bool b = pack_range_scalar(src, n_bases, w0, w1, dst);
#endif
        if (b) bad.store(1);
    });
    return bad.load() != 0;
}

void host_unpack2bit(HostPool& pool, const uint64_t* src, uint64_t n_bases, uint8_t* dst) {
    const uint64_t per = 1 << 20;                                     // bases per task, a multiple of 32
    pool.run((n_bases + per - 1) / per, [&](uint64_t t) {
        const uint64_t b0 = t * per, n = std::min(n_bases - b0, per);
#if defined(__x86_64__)
        if (have_avx2()) { unpack_range_avx2(src, b0, n, dst + b0); return; }
#endif
        unpack_range_scalar(src, b0, n, dst + b0);
    });
}

}  // namespace ntc
