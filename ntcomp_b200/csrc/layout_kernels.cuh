// layout_kernels.cuh -- the derived HBM layouts (host_index.h: rank blocks, LCS words, decode jump table, prefix table) built by
// kernels from the canonical arrays (4 subset rows + LCS bytes, 1.5 B per node uploaded) instead of on host threads followed by
// a 14 B-per-node upload.  Bit for bit the arrays host_index.cpp builds (tests/test_index_build.py compares them); the host
// builders stay for tests/emul and as NTC_HOST_LAYOUTS=1.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "device_common.cuh"
#include "scan.cuh"

namespace ntc {
namespace lay {

// scan input: ones of two rows per 64-bit word
struct RowPopLoad {
    const uint64_t* ra; const uint64_t* rb;
    NTC_HD Pair64 operator()(uint64_t w) const { return Pair64{(uint64_t)NTC_POPCLL(ra[w]), (uint64_t)NTC_POPCLL(rb[w])}; }
};

struct Rows {
    const uint64_t* row[4];     // n bits each, LSB first
    const uint64_t* pre[4];     // ones before word w (row_words + 1 entries)
    uint64_t row_words;
    uint64_t C[5];
};

NTC_HD uint32_t q_bit(const uint8_t* __restrict__ lcs, uint64_t n, uint32_t k, uint64_t v) {
    if (v >= n) return 0u;
    const bool lead = lcs[v] < k - 1;
    const bool next_lead = (v + 1 == n) || lcs[v + 1] < k - 1;
    return (lead && next_lead) ? 1u : 0u;
}
NTC_HD uint64_t row_word(const Rows& r, int c, uint64_t w) { return w < r.row_words ? r.row[c][w] : 0ull; }
NTC_HD uint32_t ones_before_word(const Rows& r, int c, uint64_t w) { return (uint32_t)r.pre[c][w < r.row_words ? w : r.row_words]; }

// one rank block (64 bytes = 16 u32) per element; NODES = 32 or 64 (layout.h)
template <int NODES>
NTC_HD void rank_block_thread(const Rows& r, const uint8_t* __restrict__ lcs, uint64_t n, uint32_t k, uint32_t* __restrict__ out, uint64_t b) {
    const uint64_t base = b * NODES;
    uint64_t q = 0;
    for (uint32_t i = 0; i < (uint32_t)NODES; i++) q |= (uint64_t)q_bit(lcs, n, k, base + i) << i;
    uint32_t* e = out + b * 16;
    if (NODES == 32) {
        const uint64_t w = b >> 1;
        const uint32_t half = (uint32_t)(b & 1);
        for (int c = 0; c < 4; c++) {
            const uint64_t word = row_word(r, c, w);
            const uint32_t bits = (uint32_t)(word >> (32 * half));
            const uint32_t before = ones_before_word(r, c, w) + (half ? (uint32_t)NTC_POPC((uint32_t)word) : 0u);
            e[4 * c + 0] = (uint32_t)r.C[c] + before;
            e[4 * c + 1] = bits;
            e[4 * c + 2] = (uint32_t)q;
            e[4 * c + 3] = 0;
        }
    } else {
        for (int c = 0; c < 4; c++) {
            const uint64_t word = row_word(r, c, b);
            uint32_t* s = e + (c >> 1) * 8;                            // {count, count, bits, bits, Q} per pair of bases
            s[c & 1] = (uint32_t)r.C[c] + ones_before_word(r, c, b);
            s[2 + 2 * (c & 1)] = (uint32_t)word;
            s[3 + 2 * (c & 1)] = (uint32_t)(word >> 32);
            s[6] = (uint32_t)q;
            s[7] = (uint32_t)(q >> 32);
        }
    }
}

// LCS word i (u32): LCS[i] | min(255, i - PSV(i)) << 8 | LCS[i+1] << 16 | min(255, NSV(i+1) - (i+1)) << 24 (host_index.h)
NTC_HD uint32_t dist_left(const uint8_t* __restrict__ lcs, uint64_t i) {        // to the nearest strictly smaller LCS on the left, saturated
    const uint32_t x = lcs[i];
    if (x == 0) return 0;
    for (uint32_t d = 1; d < 255; d++) {
        if (d > i) return (uint32_t)i;                                 // never: lcs[0] == 0 < x ends the scan at d == i
        if (lcs[i - d] < x) return d;
    }
    return 255;
}
NTC_HD uint32_t dist_right(const uint8_t* __restrict__ lcs, uint64_t n, uint64_t i) {   // ... on the right; LCS[n] = 0 is the sentinel
    const uint32_t x = lcs[i];
    if (x == 0) return 0;
    for (uint32_t d = 1; d < 255; d++) {
        if (i + d >= n || lcs[i + d] < x) return d;
    }
    return 255;
}
NTC_HD void lcs_word_thread(const uint8_t* __restrict__ lcs, uint64_t n, uint32_t* __restrict__ out, uint64_t i) {
    const uint32_t lo = i < n ? ((uint32_t)lcs[i] | (dist_left(lcs, i) << 8)) : 0u;
    const uint32_t hi = i + 1 < n ? ((uint32_t)lcs[i + 1] | (dist_right(lcs, n, i + 1) << 8)) : 0u;
    out[i] = lo | (hi << 16);
}

// jump table, step 1: one inverse-LF step per node -- the j-th one of row c is the predecessor of node C[c] + j, in-label c.
// Element = (word w, base c); entry = pred (low 32 bits) | label (bits 32..)
NTC_HD void pred_init_thread(const Rows& r, uint64_t* __restrict__ jump, uint64_t e) {
    const int c = (int)(e & 3);
    const uint64_t w = e >> 2;
    if (e == 0) jump[0] = 0;                                           // the root: pred = itself, label 0
    uint64_t x = r.row[c][w];
    uint64_t j = r.C[c] + r.pre[c][w];
    while (x) {
#if defined(__CUDA_ARCH__)
        const int b = __ffsll((long long)x) - 1;
#else
        const int b = __builtin_ctzll(x);
#endif
        jump[j++] = (w * 64 + (uint64_t)b) | ((uint64_t)c << 32);
        x &= x - 1;
    }
}
// pointer doubling: after the round with step t, pred = pred^(2t) and the label field holds the last 2t bases, LAST base in the top 2 bits
NTC_HD void pred_double_thread(const uint64_t* __restrict__ in, uint64_t* __restrict__ out, uint32_t t, uint64_t v) {
    const uint64_t a = in[v];
    const uint64_t b = in[(uint32_t)a];
    const uint32_t ca = (uint32_t)(a >> 32), cb = (uint32_t)(b >> 32);
    out[v] = (uint64_t)(uint32_t)b | ((uint64_t)((ca << (2 * t)) | cb) << 32);
}

// prefix table entry x (host_index.h): interval starts after 2..8 bases, interval end after 8, from the empty match
template <int NODES>
NTC_HD uint32_t rank_from_blocks(const uint32_t* __restrict__ rb, uint32_t c, uint64_t i) {
    const uint32_t off = (uint32_t)(i % NODES);
    if (NODES == 32) {
        const uint32_t* e = rb + (i / NODES) * 16 + c * 4;
        return e[0] + (uint32_t)NTC_POPC(e[1] & ((1u << off) - 1u));
    }
    const uint32_t* e = rb + (i / NODES) * 16 + (c >> 1) * 8;
    const uint64_t bits = (uint64_t)e[2 + 2 * (c & 1)] | ((uint64_t)e[3 + 2 * (c & 1)] << 32);
    return e[c & 1] + (uint32_t)NTC_POPCLL(bits & ((1ull << off) - 1ull));
}
template <int NODES>
NTC_HD void prefix_entry_thread(const uint32_t* __restrict__ rb, uint64_t n, uint32_t* __restrict__ out, uint64_t x) {
    uint32_t s = 0, e = (uint32_t)n, w[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    bool ok = true;
    for (uint32_t l = 1; l <= 8 && ok; l++) {
        const uint32_t c = (uint32_t)(x >> (2 * (l - 1))) & 3u;
        s = rank_from_blocks<NODES>(rb, c, s);
        e = rank_from_blocks<NODES>(rb, c, e);
        ok = s < e;
        if (l >= 2) w[l - 2] = s;
    }
    if (ok) w[7] = e;
    for (int i = 0; i < 8; i++) out[x * 8 + i] = ok ? w[i] : 0u;
}

#if defined(__CUDACC__)
template <int NODES>
__global__ void __launch_bounds__(256) rank_blocks_kernel(Rows r, const uint8_t* __restrict__ lcs, uint64_t n, uint32_t k, uint32_t* __restrict__ out,
                                                          uint64_t n_blocks) {
    const uint64_t b = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (b < n_blocks) rank_block_thread<NODES>(r, lcs, n, k, out, b);
}
static __global__ void __launch_bounds__(256) lcs_words_kernel(const uint8_t* __restrict__ lcs, uint64_t n, uint32_t* __restrict__ out) {
    const uint64_t i = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (i <= n) lcs_word_thread(lcs, n, out, i);
}
static __global__ void __launch_bounds__(256) pred_init_kernel(Rows r, uint64_t* __restrict__ jump) {
    const uint64_t e = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (e < r.row_words * 4) pred_init_thread(r, jump, e);
}
static __global__ void __launch_bounds__(256) pred_double_kernel(const uint64_t* __restrict__ in, uint64_t* __restrict__ out, uint32_t t, uint64_t n) {
    const uint64_t v = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (v < n) pred_double_thread(in, out, t, v);
}
template <int NODES>
__global__ void __launch_bounds__(256) prefix_table_kernel(const uint32_t* __restrict__ rb, uint64_t n, uint32_t* __restrict__ out) {
    const uint64_t x = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (x < (1u << 16)) prefix_entry_thread<NODES>(rb, n, out, x);
}
#endif

}  // namespace lay
}  // namespace ntc
