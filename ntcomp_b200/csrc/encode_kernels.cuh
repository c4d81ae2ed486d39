// encode_kernels.cuh -- the encode hot path as sm_100a kernels.
//
//   pack2bit_kernel   ASCII reads -> 2-bit stream, flags non-ACGT bytes            (bitnuc::as_2bit order)
//   ms_kernel         k-bounded matching statistics, one read per thread           (StreamingIndex::
//                     matching_statistics as called at src/lib.rs:172-173) + record count per read
//   emit_kernel       right-to-left greedy segmentation (src/lib.rs:175-219) fused with record
//                     packing (src/encode.rs:144-160), writing at scanned offsets
//
// Every rank probe is ONE aligned 16-byte load (one L2 sector): the entry carries the cumulative
// count (with C[c] folded in), 48 row bits and 48 Q bits.  left_extend_kmer's 4 x k-char search
// per base (src/lib.rs:106-114) is replaced by the Q bit of the node reached one step earlier
// (SURVEY §8a A3), and access_kmer on the encode side by the read's own bases (A4/A5).
//
// Kernel bodies are per-thread NTC_HD functions (`*_thread`) so tests/emul can run them on the CPU.
#pragma once
#include "device_common.cuh"

namespace ntc {

constexpr int kPackThreads = 256;
constexpr int kMsThreads = 128;
constexpr uint32_t kLiteralMax = 11;  // bitnuc_threshold, src/encode.rs:137

// ------------------------------------------------------------------------------------------
// ASCII -> 2 bit.  Thread w packs bases [32w, 32w+32) of the chunk that starts at its first read.
// ------------------------------------------------------------------------------------------
NTC_HD uint32_t ascii_code(uint32_t b, uint32_t* bad) {
    uint32_t x = (b >> 1) & 3u;          // A->0 C->1 G->3 T->2
    uint32_t code = x ^ (x >> 1);        // A0 C1 G2 T3
    uint32_t expect = (0x54474341u >> (code * 8)) & 0xFFu;
    *bad |= (expect != b);
    return code;
}

NTC_HD void pack2bit_thread(const uint8_t* __restrict__ bases, const uint64_t* __restrict__ offsets, uint64_t n_bases,
                            uint64_t* __restrict__ packed, uint32_t* __restrict__ status, uint64_t w) {
    const uint8_t* __restrict__ src = bases + NTC_LDG(offsets);
    uint64_t p0 = w * 32;
    if (p0 >= n_bases) return;
    uint64_t out = 0;
    uint32_t bad = 0;
    const uint8_t* s = src + p0;
    if (p0 + 32 <= n_bases && ((uintptr_t)s & 15) == 0) {
        uint4 v0 = NTC_LDG((const uint4*)s), v1 = NTC_LDG((const uint4*)s + 1);
        uint32_t q[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
        for (int i = 0; i < 8; i++) {
#pragma unroll
            for (int j = 0; j < 4; j++) {
                uint32_t code = ascii_code((q[i] >> (8 * j)) & 0xFFu, &bad);
                out |= (uint64_t)code << (2 * (4 * i + j));
            }
        }
    } else {
        uint32_t m = (uint32_t)tmin<uint64_t>(32, n_bases - p0);
        for (uint32_t j = 0; j < m; j++) out |= (uint64_t)ascii_code(s[j], &bad) << (2 * j);
    }
    packed[w] = out;
    if (bad) NTC_ATOMIC_OR(status, (uint32_t)ST_BAD_BASE);
}

__global__ void __launch_bounds__(kPackThreads) pack2bit_kernel(const uint8_t* __restrict__ bases,
                                                                const uint64_t* __restrict__ offsets, uint64_t n_bases,
                                                                uint64_t* __restrict__ packed, uint32_t* __restrict__ status) {
    pack2bit_thread(bases, offsets, n_bases, packed, status, (uint64_t)blockIdx.x * kPackThreads + threadIdx.x);
}

// ------------------------------------------------------------------------------------------
// segmentation of one read from its scratch entries; calls sink(L, start, end_pos, d_orig)
// for each record in reference order (read END first).  Returns the record count.
// ------------------------------------------------------------------------------------------
template <typename Sink>
NTC_HD uint32_t segment_read(const uint64_t* __restrict__ sc, uint32_t len, uint32_t k, Sink&& sink) {
    uint32_t i = len, cnt = 0;
    while (i > 0) {
        uint64_t ent = sc[i - 1];
        uint32_t d = entry_d(ent), v = entry_s(ent);
        if (d == k && i > k + 1) {                                    // src/lib.rs:180
            uint32_t ext = tmin(i - k - 1, entry_urun(sc[i - 2]));    // left_extend_kmer, src/lib.rs:186
            uint32_t new_len = k + ext, m = new_len, old_i = i;
            for (;;) {                                                // hop loop, src/lib.rs:193-203
                uint32_t dd = entry_d(sc[i - 1]);
                if (dd < m) { m -= dd; i -= dd; } else break;
            }
            uint32_t L = new_len - (m - 1);                           // src/lib.rs:199
            i -= 1;                                                   // src/lib.rs:213-214
            sink(L, v, old_i, k);
        } else {                                                      // src/lib.rs:204-211
            sink(d, v, i, d);
            i -= d;
        }
        cnt++;
    }
    return cnt;
}

// ------------------------------------------------------------------------------------------
// matching statistics + record count.  One thread per read.
// ------------------------------------------------------------------------------------------
struct CountSink {
    uint32_t big = 0;
    NTC_HD void operator()(uint32_t L, uint32_t, uint32_t, uint32_t) { big |= (L >= (1u << 24)); }
};

NTC_HD void ms_thread(const DevIndex& ix, const uint64_t* __restrict__ packed, const uint64_t* __restrict__ offsets, uint64_t n_reads,
                      uint64_t* __restrict__ scratch, uint32_t* __restrict__ counts, uint32_t* __restrict__ status, uint64_t r) {
    if (r >= n_reads) return;
    const uint64_t base0 = NTC_LDG(offsets);
    const uint64_t a = offsets[r] - base0;
    const uint32_t len = (uint32_t)(offsets[r + 1] - offsets[r]);
    if (len == 0) {                                                   // EncodeError, src/lib.rs:224
        if (counts) counts[r] = 0;
        NTC_ATOMIC_OR(status, (uint32_t)ST_EMPTY_READ);
        return;
    }
    const uint32_t n = ix.n, k = ix.k;
    uint64_t* __restrict__ sc = scratch + a;
    uint32_t s = 0, e = n, d = 0, urun = 0;
    bool absent = false;
    uint64_t word = 0;
    for (uint32_t j = 0; j < len; j++) {
        uint64_t p = a + j;
        if (j == 0 || (p & 31) == 0) word = NTC_LDG(packed + (p >> 5));
        const uint32_t c = (uint32_t)(word >> ((p & 31) * 2)) & 3u;
        const uint32_t prev_s = s, prev_d = d;
        bool fresh = true;
        for (;;) {
            uint32_t bs = s / 48u, os = s - bs * 48u;
            uint4 es = ldg_entry(ix.rank + (size_t)bs * 4 + c);
            if (fresh) {
                // the block of the previous interval start also holds its Q bit: finish entry j-1
                if (j > 0) {
                    urun = (prev_d == k && entry_qbit(es, os)) ? tmin(urun + 1u, 0xFFFFFFu) : 0u;
                    sc[j - 1] = make_entry(prev_s, prev_d, urun);
                }
                fresh = false;
            }
            uint32_t s2 = entry_rank(es, os);
            uint32_t be = e / 48u, oe = e - be * 48u;
            uint32_t e2;
            if (be == bs) e2 = entry_rank(es, oe);
            else { uint4 ee = ldg_entry(ix.rank + (size_t)be * 4 + c); e2 = entry_rank(ee, oe); }
            if (s2 < e2) { s = s2; e = e2; d = tmin(d + 1u, k); break; }  // extend_right succeeded
            if (s == 0 && e == n) { d = 0; break; }                        // base absent from the index
            // contract_left until the interval can change: for targets above max(LCS[s], LCS[e])
            // the reference loop re-probes the same empty interval (identical (d, I) result)
            int ls = s > 0 ? (int)NTC_LDG(ix.lcs + s) : -1;
            int le = e < n ? (int)NTC_LDG(ix.lcs + e) : -1;
            uint32_t t = (uint32_t)tmin((int)d - 1, tmax(ls, le));
            while (s > 0 && NTC_LDG(ix.lcs + s) >= t) s--;
            while (e < n && NTC_LDG(ix.lcs + e) >= t) e++;
            d = t;
        }
        absent |= (d == 0);
    }
    sc[len - 1] = make_entry(s, d, 0);
    if (absent) {                                                     // lib.rs:207 would underflow (Q3)
        if (counts) counts[r] = 0;
        NTC_ATOMIC_OR(status, (uint32_t)ST_BAD_BASE);
        return;
    }
    if (counts) {
        CountSink sink;
        counts[r] = segment_read(sc, len, k, sink);
        if (sink.big) NTC_ATOMIC_OR(status, (uint32_t)ST_LEN24);      // src/lib.rs:226
    }
}

__global__ void __launch_bounds__(kMsThreads) ms_kernel(DevIndex ix, const uint64_t* __restrict__ packed,
                                                        const uint64_t* __restrict__ offsets, uint64_t n_reads,
                                                        uint64_t* __restrict__ scratch, uint32_t* __restrict__ counts,
                                                        uint32_t* __restrict__ status) {
    ms_thread(ix, packed, offsets, n_reads, scratch, counts, status, (uint64_t)blockIdx.x * kMsThreads + threadIdx.x);
}

// ------------------------------------------------------------------------------------------
// emit: MODE 0 = packed u64 records (encode_dictionary . encode_sequence),
//       MODE 1 = (len, start, end) triples (encode_sequence alone)
// ------------------------------------------------------------------------------------------
template <int MODE>
struct EmitSink {
    const DevIndex& ix;
    const uint64_t* __restrict__ packed;
    uint64_t* __restrict__ out;
    uint64_t a, o, first;
    NTC_HD void operator()(uint32_t L, uint32_t v, uint32_t end_pos, uint32_t d_orig) {
        if (MODE == 0) {
            uint64_t word;
            if (L > kLiteralMax) {                                    // src/encode.rs:146-149
                word = (uint64_t)v | ((uint64_t)(L & 0xFFFFFFu) << 32) | (first << 56);
            } else {                                                  // src/encode.rs:150-156
                uint64_t lit = packed_slice(packed, a + end_pos - L, L);
                word = lit | (((first + 2) | ((uint64_t)L << 2)) << 56);
            }
            out[o] = word;
        } else {
            uint32_t end = v + 1;                                     // all nodes sharing the d-suffix
            while (end < ix.n && NTC_LDG(ix.lcs + end) >= d_orig) end++;
            out[3 * o] = L; out[3 * o + 1] = v; out[3 * o + 2] = end;
        }
        o++;
        first = 0;
    }
};

template <int MODE>
NTC_HD void emit_thread(const DevIndex& ix, const uint64_t* __restrict__ packed, const uint64_t* __restrict__ offsets, uint64_t n_reads,
                        const uint64_t* __restrict__ scratch, const uint32_t* __restrict__ counts, const uint64_t* __restrict__ rec_offsets,
                        uint64_t* __restrict__ out, uint64_t cap, uint32_t* __restrict__ status, uint64_t r) {
    if (r >= n_reads) return;
    if (counts[r] == 0) return;
    const uint64_t base0 = NTC_LDG(offsets);
    const uint64_t a = offsets[r] - base0;
    const uint32_t len = (uint32_t)(offsets[r + 1] - offsets[r]);
    if (rec_offsets[r + 1] > cap) { NTC_ATOMIC_OR(status, (uint32_t)ST_CAPACITY); return; }
    EmitSink<MODE> sink{ix, packed, out, a, rec_offsets[r], 1};
    segment_read(scratch + a, len, ix.k, sink);
}

template <int MODE>
__global__ void __launch_bounds__(kMsThreads) emit_kernel(DevIndex ix, const uint64_t* __restrict__ packed,
                                                          const uint64_t* __restrict__ offsets, uint64_t n_reads,
                                                          const uint64_t* __restrict__ scratch, const uint32_t* __restrict__ counts,
                                                          const uint64_t* __restrict__ rec_offsets, uint64_t* __restrict__ out,
                                                          uint64_t cap, uint32_t* __restrict__ status) {
    emit_thread<MODE>(ix, packed, offsets, n_reads, scratch, counts, rec_offsets, out, cap, status,
                      (uint64_t)blockIdx.x * kMsThreads + threadIdx.x);
}

// encode_dictionary alone (src/encode.rs:129-166): triples -> record words, literal bases from the
// jump table (the last 16 bases of node `start`), `first` on entry 0.
NTC_HD void dict_pack_thread(const DevIndex& ix, const uint64_t* __restrict__ triples, uint64_t n, uint64_t* __restrict__ out,
                             uint32_t* __restrict__ status, uint64_t i) {
    if (i >= n) return;
    uint64_t L = triples[3 * i], v = triples[3 * i + 1];
    uint64_t first = i == 0;
    uint64_t word;
    if (L > kLiteralMax) {
        word = (v & 0xFFFFFFFFull) | ((L & 0xFFFFFFull) << 32) | (first << 56);
    } else {
        if (v >= ix.n) { NTC_ATOMIC_OR(status, (uint32_t)ST_BAD_RECORD); out[i] = 0; return; }
        uint64_t chars16 = NTC_LDG(ix.jump + v) >> 32;
        uint64_t lit = L ? (chars16 >> (2 * (16 - L))) : 0;
        word = lit | (((first + 2) | (L << 2)) << 56);
    }
    out[i] = word;
}

__global__ void __launch_bounds__(256) dict_pack_kernel(DevIndex ix, const uint64_t* __restrict__ triples, uint64_t n,
                                                        uint64_t* __restrict__ out, uint32_t* __restrict__ status) {
    dict_pack_thread(ix, triples, n, out, status, (uint64_t)blockIdx.x * 256 + threadIdx.x);
}

// per-base (d, start) dump for the matching-statistics parity entry point
__global__ void __launch_bounds__(256) ms_dump_kernel(const uint64_t* __restrict__ scratch, uint64_t n_bases,
                                                      uint8_t* __restrict__ d_out, uint32_t* __restrict__ s_out) {
    uint64_t i = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= n_bases) return;
    uint64_t e = scratch[i];
    d_out[i] = (uint8_t)entry_d(e);
    s_out[i] = entry_s(e);
}

}  // namespace ntc
