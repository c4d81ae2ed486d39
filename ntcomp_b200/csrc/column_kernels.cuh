// column_kernels.cuh -- GPU column codec, encode direction (row f1 of SURVEY §8): the record stream of a BATCH of
// 65,536-read blocks is split into ntcomp's four columns (split_encoded_dictionary, src/encode.rs:168-229) and every
// column of every block is bit-packed with the reference's integer codes (Rice / minimal binary, src/encode.rs:59-94;
// dsi-bitstream conventions: MSB-first bit stream in big-endian u64 words, unary = zeros then a one) without the
// records leaving HBM.  DEFLATE and the 32-byte headers stay on the host (container.cpp).
//
// No per-value global atomics anywhere (the first version, r1, did one atomicAdd per record on a single address, one
// atomicOr per literal base and one or two per coded value):
//   * column values are scattered to their prefix-summed ranks with plain stores, as u32 / u32 / u8 / u64;
//   * max colex / sum len / sum flags / max literal word are block-wide shuffle reductions, one atomic per CTA;
//   * literal words and code words are GATHERED: one thread per OUTPUT word finds the first record / code that overlaps
//     it by binary search in the prefix sums (L2 resident), assembles the whole 64-bit word in registers with shifts
//     and writes it once, byte-swapped, with a plain store.  Long unary runs (a Rice code of thousands of zeros) cost
//     nothing: words inside the run find no code tail and store zero.
// All (block, column) pairs of a batch -- "segments" -- go through ONE code-length scan and ONE packing launch.
// Kernel bodies are per-thread NTC_HD functions so tests/emul can run them on the CPU.
#pragma once
#include "device_common.cuh"
#include "scan.cuh"

namespace ntc {

// scan input for the split: a = 1 for reference records, b = literal bases of the record
struct SplitLoad {
    const uint64_t* rec;
    NTC_HD Pair64 operator()(uint64_t i) const {
        uint32_t flag = (uint32_t)(NTC_LDG(rec + i) >> 56);
        if (flag & 2u) return Pair64{0, tmin<uint32_t>(flag >> 2, 32u)};
        return Pair64{1, 0};
    }
};

// ---- integer codes -------------------------------------------------------------------------------
struct CodeParams {
    int rice;          // 1 Rice, 0 minimal binary
    uint32_t b;        // Rice: log2 of the modulus
    uint32_t l;        // minimal binary: floor(log2(upper))
    uint64_t limit;    // minimal binary: 2^(l+1) - upper
};
NTC_HD uint64_t code_bits(const CodeParams& p, uint64_t v) {
    if (p.rice) return (v >> p.b) + 1 + p.b;
    return (v + 1 < p.limit) ? p.l : p.l + 1;
}

// ---- batch geometry ------------------------------------------------------------------------------
// per-block statistics, filled on the device and read back once per batch (the code parameters are computed on the
// host: rice_parameter() uses the f64 log / exp of the reference's dsi-bitstream call, src/encode.rs:62-63)
struct BlkStats {
    uint64_t n_rec, n_ref, lit_bases, max_colex, sum_len, sum_flags, max_lit, lit_word_base;
};

// the four column value arrays of a batch
struct ColArrays {
    uint32_t* colex;   // [reference records of the batch] in record order
    uint32_t* len;     // same ranks
    uint8_t* flags;    // [records of the batch]
    uint64_t* lits;    // literal words, block b's region starts at BlkStats::lit_word_base
};
NTC_HD uint64_t col_value(const ColArrays& a, uint32_t q, uint64_t i) {
    return q == 0 ? (uint64_t)a.colex[i] : q == 1 ? (uint64_t)a.len[i] : q == 2 ? (uint64_t)a.flags[i] : a.lits[i];
}

// one (block, column) pair: the unit that has its own parameters and its own output bit stream
struct ColSeg {
    uint64_t val_base;    // first value in the column's array
    uint64_t n_values;
    uint64_t flat_base;   // first value in the flattened code-length space of the batch
    uint64_t bit_base;    // device: prefix-summed code length at flat_base
    uint64_t word_base;   // device: first output word
    uint64_t n_words;     // device: ceil(bits / 64)
    CodeParams p;
    uint32_t q;           // column 0 colex, 1 len, 2 flags, 3 literal words
    uint32_t pad;
};

// blk_rec[b] = first record of block b in the batch's stream, b = 0 .. n_blocks (blocks of block_reads reads,
// src/main.rs:152,168-177)
NTC_HD void blk_bounds_thread(const uint64_t* __restrict__ rec_off, uint64_t n_reads, uint64_t block_reads, uint64_t n_blocks,
                              uint64_t* __restrict__ blk_rec, uint64_t b) {
    if (b > n_blocks) return;
    const uint64_t r = tmin<uint64_t>(n_reads, b * block_reads);
    blk_rec[b] = rec_off[r] - rec_off[0];
}

// ---- split (src/encode.rs:175-209) ------------------------------------------------------------------
struct SplitContrib {
    uint64_t max_colex, sum_len, sum_flags;
};
// record i: flags[i]; reference records also colex / len at their rank.  Returns the record's share of the reductions.
NTC_HD SplitContrib split_thread(const uint64_t* __restrict__ rec, const uint64_t* __restrict__ ref_idx, const ColArrays& out, uint64_t i) {
    const uint64_t w = NTC_LDG(rec + i);
    const uint32_t flag = (uint32_t)(w >> 56);
    out.flags[i] = (uint8_t)flag;
    SplitContrib c{0, 0, flag};
    if ((flag & 2u) == 0) {
        const uint64_t r = ref_idx[i];
        const uint32_t cx = (uint32_t)w, ln = (uint32_t)(w >> 32) & 0xFFFFFFu;
        out.colex[r] = cx;
        out.len[r] = ln;
        c.max_colex = cx;
        c.sum_len = ln;
    }
    return c;
}

// ---- literal repack, 31 bases per word (src/encode.rs:211-226) ----------------------------------------
// word w of a block whose records are [rs, re) and whose literal bases start at lit_off[rs]: bases [31 w, 31 w + 31)
// of the block's literal base sequence, base j of the word at bits 2j..2j+1.  lit_off[i] = literal bases before record i.
NTC_HD uint64_t lit_word_thread(const uint64_t* __restrict__ rec, const uint64_t* __restrict__ lit_off, uint64_t rs, uint64_t re,
                                uint64_t w) {
    const uint64_t target = lit_off[rs] + 31 * w;
    if (target >= lit_off[re]) return 0;
    // last record i in [rs, re) with lit_off[i] <= target; the records after it up to the next increase hold no
    // literal bases, so record i is the literal record that contains base `target`
    uint64_t lo = rs, hi = re;                                         // invariant: lit_off[lo] <= target < lit_off[hi]
    while (hi - lo > 1) {
        const uint64_t mid = lo + ((hi - lo) >> 1);
        if (lit_off[mid] <= target) lo = mid; else hi = mid;
    }
    uint64_t word = 0, pos = target;
    uint32_t got = 0;
    for (uint64_t i = lo; i < re && got < 31; i++) {
        const uint64_t r = NTC_LDG(rec + i);
        const uint32_t flag = (uint32_t)(r >> 56);
        if ((flag & 2u) == 0) continue;
        const uint32_t L = tmin<uint32_t>(flag >> 2, 32u);
        const uint32_t skip = (uint32_t)(pos - lit_off[i]);            // > 0 only for the first record
        if (skip >= L) continue;
        const uint32_t take = tmin(L - skip, 31u - got);
        const uint64_t payload = r & 0x00FFFFFFFFFFFFFFull;            // 7 bytes; bases 28..31 read as A (src/encode.rs:215-216)
        const uint64_t bits = (payload >> (2 * skip)) & ((take >= 32u) ? ~0ull : ((1ull << (2 * take)) - 1ull));
        word |= bits << (2 * got);
        got += take;
        pos += take;
    }
    return word;
}

// ---- flattened code-length space -----------------------------------------------------------------------
// segment of flat value index i: last s with segs[s].flat_base <= i (segments are sorted by flat_base; empty ones share
// their successor's base and are skipped by taking the LAST match)
NTC_HD uint32_t seg_of_flat(const ColSeg* __restrict__ segs, uint32_t n_segs, uint64_t i) {
    uint32_t lo = 0, hi = n_segs;                                       // segs[lo].flat_base <= i < segs[hi].flat_base (hi = n: sentinel)
    while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (segs[mid].flat_base <= i) lo = mid; else hi = mid;
    }
    return lo;
}
struct FlatCodeLen {
    ColArrays arr;
    const ColSeg* segs;
    uint32_t n_segs;
    NTC_HD Pair64 operator()(uint64_t i) const {
        const ColSeg& s = segs[seg_of_flat(segs, n_segs, i)];
        return Pair64{code_bits(s.p, col_value(arr, s.q, s.val_base + (i - s.flat_base))), 0};
    }
};

// after the scan: bit_base, n_words and word_base of every segment (+ the batch's total); one thread does it all
// (a batch has at most a few hundred segments)
NTC_HD void seg_finish_serial(ColSeg* segs, uint32_t n_segs, const uint64_t* __restrict__ bit_off, uint64_t* __restrict__ total_words) {
    uint64_t words = 0;
    for (uint32_t s = 0; s < n_segs; s++) {
        const uint64_t b0 = bit_off[segs[s].flat_base], b1 = bit_off[segs[s].flat_base + segs[s].n_values];
        segs[s].bit_base = b0;
        segs[s].n_words = (b1 - b0 + 63) / 64;
        segs[s].word_base = words;
        words += segs[s].n_words;
    }
    *total_words = words;
}

// ---- packing: one thread per output word -----------------------------------------------------------------
// places the `clen` (<= 64) low bits of `code`, MSB first, at bit position rel (may be negative / beyond 63) of a
// 64-bit MSB-first window
NTC_HD uint64_t place_bits(uint64_t code, uint32_t clen, int64_t rel) {
    if (clen == 0 || rel >= 64 || rel + (int64_t)clen <= 0) return 0;
    if (rel < 0) {                                                      // the first -rel bits belong to the previous word
        clen -= (uint32_t)(-rel);
        code &= clen >= 64 ? ~0ull : ((1ull << clen) - 1ull);
        rel = 0;
    }
    const int64_t room = 64 - rel;                                      // >= 1
    if ((int64_t)clen <= room) return code << (room - clen);
    return code >> (clen - room);                                       // the tail belongs to the next word
}

// output word wg of the batch (words of all segments back to back); returns the word in stream order (MSB first);
// the caller stores it byte-swapped (dsi-bitstream BE)
NTC_HD uint64_t pack_word_thread(const ColArrays& arr, const ColSeg* __restrict__ segs, uint32_t n_segs,
                                 const uint64_t* __restrict__ bit_off, uint64_t wg) {
    uint32_t lo = 0, hi = n_segs;                                       // last segment with word_base <= wg and words of its own
    while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (segs[mid].word_base <= wg) lo = mid; else hi = mid;
    }
    const ColSeg& s = segs[lo];
    const uint64_t w = wg - s.word_base;
    if (w >= s.n_words) return 0;
    const uint64_t* __restrict__ bo = bit_off + s.flat_base;            // bo[j] = first bit of code j (batch-wide prefix)
    const uint64_t B0 = s.bit_base + 64 * w;
    uint64_t a = 0, b = s.n_values;                                     // last code j with bo[j] <= B0 (bo[0] = bit_base <= B0)
    while (b - a > 1) {
        const uint64_t mid = a + ((b - a) >> 1);
        if (bo[mid] <= B0) a = mid; else b = mid;
    }
    uint64_t word = 0;
    for (uint64_t j = a; j < s.n_values; j++) {
        const uint64_t start = bo[j];
        if (start >= B0 + 64) break;
        const uint64_t v = col_value(arr, s.q, s.val_base + j);
        if (s.p.rice) {                                                 // unary(v >> b) + b low bits; the zeros are already there
            const uint64_t tail = (1ull << s.p.b) | (s.p.b ? (v & ((1ull << s.p.b) - 1ull)) : 0ull);
            word |= place_bits(tail, s.p.b + 1u, (int64_t)(start + (v >> s.p.b)) - (int64_t)B0);
        } else {
            const uint64_t x = v + 1;                                   // src/encode.rs:88
            if (x < s.p.limit) word |= place_bits(x, s.p.l, (int64_t)start - (int64_t)B0);
            else word |= place_bits(x + s.p.limit, s.p.l + 1u, (int64_t)start - (int64_t)B0);
        }
    }
    return word;
}

NTC_HD uint64_t bswap64(uint64_t x) {
#if defined(__CUDA_ARCH__)
    const uint32_t lo = (uint32_t)x, hi = (uint32_t)(x >> 32);
    return ((uint64_t)__byte_perm(lo, 0, 0x0123) << 32) | __byte_perm(hi, 0, 0x0123);
#else
    return __builtin_bswap64(x);
#endif
}

#if defined(__CUDACC__)
// ---- kernels ----------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(64) blk_bounds_kernel(const uint64_t* __restrict__ rec_off, uint64_t n_reads, uint64_t block_reads,
                                                        uint64_t n_blocks, uint64_t* __restrict__ blk_rec) {
    blk_bounds_thread(rec_off, n_reads, block_reads, n_blocks, blk_rec, (uint64_t)blockIdx.x * 64 + threadIdx.x);
}

// block-wide reductions (256 threads): shuffle inside the warps, one shared slot per warp, thread 0 returns the result
__device__ __forceinline__ uint64_t block_sum_256(uint64_t v, uint64_t* slots) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) slots[threadIdx.x >> 5] = v;
    __syncthreads();
    uint64_t t = 0;
    if (threadIdx.x == 0) for (int i = 0; i < 8; i++) t += slots[i];
    __syncthreads();
    return t;
}
__device__ __forceinline__ uint64_t block_max_256(uint64_t v, uint64_t* slots) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { const uint64_t u = __shfl_down_sync(0xffffffffu, v, o); v = u > v ? u : v; }
    if ((threadIdx.x & 31) == 0) slots[threadIdx.x >> 5] = v;
    __syncthreads();
    uint64_t t = 0;
    if (threadIdx.x == 0) for (int i = 0; i < 8; i++) t = slots[i] > t ? slots[i] : t;
    __syncthreads();
    return t;
}

// grid (tiles of the largest block, n_blocks): CTA (x, y) handles records blk_rec[y] + 256 x .. of block y
__global__ void __launch_bounds__(256) split_kernel(const uint64_t* __restrict__ rec, const uint64_t* __restrict__ blk_rec,
                                                    const uint64_t* __restrict__ ref_idx, ColArrays out, BlkStats* __restrict__ stats) {
    __shared__ uint64_t slots[8];
    const uint64_t rs = blk_rec[blockIdx.y], re = blk_rec[blockIdx.y + 1];
    const uint64_t i = rs + (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (rs + (uint64_t)blockIdx.x * 256 >= re) return;                 // whole CTA beyond the block
    SplitContrib c{0, 0, 0};
    if (i < re) c = split_thread(rec, ref_idx, out, i);
    const uint64_t mx = block_max_256(c.max_colex, slots);
    const uint64_t sl = block_sum_256(c.sum_len, slots);
    const uint64_t sf = block_sum_256(c.sum_flags, slots);
    if (threadIdx.x == 0) {
        BlkStats* s = stats + blockIdx.y;
        atomicMax((unsigned long long*)&s->max_colex, (unsigned long long)mx);
        atomicAdd((unsigned long long*)&s->sum_len, (unsigned long long)sl);
        atomicAdd((unsigned long long*)&s->sum_flags, (unsigned long long)sf);
    }
}

// per-block counts from the split scan + the literal word regions; one thread (a batch has a few blocks)
__global__ void blk_stats_kernel(const uint64_t* __restrict__ blk_rec, const uint64_t* __restrict__ ref_idx,
                                 const uint64_t* __restrict__ lit_off, uint32_t n_blocks, BlkStats* __restrict__ stats) {
    if (blockIdx.x || threadIdx.x) return;
    uint64_t words = 0;
    for (uint32_t b = 0; b < n_blocks; b++) {
        const uint64_t rs = blk_rec[b], re = blk_rec[b + 1];
        stats[b].n_rec = re - rs;
        stats[b].n_ref = ref_idx[re] - ref_idx[rs];
        stats[b].lit_bases = lit_off[re] - lit_off[rs];
        stats[b].lit_word_base = words;
        words += (stats[b].lit_bases + 30) / 31;
    }
}

// grid (word tiles of the largest block, n_blocks)
__global__ void __launch_bounds__(256) lit_pack_kernel(const uint64_t* __restrict__ rec, const uint64_t* __restrict__ blk_rec,
                                                       const uint64_t* __restrict__ lit_off, uint64_t* __restrict__ lits,
                                                       BlkStats* __restrict__ stats) {
    __shared__ uint64_t slots[8];
    BlkStats* s = stats + blockIdx.y;
    const uint64_t n_words = (s->lit_bases + 30) / 31;
    const uint64_t w = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if ((uint64_t)blockIdx.x * 256 >= n_words) return;
    uint64_t word = 0;
    if (w < n_words) {
        word = lit_word_thread(rec, lit_off, blk_rec[blockIdx.y], blk_rec[blockIdx.y + 1], w);
        lits[s->lit_word_base + w] = word;
    }
    const uint64_t mx = block_max_256(word, slots);
    if (threadIdx.x == 0) atomicMax((unsigned long long*)&s->max_lit, (unsigned long long)mx);
}

__global__ void seg_finish_kernel(ColSeg* segs, uint32_t n_segs, const uint64_t* __restrict__ bit_off, uint64_t* __restrict__ total_words) {
    if (blockIdx.x || threadIdx.x) return;
    seg_finish_serial(segs, n_segs, bit_off, total_words);
}

__global__ void __launch_bounds__(256) pack_words_kernel(ColArrays arr, const ColSeg* __restrict__ segs, uint32_t n_segs,
                                                         const uint64_t* __restrict__ bit_off, const uint64_t* __restrict__ total_words,
                                                         uint64_t* __restrict__ words) {
    const uint64_t wg = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (wg >= *total_words) return;
    words[wg] = bswap64(pack_word_thread(arr, segs, n_segs, bit_off, wg));
}

__global__ void __launch_bounds__(256) bswap_kernel(unsigned long long* __restrict__ words, uint64_t n) {
    uint64_t i = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (i < n) words[i] = bswap64(words[i]);
}
#endif

}  // namespace ntc
