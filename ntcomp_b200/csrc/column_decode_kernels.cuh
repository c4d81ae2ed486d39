// column_decode_kernels.cuh -- the inverse GPU column codec (row f1 of SURVEY §8, decode direction):
// Rice / minimal-binary bit streams back to values (src/decode.rs:51-77) and the four columns back to the
// u64 record stream (zip_block_contents, src/decode.rs:102-149).  gunzip stays on the host.
//
// A variable-length code stream is sequential only through ONE small piece of state: where the first code
// of a chunk starts.  The stream is cut into chunks of kCwChunkBits bits and decoded in three passes:
//   1. table    one thread per (chunk, entry state): decode the chunk speculatively from every state it can
//               be entered in; record (exit state, codes started in the chunk)
//   2. resolve  one warp per stream: compose the per-chunk tables (lanes compose segments for every state,
//               lane 0 chains the 32 segments, lanes replay their segment) -> entry state + first value
//               index of every chunk
//   3. extract  one thread per chunk: decode again from the now known entry state, store the values
// Entry states.  Minimal binary with l = floor(log2 upper): codes are l or l+1 bits, the code that straddles
// a chunk boundary belongs to the chunk it starts in, so the state is the offset (0..l) of the first code
// start.  Rice with parameter b (unary quotient, a one, b remainder bits): S = chunk starts on a code start,
// P_k = k remainder bits of the previous chunk's last code still to skip (k = 1..b), M = that code's unary
// run is still going (its terminator is this chunk's first one bit, then b bits follow).
// Bits are MSB-first in u64 words (dsi-bitstream BE order after a byte swap of the stored words).
#pragma once
#include "column_kernels.cuh"
#include "device_common.cuh"
#include "scan.cuh"

namespace ntc {

constexpr uint32_t kCwChunkBits = 2048;
constexpr uint32_t kCwMaxStates = 66;   // Rice: b + 2 <= 65, minimal binary: l + 1 <= 64

#if defined(__CUDA_ARCH__)
#define NTC_CLZLL(x) __clzll((long long)(x))
#else
#define NTC_CLZLL(x) __builtin_clzll(x)
#endif

// the 64 bits that start at bit position p of the stream (zero-padded past the end)
NTC_HD uint64_t cw_window(const uint64_t* __restrict__ words, uint64_t n_words, uint64_t p) {
    const uint64_t w = p >> 6;
    const uint32_t sh = (uint32_t)(p & 63);
    const uint64_t hi = w < n_words ? words[w] : 0;
    if (sh == 0) return hi;
    const uint64_t lo = w + 1 < n_words ? words[w + 1] : 0;
    return (hi << sh) | (lo >> (64 - sh));
}

NTC_HD uint32_t cw_num_states(const CodeParams& p) { return p.rice ? p.b + 2u : p.l + 1u; }

struct CwNoSink {
    NTC_HD void operator()(uint64_t, uint64_t) const {}
};
struct CwStoreSink {
    uint64_t* out;
    uint64_t base, n;
    NTC_HD void operator()(uint64_t i, uint64_t v) const { if (base + i < n) out[base + i] = v; }
};

// Decodes every code that STARTS in [begin + off, begin + kCwChunkBits); sink(i, value) for the i-th of them.
// Returns exit state | count << 8.
template <typename Sink>
NTC_HD uint32_t cw_run(const CodeParams& p, const uint64_t* __restrict__ words, uint64_t n_words, uint64_t begin, uint32_t off,
                       Sink&& sink) {
    const uint64_t total = n_words * 64, end = begin + kCwChunkBits;
    uint64_t pos = begin + off;
    uint32_t count = 0, exit_state = 0;
    if (p.rice) {
        const uint32_t b = p.b;
        while (pos < end && pos < total) {
            uint64_t zeros = 0;
            bool found = false;
            while (pos + zeros < total) {                              // unary part: zeros up to the terminating one
                const uint64_t win = cw_window(words, n_words, pos + zeros);
                if (win) { zeros += (uint64_t)NTC_CLZLL(win); found = true; break; }
                zeros += 64;
            }
            const uint64_t term = pos + zeros;
            if (!found || term >= total) { exit_state = 0; break; }   // zero padding after the last code
            if (term + 1 + b > total) { exit_state = 0; break; }       // terminator found but the remainder bits are cut off: not a
                                                                       // code (the host reader rejects the column too, container.cpp)
            const uint64_t rem = b ? cw_window(words, n_words, term + 1) >> (64 - b) : 0;
            sink(count, (zeros << b) | rem);
            count++;
            pos = term + 1 + b;
            if (term >= end) { exit_state = b + 1; break; }            // M: the next chunk(s) hold the rest of the unary run
            exit_state = pos > end ? (uint32_t)(pos - end) : 0u;       // P_k or S
        }
    } else {
        const uint32_t l = p.l;
        while (pos < end) {
            if (pos + l > total) break;
            uint64_t x = cw_window(words, n_words, pos) >> (64 - l);  // l >= 1 (upper >= 2)
            uint32_t len = l;
            if (x >= p.limit) {
                if (pos + l + 1 > total) break;
                x = (cw_window(words, n_words, pos) >> (63 - l)) - p.limit;   // l + 1 <= 64 bits
                len = l + 1;
            }
            sink(count, x - 1);                                        // src/decode.rs:74
            count++;
            pos += len;
        }
        exit_state = pos > end ? (uint32_t)(pos - end) : 0u;
    }
    return exit_state | (count << 8);
}

// entry state -> run.  Shared by the table and the extract pass.
template <typename Sink>
NTC_HD uint32_t cw_chunk(const CodeParams& p, const uint64_t* __restrict__ words, uint64_t n_words, uint64_t chunk, uint32_t state,
                         Sink&& sink) {
    const uint64_t begin = chunk * kCwChunkBits;
    if (!p.rice || state <= p.b) return cw_run(p, words, n_words, begin, state, sink);   // offset = state (S = 0, P_k = k)
    // M: skip the rest of the unary run (first one bit of the chunk) and the b remainder bits after it
    const uint64_t total = n_words * 64;
    for (uint32_t z = 0; z < kCwChunkBits && begin + z < total; z += 64) {
        const uint64_t win = cw_window(words, n_words, begin + z);
        if (win == 0) continue;
        const uint32_t o = z + (uint32_t)NTC_CLZLL(win) + 1u + p.b;
        if (o >= kCwChunkBits) return o - kCwChunkBits;                // P_k / S, no code starts here
        return cw_run(p, words, n_words, begin, o, sink);
    }
    return p.b + 1u;                                                   // all zeros: still M
}

// ---- segments: every (block, column) pair of a batch is one code stream ---------------------------------
// All streams of a batch share the word array, the table, the chunk arrays and the value array; the three passes
// are ONE launch each over all of them.
struct DecSeg {
    uint64_t word_base, n_words;     // its words (MSB-first native u64 after the byte swap)
    uint64_t n_values, val_base;     // values announced by the header / where they go in the value array
    uint64_t chunk_base, n_chunks;   // its chunks in the batch's chunk arrays
    uint64_t table_base;             // its first table entry (n_chunks * ns entries)
    CodeParams p;
    uint32_t ns, pad;
};
// last segment s with key(s) <= x (segments without chunks share their successor's bases and are skipped)
template <typename Key>
NTC_HD uint32_t dec_seg_of(const DecSeg* __restrict__ segs, uint32_t n_segs, uint64_t x, Key key) {
    uint32_t lo = 0, hi = n_segs;
    while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (key(segs[mid]) <= x) lo = mid; else hi = mid;
    }
    return lo;
}
struct KeyTable { NTC_HD uint64_t operator()(const DecSeg& s) const { return s.table_base; } };
struct KeyChunk { NTC_HD uint64_t operator()(const DecSeg& s) const { return s.chunk_base; } };

NTC_HD void cw_table_thread(const DecSeg* __restrict__ segs, uint32_t n_segs, const uint64_t* __restrict__ words, uint64_t n_table,
                            uint32_t* __restrict__ table, uint64_t idx) {
    if (idx >= n_table) return;
    const DecSeg& s = segs[dec_seg_of(segs, n_segs, idx, KeyTable{})];
    const uint64_t local = idx - s.table_base;
    if (local >= s.n_chunks * s.ns) return;
    const uint64_t chunk = local / s.ns;
    const uint32_t state = (uint32_t)(local - chunk * s.ns);
    table[idx] = cw_chunk(s.p, words + s.word_base, s.n_words, chunk, state, CwNoSink{});
}

NTC_HD void cw_extract_thread(const DecSeg* __restrict__ segs, uint32_t n_segs, const uint64_t* __restrict__ words, uint64_t n_chunks_total,
                              const uint32_t* __restrict__ entry, const uint64_t* __restrict__ first_index, uint64_t* __restrict__ values,
                              uint64_t c) {
    if (c >= n_chunks_total) return;
    const DecSeg& s = segs[dec_seg_of(segs, n_segs, c, KeyChunk{})];
    const uint64_t chunk = c - s.chunk_base;
    if (chunk >= s.n_chunks || first_index[c] >= s.n_values) return;
    cw_chunk(s.p, words + s.word_base, s.n_words, chunk, entry[c], CwStoreSink{values + s.val_base, first_index[c], s.n_values});
}

// ---- resolve: entry state and first value index of every chunk -------------------------------------
// sequential statement (what the warp version below computes; tests/emul runs this one)
NTC_HD void cw_resolve_serial(const DecSeg& s, const uint32_t* __restrict__ table, uint32_t* __restrict__ entry,
                              uint64_t* __restrict__ first_index, uint64_t* __restrict__ total) {
    uint32_t state = 0;
    uint64_t cnt = 0;
    for (uint64_t t = 0; t < s.n_chunks; t++) {
        entry[s.chunk_base + t] = state;
        first_index[s.chunk_base + t] = cnt;
        const uint32_t e = table[s.table_base + t * s.ns + state];
        state = e & 0xFFu;
        cnt += e >> 8;
    }
    *total = cnt;
}

#if defined(__CUDACC__)
__global__ void __launch_bounds__(128) cw_table_kernel(const DecSeg* __restrict__ segs, uint32_t n_segs, const uint64_t* __restrict__ words,
                                                       uint64_t n_table, uint32_t* __restrict__ table) {
    cw_table_thread(segs, n_segs, words, n_table, table, (uint64_t)blockIdx.x * 128 + threadIdx.x);
}
__global__ void __launch_bounds__(128) cw_extract_kernel(const DecSeg* __restrict__ segs, uint32_t n_segs, const uint64_t* __restrict__ words,
                                                         uint64_t n_chunks_total, const uint32_t* __restrict__ entry,
                                                         const uint64_t* __restrict__ first_index, uint64_t* __restrict__ values) {
    cw_extract_thread(segs, n_segs, words, n_chunks_total, entry, first_index, values, (uint64_t)blockIdx.x * 128 + threadIdx.x);
}

// One warp per stream (grid = segments).  Lane L owns chunks [L*seg, (L+1)*seg): (A) it composes their tables for
// every entry state (independent chains, plenty of loads in flight), (B) lane 0 chains the 32 composed segments,
// (C) every lane replays its segment from its now known entry state.  A single thread walking n_chunks
// dependent table lookups would take milliseconds per stream.
__global__ void __launch_bounds__(32) cw_resolve_kernel(const DecSeg* __restrict__ segs, const uint32_t* __restrict__ table_all,
                                                        uint32_t* __restrict__ entry_all, uint64_t* __restrict__ first_all,
                                                        uint64_t* __restrict__ totals) {
    __shared__ uint32_t f_state[32][kCwMaxStates];
    __shared__ uint64_t f_count[32][kCwMaxStates];
    __shared__ uint32_t in_state[32];
    __shared__ uint64_t in_count[32];
    const DecSeg s = segs[blockIdx.x];
    const uint32_t* __restrict__ table = table_all + s.table_base;
    uint32_t* __restrict__ entry = entry_all + s.chunk_base;
    uint64_t* __restrict__ first_index = first_all + s.chunk_base;
    const uint32_t ns = s.ns, lane = threadIdx.x;
    const uint64_t n_chunks = s.n_chunks;
    if (n_chunks == 0) { if (lane == 0) totals[blockIdx.x] = 0; return; }
    const uint64_t seg = (n_chunks + 31) / 32;
    const uint64_t lo = tmin<uint64_t>(n_chunks, lane * seg), hi = tmin<uint64_t>(n_chunks, lo + seg);
    for (uint32_t st = 0; st < ns; st++) {
        uint32_t state = st;
        uint64_t cnt = 0;
        for (uint64_t t = lo; t < hi; t++) {
            const uint32_t e = __ldg(table + t * ns + state);
            state = e & 0xFFu;
            cnt += e >> 8;
        }
        f_state[lane][st] = state;
        f_count[lane][st] = cnt;
    }
    __syncwarp();
    if (lane == 0) {
        uint32_t state = 0;
        uint64_t cnt = 0;
        for (uint32_t L = 0; L < 32; L++) {
            in_state[L] = state;
            in_count[L] = cnt;
            cnt += f_count[L][state];
            state = f_state[L][state];
        }
        totals[blockIdx.x] = cnt;
    }
    __syncwarp();
    uint32_t state = in_state[lane];
    uint64_t cnt = in_count[lane];
    for (uint64_t t = lo; t < hi; t++) {
        entry[t] = state;
        first_index[t] = cnt;
        const uint32_t e = __ldg(table + t * ns + state);
        state = e & 0xFFu;
        cnt += e >> 8;
    }
}
#endif

// ---- zip: four columns -> record words (zip_block_contents, src/decode.rs:102-149) -------------------
// scan input over the flags column: a = 1 for reference records, b = literal bases of the record
struct FlagLoad {
    const uint64_t* flags;
    NTC_HD Pair64 operator()(uint64_t i) const {
        const uint32_t flag = (uint32_t)flags[i] & 0xFFu;
        if (flag & 2u) return Pair64{0, flag >> 2};
        return Pair64{1, 0};
    }
};

enum : uint32_t {
    ZIP_ERR_REF = 1u << 0,      // more reference records than colex values (host: -2)
    ZIP_ERR_LIT = 1u << 1,      // literal bases beyond the literal column (src/decode.rs:138 panics; Q5) (host: -3)
    ZIP_ERR_LEN = 1u << 2,      // literal longer than 32 bases: bitnuc::as_2bit rejects it (host: -4)
};

// the columns of one block inside the batch's value array
struct ZipBlock {
    uint64_t rec_base, n_rec;        // its records in the batch's record stream (= its flags, in the flags region)
    uint64_t colex_base, len_base, n_ref;
    uint64_t lits_base, n_lit_words;
};

// record i of the batch.  flags = the batch's flags region (record order), ref_idx / lit_off = exclusive scans over it.
NTC_HD void zip_thread(const uint64_t* __restrict__ values, const uint64_t* __restrict__ flags, const ZipBlock* __restrict__ blocks,
                       uint32_t n_blocks, const uint64_t* __restrict__ ref_idx, const uint64_t* __restrict__ lit_off, int full_last_word,
                       uint64_t* __restrict__ rec, uint32_t* __restrict__ err, uint64_t i, uint64_t n) {
    if (i >= n) return;
    uint32_t lo = 0, hi = n_blocks;                                    // last block with rec_base <= i that has records
    while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (blocks[mid].rec_base <= i) lo = mid; else hi = mid;
    }
    const ZipBlock& b = blocks[lo];
    const uint64_t flag = flags[i] & 0xFFu;
    uint64_t w = 0;
    if ((flag & 2u) == 0) {
        const uint64_t r = ref_idx[i] - ref_idx[b.rec_base];
        if (r >= b.n_ref) { NTC_ATOMIC_OR(err, (uint32_t)ZIP_ERR_REF); rec[i] = 0; return; }
        w = (values[b.colex_base + r] & 0xFFFFFFFFull) | ((values[b.len_base + r] & 0xFFFFFFull) << 32);
    } else {
        const uint64_t L = flag >> 2, lit_total = lit_off[b.rec_base + b.n_rec] - lit_off[b.rec_base];
        if (L > 32) { NTC_ATOMIC_OR(err, (uint32_t)ZIP_ERR_LEN); rec[i] = 0; return; }
        uint64_t last_held = lit_total % 31;                           // src/decode.rs:114-118
        if (full_last_word && last_held == 0) last_held = 31;
        uint64_t p = lit_off[i] - lit_off[b.rec_base];
        for (uint64_t j = 0; j < L; j++, p++) {
            const uint64_t wi = p / 31, off = p % 31;
            const uint64_t held = wi + 1 == b.n_lit_words ? last_held : 31;
            if (wi >= b.n_lit_words || off >= held) { NTC_ATOMIC_OR(err, (uint32_t)ZIP_ERR_LIT); rec[i] = 0; return; }
            w |= ((values[b.lits_base + wi] >> (2 * off)) & 3ull) << (2 * j);
        }
        w &= 0x00FFFFFFFFFFFFFFull;                                    // only 7 bytes are kept (src/decode.rs:142)
    }
    rec[i] = w | (flag << 56);
}

// Records in front of a block's first `first` flag belong to no sequence: the reference decodes every block by itself
// and drops them (src/lib.rs:308-317,363-365).  When the blocks of a batch are decoded as ONE stream they would be
// glued to the previous block's last sequence, so they are replaced by empty literal records (flag 2, no bases).
NTC_HD uint64_t orphan_record() { return 2ull << 56; }

#if defined(__CUDACC__)
__global__ void __launch_bounds__(256) zip_kernel(const uint64_t* __restrict__ values, const uint64_t* __restrict__ flags,
                                                  const ZipBlock* __restrict__ blocks, uint32_t n_blocks, const uint64_t* __restrict__ ref_idx,
                                                  const uint64_t* __restrict__ lit_off, int full_last_word, uint64_t* __restrict__ rec,
                                                  uint32_t* __restrict__ err, uint64_t n) {
    zip_thread(values, flags, blocks, n_blocks, ref_idx, lit_off, full_last_word, rec, err, (uint64_t)blockIdx.x * 256 + threadIdx.x, n);
}
// one CTA per block: blanks the records before the block's first `first` flag (almost always none: a block written by
// an encoder starts with the END segment of a read)
__global__ void __launch_bounds__(256) blk_trim_kernel(const ZipBlock* __restrict__ blocks, uint64_t* __restrict__ rec) {
    __shared__ uint32_t first_in_warp[8];
    const ZipBlock b = blocks[blockIdx.x];
    for (uint64_t base = 0; base < b.n_rec; base += 256) {
        const uint64_t j = base + threadIdx.x;
        const bool opens = j < b.n_rec && ((rec[b.rec_base + j] >> 56) & 1u);
        const unsigned ball = __ballot_sync(0xffffffffu, opens);
        if ((threadIdx.x & 31) == 0) first_in_warp[threadIdx.x >> 5] = ball ? (uint32_t)__ffs((int)ball) - 1u + (threadIdx.x & ~31u) : 0xFFFFFFFFu;
        __syncthreads();
        uint32_t first = 0xFFFFFFFFu;                                  // first opener of this stripe of 256 records
        for (int wv = 0; wv < 8; wv++) first = first_in_warp[wv] < first ? first_in_warp[wv] : first;
        if (j < b.n_rec && threadIdx.x < first) rec[b.rec_base + j] = orphan_record();
        __syncthreads();
        if (first != 0xFFFFFFFFu) return;
    }
}
#endif

}  // namespace ntc
