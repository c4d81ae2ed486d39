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

NTC_HD void cw_table_thread(const CodeParams& p, const uint64_t* __restrict__ words, uint64_t n_words, uint64_t n_chunks,
                            uint32_t* __restrict__ table, uint64_t idx) {
    const uint32_t ns = cw_num_states(p);
    if (idx >= n_chunks * ns) return;
    const uint64_t chunk = idx / ns;
    const uint32_t state = (uint32_t)(idx - chunk * ns);
    table[idx] = cw_chunk(p, words, n_words, chunk, state, CwNoSink{});
}
__global__ void __launch_bounds__(128) cw_table_kernel(CodeParams p, const uint64_t* __restrict__ words, uint64_t n_words,
                                                       uint64_t n_chunks, uint32_t* __restrict__ table) {
    cw_table_thread(p, words, n_words, n_chunks, table, (uint64_t)blockIdx.x * 128 + threadIdx.x);
}

NTC_HD void cw_extract_thread(const CodeParams& p, const uint64_t* __restrict__ words, uint64_t n_words, uint64_t n_chunks,
                              const uint32_t* __restrict__ entry, const uint64_t* __restrict__ first_index, uint64_t* __restrict__ out,
                              uint64_t n, uint64_t chunk) {
    if (chunk >= n_chunks) return;
    if (first_index[chunk] >= n) return;
    cw_chunk(p, words, n_words, chunk, entry[chunk], CwStoreSink{out, first_index[chunk], n});
}
__global__ void __launch_bounds__(128) cw_extract_kernel(CodeParams p, const uint64_t* __restrict__ words, uint64_t n_words,
                                                         uint64_t n_chunks, const uint32_t* __restrict__ entry,
                                                         const uint64_t* __restrict__ first_index, uint64_t* __restrict__ out, uint64_t n) {
    cw_extract_thread(p, words, n_words, n_chunks, entry, first_index, out, n, (uint64_t)blockIdx.x * 128 + threadIdx.x);
}

// ---- resolve: entry state and first value index of every chunk -------------------------------------
// sequential statement (what the warp version below computes; tests/emul runs this one)
NTC_HD void cw_resolve_serial(const uint32_t* __restrict__ table, uint32_t ns, uint64_t n_chunks, uint32_t* __restrict__ entry,
                              uint64_t* __restrict__ first_index, uint64_t* __restrict__ total) {
    uint32_t state = 0;
    uint64_t cnt = 0;
    for (uint64_t t = 0; t < n_chunks; t++) {
        entry[t] = state;
        first_index[t] = cnt;
        const uint32_t e = table[t * ns + state];
        state = e & 0xFFu;
        cnt += e >> 8;
    }
    *total = cnt;
}

#if defined(__CUDACC__)
// One warp per stream.  Lane L owns chunks [L*seg, (L+1)*seg): (A) it composes their tables for every entry
// state (independent chains, plenty of loads in flight), (B) lane 0 chains the 32 composed segments,
// (C) every lane replays its segment from its now known entry state.  A single thread walking n_chunks
// dependent table lookups would take milliseconds per stream.
__global__ void __launch_bounds__(32) cw_resolve_kernel(const uint32_t* __restrict__ table, uint32_t ns, uint64_t n_chunks,
                                                        uint32_t* __restrict__ entry, uint64_t* __restrict__ first_index,
                                                        uint64_t* __restrict__ total) {
    __shared__ uint32_t f_state[32][kCwMaxStates];
    __shared__ uint64_t f_count[32][kCwMaxStates];
    __shared__ uint32_t in_state[32];
    __shared__ uint64_t in_count[32];
    const uint32_t lane = threadIdx.x;
    const uint64_t seg = (n_chunks + 31) / 32;
    const uint64_t lo = tmin<uint64_t>(n_chunks, lane * seg), hi = tmin<uint64_t>(n_chunks, lo + seg);
    for (uint32_t s = 0; s < ns; s++) {
        uint32_t state = s;
        uint64_t cnt = 0;
        for (uint64_t t = lo; t < hi; t++) {
            const uint32_t e = __ldg(table + t * ns + state);
            state = e & 0xFFu;
            cnt += e >> 8;
        }
        f_state[lane][s] = state;
        f_count[lane][s] = cnt;
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
        *total = cnt;
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

// lit_total = literal bases of the whole block (scan total); full_last_word: see NTC_BLOCK_FULL_LAST_LITERAL_WORD
NTC_HD void zip_thread(const uint64_t* __restrict__ colex, const uint64_t* __restrict__ len, uint64_t n_ref,
                       const uint64_t* __restrict__ flags, uint64_t n, const uint64_t* __restrict__ lits, uint64_t n_lit_words,
                       const uint64_t* __restrict__ ref_idx, const uint64_t* __restrict__ lit_off, const uint64_t* __restrict__ scan_totals,
                       int full_last_word, uint64_t* __restrict__ rec, uint32_t* __restrict__ err, uint64_t i) {
    if (i >= n) return;
    const uint64_t flag = flags[i] & 0xFFu;
    uint64_t w = 0;
    if ((flag & 2u) == 0) {
        const uint64_t r = ref_idx[i];
        if (r >= n_ref) { NTC_ATOMIC_OR(err, (uint32_t)ZIP_ERR_REF); return; }
        w = (colex[r] & 0xFFFFFFFFull) | ((len[r] & 0xFFFFFFull) << 32);
    } else {
        const uint64_t L = flag >> 2, lit_total = scan_totals[1];
        if (L > 32) { NTC_ATOMIC_OR(err, (uint32_t)ZIP_ERR_LEN); return; }
        uint64_t last_held = lit_total % 31;                           // src/decode.rs:114-118
        if (full_last_word && last_held == 0) last_held = 31;
        uint64_t p = lit_off[i];
        for (uint64_t j = 0; j < L; j++, p++) {
            const uint64_t wi = p / 31, off = p % 31;
            const uint64_t held = wi + 1 == n_lit_words ? last_held : 31;
            if (wi >= n_lit_words || off >= held) { NTC_ATOMIC_OR(err, (uint32_t)ZIP_ERR_LIT); return; }
            w |= ((lits[wi] >> (2 * off)) & 3ull) << (2 * j);
        }
        w &= 0x00FFFFFFFFFFFFFFull;                                    // only 7 bytes are kept (src/decode.rs:142)
    }
    rec[i] = w | (flag << 56);
}
__global__ void __launch_bounds__(256) zip_kernel(const uint64_t* __restrict__ colex, const uint64_t* __restrict__ len, uint64_t n_ref,
                                                  const uint64_t* __restrict__ flags, uint64_t n, const uint64_t* __restrict__ lits,
                                                  uint64_t n_lit_words, const uint64_t* __restrict__ ref_idx,
                                                  const uint64_t* __restrict__ lit_off, const uint64_t* __restrict__ scan_totals,
                                                  int full_last_word, uint64_t* __restrict__ rec, uint32_t* __restrict__ err) {
    zip_thread(colex, len, n_ref, flags, n, lits, n_lit_words, ref_idx, lit_off, scan_totals, full_last_word, rec, err,
               (uint64_t)blockIdx.x * 256 + threadIdx.x);
}

}  // namespace ntc
