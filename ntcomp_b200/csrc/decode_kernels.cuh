// decode_kernels.cuh -- the decode hot path as sm_100a kernels (replaces decode_sequence,
// src/lib.rs:254-318, and the FASTA writer loop of src/main.rs:201-209).
//
// Every reference record is a plain backward walk from its colex rank: access_kmer's
// k selects and left_extend_kmer2's 4 x search per extended base collapse into inverse-LF steps,
// and the jump table takes 16 of those steps per 8-byte load (pred^16 | last 16 bases).
// Kernel bodies are per-thread NTC_HD functions (`*_thread`) so tests/emul can run them on the CPU.
#pragma once
#include "device_common.cuh"
#include "scan.cuh"

namespace ntc {

NTC_HD uint32_t record_len(uint64_t w) {
    uint32_t flag = (uint32_t)(w >> 56);
    return (flag & 2u) ? (flag >> 2) : (uint32_t)(w >> 32) & 0xFFFFFFu;
}
NTC_HD uint32_t record_first(uint64_t w) { return (uint32_t)(w >> 56) & 1u; }

// scan input: a = bases of the record, b = 1 on the first record of a sequence
struct RecordLoad {
    const uint64_t* rec;
    NTC_HD Pair64 operator()(uint64_t i) const {
        uint64_t w = NTC_LDG(rec + i);
        uint32_t L = record_len(w);
        if (((w >> 56) & 2u) && L > 32u) L = 0;  // bitnuc::from_2bit rejects > 32, result ignored (src/lib.rs:303)
        return Pair64{L, record_first(w)};
    }
};

// number of decimal digits of x >= 1
NTC_HD uint32_t dec_digits(uint64_t x) {
    uint32_t d = 1;
    while (x >= 10) { x /= 10; d++; }
    return d;
}
// sum of dec_digits(i) for i in [1, x]
NTC_HD uint64_t digits_prefix(uint64_t x) {
    uint64_t total = 0, lo = 1;
    uint32_t d = 1;
    while (lo <= x) {
        uint64_t hi = lo * 10 - 1;
        if (hi > x || hi < lo) hi = x;
        total += (hi - lo + 1) * d;
        if (hi == x) break;
        lo *= 10;
        d++;
    }
    return total;
}
// bytes of FASTA text before the bases of sequence `sid` (0-based in this batch), excluding bases:
// headers ">seq.N\n" of sequences 0..sid and the '\n' closing sequences 0..sid-1
NTC_HD uint64_t fasta_overhead(uint64_t first_no, uint64_t sid) {
    uint64_t digits = digits_prefix(first_no + sid) - digits_prefix(first_no - 1);
    return (sid + 1) * 6 + digits + sid;
}

// seq_first[j] = index of the record that opens sequence j: noted by the record scan's last pass (scan_down_kernel's
// post hook; a kernel of its own until r1 v12 -- 0.18 ms per 39 M records for re-reading the records and their prefixes)
struct SeqFirstPost {
    uint64_t* seq_first;
    NTC_HD void operator()(uint64_t i, const Pair64& prefix, const Pair64& v) const {
        if (v.b) seq_first[prefix.b] = i;
    }
};

// ---- per-sequence layout ---------------------------------------------------------------------
// One thread per sequence j (plus one for the end marker): everything the walk needs about a sequence
// collapses into ONE word, seq_endkey[j] = (output position of the sequence's first base) + B_j, where
// [A_j, B_j) are the sequence's bases in record order coordinates (base_off); the segment of record i
// then ends at seq_endkey[sid] - base_off[i].  FASTA header digits are counted here, per sequence, not
// per record (ncu r1: digit arithmetic per record was ~1/3 of walk_kernel's instructions).
// counts[0] = n_seqs, counts[1] = output bytes; seq_offsets (optional) = bases before sequence j + carry.
NTC_HD void seq_layout_thread(const uint64_t* __restrict__ base_off, const uint64_t* __restrict__ seq_first,
                              const uint64_t* __restrict__ totals /* [bases, n_seqs] */, uint64_t n_rec,
                              uint64_t* __restrict__ seq_endkey, uint64_t* __restrict__ seq_offsets, uint64_t cap_seqs, uint64_t off_carry,
                              uint64_t first_no, int fasta, uint64_t cap_bytes, uint64_t* __restrict__ counts,
                              uint32_t* __restrict__ status, uint64_t j) {
    const uint64_t n_seqs = totals[1];
    if (j > n_seqs) return;
    const uint64_t base0 = n_seqs ? base_off[seq_first[0]] : 0;
    if (j == 0) {
        uint64_t n_bases = n_seqs ? base_off[n_rec] - base0 : 0;
        uint64_t bytes = fasta ? (n_seqs ? n_bases + fasta_overhead(first_no, n_seqs - 1) + 1 : 0) : n_bases;
        counts[0] = n_seqs;
        counts[1] = bytes;
        if (bytes > cap_bytes || (seq_offsets && n_seqs > cap_seqs)) NTC_ATOMIC_OR(status, (uint32_t)ST_CAPACITY);
    }
    const uint64_t A = j < n_seqs ? base_off[seq_first[j]] : base_off[n_rec];
    if (seq_offsets && j <= cap_seqs) seq_offsets[j] = A - base0 + off_carry;
    if (j == n_seqs) return;
    const uint64_t B = j + 1 < n_seqs ? base_off[seq_first[j + 1]] : base_off[n_rec];
    const uint64_t lo = A - base0 + (fasta ? fasta_overhead(first_no, j) : 0);
    seq_endkey[j] = lo + B;
}
__global__ void __launch_bounds__(256) seq_layout_kernel(const uint64_t* __restrict__ base_off, const uint64_t* __restrict__ seq_first,
                                                         const uint64_t* __restrict__ totals, uint64_t n_rec,
                                                         uint64_t* __restrict__ seq_endkey, uint64_t* __restrict__ seq_offsets,
                                                         uint64_t cap_seqs, uint64_t off_carry, uint64_t first_no, int fasta,
                                                         uint64_t cap_bytes, uint64_t* __restrict__ counts, uint32_t* __restrict__ status) {
    seq_layout_thread(base_off, seq_first, totals, n_rec, seq_endkey, seq_offsets, cap_seqs, off_carry, first_no, fasta, cap_bytes, counts,
                      status, (uint64_t)blockIdx.x * 256 + threadIdx.x);
}

// ---- leftward write combiner ---------------------------------------------------------------
// A record's bases come out right-to-left, 16 per jump-table load.  Bytes are merged into aligned
// 8-byte words so the interior of a segment is written with full u64 stores; only the two edges
// of a segment (which share words with neighbouring segments / headers) use 4/2/1-byte stores.
// (First version, ncu r1: 1.59 G single-byte stores made walk_kernel L1TEX-bound.)
// (Rejected, profiles/r2/ab_walk_stream.txt: evict_first text stores -- 2.22 vs 1.99 ms on C3: a text sector is completed by
// several threads at different times, and a partially written evict_first sector leaves L2 before the rest arrives.)
NTC_HD void store_lanes(uint8_t* word_addr, uint64_t w, uint32_t lo, uint32_t hi, uint64_t pol) {  // lanes [lo, hi) of the aligned word
    (void)pol;
    if (lo == 0 && hi == 8) { *(uint64_t*)word_addr = w; return; }
    uint32_t l = lo;
    while (l < hi) {
        if ((l & 3u) == 0 && l + 4 <= hi) { *(uint32_t*)(word_addr + l) = (uint32_t)(w >> (8 * l)); l += 4; }
        else if ((l & 1u) == 0 && l + 2 <= hi) { *(uint16_t*)(word_addr + l) = (uint16_t)(w >> (8 * l)); l += 2; }
        else { word_addr[l] = (uint8_t)(w >> (8 * l)); l += 1; }
    }
}
NTC_HD void store_word(uint8_t* addr, uint64_t w, uint64_t pol) {
    (void)pol;
    *(uint64_t*)addr = w;
}

// Complete words leave in descending address order; the upper half of a 16-byte unit waits in `hold` for the lower
// half, and the two go out as one 16-byte store (ncu r1 v14: after the cost sort walk_kernel ran at the L1TEX sector
// rate, 298 M store sectors for 51 M sectors of text -- every 8-byte store of a lane is a sector access of its own).
struct LeftWriter {
    uint8_t* out;   // 16-byte aligned base of the output buffer
    uint64_t p;     // bytes [p, segment end) are done; pending bytes of the word holding p sit in acc
    uint64_t acc;   // lanes [p & 7, top) of the aligned word at p & ~7
    uint64_t hold;  // complete word at hold_pos (an upper half) waiting for its lower half
    uint64_t pol;   // L2 policy of the text stores (evict_first)
    uint32_t top;
    bool have;

    NTC_HD void init(uint8_t* base, uint64_t end, uint64_t policy) { out = base; p = end; acc = 0; hold = 0; have = false; top = (uint32_t)(end & 7); pol = policy; }
    // lanes [0, hi) of the aligned word at byte position pos (a multiple of 8)
    NTC_HD void put(uint64_t pos, uint64_t w, uint32_t hi) {
        if (hi != 8) { store_lanes(out + pos, w, 0, hi, pol); return; } // right edge of the segment: nothing is held yet
        if (pos & 8) {
            if (have) store_word(out + pos + 8, hold, pol);            // (not reached: words arrive 8 bytes apart)
            hold = w;
            have = true;
        } else if (have) {
            *(uint64_t*)(out + pos) = w;
            *(uint64_t*)(out + pos + 8) = hold;
            have = false;
        } else {
            store_word(out + pos, w, pol);
        }
    }
    // x = the 8 bytes for positions [p-8, p)
    NTC_HD void push8(uint64_t x) {
        uint32_t q = (uint32_t)(p & 7);
        if (q == 0) {
            put(p - 8, x, 8);
        } else {
            uint32_t sh = 8 * q;
            put(p & ~7ull, acc | (x >> (64 - sh)), top);
            acc = x << sh;
            top = 8;
        }
        p -= 8;
    }
    // x holds n < 8 bytes for positions [p-n, p) in its UPPER lanes (lower lanes are ignored); more may follow
    NTC_HD void push_n(uint64_t x, uint32_t n) {
        if (n == 0) return;
        x &= ~0ull << (8 * (8 - n));
        const uint32_t q = (uint32_t)(p & 7);
        if (q == 0) { acc = x; top = 8; p -= n; return; }             // opens a new word: lanes [8-n, 8) pending
        const uint32_t sh = 8 * q;
        const uint64_t w = acc | (x >> (64 - sh));
        if (n < q) { acc = w; p -= n; return; }                        // lanes [q-n, top) pending
        put(p & ~7ull, w, top);                                        // the word holding p is complete
        acc = x << sh;                                                 // the other n-q bytes open the next word
        top = 8;
        p -= n;
    }
    // writes what is pending; the writer's bytes end at p
    NTC_HD void finish() {
        const uint32_t q = (uint32_t)(p & 7);
        if (have) { store_word(out + ((p + 7) & ~7ull), hold, pol); have = false; }   // the held word is the one right of p's word
        if (q) store_lanes(out + (p & ~7ull), acc, q, top, pol);
    }
};

// ASCII of four 2-bit codes (low byte of x), first code in the low byte of the result
NTC_HD uint32_t ascii4(uint32_t x) {
#if defined(__CUDA_ARCH__)
    uint32_t sel = (x & 3u) | ((x & 0xCu) << 2) | ((x & 0x30u) << 4) | ((x & 0xC0u) << 6);
    return __byte_perm(0x54474341u, 0u, sel);
#else
    uint32_t r = 0;
    for (int i = 0; i < 4; i++) r |= (uint32_t)code_ascii((x >> (2 * i)) & 3u) << (8 * i);
    return r;
#endif
}
NTC_HD uint64_t ascii8(uint32_t codes16) {  // eight 2-bit codes in the low 16 bits
    return (uint64_t)ascii4(codes16 & 0xFFu) | ((uint64_t)ascii4((codes16 >> 8) & 0xFFu) << 32);
}

// emits the last `t` (<= 16) bases of `codes` (base i of the 16-base window at bits 2i..2i+1); the caller
// ends the segment with finish()
NTC_HD void emit_codes(LeftWriter& w, uint32_t codes, uint32_t t) {
    uint64_t hi = ascii8(codes >> 16), lo = ascii8(codes & 0xFFFFu);
    if (t == 16) { w.push8(hi); w.push8(lo); return; }
    if (t >= 8) { w.push8(hi); w.push_n(lo, t - 8); }
    else w.push_n(hi, t);
}

// FASTA framing (src/main.rs:204-206): the text between the bases of sequence j-1 and those of sequence j,
// "\n>seq.N\n" (no leading newline for j = 0), lies immediately left of sequence j's START segment, so the
// thread of that segment -- the last record of the sequence -- keeps writing leftwards: digits come out least
// significant first, and the bytes merge into the same aligned word stores as the bases.
// (Earlier versions: byte stores from the END segment's thread, then a kernel of its own with one thread per
// sequence -- 0.76 ms per 10 M sequences, most of it partial-sector writes.)
NTC_HD void emit_header(LeftWriter& lw, uint64_t no, bool leading_newline) {
    uint64_t acc = 0;
    uint32_t n = 0;
    auto put = [&](uint32_t b) {
        acc |= (uint64_t)b << (8 * (7 - n));
        if (++n == 8) { lw.push8(acc); acc = 0; n = 0; }
    };
    put('\n');
    if (no <= 0xFFFFFFFFull) { uint32_t x = (uint32_t)no; do { put('0' + x % 10u); x /= 10u; } while (x); }
    else do { put('0' + (uint32_t)(no % 10)); no /= 10; } while (no);
    put('.'); put('q'); put('e'); put('s'); put('>');
    if (leading_newline) put('\n');
    lw.push_n(acc, n);
}

// What a record's thread needs, gathered with record i on thread i (coalesced loads) before the block's records are
// re-dealt by cost: with the metadata loaded after the deal, every lane touched its own sectors of rec / base_off /
// seq_idx (ncu r1 v14: 284 M load sectors, ~100 M of them jump-table entries).
struct WalkItem {
    uint64_t w;        // record word
    uint64_t seg_end;  // output position just right of the record's last base
    uint64_t sid;      // sequence (0-based in this batch)
    uint32_t L;        // bases
    uint32_t flags;    // bit 0: something to write, bit 1: START segment of its sequence (writes the FASTA framing)
};

NTC_HD WalkItem walk_prepare(const uint64_t* __restrict__ rec, uint64_t n_rec, const uint64_t* __restrict__ base_off,
                             const uint64_t* __restrict__ seq_idx, const uint64_t* __restrict__ seq_endkey, const uint64_t* __restrict__ totals,
                             int fasta, uint8_t* __restrict__ out, uint64_t cap_bytes, const uint64_t* __restrict__ counts, uint64_t i) {
    WalkItem it{0, 0, 0, 0, 0};
    if (i >= n_rec) return it;
    if (counts[1] > cap_bytes) return it;  // capacity error already flagged
    const uint64_t w = NTC_LDG(rec + i);
    const uint32_t flag = (uint32_t)(w >> 56);
    const uint64_t incl = seq_idx[i] + (flag & 1u);  // sequences opened up to and including this record
    if (incl == 0) return it;                         // records before the first `first` flag are dropped (src/lib.rs:308-317)
    const uint64_t sid = incl - 1;
    const uint64_t bo = base_off[i];
    const uint32_t L = (uint32_t)(base_off[i + 1] - bo);
    // START segment of its sequence = last record before the next `first` flag
    const bool closes = fasta && (i + 1 == n_rec || record_first(NTC_LDG(rec + i + 1)));
    if (fasta && (flag & 1u) && sid + 1 == totals[1]) out[counts[1] - 1] = '\n';   // the stream's final newline
    if (L == 0 && !closes) return it;
    it.w = w;
    it.seg_end = seq_endkey[sid] - bo;  // records of a sequence run from its END to its START: (sequence end) - (bases before it)
    it.sid = sid;
    it.L = L;
    it.flags = 1u | (closes ? 2u : 0u);
    return it;
}

// walk / unpack one record and store its ASCII bases (+ the FASTA framing left of a START segment)
NTC_HD void walk_emit(const DevIndex& ix, const WalkItem& it, uint64_t first_no, uint8_t* __restrict__ out, uint32_t* __restrict__ status) {
    if (!(it.flags & 1u)) return;
    const uint64_t w = it.w;
    const uint32_t flag = (uint32_t)(w >> 56), L = it.L;
    // the output buffer may be misaligned by the caller: fold the misalignment into the position
    const uint64_t mis = (uint64_t)((uintptr_t)out & 15u);
    LeftWriter lw;
    lw.init(out - mis, it.seg_end + mis, ix.pol_stream);

    if (flag & 2u) {  // literal record (<= 32 bases), src/lib.rs:295-306
        uint64_t bits = L >= 32 ? (w & 0x00FFFFFFFFFFFFFFull) : (w & ((1ull << (2 * L)) - 1ull));   // 56 payload bits + zeros
        if (L > 16) {
            emit_codes(lw, (uint32_t)(bits >> (2 * (L - 16))), 16);           // last 16 bases
            uint32_t rest = L - 16;                                           // first `rest` bases, to the upper lanes of a window
            emit_codes(lw, (uint32_t)(bits << (2 * (16 - rest))), rest);
        } else if (L) {
            emit_codes(lw, (uint32_t)(bits << (2 * (16 - L))), L);
        }
    } else {
        uint32_t v = (uint32_t)w;
        if (v >= ix.n) { NTC_ATOMIC_OR(status, (uint32_t)ST_BAD_RECORD); return; }
        const uint64_t keep = ix.pol_keep;            // the jump table must survive the output stream in L2
        uint32_t remaining = L;
        while (remaining > 0) {
            uint64_t ent = ld8_keep(ix.jump + v, keep);
            uint32_t t = tmin(remaining, 16u);
            remaining -= t;
            emit_codes(lw, (uint32_t)(ent >> 32), t);
            v = (uint32_t)ent;
        }
    }
    if (it.flags & 2u) emit_header(lw, first_no + it.sid, it.sid != 0);
    lw.finish();
}

// one thread per record (the CPU emulation's entry point; the kernel deals prepared items to threads by cost)
NTC_HD void walk_thread(const DevIndex& ix, const uint64_t* __restrict__ rec, uint64_t n_rec, const uint64_t* __restrict__ base_off,
                        const uint64_t* __restrict__ seq_idx, const uint64_t* __restrict__ seq_endkey, const uint64_t* __restrict__ totals,
                        uint64_t first_no, int fasta, uint8_t* __restrict__ out, uint64_t cap_bytes, const uint64_t* __restrict__ counts,
                        uint32_t* __restrict__ status, uint64_t i) {
    const WalkItem it = walk_prepare(rec, n_rec, base_off, seq_idx, seq_endkey, totals, fasta, out, cap_bytes, counts, i);
    walk_emit(ix, it, first_no, out, status);
}

// Work assignment inside a block.  A record costs one dependent jump-table load + two word pushes per 16 bases, and the
// lengths of neighbouring records differ by an order of magnitude (150 bp reads at 1 %: 92, 58, a 1-base literal, ...), so
// with record i on thread i every warp ran as long as its longest record (ncu r1 v12: 8.9 active threads per
// instruction, issue bound).  The block's records are therefore counting-sorted by their number of 16-base steps
// (descending, 16 buckets, shared-memory histogram) and handed to the threads in that order: warps get records of
// (nearly) equal cost, the heaviest warps start first.  Each record is still written by one thread, so the
// output bytes do not depend on the order.
#ifndef NTC_WALK_THREADS
#define NTC_WALK_THREADS 256
#endif
constexpr int kWalkThreads = NTC_WALK_THREADS;
constexpr uint32_t kWalkBuckets = 16;

NTC_HD uint32_t walk_cost_key(uint64_t w) {   // 0 for literal records, else min(ceil(L / 16), 15)
    if ((w >> 56) & 2u) return 0;
    const uint32_t L = (uint32_t)(w >> 32) & 0xFFFFFFu;
    return tmin((L + 15u) >> 4, kWalkBuckets - 1u);
}

__global__ void __launch_bounds__(kWalkThreads) walk_kernel(DevIndex ix, const uint64_t* __restrict__ rec, uint64_t n_rec,
                                                   const uint64_t* __restrict__ base_off, const uint64_t* __restrict__ seq_idx,
                                                   const uint64_t* __restrict__ seq_endkey, const uint64_t* __restrict__ totals,
                                                   uint64_t first_no, int fasta, uint8_t* __restrict__ out, uint64_t cap_bytes,
                                                   const uint64_t* __restrict__ counts, uint32_t* __restrict__ status) {
    const uint64_t i0 = (uint64_t)blockIdx.x * kWalkThreads;
    const uint32_t tid = threadIdx.x;
    // record i0 + tid on thread tid: coalesced loads of everything the walk needs
    const WalkItem mine = walk_prepare(rec, n_rec, base_off, seq_idx, seq_endkey, totals, fasta, out, cap_bytes, counts, i0 + tid);
    __shared__ uint32_t s_hist[kWalkBuckets];
    __shared__ uint64_t s_w[kWalkThreads], s_end[kWalkThreads], s_sid[kWalkThreads];
    __shared__ uint32_t s_L[kWalkThreads], s_flags[kWalkThreads];
    if (tid < kWalkBuckets) s_hist[tid] = 0;
    __syncthreads();
    const uint32_t bucket = kWalkBuckets - 1u - ((mine.flags & 1u) ? walk_cost_key(mine.w) : 0u);   // heaviest first
    const uint32_t pos = atomicAdd(&s_hist[bucket], 1u);
    __syncthreads();
    uint32_t slot = pos;
    for (uint32_t b = 0; b < bucket; b++) slot += s_hist[b];
    s_w[slot] = mine.w; s_end[slot] = mine.seg_end; s_sid[slot] = mine.sid; s_L[slot] = mine.L; s_flags[slot] = mine.flags;
    __syncthreads();
    const WalkItem it{s_w[tid], s_end[tid], s_sid[tid], s_L[tid], s_flags[tid]};
    walk_emit(ix, it, first_no, out, status);
}

}  // namespace ntc
