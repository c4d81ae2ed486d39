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
#include "scan.cuh"

namespace ntc {

constexpr int kPackThreads = 256;
#ifndef NTC_MS_THREADS
#define NTC_MS_THREADS 128          // threads per ms_kernel block; 64 (twice the resident blocks) measured in profiles/r2/ab_ms_threads.txt
#endif
constexpr int kMsThreads = NTC_MS_THREADS;
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

// four ASCII bases in one 32-bit word -> four 2-bit codes in the low byte; *bad |= any byte that is not A, C, G or T
// (first version: byte by byte, ~8 instructions per base -- pack2bit_kernel was instruction bound at 36% of HBM)
NTC_HD uint32_t ascii4_codes(uint32_t w, uint32_t* bad) {
    const uint32_t x = (w >> 1) & 0x03030303u;                       // A->0 C->1 G->3 T->2 in every byte
    const uint32_t c = x ^ ((x >> 1) & 0x01010101u);                 // A0 C1 G2 T3
#if defined(__CUDA_ARCH__)
    const uint32_t sel = (c & 0x3u) | ((c >> 4) & 0x30u) | ((c >> 8) & 0x300u) | ((c >> 12) & 0x3000u);
    *bad |= (__byte_perm(0x54474341u, 0u, sel) != w);                // the letters these codes stand for
#else
    for (int j = 0; j < 4; j++) *bad |= (((0x54474341u >> (8 * ((c >> (8 * j)) & 3u))) & 0xFFu) != ((w >> (8 * j)) & 0xFFu));
#endif
    return (c | (c >> 6) | (c >> 12) | (c >> 18)) & 0xFFu;
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
        for (int i = 0; i < 8; i++) out |= (uint64_t)ascii4_codes(q[i], &bad) << (8 * i);
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
// Length (capped at `limit`) of the run of "unique predecessor" positions ending at q.  A read longer
// than the tile length T is processed as independent tiles of T positions (ms_thread), each counting
// its runs from its own first position; a run that reaches the start of its tile continues with the
// run ending at the last position of the previous tile.
NTC_HD uint32_t unique_run(const uint64_t* sc, uint32_t q, uint32_t T, uint32_t limit) {
    uint32_t run = entry_urun(sc[q]);
    uint32_t total = run;
    while (total < limit) {
        uint32_t p0 = q - q % T;                                      // first position of q's tile
        if (p0 == 0 || run != q - p0 + 1) break;
        q = p0 - 1;
        run = entry_urun(sc[q]);
        total += run;
    }
    return tmin(total, limit);
}

// one record of the greedy cover: the cursor i (exclusive end, > 0) moves to the next record's end.  The step depends on
// nothing but i and the scratch entries, so two walks that ever share a cursor position coincide from there on.
// SPEC (the speculative walks of seg_spec_thread): a step whose left extension exceeds kSpecExtCap is not taken -- returns false
// with i untouched -- so that a walk started in the middle of a very long match costs a bounded number of loads.
constexpr uint32_t kSpecExtCap = 4096;
template <bool SPEC = false, typename Sink>
NTC_HD bool segment_step(const uint64_t* sc, uint32_t& i, uint32_t k, uint32_t T, Sink&& sink) {
    uint64_t ent = sc[i - 1];
    uint32_t d = entry_d(ent), v = entry_s(ent);
    if (d == k && i > k + 1) {                                        // src/lib.rs:180
        const uint32_t max_ext = i - k - 1;
        uint32_t ext = unique_run(sc, i - 2, T, SPEC ? tmin(max_ext, kSpecExtCap) : max_ext);   // left_extend_kmer, src/lib.rs:186
        if (SPEC && ext == kSpecExtCap && max_ext > kSpecExtCap) return false;
        uint32_t new_len = k + ext, m = new_len, old_i = i;
        for (;;) {                                                    // hop loop, src/lib.rs:193-203
            uint32_t dd = entry_d(sc[i - 1]);
            if (dd < m) { m -= dd; i -= dd; } else break;
        }
        uint32_t L = new_len - (m - 1);                               // src/lib.rs:199
        i -= 1;                                                       // src/lib.rs:213-214
        sink(L, v, old_i, k);
    } else {                                                          // src/lib.rs:204-211
        sink(d, v, i, d);
        i -= d;
    }
    return true;
}
template <typename Sink>
NTC_HD uint32_t segment_read(const uint64_t* sc,   // no __restrict__: CountSink parks records in dead entries
                              uint32_t len, uint32_t k, uint32_t T, Sink&& sink) {
    uint32_t i = len, cnt = 0;
    while (i > 0) {
        segment_step(sc, i, k, T, sink);
        cnt++;
    }
    return cnt;
}

// ------------------------------------------------------------------------------------------
// matching statistics + record count.  One thread per read.
//
// Flat loop, one memory round trip per iteration for every thread of the warp: the rank entry of
// the interval start (plus the entry of the interval end when it falls in another block) and,
// speculatively, the LCS words of both interval ends.  If extend_right fails, contract_left is pure
// arithmetic on the words already loaded -- LCS value plus the precomputed distances to the previous /
// next strictly smaller LCS value -- and the same base is probed again in the next iteration.
//   ncu r1 v1: inner retry loop + byte-wise LCS scans = 60% of issued instructions at ~1 thread/warp
//   ncu r1 v2: vectorised scans in a separate state   = 51% of issued instructions at ~2 threads/warp
// ------------------------------------------------------------------------------------------
// Sink of the counting pass.  With `dst` set it also builds the record words (src/encode.rs:144-160) and parks record c
// of the read at dst[-c], dst = the read's last scratch entry: after c+1 records the segmentation has consumed at
// least c+1 bases and never again reads an entry at or right of its cursor, so the slot is dead by then.  emit_kernel<0>
// is then a gather of `count` contiguous words per read instead of a second segmentation over ~12 scattered
// sectors (ncu r1 v12: emit_kernel 0.87 ms, 5.1 GB of DRAM reads, 76 % DRAM busy).
struct CountSink {
    const uint64_t* packed;   // generic pointer: the block's staged reads (shared memory) or the packed stream
    uint64_t a;
    uint64_t* dst;            // null: count only (triples mode segments again in emit_kernel<1>)
    uint64_t first = 1;
    uint32_t big = 0;
    NTC_HD void operator()(uint32_t L, uint32_t v, uint32_t end_pos, uint32_t) {
        big |= (L >= (1u << 24));
        if (!dst) return;
        uint64_t word;
        if (L > kLiteralMax) {                                        // src/encode.rs:146-149
            word = (uint64_t)v | ((uint64_t)(L & 0xFFFFFFu) << 32) | (first << 56);
        } else {                                                      // src/encode.rs:150-156
            const uint64_t p = a + end_pos - L;
            const uint32_t sh = (uint32_t)(p & 31) * 2;
            uint64_t lit = packed[p >> 5] >> sh;
            if (sh && (p & 31) + L > 32) lit |= packed[(p >> 5) + 1] << (64 - sh);
            lit &= (1ull << (2 * L)) - 1ull;                          // L <= 11
            word = lit | (((first + 2) | ((uint64_t)L << 2)) << 56);
        }
        *dst-- = word;
        first = 0;
    }
};

// scratch position of read r: its entries start on a 32-byte sector so that groups of four 8-byte
// entries are written as one full sector (no read-for-fill in L2/DRAM)
NTC_HD uint64_t scratch_base(uint64_t a, uint64_t r) { return (a + 4 * r) & ~3ull; }
inline uint64_t scratch_entries(uint64_t n_bases, uint64_t n_reads) { return n_bases + 4 * n_reads + 8; }

// LCS word i: bits 0-7 LCS[i], 8-15 min(255, i - PSV(i)), 16-23 LCS[i+1], 24-31 min(255, NSV(i+1) - (i+1)),
// PSV / NSV = nearest position left / right whose LCS is strictly smaller.  A singleton interval
// [s, s+1) -- the common case -- finds everything contract_left needs in the single word s; a wider
// interval [s, e) reads words s and e-1.
NTC_HD uint32_t lcsx_lcs(uint32_t x) { return x & 0xFFu; }

// ---- tiles ------------------------------------------------------------------------------------
// MS at position j depends only on the k-1 bases before it, so a read longer
// than the tile length T is cut into tiles of T positions; the thread of tile t > 0 starts k-1 bases
// early from the empty match (warm-up, nothing stored) and is exact from the tile's first position on.
// A 10 kbp read (BASELINE config C5) is then ~20 independent chains instead of one.  T is a multiple
// of 4 (full-sector scratch stores) and >= kTileMin >= k_max - 1.
constexpr uint32_t kTileMin = 256;
constexpr uint32_t kDescTileBits = 24;   // item descriptor: read << 24 | tile
NTC_HD uint32_t tiles_of(uint32_t len, uint32_t T) { return len <= T ? 1u : (len + T - 1u) / T; }

// scan input: tiles per read
struct TileLoad {
    const uint64_t* offsets;
    uint32_t T;
    NTC_HD Pair64 operator()(uint64_t r) const { return Pair64{tiles_of((uint32_t)(offsets[r + 1] - offsets[r]), T), 0}; }
};

// work-item descriptors, written by the tile scan's last pass (scan_down_kernel's post hook): read r's tiles are items
// [prefix.a, prefix.a + tiles)
struct TileDescPost {
    uint64_t* desc;
    NTC_HD void operator()(uint64_t r, const Pair64& prefix, const Pair64& v) const {
        for (uint64_t t = 0; t < v.a; t++) desc[prefix.a + t] = (r << kDescTileBits) | t;
    }
};

// `stage` = 4 x blockDim u64 staging slots of this thread's block (shared memory on the GPU),
// indexed [slot * stride + lane]
template <int NODES>
NTC_HD void ms_thread(const DevIndex& ix, const uint64_t* packed, const uint64_t* __restrict__ offsets, uint64_t item_desc, uint32_t T,
                      uint64_t* __restrict__ scratch, uint32_t* __restrict__ counts, uint32_t* __restrict__ status,
                      uint64_t* stage, uint32_t stride, bool keep_records) {
    const uint64_t r = item_desc >> kDescTileBits;
    const uint32_t tile = (uint32_t)item_desc & ((1u << kDescTileBits) - 1u);
    const uint64_t base0 = NTC_LDG(offsets);
    const uint64_t a = offsets[r] - base0;
    const uint64_t len64 = offsets[r + 1] - offsets[r];
    const uint32_t len = (uint32_t)len64;
    if (len == 0 || (len64 >> 32)) {                                  // EncodeError, src/lib.rs:224; or unsupported length
        if (counts) counts[r] = 0;
        NTC_ATOMIC_OR(status, (uint32_t)(len64 >> 32 ? ST_READ_TOO_LONG : ST_EMPTY_READ));
        return;
    }
    const uint32_t n = ix.n, k = ix.k;
    const bool multi = len > T;                                       // counted by count_kernel once every tile is done
    const uint32_t p0 = tile * T;                                     // first stored position
    const uint32_t p1 = multi ? tmin(len, p0 + T) : len;              // end of the tile
    const uint64_t keep = ix.pol_keep;
    uint64_t* __restrict__ sc = scratch + scratch_base(a, r);
    uint32_t s = 0, e = n, d = 0, urun = 0;
    uint32_t j = tile ? p0 - (k - 1u) : 0u;                           // warm-up start
    bool fresh = true, force = false;
    uint32_t dmin = 1;                                                // 0 once a base was absent from the index
    // the read's bases leave `word` two bits at a time, 16 per 32-bit word (ncu r1 v10: 10% of the loop's instructions
    // were 64-bit position arithmetic and variable shifts in the base fetch; v11 used a 64-bit shift register)
    const uint32_t* __restrict__ wp = (const uint32_t*)packed + ((a + j) >> 4);
    uint32_t left = 16u - (uint32_t)((a + j) & 15);
    uint32_t word = *wp >> (((uint32_t)(a + j) & 15u) * 2u);
    uint32_t c = word & 3u;
    // the first 8 bases of a read (or of a tile's warm-up) start from the empty match: their 8 iterations -- 16 rank
    // probes in different blocks, the interval is wide -- are one 32-byte prefix-table entry (host_index.h)
    if (ix.pfx && p1 - j >= 8u) {
        const uint32_t pre = (left >= 8u ? word : (word | (wp[1] << (2u * left)))) & 0xFFFFu;
        const uint4 lo = ld16_keep(ix.pfx + 2u * pre, keep), hi = ld16_keep(ix.pfx + 2u * pre + 1u, keep);
        if (hi.w > hi.z) {                                            // every prefix of these 8 bases is in the index
            if (tile == 0) {                                          // entries 0..3: one sector; 4..6 wait for entry 7 in the stage
                const uint32_t c0 = pre & 3u;   // select chain: indexing the by-value parameter struct would copy it to local memory
                const uint32_t s1 = c0 == 0u ? ix.c_start[0] : c0 == 1u ? ix.c_start[1] : c0 == 2u ? ix.c_start[2] : ix.c_start[3];
                st32_stream(sc, make_entry(s1, 1, 0), make_entry(lo.x, 2, 0), make_entry(lo.y, 3, 0),
                            make_entry(lo.z, 4, 0), ix.pol_stream);
                stage[0] = make_entry(lo.w, 5, 0);
                stage[stride] = make_entry(hi.x, 6, 0);
                stage[2 * stride] = make_entry(hi.y, 7, 0);
            }
            s = hi.z; e = hi.w; d = 8; j += 8;
            if (left > 8u) { word >>= 16; left -= 8u; }
            else {
                const uint32_t used = 8u - left;                      // bases taken from the next word
                word = j < p1 ? *++wp >> (2u * used) : 0u;
                left = 16u - used;
            }
            c = word & 3u;
        }
    }
    while (j < p1) {
        // a singleton interval [s, s+1) -- the common case once the match is longer than ~log4(n) -- needs one
        // entry only, also when s is the last node of its block: rank(s+1) = rank(s) + row bit of s
        const bool single = e == s + 1u;
        typedef RankL<NODES> RL;
        const uint32_t bs = RL::block(s), os = RL::offset(s);
        const uint32_t be = RL::block(e), oe = RL::offset(e);
        const bool other = !single && be != bs;                       // the interval end lies in another block
        // LCS words of the interval ends: loaded speculatively only while d is in the window where
        // extensions of chance matches fail (around log4(n_sets), set by the host), on demand otherwise
        // (ncu r1 v3: unconditional speculation doubled the sector traffic; v5: d < 20 still spent 39% of
        // all load sectors on it)
        const bool spec = force || (d - ix.spec_lo < ix.spec_span);   // unsigned: spec_lo <= d < spec_lo + spec_span
        // all loads of the iteration issue back to back, before any of their results is used
        const typename RL::Ent va = RL::ld_ent(ix.rank, bs, c, keep);
        const typename RL::Part vb = RL::ld_part_if(ix.rank, be, c, keep, other);
        const uint32_t xs = ld4_keep_if(ix.lcsx + s, keep, spec);
        uint32_t xe = ld4_keep_if(ix.lcsx + (e - 1u), keep, spec && !single);
        xe = single ? xs : xe;
        const uint32_t s2 = RL::rank(va, os);
        const uint32_t e2 = single ? s2 + RL::bit(va, os) : RL::rank_sel(va, vb, other, oe);
        if (fresh) {
            // the block of the interval start (unchanged since the last successful step) also holds its Q bit:
            // finish entry j-1, whose (start, d) are still in s and d
            if (j > p0) {
                urun = (d == k && RL::qbit(va, os)) ? tmin(urun + 1u, 0xFFFFFFu) : 0u;
                const uint32_t slot = (j - 1u) & 3u;
                stage[slot * stride] = make_entry(s, d, urun);
                if (slot == 3u) {                                     // one full 32-byte sector
                    st32_stream(sc + (j - 4u), stage[0], stage[stride], stage[2 * stride], stage[3 * stride], ix.pol_stream);
                }
            }
            fresh = false;
        }
        if (s2 >= e2 && !(s == 0 && e == n)) {
            // contract_left (sbwt LcsArray): the interval is maximal for its level d, so both boundary
            // LCS values are < d and the reference's loop -- which re-probes the same empty interval for
            // every target above max(LCS[s], LCS[e]) -- ends up at that level with the interval widened
            // to the nearest strictly smaller LCS on the side(s) that hold the maximum.
            // no LCS words at hand (failure outside the speculation window, e.g. the first mismatch of
            // a long match): do NOT load them here -- a dependent load on this divergent path stalls the
            // whole warp for a memory round trip (ncu r1 v6) -- re-run the iteration with the load forced
            if (!spec) { force = true; continue; }   // (a B200 sweep found on-demand loads here no faster for any window)
            force = false;
            const int ls = s > 0 ? (int)lcsx_lcs(xs) : -1;              // LCS[s]   (low half of word s)
            const int le = e < n ? (int)((xe >> 16) & 0xFFu) : -1;      // LCS[e]   (high half of word e-1)
            const int t = tmax(ls, le);
            if (t <= 0) { s = 0; e = n; d = 0; continue; }             // LCS >= 0 everywhere
            if (ls == t) {
                uint32_t dl = (xs >> 8) & NTC_LCS_DIST_SAT;
                s -= dl;
                if (dl == NTC_LCS_DIST_SAT) while (s > 0 && lcsx_lcs(NTC_LDG(ix.lcsx + s)) >= (uint32_t)t) s--;
            }
            if (le == t) {
                uint32_t dr = xe >> 24;
                e += dr;
                if (dr == NTC_LCS_DIST_SAT) while (e < n && lcsx_lcs(NTC_LDG(ix.lcsx + e)) >= (uint32_t)t) e++;
            }
            d = (uint32_t)t;
            continue;
        }
        const bool ok = s2 < e2;                                      // extend_right succeeded; else the base is absent from the index
        s = ok ? s2 : s;
        e = ok ? e2 : e;
        d = ok ? tmin(d + 1u, k) : 0u;
        dmin = tmin(dmin, d);
        fresh = true;
        j++;
        word >>= 2;
        if (--left == 0) { word = j < p1 ? *++wp : 0; left = 16; }
        c = word & 3u;
    }
    // tail: entries of the last, partial group of four + the tile's final position (whose start and d are in s, d)
    {
        uint32_t first_unflushed = (p1 - 1) & ~3u;                     // entries [first_unflushed, p1-1) are staged (>= p0: p0 % 4 == 0)
        for (uint32_t q = first_unflushed; q + 1 < p1; q++) sc[q] = stage[(q & 3u) * stride];
        uint32_t last_run = 0;                                         // urun of a read's last position is never read (ext looks at i-2)
        if (p1 < len) {                                                // ... but the next tile's runs may continue this one
            const typename RankL<NODES>::Ent vq = RankL<NODES>::ld_ent(ix.rank, RankL<NODES>::block(s), 0u, keep);
            last_run = (d == k && RankL<NODES>::qbit(vq, RankL<NODES>::offset(s))) ? tmin(urun + 1u, 0xFFFFFFu) : 0u;
        }
        sc[p1 - 1] = make_entry(s, d, last_run);
    }
    if (dmin == 0) {                                                  // lib.rs:207 would underflow (Q3)
        if (counts && !multi) counts[r] = 0;
        NTC_ATOMIC_OR(status, (uint32_t)ST_BAD_BASE);
        return;
    }
    if (counts && !multi) {
        CountSink sink{packed, a, keep_records ? sc + (len - 1u) : nullptr};
        counts[r] = segment_read(sc, len, k, T, sink);
        if (sink.big) NTC_ATOMIC_OR(status, (uint32_t)ST_LEN24);      // src/lib.rs:226
    }
}

// record count of a read that was processed as several tiles (single-tile reads are counted by ms_thread)
NTC_HD void count_thread(const DevIndex& ix, const uint64_t* __restrict__ packed, const uint64_t* __restrict__ offsets, uint64_t n_reads,
                         uint32_t T, uint64_t* __restrict__ scratch, uint32_t* __restrict__ counts, uint32_t* __restrict__ status,
                         bool keep_records, uint64_t r) {
    if (r >= n_reads) return;
    const uint32_t len = (uint32_t)(offsets[r + 1] - offsets[r]);
    if (len <= T) return;
    if (*status & (ST_BAD_BASE | ST_READ_TOO_LONG)) { counts[r] = 0; return; }   // some d == 0: segmentation would not terminate (Q3)
    const uint64_t a = offsets[r] - NTC_LDG(offsets);
    uint64_t* sc = scratch + scratch_base(a, r);
    CountSink sink{packed, a, keep_records ? sc + (len - 1u) : nullptr};
    counts[r] = segment_read(sc, len, ix.k, T, sink);
    if (sink.big) NTC_ATOMIC_OR(status, (uint32_t)ST_LEN24);
}

// ---- long reads: the segmentation cut into segments ---------------------------------------------------------------
// count_thread walks a 10 kbp read as ONE chain of ~600 dependent, scattered scratch loads (C5: 5.4 ms for 250 k reads, an eighth
// of the step).  The cover is a deterministic walk over cursor positions (segment_step), and walks that meet stay together -- they
// meet at the first mismatch below both, where every match ends -- so the walk is cut into segments of S cursor positions (512 .. 16,384):
//   A  seg_spec_thread, one thread per segment: a SPECULATIVE walk from the segment's top cursor to its bottom; remembers the
//      cursors it visited (one bit each), how many records that was and where it left the segment;
//   B  seg_join_thread, one thread per read: the TRUE walk, top down -- inside a segment it steps until it stands on a cursor the
//      speculative walk visited (usually within a record or two), then takes that walk's count and exit; notes where the true
//      walk entered each segment and how many records came before;
//   C  seg_emit_thread, one thread per segment, after the record offsets are scanned: walks again from the true entry and
//      writes the record words straight to their final place (these reads park nothing and emit_kernel<0> skips them).
// Same records as count_thread + emit_kernel<0>, in 20 x as many independent chains.
// Segment length S: a multiple of 64 (one bitmap word per 64 cursors), chosen per batch (seg_len_for): 512 for reads up to a few
// 10 kbp; longer for contig-sized sequences, where the per-read join (one step per segment, ~1.2 us each on B200) would
// otherwise dominate -- 5 Mbp in one read: 12 ms of join at S = 512.
constexpr uint32_t kSegLenMin = 512, kSegLenMax = 16384;
inline uint32_t seg_len_for(uint64_t n_bases, uint64_t n_reads) {
    const uint64_t mean = n_reads ? n_bases / n_reads : 0;
    uint32_t S = kSegLenMin;
    while (S < kSegLenMax && (uint64_t)S * S < mean * 32) S *= 2;    // join ~ len / S steps per read, walk ~ S / 32 steps per segment
    return S;
}
struct SegLoad {             // scan input: segments of a read (only reads longer than the tile length take this path)
    const uint64_t* offsets;
    uint32_t T, S;
    NTC_HD Pair64 operator()(uint64_t r) const {
        const uint64_t len = offsets[r + 1] - offsets[r];
        return Pair64{len > T && !(len >> 32) ? (len + S - 1) / S : 0, 0};
    }
};
struct NullSink {
    NTC_HD void operator()(uint32_t, uint32_t, uint32_t, uint32_t) const {}
};
// seg_info: 4 x u32 per segment = records of the speculative walk, its exit cursor (below the segment, or the cursor inside it at
// which the walk gave up: not marked, the true walk steps on from there), the true walk's entry cursor (0: the true walk jumps
// over this segment), records of the read before that entry
NTC_HD void seg_spec_thread(const DevIndex& ix, const uint64_t* __restrict__ offsets, uint32_t T, uint32_t S, const uint64_t* __restrict__ scratch,
                            const uint64_t* __restrict__ desc, uint64_t* __restrict__ seg_bits, uint32_t* __restrict__ seg_info,
                            const uint32_t* __restrict__ status, uint64_t gi) {
    const uint64_t r = desc[gi] >> kDescTileBits;
    const uint32_t g = (uint32_t)desc[gi] & ((1u << kDescTileBits) - 1u);
    const uint64_t a = offsets[r] - NTC_LDG(offsets);
    const uint32_t len = (uint32_t)(offsets[r + 1] - offsets[r]);
    const uint64_t* sc = scratch + scratch_base(a, r);
    const uint32_t lo = g * S, top = tmin(len, lo + S);
    uint64_t* bits = seg_bits + gi * (S / 64);
    uint32_t cnt = 0, i = top;
    if (*status & (ST_BAD_BASE | ST_READ_TOO_LONG)) i = lo;           // some d == 0: the walk would not move (Q3)
    int w = (int)(S / 64) - 1;                                        // the walk descends: one bitmap word at a time
    uint64_t cur = 0;
    while (i > lo) {
        const uint32_t b = i - lo - 1;
        while ((int)(b >> 6) < w) { bits[w] = cur; cur = 0; w--; }
        if (!segment_step<true>(sc, i, ix.k, T, NullSink())) break;   // too long a step to take on speculation: the true walk goes on from here
        cur |= 1ull << (b & 63);
        cnt++;
    }
    for (; w >= 0; w--) { bits[w] = cur; cur = 0; }
    seg_info[4 * gi + 0] = cnt;
    seg_info[4 * gi + 1] = i;
    seg_info[4 * gi + 2] = 0;
    seg_info[4 * gi + 3] = 0;
}
NTC_HD void seg_join_thread(const DevIndex& ix, const uint64_t* __restrict__ offsets, uint64_t n_reads, uint32_t T, uint32_t S,
                            const uint64_t* __restrict__ scratch, const uint64_t* __restrict__ seg_off, const uint64_t* __restrict__ seg_bits,
                            uint32_t* __restrict__ seg_info, uint32_t* __restrict__ counts, uint32_t* __restrict__ status, uint64_t r) {
    if (r >= n_reads) return;
    const uint64_t len64 = offsets[r + 1] - offsets[r];
    const uint32_t len = (uint32_t)len64;
    if (len64 <= T || (len64 >> 32)) return;
    if (*status & (ST_BAD_BASE | ST_READ_TOO_LONG)) { counts[r] = 0; return; }
    const uint64_t a = offsets[r] - NTC_LDG(offsets);
    const uint64_t* sc = scratch + scratch_base(a, r);
    const uint64_t s0 = seg_off[r];
    uint32_t i = len, cnt = 0;
    while (i > 0) {
        const uint32_t g = (i - 1) / S, lo = g * S;
        const uint64_t gi = s0 + g;
        const uint64_t* bits = seg_bits + gi * (S / 64);
        seg_info[4 * gi + 2] = i;
        seg_info[4 * gi + 3] = cnt;
        while (i > lo) {
            const uint32_t b = i - lo - 1;
            if ((bits[b >> 6] >> (b & 63)) & 1ull) {                  // the speculative walk stood here: same walk from now on
                // records of that walk from here on = all of them minus the cursors it visited above this one; the walks meet
                // near the top of the segment, so that is a word or two (counting the words below cost 17 us per segment at S = 16,384)
                uint32_t above = (b & 63) == 63 ? 0u : (uint32_t)NTC_POPCLL(bits[b >> 6] & (~0ull << ((b & 63) + 1)));
                for (uint32_t w = (b >> 6) + 1; w < S / 64; w++) above += (uint32_t)NTC_POPCLL(bits[w]);
                cnt += seg_info[4 * gi + 0] - above;
                i = seg_info[4 * gi + 1];                             // below the segment, or the cursor at which that walk gave up
                continue;
            }
            segment_step(sc, i, ix.k, T, NullSink());
            cnt++;
        }
    }
    counts[r] = cnt;
}
struct SegEmitSink {
    const uint64_t* __restrict__ packed;
    uint64_t* __restrict__ out;
    uint64_t a, first;
    uint32_t big = 0;
    NTC_HD void operator()(uint32_t L, uint32_t v, uint32_t end_pos, uint32_t) {
        big |= (L >= (1u << 24));
        uint64_t word;
        if (L > kLiteralMax) word = (uint64_t)v | ((uint64_t)(L & 0xFFFFFFu) << 32) | (first << 56);       // src/encode.rs:146-149
        else word = packed_slice(packed, a + end_pos - L, L) | (((first + 2) | ((uint64_t)L << 2)) << 56);  // src/encode.rs:150-156
        *out++ = word;
        first = 0;
    }
};
NTC_HD void seg_emit_thread(const DevIndex& ix, const uint64_t* __restrict__ packed, const uint64_t* __restrict__ offsets, uint32_t T, uint32_t S,
                            const uint64_t* __restrict__ scratch, const uint64_t* __restrict__ desc, const uint32_t* __restrict__ seg_info,
                            const uint32_t* __restrict__ counts, const uint64_t* __restrict__ rec_offsets, const uint64_t* __restrict__ rel,
                            uint64_t* __restrict__ out, uint64_t cap, uint32_t* __restrict__ status, uint64_t gi) {
    uint32_t i = seg_info[4 * gi + 2];
    if (i == 0) return;                                               // the true walk never stood in this segment
    const uint64_t r = desc[gi] >> kDescTileBits;
    const uint32_t g = (uint32_t)desc[gi] & ((1u << kDescTileBits) - 1u);
    if (counts[r] == 0) return;
    const uint64_t rel0 = rel ? NTC_LDG(rel) : 0;
    if (rec_offsets[r + 1] - rel0 > cap) { NTC_ATOMIC_OR(status, (uint32_t)ST_CAPACITY); return; }
    const uint64_t a = offsets[r] - NTC_LDG(offsets);
    const uint64_t* sc = scratch + scratch_base(a, r);
    const uint32_t before = seg_info[4 * gi + 3], lo = g * S;
    SegEmitSink sink{packed, out + (rec_offsets[r] - rel0) + before, a, before == 0 ? 1ull : 0ull};
    while (i > lo) segment_step(sc, i, ix.k, T, sink);
    if (sink.big) NTC_ATOMIC_OR(status, (uint32_t)ST_LEN24);
}

#if defined(__CUDACC__)
// 1-D bulk TMA (cp.async.bulk) of `bytes` (multiple of 16) from global to shared memory, completion
// on an mbarrier; issued by one thread.
__device__ __forceinline__ void tma_load_1d(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
    uint32_t dst = (uint32_t)__cvta_generic_to_shared(smem_dst);
    uint32_t b = (uint32_t)__cvta_generic_to_shared(bar);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(gmem_src),
                 "r"(bytes), "r"(b)
                 : "memory");
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    uint32_t b = (uint32_t)__cvta_generic_to_shared(bar);
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(b), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t b = (uint32_t)__cvta_generic_to_shared(bar);
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(b),
        "r"(parity)
        : "memory");
}
#endif

constexpr uint32_t kMsStageWordsMax = 1024;  // at most 8 KB of packed reads per block (128 reads x up to 256 bases); longer reads load from global
// The stage buffer is dynamic shared memory sized for the batch's mean read length (api.cu): 4.75 KB for 150 bp reads.  With
// 9.8 KB per block twelve blocks fit one SM beside a 124 KB L1 -- the kernel needs ~100 KB of L1 (profiles/r2/ab_ms_scratch_on_chip.txt)
// -- and twelve blocks of 40 registers measure 11.49 against 11.69 ms for ten of 48 (profiles/r2/ab_ms_blocks.txt).  Batches of longer
// reads take the ten-block instantiation: twelve blocks with an 8 KB stage each would push the L1 below that size (12.55 ms, r1).
template <int MIN_BLOCKS, int NODES>
__global__ void __launch_bounds__(kMsThreads, MIN_BLOCKS * (128 / kMsThreads)) ms_kernel(DevIndex ix, const uint64_t* __restrict__ packed,
                                                        const uint64_t* __restrict__ offsets, const uint64_t* __restrict__ desc,
                                                        const uint64_t* __restrict__ n_items, uint32_t T,
                                                        uint64_t* __restrict__ scratch, uint32_t* __restrict__ counts,
                                                        uint32_t* __restrict__ status, int keep_records, uint32_t stage_words) {
    extern __shared__ __align__(16) uint64_t s_packed[];               // stage_words words
    __shared__ uint64_t s_stage[4 * kMsThreads];
    __shared__ __align__(8) uint64_t s_bar;
    // the grid covers an upper bound of the item count (n_reads + n_bases / T); the exact count is on the device
    const uint64_t total = *n_items;
    const uint64_t i0 = (uint64_t)blockIdx.x * kMsThreads;
    if (i0 >= total) return;
    const uint64_t i1 = i0 + kMsThreads < total ? i0 + kMsThreads : total;
    // the block's items are one contiguous run of the packed stream (tiles of one read overlap by their
    // warm-up): stage it with one bulk TMA copy when it fits (always for short reads)
    const uint64_t base0 = offsets[0];
    const uint64_t da = desc[i0], db = desc[i1 - 1];
    const uint64_t ra = da >> kDescTileBits, rb = db >> kDescTileBits;
    const uint32_t ta = (uint32_t)da & ((1u << kDescTileBits) - 1u), tb = (uint32_t)db & ((1u << kDescTileBits) - 1u);
    const uint64_t len_b = offsets[rb + 1] - offsets[rb];
    const uint64_t lo = offsets[ra] - base0 + (ta ? (uint64_t)ta * T - (ix.k - 1u) : 0);
    const uint64_t hi = offsets[rb] - base0 + (len_b > T ? tmin<uint64_t>(len_b, ((uint64_t)tb + 1) * T) : len_b);
    const uint64_t w_lo = (lo >> 5) & ~1ull;                                 // 16-byte aligned
    const uint64_t w_hi = ((hi >> 5) + 2) & ~1ull;                           // exclusive, 16-byte aligned (buffer is padded)
    const bool staged = (w_hi - w_lo) <= stage_words;
    if (staged) {
        if (threadIdx.x == 0) mbar_init(&s_bar, 1);
        __syncthreads();
        if (threadIdx.x == 0) tma_load_1d(s_packed, packed + w_lo, (uint32_t)((w_hi - w_lo) * 8), &s_bar);
        mbar_wait(&s_bar, 0);
    }
    const uint64_t* pk = staged ? (s_packed - w_lo) : packed;                // generic pointer: shared or global
    const uint64_t i = i0 + threadIdx.x;
    if (i >= i1) return;
    ms_thread<NODES>(ix, pk, offsets, desc[i], T, scratch, counts, status, s_stage + threadIdx.x, kMsThreads, keep_records != 0);
}

__global__ void __launch_bounds__(256) count_kernel(DevIndex ix, const uint64_t* __restrict__ packed, const uint64_t* __restrict__ offsets,
                                                    uint64_t n_reads, uint32_t T, uint64_t* __restrict__ scratch,
                                                    uint32_t* __restrict__ counts, uint32_t* __restrict__ status, int keep_records) {
    count_thread(ix, packed, offsets, n_reads, T, scratch, counts, status, keep_records != 0, (uint64_t)blockIdx.x * 256 + threadIdx.x);
}

__global__ void __launch_bounds__(256) seg_spec_kernel(DevIndex ix, const uint64_t* __restrict__ offsets, uint32_t T, uint32_t S,
                                                       const uint64_t* __restrict__ scratch, const uint64_t* __restrict__ desc,
                                                       const uint64_t* __restrict__ n_segs, uint64_t* __restrict__ seg_bits,
                                                       uint32_t* __restrict__ seg_info, const uint32_t* __restrict__ status) {
    const uint64_t gi = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (gi < *n_segs) seg_spec_thread(ix, offsets, T, S, scratch, desc, seg_bits, seg_info, status, gi);
}
__global__ void __launch_bounds__(256) seg_join_kernel(DevIndex ix, const uint64_t* __restrict__ offsets, uint64_t n_reads, uint32_t T, uint32_t S,
                                                       const uint64_t* __restrict__ scratch, const uint64_t* __restrict__ seg_off,
                                                       const uint64_t* __restrict__ seg_bits, uint32_t* __restrict__ seg_info,
                                                       uint32_t* __restrict__ counts, uint32_t* __restrict__ status) {
    seg_join_thread(ix, offsets, n_reads, T, S, scratch, seg_off, seg_bits, seg_info, counts, status, (uint64_t)blockIdx.x * 256 + threadIdx.x);
}
__global__ void __launch_bounds__(256) seg_emit_kernel(DevIndex ix, const uint64_t* __restrict__ packed, const uint64_t* __restrict__ offsets,
                                                       uint32_t T, uint32_t S, const uint64_t* __restrict__ scratch, const uint64_t* __restrict__ desc,
                                                       const uint64_t* __restrict__ n_segs, const uint32_t* __restrict__ seg_info,
                                                       const uint32_t* __restrict__ counts, const uint64_t* __restrict__ rec_offsets,
                                                       const uint64_t* __restrict__ rel, uint64_t* __restrict__ out, uint64_t cap,
                                                       uint32_t* __restrict__ status) {
    const uint64_t gi = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (gi < *n_segs) seg_emit_thread(ix, packed, offsets, T, S, scratch, desc, seg_info, counts, rec_offsets, rel, out, cap, status, gi);
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
            while (end < ix.n && lcsx_lcs(NTC_LDG(ix.lcsx + end)) >= d_orig) end++;
            out[3 * o] = L; out[3 * o + 1] = v; out[3 * o + 2] = end;
        }
        o++;
        first = 0;
    }
};

template <int MODE>
NTC_HD void emit_thread(const DevIndex& ix, const uint64_t* __restrict__ packed, const uint64_t* __restrict__ offsets, uint64_t n_reads,
                        const uint64_t* __restrict__ scratch, const uint32_t* __restrict__ counts, const uint64_t* __restrict__ rec_offsets,
                        const uint64_t* __restrict__ rel, uint64_t* __restrict__ out, uint64_t cap, uint32_t* __restrict__ status, uint32_t T,
                        uint64_t r, bool segs = false) {
    // rec_offsets are absolute (a carry from earlier chunks is folded in); `out` is this chunk's
    // buffer, so positions are taken relative to *rel (the number of records before the chunk)
    if (r >= n_reads) return;
    if (counts[r] == 0) return;
    const uint64_t rel0 = rel ? NTC_LDG(rel) : 0;
    const uint64_t base0 = NTC_LDG(offsets);
    const uint64_t a = offsets[r] - base0;
    const uint32_t len = (uint32_t)(offsets[r + 1] - offsets[r]);
    if (MODE == 0 && segs && len > T) return;                         // written by seg_emit_kernel, nothing parked
    if (rec_offsets[r + 1] - rel0 > cap) { NTC_ATOMIC_OR(status, (uint32_t)ST_CAPACITY); return; }
    const uint64_t* sc = scratch + scratch_base(a, r);
    if (MODE == 0) {
        // the counting pass left record c at entry len-1-c (CountSink): copy them out in order
        const uint32_t cnt = counts[r];
        const uint64_t* src = sc + (len - 1u);
        uint64_t* dst = out + (rec_offsets[r] - rel0);
        for (uint32_t c = 0; c < cnt; c++) dst[c] = src[-(int64_t)c];
    } else {
        EmitSink<MODE> sink{ix, packed, out, a, rec_offsets[r] - rel0, 1};
        segment_read(sc, len, ix.k, T, sink);
    }
}

template <int MODE>
__global__ void __launch_bounds__(kMsThreads) emit_kernel(DevIndex ix, const uint64_t* __restrict__ packed,
                                                          const uint64_t* __restrict__ offsets, uint64_t n_reads,
                                                          const uint64_t* __restrict__ scratch, const uint32_t* __restrict__ counts,
                                                          const uint64_t* __restrict__ rec_offsets, const uint64_t* __restrict__ rel,
                                                          uint64_t* __restrict__ out, uint64_t cap, uint32_t* __restrict__ status, uint32_t T,
                                                          int segs) {
    emit_thread<MODE>(ix, packed, offsets, n_reads, scratch, counts, rec_offsets, rel, out, cap, status, T,
                      (uint64_t)blockIdx.x * kMsThreads + threadIdx.x, segs != 0);
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

// per-base (d, start) dump for the matching-statistics parity entry point (diagnostic, one thread per read)
__global__ void __launch_bounds__(256) ms_dump_kernel(const uint64_t* __restrict__ scratch, const uint64_t* __restrict__ offsets,
                                                      uint64_t n_reads, uint8_t* __restrict__ d_out, uint32_t* __restrict__ s_out) {
    uint64_t r = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (r >= n_reads) return;
    const uint64_t a = offsets[r] - offsets[0];
    const uint32_t len = (uint32_t)(offsets[r + 1] - offsets[r]);
    const uint64_t* sc = scratch + scratch_base(a, r);
    for (uint32_t j = 0; j < len; j++) {
        d_out[a + j] = (uint8_t)entry_d(sc[j]);
        s_out[a + j] = entry_s(sc[j]);
    }
}

}  // namespace ntc
