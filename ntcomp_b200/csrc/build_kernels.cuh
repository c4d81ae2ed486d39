// build_kernels.cuh -- GPU SBWT + LCS construction (`ntcomp build`, src/main.rs:97-140, i.e. kbo::build
// with add_revcomp as called at src/main.rs:111-134).  Per-element bodies are NTC_HD functors so that tests/emul runs the
// very same code on the CPU; the device-only pieces are the for_each launcher and the radix sort at the bottom.
//
// Representation.  Both strands and every input sequence live in ONE 2-bit text X (32 characters per u64, character q at
// bits 2(q % 32) of word q / 32):  [kPad invalid] [forward: each sequence followed by one separator] [reverse complement
// of the forward part] [invalid tail].  A string of the construction -- a k-mer, a (k-1)-mer or a '$'-padded dummy node --
// is never materialised: it is a pair (vs, lo), "kk characters starting at text position vs, every position below lo
// reads as A".  A k-mer that starts at p is (p, p); the dummy node $^(k-l) x[0..l) of a k-mer x at p is (p - (k - l), p).
// With '$' read as 'A' and ties broken by length (more '$' first) the integer order of these strings, last character
// most significant, is the SBWT's colex order ($ < A < C < G < T): index_build.cpp uses the same trick on the host.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "device_common.cuh"
#include "scan.cuh"

#if defined(__CUDA_ARCH__)
#define NTC_CLZLL(x) __clzll((long long)(x))
#define NTC_ATOMIC_ADD32(p, v) atomicAdd((p), (v))
#define NTC_ATOMIC_OR64(p, v) atomicOr((unsigned long long*)(p), (unsigned long long)(v))
#else
#define NTC_CLZLL(x) __builtin_clzll(x)
#define NTC_ATOMIC_ADD32(p, v) (*(p) += (v))
#define NTC_ATOMIC_OR64(p, v) (*(p) |= (v))
#endif

namespace ntc {
namespace bld {

constexpr uint32_t kPad = 256;           // >= kMaxK: a dummy's virtual start p - (k - l) never underflows
constexpr uint32_t kTail = 320;          // invalid characters after the last strand: every window read stays inside
constexpr uint32_t kNone = 0xFFFFFFFFu;

// characters [c0, c0 + nch) (nch <= 32) of the string (vs, lo): character c0 + j at bits 2j
NTC_HD uint64_t str_chars(const uint64_t* __restrict__ packed, uint32_t vs, uint32_t lo, uint32_t c0, uint32_t nch) {
    const uint64_t q = (uint64_t)vs + c0;
    uint64_t v = packed_slice(packed, q, nch);
    if ((uint64_t)lo > q) {
        const uint64_t z = (uint64_t)lo - q;
        v = z >= nch ? 0ull : (v >> (2 * z)) << (2 * z);
    }
    return v;
}
// number of equal characters of two kk-character strings, counted from the LAST character downwards
NTC_HD uint32_t common_suffix(const uint64_t* __restrict__ packed, uint32_t kk, uint32_t avs, uint32_t alo, uint32_t bvs, uint32_t blo) {
    for (int w = (int)((kk + 31) / 32) - 1; w >= 0; w--) {
        const uint32_t c0 = 32u * (uint32_t)w, nch = tmin(32u, kk - c0);
        const uint64_t x = str_chars(packed, avs, alo, c0, nch) ^ str_chars(packed, bvs, blo, c0, nch);
        if (x) {
            const uint32_t top_bit = 63u - (uint32_t)NTC_CLZLL(x);
            return kk - 1 - (c0 + top_bit / 2);
        }
    }
    return kk;
}
// colex comparison: -1 / 0 / +1
NTC_HD int cmp_colex(const uint64_t* __restrict__ packed, uint32_t kk, uint32_t avs, uint32_t alo, uint32_t bvs, uint32_t blo) {
    for (int w = (int)((kk + 31) / 32) - 1; w >= 0; w--) {
        const uint32_t c0 = 32u * (uint32_t)w, nch = tmin(32u, kk - c0);
        const uint64_t a = str_chars(packed, avs, alo, c0, nch), b = str_chars(packed, bvs, blo, c0, nch);
        if (a != b) return a < b ? -1 : 1;
    }
    return 0;
}
// radix digit t of a k-character string: characters [4t, 4t + 4)
NTC_HD uint32_t str_digit(const uint64_t* __restrict__ packed, uint32_t k, uint32_t vs, uint32_t lo, uint32_t t) {
    const uint32_t c0 = 4u * t;
    return (uint32_t)str_chars(packed, vs, lo, c0, tmin(4u, k - c0));
}
NTC_HD int base_code(uint8_t b) { return b == 'A' ? 0 : b == 'C' ? 1 : b == 'G' ? 2 : b == 'T' ? 3 : -1; }

// ---- 1. text -------------------------------------------------------------------------------------------
// one element = 32 characters of X: the packed word and a mask of invalid characters (separator, non-ACGT, padding)
struct PackText {
    const uint8_t* fwd;   // forward strand, ASCII: every sequence followed by one separator byte
    uint64_t len;         // bytes of fwd
    uint32_t revcomp;
    uint64_t* packed;
    uint32_t* inv;
    NTC_HD void operator()(uint64_t w) const {
        uint64_t word = 0;
        uint32_t bad = 0;
        for (uint32_t j = 0; j < 32; j++) {
            const uint64_t q = w * 32 + j;
            int c = -1;
            if (q >= kPad) {
                const uint64_t o = q - kPad;
                if (o < len) c = base_code(NTC_LDG(fwd + o));
                else if (revcomp && o < 2 * len) { c = base_code(NTC_LDG(fwd + (2 * len - 1 - o))); if (c >= 0) c = 3 - c; }
            }
            if (c < 0) bad |= 1u << j; else word |= (uint64_t)c << (2 * j);
        }
        packed[w] = word;
        inv[w] = bad;
    }
};
// vstart[p] = 1 when the k characters from p are all valid (a k-mer of one strand of one sequence starts here)
struct ValidStart {
    const uint32_t* inv;
    uint32_t k;
    uint64_t n_chars;     // multiple of 32; windows that would leave the text are invalid
    uint8_t* vstart;
    NTC_HD void operator()(uint64_t p) const {
        bool ok = p + k <= n_chars;
        for (uint64_t q = p; ok && q < p + k;) {
            const uint32_t off = (uint32_t)(q & 31), take = tmin<uint32_t>(32u - off, (uint32_t)(p + k - q));
            const uint32_t m = (take >= 32 ? 0xFFFFFFFFu : ((1u << take) - 1u)) << off;
            if (NTC_LDG(inv + (q >> 5)) & m) ok = false;
            q += take;
        }
        vstart[p] = ok ? 1 : 0;
    }
};
struct FlagLoad {            // scan input: a byte array of 0/1
    const uint8_t* f;
    NTC_HD Pair64 operator()(uint64_t i) const { return Pair64{f[i], 0}; }
};
struct CompactPos {          // scan post hook: the positions whose flag is set, in order
    uint32_t* out;
    NTC_HD void operator()(uint64_t i, const Pair64& pre, const Pair64& v) const { if (v.a) out[pre.a] = (uint32_t)i; }
};
struct Fill32 {
    uint32_t* p; uint32_t v;
    NTC_HD void operator()(uint64_t i) const { p[i] = v; }
};
struct Iota32 {
    uint32_t* p;
    NTC_HD void operator()(uint64_t i) const { p[i] = (uint32_t)i; }
};

// ---- 2. sorted k-mer occurrences -> distinct k-mers -------------------------------------------------------
struct RealDigit {           // radix digit of the k-mer that starts at text position p
    const uint64_t* packed; uint32_t k;
    NTC_HD uint32_t operator()(uint32_t p, uint32_t t) const { return str_digit(packed, k, p, p, t); }
};
struct AdjReal {             // head[i] = occurrence i opens a new distinct k-mer
    const uint64_t* packed; const uint32_t* sorted; uint32_t k; uint8_t* head;
    NTC_HD void operator()(uint64_t i) const {
        if (i == 0) { head[0] = 1; return; }
        const uint32_t a = sorted[i - 1], b = sorted[i];
        head[i] = common_suffix(packed, k, a, a, b, b) < k ? 1 : 0;
    }
};
struct RankPost {            // distinct rank of every occurrence; first occurrence of every distinct k-mer
    const uint32_t* sorted; uint32_t* rank_of_pos; uint32_t* uniq;
    NTC_HD void operator()(uint64_t i, const Pair64& pre, const Pair64& v) const {
        const uint32_t r = (uint32_t)(pre.a + v.a) - 1u;
        rank_of_pos[sorted[i]] = r;
        if (v.a) uniq[r] = sorted[i];
    }
};
// The occurrences are first sorted by their LAST characters only (the most significant radix digits); occurrences whose
// neighbour agrees on all of those characters are "tied" and get a full-key sort among themselves afterwards.
struct TieFlag {
    const uint64_t* packed; const uint32_t* sorted; uint32_t k; uint32_t top_chars; uint64_t m; uint8_t* tie;
    NTC_HD void operator()(uint64_t i) const {
        const uint32_t b = sorted[i];
        bool t = false;
        if (i > 0) { const uint32_t a = sorted[i - 1]; t = common_suffix(packed, k, a, a, b, b) >= top_chars; }
        if (!t && i + 1 < m) { const uint32_t c = sorted[i + 1]; t = common_suffix(packed, k, b, b, c, c) >= top_chars; }
        tie[i] = t ? 1 : 0;
    }
};
struct TiePost {             // slots and values of the tied occurrences, in slot order
    const uint32_t* sorted; uint32_t* tslot; uint32_t* tval;
    NTC_HD void operator()(uint64_t i, const Pair64& pre, const Pair64& v) const {
        if (v.a) { tslot[pre.a] = (uint32_t)i; tval[pre.a] = sorted[i]; }
    }
};
struct TieScatter {          // tied slots, ascending, receive the tied occurrences in full-key order
    const uint32_t* tslot; const uint32_t* tval; uint32_t* sorted;
    NTC_HD void operator()(uint64_t j) const { sorted[tslot[j]] = tval[j]; }
};
// ---- 3. predecessors and sources -------------------------------------------------------------------------
// An occurrence at p whose left neighbour p-1 starts a k-mer too: that k-mer's (k-1)-suffix is this k-mer's (k-1)-prefix,
// so its node is in the predecessor group.  (Several occurrences may write; any of them names the same group.)
struct PredFromText {
    const uint32_t* sorted; const uint8_t* vstart; const uint32_t* rank_of_pos; uint32_t* predrank;
    NTC_HD void operator()(uint64_t i) const {
        const uint32_t p = sorted[i];
        if (vstart[p - 1]) predrank[rank_of_pos[p]] = rank_of_pos[p - 1];
    }
};
// A k-mer seen only at the start of a sequence (or after a non-ACGT byte) may still have a predecessor elsewhere:
// binary search for a distinct k-mer whose (k-1)-suffix equals its (k-1)-prefix.  None: it is a source.
struct PredBySearch {
    const uint64_t* packed; const uint32_t* uniq; uint32_t nr; uint32_t k; uint32_t* predrank; uint8_t* is_source;
    NTC_HD void operator()(uint64_t r) const {
        if (predrank[r] != kNone) { is_source[r] = 0; return; }
        const uint32_t p = uniq[r];
        uint32_t lo = 0, hi = nr;
        while (lo < hi) {                                   // first g with suffix(g) >= prefix(r)
            const uint32_t mid = lo + (hi - lo) / 2, u = uniq[mid] + 1;
            if (cmp_colex(packed, k - 1, u, u, p, p) < 0) lo = mid + 1; else hi = mid;
        }
        bool found = false;
        if (lo < nr) { const uint32_t u = uniq[lo] + 1; found = cmp_colex(packed, k - 1, u, u, p, p) == 0; }
        if (found) predrank[r] = lo;
        is_source[r] = found ? 0 : 1;
    }
};
struct SourcePost {
    const uint32_t* uniq; uint32_t* srcpos; uint32_t* src_of;
    NTC_HD void operator()(uint64_t r, const Pair64& pre, const Pair64& v) const {
        if (v.a) { srcpos[pre.a] = uniq[r]; src_of[r] = (uint32_t)pre.a; }
    }
};
// ---- 4. dummy nodes -------------------------------------------------------------------------------------------
// id 0 = root ($^k); id 1 + (l-1) * nsrc + s = the first l characters of source s behind k-l '$' (1 <= l <= k-1).
// Ids ascend with l, so a stable sort on the '$'-as-'A' key leaves equal keys ordered by length.
struct DummyDesc {
    const uint32_t* srcpos; uint32_t nsrc; uint32_t k;
    NTC_HD void get(uint32_t id, uint32_t* vs, uint32_t* lo, uint32_t* len) const {
        if (id == 0) { *vs = 0; *lo = kPad; *len = 0; return; }       // every character below kPad is padding: all 'A'
        const uint32_t l = (id - 1) / nsrc + 1, s = (id - 1) % nsrc, p = srcpos[s];
        *vs = p - (k - l); *lo = p; *len = l;
    }
};
struct DummyDigit {
    const uint64_t* packed; DummyDesc d;
    NTC_HD uint32_t operator()(uint32_t id, uint32_t t) const {
        uint32_t vs, lo, len;
        d.get(id, &vs, &lo, &len);
        return str_digit(packed, d.k, vs, lo, t);
    }
};
struct AdjDummy {
    const uint64_t* packed; const uint32_t* sorted; DummyDesc d; uint8_t* head;
    NTC_HD void operator()(uint64_t i) const {
        if (i == 0) { head[0] = 1; return; }
        uint32_t avs, alo, alen, bvs, blo, blen;
        d.get(sorted[i - 1], &avs, &alo, &alen);
        d.get(sorted[i], &bvs, &blo, &blen);
        head[i] = (alen != blen || common_suffix(packed, d.k, avs, alo, bvs, blo) < d.k) ? 1 : 0;
    }
};
struct DummyPost {           // slot (index among the distinct dummies) of every id; one id per slot
    const uint32_t* sorted; uint32_t* dslot; uint32_t* dn_id;
    NTC_HD void operator()(uint64_t i, const Pair64& pre, const Pair64& v) const {
        const uint32_t j = (uint32_t)(pre.a + v.a) - 1u;
        dslot[sorted[i]] = j;
        if (v.a) dn_id[j] = sorted[i];
    }
};
// ---- 5. merge -------------------------------------------------------------------------------------------------
// distinct dummy j goes in front of the first distinct k-mer whose key is >= its own (equal bits: the dummy is shorter)
struct DummyInsert {
    const uint64_t* packed; const uint32_t* uniq; uint32_t nr; const uint32_t* dn_id; DummyDesc d; uint32_t* ip; uint32_t* cnt;
    NTC_HD void operator()(uint64_t j) const {
        uint32_t vs, dlo, len;
        d.get(dn_id[j], &vs, &dlo, &len);
        uint32_t lo = 0, hi = nr;
        while (lo < hi) {
            const uint32_t mid = lo + (hi - lo) / 2, u = uniq[mid];
            if (cmp_colex(packed, d.k, u, u, vs, dlo) < 0) lo = mid + 1; else hi = mid;
        }
        ip[j] = lo;
        NTC_ATOMIC_ADD32(cnt + lo, 1u);
    }
};
struct CountLoad32 {
    const uint32_t* c;
    NTC_HD Pair64 operator()(uint64_t i) const { return Pair64{c[i], 0}; }
};
struct PlaceReal {           // scan post hook over cnt: node index of distinct k-mer r = r + dummies inserted at or before r
    const uint32_t* uniq; uint32_t k; uint32_t* final_real; uint32_t* node_vs; uint8_t* node_len;
    NTC_HD void operator()(uint64_t r, const Pair64& pre, const Pair64& v) const {
        const uint32_t f = (uint32_t)(r + pre.a + v.a);
        final_real[r] = f;
        node_vs[f] = uniq[r];
        node_len[f] = (uint8_t)k;
    }
};
struct PlaceDummy {
    const uint32_t* dn_id; const uint32_t* ip; DummyDesc d; uint32_t* final_d; uint32_t* node_vs; uint8_t* node_len;
    NTC_HD void operator()(uint64_t j) const {
        uint32_t vs, lo, len;
        d.get(dn_id[j], &vs, &lo, &len);
        const uint32_t f = (uint32_t)j + ip[j];
        final_d[j] = f;
        node_vs[f] = vs;
        node_len[f] = (uint8_t)len;
    }
};
// ---- 6. LCS and subset rows ---------------------------------------------------------------------------------------
struct NodeLcs {             // '$' matches nothing: the common suffix stops at the shorter node's '$'
    const uint64_t* packed; const uint32_t* node_vs; const uint8_t* node_len; uint32_t k; uint8_t* lcs;
    NTC_HD void operator()(uint64_t v) const {
        if (v == 0) { lcs[0] = 0; return; }
        const uint32_t la = node_len[v - 1], lb = node_len[v];
        const uint32_t avs = node_vs[v - 1], bvs = node_vs[v];
        const uint32_t common = common_suffix(packed, k, avs, avs + (k - la), bvs, bvs + (k - lb));
        lcs[v] = (uint8_t)tmin(common, tmin(la, lb));
    }
};
NTC_HD uint32_t group_leader(const uint8_t* lcs, uint32_t k, uint32_t u) {   // first node of u's (k-1)-suffix group
    while (u > 0 && lcs[u] >= k - 1) u--;
    return u;
}
NTC_HD void set_edge(uint64_t* rows, uint64_t row_words, uint32_t c, uint32_t g) {
    NTC_ATOMIC_OR64(rows + (uint64_t)c * row_words + (g >> 6), 1ull << (g & 63));
}
// every non-root node has one stored in-edge: label = its last character, kept on the leader of its predecessor group
struct RowReal {
    const uint64_t* packed; const uint32_t* uniq; const uint32_t* predrank; const uint32_t* final_real; const uint32_t* src_of;
    const uint32_t* dslot; const uint32_t* final_d; const uint8_t* lcs; uint32_t nsrc; uint32_t k; uint64_t* rows; uint64_t row_words;
    NTC_HD void operator()(uint64_t r) const {
        const uint32_t pr = predrank[r];
        const uint32_t u = pr != kNone ? final_real[pr] : final_d[dslot[1u + (k - 2) * nsrc + src_of[r]]];
        set_edge(rows, row_words, packed_base(packed, (uint64_t)uniq[r] + k - 1), group_leader(lcs, k, u));
    }
};
struct RowDummy {
    const uint64_t* packed; const uint32_t* dn_id; const uint32_t* dslot; const uint32_t* final_d; const uint8_t* lcs; DummyDesc d;
    uint64_t* rows; uint64_t row_words;
    NTC_HD void operator()(uint64_t j) const {
        const uint32_t id = dn_id[j];
        if (id == 0) return;                                          // the root has no in-edge
        uint32_t vs, lo, len;
        d.get(id, &vs, &lo, &len);
        const uint32_t pred_id = len == 1 ? 0u : id - d.nsrc;
        set_edge(rows, row_words, packed_base(packed, (uint64_t)lo + len - 1), group_leader(lcs, d.k, final_d[dslot[pred_id]]));
    }
};

// ======================================= device-only part =========================================================
#if defined(__CUDACC__)
template <typename F>
__global__ void __launch_bounds__(256) for_each_kernel(F f, uint64_t n) {
    const uint64_t i = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (i < n) f(i);
}

// LSD radix sort of u32 elements, 8-bit digits supplied by a functor (digit(elem, t), t = 0 least significant): the keys
// (k-mers of up to 255 characters) are never materialised -- a pass gathers 4 characters per element from the 2-bit text,
// which is a few MB and stays in L2, and moves 4 bytes per element.  Three launches per pass: per-block digit histogram
// (which also leaves every element's digit in a byte array, so the scatter does not gather again), scan over blocks per
// digit, stable scatter.  What bounds a pass is random 32-byte sectors through L1TEX (one per gathered digit, one per
// scattered store), the same roofline as ms_kernel: the scatter therefore orders its tile in shared memory first and writes
// runs of equal digits, ~2 sectors per 16 elements instead of one each (first version: 211 us per pass over 10 M elements,
// profiles/r2/launches_index_build_gpu.txt).
constexpr int kSortThreads = 256;
constexpr int kSortRounds = 16;
constexpr int kSortTile = kSortThreads * kSortRounds;      // 4096 elements per block, 512 per warp

template <typename D>
__global__ void __launch_bounds__(kSortThreads) radix_hist_kernel(const uint32_t* __restrict__ in, uint64_t n, uint32_t t, D digit,
                                                                  uint8_t* __restrict__ dig, uint32_t* __restrict__ block_hist, uint32_t n_blocks) {
    __shared__ uint32_t h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const uint64_t base = (uint64_t)blockIdx.x * kSortTile;
#pragma unroll 4
    for (int r = 0; r < kSortRounds; r++) {
        const uint64_t i = base + (uint64_t)r * kSortThreads + threadIdx.x;
        if (i < n) {
            const uint32_t d = digit(in[i], t);
            dig[i] = (uint8_t)d;
            atomicAdd(&h[d], 1u);
        }
    }
    __syncthreads();
    block_hist[(uint64_t)threadIdx.x * n_blocks + blockIdx.x] = h[threadIdx.x];
}

__device__ __forceinline__ uint32_t block_exclusive_u32(uint32_t v, uint32_t* total) {   // 256 threads
    __shared__ uint32_t ws[8];
    __shared__ uint32_t tot;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
    if (lane == 31) ws[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        const uint32_t w = lane < 8 ? ws[lane] : 0u;
        uint32_t winc = w;
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) { const uint32_t u = __shfl_up_sync(0xffffffffu, winc, o); if (lane >= o) winc += u; }
        if (lane < 8) ws[lane] = winc - w;
        if (lane == 7) tot = winc;
    }
    __syncthreads();
    const uint32_t r = ws[warp] + inc - v;
    *total = tot;
    __syncthreads();
    return r;
}
// block d: exclusive scan of digit d's per-block counts in place; totals[d] = their sum
static __global__ void __launch_bounds__(kSortThreads) radix_scan_kernel(uint32_t* __restrict__ block_hist, uint32_t n_blocks, uint32_t* __restrict__ totals) {
    uint32_t* row = block_hist + (uint64_t)blockIdx.x * n_blocks;
    const uint32_t per = (n_blocks + kSortThreads - 1) / kSortThreads;
    const uint32_t lo = tmin(n_blocks, per * threadIdx.x), hi = tmin(n_blocks, lo + per);
    uint32_t acc = 0;
    for (uint32_t i = lo; i < hi; i++) acc += row[i];
    uint32_t total;
    uint32_t run = block_exclusive_u32(acc, &total);
    for (uint32_t i = lo; i < hi; i++) { const uint32_t v = row[i]; row[i] = run; run += v; }
    if (threadIdx.x == 0) totals[blockIdx.x] = total;
}
// stable scatter: a warp owns 512 consecutive elements and ranks them 32 at a time (lanes with equal digits form a group, the
// lowest lane of a group bumps the warp's counter -- the groups come from eight ballots); the per-warp counters become positions in the block's tile ordered
// by digit, the tile is put in that order in shared memory, and thread j writes tile element j: neighbours with the same
// digit go to neighbouring addresses
__global__ void __launch_bounds__(kSortThreads) radix_scatter_kernel(const uint32_t* __restrict__ in, const uint8_t* __restrict__ dig,
                                                                     uint32_t* __restrict__ out, uint64_t n, const uint32_t* __restrict__ block_hist,
                                                                     uint32_t n_blocks, const uint32_t* __restrict__ totals) {
    __shared__ uint32_t cnt[kSortThreads / 32][256];
    __shared__ uint32_t gbase[256];          // first output position of the block's run of digit d, minus the run's position in the tile
    __shared__ uint32_t s_val[kSortTile];
    __shared__ uint8_t s_dig[kSortTile];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int w = 0; w < kSortThreads / 32; w++) cnt[w][threadIdx.x] = 0;
    uint32_t total;
    const uint32_t before = block_exclusive_u32(totals[threadIdx.x], &total);       // elements with a smaller digit, all blocks
    const uint32_t first_out = before + block_hist[(uint64_t)threadIdx.x * n_blocks + blockIdx.x];
    const uint64_t tbase = (uint64_t)blockIdx.x * kSortTile;
    const uint64_t wbase = tbase + (uint64_t)warp * (kSortRounds * 32);
    uint32_t val[kSortRounds], key[kSortRounds];                     // key = digit << 16 | rank inside the warp's digit group
    const uint32_t lt = (1u << lane) - 1u;
    // all 16 loads of the thread first: the ranking rounds below are separated by warp barriers, which the loads may not cross
#pragma unroll
    for (int r = 0; r < kSortRounds; r++) {
        const uint64_t i = wbase + (uint64_t)r * 32 + lane;
        const bool ok = i < n;
        val[r] = ok ? in[i] : 0u;
        key[r] = ok ? (uint32_t)dig[i] : 0xFFFFFFFFu;
    }
#pragma unroll
    for (int r = 0; r < kSortRounds; r++) {
        const bool ok = key[r] != 0xFFFFFFFFu;
        const uint32_t d = ok ? key[r] : 0u;
        // lanes holding the same digit: eight ballots, one per bit (match.any measured the same)
        uint32_t peers = __ballot_sync(0xffffffffu, ok);
#pragma unroll
        for (int bit = 0; bit < 8; bit++) {
            const uint32_t has = __ballot_sync(0xffffffffu, (d >> bit) & 1u);
            peers &= ((d >> bit) & 1u) ? has : ~has;
        }
        uint32_t rank = 0;
        if (ok) rank = cnt[warp][d] + (uint32_t)__popc(peers & lt);
        __syncwarp();
        if (ok && (peers & lt) == 0) cnt[warp][d] += (uint32_t)__popc(peers);
        __syncwarp();
        if (ok) key[r] = d << 16 | rank;
    }
    __syncthreads();
    {
        uint32_t mine = 0;                                            // elements of digit threadIdx.x in this tile
#pragma unroll
        for (int w = 0; w < kSortThreads / 32; w++) mine += cnt[w][threadIdx.x];
        uint32_t run = block_exclusive_u32(mine, &total);             // the digit's position in the tile
        gbase[threadIdx.x] = first_out - run;
#pragma unroll
        for (int w = 0; w < kSortThreads / 32; w++) { const uint32_t c = cnt[w][threadIdx.x]; cnt[w][threadIdx.x] = run; run += c; }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kSortRounds; r++)
        if (key[r] != 0xFFFFFFFFu) {
            const uint32_t p = cnt[warp][key[r] >> 16] + (key[r] & 0xFFFFu);
            s_val[p] = val[r];
            s_dig[p] = (uint8_t)(key[r] >> 16);
        }
    __syncthreads();
    const uint32_t in_tile = (uint32_t)tmin<uint64_t>(kSortTile, n - tbase);
#pragma unroll 4
    for (uint32_t j = threadIdx.x; j < in_tile; j += kSortThreads) out[gbase[s_dig[j]] + j] = s_val[j];
}
#endif  // __CUDACC__

}  // namespace bld
}  // namespace ntc
