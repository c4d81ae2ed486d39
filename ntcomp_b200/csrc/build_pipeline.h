// build_pipeline.h -- the order of operations of the GPU SBWT + LCS construction, written once against an executor X
// (for_each / scan / sort / alloc / fetch): index_build_gpu.cu runs it with kernels on a CUDA stream, tests/emul with
// loops on the CPU, so the "not gpu" suite checks the very same functors against index_build.cpp and the oracle builder.
// Stands in for kbo::build(&seqs, BuildOpts{k, add_revcomp, build_select: true, ..}) -- src/main.rs:111-134.
#pragma once
#include <cstdint>
#include <vector>

#include "build_kernels.cuh"
#include "host_index.h"

namespace ntc {
namespace bld {

constexpr uint32_t kTopDigits = 8;       // radix digits (4 characters each) of the first, partial-key sort

struct BuildStats {
    uint64_t text_chars = 0, occurrences = 0, distinct = 0, sources = 0, dummies = 0, nodes = 0, sort_passes = 0, tied = 0;
};

// X provides:
//   template <class T> T* alloc(uint64_t n)                (device memory, freed by X's destructor; nullptr => x.failed())
//   void zero(void* p, uint64_t bytes)
//   template <class F> void for_each(uint64_t n, F f)
//   template <class L, class P> uint64_t scan(L load, uint64_t n, P post)      exclusive sum of load(i).a; returns the total
//   template <class D> uint32_t* sort(uint32_t* a, uint32_t* b, uint64_t n, uint32_t t_lo, uint32_t t_hi, D digit)
//                                                          stable LSD passes over digits t_lo .. t_hi-1; returns a or b
//   void download(void* host, const void* dev, uint64_t bytes)
//   bool failed()
template <class X>
int32_t sbwt_build(X& x, const uint8_t* d_fwd, uint64_t fwd_len, uint32_t k, bool revcomp, ntc_index** out, BuildStats* stats) {
    const uint64_t strands = revcomp ? 2 : 1;
    const uint64_t n_chars = ((uint64_t)kPad + strands * fwd_len + kTail + 31) & ~31ull;
    if (n_chars >= (1ull << 32)) { set_error("input too large: text positions are 32-bit"); return 9; }
    const uint64_t n_words = n_chars / 32;
    const uint32_t n_digits = (k + 3) / 4;

    // 1. text, k-mer starts
    uint64_t* packed = x.template alloc<uint64_t>(n_words + 2);
    uint32_t* inv = x.template alloc<uint32_t>(n_words);
    uint8_t* vstart = x.template alloc<uint8_t>(n_chars);
    if (x.failed()) return 2;
    x.zero(packed + n_words, 16);
    x.for_each(n_words, PackText{d_fwd, fwd_len, revcomp ? 1u : 0u, packed, inv});
    x.for_each(n_chars, ValidStart{inv, k, n_chars, vstart});
    uint32_t* occ_a = x.template alloc<uint32_t>(n_chars);
    uint32_t* occ_b = x.template alloc<uint32_t>(n_chars);
    if (x.failed()) return 2;
    const uint64_t m = x.scan(FlagLoad{vstart}, n_chars, CompactPos{occ_a});

    // 2. sort the occurrences by k-mer (colex), distinct k-mers.  k > 32: the 8 most significant digits first (the last
    //    29..32 characters, which tell almost all k-mers of a genome apart), then a full-key sort of the few occurrences that
    //    still tie with a neighbour, written back into the slots they came from -- 8 passes over everything instead of ceil(k/4).
    //    Stable passes both times, so the result is the full LSD sort's, order of equal k-mers included.
    uint8_t* head = x.template alloc<uint8_t>(n_chars);
    if (x.failed()) return 2;
    uint32_t* sorted = occ_a;
    uint64_t n_tied = 0, passes = 0;
    if (m > 1) {
        const uint32_t t_lo = n_digits > kTopDigits ? n_digits - kTopDigits : 0;
        sorted = x.sort(occ_a, occ_b, m, t_lo, n_digits, RealDigit{packed, k});
        passes += n_digits - t_lo;
        if (t_lo > 0) {
            uint32_t* spare = sorted == occ_a ? occ_b : occ_a;       // values of the tied occurrences
            x.for_each(m, TieFlag{packed, sorted, k, k - 4 * t_lo, m, head});
            n_tied = x.scan(FlagLoad{head}, m, NoPost());           // count first: the slot list is sized to it
            if (n_tied) {
                uint32_t* tslot = x.template alloc<uint32_t>(n_tied);
                uint32_t* ttmp = x.template alloc<uint32_t>(n_tied);
                if (x.failed()) return 2;
                x.scan(FlagLoad{head}, m, TiePost{sorted, tslot, spare});
                uint32_t* tv = x.sort(spare, ttmp, n_tied, 0, n_digits, RealDigit{packed, k});
                passes += n_digits;
                x.for_each(n_tied, TieScatter{tslot, tv, sorted});
            }
        }
    }
    uint32_t* rank_of_pos = x.template alloc<uint32_t>(n_chars);
    uint32_t* uniq = x.template alloc<uint32_t>(m + 1);
    if (x.failed()) return 2;
    x.for_each(m, AdjReal{packed, sorted, k, head});
    const uint64_t nr = x.scan(FlagLoad{head}, m, RankPost{sorted, rank_of_pos, uniq});

    // 3. predecessors, sources
    uint32_t* predrank = x.template alloc<uint32_t>(nr + 1);
    uint8_t* is_source = x.template alloc<uint8_t>(nr + 1);
    uint32_t* src_of = x.template alloc<uint32_t>(nr + 1);
    uint32_t* srcpos = x.template alloc<uint32_t>(nr + 1);
    if (x.failed()) return 2;
    x.for_each(nr, Fill32{predrank, kNone});
    x.for_each(m, PredFromText{sorted, vstart, rank_of_pos, predrank});
    x.for_each(nr, PredBySearch{packed, uniq, (uint32_t)nr, k, predrank, is_source});
    const uint64_t nsrc = x.scan(FlagLoad{is_source}, nr, SourcePost{uniq, srcpos, src_of});

    // 4. dummy nodes
    const uint64_t m0 = 1 + nsrc * (uint64_t)(k - 1);
    if (m0 + nr >= (1ull << 32)) { set_error("n_sets >= 2^32 (colex ranks are stored as u32, src/encode.rs:147)"); return 9; }
    const DummyDesc dd{srcpos, (uint32_t)(nsrc ? nsrc : 1), k};
    uint32_t* did_a = x.template alloc<uint32_t>(m0);
    uint32_t* did_b = x.template alloc<uint32_t>(m0);
    uint8_t* dhead = x.template alloc<uint8_t>(m0);
    uint32_t* dslot = x.template alloc<uint32_t>(m0);
    uint32_t* dn_id = x.template alloc<uint32_t>(m0);
    if (x.failed()) return 2;
    x.for_each(m0, Iota32{did_a});
    uint32_t* dsorted = m0 > 1 ? x.sort(did_a, did_b, m0, 0, n_digits, DummyDigit{packed, dd}) : did_a;
    if (m0 > 1) passes += n_digits;
    x.for_each(m0, AdjDummy{packed, dsorted, dd, dhead});
    const uint64_t nd = x.scan(FlagLoad{dhead}, m0, DummyPost{dsorted, dslot, dn_id});

    // 5. merge the two ordered lists
    const uint64_t n = nr + nd;
    uint32_t* ip = x.template alloc<uint32_t>(nd);
    uint32_t* cnt = x.template alloc<uint32_t>(nr + 1);
    uint32_t* final_real = x.template alloc<uint32_t>(nr + 1);
    uint32_t* final_d = x.template alloc<uint32_t>(nd);
    uint32_t* node_vs = x.template alloc<uint32_t>(n);
    uint8_t* node_len = x.template alloc<uint8_t>(n);
    uint8_t* lcs = x.template alloc<uint8_t>(n);
    const uint64_t row_words = (n + 63) / 64;
    uint64_t* rows = x.template alloc<uint64_t>(4 * row_words);
    if (x.failed()) return 2;
    x.zero(cnt, (nr + 1) * 4);
    x.zero(rows, 4 * row_words * 8);
    x.for_each(nd, DummyInsert{packed, uniq, (uint32_t)nr, dn_id, dd, ip, cnt});
    x.scan(CountLoad32{cnt}, nr, PlaceReal{uniq, k, final_real, node_vs, node_len});
    x.for_each(nd, PlaceDummy{dn_id, ip, dd, final_d, node_vs, node_len});

    // 6. LCS, subset rows
    x.for_each(n, NodeLcs{packed, node_vs, node_len, k, lcs});
    x.for_each(nr, RowReal{packed, uniq, predrank, final_real, src_of, dslot, final_d, lcs, (uint32_t)nsrc, k, rows, row_words});
    x.for_each(nd, RowDummy{packed, dn_id, dslot, final_d, lcs, dd, rows, row_words});

    ntc_index* idx = new ntc_index();
    idx->k = k;
    idx->n = n;
    for (int c = 0; c < 4; c++) {
        idx->rows[c].resize(row_words);
        x.download(idx->rows[c].data(), rows + (uint64_t)c * row_words, row_words * 8);
    }
    idx->lcs.resize(n);
    x.download(idx->lcs.data(), lcs, n);
    if (x.failed()) { delete idx; return 2; }
    const int32_t rc = finalize_index(idx);
    if (rc) { delete idx; return rc; }
    if (stats) {
        stats->text_chars = n_chars; stats->occurrences = m; stats->distinct = nr; stats->sources = nsrc;
        stats->dummies = nd; stats->nodes = n; stats->sort_passes = passes; stats->tied = n_tied;
    }
    *out = idx;
    return 0;
}

}  // namespace bld
}  // namespace ntc
