// oracle/oracle_index.h -- TEST INFRASTRUCTURE ONLY (CPU oracle; never linked into the product).
//
// PARITY UNPINNED: the reference (tmaklin/ntcomp) holds no golden vectors for this path
// (tests/fasta_data.rs:93 asserts only the round trip) and its index arithmetic lives in the
// un-vendored crates sbwt 0.3.11 / simple-sds-sbwt 0.3.2 / kbo 0.5.1 (Cargo.lock:1358,1446,740),
// which cannot be built here (no cargo/rustc).  This file restates the *published* SBWT
// definition (Alanko, Puglisi, Vuohtoniemi 2023) and the sbwt crate's query semantics:
//
//   * node set  = distinct k-mers of all inputs (split at non-ACGT) [+ reverse complements]
//                 + '$'-padded proper prefixes of every k-mer lacking a predecessor + root $^k
//   * order     = colexicographic, $ < A < C < G < T
//   * row c bit i = node i is the first node of its (k-1)-suffix group and node[1..k)+c is a node
//   * C[0] = 1, C[c+1] = C[c] + ones(row c);  C[4] == n
//   * rank_c(i) = ones of row c in [0,i);  select_c(j) = position of the j-th (0-based) one
//   * LCS[i]    = longest common suffix of nodes i-1 and i ('$' never matches), LCS[0] = 0
//
// Call sites in the reference that rely on these: src/lib.rs:106,108,140,142,171-173,181,286,291;
// src/encode.rs:143,151.
#pragma once
#include <cstdint>
#include <cstddef>
#include <string>
#include <vector>

namespace orc {

constexpr int kMaxK = 255;  // LCS and d fit one byte; reference test uses k = 255 (tests/fasta_data.rs:50)

struct Index {
    int k = 0;
    uint64_t n = 0;                       // number of nodes (sets)
    std::vector<uint64_t> rows[4];        // n bits each, LSB-first in 64-bit words (+1 pad word)
    std::vector<uint32_t> rank_cum[4];    // ones before each 64-bit word
    std::vector<uint32_t> ones[4];        // select support: positions of ones
    uint64_t C[5] = {0, 0, 0, 0, 0};      // C[4] == n
    std::vector<uint8_t> lcs;             // n entries

    // --- query primitives (sbwt crate semantics, see header comment) ---
    inline uint64_t rank(int c, uint64_t i) const {
        uint64_t w = i >> 6, b = i & 63;
        uint64_t r = rank_cum[c][w];
        if (b) r += (uint64_t)__builtin_popcountll(rows[c][w] & ((1ull << b) - 1));
        return r;
    }
    inline bool bit(int c, uint64_t i) const { return (rows[c][i >> 6] >> (i & 63)) & 1; }
    inline uint64_t select(int c, uint64_t j) const { return ones[c][j]; }
    // in-label of node v: 0..3, or -1 for the root ($)
    inline int inlabel(uint64_t v) const {
        if (v == 0) return -1;
        int c = 0;
        while (c < 3 && C[c + 1] <= v) c++;
        return c;
    }
    // inverse LF step; root maps to itself
    inline uint64_t pred(uint64_t v) const {
        int c = inlabel(v);
        if (c < 0) return 0;
        return select(c, v - C[c]);
    }
    void finalize();                      // builds rank_cum / ones / C from rows
};

inline int base_code(uint8_t b) {
    switch (b) { case 'A': return 0; case 'C': return 1; case 'G': return 2; case 'T': return 3; default: return -1; }
}

// Builds the index from sequences (ASCII, upper case; non-ACGT splits a sequence).
// Mirrors kbo::build(&seqs, BuildOpts{k, add_revcomp, build_select=true,..}) as called at
// src/main.rs:111-134 and tests/fasta_data.rs:49-56.
Index* build_index(const uint8_t* const* seqs, const uint64_t* lens, uint64_t n_seqs, int k, bool add_revcomp,
                   std::string* err);

// Wraps externally produced arrays (used to run the oracle on an index flattened elsewhere).
Index* index_from_arrays(int k, uint64_t n, const uint64_t* const rows[4], const uint8_t* lcs, std::string* err);

// extend_right / search / matching statistics / access_kmer (sbwt crate; SURVEY §8a A1,A4,A8)
struct Interval { uint64_t s, e; };
inline Interval extend_right(const Index& ix, Interval I, int c) {
    if (c < 0) return Interval{0, 0};
    return Interval{ix.C[c] + ix.rank(c, I.s), ix.C[c] + ix.rank(c, I.e)};
}
inline Interval contract_left(const Index& ix, Interval I, uint64_t target) {
    uint64_t s = I.s, e = I.e;
    while (s > 0 && ix.lcs[s] >= target) s--;
    while (e < ix.n && ix.lcs[e] >= target) e++;
    return Interval{s, e};
}
// returns false when the pattern does not occur (Option::None)
bool search(const Index& ix, const uint8_t* pat, uint64_t len, Interval* out);
void matching_statistics(const Index& ix, const uint8_t* read, uint64_t len, uint32_t* d, uint64_t* s, uint64_t* e);
void access_kmer(const Index& ix, uint64_t v, uint8_t* out /*k bytes, '$'-padded*/);

}  // namespace orc
