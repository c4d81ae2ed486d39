// host_index.h -- host-side flat SBWT + LCS (the canonical arrays) and the derived HBM layouts.
// Replaces sbwt::SbwtIndexVariant::SubsetMatrix + sbwt::LcsArray as used at src/lib.rs:20-22,170,263.
#pragma once
#include <cstdint>
#include <string>
#include <vector>
#include "layout.h"

struct ntc_index {
    uint32_t k = 0;
    uint64_t n = 0;                 // n_sets
    std::vector<uint64_t> rows[4];  // n bits each, LSB-first, A,C,G,T
    uint64_t C[5] = {0, 0, 0, 0, 0};
    std::vector<uint8_t> lcs;       // n bytes
};

namespace ntc {

constexpr uint32_t kMinK = 12;   // len <= 11 literals index kmer[k-len..k) (src/encode.rs:152): needs k >= 11;
                                 // the decode jump table packs 16 bases, the bucketed builder uses 8
constexpr uint32_t kMaxK = 255;  // d and LCS are stored in one byte
// nodes per 64-byte rank block (layout.h): 64 for indexes of NTC_RANK64_MIN_NODES nodes or more, else 32; NTC_RANK_NODES=32|64 overrides
uint32_t rank_nodes_for(uint64_t n_sets);

void set_error(const std::string& msg);

// recomputes C from rows and validates (edge count == n-1, k range, n < 2^32)
int32_t finalize_index(ntc_index* idx);

int32_t build_index(const uint8_t* const* seqs, const uint64_t* lens, uint64_t n_seqs, uint32_t k, bool add_revcomp,
                    int n_threads, ntc_index** out);

// ---- derived device layouts (built on the host, uploaded by ntc_ctx_create) ----

// Rank blocks, 64 bytes each.  nodes == 32: 4 x 16-byte entries (A,C,G,T), entry =
//   u32 C[c] + ones of this row before the block | 32 row bits | 32 Q bits | spare word;  nodes == 64: two sectors (layout.h)
// Q[v] = node v is alone in its (k-1)-suffix group  <=>  LCS[v] < k-1 and (v+1 == n or LCS[v+1] < k-1).
// One rank probe = one aligned 16-byte load = one 32-byte L2 sector.
void build_rank_blocks(const ntc_index& idx, std::vector<uint32_t>* out /* 16 u32 per block */, int n_threads, uint32_t nodes = 32);

// LCS words for O(1) contract_left: u32 word i = LCS[i] | min(255, i - PSV(i)) << 8 | LCS[i+1] << 16 |
// min(255, NSV(i+1) - (i+1)) << 24, PSV / NSV = nearest position left / right with a strictly smaller
// LCS (LCS[n_sets] = 0 is the sentinel).  Saturated distances fall back to a scan on the device.
void build_lcs_words(const ntc_index& idx, std::vector<ntc_lcsx_t>* out);
void spec_window(uint64_t n_sets, uint32_t* lo, uint32_t* span);

// Prefix table for the first kPrefixBases bases of a read (or of a tile's warm-up): entry x (base i of the prefix at
// bits 2i..2i+1) = 8 x u32 = interval starts after 2..8 bases, interval end after 8 bases -- what the first 8
// iterations of matching_statistics compute from the empty match (16 rank probes) in one 32-byte sector.  All
// zero when some prefix of x is absent from the index (the kernel then takes the ordinary path).
// c_start[c] = interval start after the single base c.
constexpr uint32_t kPrefixBases = 8;
void build_prefix_table(const std::vector<uint32_t>& rank_blocks, uint64_t n, std::vector<uint32_t>* out, uint32_t c_start[4], uint32_t nodes = 32);

// Decode jump table: one u64 per node = pred^16(v) (low 32 bits) | last 16 bases of node v's k-mer
// (high 32 bits, base at string position k-16+i in bits 2i..2i+1).  Root and '$' positions code 0.
void build_jump_table(const ntc_index& idx, std::vector<uint64_t>* out, int n_threads);

}  // namespace ntc
