// column_plan.h -- host-side planning of the GPU column codec (api_columns.inc): which (block, column) segments a batch
// has, where their values / words / chunks live, and the code parameters from the per-block statistics.  Pure host
// arithmetic, shared with tests/emul so the CPU suite exercises the same planning as the GPU path.
#pragma once
#include <algorithm>
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/ntcomp_b200.h"
#include "column_decode_kernels.cuh"
#include "column_kernels.cuh"
#include "container.h"

namespace ntc {

struct HostSeg {       // what the host keeps about one (block, column) pair of a coded batch
    uint64_t n_values = 0, n_words = 0, param = 0, word_base = 0;
};
struct HostCol {       // one coded column as the host hands it to the decoder: header fields + where its words are
    uint64_t n_values, n_words, param, word_offset;
};

// rice_encode / minimal_binary_encode parameters of one column from its statistics (src/encode.rs:62-63,80-82)
inline CodeParams rice_params(uint64_t n, uint64_t sum, uint64_t* param) {
    *param = rice_parameter(n, sum);
    return CodeParams{1, (uint32_t)*param, 0, 0};
}
inline CodeParams minbin_params(uint64_t mx, uint64_t* param) {
    *param = mx + 2;                                                               // src/encode.rs:82
    unsigned l; uint64_t limit;
    minimal_binary_shape(*param, &l, &limit);
    return CodeParams{0, 0, l, limit};
}

// Encode direction: segments 4 b + q of a batch from the statistics the device gathered.  Parameters come from the very
// arithmetic of container.cpp / the reference (f64 logs).  A block without a reference record or without a literal base
// cannot be coded (Q4): status NTC_ERR_EMPTY_COLUMN, no values.  Returns the flattened value count; *words_bound = an upper
// bound of the coded words (grid size of the packing launch).
inline uint64_t plan_encode(const BlkStats* h_stats, const std::vector<uint64_t>& blk_rec, std::vector<ColSeg>* segs, std::vector<HostSeg>* hseg,
                            std::vector<int32_t>* status, uint64_t* words_bound) {
    const uint32_t n_blocks = (uint32_t)(blk_rec.size() - 1);
    segs->assign((size_t)n_blocks * 4, ColSeg());
    hseg->assign((size_t)n_blocks * 4, HostSeg());
    status->assign(n_blocks, NTC_OK);
    uint64_t flat = 0, ref_base = 0, bits_bound = 0;
    for (uint32_t b = 0; b < n_blocks; b++) {
        const BlkStats& s = h_stats[b];
        const uint64_t n_lw = (s.lit_bases + 30) / 31;
        const bool dropped = s.n_ref == 0 || n_lw == 0;                             // Q4: minimal_binary_encode fails on an empty column
        if (dropped) (*status)[b] = s.n_rec == 0 ? NTC_ERR_EMPTY_READ : NTC_ERR_EMPTY_COLUMN;
        const uint64_t cnt[4] = {s.n_ref, s.n_ref, s.n_rec, n_lw};
        const uint64_t base[4] = {ref_base, ref_base, blk_rec[b], s.lit_word_base};
        uint64_t param[4] = {0, 0, 0, 0};
        CodeParams cp[4];
        cp[0] = minbin_params(s.max_colex, &param[0]);
        cp[1] = rice_params(s.n_ref, s.sum_len, &param[1]);
        cp[2] = rice_params(s.n_rec, s.sum_flags, &param[2]);
        cp[3] = minbin_params(s.max_lit, &param[3]);
        const uint64_t sums[4] = {0, s.sum_len, s.sum_flags, 0};
        for (uint32_t q = 0; q < 4; q++) {
            ColSeg& g = (*segs)[(size_t)b * 4 + q];
            g = ColSeg{base[q], dropped ? 0 : cnt[q], flat, 0, 0, 0, cp[q], q, 0};
            HostSeg& h = (*hseg)[(size_t)b * 4 + q];
            h.n_values = g.n_values;
            h.param = param[q];
            flat += g.n_values;
            bits_bound += cp[q].rice ? (sums[q] >> cp[q].b) + g.n_values * (cp[q].b + 1) : g.n_values * (cp[q].l + 1);
            bits_bound += 64;
        }
        ref_base += s.n_ref;
    }
    *words_bound = bits_bound / 64 + (uint64_t)n_blocks * 4 + 1;
    return flat;
}

struct DecodePlan {
    std::vector<DecSeg> segs;          // 4 b + q
    std::vector<ZipBlock> zb;
    std::vector<uint64_t> blk_rec;     // n_blocks + 1
    uint64_t region[4] = {0, 0, 0, 0}; // value array, column-major: the flags of all blocks are one region in record order
    uint64_t n_values = 0, chunks = 0, table = 0, n_rec = 0;
};

// Decode direction: validates the header fields that come from the file (the same checks as the host reader,
// container.cpp) and lays out words / values / chunks / tables.  NTC_OK, or an error code with *err set.
inline int32_t plan_decode(const HostCol* cols, uint32_t n_blocks, uint64_t n_words_total, DecodePlan* plan, std::string* err) {
    const int rice[4] = {0, 1, 1, 0};                                               // src/lib.rs:358-361
    uint64_t cnt[4] = {0, 0, 0, 0};
    for (uint32_t b = 0; b < n_blocks; b++)
        for (int q = 0; q < 4; q++) {
            const HostCol& c = cols[(size_t)b * 4 + q];
            if (c.word_offset > n_words_total || c.n_words > n_words_total - c.word_offset) { *err = "column words outside the word buffer"; return NTC_ERR_INVALID_ARG; }
            if (c.n_values > c.n_words * 64) {                                      // every code is at least one bit
                *err = "malformed sub-block " + std::to_string(q) + " (code -6): more values announced than bits present";
                return NTC_ERR_IO;
            }
            if (rice[q] ? c.param > 63 : c.param < 2) { *err = rice[q] ? "Rice parameter out of range" : "minimal-binary bound out of range"; return NTC_ERR_IO; }
            cnt[q] += c.n_values;
        }
    plan->region[0] = 0; plan->region[1] = cnt[0]; plan->region[2] = cnt[0] + cnt[1]; plan->region[3] = cnt[0] + cnt[1] + cnt[2];
    plan->n_values = plan->region[3] + cnt[3];
    plan->segs.assign((size_t)n_blocks * 4, DecSeg());
    plan->zb.assign(n_blocks, ZipBlock());
    plan->blk_rec.assign(n_blocks + 1, 0);
    uint64_t at[4] = {0, 0, 0, 0}, chunks = 0, table = 0, n_rec = 0;
    for (uint32_t b = 0; b < n_blocks; b++) {
        const HostCol* c = cols + (size_t)b * 4;
        if (c[0].n_values != c[1].n_values) { *err = "columns do not zip (code -1; src/decode.rs:108)"; return NTC_ERR_IO; }
        for (int q = 0; q < 4; q++) {
            CodeParams p;
            if (rice[q]) p = CodeParams{1, (uint32_t)c[q].param, 0, 0};
            else { unsigned l; uint64_t limit; minimal_binary_shape(c[q].param, &l, &limit); p = CodeParams{0, 0, l, limit}; }
            const uint64_t n_chunks = c[q].n_values ? (c[q].n_words * 64 + kCwChunkBits - 1) / kCwChunkBits : 0;
            const uint32_t ns = cw_num_states(p);
            plan->segs[(size_t)b * 4 + q] = DecSeg{c[q].word_offset, c[q].n_words, c[q].n_values, plan->region[q] + at[q], chunks, n_chunks, table, p, ns, 0};
            chunks += n_chunks;
            table += n_chunks * ns;
        }
        plan->zb[b] = ZipBlock{n_rec, c[2].n_values, plan->region[0] + at[0], plan->region[1] + at[1], c[0].n_values, plan->region[3] + at[3], c[3].n_values};
        plan->blk_rec[b] = n_rec;
        n_rec += c[2].n_values;
        for (int q = 0; q < 4; q++) at[q] += c[q].n_values;
    }
    plan->blk_rec[n_blocks] = n_rec;
    plan->n_rec = n_rec;
    plan->chunks = chunks;
    plan->table = table;
    return NTC_OK;
}

}  // namespace ntc
