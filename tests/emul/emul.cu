// tests/emul/emul.cu -- TEST INFRASTRUCTURE: runs the product's per-thread kernel bodies
// (ntcomp_b200/csrc/*_kernels.cuh, NTC_HD functions) on the CPU, one loop iteration per CUDA thread,
// so kernel LOGIC is checked against the oracle in the "not gpu" suite before GPU time is spent.
// Not part of the product: libntcomp_b200.so never links or calls this file.
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../ntcomp_b200/csrc/build_pipeline.h"
#include "../../ntcomp_b200/csrc/column_decode_kernels.cuh"
#include "../../ntcomp_b200/csrc/column_kernels.cuh"
#include "../../ntcomp_b200/csrc/column_plan.h"
#include "../../ntcomp_b200/csrc/container.h"
#include "../../ntcomp_b200/csrc/decode_kernels.cuh"
#include "../../ntcomp_b200/csrc/encode_kernels.cuh"
#include "../../ntcomp_b200/csrc/host_index.h"
#include "../../ntcomp_b200/csrc/layout_kernels.cuh"

using namespace ntc;

struct EmulIndex {
    uint32_t tile_len = 1u << 22;
    std::vector<uint32_t> rank;
    std::vector<uint64_t> jump;
    std::vector<ntc_lcsx_t> lcs;
    std::vector<uint32_t> pfx;
    DevIndex dix;
};

extern "C" {

void* emul_index_create(const ntc_index* idx) {
    EmulIndex* e = new EmulIndex();
    const uint32_t nodes = rank_nodes_for(idx->n);                   // as ntc_ctx_create; NTC_RANK_NODES=64 runs the 64-node layout on small cases
    build_rank_blocks(*idx, &e->rank, 2, nodes);
    e->dix.rank_nodes = nodes;
    build_jump_table(*idx, &e->jump, 2);
    build_lcs_words(*idx, &e->lcs);
    e->dix.rank = (const uint4*)e->rank.data();
    e->dix.lcsx = e->lcs.data();
    e->dix.jump = e->jump.data();
    e->dix.pfx = nullptr;
    if (idx->k >= 10) {                                              // as ntc_ctx_create
        build_prefix_table(e->rank, idx->n, &e->pfx, e->dix.c_start, nodes);
        e->dix.pfx = (const uint4*)e->pfx.data();
    }
    e->dix.n = (uint32_t)idx->n;
    e->dix.k = idx->k;
    spec_window(idx->n, &e->dix.spec_lo, &e->dix.spec_span);
    return e;
}
void emul_index_free(void* p) { delete (EmulIndex*)p; }
void emul_set_tile_len(void* p, uint32_t T) { ((EmulIndex*)p)->tile_len = T ? (T + 3u) & ~3u : 1u << 22; }

// state of the segment path (long reads in record mode; api.cu: ms_pipeline / encode_pipeline)
struct EmulSegs {
    bool live = false;
    uint32_t S = 0;
    std::vector<uint64_t> off, desc, bits;
    std::vector<uint32_t> info;
};

static void run_ms(const EmulIndex* e, const uint8_t* bases, const uint64_t* offsets, uint64_t n_reads, std::vector<uint64_t>* packed,
                   std::vector<uint64_t>* scratch, uint64_t** scratch_aligned, uint32_t* counts, uint32_t* status, bool keep_records = false,
                   EmulSegs* segs = nullptr) {
    uint64_t nb = offsets[n_reads] - offsets[0];
    packed->assign(nb / 32 + 6, 0);
    scratch->assign(scratch_entries(nb, n_reads) + 4, 0);
    for (uint64_t w = 0; w < (nb + 31) / 32; w++) pack2bit_thread(bases, offsets, nb, packed->data(), status, w);
    uint64_t* sc = scratch->data();
    sc += (4 - (((uintptr_t)sc / 8) & 3)) & 3;                       // 32-byte aligned, like cudaMalloc
    *scratch_aligned = sc;
    // the host side of ms_pipeline (api.cu): tiles per read -> exclusive scan -> item descriptors -> items -> counts
    const uint32_t T = e->tile_len;
    std::vector<uint64_t> tile_off(n_reads + 1, 0);
    TileLoad tl{offsets, T};
    for (uint64_t r = 0; r < n_reads; r++) tile_off[r + 1] = tile_off[r] + tl(r).a;
    std::vector<uint64_t> desc(tile_off[n_reads] + 1, ~0ull);
    TileDescPost tpost{desc.data()};
    for (uint64_t r = 0; r < n_reads; r++) tpost(r, Pair64{tile_off[r], 0}, tl(r));
    uint64_t stage[4];
    // items in reverse order: tiles of a read must not depend on the order in which they run
    for (uint64_t i = tile_off[n_reads]; i-- > 0;) {
        if (e->dix.rank_nodes == 64) ms_thread<64>(e->dix, packed->data(), offsets, desc[i], T, sc, counts, status, stage, 1, keep_records);
        else ms_thread<32>(e->dix, packed->data(), offsets, desc[i], T, sc, counts, status, stage, 1, keep_records);
    }
    if (!counts) return;
    if (keep_records && segs && nb > T && !getenv("NTC_SEG_SERIAL")) {
        const char* fs = getenv("NTC_SEG_LEN");
        const uint32_t S = fs ? (uint32_t)atoi(fs) & ~63u : seg_len_for(nb, n_reads);
        segs->S = S;
        SegLoad sl{offsets, T, S};
        segs->off.assign(n_reads + 1, 0);
        for (uint64_t r = 0; r < n_reads; r++) segs->off[r + 1] = segs->off[r] + sl(r).a;
        const uint64_t n_segs = segs->off[n_reads];
        segs->desc.assign(n_segs + 1, ~0ull);
        TileDescPost spost{segs->desc.data()};
        for (uint64_t r = 0; r < n_reads; r++) spost(r, Pair64{segs->off[r], 0}, sl(r));
        segs->bits.assign((n_segs + 1) * (S / 64), ~0ull);          // garbage, as fresh device memory would be
        segs->info.assign((n_segs + 1) * 4, 0xDEADBEEFu);
        for (uint64_t gi = n_segs; gi-- > 0;) seg_spec_thread(e->dix, offsets, T, S, sc, segs->desc.data(), segs->bits.data(), segs->info.data(), status, gi);
        for (uint64_t r = 0; r < n_reads; r++) seg_join_thread(e->dix, offsets, n_reads, T, S, sc, segs->off.data(), segs->bits.data(), segs->info.data(), counts, status, r);
        segs->live = true;
        return;
    }
    for (uint64_t r = 0; r < n_reads; r++) count_thread(e->dix, packed->data(), offsets, n_reads, T, sc, counts, status, keep_records, r);
}

// mode 0 records, 1 triples.  Returns the accumulated status bits.
uint32_t emul_encode(void* ep, const uint8_t* bases, const uint64_t* offsets, uint64_t n_reads, int mode, uint64_t* out, uint64_t cap,
                     uint64_t* n_records, uint64_t* rec_offsets) {
    const EmulIndex* e = (const EmulIndex*)ep;
    uint32_t status = 0;
    std::vector<uint64_t> packed, scratch;
    std::vector<uint32_t> counts(n_reads + 1, 0);
    uint64_t* sc = nullptr;
    EmulSegs segs;
    run_ms(e, bases, offsets, n_reads, &packed, &scratch, &sc, counts.data(), &status, mode == 0, &segs);
    std::vector<uint64_t> roff(n_reads + 1, 0);
    for (uint64_t r = 0; r < n_reads; r++) roff[r + 1] = roff[r] + counts[r];   // exclusive scan (scan.cuh on the GPU)
    for (uint64_t r = 0; r < n_reads; r++) {
        if (mode == 0) emit_thread<0>(e->dix, packed.data(), offsets, n_reads, sc, counts.data(), roff.data(), nullptr, out, cap, &status, e->tile_len, r, segs.live);
        else emit_thread<1>(e->dix, packed.data(), offsets, n_reads, sc, counts.data(), roff.data(), nullptr, out, cap, &status, e->tile_len, r);
    }
    if (mode == 0 && segs.live)
        for (uint64_t gi = 0; gi < segs.off[n_reads]; gi++)
            seg_emit_thread(e->dix, packed.data(), offsets, e->tile_len, segs.S, sc, segs.desc.data(), segs.info.data(), counts.data(), roff.data(), nullptr, out, cap, &status, gi);
    *n_records = roff[n_reads];
    if (rec_offsets) memcpy(rec_offsets, roff.data(), (n_reads + 1) * 8);
    return status;
}

uint32_t emul_ms(void* ep, const uint8_t* bases, const uint64_t* offsets, uint64_t n_reads, uint8_t* d_out, uint32_t* s_out) {
    const EmulIndex* e = (const EmulIndex*)ep;
    uint32_t status = 0;
    std::vector<uint64_t> packed, scratch;
    uint64_t* sc = nullptr;
    run_ms(e, bases, offsets, n_reads, &packed, &scratch, &sc, nullptr, &status);
    for (uint64_t r = 0; r < n_reads; r++) {
        uint64_t a = offsets[r] - offsets[0];
        const uint64_t* rs = sc + scratch_base(a, r);
        for (uint64_t j = 0; j < offsets[r + 1] - offsets[r]; j++) { d_out[a + j] = (uint8_t)entry_d(rs[j]); s_out[a + j] = entry_s(rs[j]); }
    }
    return status;
}

uint32_t emul_dict_pack(void* ep, const uint64_t* triples, uint64_t n, uint64_t* out) {
    const EmulIndex* e = (const EmulIndex*)ep;
    uint32_t status = 0;
    for (uint64_t i = 0; i < n; i++) dict_pack_thread(e->dix, triples, n, out, &status, i);
    return status;
}

uint32_t emul_decode(void* ep, const uint64_t* rec, uint64_t n_rec, uint64_t first_no, int fasta, uint8_t* out, uint64_t cap,
                     uint64_t* seq_offsets, uint64_t cap_seqs, uint64_t* counts) {
    const EmulIndex* e = (const EmulIndex*)ep;
    uint32_t status = 0;
    counts[0] = counts[1] = 0;
    if (n_rec == 0) { if (seq_offsets) seq_offsets[0] = 0; return 0; }
    std::vector<uint64_t> base_off(n_rec + 1), seq_idx(n_rec + 1), seq_first(n_rec + 1, 0);
    RecordLoad load{rec};
    uint64_t a = 0, b = 0;
    SeqFirstPost post{seq_first.data()};
    for (uint64_t i = 0; i < n_rec; i++) { base_off[i] = a; seq_idx[i] = b; Pair64 p = load(i); post(i, Pair64{a, b}, p); a += p.a; b += p.b; }
    base_off[n_rec] = a; seq_idx[n_rec] = b;
    uint64_t totals[2] = {a, b};
    std::vector<uint64_t> seq_endkey(n_rec + 1, 0);
    for (uint64_t j = 0; j <= n_rec; j++)
        seq_layout_thread(base_off.data(), seq_first.data(), totals, n_rec, seq_endkey.data(), seq_offsets, cap_seqs, 0, first_no, fasta, cap,
                          counts, &status, j);
    for (uint64_t i = 0; i < n_rec; i++)
        walk_thread(e->dix, rec, n_rec, base_off.data(), seq_idx.data(), seq_endkey.data(), totals, first_no, fasta, out, cap, counts, &status, i);
    return status;
}

// GPU column codec bodies (column_kernels.cuh) on the CPU, with the host planning of column_plan.h: the records of a
// batch cut into blocks at blk_rec[0..n_blocks] -> coded words of all (block, column) segments back to back in `words`;
// seg_out[4 * (4 b + q) ..] = n_values, n_words, param, word_base; status_out[b] as ntc_coded_block::status.
// Returns 0, -2 when `words` is too small.
int emul_batch_columns(const uint64_t* rec, const uint64_t* blk_rec_in, uint32_t n_blocks, uint64_t* words, uint64_t cap, uint64_t* seg_out,
                       int32_t* status_out, uint64_t* total_out) {
    std::vector<uint64_t> blk_rec(blk_rec_in, blk_rec_in + n_blocks + 1);
    const uint64_t n = blk_rec[n_blocks];
    std::vector<uint64_t> ref_idx(n + 1), lit_off(n + 1);
    std::vector<uint32_t> colex(n + 1), len(n + 1);
    std::vector<uint8_t> flags(n + 1);
    std::vector<uint64_t> lits(n * 32 / 31 + n_blocks + 2, 0);
    SplitLoad load{rec};
    uint64_t a = 0, b = 0;
    for (uint64_t i = 0; i < n; i++) { ref_idx[i] = a; lit_off[i] = b; Pair64 p = load(i); a += p.a; b += p.b; }
    ref_idx[n] = a; lit_off[n] = b;
    ColArrays arr{colex.data(), len.data(), flags.data(), lits.data()};
    std::vector<BlkStats> stats(n_blocks, BlkStats{0, 0, 0, 0, 0, 0, 0, 0});
    uint64_t lw = 0;
    for (uint32_t k = 0; k < n_blocks; k++) {                          // split_kernel + blk_stats_kernel + lit_pack_kernel
        const uint64_t rs = blk_rec[k], re = blk_rec[k + 1];
        for (uint64_t i = re; i-- > rs;) {
            SplitContrib c = split_thread(rec, ref_idx.data(), arr, i);
            stats[k].max_colex = std::max(stats[k].max_colex, c.max_colex);
            stats[k].sum_len += c.sum_len;
            stats[k].sum_flags += c.sum_flags;
        }
        stats[k].n_rec = re - rs;
        stats[k].n_ref = ref_idx[re] - ref_idx[rs];
        stats[k].lit_bases = lit_off[re] - lit_off[rs];
        stats[k].lit_word_base = lw;
        const uint64_t nw = (stats[k].lit_bases + 30) / 31;
        for (uint64_t w = nw; w-- > 0;) {
            lits[lw + w] = lit_word_thread(rec, lit_off.data(), rs, re, w);
            stats[k].max_lit = std::max(stats[k].max_lit, lits[lw + w]);
        }
        lw += nw;
    }
    std::vector<ColSeg> segs;
    std::vector<HostSeg> hseg;
    std::vector<int32_t> status;
    uint64_t bound = 0;
    const uint64_t flat = plan_encode(stats.data(), blk_rec, &segs, &hseg, &status, &bound);
    const uint32_t n_segs = n_blocks * 4;
    std::vector<uint64_t> bit_off(flat + 1, 0);
    FlatCodeLen cl{arr, segs.data(), n_segs};
    for (uint64_t i = 0; i < flat; i++) bit_off[i + 1] = bit_off[i] + cl(i).a;
    uint64_t total = 0;
    seg_finish_serial(segs.data(), n_segs, bit_off.data(), &total);
    if (total > bound) return -3;
    if (total > cap) return -2;
    for (uint64_t w = total; w-- > 0;) words[w] = bswap64(pack_word_thread(arr, segs.data(), n_segs, bit_off.data(), w));
    for (uint32_t s2 = 0; s2 < n_segs; s2++) {
        seg_out[4 * s2] = hseg[s2].n_values; seg_out[4 * s2 + 1] = segs[s2].n_words; seg_out[4 * s2 + 2] = hseg[s2].param; seg_out[4 * s2 + 3] = segs[s2].word_base;
    }
    for (uint32_t k = 0; k < n_blocks; k++) status_out[k] = status[k];
    *total_out = total;
    return 0;
}

// GPU column decoder bodies (column_decode_kernels.cuh) + plan_decode on the CPU: cols[4 * (4 b + q) ..] = n_values, n_words,
// param, word_offset; words as stored (big-endian u64).  Returns 0 and the records of all blocks (concatenated), or the
// C-ABI error code (7 = NTC_ERR_IO ...); *detail: 1 stream ends early, 2.. zip error bits << 1.
int emul_columns_decode(const uint64_t* cols_in, uint32_t n_blocks, const uint64_t* words_be, uint64_t n_words, int full_last_word, int trim,
                        uint64_t* rec, uint64_t cap, uint64_t* n_rec_out, uint32_t* detail) {
    std::vector<HostCol> cols((size_t)n_blocks * 4);
    for (size_t i = 0; i < cols.size(); i++) cols[i] = HostCol{cols_in[4 * i], cols_in[4 * i + 1], cols_in[4 * i + 2], cols_in[4 * i + 3]};
    DecodePlan plan;
    std::string err;
    *detail = 0;
    *n_rec_out = 0;
    int32_t rc = plan_decode(cols.data(), n_blocks, n_words, &plan, &err);
    if (rc) return rc;
    if (plan.n_rec > cap) return 3;
    const uint32_t n_segs = n_blocks * 4;
    std::vector<uint64_t> w(n_words + 2, 0);
    for (uint64_t i = 0; i < n_words; i++) w[i] = __builtin_bswap64(words_be[i]);      // bswap_kernel
    std::vector<uint32_t> table(plan.table + 1), entry(plan.chunks + 1);
    std::vector<uint64_t> first(plan.chunks + 1), values(plan.n_values + 1, 0), totals(n_segs, 0);
    for (uint64_t i = plan.table; i-- > 0;) cw_table_thread(plan.segs.data(), n_segs, w.data(), plan.table, table.data(), i);
    for (uint32_t s2 = 0; s2 < n_segs; s2++) cw_resolve_serial(plan.segs[s2], table.data(), entry.data(), first.data(), &totals[s2]);
    for (uint64_t c = plan.chunks; c-- > 0;) cw_extract_thread(plan.segs.data(), n_segs, w.data(), plan.chunks, entry.data(), first.data(), values.data(), c);
    for (uint32_t s2 = 0; s2 < n_segs; s2++) if (totals[s2] < plan.segs[s2].n_values) { *detail = 1; return 7; }
    const uint64_t n = plan.n_rec;
    const uint64_t* flags = values.data() + plan.region[2];
    std::vector<uint64_t> ref_idx(n + 1), lit_off(n + 1);
    FlagLoad load{flags};
    uint64_t a = 0, b = 0;
    for (uint64_t i = 0; i < n; i++) { ref_idx[i] = a; lit_off[i] = b; Pair64 pr = load(i); a += pr.a; b += pr.b; }
    ref_idx[n] = a; lit_off[n] = b;
    uint32_t zerr = 0;
    for (uint64_t i = n; i-- > 0;)
        zip_thread(values.data(), flags, plan.zb.data(), n_blocks, ref_idx.data(), lit_off.data(), full_last_word, rec, &zerr, i, n);
    if (zerr) { *detail = zerr << 1; return 7; }
    if (trim)                                                            // blk_trim_kernel
        for (uint32_t k = 0; k < n_blocks; k++)
            for (uint64_t j = plan.blk_rec[k]; j < plan.blk_rec[k + 1] && ((rec[j] >> 56) & 1u) == 0; j++) rec[j] = orphan_record();
    *n_rec_out = n;
    return 0;
}


// one code stream through the decoder bodies (table / resolve / extract with a single segment): returns the number of
// codes found (values beyond n are not stored).  -1: bad parameter.
int64_t emul_column_decode(const uint64_t* words_be, uint64_t n_words, uint64_t n, int rice, uint64_t param, uint64_t* out) {
    CodeParams p;
    if (rice) { if (param > 63) return -1; p = CodeParams{1, (uint32_t)param, 0, 0}; }
    else { if (param < 2) return -1; unsigned l; uint64_t limit; minimal_binary_shape(param, &l, &limit); p = CodeParams{0, 0, l, limit}; }
    if (n_words == 0) return 0;
    std::vector<uint64_t> w(n_words);
    for (uint64_t i = 0; i < n_words; i++) w[i] = __builtin_bswap64(words_be[i]);
    const uint64_t n_chunks = (n_words * 64 + kCwChunkBits - 1) / kCwChunkBits;
    const uint32_t ns = cw_num_states(p);
    DecSeg seg{0, n_words, n, 0, 0, n_chunks, 0, p, ns, 0};
    std::vector<uint32_t> table(n_chunks * ns), entry(n_chunks);
    std::vector<uint64_t> first(n_chunks);
    for (uint64_t i = n_chunks * ns; i-- > 0;) cw_table_thread(&seg, 1, w.data(), n_chunks * ns, table.data(), i);
    uint64_t total = 0;
    cw_resolve_serial(seg, table.data(), entry.data(), first.data(), &total);
    for (uint64_t c = n_chunks; c-- > 0;) cw_extract_thread(&seg, 1, w.data(), n_chunks, entry.data(), first.data(), out, c);
    return (int64_t)total;
}

}  // extern "C"

// ---- GPU index construction (build_pipeline.h) with loops in place of launches -------------------------------------
// for_each = one call per element in order, scan = a serial exclusive sum calling the post hook, sort = one stable
// counting sort per digit with the product's digit functor (what the radix kernels compute, pass for pass).
namespace {
struct CpuExec {
    std::vector<void*> owned;
    bool failed() const { return false; }
    template <class T> T* alloc(uint64_t n) { void* p = calloc(n ? n : 1, sizeof(T)); owned.push_back(p); return (T*)p; }
    void zero(void* p, uint64_t bytes) { memset(p, 0, bytes); }
    template <class F> void for_each(uint64_t n, F f) { for (uint64_t i = 0; i < n; i++) f(i); }
    template <class L, class P> uint64_t scan(L load, uint64_t n, P post) {
        Pair64 run{0, 0};
        for (uint64_t i = 0; i < n; i++) { const Pair64 v = load(i); post(i, run, v); run = run + v; }
        return run.a;
    }
    template <class D> uint32_t* sort(uint32_t* a, uint32_t* b, uint64_t n, uint32_t t_lo, uint32_t t_hi, D digit) {
        for (uint32_t t = t_lo; t < t_hi; t++) {
            uint64_t pos[257] = {0};
            for (uint64_t i = 0; i < n; i++) pos[digit(a[i], t) + 1]++;
            for (int d = 0; d < 256; d++) pos[d + 1] += pos[d];
            for (uint64_t i = 0; i < n; i++) b[pos[digit(a[i], t)]++] = a[i];
            uint32_t* s2 = a; a = b; b = s2;
        }
        return a;
    }
    void download(void* host, const void* dev, uint64_t bytes) { memcpy(host, dev, bytes); }
    ~CpuExec() { for (void* p : owned) free(p); }
};
}  // namespace

extern "C" int32_t emul_index_build(const uint8_t* const* seqs, const uint64_t* lens, uint64_t n_seqs, uint32_t k, int32_t add_revcomp,
                                    ntc_index** out, uint64_t* stats /* [7], nullable */) {
    *out = nullptr;
    uint64_t fwd_len = 0;
    for (uint64_t s2 = 0; s2 < n_seqs; s2++) fwd_len += lens[s2] + 1;
    std::vector<uint8_t> fwd(fwd_len ? fwd_len : 1);
    uint64_t o = 0;
    for (uint64_t s2 = 0; s2 < n_seqs; s2++) { if (lens[s2]) memcpy(fwd.data() + o, seqs[s2], lens[s2]); o += lens[s2]; fwd[o++] = '\n'; }
    CpuExec x;
    bld::BuildStats bs;
    const int32_t rc = bld::sbwt_build(x, fwd.data(), fwd_len, k, add_revcomp != 0, out, &bs);
    if (stats) { stats[0] = bs.text_chars; stats[1] = bs.occurrences; stats[2] = bs.distinct; stats[3] = bs.sources; stats[4] = bs.dummies; stats[5] = bs.nodes; stats[6] = bs.sort_passes; }
    return rc;
}

// ---- derived layouts: the kernels' per-element bodies (layout_kernels.cuh) in loops against the host builders (host_index.cpp) ----
// returns 0 when all four arrays are identical for both rank-block layouts, else a code naming the first difference
extern "C" int32_t emul_layouts_check(const ntc_index* idx) {
    const uint64_t n = idx->n, row_words = (n + 63) / 64;
    std::vector<uint64_t> pre[4];
    lay::Rows rows;
    rows.row_words = row_words;
    for (int c = 0; c < 4; c++) {
        pre[c].assign(row_words + 1, 0);
        for (uint64_t w = 0; w < row_words; w++) pre[c][w + 1] = pre[c][w] + (uint64_t)__builtin_popcountll(idx->rows[c][w]);   // the scan
        rows.row[c] = idx->rows[c].data();
        rows.pre[c] = pre[c].data();
    }
    for (int c = 0; c < 5; c++) rows.C[c] = idx->C[c];
    for (uint32_t nodes : {32u, 64u}) {
        std::vector<uint32_t> want, got, want_pfx, got_pfx((size_t)(1u << 16) * 8, 0xDEADBEEFu);
        uint32_t c_start[4];
        build_rank_blocks(*idx, &want, 2, nodes);
        const uint64_t n_blocks = n / nodes + 1;
        got.assign(n_blocks * 16, 0xDEADBEEFu);
        for (uint64_t b = 0; b < n_blocks; b++) {
            if (nodes == 32) lay::rank_block_thread<32>(rows, idx->lcs.data(), n, idx->k, got.data(), b);
            else lay::rank_block_thread<64>(rows, idx->lcs.data(), n, idx->k, got.data(), b);
        }
        if (got != want) return 10 + (int32_t)nodes;
        build_prefix_table(want, n, &want_pfx, c_start, nodes);
        for (uint64_t x = 0; x < (1u << 16); x++) {
            if (nodes == 32) lay::prefix_entry_thread<32>(got.data(), n, got_pfx.data(), x);
            else lay::prefix_entry_thread<64>(got.data(), n, got_pfx.data(), x);
        }
        if (got_pfx != want_pfx) return 20 + (int32_t)nodes;
        for (int c = 0; c < 4; c++) if (c_start[c] != (uint32_t)idx->C[c]) return 30;
    }
    {
        std::vector<ntc_lcsx_t> want;
        build_lcs_words(*idx, &want);
        std::vector<uint32_t> got(n + 1, 0xDEADBEEFu);
        for (uint64_t i = 0; i <= n; i++) lay::lcs_word_thread(idx->lcs.data(), n, got.data(), i);
        if (want.size() != got.size() || memcmp(want.data(), got.data(), got.size() * 4) != 0) return 40;
    }
    {
        std::vector<uint64_t> want, a(n, ~0ull), b(n, ~0ull);
        build_jump_table(*idx, &want, 2);
        for (uint64_t e = 0; e < row_words * 4; e++) lay::pred_init_thread(rows, a.data(), e);
        for (uint32_t t = 1; t < 16; t *= 2) {
            for (uint64_t v = 0; v < n; v++) lay::pred_double_thread(a.data(), b.data(), t, v);
            a.swap(b);
        }
        if (a != want) return 50;
    }
    return 0;
}
