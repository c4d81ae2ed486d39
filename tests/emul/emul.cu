// tests/emul/emul.cu -- TEST INFRASTRUCTURE: runs the product's per-thread kernel bodies
// (ntcomp_b200/csrc/*_kernels.cuh, NTC_HD functions) on the CPU, one loop iteration per CUDA thread,
// so kernel LOGIC is checked against the oracle in the "not gpu" suite before GPU time is spent.
// Not part of the product: libntcomp_b200.so never links or calls this file.
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../ntcomp_b200/csrc/column_decode_kernels.cuh"
#include "../../ntcomp_b200/csrc/column_kernels.cuh"
#include "../../ntcomp_b200/csrc/container.h"
#include "../../ntcomp_b200/csrc/decode_kernels.cuh"
#include "../../ntcomp_b200/csrc/encode_kernels.cuh"
#include "../../ntcomp_b200/csrc/host_index.h"

using namespace ntc;

struct EmulIndex {
    uint32_t tile_len = 1u << 22;
    std::vector<uint32_t> rank;
    std::vector<uint64_t> jump;
    std::vector<uint32_t> lcs;
    std::vector<uint32_t> pfx;
    DevIndex dix;
};

extern "C" {

void* emul_index_create(const ntc_index* idx) {
    EmulIndex* e = new EmulIndex();
    build_rank_blocks(*idx, &e->rank, 2);
    build_jump_table(*idx, &e->jump, 2);
    build_lcs_words(*idx, &e->lcs);
    e->dix.rank = (const uint4*)e->rank.data();
    e->dix.lcsx = e->lcs.data();
    e->dix.jump = e->jump.data();
    e->dix.pfx = nullptr;
    if (idx->k >= 10) {                                              // as ntc_ctx_create
        build_prefix_table(e->rank, idx->n, &e->pfx, e->dix.c_start);
        e->dix.pfx = (const uint4*)e->pfx.data();
    }
    e->dix.n = (uint32_t)idx->n;
    e->dix.k = idx->k;
    spec_window(idx->n, &e->dix.spec_lo, &e->dix.spec_span);
    return e;
}
void emul_index_free(void* p) { delete (EmulIndex*)p; }
void emul_set_tile_len(void* p, uint32_t T) { ((EmulIndex*)p)->tile_len = T ? (T + 3u) & ~3u : 1u << 22; }

static void run_ms(const EmulIndex* e, const uint8_t* bases, const uint64_t* offsets, uint64_t n_reads, std::vector<uint64_t>* packed,
                   std::vector<uint64_t>* scratch, uint64_t** scratch_aligned, uint32_t* counts, uint32_t* status, bool keep_records = false) {
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
    for (uint64_t i = tile_off[n_reads]; i-- > 0;) ms_thread(e->dix, packed->data(), offsets, desc[i], T, sc, counts, status, stage, 1, keep_records);
    if (counts) for (uint64_t r = 0; r < n_reads; r++) count_thread(e->dix, packed->data(), offsets, n_reads, T, sc, counts, status, keep_records, r);
}

// mode 0 records, 1 triples.  Returns the accumulated status bits.
uint32_t emul_encode(void* ep, const uint8_t* bases, const uint64_t* offsets, uint64_t n_reads, int mode, uint64_t* out, uint64_t cap,
                     uint64_t* n_records, uint64_t* rec_offsets) {
    const EmulIndex* e = (const EmulIndex*)ep;
    uint32_t status = 0;
    std::vector<uint64_t> packed, scratch;
    std::vector<uint32_t> counts(n_reads + 1, 0);
    uint64_t* sc = nullptr;
    run_ms(e, bases, offsets, n_reads, &packed, &scratch, &sc, counts.data(), &status, mode == 0);
    std::vector<uint64_t> roff(n_reads + 1, 0);
    for (uint64_t r = 0; r < n_reads; r++) roff[r + 1] = roff[r] + counts[r];   // exclusive scan (scan.cuh on the GPU)
    for (uint64_t r = 0; r < n_reads; r++) {
        if (mode == 0) emit_thread<0>(e->dix, packed.data(), offsets, n_reads, sc, counts.data(), roff.data(), nullptr, out, cap, &status, e->tile_len, r);
        else emit_thread<1>(e->dix, packed.data(), offsets, n_reads, sc, counts.data(), roff.data(), nullptr, out, cap, &status, e->tile_len, r);
    }
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

// GPU column codec bodies (column_kernels.cuh) on the CPU: returns the four coded columns back to back
// in `words` (sizes in n_words[4], parameters in params[4]); -1 when a column is empty (Q4).
int emul_block_columns(const uint64_t* rec, uint64_t n, uint64_t* words, uint64_t cap, uint64_t* n_words, uint64_t* params) {
    std::vector<uint64_t> ref_idx(n + 1), lit_off(n + 1), colex(n), len(n), flags(n);
    std::vector<unsigned long long> lits(n * 32 / 31 + 2, 0), red(3, 0);
    SplitLoad load{rec};
    uint64_t a = 0, b = 0;
    for (uint64_t i = 0; i < n; i++) { ref_idx[i] = a; lit_off[i] = b; Pair64 p = load(i); a += p.a; b += p.b; }
    for (uint64_t i = 0; i < n; i++) split_thread(rec, n, ref_idx.data(), lit_off.data(), colex.data(), len.data(), flags.data(), lits.data(), red.data(), i);
    const uint64_t n_ref = a, n_lw = (b + 30) / 31;
    if (n_ref == 0 || n_lw == 0) return -1;
    uint64_t max_lit = 0;
    for (uint64_t i = 0; i < n_lw; i++) max_lit = std::max<uint64_t>(max_lit, lits[i]);
    const uint64_t* vals[4] = {colex.data(), len.data(), flags.data(), (const uint64_t*)lits.data()};
    const uint64_t cnt[4] = {n_ref, n_ref, n, n_lw};
    CodeParams cp[4];
    unsigned l; uint64_t limit;
    minimal_binary_shape(red[0] + 2, &l, &limit); cp[0] = CodeParams{0, 0, l, limit}; params[0] = red[0] + 2;
    params[1] = rice_parameter(n_ref, red[1]); cp[1] = CodeParams{1, (uint32_t)params[1], 0, 0};
    params[2] = rice_parameter(n, red[2]); cp[2] = CodeParams{1, (uint32_t)params[2], 0, 0};
    minimal_binary_shape(max_lit + 2, &l, &limit); cp[3] = CodeParams{0, 0, l, limit}; params[3] = max_lit + 2;
    uint64_t pos = 0;
    for (int q = 0; q < 4; q++) {
        std::vector<uint64_t> off(cnt[q] + 1);
        uint64_t bits = 0;
        for (uint64_t i = 0; i < cnt[q]; i++) { off[i] = bits; bits += code_bits(cp[q], vals[q][i]); }
        std::vector<unsigned long long> w((bits + 63) / 64 + 2, 0);
        for (uint64_t i = 0; i < cnt[q]; i++) pack_thread(cp[q], vals[q], cnt[q], off.data(), w.data(), i);
        n_words[q] = (bits + 63) / 64;
        if (pos + n_words[q] > cap) return -2;
        for (uint64_t i = 0; i < n_words[q]; i++) words[pos + i] = __builtin_bswap64(w[i]);
        pos += n_words[q];
    }
    return 0;
}


// GPU column decoder bodies (column_decode_kernels.cuh) on the CPU: words as stored (big-endian u64), n values
// expected; returns the number of codes found (values beyond n are not stored).  -1: bad parameter.
int64_t emul_column_decode(const uint64_t* words_be, uint64_t n_words, uint64_t n, int rice, uint64_t param, uint64_t* out) {
    CodeParams p;
    if (rice) { if (param > 63) return -1; p = CodeParams{1, (uint32_t)param, 0, 0}; }
    else { if (param < 2) return -1; unsigned l; uint64_t limit; minimal_binary_shape(param, &l, &limit); p = CodeParams{0, 0, l, limit}; }
    if (n_words == 0) return 0;
    std::vector<uint64_t> w(n_words);
    for (uint64_t i = 0; i < n_words; i++) w[i] = __builtin_bswap64(words_be[i]);      // bswap_kernel
    const uint64_t n_chunks = (n_words * 64 + kCwChunkBits - 1) / kCwChunkBits;
    const uint32_t ns = cw_num_states(p);
    std::vector<uint32_t> table(n_chunks * ns), entry(n_chunks);
    std::vector<uint64_t> first(n_chunks);
    for (uint64_t i = n_chunks * ns; i-- > 0;) cw_table_thread(p, w.data(), n_words, n_chunks, table.data(), i);
    uint64_t total = 0;
    cw_resolve_serial(table.data(), ns, n_chunks, entry.data(), first.data(), &total);
    for (uint64_t c = n_chunks; c-- > 0;) cw_extract_thread(p, w.data(), n_words, n_chunks, entry.data(), first.data(), out, n, c);
    return (int64_t)total;
}

// zip_thread over a block's four columns; returns the error bits (0 = ok)
uint32_t emul_zip(const uint64_t* colex, const uint64_t* len, uint64_t n_ref, const uint64_t* flags, uint64_t n, const uint64_t* lits,
                  uint64_t n_lit_words, int full_last_word, uint64_t* rec) {
    std::vector<uint64_t> ref_idx(n + 1), lit_off(n + 1);
    FlagLoad load{flags};
    uint64_t a = 0, b = 0;
    for (uint64_t i = 0; i < n; i++) { ref_idx[i] = a; lit_off[i] = b; Pair64 pr = load(i); a += pr.a; b += pr.b; }
    uint64_t totals[2] = {a, b};
    uint32_t err = 0;
    for (uint64_t i = 0; i < n; i++)
        zip_thread(colex, len, n_ref, flags, n, lits, n_lit_words, ref_idx.data(), lit_off.data(), totals, full_last_word, rec, &err, i);
    return err;
}

}  // extern "C"
