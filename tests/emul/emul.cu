// tests/emul/emul.cu -- TEST INFRASTRUCTURE: runs the product's per-thread kernel bodies
// (ntcomp_b200/csrc/*_kernels.cuh, NTC_HD functions) on the CPU, one loop iteration per CUDA thread,
// so kernel LOGIC is checked against the oracle in the "not gpu" suite before GPU time is spent.
// Not part of the product: libntcomp_b200.so never links or calls this file.
#include <cstring>
#include <vector>

#include "../../ntcomp_b200/csrc/decode_kernels.cuh"
#include "../../ntcomp_b200/csrc/encode_kernels.cuh"
#include "../../ntcomp_b200/csrc/host_index.h"

using namespace ntc;

struct EmulIndex {
    std::vector<uint32_t> rank;
    std::vector<uint64_t> jump;
    std::vector<uint32_t> lcs;
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
    e->dix.n = (uint32_t)idx->n;
    e->dix.k = idx->k;
    spec_window(idx->n, &e->dix.spec_lo, &e->dix.spec_span);
    return e;
}
void emul_index_free(void* p) { delete (EmulIndex*)p; }

static void run_ms(const EmulIndex* e, const uint8_t* bases, const uint64_t* offsets, uint64_t n_reads, std::vector<uint64_t>* packed,
                   std::vector<uint64_t>* scratch, uint64_t** scratch_aligned, uint32_t* counts, uint32_t* status) {
    uint64_t nb = offsets[n_reads] - offsets[0];
    packed->assign(nb / 32 + 6, 0);
    scratch->assign(scratch_entries(nb, n_reads) + 4, 0);
    for (uint64_t w = 0; w < (nb + 31) / 32; w++) pack2bit_thread(bases, offsets, nb, packed->data(), status, w);
    uint64_t* sc = scratch->data();
    sc += (4 - (((uintptr_t)sc / 8) & 3)) & 3;                       // 32-byte aligned, like cudaMalloc
    *scratch_aligned = sc;
    uint64_t stage[4];
    for (uint64_t r = 0; r < n_reads; r++) ms_thread(e->dix, packed->data(), offsets, n_reads, sc, counts, status, r, stage, 1);
}

// mode 0 records, 1 triples.  Returns the accumulated status bits.
uint32_t emul_encode(void* ep, const uint8_t* bases, const uint64_t* offsets, uint64_t n_reads, int mode, uint64_t* out, uint64_t cap,
                     uint64_t* n_records, uint64_t* rec_offsets) {
    const EmulIndex* e = (const EmulIndex*)ep;
    uint32_t status = 0;
    std::vector<uint64_t> packed, scratch;
    std::vector<uint32_t> counts(n_reads + 1, 0);
    uint64_t* sc = nullptr;
    run_ms(e, bases, offsets, n_reads, &packed, &scratch, &sc, counts.data(), &status);
    std::vector<uint64_t> roff(n_reads + 1, 0);
    for (uint64_t r = 0; r < n_reads; r++) roff[r + 1] = roff[r] + counts[r];   // exclusive scan (scan.cuh on the GPU)
    for (uint64_t r = 0; r < n_reads; r++) {
        if (mode == 0) emit_thread<0>(e->dix, packed.data(), offsets, n_reads, sc, counts.data(), roff.data(), nullptr, out, cap, &status, r);
        else emit_thread<1>(e->dix, packed.data(), offsets, n_reads, sc, counts.data(), roff.data(), nullptr, out, cap, &status, r);
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
    for (uint64_t i = 0; i < n_rec; i++) { base_off[i] = a; seq_idx[i] = b; Pair64 p = load(i); a += p.a; b += p.b; }
    base_off[n_rec] = a; seq_idx[n_rec] = b;
    uint64_t totals[2] = {a, b};
    for (uint64_t i = 0; i < n_rec; i++) seq_starts_thread(rec, n_rec, seq_idx.data(), seq_first.data(), i);
    for (uint64_t j = 0; j <= n_rec; j++)
        seq_offsets_thread(base_off.data(), seq_first.data(), totals, n_rec, seq_offsets, cap_seqs, first_no, fasta, cap, counts, &status, j);
    for (uint64_t i = 0; i < n_rec; i++)
        walk_thread(e->dix, rec, n_rec, base_off.data(), seq_idx.data(), seq_first.data(), totals, first_no, fasta, out, cap, counts, &status, i);
    return status;
}

}  // extern "C"
