// tests/emul/asan_driver.cu -- TEST INFRASTRUCTURE: runs the kernel bodies (via emul.cu) over random
// inputs in a stand-alone executable built with AddressSanitizer, so out-of-bounds reads/writes in the
// per-thread code show up on the CPU (compute-sanitizer is closed on the GPU pool).  Buffers are sized
// exactly like the GPU workspaces.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>

#include "../../include/ntcomp_b200.h"

extern "C" {
void* emul_index_create(const ntc_index* idx);
void emul_index_free(void* p);
void emul_set_tile_len(void* p, uint32_t T);
uint32_t emul_encode(void* ep, const uint8_t* bases, const uint64_t* offsets, uint64_t n_reads, int mode, uint64_t* out, uint64_t cap,
                     uint64_t* n_records, uint64_t* rec_offsets);
uint32_t emul_decode(void* ep, const uint64_t* rec, uint64_t n_rec, uint64_t first_no, int fasta, uint8_t* out, uint64_t cap,
                     uint64_t* seq_offsets, uint64_t cap_seqs, uint64_t* counts);
int emul_batch_columns(const uint64_t* rec, const uint64_t* blk_rec_in, uint32_t n_blocks, uint64_t* words, uint64_t cap, uint64_t* seg_out,
                       int32_t* status_out, uint64_t* total_out);
int emul_columns_decode(const uint64_t* cols_in, uint32_t n_blocks, const uint64_t* words_be, uint64_t n_words, int full_last_word, int trim,
                        uint64_t* rec, uint64_t cap, uint64_t* n_rec_out, uint32_t* detail);
}

extern "C" int32_t emul_index_build(const uint8_t* const* seqs, const uint64_t* lens, uint64_t n_seqs, uint32_t k, int32_t add_revcomp,
                                    ntc_index** out, uint64_t* stats);

// the GPU index construction (build_pipeline.h) with exactly-sized work arrays: many short sequences, separators, k up to 255
static int check_builder(std::mt19937_64& rng) {
    const char* alpha = "ACGTN";
    for (uint32_t k : {12u, 31u, 33u, 91u, 255u}) {
        std::vector<std::vector<uint8_t>> store;
        for (int i = 0; i < 12; i++) {
            std::vector<uint8_t> s(rng() % (3 * k + 40));
            for (auto& b : s) b = (uint8_t)alpha[rng() % (i % 4 == 0 ? 5 : 4)];
            store.push_back(s);
        }
        store.push_back(store[1]);
        std::vector<const uint8_t*> ptrs;
        std::vector<uint64_t> lens;
        for (auto& s : store) { ptrs.push_back(s.data()); lens.push_back(s.size()); }
        for (int rc = 0; rc < 2; rc++) {
            ntc_index *a = nullptr, *b = nullptr;
            if (emul_index_build(ptrs.data(), lens.data(), ptrs.size(), k, rc, &a, nullptr)) return 20;
            if (ntc_index_build(ptrs.data(), lens.data(), ptrs.size(), k, rc, 2, &b)) return 21;
            const uint64_t n = ntc_index_n_sets(a);
            if (n != ntc_index_n_sets(b)) return 22;
            std::vector<uint8_t> la(n), lb(n);
            ntc_index_get_lcs(a, la.data());
            ntc_index_get_lcs(b, lb.data());
            if (la != lb) return 23;
            std::vector<uint64_t> ra((n + 63) / 64), rb((n + 63) / 64);
            for (int c = 0; c < 4; c++) {
                ntc_index_get_row(a, c, ra.data());
                ntc_index_get_row(b, c, rb.data());
                if (ra != rb) return 24;
            }
            ntc_index_free(a);
            ntc_index_free(b);
        }
    }
    return 0;
}

int main() {
    std::mt19937_64 rng(7);
    const char* alpha = "ACGT";
    if (int rc = check_builder(rng)) { fprintf(stderr, "index construction check failed: %d\n", rc); return rc; }
    for (int round = 0; round < 6; round++) {
        const uint32_t k = round % 3 == 0 ? 12 : (round % 3 == 1 ? 31 : 91);
        std::vector<uint8_t> g(3000 + 2000 * round);
        for (auto& b : g) b = (uint8_t)alpha[rng() & 3];
        const uint8_t* seqs[1] = {g.data()};
        uint64_t lens[1] = {g.size()};
        ntc_index* idx = nullptr;
        if (ntc_index_build(seqs, lens, 1, k, 1, 2, &idx)) { fprintf(stderr, "build failed: %s\n", ntc_last_error()); return 2; }
        void* e = emul_index_create(idx);
        if (round & 1) emul_set_tile_len(e, 256);                     // reads of up to 700 bases become several tiles: the segment path (seg_*_thread)
        std::vector<uint8_t> bases;
        std::vector<uint64_t> off(1, 0);
        for (int r = 0; r < 300; r++) {
            uint64_t L = 1 + rng() % (r % 7 == 0 ? 700 : 160);
            L = std::min<uint64_t>(L, g.size());
            uint64_t a = rng() % (g.size() - L + 1);
            for (uint64_t j = 0; j < L; j++) {
                uint8_t b = g[a + j];
                if (rng() % 100 < (uint64_t)(round * 3)) b = (uint8_t)alpha[rng() & 3];
                bases.push_back(b);
            }
            off.push_back(bases.size());
        }
        const uint64_t n_reads = off.size() - 1, nb = bases.size();
        std::vector<uint64_t> rec(nb), roff(n_reads + 1), tri(3 * nb);
        uint64_t n_rec = 0, n_tri = 0;
        if (emul_encode(e, bases.data(), off.data(), n_reads, 0, rec.data(), nb, &n_rec, roff.data())) return 3;
        if (emul_encode(e, bases.data(), off.data(), n_reads, 1, tri.data(), nb, &n_tri, nullptr)) return 4;
        rec.resize(n_rec);   // exact-size buffers: one element past the end is poisoned
        for (uint64_t first_no : {1ull, 99ull, 99999ull}) {
            uint64_t nbases = 0, nseqs = 0, fbytes = 0;
            ntc_decode_bound(rec.data(), n_rec, first_no, &nbases, &nseqs, &fbytes);
            std::vector<uint8_t> out(nbases), text(fbytes);
            std::vector<uint64_t> soff(nseqs + 1);
            uint64_t counts[2];
            if (emul_decode(e, rec.data(), n_rec, 1, 0, out.data(), nbases, soff.data(), nseqs, counts)) return 5;
            if (counts[1] != nb || memcmp(out.data(), bases.data(), nb)) { fprintf(stderr, "round trip mismatch\n"); return 6; }
            if (emul_decode(e, rec.data(), n_rec, first_no, 1, text.data(), fbytes, nullptr, 0, counts)) return 7;
            if (counts[1] != fbytes) return 8;
        }
        // column codec, both directions, over a batch of three blocks cut at read boundaries
        const uint64_t cut[4] = {0, roff[n_reads / 3], roff[2 * n_reads / 3], n_rec};
        std::vector<uint64_t> words(4 * n_rec + 64), seg(48), cols(48);
        int32_t status[3];
        uint64_t total = 0;
        if (emul_batch_columns(rec.data(), cut, 3, words.data(), words.size(), seg.data(), status, &total)) return 9;
        words.resize(total);                                          // exact size again
        bool all_ok = true;
        for (int b = 0; b < 3; b++) all_ok = all_ok && status[b] == 0;
        if (all_ok) {
            std::vector<uint64_t> back(n_rec);
            uint64_t n_back = 0;
            uint32_t detail = 0;
            int rc = emul_columns_decode(seg.data(), 3, words.data(), total, 1, 0, back.data(), n_rec, &n_back, &detail);
            if (rc || n_back != n_rec || memcmp(back.data(), rec.data(), n_rec * 8)) { fprintf(stderr, "column round trip mismatch rc=%d\n", rc); return 10; }
        }
        emul_index_free(e);
        ntc_index_free(idx);
    }
    printf("ASAN_DRIVER_OK\n");
    return 0;
}
