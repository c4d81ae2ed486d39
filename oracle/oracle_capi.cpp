// oracle/oracle_capi.cpp -- TEST INFRASTRUCTURE ONLY.  PARITY UNPINNED (see oracle_index.h).
// Flat C entry points so tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
// reference legs can drive the oracle through ctypes.  Nothing in ntcomp_b200/ may link this.
#include <atomic>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "oracle_codec.h"
#include "oracle_container.h"
#include "oracle_index.h"

using namespace orc;

static thread_local std::string g_err;

extern "C" {

const char* orc_last_error() { return g_err.c_str(); }

void* orc_build(const uint8_t* bases, const uint64_t* offsets, uint64_t n_seqs, int k, int add_revcomp) {
    std::vector<const uint8_t*> ptrs(n_seqs);
    std::vector<uint64_t> lens(n_seqs);
    for (uint64_t i = 0; i < n_seqs; i++) { ptrs[i] = bases + offsets[i]; lens[i] = offsets[i + 1] - offsets[i]; }
    return build_index(ptrs.data(), lens.data(), n_seqs, k, add_revcomp != 0, &g_err);
}

void* orc_from_arrays(int k, uint64_t n, const uint64_t* ra, const uint64_t* rc, const uint64_t* rg, const uint64_t* rt,
                      const uint8_t* lcs) {
    const uint64_t* rows[4] = {ra, rc, rg, rt};
    return index_from_arrays(k, n, rows, lcs, &g_err);
}

void orc_free(void* h) { delete (Index*)h; }
uint64_t orc_n(const void* h) { return ((const Index*)h)->n; }
int orc_k(const void* h) { return ((const Index*)h)->k; }
void orc_get_C(const void* h, uint64_t* out) { memcpy(out, ((const Index*)h)->C, 5 * sizeof(uint64_t)); }
void orc_get_row(const void* h, int c, uint64_t* out) {
    const Index* ix = (const Index*)h;
    memcpy(out, ix->rows[c].data(), ((ix->n + 63) / 64) * 8);
}
void orc_get_lcs(const void* h, uint8_t* out) {
    const Index* ix = (const Index*)h;
    memcpy(out, ix->lcs.data(), ix->n);
}

void orc_ms(const void* h, const uint8_t* read, uint64_t len, uint32_t* d, uint64_t* s, uint64_t* e) {
    matching_statistics(*(const Index*)h, read, len, d, s, e);
}
int orc_search(const void* h, const uint8_t* pat, uint64_t len, uint64_t* out2) {
    Interval I;
    if (!search(*(const Index*)h, pat, len, &I)) return 0;
    out2[0] = I.s; out2[1] = I.e;
    return 1;
}
void orc_access_kmer(const void* h, uint64_t v, uint8_t* out) { access_kmer(*(const Index*)h, v, out); }

// mode 0 = literal restatement, 1 = fast equivalent.  out = cap triples (len,start,end).
int64_t orc_encode_sequence(const void* h, const uint8_t* read, uint64_t len, int mode, uint64_t* out, uint64_t cap) {
    std::vector<DictEntry> dict;
    int64_t rc = mode == 0 ? literal_encode_sequence(*(const Index*)h, read, len, &dict)
                           : fast_encode_sequence(*(const Index*)h, read, len, &dict);
    if (rc < 0) return rc;
    if (dict.size() > cap) return ERR_CAPACITY;
    for (size_t i = 0; i < dict.size(); i++) { out[3 * i] = dict[i].len; out[3 * i + 1] = dict[i].start; out[3 * i + 2] = dict[i].end; }
    return rc;
}

int64_t orc_encode_dictionary(const void* h, const uint64_t* triples, uint64_t n, uint64_t* out) {
    std::vector<DictEntry> dict(n);
    for (uint64_t i = 0; i < n; i++) dict[i] = DictEntry{triples[3 * i], triples[3 * i + 1], triples[3 * i + 2]};
    std::vector<uint64_t> words;
    int64_t rc = literal_encode_dictionary(*(const Index*)h, dict.data(), n, &words);
    if (rc < 0) return rc;
    memcpy(out, words.data(), words.size() * 8);
    return rc;
}

// flat_map(encode_dictionary . encode_sequence) over a batch of reads (src/main.rs:166,169).
// rec_offsets has n_reads+1 entries.  Returns total records, or the first (lowest read index) error.
int64_t orc_encode_reads(const void* h, const uint8_t* bases, const uint64_t* offsets, uint64_t n_reads, int mode,
                         uint64_t* records, uint64_t cap, uint64_t* rec_offsets, int threads) {
    const Index& ix = *(const Index*)h;
    if (threads < 1) threads = 1;
    std::vector<std::vector<uint64_t>> per(n_reads);
    std::vector<int64_t> status(n_reads, 0);
    std::atomic<uint64_t> next(0);
    auto work = [&]() {
        std::vector<DictEntry> dict;
        for (;;) {
            uint64_t r = next.fetch_add(1);
            if (r >= n_reads) break;
            const uint8_t* read = bases + offsets[r];
            uint64_t len = offsets[r + 1] - offsets[r];
            int64_t rc;
            if (mode == 0) {
                rc = literal_encode_sequence(ix, read, len, &dict);
                if (rc >= 0) rc = literal_encode_dictionary(ix, dict.data(), dict.size(), &per[r]);
            } else {
                rc = fast_encode_read(ix, read, len, &per[r]);
            }
            status[r] = rc;
        }
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < threads; t++) pool.emplace_back(work);
    work();
    for (auto& t : pool) t.join();
    uint64_t total = 0;
    for (uint64_t r = 0; r < n_reads; r++) {
        if (status[r] < 0) return status[r];
        if (rec_offsets) rec_offsets[r] = total;
        if (total + per[r].size() > cap) return ERR_CAPACITY;
        memcpy(records + total, per[r].data(), per[r].size() * 8);
        total += per[r].size();
    }
    if (rec_offsets) rec_offsets[n_reads] = total;
    return (int64_t)total;
}

// decode_sequence over one record stream.  seq_offsets gets n_seqs+1 entries (first = 0).
int64_t orc_decode(const void* h, const uint64_t* records, uint64_t n, int mode, uint8_t* bases, uint64_t cap,
                   uint64_t* seq_offsets, uint64_t cap_seqs) {
    std::vector<uint8_t> b;
    std::vector<uint64_t> off;
    off.push_back(0);
    int64_t rc = mode == 0 ? literal_decode_sequence(*(const Index*)h, records, n, &b, &off)
                           : fast_decode_sequence(*(const Index*)h, records, n, &b, &off);
    if (rc < 0) return rc;
    if (b.size() > cap || off.size() > cap_seqs + 1) return ERR_CAPACITY;
    memcpy(bases, b.data(), b.size());
    memcpy(seq_offsets, off.data(), off.size() * 8);
    return rc;
}

// ---- container ------------------------------------------------------------------------------
// codes one column: rice != 0 -> Rice, else minimal binary.  Returns number of words or -1.
int64_t orc_code_column(const uint64_t* ints, uint64_t n, int rice, uint64_t* words, uint64_t cap, uint64_t* param) {
    std::vector<uint64_t> v(ints, ints + n), w;
    if (!code_column(v, rice != 0, &w, param)) return -1;
    if (w.size() > cap) return ERR_CAPACITY;
    memcpy(words, w.data(), w.size() * 8);
    return (int64_t)w.size();
}
int64_t orc_decode_column(const uint64_t* words, uint64_t n_words, uint64_t n, int rice, uint64_t param, uint64_t* out) {
    std::vector<uint64_t> v;
    bool ok = rice ? rice_decode(words, n_words, n, param, &v) : minimal_binary_decode(words, n_words, n, param, &v);
    if (!ok) return -1;
    memcpy(out, v.data(), v.size() * 8);
    return (int64_t)v.size();
}
// split_encoded_dictionary: sizes[4] receive column lengths; each out buffer must hold n entries.
int64_t orc_split(const uint64_t* rec, uint64_t n, uint64_t* colex, uint64_t* len, uint64_t* flags, uint64_t* lits,
                  uint64_t* sizes) {
    Columns c;
    if (!split_encoded_dictionary(rec, n, &c)) return -1;
    memcpy(colex, c.colex.data(), c.colex.size() * 8);
    memcpy(len, c.len.data(), c.len.size() * 8);
    memcpy(flags, c.flags.data(), c.flags.size() * 8);
    memcpy(lits, c.literals.data(), c.literals.size() * 8);
    sizes[0] = c.colex.size(); sizes[1] = c.len.size(); sizes[2] = c.flags.size(); sizes[3] = c.literals.size();
    return 0;
}
// write_block_to: returns bytes written into out, -1 when the reference would return Err (Q4).
int64_t orc_write_block(const uint64_t* rec, uint64_t n, uint64_t num_records, uint8_t* out, uint64_t cap) {
    std::vector<uint8_t> sink;
    if (!write_block(rec, n, num_records, &sink)) return -1;
    if (sink.size() > cap) return ERR_CAPACITY;
    memcpy(out, sink.data(), sink.size());
    return (int64_t)sink.size();
}
// decode_block minus decode_sequence: returns bytes consumed (0 = EOF, <0 error); headers = 4 x
// {block_size,num_records,num_u64,encoded_size,rice_param,bitpacker_exponent}.
int64_t orc_read_block(const uint8_t* data, uint64_t size, uint64_t* rec, uint64_t cap, uint64_t* n_rec, uint64_t* headers) {
    std::vector<uint64_t> r;
    SubHeader h[4];
    int64_t used = read_block(data, size, &r, h, nullptr);
    if (used <= 0) return used;
    if (r.size() > cap) return ERR_CAPACITY;
    memcpy(rec, r.data(), r.size() * 8);
    *n_rec = r.size();
    if (headers) for (int q = 0; q < 4; q++) {
        headers[6 * q + 0] = h[q].block_size; headers[6 * q + 1] = h[q].num_records; headers[6 * q + 2] = h[q].num_u64;
        headers[6 * q + 3] = h[q].encoded_size; headers[6 * q + 4] = h[q].rice_param; headers[6 * q + 5] = h[q].bitpacker_exponent;
    }
    return used;
}

}  // extern "C"
