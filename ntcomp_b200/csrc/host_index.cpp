// host_index.cpp -- flat index bookkeeping, own file format, and the HBM layouts derived from the
// canonical arrays (rank blocks with inline counts + Q bits; 16-step decode jump table).
#include "host_index.h"

#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <thread>

#include "../../include/ntcomp_b200.h"

namespace ntc {

static thread_local std::string g_error;
void set_error(const std::string& msg) { g_error = msg; }
const char* last_error() { return g_error.c_str(); }

namespace {
template <typename F>
void parallel_chunks(uint64_t n, uint64_t chunk, int n_threads, F f) {  // f(lo, hi)
    uint64_t n_chunks = (n + chunk - 1) / chunk;
    std::atomic<uint64_t> next(0);
    auto work = [&]() {
        for (;;) {
            uint64_t c = next.fetch_add(1);
            if (c >= n_chunks) break;
            f(c * chunk, std::min(n, (c + 1) * chunk));
        }
    };
    if (n_threads < 1) n_threads = (int)std::max(1u, std::thread::hardware_concurrency());
    std::vector<std::thread> pool;
    for (int t = 1; t < n_threads; t++) pool.emplace_back(work);
    work();
    for (auto& t : pool) t.join();
}
inline bool row_bit(const ntc_index& idx, int c, uint64_t i) { return (idx.rows[c][i >> 6] >> (i & 63)) & 1; }
}  // namespace

int32_t finalize_index(ntc_index* idx) {
    if (idx->k < kMinK || idx->k > kMaxK) { set_error("k must be in [12,255]"); return NTC_ERR_UNSUPPORTED; }
    if (idx->n == 0 || idx->n >= (1ull << 32)) { set_error("n_sets must be in [1, 2^32)"); return NTC_ERR_UNSUPPORTED; }
    const uint64_t words = (idx->n + 63) / 64;
    idx->C[0] = 1;
    for (int c = 0; c < 4; c++) {
        if (idx->rows[c].size() < words) { set_error("row too short"); return NTC_ERR_INVALID_ARG; }
        idx->rows[c].resize(words);
        if (idx->n & 63) idx->rows[c][words - 1] &= (1ull << (idx->n & 63)) - 1;
        uint64_t ones = 0;
        for (uint64_t w = 0; w < words; w++) ones += (uint64_t)__builtin_popcountll(idx->rows[c][w]);
        idx->C[c + 1] = idx->C[c] + ones;
    }
    if (idx->C[4] != idx->n) { set_error("subset rows must hold exactly n_sets-1 edges"); return NTC_ERR_INVALID_ARG; }
    if (idx->lcs.size() != idx->n) { set_error("lcs must have n_sets entries"); return NTC_ERR_INVALID_ARG; }
    return NTC_OK;
}

uint32_t rank_nodes_for(uint64_t n_sets) {
    if (const char* env = getenv("NTC_RANK_NODES")) { const int v = atoi(env); if (v == 32 || v == 64) return (uint32_t)v; }
    return n_sets >= NTC_RANK64_MIN_NODES ? 64u : 32u;
}

void build_prefix_table(const std::vector<uint32_t>& rb, uint64_t n, std::vector<uint32_t>* out, uint32_t c_start[4], uint32_t kBlockNodes) {
    // the same arithmetic as the device's extend_right: entry count + popcount of the row bits below the offset
    auto rank = [&](uint32_t c, uint64_t i) -> uint32_t {
        const uint32_t off = (uint32_t)(i % kBlockNodes);
        if (kBlockNodes == 32) {
            const uint32_t* e = rb.data() + (i / kBlockNodes) * 16 + c * 4;
            return e[0] + (uint32_t)__builtin_popcount(e[1] & ((1u << off) - 1u));
        }
        const uint32_t* e = rb.data() + (i / kBlockNodes) * 16 + (c >> 1) * 8;      // the sector of bases {A,C} or {G,T}
        uint64_t bits;
        memcpy(&bits, e + 2 + 2 * (c & 1), 8);
        return e[c & 1] + (uint32_t)__builtin_popcountll(bits & ((1ull << off) - 1ull));
    };
    const uint32_t n_entries = 1u << (2 * kPrefixBases);
    out->assign((size_t)n_entries * 8, 0);
    for (uint32_t c = 0; c < 4; c++) c_start[c] = rank(c, 0);
    for (uint32_t x = 0; x < n_entries; x++) {
        uint32_t s = 0, e = (uint32_t)n, w[8];
        bool ok = true;
        for (uint32_t l = 1; l <= kPrefixBases && ok; l++) {
            const uint32_t c = (x >> (2 * (l - 1))) & 3u;
            s = rank(c, s);
            e = rank(c, e);
            ok = s < e;
            if (l >= 2) w[l - 2] = s;
        }
        if (!ok) continue;
        w[7] = e;
        for (int i = 0; i < 8; i++) (*out)[(size_t)x * 8 + i] = w[i];
    }
}

void build_rank_blocks(const ntc_index& idx, std::vector<uint32_t>* out, int n_threads, uint32_t kBlockNodes) {
    const uint64_t n = idx.n;
    const uint32_t k = idx.k;
    const uint64_t n_blocks = n / kBlockNodes + 1;  // position n (interval end) must be addressable
    out->assign(n_blocks * 16, 0);
    // running counts at block starts: one sequential popcount pass per row
    std::vector<uint32_t> cum[4];
    for (int c = 0; c < 4; c++) {
        cum[c].resize(idx.rows[c].size() + 1);
        uint32_t acc = 0;
        for (size_t w = 0; w < idx.rows[c].size(); w++) { cum[c][w] = acc; acc += (uint32_t)__builtin_popcountll(idx.rows[c][w]); }
        cum[c][idx.rows[c].size()] = acc;
    }
    auto rank = [&](int c, uint64_t i) -> uint32_t {
        if (i >= n) return cum[c][idx.rows[c].size()];
        uint64_t w = i >> 6, b = i & 63;
        uint32_t r = cum[c][w];
        if (b) r += (uint32_t)__builtin_popcountll(idx.rows[c][w] & ((1ull << b) - 1));
        return r;
    };
    auto qbit = [&](uint64_t v) -> uint64_t {
        if (v >= n) return 0;
        bool lead = idx.lcs[v] < k - 1;
        bool next_lead = (v + 1 == n) || idx.lcs[v + 1] < k - 1;
        return (lead && next_lead) ? 1 : 0;
    };
    parallel_chunks(n_blocks, 4096, n_threads, [&](uint64_t lo, uint64_t hi) {
        for (uint64_t b = lo; b < hi; b++) {
            uint64_t base = b * kBlockNodes;
            uint64_t q = 0;
            for (uint32_t i = 0; i < kBlockNodes; i++) q |= qbit(base + i) << i;
            for (int c = 0; c < 4; c++) {
                uint64_t bits = 0;
                for (uint32_t i = 0; i < kBlockNodes && base + i < n; i++) bits |= (uint64_t)row_bit(idx, c, base + i) << i;
                const uint32_t cnt = (uint32_t)idx.C[c] + rank(c, base);  // C[c] folded in: extend_right is entry count + popcount
                if (kBlockNodes == 32) {
                    uint32_t* e = out->data() + b * 16 + c * 4;
                    e[0] = cnt;
                    e[1] = (uint32_t)bits;
                    e[2] = (uint32_t)q;
                    e[3] = 0;
                } else {
                    uint32_t* e = out->data() + b * 16 + (c >> 1) * 8;    // {count, count, bits, bits, Q} per pair of bases (layout.h)
                    e[c & 1] = cnt;
                    memcpy(e + 2 + 2 * (c & 1), &bits, 8);
                    memcpy(e + 6, &q, 8);
                }
            }
        }
    });
}

void build_lcs_words(const ntc_index& idx, std::vector<ntc_lcsx_t>* out) {
    const uint64_t n = idx.n;
    std::vector<uint8_t> dl(n + 1, 0), dr(n + 1, 0);
    std::vector<uint32_t> stack;  // indices with strictly increasing LCS
    for (uint64_t i = 0; i < n; i++) {
        while (!stack.empty() && idx.lcs[stack.back()] >= idx.lcs[i]) stack.pop_back();
        uint64_t psv = stack.empty() ? 0 : stack.back();
        dl[i] = idx.lcs[i] ? (uint8_t)std::min<uint64_t>(255, i - psv) : 0;
        stack.push_back((uint32_t)i);
    }
    stack.clear();
    for (uint64_t i = n; i-- > 0;) {
        while (!stack.empty() && idx.lcs[stack.back()] >= idx.lcs[i]) stack.pop_back();
        uint64_t nsv = stack.empty() ? n : stack.back();  // LCS[n] = 0 sentinel is smaller than any LCS >= 1
        dr[i] = idx.lcs[i] ? (uint8_t)std::min<uint64_t>(255, nsv - i) : 0;
        stack.push_back((uint32_t)i);
    }
    out->assign(n + 1, 0);
    for (uint64_t i = 0; i <= n; i++) {
        uint32_t lo = i < n ? ((uint32_t)idx.lcs[i] | ((uint32_t)dl[i] << 8)) : 0u;
        uint32_t hi = i + 1 < n ? ((uint32_t)idx.lcs[i + 1] | ((uint32_t)dr[i + 1] << 8)) : 0u;
        (*out)[i] = lo | (hi << 16);
    }
}

// window of match lengths in which extensions of chance matches keep failing: around log4(n_sets)
void spec_window(uint64_t n_sets, uint32_t* lo, uint32_t* span) {
    double l4 = std::log((double)std::max<uint64_t>(n_sets, 4)) / std::log(4.0);
    // swept on B200 (tools/sweep_spec.py, C3 shape, n = 10^7, log4 n = 11.6; profiles/r1/sweep_spec_v12.txt): with all
    // loads of an iteration issued together (v12) the window hardly matters any more -- [11,14) 4.77 ms, [10,16) 4.91,
    // no speculation at all 4.91, always 7.62 -- so it is kept narrow
    int a = (int)std::floor(l4);
    *lo = (uint32_t)std::max(1, a);
    *span = 3;
}

void build_jump_table(const ntc_index& idx, std::vector<uint64_t>* out, int n_threads) {
    const uint64_t n = idx.n;
    // one inverse-LF step: pred[C[c] + j] = position of the j-th one of row c; in-label = c
    std::vector<uint32_t> pred(n), tmp_p(n);
    std::vector<uint32_t> chars(n), tmp_c(n);
    pred[0] = 0;
    chars[0] = 0;
    for (int c = 0; c < 4; c++) {
        uint64_t j = idx.C[c];
        for (size_t w = 0; w < idx.rows[c].size(); w++) {
            uint64_t x = idx.rows[c][w];
            while (x) {
                int b = __builtin_ctzll(x);
                pred[j] = (uint32_t)(w * 64 + b);
                chars[j] = (uint32_t)c;
                j++;
                x &= x - 1;
            }
        }
    }
    // pointer doubling: after the round with step t, pred = pred^(2t) and chars holds the last 2t
    // bases with the LAST base in the TOP 2 bits of a 2*(2t)-bit field.
    for (int t = 1; t < 16; t *= 2) {
        parallel_chunks(n, 1 << 16, n_threads, [&](uint64_t lo, uint64_t hi) {
            for (uint64_t v = lo; v < hi; v++) {
                uint32_t u = pred[v];
                tmp_p[v] = pred[u];
                tmp_c[v] = (chars[v] << (2 * t)) | chars[u];  // earlier bases (from u) go below
            }
        });
        pred.swap(tmp_p);
        chars.swap(tmp_c);
    }
    out->resize(n);
    parallel_chunks(n, 1 << 16, n_threads, [&](uint64_t lo, uint64_t hi) {
        for (uint64_t v = lo; v < hi; v++) (*out)[v] = (uint64_t)pred[v] | ((uint64_t)chars[v] << 32);
    });
}

}  // namespace ntc

// ---------------------------------------------------------------------------------------------
// C ABI: index management
// ---------------------------------------------------------------------------------------------
using namespace ntc;

extern "C" {

const char* ntc_last_error(void) { return ntc::last_error(); }
const char* ntc_version(void) { return "ntcomp_b200 0.2 (sm_100a)"; }

int32_t ntc_index_build(const uint8_t* const* seqs, const uint64_t* lens, uint64_t n_seqs, uint32_t k, int32_t add_revcomp,
                        int32_t n_threads, ntc_index** out) {
    if (!out || (n_seqs && (!seqs || !lens))) { set_error("null argument"); return NTC_ERR_INVALID_ARG; }
    *out = nullptr;
    try {
        return build_index(seqs, lens, n_seqs, k, add_revcomp != 0, n_threads, out);
    } catch (const std::exception& e) {
        set_error(std::string("index build failed: ") + e.what());
        return NTC_ERR_INVALID_ARG;
    }
}

int32_t ntc_index_from_arrays(uint32_t k, uint64_t n_sets, const uint64_t* row_a, const uint64_t* row_c, const uint64_t* row_g,
                              const uint64_t* row_t, const uint8_t* lcs, ntc_index** out) {
    if (!out || !row_a || !row_c || !row_g || !row_t || !lcs) { set_error("null argument"); return NTC_ERR_INVALID_ARG; }
    *out = nullptr;
    try {
        ntc_index* idx = new ntc_index();
        idx->k = k;
        idx->n = n_sets;
        const uint64_t words = (n_sets + 63) / 64;
        const uint64_t* rows[4] = {row_a, row_c, row_g, row_t};
        for (int c = 0; c < 4; c++) idx->rows[c].assign(rows[c], rows[c] + words);
        idx->lcs.assign(lcs, lcs + n_sets);
        int32_t rc = finalize_index(idx);
        if (rc) { delete idx; return rc; }
        *out = idx;
        return NTC_OK;
    } catch (const std::exception& e) {
        set_error(e.what());
        return NTC_ERR_INVALID_ARG;
    }
}

static const char kMagic[8] = {'N', 'T', 'C', 'B', '2', '0', '0', '1'};

int32_t ntc_index_save(const ntc_index* idx, const char* path) {
    if (!idx || !path) { set_error("null argument"); return NTC_ERR_INVALID_ARG; }
    FILE* f = fopen(path, "wb");
    if (!f) { set_error(std::string("cannot open for writing: ") + path); return NTC_ERR_IO; }
    uint64_t hdr[2] = {idx->k, idx->n};
    bool ok = fwrite(kMagic, 1, 8, f) == 8 && fwrite(hdr, 8, 2, f) == 2;
    const uint64_t words = (idx->n + 63) / 64;
    for (int c = 0; c < 4 && ok; c++) ok = fwrite(idx->rows[c].data(), 8, words, f) == words;
    ok = ok && fwrite(idx->lcs.data(), 1, idx->n, f) == idx->n;
    ok = (fclose(f) == 0) && ok;
    if (!ok) { set_error("short write"); return NTC_ERR_IO; }
    return NTC_OK;
}

int32_t ntc_index_load(const char* path, ntc_index** out) {
    if (!out || !path) { set_error("null argument"); return NTC_ERR_INVALID_ARG; }
    *out = nullptr;
    FILE* f = fopen(path, "rb");
    if (!f) { set_error(std::string("cannot open: ") + path); return NTC_ERR_IO; }
    char magic[8];
    uint64_t hdr[2];
    if (fread(magic, 1, 8, f) != 8 || memcmp(magic, kMagic, 8) != 0 || fread(hdr, 8, 2, f) != 2 || hdr[1] == 0 ||
        hdr[1] >= (1ull << 32)) {
        fclose(f);
        set_error("not an ntcomp_b200 index file");
        return NTC_ERR_IO;
    }
    ntc_index* idx = new ntc_index();
    idx->k = (uint32_t)hdr[0];
    idx->n = hdr[1];
    const uint64_t words = (idx->n + 63) / 64;
    bool ok = true;
    for (int c = 0; c < 4 && ok; c++) {
        idx->rows[c].resize(words);
        ok = fread(idx->rows[c].data(), 8, words, f) == words;
    }
    idx->lcs.resize(idx->n);
    ok = ok && fread(idx->lcs.data(), 1, idx->n, f) == idx->n;
    fclose(f);
    if (!ok) { delete idx; set_error("truncated index file"); return NTC_ERR_IO; }
    int32_t rc = finalize_index(idx);
    if (rc) { delete idx; return rc; }
    *out = idx;
    return NTC_OK;
}

uint32_t ntc_index_k(const ntc_index* idx) { return idx ? idx->k : 0; }
uint64_t ntc_index_n_sets(const ntc_index* idx) { return idx ? idx->n : 0; }
void ntc_index_get_C(const ntc_index* idx, uint64_t out[5]) { memcpy(out, idx->C, 5 * sizeof(uint64_t)); }
void ntc_index_get_row(const ntc_index* idx, int32_t c, uint64_t* out) {
    memcpy(out, idx->rows[c].data(), ((idx->n + 63) / 64) * 8);
}
void ntc_index_get_lcs(const ntc_index* idx, uint8_t* out) { memcpy(out, idx->lcs.data(), idx->n); }
void ntc_index_free(ntc_index* idx) { delete idx; }

uint64_t ntc_encode_bound(uint64_t n_bases) { return n_bases; }

}  // extern "C"
