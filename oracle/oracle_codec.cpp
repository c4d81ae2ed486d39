// oracle/oracle_codec.cpp -- TEST INFRASTRUCTURE ONLY.  PARITY UNPINNED (see oracle_index.h).
#include "oracle_codec.h"

#include <algorithm>
#include <cstring>

namespace orc {

static const uint8_t kAlphabet[4] = {'A', 'C', 'G', 'T'};  // sbwt.alphabet()

bool as_2bit(const uint8_t* seq, uint64_t len, uint64_t* out) {
    if (len > 32) return false;
    uint64_t w = 0;
    for (uint64_t j = 0; j < len; j++) {
        int c = base_code(seq[j]);
        if (c < 0) return false;
        w |= (uint64_t)c << (2 * j);
    }
    *out = w;
    return true;
}

bool from_2bit(uint64_t word, uint64_t len, std::vector<uint8_t>* out) {
    if (len > 32) return false;
    for (uint64_t j = 0; j < len; j++) out->push_back(kAlphabet[(word >> (2 * j)) & 3]);
    return true;
}

// ---------------------------------------------------------------------------------------------
// literal restatement
// ---------------------------------------------------------------------------------------------

// left_extend_kmer (src/lib.rs:94-128) when ref_nts != nullptr, left_extend_kmer2 (:130-161) otherwise.
static std::vector<uint8_t> left_extend(const Index& ix, const std::vector<uint8_t>& kmer_start, const uint8_t* ref_nts,
                                        uint64_t ref_len, uint64_t max_extension_len) {
    uint64_t ext = 0;
    std::vector<uint8_t> kmer = kmer_start;
    while (ext < max_extension_len) {
        std::vector<std::pair<std::vector<uint8_t>, Interval>> found;
        for (int a = 0; a < 4; a++) {
            std::vector<uint8_t> cand;
            cand.push_back(kAlphabet[a]);
            cand.insert(cand.end(), kmer.begin(), kmer.begin() + (kmer.size() - (ext + 1)));
            Interval r;
            if (search(ix, cand.data(), cand.size(), &r)) found.emplace_back(std::move(cand), r);
        }
        if (found.empty()) break;
        bool seq_matches = true;
        if (ref_nts) seq_matches = found[0].first[0] == ref_nts[ref_len - kmer.size() - 1];
        if (seq_matches && found.size() == 1 && found[0].second.e - found[0].second.s == 1) {
            kmer.insert(kmer.begin(), found[0].first[0]);
        } else {
            break;
        }
        ext++;
    }
    return kmer;
}

int64_t literal_encode_sequence(const Index& ix, const uint8_t* read, uint64_t n, std::vector<DictEntry>* out) {
    out->clear();
    const uint64_t k = (uint64_t)ix.k;
    std::vector<uint32_t> d(n);
    std::vector<uint64_t> s(n), e(n);
    matching_statistics(ix, read, n, d.data(), s.data(), e.data());
    for (uint64_t j = 0; j < n; j++) if (d[j] == 0) return ERR_ABSENT_BASE;  // lib.rs:207 underflow (Q3)
    std::vector<uint64_t> res_len(d.begin(), d.end());

    uint64_t i = n;
    std::vector<uint64_t> kept;
    std::vector<uint8_t> kmer(k);
    while (i > 0) {
        if (res_len[i - 1] == k && i > k + 1) {                       // lib.rs:180
            access_kmer(ix, s[i - 1], kmer.data());                   // lib.rs:181
            if (memcmp(kmer.data(), read + (i - k), k) != 0) return ERR_ASSERT;  // lib.rs:183
            std::vector<uint8_t> ext = left_extend(ix, kmer, read, i, i - k - 1);  // lib.rs:186
            uint64_t match_len = ext.size();
            uint64_t old_i = i;
            for (;;) {                                                // lib.rs:193-203
                if (i == 0) return ERR_ASSERT;
                if (res_len[i - 1] < match_len) {
                    uint64_t old_ms = res_len[i - 1];
                    i -= old_ms;
                    match_len -= old_ms;
                } else {
                    res_len[old_i - 1] = ext.size() - (match_len - 1);
                    kept.push_back(old_i - 1);
                    break;
                }
            }
        } else {                                                      // lib.rs:204-211
            kept.push_back(i - 1);
            if (i > res_len[i - 1]) i -= res_len[i - 1] - 1;
            else break;
        }
        if (i > 0) i -= 1; else break;                                // lib.rs:213-217
    }
    uint64_t total = 0, mx = 0;
    for (uint64_t x : kept) {
        out->push_back(DictEntry{res_len[x], s[x], e[x]});
        total += res_len[x];
        mx = std::max(mx, res_len[x]);
    }
    if (out->empty()) return ERR_EMPTY;                               // lib.rs:224
    if (mx >= 16777216) return ERR_LEN24;                             // lib.rs:226
    if (total != n) return ERR_ASSERT;                                // lib.rs:227
    return (int64_t)out->size();
}

int64_t literal_encode_dictionary(const Index& ix, const DictEntry* dict, uint64_t n, std::vector<uint64_t>* out) {
    if (n == 0) return ERR_EMPTY;                                     // encode.rs:133
    const uint64_t k = (uint64_t)ix.k;
    const uint64_t bitnuc_threshold = 11;                             // encode.rs:137
    bool first = true;
    std::vector<uint8_t> kmer(k);
    for (uint64_t r = 0; r < n; r++) {
        uint8_t arr[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (dict[r].len > bitnuc_threshold) {                         // encode.rs:146-149
            uint32_t colex = (uint32_t)dict[r].start;
            uint32_t len32 = (uint32_t)dict[r].len;
            memcpy(arr, &colex, 4);
            memcpy(arr + 4, &len32, 3);
            arr[7] = first ? 1 : 0;
        } else {                                                      // encode.rs:150-156
            if (dict[r].len > k) return ERR_ASSERT;                   // k - len underflows at encode.rs:152
            access_kmer(ix, dict[r].start, kmer.data());
            uint64_t w;
            if (!as_2bit(kmer.data() + (k - dict[r].len), dict[r].len, &w)) return ERR_ASSERT;
            memcpy(arr, &w, 7);
            arr[7] = (uint8_t)(((first ? 1 : 0) + 2) | ((uint8_t)dict[r].len << 2));
        }
        uint64_t word;
        memcpy(&word, arr, 8);
        out->push_back(word);
        first = false;
    }
    return (int64_t)n;
}

int64_t literal_decode_sequence(const Index& ix, const uint64_t* rec, uint64_t n, std::vector<uint8_t>* bases,
                                std::vector<uint64_t>* offsets) {
    const uint64_t k = (uint64_t)ix.k;
    std::vector<std::vector<uint8_t>> sequences;
    std::vector<uint8_t> sequence;
    std::vector<uint8_t> kmer(k);
    for (uint64_t r = n; r-- > 0;) {                                  // lib.rs:266 (.rev())
        uint8_t bytes[8];
        memcpy(bytes, &rec[r], 8);
        uint8_t flag = bytes[7];
        bool first = (flag & 1) == 1;
        if ((flag & 2) == 0) {                                        // lib.rs:273-294
            uint32_t colex = 0, suffix_len = 0;
            memcpy(&colex, bytes, 4);
            memcpy(&suffix_len, bytes + 4, 3);
            if (colex >= ix.n) return ERR_ASSERT;
            access_kmer(ix, colex, kmer.data());
            std::vector<uint8_t> km = kmer;
            if (suffix_len > k) {
                km = left_extend(ix, kmer, nullptr, 0, suffix_len - k);
                if (km.size() != suffix_len) return ERR_ASSERT;       // lib.rs:288
            }
            sequence.insert(sequence.end(), km.end() - suffix_len, km.end());
        } else {                                                      // lib.rs:295-306
            uint64_t length = (flag & 0xFC) >> 2;
            uint64_t w = 0;
            memcpy(&w, bytes, 7);
            from_2bit(w, length, &sequence);                          // error ignored, lib.rs:303
        }
        if (first) {                                                  // lib.rs:308-312
            sequences.push_back(sequence);
            sequence.clear();
        }
    }
    for (size_t q = sequences.size(); q-- > 0;) {                     // lib.rs:317
        bases->insert(bases->end(), sequences[q].begin(), sequences[q].end());
        offsets->push_back(bases->size());
    }
    return (int64_t)sequences.size();
}

// ---------------------------------------------------------------------------------------------
// fast equivalents
// ---------------------------------------------------------------------------------------------

namespace {
struct FastRec { uint64_t len, start, end, end_pos; };

int64_t fast_segments(const Index& ix, const uint8_t* read, uint64_t n, std::vector<FastRec>* out) {
    out->clear();
    if (n == 0) return ERR_EMPTY;
    const uint64_t k = (uint64_t)ix.k;
    std::vector<uint32_t> d(n);
    std::vector<uint64_t> s(n), e(n);
    matching_statistics(ix, read, n, d.data(), s.data(), e.data());
    for (uint64_t j = 0; j < n; j++) if (d[j] == 0) return ERR_ABSENT_BASE;
    auto uniq = [&](uint64_t p) -> bool {
        if (d[p] != k) return false;
        uint64_t v = s[p];
        if (ix.lcs[v] >= k - 1) return false;
        return v + 1 == ix.n || ix.lcs[v + 1] < k - 1;
    };
    uint64_t i = n;
    while (i > 0) {
        uint64_t di = d[i - 1];
        if (di == k && i > k + 1) {
            uint64_t ext = 0;
            while (ext < i - k - 1 && uniq(i - 2 - ext)) ext++;
            uint64_t new_len = k + ext, m = new_len, old_i = i;
            while (d[i - 1] < m) { m -= d[i - 1]; i -= d[i - 1]; }
            uint64_t L = new_len - (m - 1);
            i -= 1;
            out->push_back(FastRec{L, s[old_i - 1], e[old_i - 1], old_i});
        } else {
            out->push_back(FastRec{di, s[i - 1], e[i - 1], i});
            i -= di;
        }
    }
    for (auto& r : *out) if (r.len >= 16777216) return ERR_LEN24;
    return (int64_t)out->size();
}
}  // namespace

int64_t fast_encode_sequence(const Index& ix, const uint8_t* read, uint64_t n, std::vector<DictEntry>* out) {
    std::vector<FastRec> recs;
    int64_t rc = fast_segments(ix, read, n, &recs);
    out->clear();
    if (rc < 0) return rc;
    for (auto& r : recs) out->push_back(DictEntry{r.len, r.start, r.end});
    return rc;
}

int64_t fast_encode_read(const Index& ix, const uint8_t* read, uint64_t n, std::vector<uint64_t>* out) {
    std::vector<FastRec> recs;
    int64_t rc = fast_segments(ix, read, n, &recs);
    if (rc < 0) return rc;
    uint64_t first = 1;
    for (auto& r : recs) {
        uint64_t word;
        if (r.len > 11) {
            word = (uint64_t)(uint32_t)r.start | ((r.len & 0xFFFFFF) << 32) | (first << 56);
        } else {
            uint64_t w = 0;
            as_2bit(read + (r.end_pos - r.len), r.len, &w);
            word = (w & 0x00FFFFFFFFFFFFFFull) | (((first + 2) | (r.len << 2)) << 56);
        }
        out->push_back(word);
        first = 0;
    }
    return rc;
}

int64_t fast_decode_sequence(const Index& ix, const uint64_t* rec, uint64_t n, std::vector<uint8_t>* bases,
                             std::vector<uint64_t>* offsets) {
    // sequences are delimited by bit0 on their first record in stream order; records before the
    // first flagged record never reach `sequences.push` in the reference and are dropped.
    uint64_t r = 0;
    while (r < n && ((rec[r] >> 56) & 1) == 0) r++;
    int64_t n_seqs = 0;
    while (r < n) {
        uint64_t q = r + 1;
        while (q < n && ((rec[q] >> 56) & 1) == 0) q++;
        for (uint64_t t = q; t-- > r;) {  // last record of the group holds the leftmost segment
            uint64_t word = rec[t];
            uint8_t flag = (uint8_t)(word >> 56);
            if ((flag & 2) == 0) {
                uint64_t v = word & 0xFFFFFFFFull, len = (word >> 32) & 0xFFFFFF;
                if (v >= ix.n) return ERR_ASSERT;
                size_t base = bases->size();
                bases->resize(base + len);
                for (uint64_t x = 0; x < len; x++) {
                    int c = ix.inlabel(v);
                    (*bases)[base + len - 1 - x] = c < 0 ? (uint8_t)'$' : kAlphabet[c];
                    v = ix.pred(v);
                }
            } else {
                from_2bit(word & 0x00FFFFFFFFFFFFFFull, (flag & 0xFC) >> 2, bases);
            }
        }
        offsets->push_back(bases->size());
        n_seqs++;
        r = q;
    }
    return n_seqs;
}

}  // namespace orc
