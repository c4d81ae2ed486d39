// oracle/oracle_index.cpp -- TEST INFRASTRUCTURE ONLY.  PARITY UNPINNED (see oracle_index.h).
// Straightforward (sort + binary search) SBWT/LCS construction and the query primitives.
#include "oracle_index.h"

#include <algorithm>
#include <cstring>

namespace orc {

namespace {

// A node label: the non-'$' characters 2-bit packed so that integer order == colex order.
// Character at string position i sits at bits 2i..2i+1 ('$' positions hold 0); `len` counts the
// real (non-'$') characters, which occupy string positions k-len..k-1.
template <int W>
struct Key {
    uint64_t w[W];
    uint16_t len;
};

template <int W>
inline int cmp_words(const Key<W>& a, const Key<W>& b) {
    for (int i = W - 1; i >= 0; i--) {
        if (a.w[i] != b.w[i]) return a.w[i] < b.w[i] ? -1 : 1;
    }
    return 0;
}
template <int W>
inline bool key_less(const Key<W>& a, const Key<W>& b) {
    int c = cmp_words<W>(a, b);
    if (c) return c < 0;
    return a.len < b.len;  // equal packed bits: more '$' padding sorts first
}
template <int W>
inline bool key_eq(const Key<W>& a, const Key<W>& b) { return cmp_words<W>(a, b) == 0 && a.len == b.len; }

template <int W>
inline void shl2(Key<W>& a) {  // multiply by 4
    for (int i = W - 1; i > 0; i--) a.w[i] = (a.w[i] << 2) | (a.w[i - 1] >> 62);
    a.w[0] <<= 2;
}
template <int W>
inline void shr2(Key<W>& a) {
    for (int i = 0; i < W - 1; i++) a.w[i] = (a.w[i] >> 2) | (a.w[i + 1] << 62);
    a.w[W - 1] >>= 2;
}
template <int W>
inline void mask_k(Key<W>& a, int k) {  // keep the low 2k bits
    int bits = 2 * k;
    for (int i = 0; i < W; i++) {
        int lo = i * 64;
        if (bits >= lo + 64) continue;
        if (bits <= lo) a.w[i] = 0;
        else a.w[i] &= ((1ull << (bits - lo)) - 1);
    }
}
template <int W>
inline void set_char(Key<W>& a, int pos, uint64_t c) { a.w[pos >> 5] |= c << ((pos & 31) * 2); }
template <int W>
inline int get_char(const Key<W>& a, int pos) { return (int)((a.w[pos >> 5] >> ((pos & 31) * 2)) & 3); }

// number of equal characters counted from string position k-1 downwards
template <int W>
inline int common_suffix_chars(const Key<W>& a, const Key<W>& b, int k) {
    int n = 0;
    for (int p = k - 1; p >= 0; p--) {
        if (get_char<W>(a, p) != get_char<W>(b, p)) break;
        n++;
    }
    return n;
}

template <int W>
Index* build_impl(const uint8_t* const* seqs, const uint64_t* lens, uint64_t n_seqs, int k, bool add_revcomp) {
    typedef Key<W> K;
    std::vector<K> kmers;
    // 1. all k-mers of both strands
    for (uint64_t si = 0; si < n_seqs; si++) {
        const uint8_t* s = seqs[si];
        uint64_t L = lens[si];
        K fw, rc;
        memset(&fw, 0, sizeof(K));
        memset(&rc, 0, sizeof(K));
        uint64_t run = 0;
        for (uint64_t j = 0; j < L; j++) {
            int c = base_code(s[j]);
            if (c < 0) { run = 0; memset(&fw, 0, sizeof(K)); memset(&rc, 0, sizeof(K)); continue; }
            shr2<W>(fw);
            set_char<W>(fw, k - 1, (uint64_t)c);
            shl2<W>(rc);
            rc.w[0] |= (uint64_t)(3 - c);
            mask_k<W>(rc, k);
            run++;
            if (run >= (uint64_t)k) {
                fw.len = (uint16_t)k; rc.len = (uint16_t)k;
                kmers.push_back(fw);
                if (add_revcomp) kmers.push_back(rc);
            }
        }
    }
    std::sort(kmers.begin(), kmers.end(), key_less<W>);
    kmers.erase(std::unique(kmers.begin(), kmers.end(), key_eq<W>), kmers.end());

    // 2. k-mers without a predecessor: no k-mer y with y[1..k) == x[0..k-1)
    std::vector<K> dummies;
    {
        K root; memset(&root, 0, sizeof(K)); root.len = 0;
        dummies.push_back(root);
    }
    for (size_t i = 0; i < kmers.size(); i++) {
        K p = kmers[i];
        shl2<W>(p); mask_k<W>(p, k);        // x[0..k-1) moved to positions 1..k-1, position 0 cleared
        p.len = 0;                          // smallest key with these bits
        auto it = std::lower_bound(kmers.begin(), kmers.end(), p, key_less<W>);
        bool has_pred = false;
        if (it != kmers.end()) {
            K y = *it; y.w[0] &= ~3ull;     // drop y's first character
            has_pred = cmp_words<W>(y, p) == 0;
        }
        if (has_pred) continue;
        // 3. '$'-padded proper prefixes: $^j x[0..k-j), j = 1..k-1 (root already present)
        K d = kmers[i];
        for (int j = 1; j < k; j++) {
            shl2<W>(d); mask_k<W>(d, k);
            d.len = (uint16_t)(k - j);
            dummies.push_back(d);
        }
    }
    std::sort(dummies.begin(), dummies.end(), key_less<W>);
    dummies.erase(std::unique(dummies.begin(), dummies.end(), key_eq<W>), dummies.end());

    // 4. merge
    std::vector<K> nodes(kmers.size() + dummies.size());
    std::merge(kmers.begin(), kmers.end(), dummies.begin(), dummies.end(), nodes.begin(), key_less<W>);
    kmers.clear(); kmers.shrink_to_fit();
    dummies.clear(); dummies.shrink_to_fit();

    Index* ix = new Index();
    ix->k = k;
    ix->n = nodes.size();
    uint64_t n = ix->n;
    uint64_t words = n / 64 + 2;
    for (int c = 0; c < 4; c++) ix->rows[c].assign(words, 0);
    ix->lcs.assign(n, 0);

    // 5. edges: every non-root node v has exactly one stored in-edge, from the leader (colex-first
    //    member) of the group of nodes whose (k-1)-suffix equals v's (k-1)-prefix.
    for (uint64_t vi = 1; vi < n; vi++) {
        const K& v = nodes[vi];
        int c = get_char<W>(v, k - 1);
        K p = v;
        shl2<W>(p); mask_k<W>(p, k);
        p.len = (uint16_t)(v.len - 1);      // k-1 for real k-mers, exact dummy length otherwise
        auto it = std::lower_bound(nodes.begin(), nodes.end(), p, key_less<W>);
        // leader must exist and belong to that group
        uint64_t u = (uint64_t)(it - nodes.begin());
        if (u >= n) { delete ix; return nullptr; }
        K y = nodes[u]; y.w[0] &= ~3ull;
        uint16_t ylen = nodes[u].len < (uint16_t)(k - 1) ? nodes[u].len : (uint16_t)(k - 1);
        if (cmp_words<W>(y, p) != 0 || ylen != p.len) { delete ix; return nullptr; }
        ix->rows[c][u >> 6] |= 1ull << (u & 63);
    }
    // 7. LCS
    for (uint64_t i = 1; i < n; i++) {
        int common = common_suffix_chars<W>(nodes[i - 1], nodes[i], k);
        int lim = std::min<int>(nodes[i - 1].len, nodes[i].len);
        ix->lcs[i] = (uint8_t)std::min(common, lim);
    }
    ix->finalize();
    return ix;
}

}  // namespace

void Index::finalize() {
    uint64_t words = rows[0].size();
    C[0] = 1;
    for (int c = 0; c < 4; c++) {
        rank_cum[c].assign(words + 1, 0);
        ones[c].clear();
        uint32_t acc = 0;
        for (uint64_t w = 0; w < words; w++) {
            rank_cum[c][w] = acc;
            uint64_t x = rows[c][w];
            acc += (uint32_t)__builtin_popcountll(x);
            while (x) {
                int b = __builtin_ctzll(x);
                ones[c].push_back((uint32_t)(w * 64 + b));
                x &= x - 1;
            }
        }
        rank_cum[c][words] = acc;
        C[c + 1] = C[c] + acc;
    }
}

Index* build_index(const uint8_t* const* seqs, const uint64_t* lens, uint64_t n_seqs, int k, bool add_revcomp,
                   std::string* err) {
    if (k < 2 || k > kMaxK) { if (err) *err = "k out of range [2,255]"; return nullptr; }
    int W = (2 * k + 63) / 64;
    Index* ix = nullptr;
    if (W <= 1) ix = build_impl<1>(seqs, lens, n_seqs, k, add_revcomp);
    else if (W == 2) ix = build_impl<2>(seqs, lens, n_seqs, k, add_revcomp);
    else if (W == 3) ix = build_impl<3>(seqs, lens, n_seqs, k, add_revcomp);
    else if (W == 4) ix = build_impl<4>(seqs, lens, n_seqs, k, add_revcomp);
    else ix = build_impl<8>(seqs, lens, n_seqs, k, add_revcomp);
    if (!ix && err) *err = "internal: predecessor group not found";
    if (ix && ix->C[4] != ix->n) { if (err) *err = "internal: edge count != n-1"; delete ix; return nullptr; }
    return ix;
}

Index* index_from_arrays(int k, uint64_t n, const uint64_t* const rows[4], const uint8_t* lcs, std::string* err) {
    if (k < 2 || k > kMaxK) { if (err) *err = "k out of range"; return nullptr; }
    Index* ix = new Index();
    ix->k = k; ix->n = n;
    uint64_t words = n / 64 + 2, src_words = (n + 63) / 64;
    for (int c = 0; c < 4; c++) {
        ix->rows[c].assign(words, 0);
        memcpy(ix->rows[c].data(), rows[c], src_words * 8);
        if (n & 63) ix->rows[c][src_words - 1] &= (1ull << (n & 63)) - 1;
    }
    ix->lcs.assign(lcs, lcs + n);
    ix->finalize();
    if (ix->C[4] != n) { if (err) *err = "rows hold != n-1 edges"; delete ix; return nullptr; }
    return ix;
}

bool search(const Index& ix, const uint8_t* pat, uint64_t len, Interval* out) {
    Interval I{0, ix.n};
    for (uint64_t i = 0; i < len; i++) {
        int c = base_code(pat[i]);
        if (c < 0) return false;
        I = extend_right(ix, I, c);
        if (I.s >= I.e) return false;
    }
    *out = I;
    return true;
}

// StreamingIndex::matching_statistics (sbwt crate) as called at src/lib.rs:172-173; SURVEY §8a A1.
void matching_statistics(const Index& ix, const uint8_t* read, uint64_t len, uint32_t* d_out, uint64_t* s_out,
                         uint64_t* e_out) {
    uint64_t d = 0;
    Interval I{0, ix.n};
    for (uint64_t j = 0; j < len; j++) {
        int c = base_code(read[j]);
        Interval Ic = extend_right(ix, I, c);
        while (d > 0 && Ic.s >= Ic.e) {
            I = contract_left(ix, I, d - 1);
            d -= 1;
            Ic = extend_right(ix, I, c);
        }
        if (Ic.s < Ic.e) {
            I = Ic;
            d = std::min<uint64_t>(d + 1, (uint64_t)ix.k);
        }
        d_out[j] = (uint32_t)d;
        s_out[j] = I.s;
        e_out[j] = I.e;
    }
}

// SbwtIndex::access_kmer (src/lib.rs:181,286,291; src/encode.rs:151); SURVEY §8a A4.
void access_kmer(const Index& ix, uint64_t v, uint8_t* out) {
    static const char kAlpha[4] = {'A', 'C', 'G', 'T'};
    for (int t = 0; t < ix.k; t++) {
        int c = ix.inlabel(v);
        out[ix.k - 1 - t] = c < 0 ? (uint8_t)'$' : (uint8_t)kAlpha[c];
        v = ix.pred(v);
    }
}

}  // namespace orc
