// index_build.cpp -- host SBWT + LCS construction (offline preprocessing, stays on the CPU as
// north_star says).  Stands in for kbo::build as called at src/main.rs:111-134 with
// add_revcomp = true, build_select = true (the select side becomes the decode jump table).
//
// Method: 2-bit pack every k-mer of both strands so that integer order == colex order, bucket by
// the last 8 characters, sort buckets in parallel, then derive dummy nodes, subset rows and LCS
// with linear merges (nodes ending in c, ordered, have ordered (k-1)-prefixes).
#include <algorithm>
#include <atomic>
#include <cstring>
#include <thread>

#include "host_index.h"

namespace ntc {
namespace {

template <int W>
struct Kmer {
    uint64_t w[W];
};

template <int W>
inline int cmp(const Kmer<W>& a, const Kmer<W>& b) {
    for (int i = W - 1; i >= 0; i--)
        if (a.w[i] != b.w[i]) return a.w[i] < b.w[i] ? -1 : 1;
    return 0;
}
template <int W>
struct Less {
    bool operator()(const Kmer<W>& a, const Kmer<W>& b) const { return cmp<W>(a, b) < 0; }
};
template <int W>
inline Kmer<W> shl2(Kmer<W> a, const uint64_t* mask) {  // drop the last character, shift the rest up
    for (int i = W - 1; i > 0; i--) a.w[i] = ((a.w[i] << 2) | (a.w[i - 1] >> 62)) & mask[i];
    a.w[0] = (a.w[0] << 2) & mask[0];
    return a;
}
template <int W>
inline Kmer<W> shr2(Kmer<W> a) {
    for (int i = 0; i < W - 1; i++) a.w[i] = (a.w[i] >> 2) | (a.w[i + 1] << 62);
    a.w[W - 1] >>= 2;
    return a;
}

inline int code_of(uint8_t b) {
    switch (b) { case 'A': return 0; case 'C': return 1; case 'G': return 2; case 'T': return 3; default: return -1; }
}

template <typename F>
void parallel_for(uint64_t n_items, int n_threads, F f) {  // f(item) with dynamic scheduling
    std::atomic<uint64_t> next(0);
    auto work = [&]() { for (;;) { uint64_t i = next.fetch_add(1); if (i >= n_items) break; f(i); } };
    std::vector<std::thread> pool;
    for (int t = 1; t < n_threads; t++) pool.emplace_back(work);
    work();
    for (auto& t : pool) t.join();
}

template <int W>
struct Builder {
    typedef Kmer<W> K;
    uint32_t k;
    int threads;
    uint64_t mask[W];
    int top_shift;  // bit position (within word W-1) of the last character

    Builder(uint32_t k_, int threads_) : k(k_), threads(threads_) {
        int bits = 2 * (int)k;
        for (int i = 0; i < W; i++) {
            int lo = i * 64;
            mask[i] = bits >= lo + 64 ? ~0ull : (bits <= lo ? 0ull : ((1ull << (bits - lo)) - 1));
        }
        top_shift = (2 * ((int)k - 1)) % 64;
    }
    inline int top_word() const { return (2 * ((int)k - 1)) / 64; }
    inline int last_char(const K& x) const { return (int)((x.w[top_word()] >> top_shift) & 3); }
    inline uint32_t bucket16(const K& x) const {  // last 8 characters as a 16-bit number
        int hi = 2 * (int)k;                       // exclusive top bit
        int lo = hi - 16;
        int wl = lo / 64, off = lo % 64;
        uint64_t v = x.w[wl] >> off;
        if (off > 48 && wl + 1 < W) v |= x.w[wl + 1] << (64 - off);
        return (uint32_t)(v & 0xFFFF);
    }
    // number of equal characters from the last position downwards
    inline int common_suffix(const K& a, const K& b) const {
        for (int i = W - 1; i >= 0; i--) {
            uint64_t x = a.w[i] ^ b.w[i];
            if (x) {
                int top_bit = 63 - __builtin_clzll(x);       // highest differing bit in this word
                int diff_char = (i * 64 + top_bit) / 2;      // string position of the first mismatch
                return (int)k - 1 - diff_char;
            }
        }
        return (int)k;
    }

    // ---- 1. enumerate k-mers of both strands ----
    void collect(const uint8_t* const* seqs, const uint64_t* lens, uint64_t n_seqs, bool add_revcomp, std::vector<K>* out) {
        // split every sequence into jobs of ~1 Mbp of k-mer end positions
        struct Job { uint64_t seq, a, b; };
        std::vector<Job> jobs;
        const uint64_t step = 1 << 20;
        for (uint64_t s = 0; s < n_seqs; s++)
            for (uint64_t a = 0; a < lens[s]; a += step) jobs.push_back(Job{s, a, std::min(lens[s], a + step)});
        std::vector<std::vector<K>> parts(jobs.size());
        parallel_for(jobs.size(), threads, [&](uint64_t ji) {
            const Job& jb = jobs[ji];
            const uint8_t* s = seqs[jb.seq];
            std::vector<K>& dst = parts[ji];
            K fw, rc;
            memset(&fw, 0, sizeof(K)); memset(&rc, 0, sizeof(K));
            uint64_t run = 0;
            uint64_t begin = jb.a >= (uint64_t)(k - 1) ? jb.a - (k - 1) : 0;  // warm-up so windows ending in [a,b) are complete
            for (uint64_t j = begin; j < jb.b; j++) {
                int c = code_of(s[j]);
                if (c < 0) { run = 0; continue; }
                fw = shr2<W>(fw);
                fw.w[top_word()] |= (uint64_t)c << top_shift;
                rc = shl2<W>(rc, mask);
                rc.w[0] |= (uint64_t)(3 - c);
                run++;
                if (run >= k && j >= jb.a) {
                    dst.push_back(fw);
                    if (add_revcomp) dst.push_back(rc);
                }
            }
        });
        uint64_t total = 0;
        for (auto& p : parts) total += p.size();
        out->reserve(total);
        for (auto& p : parts) { out->insert(out->end(), p.begin(), p.end()); std::vector<K>().swap(p); }
    }

    // ---- 2. bucketed parallel sort + dedup ----
    void sort_unique(std::vector<K>* v) {
        const uint32_t NB = 1 << 16;
        std::vector<uint64_t> cnt(NB + 1, 0);
        for (auto& x : *v) cnt[bucket16(x) + 1]++;
        for (uint32_t b = 0; b < NB; b++) cnt[b + 1] += cnt[b];
        std::vector<K> tmp(v->size());
        {
            std::vector<uint64_t> pos(cnt.begin(), cnt.end() - 1);
            for (auto& x : *v) tmp[pos[bucket16(x)]++] = x;
        }
        std::vector<uint64_t> uniq_cnt(NB, 0);
        parallel_for(NB, threads, [&](uint64_t b) {
            K* lo = tmp.data() + cnt[b];
            K* hi = tmp.data() + cnt[b + 1];
            std::sort(lo, hi, Less<W>());
            K* e = std::unique(lo, hi, [](const K& a, const K& c) { return cmp<W>(a, c) == 0; });
            uniq_cnt[b] = (uint64_t)(e - lo);
        });
        uint64_t total = 0;
        for (uint32_t b = 0; b < NB; b++) total += uniq_cnt[b];
        v->resize(total);
        uint64_t o = 0;
        for (uint32_t b = 0; b < NB; b++) {
            memcpy((void*)(v->data() + o), (const void*)(tmp.data() + cnt[b]), uniq_cnt[b] * sizeof(K));
            o += uniq_cnt[b];
        }
    }

    int32_t run(const uint8_t* const* seqs, const uint64_t* lens, uint64_t n_seqs, bool add_revcomp, ntc_index** out) {
        std::vector<K> real;
        collect(seqs, lens, n_seqs, add_revcomp, &real);
        sort_unique(&real);
        const uint64_t nr = real.size();

        // ---- 3. sources: k-mers whose (k-1)-prefix is no k-mer's (k-1)-suffix ----
        // nodes ending in c are one contiguous colex range whose prefixes are ordered -> linear merge
        uint64_t crange[5];
        crange[0] = 0;
        for (int c = 0; c < 4; c++) {
            K probe; memset(&probe, 0, sizeof(K));
            probe.w[top_word()] = (uint64_t)(c + 1) << top_shift;  // first key whose last char is > c (c == 3 overflows past mask)
            crange[c + 1] = c == 3 ? nr : (uint64_t)(std::lower_bound(real.begin(), real.end(), probe, Less<W>()) - real.begin());
        }
        std::vector<uint64_t> sources[4];
        parallel_for(4, threads, [&](uint64_t c) {
            uint64_t g = 0;
            for (uint64_t i = crange[c]; i < crange[c + 1]; i++) {
                K p = shl2<W>(real[i], mask);
                for (; g < nr; g++) {
                    K y = real[g]; y.w[0] &= ~3ull;
                    if (cmp<W>(y, p) >= 0) break;
                }
                bool has_pred = false;
                if (g < nr) { K y = real[g]; y.w[0] &= ~3ull; has_pred = cmp<W>(y, p) == 0; }
                if (!has_pred) sources[c].push_back(i);
            }
        });

        // ---- 4. dummy nodes: root + $^j x[0..k-j) for every source x ----
        struct Dummy { K key; uint16_t len; };
        std::vector<Dummy> dummies;
        { Dummy r; memset(&r, 0, sizeof(r)); dummies.push_back(r); }
        for (int c = 0; c < 4; c++)
            for (uint64_t i : sources[c]) {
                K d = real[i];
                for (uint32_t j = 1; j < k; j++) {
                    d = shl2<W>(d, mask);
                    Dummy dm; dm.key = d; dm.len = (uint16_t)(k - j);
                    dummies.push_back(dm);
                }
            }
        auto dless = [](const Dummy& a, const Dummy& b) { int c = cmp<W>(a.key, b.key); return c ? c < 0 : a.len < b.len; };
        std::sort(dummies.begin(), dummies.end(), dless);
        dummies.erase(std::unique(dummies.begin(), dummies.end(),
                                  [](const Dummy& a, const Dummy& b) { return cmp<W>(a.key, b.key) == 0 && a.len == b.len; }),
                      dummies.end());

        // ---- 5. merge into the final colex order (key, len); equal bits: shorter (more '$') first ----
        const uint64_t n = nr + dummies.size();
        if (n >= (1ull << 32)) { set_error("n_sets >= 2^32 (colex ranks are stored as u32, src/encode.rs:147)"); return 9; }
        std::vector<K> keys(n);
        std::vector<uint16_t> klen(n);
        {
            uint64_t a = 0, b = 0, o = 0;
            while (a < nr || b < dummies.size()) {
                bool take_dummy;
                if (a >= nr) take_dummy = true;
                else if (b >= dummies.size()) take_dummy = false;
                else take_dummy = cmp<W>(dummies[b].key, real[a]) <= 0;  // equal bits: dummy (len < k) first
                if (take_dummy) { keys[o] = dummies[b].key; klen[o] = dummies[b].len; b++; }
                else { keys[o] = real[a]; klen[o] = (uint16_t)k; a++; }
                o++;
            }
        }
        std::vector<K>().swap(real);

        ntc_index* idx = new ntc_index();
        idx->k = k;
        idx->n = n;
        const uint64_t words = (n + 63) / 64;
        for (int c = 0; c < 4; c++) idx->rows[c].assign(words, 0);
        idx->lcs.assign(n, 0);

        // ---- 6. subset rows: every non-root node has one stored in-edge, on the leader of the
        //         group of nodes whose (k-1)-suffix equals its (k-1)-prefix ----
        uint64_t nrange[5];
        nrange[0] = 1;  // node 0 is the root
        for (int c = 0; c < 4; c++) {
            uint64_t lo = nrange[c], hi = n;
            while (lo < hi) {  // first non-root node whose last char is > c
                uint64_t mid = (lo + hi) / 2;
                if (last_char(keys[mid]) > c) hi = mid; else lo = mid + 1;
            }
            nrange[c + 1] = lo;
        }
        std::atomic<int> bad(0);
        parallel_for(4, threads, [&](uint64_t c) {
            uint64_t g = 0;
            for (uint64_t v = nrange[c]; v < nrange[c + 1]; v++) {
                K p = shl2<W>(keys[v], mask);
                uint16_t plen = (uint16_t)(klen[v] - 1);
                for (; g < n; g++) {
                    K y = keys[g]; y.w[0] &= ~3ull;
                    uint16_t ylen = std::min<uint16_t>(klen[g], (uint16_t)(k - 1));
                    int q = cmp<W>(y, p);
                    if (q > 0 || (q == 0 && ylen >= plen)) break;
                }
                bool ok = false;
                if (g < n) {
                    K y = keys[g]; y.w[0] &= ~3ull;
                    ok = cmp<W>(y, p) == 0 && std::min<uint16_t>(klen[g], (uint16_t)(k - 1)) == plen;
                }
                if (!ok) { bad.store(1); return; }
                idx->rows[c][g >> 6] |= 1ull << (g & 63);
            }
        });
        if (bad.load()) { delete idx; set_error("internal: predecessor group missing"); return 1; }

        // ---- 7. LCS ----
        const uint64_t chunk = 1 << 16;
        parallel_for((n + chunk - 1) / chunk, threads, [&](uint64_t ci) {
            uint64_t lo = std::max<uint64_t>(1, ci * chunk), hi = std::min(n, (ci + 1) * chunk);
            for (uint64_t i = lo; i < hi; i++) {
                int common = common_suffix(keys[i - 1], keys[i]);
                int lim = std::min<int>(klen[i - 1], klen[i]);
                idx->lcs[i] = (uint8_t)std::min(common, lim);
            }
        });
        int32_t rc = finalize_index(idx);
        if (rc) { delete idx; return rc; }
        *out = idx;
        return 0;
    }
};

}  // namespace

int32_t build_index(const uint8_t* const* seqs, const uint64_t* lens, uint64_t n_seqs, uint32_t k, bool add_revcomp,
                    int n_threads, ntc_index** out) {
    if (k < kMinK || k > kMaxK) { set_error("k must be in [12,255]"); return 9; }
    if (n_threads < 1) n_threads = (int)std::max(1u, std::thread::hardware_concurrency());
    int W = (2 * (int)k + 63) / 64;
    switch (W) {
        case 1: return Builder<1>(k, n_threads).run(seqs, lens, n_seqs, add_revcomp, out);
        case 2: return Builder<2>(k, n_threads).run(seqs, lens, n_seqs, add_revcomp, out);
        case 3: return Builder<3>(k, n_threads).run(seqs, lens, n_seqs, add_revcomp, out);
        case 4: return Builder<4>(k, n_threads).run(seqs, lens, n_seqs, add_revcomp, out);
        case 5: return Builder<5>(k, n_threads).run(seqs, lens, n_seqs, add_revcomp, out);
        case 6: return Builder<6>(k, n_threads).run(seqs, lens, n_seqs, add_revcomp, out);
        case 7: return Builder<7>(k, n_threads).run(seqs, lens, n_seqs, add_revcomp, out);
        default: return Builder<8>(k, n_threads).run(seqs, lens, n_seqs, add_revcomp, out);
    }
}

}  // namespace ntc
