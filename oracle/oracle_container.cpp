// oracle/oracle_container.cpp -- TEST INFRASTRUCTURE ONLY.  PARITY UNPINNED (see oracle_index.h).
// CPU restatement of ntcomp's block container:
//   split_encoded_dictionary  src/encode.rs:168-229     zip_block_contents  src/decode.rs:102-149
//   rice_encode / minimal_binary_encode  src/encode.rs:59-94   (+ decode.rs:51-77)
//   compress_block  src/encode.rs:96-127                 decompress_block    src/decode.rs:79-100
//   write_block_to  src/lib.rs:232-252                   decode_block        src/lib.rs:320-368
// Third-party semantics restated (not under /root/reference): dsi-bitstream 0.5.0 (BE bit writer over
// u64 words, unary = n zeros then a one, Rice, minimal binary, rice::log2_b), bincode 2.0.1 fixed-int
// little-endian headers, flate2 1.1.2 gzip member (mtime 0, xfl 0, os 255).  DEFLATE payload bytes
// come from system zlib here, NOT zlib-rs: payload byte identity with the reference is unpinned and
// parity is defined on the inflated words.
#include "oracle_container.h"

#include <zlib.h>

#include <algorithm>
#include <cmath>
#include <cstring>

namespace orc {

// ---- dsi-bitstream BufBitWriter<BE, MemWordWriterVec<u64>> -------------------------------------
struct BitWriterBE {
    std::vector<uint64_t> words;  // each stored as to_be(): byte order in memory == stream order
    uint64_t buf = 0;
    int used = 0;                 // bits currently in buf (MSB-first)
    void write_bits(uint64_t v, int nbits) {
        if (nbits == 0) return;                       // v < 2^nbits, nbits in 1..64
        int room = 64 - used;
        if (nbits <= room) {
            buf |= (nbits == 64) ? v : (v << (room - nbits));
            used += nbits;
            if (used == 64) { words.push_back(__builtin_bswap64(buf)); buf = 0; used = 0; }
        } else {
            int rest = nbits - room;                  // 1..63
            buf |= v >> rest;
            words.push_back(__builtin_bswap64(buf));
            buf = v << (64 - rest);
            used = rest;
        }
    }
    void write_unary(uint64_t n) {
        while (n >= 64) { write_bits(0, 64); n -= 64; }
        write_bits(1, (int)n + 1);
    }
    void flush() { if (used) { words.push_back(__builtin_bswap64(buf)); buf = 0; used = 0; } }
};

struct BitReaderBE {
    const uint64_t* words; uint64_t n_words; uint64_t pos = 0;  // bit position
    bool read_bit(int* b) {
        uint64_t w = pos >> 6;
        if (w >= n_words) return false;
        uint64_t x = __builtin_bswap64(words[w]);
        *b = (int)((x >> (63 - (pos & 63))) & 1);
        pos++;
        return true;
    }
    bool read_bits(int nbits, uint64_t* v) {
        uint64_t r = 0;
        for (int i = 0; i < nbits; i++) { int b; if (!read_bit(&b)) return false; r = (r << 1) | (uint64_t)b; }
        *v = r;
        return true;
    }
    bool read_unary(uint64_t* v) {
        uint64_t r = 0;
        for (;;) { int b; if (!read_bit(&b)) return false; if (b) break; r++; }
        *v = r;
        return true;
    }
};

// dsi_bitstream::codes::rice::log2_b(p) = ceil(log2(-ln(phi) / ln(1-p))) as usize (saturating cast)
uint64_t rice_log2_b(double p) {
    double v = std::ceil(std::log2(-std::log((std::sqrt(5.0) + 1.0) / 2.0) / std::log1p(-p)));
    if (!(v > 0.0)) return 0;  // NaN, -inf, negatives saturate to 0
    if (v > 63.0) return 63;
    return (uint64_t)v;
}

bool rice_encode(const std::vector<uint64_t>& ints, std::vector<uint64_t>* words, uint64_t* param) {
    double sum = 0.0;
    uint64_t isum = 0;
    for (uint64_t x : ints) isum += x;
    sum = (double)isum;
    double inv_mean = std::exp(std::log((double)ints.size()) - std::log(sum));   // encode.rs:62
    uint64_t b = rice_log2_b(inv_mean);
    BitWriterBE w;
    for (uint64_t x : ints) {
        w.write_unary(x >> b);
        if (b) w.write_bits(x & ((1ull << b) - 1), (int)b);
    }
    w.flush();
    *words = std::move(w.words);
    *param = b;
    return true;
}

static inline void minimal_binary_params(uint64_t max, int* l, uint64_t* limit) {
    *l = 63 - __builtin_clzll(max);
    *limit = ((1ull << *l) << 1) - max;
}

bool minimal_binary_encode(const std::vector<uint64_t>& ints, std::vector<uint64_t>* words, uint64_t* param) {
    if (ints.empty()) return false;                                             // encode.rs:80
    uint64_t mx = *std::max_element(ints.begin(), ints.end());
    uint64_t max = mx + 2;                                                      // encode.rs:82
    int l; uint64_t limit;
    minimal_binary_params(max, &l, &limit);
    BitWriterBE w;
    for (uint64_t x : ints) {
        uint64_t n = x + 1;                                                     // encode.rs:88
        if (n < limit) { if (l) w.write_bits(n, l); }
        else w.write_bits(n + limit, l + 1);
    }
    w.flush();
    *words = std::move(w.words);
    *param = max;
    return true;
}

bool rice_decode(const uint64_t* words, uint64_t n_words, uint64_t n, uint64_t b, std::vector<uint64_t>* out) {
    BitReaderBE r{words, n_words};
    for (uint64_t i = 0; i < n; i++) {
        uint64_t q, rem = 0;
        if (!r.read_unary(&q)) return false;
        if (b && !r.read_bits((int)b, &rem)) return false;
        out->push_back((q << b) | rem);
    }
    return true;
}

bool minimal_binary_decode(const uint64_t* words, uint64_t n_words, uint64_t n, uint64_t max, std::vector<uint64_t>* out) {
    if (max == 0) return false;
    int l; uint64_t limit;
    minimal_binary_params(max, &l, &limit);
    BitReaderBE r{words, n_words};
    for (uint64_t i = 0; i < n; i++) {
        uint64_t v = 0;
        if (l && !r.read_bits(l, &v)) return false;
        if (v >= limit) { int b; if (!r.read_bit(&b)) return false; v = ((v << 1) | (uint64_t)b) - limit; }
        out->push_back(v - 1);                                                  // decode.rs:74
    }
    return true;
}

// ---- split / zip --------------------------------------------------------------------------------
bool split_encoded_dictionary(const uint64_t* rec, uint64_t n, Columns* c) {
    if (n == 0) return false;                                                   // encode.rs:171
    std::vector<uint8_t> lit;
    for (uint64_t i = 0; i < n; i++) {
        uint8_t flag = (uint8_t)(rec[i] >> 56);
        if ((flag & 2) == 0) {
            c->colex.push_back(rec[i] & 0xFFFFFFFFull);
            c->len.push_back((rec[i] >> 32) & 0xFFFFFFull);
        } else {
            uint64_t length = (flag & 0xFC) >> 2;
            const uint64_t payload = rec[i] & 0x00FFFFFFFFFFFFFFull;               // arr1[0..7], encode.rs:215-216
            for (uint64_t j = 0; j < length && j < 32; j++) lit.push_back((uint8_t)((payload >> (2 * j)) & 3));
        }
        c->flags.push_back(flag);
    }
    for (size_t p = 0; p < lit.size(); p += 31) {                               // encode.rs:224
        uint64_t w = 0;
        for (size_t j = 0; j < 31 && p + j < lit.size(); j++) w |= (uint64_t)lit[p + j] << (2 * j);
        c->literals.push_back(w);
    }
    return true;
}

int zip_block_contents(const Columns& c, std::vector<uint64_t>* rec) {
    if (c.colex.size() != c.len.size()) return -1;                              // decode.rs:108
    uint64_t bitnuc_len = 0;
    for (uint64_t f : c.flags) bitnuc_len += ((f & 0xFF) & 0xFC) >> 2;
    std::vector<uint8_t> vals;
    for (size_t idx = 0; idx < c.literals.size(); idx++) {
        uint64_t len = idx + 1 == c.literals.size() ? bitnuc_len % 31 : 31;     // decode.rs:114-118 (Q5)
        for (uint64_t j = 0; j < len; j++) vals.push_back((uint8_t)((c.literals[idx] >> (2 * j)) & 3));
    }
    size_t i = 0, j = 0;
    for (uint64_t f : c.flags) {
        uint8_t flag = (uint8_t)f;
        uint64_t word = 0;
        if ((flag & 2) == 0) {
            if (i >= c.colex.size()) return -2;
            word = (c.colex[i] & 0xFFFFFFFFull) | ((c.len[i] & 0xFFFFFFull) << 32);
            i++;
        } else {
            uint64_t length = (flag & 0xFC) >> 2;
            if (j + length > vals.size()) return -3;                            // decode.rs:138 panics (Q5)
            if (length > 32) return -4;
            for (uint64_t t = 0; t < length; t++) word |= (uint64_t)vals[j + t] << (2 * t);
            word &= 0x00FFFFFFFFFFFFFFull;
            j += length;
        }
        word |= (uint64_t)flag << 56;
        rec->push_back(word);
    }
    return 0;
}

// ---- gzip member (flate2 GzEncoder, Compression::default()) ------------------------------------
static bool gzip_bytes(const uint8_t* in, size_t n, std::vector<uint8_t>* out) {
    z_stream zs;
    memset(&zs, 0, sizeof(zs));
    if (deflateInit2(&zs, 6, Z_DEFLATED, 15 + 16, 8, Z_DEFAULT_STRATEGY) != Z_OK) return false;
    gz_header hd;
    memset(&hd, 0, sizeof(hd));
    hd.os = 255;
    deflateSetHeader(&zs, &hd);
    out->resize(deflateBound(&zs, n) + 32);
    zs.next_in = const_cast<Bytef*>(in);
    zs.avail_in = (uInt)n;
    zs.next_out = out->data();
    zs.avail_out = (uInt)out->size();
    int rc = deflate(&zs, Z_FINISH);
    deflateEnd(&zs);
    if (rc != Z_STREAM_END) return false;
    out->resize(zs.total_out);
    return true;
}

static bool gunzip_bytes(const uint8_t* in, size_t n, std::vector<uint8_t>* out) {
    z_stream zs;
    memset(&zs, 0, sizeof(zs));
    if (inflateInit2(&zs, 15 + 16) != Z_OK) return false;
    zs.next_in = const_cast<Bytef*>(in);
    zs.avail_in = (uInt)n;
    uint8_t buf[65536];
    int rc;
    do {
        zs.next_out = buf; zs.avail_out = sizeof(buf);
        rc = inflate(&zs, Z_NO_FLUSH);
        if (rc != Z_OK && rc != Z_STREAM_END) { inflateEnd(&zs); return false; }
        out->insert(out->end(), buf, buf + (sizeof(buf) - zs.avail_out));
    } while (rc != Z_STREAM_END);
    inflateEnd(&zs);
    return true;
}

// BlockHeader, src/lib.rs:37-47, bincode fixed-int LE, 32 bytes
static void put_header(const SubHeader& h, std::vector<uint8_t>* out) {
    uint8_t b[32];
    memset(b, 0, 32);
    memcpy(b + 0, &h.block_size, 4);
    memcpy(b + 4, &h.num_records, 4);
    memcpy(b + 8, &h.num_u64, 4);
    memcpy(b + 12, &h.encoded_size, 4);
    memcpy(b + 16, &h.rice_param, 8);
    b[24] = h.bitpacker_exponent;
    out->insert(out->end(), b, b + 32);
}
static void get_header(const uint8_t* b, SubHeader* h) {
    memcpy(&h->block_size, b + 0, 4);
    memcpy(&h->num_records, b + 4, 4);
    memcpy(&h->num_u64, b + 8, 4);
    memcpy(&h->encoded_size, b + 12, 4);
    memcpy(&h->rice_param, b + 16, 8);
    h->bitpacker_exponent = b[24];
}

bool code_column(const std::vector<uint64_t>& ints, bool rice, std::vector<uint64_t>* words, uint64_t* param) {
    return rice ? rice_encode(ints, words, param) : minimal_binary_encode(ints, words, param);
}

static bool compress_block(const std::vector<uint64_t>& ints, uint64_t num_records, bool rice, std::vector<uint8_t>* out) {
    std::vector<uint64_t> words;
    uint64_t param = 0;
    if (!code_column(ints, rice, &words, &param)) return false;
    std::vector<uint8_t> gz;
    if (!gzip_bytes((const uint8_t*)words.data(), words.size() * 8, &gz)) return false;
    SubHeader h;
    h.block_size = (uint32_t)gz.size();
    h.num_records = (uint32_t)num_records;
    h.num_u64 = (uint32_t)ints.size();
    h.encoded_size = (uint32_t)words.size();
    h.rice_param = param;
    h.bitpacker_exponent = 8;                                                   // encode.rs:118
    put_header(h, out);
    out->insert(out->end(), gz.begin(), gz.end());
    return true;
}

// write_block_to: nothing is written unless all four columns code successfully (Q4).
bool write_block(const uint64_t* rec, uint64_t n, uint64_t num_records, std::vector<uint8_t>* sink) {
    Columns c;
    if (!split_encoded_dictionary(rec, n, &c)) return false;
    std::vector<uint8_t> b1, b2, b3, b4;
    if (!compress_block(c.colex, num_records, false, &b1)) return false;        // lib.rs:242
    if (!compress_block(c.len, num_records, true, &b2)) return false;
    if (!compress_block(c.flags, num_records, true, &b3)) return false;
    if (!compress_block(c.literals, num_records, false, &b4)) return false;
    sink->insert(sink->end(), b1.begin(), b1.end());
    sink->insert(sink->end(), b2.begin(), b2.end());
    sink->insert(sink->end(), b3.begin(), b3.end());
    sink->insert(sink->end(), b4.begin(), b4.end());
    return true;
}

// decode_block up to (not including) decode_sequence: returns bytes consumed, 0 on EOF, <0 on error.
int64_t read_block(const uint8_t* data, uint64_t size, std::vector<uint64_t>* rec, SubHeader hdrs[4],
                   std::vector<uint64_t> inflated[4]) {
    if (size < 32) return 0;                                                    // lib.rs:327 read_exact fails
    uint64_t pos = 0;
    Columns c;
    std::vector<uint64_t>* cols[4] = {&c.colex, &c.len, &c.flags, &c.literals};
    for (int q = 0; q < 4; q++) {
        if (pos + 32 > size) return -1;
        SubHeader h;
        get_header(data + pos, &h);
        pos += 32;
        if (pos + h.block_size > size) return -2;
        std::vector<uint8_t> raw;
        if (!gunzip_bytes(data + pos, h.block_size, &raw)) return -3;
        pos += h.block_size;
        if (raw.size() % 8) return -4;
        std::vector<uint64_t> words(raw.size() / 8);
        memcpy(words.data(), raw.data(), raw.size());
        if (words.size() != h.encoded_size) return -5;                          // decode.rs:90
        bool ok = (q == 1 || q == 2) ? rice_decode(words.data(), words.size(), h.num_u64, h.rice_param, cols[q])
                                     : minimal_binary_decode(words.data(), words.size(), h.num_u64, h.rice_param, cols[q]);
        if (!ok) return -6;
        if (hdrs) hdrs[q] = h;
        if (inflated) inflated[q] = std::move(words);
    }
    int z = zip_block_contents(c, rec);
    if (z < 0) return -10 + z;
    return (int64_t)pos;
}

}  // namespace orc
