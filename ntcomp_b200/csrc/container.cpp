// container.cpp -- host side of ntcomp's block container ("next" row f1 of SURVEY §8, host first):
//   column split          src/encode.rs:168-229      zip            src/decode.rs:102-149
//   Rice / minimal binary src/encode.rs:59-94        decoders       src/decode.rs:51-77
//   32-byte headers       src/lib.rs:37-47,75-92     gzip members   src/encode.rs:49-57, decode.rs:41-49
//   write_block_to        src/lib.rs:232-252         decode_block   src/lib.rs:320-363 (up to decode_sequence)
// Bit-stream conventions restate dsi-bitstream 0.5.0 (BufBitWriter<BE, MemWordWriterVec<u64>>): bits are
// written MSB first into u64 words that are stored big-endian; unary(n) = n zeros then a one.
// DEFLATE comes from the system zlib (level 6, gzip wrapper, mtime 0, OS 255); the reference uses
// zlib-rs, so payload BYTES may differ while every member inflates to the same words.
#include <zlib.h>

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/ntcomp_b200.h"
#include "container.h"

namespace ntc {
void set_error(const std::string& msg);

// ---- MSB-first bit packer over big-endian u64 words ------------------------------------------
namespace {
struct BitPacker {
    std::vector<uint64_t>& words;
    uint64_t acc = 0;   // bits collected for the current word, left-aligned
    unsigned fill = 0;  // number of valid bits in acc
    explicit BitPacker(std::vector<uint64_t>& w) : words(w) {}
    inline void put(uint64_t value, unsigned nbits) {  // nbits <= 64, value < 2^nbits
        while (nbits) {
            unsigned room = 64 - fill, take = nbits < room ? nbits : room;
            uint64_t part = take == 64 ? value : ((value >> (nbits - take)) & ((1ull << take) - 1));
            acc |= take == 64 ? part : (part << (room - take));
            fill += take;
            nbits -= take;
            if (fill == 64) { words.push_back(__builtin_bswap64(acc)); acc = 0; fill = 0; }
        }
    }
    inline void zeros(uint64_t n) {
        while (n) { unsigned take = (unsigned)std::min<uint64_t>(n, 64); put(0, take); n -= take; }
    }
    void finish() { if (fill) { words.push_back(__builtin_bswap64(acc)); acc = 0; fill = 0; } }
};

// MSB-first reader over big-endian u64 words: a 64-bit window refilled word-wise, unary runs by clz
struct BitCursor {
    const uint64_t* words; uint64_t n_words; uint64_t bit = 0;
    inline uint64_t word(uint64_t w) const { return w < n_words ? __builtin_bswap64(words[w]) : 0; }
    // the next 64 bits of the stream, left-aligned (zero-padded past the end)
    inline uint64_t window() const {
        const uint64_t w = bit >> 6;
        const unsigned sh = (unsigned)(bit & 63);
        uint64_t hi = word(w);
        return sh ? (hi << sh) | (word(w + 1) >> (64 - sh)) : hi;
    }
    inline bool take(unsigned nbits, uint64_t* v) {                    // nbits <= 64
        if (nbits == 0) { *v = 0; return true; }
        if (bit + nbits > n_words * 64) return false;
        *v = window() >> (64 - nbits);
        bit += nbits;
        return true;
    }
    inline bool unary(uint64_t* q) {                                   // zeros up to and including the terminating one
        uint64_t zeros = 0;
        for (;;) {
            if (bit >= n_words * 64) return false;
            uint64_t win = window();
            if (win) {
                unsigned z = (unsigned)__builtin_clzll(win);
                if (bit + z >= n_words * 64) return false;
                zeros += z;
                bit += z + 1;
                *q = zeros;
                return true;
            }
            uint64_t room = n_words * 64 - bit;
            uint64_t step = room < 64 ? room : 64;
            zeros += step;
            bit += step;
        }
    }
};
}  // namespace

// dsi_bitstream::codes::rice::log2_b: ceil(log2(-ln(golden ratio) / ln(1 - p))), saturating cast to usize
uint64_t rice_parameter(uint64_t n_values, uint64_t sum) {
    double p = std::exp(std::log((double)n_values) - std::log((double)sum));     // src/encode.rs:62
    double v = std::ceil(std::log2(-std::log((std::sqrt(5.0) + 1.0) / 2.0) / std::log1p(-p)));
    if (!(v > 0.0)) return 0;
    return v > 63.0 ? 63 : (uint64_t)v;
}

void minimal_binary_shape(uint64_t upper, unsigned* l, uint64_t* limit) {
    *l = 63 - (unsigned)__builtin_clzll(upper);
    *limit = ((1ull << *l) << 1) - upper;
}

bool code_column_host(const uint64_t* v, uint64_t n, int rice, std::vector<uint64_t>* words, uint64_t* param) {
    words->clear();
    BitPacker bp(*words);
    if (rice) {
        uint64_t sum = 0;
        for (uint64_t i = 0; i < n; i++) sum += v[i];
        const uint64_t b = rice_parameter(n, sum);
        for (uint64_t i = 0; i < n; i++) {
            bp.zeros(v[i] >> b);
            bp.put(1, 1);
            if (b) bp.put(v[i] & ((1ull << b) - 1), (unsigned)b);
        }
        *param = b;
    } else {
        if (n == 0) return false;                                                  // src/encode.rs:80 (Q4)
        uint64_t mx = 0;
        for (uint64_t i = 0; i < n; i++) mx = std::max(mx, v[i]);
        const uint64_t upper = mx + 2;                                             // src/encode.rs:82
        unsigned l; uint64_t limit;
        minimal_binary_shape(upper, &l, &limit);
        for (uint64_t i = 0; i < n; i++) {
            uint64_t x = v[i] + 1;                                                 // src/encode.rs:88
            if (x < limit) { if (l) bp.put(x, l); } else bp.put(x + limit, l + 1);
        }
        *param = upper;
    }
    bp.finish();
    return true;
}

bool decode_column_host(const uint64_t* words, uint64_t n_words, uint64_t n, int rice, uint64_t param, std::vector<uint64_t>* out) {
    BitCursor cur{words, n_words};
    out->clear();
    // header fields come from the file: the same checks as ntc_block_read_gpu (api_columns.inc) before anything is sized by
    // them.  Every code is at least one bit, so n values need n bits (minimal binary with upper bound 2 or 3 has 1-bit
    // codes; a bound below 2 would mean 0-bit codes and `x - 1` underflowing at src/decode.rs:74); a Rice parameter above
    // 63 cannot be a shift count.
    if (n > n_words * 64) return false;
    if (rice ? param > 63 : param < 2) return false;
    out->reserve(n);
    if (rice) {
        for (uint64_t i = 0; i < n; i++) {
            uint64_t q = 0, rem = 0;
            if (!cur.unary(&q)) return false;
            if (param && !cur.take((unsigned)param, &rem)) return false;
            out->push_back((q << param) | rem);
        }
    } else {
        unsigned l; uint64_t limit;
        minimal_binary_shape(param, &l, &limit);
        for (uint64_t i = 0; i < n; i++) {
            uint64_t x = 0;
            if (l && !cur.take(l, &x)) return false;
            if (x >= limit) { uint64_t b; if (!cur.take(1, &b)) return false; x = ((x << 1) | b) - limit; }
            out->push_back(x - 1);                                                 // src/decode.rs:74
        }
    }
    return true;
}

// ---- column split / zip ------------------------------------------------------------------------
void split_records(const uint64_t* rec, uint64_t n, BlockColumns* c) {
    c->colex.clear(); c->len.clear(); c->flags.clear(); c->literals.clear();
    uint64_t cur = 0;   // literal word being filled, 31 bases per word (src/encode.rs:224)
    unsigned held = 0;
    for (uint64_t i = 0; i < n; i++) {
        const uint64_t w = rec[i];
        const uint32_t flag = (uint32_t)(w >> 56);
        c->flags.push_back(flag);
        if ((flag & 2u) == 0) {
            c->colex.push_back(w & 0xFFFFFFFFull);
            c->len.push_back((w >> 32) & 0xFFFFFFull);
        } else {
            unsigned L = std::min<unsigned>(flag >> 2, 32);
            const uint64_t payload = w & 0x00FFFFFFFFFFFFFFull;                    // 7 bytes; bases 28..31 read as A
            for (unsigned j = 0; j < L; j++) {
                cur |= ((payload >> (2 * j)) & 3ull) << (2 * held);
                if (++held == 31) { c->literals.push_back(cur); cur = 0; held = 0; }
            }
        }
    }
    if (held) c->literals.push_back(cur);
}

int zip_columns(const BlockColumns& c, std::vector<uint64_t>* rec, bool full_last_word) {
    if (c.colex.size() != c.len.size()) return -1;                                // src/decode.rs:108
    uint64_t lit_bases = 0;
    for (uint64_t f : c.flags) lit_bases += (f & 0xFC) >> 2;
    // src/decode.rs:114-118: the last literal word is taken to hold lit_bases % 31 bases (Q5)
    auto base_at = [&](uint64_t pos, bool* ok) -> uint64_t {
        uint64_t wi = pos / 31, off = pos % 31;
        uint64_t in_word = wi + 1 == c.literals.size() ? lit_bases % 31 : 31;
        // full_last_word: a block whose literal bases are a non-zero multiple of 31 has a full last word; the
        // reference reads it as empty and panics at src/decode.rs:138 (Q5) -- the file pipeline decodes it
        if (full_last_word && in_word == 0) in_word = 31;
        if (wi >= c.literals.size() || off >= in_word) { *ok = false; return 0; }
        return (c.literals[wi] >> (2 * off)) & 3ull;
    };
    uint64_t ri = 0, lp = 0;
    rec->clear();
    for (uint64_t f : c.flags) {
        const uint64_t flag = f & 0xFF;
        uint64_t w = 0;
        if ((flag & 2) == 0) {
            if (ri >= c.colex.size()) return -2;
            w = (c.colex[ri] & 0xFFFFFFFFull) | ((c.len[ri] & 0xFFFFFFull) << 32);
            ri++;
        } else {
            uint64_t L = flag >> 2;
            if (L > 32) return -4;                                                // as_2bit rejects > 32 bases (decode.rs:140 unwrap)
            bool ok = true;
            for (uint64_t j = 0; j < L; j++) w |= base_at(lp + j, &ok) << (2 * j);
            if (!ok) return -3;                                                   // src/decode.rs:138 panics (Q5)
            w &= 0x00FFFFFFFFFFFFFFull;                                            // only 7 bytes are kept (decode.rs:142)
            lp += L;
        }
        rec->push_back(w | (flag << 56));
    }
    return 0;
}

// ---- gzip ----------------------------------------------------------------------------------------
// flate2's Compression::default() is level 6 (src/encode.rs:53); NTC_GZIP_LEVEL (1..9) trades size for speed in experiments
static int gzip_level() {
    static const int level = [] {
        const char* env = getenv("NTC_GZIP_LEVEL");
        const int v = env ? atoi(env) : 6;
        return v >= 1 && v <= 9 ? v : 6;
    }();
    return level;
}
static bool gzip_member(const uint8_t* in, size_t n, std::vector<uint8_t>* out) {
    z_stream zs;
    memset(&zs, 0, sizeof(zs));
    if (deflateInit2(&zs, gzip_level(), Z_DEFLATED, 15 + 16, 8, Z_DEFAULT_STRATEGY) != Z_OK) return false;
    gz_header hd;
    memset(&hd, 0, sizeof(hd));
    hd.os = 255;
    deflateSetHeader(&zs, &hd);
    out->resize(deflateBound(&zs, n) + 64);
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
// limit: stop once more than this many bytes came out (the header's encoded_size says how many there must be; a
// hostile member must not inflate without bound)
static bool gunzip_member(const uint8_t* in, size_t n, std::vector<uint8_t>* out, uint64_t limit) {
    z_stream zs;
    memset(&zs, 0, sizeof(zs));
    if (inflateInit2(&zs, 15 + 16) != Z_OK) return false;
    zs.next_in = const_cast<Bytef*>(in);
    zs.avail_in = (uInt)n;
    out->clear();
    std::vector<uint8_t> buf(1 << 16);
    int rc;
    do {
        zs.next_out = buf.data();
        zs.avail_out = (uInt)buf.size();
        rc = inflate(&zs, Z_NO_FLUSH);
        if (rc != Z_OK && rc != Z_STREAM_END) { inflateEnd(&zs); return false; }
        out->insert(out->end(), buf.data(), buf.data() + (buf.size() - zs.avail_out));
        if (out->size() > limit) { inflateEnd(&zs); return false; }
        if (rc == Z_OK && zs.avail_in == 0 && zs.avail_out != 0) { inflateEnd(&zs); return false; }  // truncated member
    } while (rc != Z_STREAM_END);
    inflateEnd(&zs);
    return true;
}

void put_sub_block(const CodedColumn& col, uint64_t num_records, std::vector<uint8_t>* sink, bool* ok) {
    std::vector<uint8_t> gz;
    if (!gzip_member((const uint8_t*)col.words.data(), col.words.size() * 8, &gz)) { *ok = false; return; }
    uint8_t h[32];
    memset(h, 0, 32);
    uint32_t v32;
    v32 = (uint32_t)gz.size(); memcpy(h + 0, &v32, 4);               // block_size
    v32 = (uint32_t)num_records; memcpy(h + 4, &v32, 4);             // num_records (as passed, Q6)
    v32 = (uint32_t)col.n_values; memcpy(h + 8, &v32, 4);            // num_u64
    v32 = (uint32_t)col.words.size(); memcpy(h + 12, &v32, 4);       // encoded_size
    memcpy(h + 16, &col.param, 8);                                   // rice_param
    h[24] = 8;                                                       // bitpacker_exponent, src/encode.rs:118
    sink->insert(sink->end(), h, h + 32);
    sink->insert(sink->end(), gz.begin(), gz.end());
}

// the same sub-block from a plain word pointer straight into a caller buffer (one gzip task of the file pipeline)
int32_t put_sub_block_raw(const uint64_t* words, uint64_t n_words, uint64_t n_values, uint64_t param, uint64_t num_records, uint8_t* out,
                          uint64_t cap, uint64_t* n_bytes) {
    *n_bytes = 0;
    if (cap < 32 + 64) return -2;
    z_stream zs;
    memset(&zs, 0, sizeof(zs));
    if (deflateInit2(&zs, gzip_level(), Z_DEFLATED, 15 + 16, 8, Z_DEFAULT_STRATEGY) != Z_OK) return -1;
    gz_header hd;
    memset(&hd, 0, sizeof(hd));
    hd.os = 255;
    deflateSetHeader(&zs, &hd);
    if (n_words * 8 > 0xFFFFFFF0ull || cap - 32 > 0xFFFFFFF0ull) { deflateEnd(&zs); return -1; }
    zs.next_in = const_cast<Bytef*>((const Bytef*)words);
    zs.avail_in = (uInt)(n_words * 8);
    zs.next_out = out + 32;
    zs.avail_out = (uInt)(cap - 32);
    const int rc = deflate(&zs, Z_FINISH);
    const uint64_t gz = zs.total_out;
    deflateEnd(&zs);
    if (rc != Z_STREAM_END) return rc == Z_OK || rc == Z_BUF_ERROR ? -2 : -1;
    uint8_t* h = out;
    memset(h, 0, 32);
    uint32_t v32;
    v32 = (uint32_t)gz; memcpy(h + 0, &v32, 4);
    v32 = (uint32_t)num_records; memcpy(h + 4, &v32, 4);
    v32 = (uint32_t)n_values; memcpy(h + 8, &v32, 4);
    v32 = (uint32_t)n_words; memcpy(h + 12, &v32, 4);
    memcpy(h + 16, &param, 8);
    h[24] = 8;
    *n_bytes = 32 + gz;
    return 0;
}

// header fields + gunzip of one sub-block, without decoding the integer codes (the GPU column decoder's input)
int64_t inflate_sub_block(const uint8_t* data, uint64_t size, std::vector<uint64_t>* words, uint64_t* n_values, uint64_t* param) {
    if (size < 32) return -1;
    uint32_t block_size, num_u64, encoded_size;
    memcpy(&block_size, data, 4);
    memcpy(&num_u64, data + 8, 4);
    memcpy(&encoded_size, data + 12, 4);
    memcpy(param, data + 16, 8);
    if (32ull + block_size > size) return -2;
    std::vector<uint8_t> raw;
    if (!gunzip_member(data + 32, block_size, &raw, (uint64_t)encoded_size * 8)) return -3;
    if (raw.size() % 8 || raw.size() / 8 != encoded_size) return -5;              // src/decode.rs:90
    words->resize(raw.size() / 8);
    if (!raw.empty()) memcpy(words->data(), raw.data(), raw.size());
    *n_values = num_u64;
    return 32 + (int64_t)block_size;
}

int64_t read_sub_block(const uint8_t* data, uint64_t size, int rice, std::vector<uint64_t>* values) {
    if (size < 32) return -1;
    uint32_t block_size, num_u64, encoded_size;
    uint64_t param;
    memcpy(&block_size, data, 4);
    memcpy(&num_u64, data + 8, 4);
    memcpy(&encoded_size, data + 12, 4);
    memcpy(&param, data + 16, 8);
    if (32ull + block_size > size) return -2;
    std::vector<uint8_t> raw;
    if (!gunzip_member(data + 32, block_size, &raw, (uint64_t)encoded_size * 8)) return -3;
    if (raw.size() % 8 || raw.size() / 8 != encoded_size) return -5;              // src/decode.rs:90
    std::vector<uint64_t> words(raw.size() / 8);
    memcpy(words.data(), raw.data(), raw.size());
    if (!decode_column_host(words.data(), words.size(), num_u64, rice, param, values)) return -6;
    return 32 + (int64_t)block_size;
}

}  // namespace ntc

using namespace ntc;

extern "C" {

uint64_t ntc_block_bound(uint64_t n_records) { return 4 * 32 + 4 * 64 + n_records * 20 + (n_records * 20) / 1000; }

// No exception crosses the C ABI: allocation failures (a hostile header can announce gigabytes) and anything else
// thrown below come back as a status code.
#define NTC_ABI_GUARD_BEGIN try {
#define NTC_ABI_GUARD_END                                                                                   \
    } catch (const std::bad_alloc&) { set_error("out of memory"); return NTC_ERR_IO;                        \
    } catch (const std::exception& e) { set_error(std::string("internal error: ") + e.what()); return NTC_ERR_IO; \
    } catch (...) { set_error("internal error"); return NTC_ERR_IO; }

int32_t ntc_column_code(const uint64_t* values, uint64_t n, int32_t rice, uint64_t* words, uint64_t cap, uint64_t* n_words,
                        uint64_t* param) {
    if ((n && !values) || !n_words || !param) { set_error("null argument"); return NTC_ERR_INVALID_ARG; }
    NTC_ABI_GUARD_BEGIN
    std::vector<uint64_t> w;
    if (!code_column_host(values, n, rice, &w, param)) { set_error("empty column (minimal_binary_encode, src/encode.rs:80)"); return NTC_ERR_EMPTY_COLUMN; }
    *n_words = w.size();
    if (w.size() > cap) { set_error("words buffer too small"); return NTC_ERR_CAPACITY; }
    if (!w.empty()) memcpy(words, w.data(), w.size() * 8);
    return NTC_OK;
    NTC_ABI_GUARD_END
}

int32_t ntc_block_write(const uint64_t* records, uint64_t n_records, uint64_t num_reads, uint8_t* out, uint64_t cap, uint64_t* n_bytes) {
    if (!records || !out || !n_bytes) { set_error("null argument"); return NTC_ERR_INVALID_ARG; }
    *n_bytes = 0;
    if (n_records == 0) { set_error("empty block (EncodeError, src/encode.rs:171)"); return NTC_ERR_EMPTY_READ; }
    NTC_ABI_GUARD_BEGIN
    BlockColumns c;
    split_records(records, n_records, &c);
    const std::vector<uint64_t>* cols[4] = {&c.colex, &c.len, &c.flags, &c.literals};
    const int rice[4] = {0, 1, 1, 0};                                               // src/lib.rs:242-245
    CodedColumn coded[4];
    for (int q = 0; q < 4; q++) {
        coded[q].n_values = cols[q]->size();
        if (!code_column_host(cols[q]->data(), cols[q]->size(), rice[q], &coded[q].words, &coded[q].param)) {
            // Q4: the reference returns Err before writing anything and main.rs:170 drops the block
            set_error("block has no literal record or no reference record: the reference drops it (src/encode.rs:80)");
            return NTC_ERR_EMPTY_COLUMN;
        }
    }
    std::vector<uint8_t> sink;
    bool ok = true;
    for (int q = 0; q < 4; q++) put_sub_block(coded[q], num_reads, &sink, &ok);
    if (!ok) { set_error("deflate failed"); return NTC_ERR_IO; }
    if (sink.size() > cap) { set_error("output buffer too small"); return NTC_ERR_CAPACITY; }
    memcpy(out, sink.data(), sink.size());
    *n_bytes = sink.size();
    return NTC_OK;
    NTC_ABI_GUARD_END
}

static int32_t block_read_impl(const uint8_t* data, uint64_t size, uint64_t* records, uint64_t cap, uint64_t* n_records, uint64_t* consumed,
                               bool full_last_word) {
    if (!n_records || !consumed || (size && !data)) { set_error("null argument"); return NTC_ERR_INVALID_ARG; }
    *n_records = 0;
    *consumed = 0;
    if (size < 32) return NTC_OK;                                                   // EOF: read_exact fails, src/lib.rs:327
    NTC_ABI_GUARD_BEGIN
    BlockColumns c;
    std::vector<uint64_t>* cols[4] = {&c.colex, &c.len, &c.flags, &c.literals};
    const int rice[4] = {0, 1, 1, 0};
    uint64_t pos = 0;
    for (int q = 0; q < 4; q++) {
        int64_t used = read_sub_block(data + pos, size - pos, rice[q], cols[q]);
        if (used < 0) { set_error("malformed sub-block " + std::to_string(q) + " (code " + std::to_string(used) + ")"); return NTC_ERR_IO; }
        pos += (uint64_t)used;
    }
    std::vector<uint64_t> rec;
    int z = zip_columns(c, &rec, full_last_word);
    if (z < 0) { set_error("columns do not zip (code " + std::to_string(z) + "; src/decode.rs:108,138)"); return NTC_ERR_IO; }
    if (rec.size() > cap) { set_error("records buffer too small"); return NTC_ERR_CAPACITY; }
    if (!rec.empty()) memcpy(records, rec.data(), rec.size() * 8);
    *n_records = rec.size();
    *consumed = pos;
    return NTC_OK;
    NTC_ABI_GUARD_END
}

int32_t ntc_block_read(const uint8_t* data, uint64_t size, uint64_t* records, uint64_t cap, uint64_t* n_records, uint64_t* consumed) {
    return block_read_impl(data, size, records, cap, n_records, consumed, false);
}

int32_t ntc_block_read_ex(const uint8_t* data, uint64_t size, uint64_t* records, uint64_t cap, uint64_t* n_records, uint64_t* consumed,
                          int32_t flags) {
    return block_read_impl(data, size, records, cap, n_records, consumed, (flags & NTC_BLOCK_FULL_LAST_LITERAL_WORD) != 0);
}

}  // extern "C"
