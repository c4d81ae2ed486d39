// pipeline.cpp -- host pipeline around the hot path ("next" row f3 of SURVEY §8): the driver loops of
// `ntcomp encode` (src/main.rs:141-181) and `ntcomp decode` (src/main.rs:183-211) as overlapped stages
//   encode: fastX parse (reader thread) -> ntc_encode_batch on the GPU -> write_block_to per 65,536-read
//           block on a thread pool (column split, Rice / minimal binary, gzip) -> in-order writer
//   decode: sub-block framing (reader) -> gunzip + Rice / minimal-binary decode + zip per block on a
//           thread pool -> ntc_decode_fasta on the GPU -> in-order writer
// Everything here is host orchestration above the C ABI of include/ntcomp_b200.h; no kernel is launched
// from this file.  needletail's parser (src/main.rs:158-163, normalize(true)) is restated for plain and
// gzip-compressed FASTA / FASTQ; bytes other than ACGT after normalisation make the encoder return
// NTC_ERR_BAD_BASE (the reference underflows at src/lib.rs:207).
#include <zlib.h>

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <exception>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/ntcomp_b200.h"

namespace ntc {
void set_error(const std::string& msg);
}
using ntc::set_error;

namespace {

constexpr uint64_t kBlockReads = 65536;            // src/main.rs:152
constexpr uint64_t kBatchBlocksDefault = 8;        // blocks handed to the GPU per call (stage overlap granularity)
constexpr uint64_t kBatchBases = 768ull << 20;     // ... or this many bases, whichever comes first

uint64_t batch_blocks() {                          // NTC_PIPELINE_BATCH_BLOCKS overrides (tests: several groups from small files)
    if (const char* env = getenv("NTC_PIPELINE_BATCH_BLOCKS")) {
        unsigned long long v = strtoull(env, nullptr, 10);
        if (v >= 1 && v <= 4096) return v;
    }
    return kBatchBlocksDefault;
}

// ---- bounded single-producer / single-consumer queue -------------------------------------------
template <typename T>
class Channel {
public:
    explicit Channel(size_t depth) : depth_(depth) {}
    bool push(T&& v) {                               // false once the consumer has closed the channel
        std::unique_lock<std::mutex> lk(m_);
        cv_.wait(lk, [&] { return q_.size() < depth_ || closed_; });
        if (closed_) return false;
        q_.push_back(std::move(v));
        cv_.notify_all();
        return true;
    }
    bool pop(T* out) {                               // false at end of stream
        std::unique_lock<std::mutex> lk(m_);
        cv_.wait(lk, [&] { return !q_.empty() || done_ || closed_; });
        if (q_.empty()) return false;
        *out = std::move(q_.front());
        q_.pop_front();
        cv_.notify_all();
        return true;
    }
    void finish() { std::lock_guard<std::mutex> lk(m_); done_ = true; cv_.notify_all(); }   // producer: no more items
    void close() { std::lock_guard<std::mutex> lk(m_); closed_ = true; q_.clear(); cv_.notify_all(); }  // consumer: stop
private:
    std::mutex m_;
    std::condition_variable cv_;
    std::deque<T> q_;
    size_t depth_;
    bool done_ = false, closed_ = false;
};

double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

// runs fn(i) for i in [0, n) on up to n_threads threads
void parallel_for(uint64_t n, int n_threads, const std::function<void(uint64_t)>& fn) {
    int t = (int)std::min<uint64_t>(n, (uint64_t)std::max(1, n_threads));
    if (t <= 1) { for (uint64_t i = 0; i < n; i++) fn(i); return; }
    std::vector<std::thread> pool;
    std::mutex m;
    uint64_t next = 0;
    std::exception_ptr failed;                       // an exception must not leave a std::thread (std::terminate)
    for (int w = 0; w < t; w++)
        pool.emplace_back([&] {
            for (;;) {
                uint64_t i;
                { std::lock_guard<std::mutex> lk(m); if (next >= n || failed) return; i = next++; }
                try { fn(i); } catch (...) { std::lock_guard<std::mutex> lk(m); if (!failed) failed = std::current_exception(); }
            }
        });
    for (auto& th : pool) th.join();
    if (failed) std::rethrow_exception(failed);
}

// body of a pipeline stage thread: whatever it throws (bad_alloc on a hostile header, ...) becomes a flag the
// calling thread turns into NTC_ERR_IO once the stages are joined; on_fail unblocks the neighbours
template <typename Body, typename OnFail>
void stage_guard(bool* fatal, Body&& body, OnFail&& on_fail) {
    try { body(); } catch (...) { *fatal = true; on_fail(); }
}

// ---- output sink: a file, or stdout for "-" -------------------------------------------------------
struct Sink {
    FILE* f = nullptr;
    bool own = false;
    bool open(const char* path) {
        if (!path || strcmp(path, "-") == 0) { f = stdout; own = false; return true; }
        f = fopen(path, "wb");
        own = true;
        return f != nullptr;
    }
    bool write(const void* p, size_t n) { return n == 0 || fwrite(p, 1, n, f) == n; }
    void close() { if (f) { fflush(f); if (own) fclose(f); f = nullptr; } }
    ~Sink() { close(); }
};

// ---- fastX reader (plain or gzip; FASTA with wrapped lines, FASTQ with 4-line or wrapped records) ----
// Reads the stream in 4 MB pieces and walks lines with memchr; sequence bytes are copied line by line
// and normalised in place (needletail normalize(true) for the bytes that can reach the encoder:
// upper-case, U -> T, white space dropped; everything else passes through and is rejected by the encoder
// unless it is A, C, G or T).
class FastxReader {
public:
    bool open(const char* path) {
        f_ = (strcmp(path, "-") == 0) ? gzdopen(0, "rb") : gzopen(path, "rb");
        if (!f_) return false;
        gzbuffer(f_, 1 << 20);
        buf_.resize(4 << 20);
        return true;
    }
    ~FastxReader() { if (f_) gzclose(f_); }
    // appends the next record's normalised sequence to *bases; 1 = got one, 0 = end of file, -1 = malformed
    int next(std::vector<uint8_t>* bases) {
        const uint8_t* ln;
        size_t n;
        do { if (!line(&ln, &n)) return 0; } while (n == 0);            // blank lines between records
        if (ln[0] == '>') {
            while (peek() >= 0 && peek() != '>') {
                if (!line(&ln, &n)) break;
                append_normalised(ln, n, bases);
            }
            return 1;
        }
        if (ln[0] == '@') {
            size_t want = 0, got = 0;                                  // quality has one byte per sequence byte
            for (;;) {                                                 // sequence lines up to the '+' separator
                if (!line(&ln, &n)) return -1;
                if (n && ln[0] == '+') break;
                want += append_normalised(ln, n, bases);
            }
            while (got < want) {
                if (!line(&ln, &n)) return -1;
                got += n;
            }
            return got == want ? 1 : -1;
        }
        return -1;
    }

private:
    gzFile f_ = nullptr;
    std::vector<uint8_t> buf_;
    size_t lo_ = 0, hi_ = 0;      // unread bytes are buf_[lo_, hi_)
    bool eof_ = false;

    bool fill() {                 // keeps the unread tail, appends more input; false when nothing was added
        if (eof_) return false;
        if (lo_ > 0) { memmove(buf_.data(), buf_.data() + lo_, hi_ - lo_); hi_ -= lo_; lo_ = 0; }
        if (hi_ == buf_.size()) buf_.resize(buf_.size() * 2);         // a single line longer than the buffer
        int got = gzread(f_, buf_.data() + hi_, (unsigned)std::min<size_t>(buf_.size() - hi_, 1u << 30));
        if (got <= 0) { eof_ = true; return false; }
        hi_ += (size_t)got;
        return true;
    }
    int peek() {                  // first byte of the next line, -1 at end of input
        if (lo_ == hi_ && !fill()) return -1;
        return buf_[lo_];
    }
    bool line(const uint8_t** p, size_t* n) {                          // next line without its "\n" / "\r\n"
        for (;;) {
            const uint8_t* nl = hi_ > lo_ ? (const uint8_t*)memchr(buf_.data() + lo_, '\n', hi_ - lo_) : nullptr;
            if (nl) {
                *p = buf_.data() + lo_;
                *n = (size_t)(nl - *p);
                lo_ += *n + 1;
                break;
            }
            if (!fill()) {                                             // last line without a newline
                if (lo_ == hi_) return false;
                *p = buf_.data() + lo_;
                *n = hi_ - lo_;
                lo_ = hi_;
                break;
            }
        }
        while (*n && ((*p)[*n - 1] == '\r')) (*n)--;
        return true;
    }
    size_t append_normalised(const uint8_t* __restrict__ s, size_t n, std::vector<uint8_t>* out) {
        const size_t at = out->size();
        out->insert(out->end(), s, s + n);                             // plain copy; upper-case ACGT lines need nothing else
        // anything to fix?  lower case / space have bit 5 set, tab and U are looked for with the zero-byte trick;
        // eight bytes per step (the byte-wise loop did not vectorise: 2 ns per base)
        uint64_t odd = 0;
        size_t i = 0;
        const uint64_t k01 = 0x0101010101010101ull, k80 = 0x8080808080808080ull;
        for (; i + 8 <= n; i += 8) {
            uint64_t w;
            memcpy(&w, s + i, 8);
            const uint64_t t = w ^ (k01 * 0x09), u = w ^ (k01 * 0x55);
            odd |= (w & (k01 * 0x20)) | ((t - k01) & ~t & k80) | ((u - k01) & ~u & k80);
        }
        for (; i < n; i++) odd |= (uint64_t)(s[i] & 0x20u) | (uint64_t)(s[i] == '\t') | (uint64_t)(s[i] == 'U');
        if (!odd) return n;
        uint8_t* d = out->data() + at;
        size_t w = 0;
        for (size_t i = 0; i < n; i++) {
            uint8_t ch = d[i];
            if (ch == ' ' || ch == '\t') continue;
            if (ch >= 'a' && ch <= 'z') ch = (uint8_t)(ch - 32);
            if (ch == 'U') ch = 'T';
            d[w++] = ch;
        }
        out->resize(at + w);
        return w;
    }
};

// one GPU call's worth of reads / one group of blocks travelling through the stages
struct ReadBatch {
    std::vector<uint8_t> bases;
    std::vector<uint64_t> offsets;   // n_reads + 1
    int err = 0;                     // reader error (malformed input)
};
struct BytesBatch {
    std::vector<std::vector<uint8_t>> parts;   // written in order
    int32_t rc = NTC_OK;
    std::string msg;
};

}  // namespace

extern "C" {

// `ntcomp encode` (src/main.rs:141-181): 32-byte zero file header, then one block per 65,536 reads.
// stats (optional): [0] reads, [1] bases, [2] records, [3] bytes written, [4] blocks written,
// [5] blocks dropped because a column was empty (the reference drops them too: src/main.rs:170, Q4).
int32_t ntc_encode_file(ntc_ctx* ctx, const char* fastx_path, const char* out_path, int32_t n_threads, uint64_t* stats) {
    if (!ctx || !fastx_path) { set_error("null argument"); return NTC_ERR_INVALID_ARG; }
    if (n_threads <= 0) n_threads = (int32_t)std::max(1u, std::thread::hardware_concurrency());
    auto reader = std::make_shared<FastxReader>();
    if (!reader->open(fastx_path)) { set_error(std::string("cannot open ") + fastx_path); return NTC_ERR_IO; }
    Sink sink;
    if (!sink.open(out_path)) { set_error(std::string("cannot open ") + (out_path ? out_path : "-")); return NTC_ERR_IO; }
    uint64_t st[6] = {0, 0, 0, 0, 0, 0};
    const uint8_t file_header[32] = {0};                               // encode_file_header(0,0,0,0), src/main.rs:154
    if (!sink.write(file_header, 32)) { set_error("write failed"); return NTC_ERR_IO; }
    st[3] = 32;

    const uint64_t kBatchBlocks = batch_blocks();
    Channel<ReadBatch> parsed(2);
    Channel<BytesBatch> coded(2);
    double t_stage[4] = {0, 0, 0, 0};                                  // busy seconds: parse, GPU, block container, write
    bool fatal = false;
    // stage 1: parse
    std::thread t_read([&] { stage_guard(&fatal, [&] {
        bool eof = false;
        size_t reserve_bases = 0;
        while (!eof) {
            const double t0 = now_s();
            ReadBatch b;
            b.bases.reserve(reserve_bases);
            b.offsets.reserve(kBatchBlocks * kBlockReads + 1);
            b.offsets.push_back(0);
            while (b.offsets.size() - 1 < kBatchBlocks * kBlockReads && b.bases.size() < kBatchBases) {
                int r = reader->next(&b.bases);
                if (r == 0) { eof = true; break; }
                if (r < 0) { b.err = 1; eof = true; break; }
                b.offsets.push_back(b.bases.size());
            }
            // every batch but the last holds whole 65,536-read blocks: one that stopped on the base budget
            // is filled up to the next block boundary
            while (!eof && (b.offsets.size() - 1) % kBlockReads != 0) {
                int r = reader->next(&b.bases);
                if (r == 0) { eof = true; break; }
                if (r < 0) { b.err = 1; eof = true; break; }
                b.offsets.push_back(b.bases.size());
            }
            reserve_bases = b.bases.size() + b.bases.size() / 16;
            t_stage[0] += now_s() - t0;
            if (b.offsets.size() > 1 || b.err) { if (!parsed.push(std::move(b))) return; }
        }
        parsed.finish();
    }, [&] { parsed.finish(); }); });
    // stage 3: write in order
    int32_t w_rc = NTC_OK;
    std::string w_msg;
    std::thread t_write([&] { stage_guard(&fatal, [&] {
        BytesBatch bb;
        while (coded.pop(&bb)) {
            if (bb.rc != NTC_OK) { w_rc = bb.rc; w_msg = bb.msg; coded.close(); return; }
            const double t0 = now_s();
            for (auto& p : bb.parts) {
                if (!sink.write(p.data(), p.size())) { w_rc = NTC_ERR_IO; w_msg = "write failed"; coded.close(); return; }
                st[3] += p.size();
            }
            t_stage[3] += now_s() - t0;
        }
    }, [&] { coded.close(); }); });
    // stage 3: the block container per 65,536-read block on the pool (its own thread, so the GPU works on the
    // next batch meanwhile)
    struct RecordBatch { std::unique_ptr<uint64_t[]> records; std::vector<uint64_t> rec_off; uint64_t n_reads = 0; };
    Channel<RecordBatch> encoded(2);
    std::thread t_code([&] { stage_guard(&fatal, [&] {
        RecordBatch rb;
        while (encoded.pop(&rb)) {
            const uint64_t n = rb.n_reads, n_blocks = (n + kBlockReads - 1) / kBlockReads;
            BytesBatch out;
            out.parts.resize(n_blocks);
            std::vector<int32_t> brc(n_blocks, NTC_OK);
            const double t0 = now_s();
            parallel_for(n_blocks, n_threads, [&](uint64_t j) {
                const uint64_t lo = j * kBlockReads, hi = std::min(n, lo + kBlockReads);
                const uint64_t r0 = rb.rec_off[lo], r1 = rb.rec_off[hi];
                std::vector<uint8_t>& buf = out.parts[j];
                buf.resize(ntc_block_bound(r1 - r0));
                uint64_t nb = 0;
                // src/main.rs:170,176: num_records = the block size argument (65,536, or the tail count)
                brc[j] = ntc_block_write(rb.records.get() + r0, r1 - r0, hi - lo, buf.data(), buf.size(), &nb);
                buf.resize(brc[j] == NTC_OK ? nb : 0);
            });
            t_stage[2] += now_s() - t0;
            for (uint64_t j = 0; j < n_blocks; j++) {
                if (brc[j] == NTC_OK) st[4]++;
                else if (brc[j] == NTC_ERR_EMPTY_COLUMN) st[5]++;       // dropped, as the reference does (Q4)
                else { out.rc = brc[j]; out.msg = "block container failed"; }
            }
            if (!coded.push(std::move(out))) { encoded.close(); return; }
        }
        coded.finish();
    }, [&] { encoded.close(); coded.finish(); }); });
    // stage 2 (this thread): the hot path on the GPU
    int32_t rc = NTC_OK;
    std::string msg;
    ReadBatch b;
    while (parsed.pop(&b)) {
        if (b.err) { rc = NTC_ERR_IO; msg = "malformed fastX record near read " + std::to_string(st[0] + b.offsets.size()); break; }
        const uint64_t n = b.offsets.size() - 1;
        RecordBatch rb;
        rb.n_reads = n;
        rb.rec_off.resize(n + 1);
        // one record per base is the worst case (ntc_encode_bound); typical streams hold one per ~35 bases, so start
        // with a sixth of that and fall back to the bound -- the buffers are plain arrays, not zero-filled vectors
        uint64_t n_rec = 0;
        const double t0 = now_s();
        for (uint64_t cap = std::max<uint64_t>(b.bases.size() / 6, 1024);; cap = std::max<uint64_t>(b.bases.size(), 1)) {
            rb.records.reset(new uint64_t[cap]);
            rc = ntc_encode_batch(ctx, b.bases.data(), b.offsets.data(), n, rb.records.get(), cap, &n_rec, rb.rec_off.data());
            if (rc != NTC_ERR_CAPACITY || cap >= b.bases.size()) break;
        }
        t_stage[1] += now_s() - t0;
        if (rc != NTC_OK) { msg = ntc_last_error(); break; }
        st[0] += n; st[1] += b.bases.size(); st[2] += n_rec;
        if (!encoded.push(std::move(rb))) break;
    }
    parsed.close();
    encoded.finish();
    t_read.join();
    t_code.join();
    t_write.join();
    sink.close();
    if (getenv("NTC_TRACE"))
        fprintf(stderr, "[ntc trace] encode_file busy seconds: parse %.3f, gpu %.3f, block container %.3f, write %.3f\n", t_stage[0], t_stage[1],
                t_stage[2], t_stage[3]);
    if (stats) memcpy(stats, st, sizeof(st));
    if (rc != NTC_OK) { set_error(msg); return rc; }
    if (w_rc != NTC_OK) { set_error(w_msg); return w_rc; }
    if (fatal) { set_error("out of memory or internal error in a pipeline stage"); return NTC_ERR_IO; }
    return NTC_OK;
}


// `ntcomp decode` (src/main.rs:183-211): 32-byte file header, then blocks until the first header that
// cannot be read (src/lib.rs:327); FASTA records ">seq.N" numbered from 1 across blocks.
// stats (optional): [0] sequences, [1] bases + header bytes written, [2] records, [3] bytes read, [4] blocks.
int32_t ntc_decode_file(ntc_ctx* ctx, const char* in_path, const char* out_path, int32_t n_threads, uint64_t* stats) {
    if (!ctx || !in_path) { set_error("null argument"); return NTC_ERR_INVALID_ARG; }
    if (n_threads <= 0) n_threads = (int32_t)std::max(1u, std::thread::hardware_concurrency());
    FILE* in = strcmp(in_path, "-") == 0 ? stdin : fopen(in_path, "rb");
    if (!in) { set_error(std::string("cannot open ") + in_path); return NTC_ERR_IO; }
    Sink sink;
    if (!sink.open(out_path)) { if (in != stdin) fclose(in); set_error(std::string("cannot open ") + (out_path ? out_path : "-")); return NTC_ERR_IO; }
    uint64_t st[6] = {0, 0, 0, 0, 0, 0};
    uint8_t file_header[32];
    if (fread(file_header, 1, 32, in) != 32) { if (in != stdin) fclose(in); set_error("truncated file header"); return NTC_ERR_IO; }
    st[3] = 32;

    const uint64_t kBatchBlocks = batch_blocks();
    struct RecBatch { std::vector<uint64_t> records; uint64_t n_blocks = 0; int32_t rc = NTC_OK; std::string msg; };
    Channel<RecBatch> unpacked(2);
    Channel<BytesBatch> texts(2);
    bool fatal = false;
    // stage 1: frame the blocks of a group, then gunzip + decode + zip them on the pool
    std::thread t_read([&] { stage_guard(&fatal, [&] {
        bool eof = false;
        while (!eof) {
            std::vector<std::vector<uint8_t>> raw;
            uint64_t payload = 0;
            while (raw.size() < kBatchBlocks && payload < (256ull << 20)) {
                std::vector<uint8_t> blk;
                bool ok = true;
                for (int q = 0; q < 4 && ok; q++) {                     // four sub-blocks, src/lib.rs:326-355
                    uint8_t h[32];
                    size_t got = fread(h, 1, 32, in);
                    if (got != 32) { ok = false; break; }              // EOF (or a torn header) ends the stream, src/lib.rs:327
                    uint32_t block_size;
                    memcpy(&block_size, h, 4);
                    // block_size comes from the file: grow the buffer as bytes really arrive (a hostile header announcing
                    // 4 GiB must not become a 4 GiB allocation)
                    blk.insert(blk.end(), h, h + 32);
                    uint64_t left = block_size;
                    while (left && ok) {
                        const size_t step = (size_t)std::min<uint64_t>(left, 16u << 20), at = blk.size();
                        blk.resize(at + step);
                        if (fread(blk.data() + at, 1, step, in) != step) ok = false;
                        left -= step;
                    }
                    if (!ok) break;
                }
                if (!ok) { eof = true; break; }
                payload += blk.size();
                raw.push_back(std::move(blk));
            }
            if (raw.empty()) break;
            RecBatch rb;
            rb.n_blocks = raw.size();
            std::vector<std::vector<uint64_t>> recs(raw.size());
            std::vector<int32_t> brc(raw.size(), NTC_OK);
            std::vector<std::string> bmsg(raw.size());
            parallel_for(raw.size(), n_threads, [&](uint64_t j) {
                // a block holds at most one record per coded flag: num_u64 of the flags column (3rd sub-block)
                uint64_t pos = 0, cap = 0;
                for (int q = 0; q < 4; q++) {
                    uint32_t bs, nu;
                    memcpy(&bs, raw[j].data() + pos, 4);
                    memcpy(&nu, raw[j].data() + pos + 8, 4);
                    if (q == 2) cap = nu;
                    pos += 32 + bs;
                }
                recs[j].resize(std::max<uint64_t>(cap, 1));
                uint64_t n_rec = 0, used = 0;
                brc[j] = ntc_block_read_ex(raw[j].data(), raw[j].size(), recs[j].data(), recs[j].size(), &n_rec, &used,
                                           NTC_BLOCK_FULL_LAST_LITERAL_WORD);   // Q5: the reference panics on 1 block in 31
                if (brc[j] != NTC_OK) bmsg[j] = ntc_last_error();
                recs[j].resize(brc[j] == NTC_OK ? n_rec : 0);
            });
            for (size_t j = 0; j < raw.size(); j++) {
                if (brc[j] != NTC_OK && rb.rc == NTC_OK) { rb.rc = brc[j]; rb.msg = bmsg[j]; }
                // the reference decodes every block by itself (decode_block -> decode_sequence, src/lib.rs:363-365): records
                // in front of a block's first `first` flag belong to no sequence and are dropped (src/lib.rs:308-317) --
                // they must not be glued to the previous block's last sequence when blocks are concatenated here
                size_t skip = 0;
                while (skip < recs[j].size() && ((recs[j][skip] >> 56) & 1u) == 0) skip++;
                rb.records.insert(rb.records.end(), recs[j].begin() + skip, recs[j].end());
                st[3] += raw[j].size();
            }
            if (!unpacked.push(std::move(rb))) return;
        }
        unpacked.finish();
    }, [&] { unpacked.finish(); }); });
    int32_t w_rc = NTC_OK;
    std::string w_msg;
    std::thread t_write([&] { stage_guard(&fatal, [&] {
        BytesBatch bb;
        while (texts.pop(&bb))
            for (auto& p : bb.parts)
                if (!sink.write(p.data(), p.size())) { w_rc = NTC_ERR_IO; w_msg = "write failed"; texts.close(); return; }
    }, [&] { texts.close(); }); });
    // stage 2 (this thread): decode_sequence + FASTA text on the GPU
    int32_t rc = NTC_OK;
    std::string msg;
    RecBatch rb;
    while (unpacked.pop(&rb)) {
        if (rb.rc != NTC_OK) { rc = rb.rc; msg = rb.msg; break; }
        st[4] += rb.n_blocks;
        st[2] += rb.records.size();
        if (rb.records.empty()) continue;
        uint64_t n_bases = 0, n_seqs = 0, fasta_bytes = 0;
        ntc_decode_bound(rb.records.data(), rb.records.size(), st[0] + 1, &n_bases, &n_seqs, &fasta_bytes);
        BytesBatch out;
        out.parts.resize(1);
        out.parts[0].resize(fasta_bytes + 64);
        uint64_t nb = 0, ns = 0;
        rc = ntc_decode_fasta(ctx, rb.records.data(), rb.records.size(), st[0] + 1, out.parts[0].data(), fasta_bytes, &nb, &ns);
        if (rc != NTC_OK) { msg = ntc_last_error(); break; }
        out.parts[0].resize(nb);
        st[0] += ns;
        st[1] += nb;
        if (!texts.push(std::move(out))) break;
    }
    unpacked.close();
    texts.finish();
    t_read.join();
    t_write.join();
    sink.close();
    if (in != stdin) fclose(in);
    if (stats) memcpy(stats, st, sizeof(st));
    if (rc != NTC_OK) { set_error(msg); return rc; }
    if (w_rc != NTC_OK) { set_error(w_msg); return w_rc; }
    if (fatal) { set_error("out of memory or internal error in a pipeline stage"); return NTC_ERR_IO; }
    return NTC_OK;
}


// `ntcomp build` front end (src/main.rs:97-134): every record of every fastX file is one input sequence of
// kbo::build (add_revcomp = true there).  Construction itself is ntc_index_build (host, offline).
int32_t ntc_index_build_from_files(const char* const* paths, uint64_t n_paths, uint32_t k, int32_t add_revcomp, int32_t n_threads,
                                   ntc_index** out) {
    if (!out || (n_paths && !paths)) { set_error("null argument"); return NTC_ERR_INVALID_ARG; }
    std::vector<uint8_t> bases;
    std::vector<uint64_t> off(1, 0);
    for (uint64_t f = 0; f < n_paths; f++) {
        FastxReader rd;
        if (!rd.open(paths[f])) { set_error(std::string("cannot open ") + paths[f]); return NTC_ERR_IO; }
        for (;;) {
            int r = rd.next(&bases);
            if (r == 0) break;
            if (r < 0) { set_error(std::string("malformed fastX record in ") + paths[f]); return NTC_ERR_IO; }
            off.push_back(bases.size());
        }
    }
    const uint64_t n = off.size() - 1;
    std::vector<const uint8_t*> ptrs(n);
    std::vector<uint64_t> lens(n);
    for (uint64_t i = 0; i < n; i++) { ptrs[i] = bases.data() + off[i]; lens[i] = off[i + 1] - off[i]; }
    return ntc_index_build(ptrs.data(), lens.data(), n, k, add_revcomp, n_threads, out);
}

}  // extern "C"
