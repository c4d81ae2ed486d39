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
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <zlib.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <exception>
#include <functional>
#include <map>
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
constexpr uint64_t kBatchBlocksDefault = 2;        // blocks handed to the GPU per call (stage overlap granularity); with the gzip tasks of several
                                                   // batches in flight: 2 -> 0.22 s, 4 -> 0.25 s, 8 -> 0.34 s per 600 Mbases (profiles/r2/cli_c3_inflight_*.json)
constexpr uint64_t kBatchBases = 768ull << 20;     // ... or this many bases, whichever comes first

uint64_t batch_blocks() {                          // NTC_PIPELINE_BATCH_BLOCKS overrides (tests: several groups from small files)
    if (const char* env = getenv("NTC_PIPELINE_BATCH_BLOCKS")) {
        unsigned long long v = strtoull(env, nullptr, 10);
        if (v >= 1 && v <= 64) return v;
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
    bool try_pop(T* out) {                           // non-blocking
        std::lock_guard<std::mutex> lk(m_);
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

// ---- fastX parsing (plain or gzip; FASTA with wrapped lines, FASTQ with 4-line or wrapped records) ----
// needletail's parser (src/main.rs:158-163, normalize(true)) restated for the bytes that can reach the encoder: upper-case,
// U -> T, white space dropped; everything else passes through and is rejected by the encoder unless it is A, C, G or T.
// ---- in-memory parsing of a piece of fastX text that holds WHOLE records ------------------------------------
// (the reader thread cuts the input at record boundaries and the pieces are parsed on the worker pool)
struct ParsedPiece {
    std::vector<uint8_t> bases;      // normalised sequences, concatenated
    std::vector<uint32_t> lens;      // one per record
    int err = 0;                     // 1 malformed record, 2 a read of 2^32 bases or more
};

inline size_t normalise_into(const uint8_t* __restrict__ s, size_t n, std::vector<uint8_t>* out) {
    const size_t at = out->size();
    out->insert(out->end(), s, s + n);
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
    for (size_t j = 0; j < n; j++) {
        uint8_t ch = d[j];
        if (ch == ' ' || ch == '\t') continue;
        if (ch >= 'a' && ch <= 'z') ch = (uint8_t)(ch - 32);
        if (ch == 'U') ch = 'T';
        d[w++] = ch;
    }
    out->resize(at + w);
    return w;
}

struct MemLines {
    const uint8_t* p;
    const uint8_t* end;
    bool line(const uint8_t** s, size_t* n) {                          // next line without its "\n" / "\r\n"
        if (p >= end) return false;
        const uint8_t* nl = (const uint8_t*)memchr(p, '\n', (size_t)(end - p));
        *s = p;
        *n = nl ? (size_t)(nl - p) : (size_t)(end - p);
        p = nl ? nl + 1 : end;
        while (*n && (*s)[*n - 1] == '\r') (*n)--;
        return true;
    }
    int peek() const { return p < end ? *p : -1; }
};

void parse_piece(const uint8_t* text, size_t n, ParsedPiece* out) {
    MemLines ml{text, text + n};
    out->bases.reserve(n / 2 + 64);
    const uint8_t* ln;
    size_t len;
    while (ml.line(&ln, &len)) {
        if (len == 0) continue;                                        // blank lines between records
        size_t got_bases = 0;
        if (ln[0] == '>') {
            while (ml.peek() >= 0 && ml.peek() != '>') {
                if (!ml.line(&ln, &len)) break;
                got_bases += normalise_into(ln, len, &out->bases);
            }
        } else if (ln[0] == '@') {
            size_t want = 0, got = 0;                                  // quality has one byte per sequence byte
            for (;;) {
                if (!ml.line(&ln, &len)) { out->err = 1; return; }
                if (len && ln[0] == '+') break;
                size_t before = out->bases.size();
                normalise_into(ln, len, &out->bases);
                got_bases += out->bases.size() - before;
                want += out->bases.size() - before;
            }
            while (got < want) {
                if (!ml.line(&ln, &len)) { out->err = 1; return; }
                got += len;
            }
            if (got != want) { out->err = 1; return; }
        } else { out->err = 1; return; }
        if (got_bases >= (1ull << 32)) { out->err = 2; return; }
        out->lens.push_back((uint32_t)got_bases);
    }
}

// Is p (the first byte of a line) the start of a record?  FASTA: '>' at a line start always is.  FASTQ: '@' may also open
// a quality line, so the 4-line shape is checked: '@' line, sequence line, '+' line, quality line of the same length, and
// the line after it starts with '@' (or the text ends).  A quality line that begins with '@' fails the check -- the line
// two further on would be a sequence, which cannot start with '+'.  Multi-line FASTQ fails everywhere: the reader then
// keeps the text together and one worker parses it serially (correct, just not parallel).
bool record_starts_at(const uint8_t* p, const uint8_t* end, int fmt) {
    if (p >= end) return false;
    if (fmt == '>') return *p == '>';
    if (*p != '@') return false;
    MemLines ml{p, end};
    const uint8_t* s;
    size_t n1, n2, n3, n4;
    if (!ml.line(&s, &n1) || !ml.line(&s, &n2) || n2 == 0 || s[0] == '@' || s[0] == '+') return false;
    if (!ml.line(&s, &n3) || n3 == 0 || s[0] != '+') return false;
    if (!ml.line(&s, &n4) || n4 != n2) return false;
    return ml.peek() < 0 || ml.peek() == '@';
}

// last record start in text[lo, n) that is not at position 0; returns 0 when there is none.  Candidates are verified
// against the text up to n_total (>= n): a mapped file shows the bytes after the window, a stream does not (n_total == n)
size_t last_record_start(const uint8_t* text, size_t n, size_t lo, int fmt, size_t n_total) {
    size_t hi = n;
    while (hi > lo + 1) {
        const void* q = memrchr(text + lo, '\n', hi - lo - 1);         // a newline that is followed by at least one byte
        if (!q) return 0;
        const size_t at = (size_t)((const uint8_t*)q - text) + 1;
        if (text[at] == (uint8_t)fmt && record_starts_at(text + at, text + n_total, fmt)) return at;
        hi = at - 1 > lo ? at - 1 : lo;
        if (hi == lo) break;
    }
    return 0;
}

// free list of buffers that keep their capacity (fresh multi-megabyte vectors cost a page fault per 4 KB on first touch,
// and concurrent faults serialise in the kernel: measured here the in-order consumer spent 0.7 s per 300 MB in them)
template <typename T>
class Recycler {
public:
    T get() {
        std::lock_guard<std::mutex> lk(m_);
        if (free_.empty()) return T();
        T v = std::move(free_.back());
        free_.pop_back();
        return v;
    }
    void put(T&& v) {
        std::lock_guard<std::mutex> lk(m_);
        if (free_.size() < 64) free_.push_back(std::move(v));
    }
private:
    std::mutex m_;
    std::vector<T> free_;
};

// counting semaphore: bounds the pieces in flight between the reader and the batch assembly
class Tokens {
public:
    explicit Tokens(size_t n) : n_(n) {}
    bool acquire() { std::unique_lock<std::mutex> lk(m_); cv_.wait(lk, [&] { return n_ > 0 || closed_; }); if (closed_) return false; n_--; return true; }
    void release() { std::lock_guard<std::mutex> lk(m_); n_++; cv_.notify_one(); }
    void close() { std::lock_guard<std::mutex> lk(m_); closed_ = true; cv_.notify_all(); }
private:
    std::mutex m_;
    std::condition_variable cv_;
    size_t n_;
    bool closed_ = false;
};

// results of a worker pool, handed on in submission order
template <typename T>
class Ordered {
public:
    void put(uint64_t seq, T&& v) {
        std::lock_guard<std::mutex> lk(m_);
        ready_.emplace(seq, std::move(v));
        cv_.notify_all();
    }
    void finish(uint64_t total) { std::lock_guard<std::mutex> lk(m_); total_ = total; done_ = true; cv_.notify_all(); }
    void abort() { std::lock_guard<std::mutex> lk(m_); aborted_ = true; cv_.notify_all(); }
    bool next(T* out) {                                               // false at end of stream / abort
        std::unique_lock<std::mutex> lk(m_);
        cv_.wait(lk, [&] { return aborted_ || ready_.count(next_) || (done_ && next_ >= total_); });
        if (aborted_ || !ready_.count(next_)) return false;
        auto it = ready_.find(next_);
        *out = std::move(it->second);
        ready_.erase(it);
        next_++;
        return true;
    }
private:
    std::mutex m_;
    std::condition_variable cv_;
    std::map<uint64_t, T> ready_;
    uint64_t next_ = 0, total_ = 0;
    bool done_ = false, aborted_ = false;
};

// fixed pool of worker threads running queued tasks; a task's exception is kept and re-thrown by wait_idle() / the destructor's caller
class TaskPool {
public:
    explicit TaskPool(int n) {
        for (int i = 0; i < std::max(1, n); i++) th_.emplace_back([this] { run(); });
    }
    ~TaskPool() {
        { std::lock_guard<std::mutex> lk(m_); stop_ = true; cv_.notify_all(); }
        for (auto& t : th_) t.join();
    }
    void submit(std::function<void()> fn) {
        std::unique_lock<std::mutex> lk(m_);
        cv_room_.wait(lk, [&] { return q_.size() < 4 * th_.size() + 4; });   // bounded: the producer waits instead of piling up memory
        q_.push_back(std::move(fn));
        cv_.notify_one();
    }
    bool failed() { std::lock_guard<std::mutex> lk(m_); return (bool)failed_; }
private:
    void run() {
        for (;;) {
            std::function<void()> fn;
            {
                std::unique_lock<std::mutex> lk(m_);
                cv_.wait(lk, [&] { return stop_ || !q_.empty(); });
                if (q_.empty()) return;
                fn = std::move(q_.front());
                q_.pop_front();
                cv_room_.notify_one();
            }
            try { fn(); } catch (...) { std::lock_guard<std::mutex> lk(m_); if (!failed_) failed_ = std::current_exception(); }
        }
    }
    std::mutex m_;
    std::condition_variable cv_, cv_room_;
    std::deque<std::function<void()>> q_;
    std::vector<std::thread> th_;
    std::exception_ptr failed_;
    bool stop_ = false;
};

// fastX input: a gzip stream / pipe read through zlib, or -- a plain regular file -- a read-only mapping the parse workers
// read in place (no copy on the reader thread; the page faults spread over the workers)
struct FastxInput {
    gzFile gz = nullptr;
    const uint8_t* map = nullptr;
    size_t map_size = 0;
    bool open(const char* path) {
        if (strcmp(path, "-") != 0 && !getenv("NTC_NO_MMAP")) {
            int fd = ::open(path, O_RDONLY);
            if (fd < 0) return false;
            struct stat sb;
            uint8_t magic[2] = {0, 0};
            if (fstat(fd, &sb) == 0 && S_ISREG(sb.st_mode) && sb.st_size > 0 && pread(fd, magic, 2, 0) == 2 && !(magic[0] == 0x1f && magic[1] == 0x8b)) {
                void* m = mmap(nullptr, (size_t)sb.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
                if (m != MAP_FAILED) {
                    madvise(m, (size_t)sb.st_size, MADV_SEQUENTIAL);
                    map = (const uint8_t*)m;
                    map_size = (size_t)sb.st_size;
                    ::close(fd);
                    return true;
                }
            }
            ::close(fd);
        }
        gz = (strcmp(path, "-") == 0) ? gzdopen(0, "rb") : gzopen(path, "rb");
        if (gz) gzbuffer(gz, 1 << 20);
        return gz != nullptr;
    }
    void close() {
        if (gz) { gzclose(gz); gz = nullptr; }
        if (map) { munmap((void*)map, map_size); map = nullptr; }
    }
    ~FastxInput() { close(); }
};

// Cuts the fastX text of `in` into pieces of whole records, parses the pieces on `pool` and hands them to
// on_piece(std::shared_ptr<ParsedPiece>) IN FILE ORDER on the calling thread; on_piece returns false to stop early and may
// keep the pointer (the piece's buffers are recycled when the last reference goes).  At most `max_in_flight` pieces exist
// at a time.  *read_busy accumulates the reader thread's busy seconds.
template <typename OnPiece>
void parallel_fastx(FastxInput& in, TaskPool& pool, size_t max_in_flight, OnPiece&& on_piece, double* read_busy, bool* fatal) {
    const size_t kPiece = 4u << 20;                                    // text per parse task
    using PiecePtr = std::shared_ptr<ParsedPiece>;
    // declared before the channel: pieces still queued there are recycled by the destructors running in reverse order
    auto in_flight = std::make_shared<Tokens>(max_in_flight);
    auto texts = std::make_shared<Recycler<std::vector<uint8_t>>>();
    auto pieces = std::make_shared<Recycler<ParsedPiece>>();
    Ordered<PiecePtr> parsed;
    std::atomic<uint64_t> parse_busy_us{0};
    std::atomic<int64_t> tasks_out{0};                                 // parse tasks submitted and not finished (they refer to this frame)
    double consume_busy = 0, consume_wait = 0;
    uint64_t seq = 0;
    bool stop = false;
    // a piece of text holding whole records goes to the pool; own (may be null) keeps a copied text alive
    auto emit = [&](const uint8_t* p, size_t n, std::shared_ptr<std::vector<uint8_t>> own) {
        if (n == 0) return;
        if (!in_flight->acquire()) { stop = true; return; }
        const uint64_t my = seq++;
        tasks_out++;
        pool.submit([p, n, own, my, &parsed, &parse_busy_us, &tasks_out, texts, pieces, in_flight] {
            struct Done { std::atomic<int64_t>& c; ~Done() { c--; } } done{tasks_out};
            const double p0 = now_s();
            ParsedPiece* raw = new ParsedPiece(pieces->get());
            raw->bases.clear(); raw->lens.clear(); raw->err = 0;
            PiecePtr pp(raw, [pieces, in_flight](ParsedPiece* q) { pieces->put(std::move(*q)); delete q; in_flight->release(); });
            parse_piece(p, n, raw);
            if (own) texts->put(std::move(*own));
            parse_busy_us += (uint64_t)((now_s() - p0) * 1e6);
            parsed.put(my, std::move(pp));
        });
    };
    std::thread t_read([&] { stage_guard(fatal, [&] {
        int fmt = 0;
        if (in.map) {
            const uint8_t* base = in.map;
            const size_t size = in.map_size;
            for (size_t i = 0; i < size; i++) if (base[i] != '\n' && base[i] != '\r' && base[i] != ' ') { fmt = base[i]; break; }
            size_t pos = 0;
            if (fmt != '>' && fmt != '@') { emit(base, std::min<size_t>(size, 1u << 16), nullptr); pos = size; }   // not fastX: the parser reports it
            while (pos < size && !stop) {
                const double t0 = now_s();
                size_t cut = 0, win = kPiece, searched = 0;
                for (;;) {
                    if (pos + win >= size) { cut = size - pos; break; }
                    cut = last_record_start(base + pos, win, searched, fmt, size - pos);
                    if (cut) break;
                    searched = win - 1;
                    win += kPiece;
                }
                if (read_busy) *read_busy += now_s() - t0;
                emit(base + pos, cut, nullptr);
                pos += cut;
            }
        } else {
            std::vector<uint8_t> buf;
            bool eof = false;
            size_t searched = 0;                                       // buf[0, searched) holds no usable cut
            auto emit_front = [&](size_t n) {                          // buf[0, n) holds whole records
                if (n == 0) return;
                auto text = std::make_shared<std::vector<uint8_t>>(texts->get());
                text->assign(buf.begin(), buf.begin() + n);
                emit(text->data(), n, text);
                buf.erase(buf.begin(), buf.begin() + n);
            };
            while (!eof && !stop) {
                const double t0 = now_s();
                const size_t at = buf.size();
                buf.resize(at + kPiece);
                int got = gzread(in.gz, buf.data() + at, (unsigned)kPiece);
                if (got <= 0) { got = 0; eof = true; }
                buf.resize(at + (size_t)got);
                if (!fmt) for (uint8_t c : buf) if (c != '\n' && c != '\r' && c != ' ') { fmt = c; break; }
                if (fmt && fmt != '>' && fmt != '@') {                 // not fastX: let the parser report it
                    if (read_busy) *read_busy += now_s() - t0;
                    break;
                }
                if (!eof && fmt && buf.size() >= kPiece) {
                    const size_t cut = last_record_start(buf.data(), buf.size(), searched, fmt, buf.size());
                    if (cut) { emit_front(cut); searched = 0; } else searched = buf.size() > 1 ? buf.size() - 1 : 0;
                }
                if (read_busy) *read_busy += now_s() - t0;
            }
            if (!stop) emit_front(buf.size());
        }
        parsed.finish(seq);
    }, [&] { parsed.abort(); }); });
    for (;;) {
        PiecePtr pp;
        const double w0 = now_s();
        if (!parsed.next(&pp)) break;
        const double w1 = now_s();
        const bool go = on_piece(pp);
        pp.reset();
        consume_wait += w1 - w0;
        consume_busy += now_s() - w1;
        if (!go) break;
    }
    parsed.abort();
    in_flight->close();
    t_read.join();
    while (tasks_out.load() > 0) std::this_thread::sleep_for(std::chrono::microseconds(200));   // early stop only: a few queued tasks
    if (getenv("NTC_TRACE"))
        fprintf(stderr, "[ntc trace] fastx (%s): %llu pieces, parse tasks %.3f s (sum over workers), in-order consumer busy %.3f s / waiting %.3f s\n",
                in.map ? "mapped file" : "stream", (unsigned long long)seq, parse_busy_us.load() * 1e-6, consume_busy, consume_wait);
}

// one GPU call's worth of reads / one group of blocks travelling through the stages
struct ReadBatch {
    uint8_t* bases = nullptr;        // page-locked (ntc_pinned_alloc): the H2D copy of the GPU stage reads it in place
    uint64_t cap = 0;                // bytes allocated
    uint64_t first_cap = 0;          // size of the first allocation (a batch of short reads fits without growing)
    uint64_t n_bases = 0, n_reads = 0;
    std::vector<uint64_t> offsets;   // n_reads + 1 once the copy tasks are done
    int err = 0;                     // reader error (1 malformed input, 2 read too long)
    // copy tasks still writing into bases / offsets
    std::mutex m;
    std::condition_variable cv;
    uint64_t pending = 0;
    void task_begin() { std::lock_guard<std::mutex> lk(m); pending++; }
    void task_end() { std::lock_guard<std::mutex> lk(m); pending--; cv.notify_all(); }
    void wait_filled() { std::unique_lock<std::mutex> lk(m); cv.wait(lk, [&] { return pending == 0; }); }
    bool reserve(uint64_t want) {    // grows (long reads); the caller holds the only writer role and no task is pending
        if (want <= cap) return true;
        wait_filled();
        uint64_t ncap = std::max<uint64_t>(std::max<uint64_t>(want + want / 4, 2 * cap), first_cap);
        void* p = nullptr;
        if (ntc_pinned_alloc(ncap, &p) != NTC_OK) return false;
        if (n_bases) memcpy(p, bases, n_bases);
        if (bases) ntc_pinned_free(bases);
        bases = (uint8_t*)p;
        cap = ncap;
        return true;
    }
    ~ReadBatch() { wait_filled(); if (bases) ntc_pinned_free(bases); }   // copy tasks hold a plain pointer to this batch
};
using BatchPtr = std::unique_ptr<ReadBatch>;
struct BytesBatch {
    std::vector<std::vector<uint8_t>> parts;   // written in order
    int32_t rc = NTC_OK;
    std::string msg;
};
struct CodedBatch {                  // what ntc_encode_columns returned for one ReadBatch
    std::vector<ntc_coded_block> blocks;
    std::vector<uint64_t> words;
};

}  // namespace

extern "C" {

// `ntcomp encode` (src/main.rs:141-181): 32-byte zero file header, then one block per 65,536 reads.
// Stages: reader (inflate / read + cut at record boundaries) -> parse pool -> batch assembly -> the hot path + column codec
// on the GPU (this thread; records never leave HBM) -> gzip + headers per block on the pool -> in-order writer.
// stats (optional): [0] reads, [1] bases, [2] records, [3] bytes written, [4] blocks written,
// [5] blocks dropped because a column was empty (the reference drops them too: src/main.rs:170, Q4).
int32_t ntc_encode_file(ntc_ctx* ctx, const char* fastx_path, const char* out_path, int32_t n_threads, uint64_t* stats) {
    if (!ctx || !fastx_path) { set_error("null argument"); return NTC_ERR_INVALID_ARG; }
    if (n_threads <= 0) n_threads = (int32_t)std::max(1u, std::thread::hardware_concurrency());
    FastxInput in;
    if (!in.open(fastx_path)) { set_error(std::string("cannot open ") + fastx_path); return NTC_ERR_IO; }
    Sink sink;
    if (!sink.open(out_path)) { set_error(std::string("cannot open ") + (out_path ? out_path : "-")); return NTC_ERR_IO; }
    uint64_t st[6] = {0, 0, 0, 0, 0, 0};
    const uint8_t file_header[32] = {0};                               // encode_file_header(0,0,0,0), src/main.rs:154
    if (!sink.write(file_header, 32)) { set_error("write failed"); return NTC_ERR_IO; }
    st[3] = 32;

    const uint64_t kBatchBlocks = batch_blocks();
    const uint64_t kBatchReads = kBatchBlocks * kBlockReads;
    double t_stage[5] = {0, 0, 0, 0, 0};                               // busy seconds: read + cut, assemble, GPU, gzip, write
    std::atomic<uint64_t> copy_us{0}, gzip_us{0};                      // summed over the pool's workers
    bool fatal = false;
    int32_t rc = NTC_OK;
    std::string msg;
    {
    TaskPool pool(n_threads);
    Channel<BatchPtr> batches(2);
    Channel<BatchPtr> spare(8);                                        // used batch buffers on their way back to the assembler
    Channel<CodedBatch> coded(2);
    Channel<BytesBatch> bytes(2);

    // stage 2: parsed pieces, in order -> batches of whole 65,536-read blocks.  This thread only decides where every piece
    // goes; the bytes are moved by copy tasks on the pool, straight into the page-locked buffer the GPU stage uploads from.
    std::thread t_batch([&] { stage_guard(&fatal, [&] {
        // this thread allocates the page-locked batch buffers: without its own current device the first cudaHostAlloc would
        // bring up a primary context on GPU 0, whatever GPU ctx lives on
        ntc_ctx_bind_thread(ctx);
        uint64_t made = 0;
        auto fresh = [&]() -> BatchPtr {
            BatchPtr n;
            if (made >= 3) { if (!spare.pop(&n)) return nullptr; }     // three buffers circulate: filling, queued, on the GPU
            else if (!spare.try_pop(&n)) {
                n.reset(new ReadBatch());
                n->first_cap = std::min<uint64_t>(kBatchReads * 152 + (1u << 20), 256ull << 20);
                made++;
            }
            n->n_bases = n->n_reads = 0;
            n->err = 0;
            n->offsets.resize(kBatchReads + 1);
            n->offsets[0] = 0;
            return n;
        };
        BatchPtr b = fresh();
        bool stopped = false;
        auto hand_on = [&]() -> bool {
            BatchPtr out = std::move(b);
            if (!batches.push(std::move(out))) return false;
            b = fresh();
            return (bool)b;
        };
        // stage 1 lives inside parallel_fastx: cut at record boundaries (reader), parse on the pool
        parallel_fastx(in, pool, (size_t)n_threads * 3 + 8, [&](const std::shared_ptr<ParsedPiece>& pp) -> bool {
            const double t0 = now_s();
            struct Timer { double& acc; double t0; ~Timer() { acc += now_s() - t0; } } timer{t_stage[1], t0};
            if (pp->err && !b->err) b->err = pp->err;
            const uint64_t n = pp->lens.size();
            uint64_t i = 0, src = 0;
            while (i < n) {
                const uint64_t take = std::min<uint64_t>(kBatchReads - b->n_reads, n - i);
                uint64_t tb = 0;
                if (i == 0 && take == n) tb = pp->bases.size();
                else for (uint64_t j = i; j < i + take; j++) tb += pp->lens[j];
                if (!b->reserve(b->n_bases + tb + 64)) { b->err = 3; break; }
                ReadBatch* dst = b.get();
                const uint64_t r0 = dst->n_reads, b0 = dst->n_bases;
                dst->task_begin();
                pool.submit([pp, dst, i, take, src, tb, r0, b0, &copy_us] {
                    const double c0 = now_s();
                    if (tb) memcpy(dst->bases + b0, pp->bases.data() + src, tb);
                    uint64_t at = b0;
                    for (uint64_t j = 0; j < take; j++) { at += pp->lens[i + j]; dst->offsets[r0 + j + 1] = at; }
                    copy_us += (uint64_t)((now_s() - c0) * 1e6);
                    dst->task_end();
                });
                dst->n_reads += take;
                dst->n_bases += tb;
                i += take;
                src += tb;
                if (dst->n_reads == kBatchReads && !hand_on()) { stopped = true; return false; }
            }
            if (b->err) { hand_on(); stopped = true; return false; }
            return true;
        }, &t_stage[0], &fatal);
        if (!stopped && b && (b->n_reads || b->err)) batches.push(std::move(b));
        if (b) b->wait_filled();
        batches.finish();
    }, [&] { batches.finish(); }); });

    // stage 4: gzip + headers per SUB-block on the pool.  4a dispatches a batch's tasks and moves on -- the tasks of several batches are in flight
    // at once, so the one big task of a block (the colex column, ~0.8 MB of coded words = ~25 ms of DEFLATE at the reference's level 6) no longer
    // holds up the pool: the first version waited for each batch of 16 tasks before dispatching the next (0.38 s per 600 Mbases on 16 threads for
    // 2.2 CPU-seconds of work).  4b collects the batches in order.
    struct Latch { std::mutex m; std::condition_variable cv; size_t done = 0; };
    struct InFlight {
        std::shared_ptr<BytesBatch> out;
        std::shared_ptr<std::vector<int32_t>> brc;
        std::shared_ptr<Latch> latch;
        size_t n_blocks = 0;
    };
    Channel<InFlight> flying(4);
    std::thread t_code([&] { stage_guard(&fatal, [&] {
        CodedBatch cb;
        while (coded.pop(&cb)) {
            InFlight f;
            f.n_blocks = cb.blocks.size();
            f.out = std::make_shared<BytesBatch>();
            f.out->parts.resize(f.n_blocks * 4);                       // one gzip task per sub-block: 4 x finer than per block
            f.brc = std::make_shared<std::vector<int32_t>>(f.n_blocks * 4, NTC_OK);
            f.latch = std::make_shared<Latch>();
            auto shared = std::make_shared<CodedBatch>(std::move(cb));
            auto out = f.out;
            auto brc = f.brc;
            auto latch = f.latch;
            for (size_t j = 0; j < f.n_blocks * 4; j++)
                pool.submit([j, out, brc, shared, latch, &gzip_us] {
                    const double g0 = now_s();
                    const ntc_coded_block& blk = shared->blocks[j / 4];
                    int32_t r = blk.status;
                    if (r == NTC_OK) {
                        const uint64_t n_w = blk.col[j % 4].n_words;
                        std::vector<uint8_t>& buf = out->parts[j];
                        buf.resize(32 + 64 + n_w * 8 + n_w / 100 + 1024);
                        uint64_t nb = 0;
                        r = ntc_sub_block_assemble(&blk, (int32_t)(j % 4), shared->words.data(), buf.data(), buf.size(), &nb);
                        buf.resize(r == NTC_OK ? nb : 0);
                    }
                    (*brc)[j] = r;
                    gzip_us += (uint64_t)((now_s() - g0) * 1e6);
                    std::lock_guard<std::mutex> lk(latch->m);
                    latch->done++;
                    latch->cv.notify_all();
                });
            if (!flying.push(std::move(f))) { coded.close(); return; }
        }
        flying.finish();
    }, [&] { coded.close(); flying.finish(); }); });
    std::thread t_collect([&] { stage_guard(&fatal, [&] {
        InFlight f;
        while (flying.pop(&f)) {
            const double t0 = now_s();
            { std::unique_lock<std::mutex> lk(f.latch->m); f.latch->cv.wait(lk, [&] { return f.latch->done == f.n_blocks * 4; }); }
            for (size_t j = 0; j < f.n_blocks; j++) {
                int32_t r = NTC_OK;
                for (int q = 0; q < 4; q++) if ((*f.brc)[j * 4 + q] != NTC_OK) r = (*f.brc)[j * 4 + q];
                if (r == NTC_OK) st[4]++;
                else {
                    for (int q = 0; q < 4; q++) f.out->parts[j * 4 + q].clear();
                    if (r == NTC_ERR_EMPTY_COLUMN) st[5]++;            // dropped, as the reference does (Q4)
                    else { f.out->rc = r; f.out->msg = "block container failed"; }
                }
            }
            t_stage[3] += now_s() - t0;
            if (!bytes.push(std::move(*f.out))) { flying.close(); return; }
        }
        bytes.finish();
    }, [&] { flying.close(); bytes.finish(); }); });

    // stage 5: write in order
    int32_t w_rc = NTC_OK;
    std::string w_msg;
    std::thread t_write([&] { stage_guard(&fatal, [&] {
        BytesBatch bb;
        while (bytes.pop(&bb)) {
            if (bb.rc != NTC_OK) { w_rc = bb.rc; w_msg = bb.msg; bytes.close(); return; }
            const double t0 = now_s();
            for (auto& p : bb.parts) {
                if (!sink.write(p.data(), p.size())) { w_rc = NTC_ERR_IO; w_msg = "write failed"; bytes.close(); return; }
                st[3] += p.size();
            }
            t_stage[4] += now_s() - t0;
        }
    }, [&] { bytes.close(); }); });

    // stage 3 (this thread): the hot path + the column codec on the GPU
    BatchPtr b;
    while (batches.pop(&b)) {
        b->wait_filled();
        const uint64_t n = b->n_reads;
        if (n) {
            const double t0 = now_s();
            CodedBatch cb;
            cb.blocks.resize((n + kBlockReads - 1) / kBlockReads);
            uint64_t n_blocks = 0, n_words = 0, n_rec = 0;
            const uint64_t* words = nullptr;
            rc = ntc_encode_columns(ctx, b->bases, b->offsets.data(), n, kBlockReads, cb.blocks.data(), cb.blocks.size(), &n_blocks, &words,
                                    &n_words, &n_rec);
            if (rc == NTC_OK) cb.words.assign(words, words + n_words);   // the pinned buffer is reused by the call after next
            t_stage[2] += now_s() - t0;
            if (rc != NTC_OK) { msg = ntc_last_error(); break; }
            st[0] += n; st[1] += b->n_bases; st[2] += n_rec;
            if (!coded.push(std::move(cb))) break;
        }
        const int berr = b->err;
        spare.push(std::move(b));                                      // the buffer goes back to the assembler
        if (berr) {
            rc = berr == 2 ? NTC_ERR_UNSUPPORTED : NTC_ERR_IO;
            msg = berr == 2 ? "a read of 2^32 bases or more is not supported"
                : berr == 3 ? "cannot page-lock a batch buffer" : "malformed fastX record near read " + std::to_string(st[0] + 1);
            break;
        }
    }
    batches.close();
    spare.close();
    coded.finish();
    t_batch.join();
    t_code.join();
    t_collect.join();
    t_write.join();
    if (pool.failed()) fatal = true;
    if (rc == NTC_OK && w_rc != NTC_OK) { rc = w_rc; msg = w_msg; }
    }
    in.close();
    sink.close();
    if (getenv("NTC_TRACE"))
        fprintf(stderr, "[ntc trace] encode_file busy seconds: cut %.3f, place %.3f, gpu %.3f, gzip (collector waiting, in order) %.3f, write %.3f | pool work: "
                "copy into batch %.3f, headers + gzip %.3f\n", t_stage[0], t_stage[1], t_stage[2], t_stage[3], t_stage[4], copy_us.load() * 1e-6,
                gzip_us.load() * 1e-6);
    if (stats) memcpy(stats, st, sizeof(st));
    if (rc != NTC_OK) { set_error(msg); return rc; }
    if (fatal) { set_error("out of memory or internal error in a pipeline stage"); return NTC_ERR_IO; }
    return NTC_OK;
}


// `ntcomp decode` (src/main.rs:183-211): 32-byte file header, then blocks until the first header that
// cannot be read (src/lib.rs:327); FASTA records ">seq.N" numbered from 1 across blocks.
// Stages: reader (frames the four sub-blocks of every block) -> gunzip per block on the pool -> the column codec + decode on
// the GPU (this thread; records never leave HBM) -> in-order writer straight from the library's pinned text buffers.
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
    struct ColBatch { std::vector<ntc_coded_block> blocks; std::vector<uint64_t> words; int32_t rc = NTC_OK; std::string msg; };
    struct TextRef { const uint8_t* p; uint64_t n; };
    bool fatal = false;
    int32_t rc = NTC_OK;
    std::string msg;
    double t_stage[3] = {0, 0, 0};                                     // busy seconds: read + gunzip, GPU, write
    {
    Channel<ColBatch> unpacked(2);
    Channel<TextRef> texts(1);
    std::mutex wm;
    std::condition_variable wcv;
    uint64_t written = 0;                                              // batches the writer is done with
    bool writer_gone = false;
    // stage 1: frame the blocks of a group, then headers + gunzip on the pool
    std::thread t_read([&] { stage_guard(&fatal, [&] {
        bool eof = false;
        while (!eof) {
            const double t0 = now_s();
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
            ColBatch cb;
            const size_t nb = raw.size();
            cb.blocks.resize(nb);
            std::vector<std::vector<uint64_t>> words(nb);
            std::vector<int32_t> brc(nb, NTC_OK);
            std::vector<std::string> bmsg(nb);
            parallel_for(nb, n_threads, [&](uint64_t j) {
                uint64_t cap = 0, pos = 0;                              // encoded_size of the four headers
                for (int q = 0; q < 4; q++) {
                    uint32_t bs, es;
                    memcpy(&bs, raw[j].data() + pos, 4);
                    memcpy(&es, raw[j].data() + pos + 12, 4);
                    cap += std::min<uint64_t>(es, (uint64_t)bs * 1100 + 64);   // a gzip member cannot expand more than ~1032 : 1
                    pos += 32 + bs;
                }
                words[j].resize(cap + 8);
                uint64_t nw = 0, used = 0;
                brc[j] = ntc_block_disassemble(raw[j].data(), raw[j].size(), &cb.blocks[j], words[j].data(), words[j].size(), &nw, &used);
                if (brc[j] != NTC_OK) bmsg[j] = ntc_last_error();
                words[j].resize(brc[j] == NTC_OK ? nw : 0);
            });
            // A block that cannot be unpacked ends the stream: the group keeps the blocks before it, which are still decoded
            // and written (the reference's `while let Ok(..)` loop writes every block before the first bad one, src/main.rs:203).
            size_t n_good = nb;
            for (size_t j = 0; j < nb; j++) {
                if (brc[j] == NTC_OK) continue;
                cb.rc = brc[j] == NTC_ERR_CAPACITY ? NTC_ERR_IO : brc[j];
                cb.msg = bmsg[j];
                n_good = j;
                eof = true;
                break;
            }
            uint64_t at = 0;
            for (size_t j = 0; j < n_good; j++) {
                for (int q = 0; q < 4; q++) cb.blocks[j].col[q].word_offset += at;
                at += words[j].size();
                st[3] += raw[j].size();
            }
            cb.blocks.resize(n_good);
            cb.words.reserve(at);
            for (size_t j = 0; j < n_good; j++) cb.words.insert(cb.words.end(), words[j].begin(), words[j].end());
            t_stage[0] += now_s() - t0;
            if (!unpacked.push(std::move(cb))) return;
        }
        unpacked.finish();
    }, [&] { unpacked.finish(); }); });
    int32_t w_rc = NTC_OK;
    std::string w_msg;
    std::thread t_write([&] { stage_guard(&fatal, [&] {
        TextRef tr;
        while (texts.pop(&tr)) {
            const double t0 = now_s();
            const bool ok = sink.write(tr.p, tr.n);   // one buffered write per batch: a single file takes ~3.5 GB/s (page-cache / tmpfs writes serialise on the inode; parallel pwrite measured no faster, r2)
            t_stage[2] += now_s() - t0;
            { std::lock_guard<std::mutex> lk(wm); written++; wcv.notify_all(); }
            if (!ok) { w_rc = NTC_ERR_IO; w_msg = "write failed"; texts.close(); { std::lock_guard<std::mutex> lk(wm); writer_gone = true; wcv.notify_all(); } return; }
        }
        { std::lock_guard<std::mutex> lk(wm); writer_gone = true; wcv.notify_all(); }
    }, [&] { { std::lock_guard<std::mutex> lk(wm); writer_gone = true; wcv.notify_all(); } texts.close(); }); });
    // stage 2 (this thread): integer decoders + zip + decode_sequence + FASTA text on the GPU
    ColBatch cb;
    uint64_t pushed = 0;
    // one call of the fused decoder over blocks [b, b + n_b) of the current group; false = stop (rc / msg say why, or the writer is gone)
    auto decode_group = [&](const ntc_coded_block* b, uint64_t n_b) -> bool {
        const double t0 = now_s();
        const uint8_t* text = nullptr;
        uint64_t nb = 0, ns = 0, nr = 0;
        rc = ntc_decode_columns_fasta(ctx, b, n_b, cb.words.data(), cb.words.size(), st[0] + 1,
                                      NTC_BLOCK_FULL_LAST_LITERAL_WORD, &text, &nb, &ns, &nr);   // Q5: the reference panics on 1 block in 31
        t_stage[1] += now_s() - t0;
        if (rc != NTC_OK) { msg = ntc_last_error(); return false; }
        st[0] += ns;
        st[1] += nb;
        st[2] += nr;
        st[4] += n_b;
        if (nb) {
            if (!texts.push(TextRef{text, nb})) return false;
            pushed++;
        }
        return true;
    };
    while (unpacked.pop(&cb)) {
        // the text of call i lives in a pinned buffer that call i + 2 reuses: wait until the writer is done with it
        if (pushed >= 2) { std::unique_lock<std::mutex> lk(wm); wcv.wait(lk, [&] { return writer_gone || written + 2 > pushed; }); }
        if (!cb.blocks.empty() && !decode_group(cb.blocks.data(), cb.blocks.size())) {
            if (rc == NTC_OK || cb.blocks.size() == 1) break;
            // A damaged block somewhere in the group.  Like the reference (src/main.rs:203), everything before the first bad
            // block is still written: go through the group block by block.  (A deliberate difference remains: the reference
            // then stops silently with exit status 0, this call returns the error.)  The failed call consumed one of the two
            // text buffers out of turn, so each call here waits for the writer to be done with everything pushed so far.
            const int32_t group_rc = rc;
            const std::string group_msg = msg;
            for (size_t j = 0; j < cb.blocks.size(); j++) {
                { std::unique_lock<std::mutex> lk(wm); wcv.wait(lk, [&] { return writer_gone || written >= pushed; }); }
                if (!decode_group(&cb.blocks[j], 1)) break;
            }
            if (rc == NTC_OK) { rc = group_rc; msg = group_msg; }
            break;
        }
        if (cb.rc != NTC_OK) { rc = cb.rc; msg = cb.msg; break; }
    }
    unpacked.close();
    texts.finish();
    t_read.join();
    t_write.join();
    if (rc == NTC_OK && w_rc != NTC_OK) { rc = w_rc; msg = w_msg; }
    }
    sink.close();
    if (in != stdin) fclose(in);
    if (getenv("NTC_TRACE"))
        fprintf(stderr, "[ntc trace] decode_file busy seconds: read+gunzip %.3f, gpu %.3f, write %.3f\n", t_stage[0], t_stage[1], t_stage[2]);
    if (stats) memcpy(stats, st, sizeof(st));
    if (rc != NTC_OK) { set_error(msg); return rc; }
    if (fatal) { set_error("out of memory or internal error in a pipeline stage"); return NTC_ERR_IO; }
    return NTC_OK;
}


// fastX file (plain or gzip; FASTA or FASTQ) -> normalised sequences, parsed the way the encode pipeline does it (reader thread
// cutting at record boundaries + parse pool).  *bases / *offsets are malloc'ed: release them with ntc_free.
int32_t ntc_fastx_load(const char* path, int32_t n_threads, uint8_t** bases_out, uint64_t** offsets_out, uint64_t* n_reads, uint64_t* n_bases) {
    if (!path || !bases_out || !offsets_out || !n_reads) { set_error("null argument"); return NTC_ERR_INVALID_ARG; }
    *bases_out = nullptr; *offsets_out = nullptr; *n_reads = 0;
    if (n_bases) *n_bases = 0;
    if (n_threads <= 0) n_threads = (int32_t)std::max(1u, std::thread::hardware_concurrency());
    FastxInput in;
    if (!in.open(path)) { set_error(std::string("cannot open ") + path); return NTC_ERR_IO; }
    std::vector<uint8_t> bases;
    std::vector<uint64_t> off(1, 0);
    int err = 0;
    bool fatal = false;
    try {
        TaskPool pool(n_threads);
        parallel_fastx(in, pool, (size_t)n_threads * 3 + 8, [&](const std::shared_ptr<ParsedPiece>& pp) -> bool {
            bases.insert(bases.end(), pp->bases.begin(), pp->bases.end());
            uint64_t at = off.back();
            for (uint32_t l : pp->lens) { at += l; off.push_back(at); }
            if (pp->err) { err = pp->err; return false; }
            return true;
        }, nullptr, &fatal);
        if (pool.failed()) fatal = true;
    } catch (...) { fatal = true; }
    in.close();
    if (fatal) { set_error("out of memory or internal error while parsing"); return NTC_ERR_IO; }
    if (err) { set_error(err == 2 ? std::string("a read of 2^32 bases or more in ") + path : std::string("malformed fastX record in ") + path + " near record " + std::to_string(off.size())); return err == 2 ? NTC_ERR_UNSUPPORTED : NTC_ERR_IO; }
    uint8_t* b = (uint8_t*)malloc(bases.size() + 1);
    uint64_t* o = (uint64_t*)malloc(off.size() * 8);
    if (!b || !o) { free(b); free(o); set_error("out of memory"); return NTC_ERR_IO; }
    if (!bases.empty()) memcpy(b, bases.data(), bases.size());
    memcpy(o, off.data(), off.size() * 8);
    *bases_out = b; *offsets_out = o; *n_reads = off.size() - 1;
    if (n_bases) *n_bases = bases.size();
    return NTC_OK;
}
void ntc_free(void* p) { free(p); }

// `ntcomp build` front end (src/main.rs:97-134): every record of every fastX file is one input sequence of
// kbo::build (add_revcomp = true there).  Construction itself is ntc_index_build (host, offline).
static int32_t build_from_files(const char* const* paths, uint64_t n_paths, uint32_t k, int32_t add_revcomp, int32_t n_threads, int32_t device,
                                ntc_index** out);
int32_t ntc_index_build_from_files(const char* const* paths, uint64_t n_paths, uint32_t k, int32_t add_revcomp, int32_t n_threads,
                                   ntc_index** out) {
    return build_from_files(paths, n_paths, k, add_revcomp, n_threads, -1, out);
}
// the same with the construction on GPU `device` (ntc_index_build_gpu): what `ntcomp-b200 build --gpu` runs
int32_t ntc_index_build_from_files_gpu(int32_t device, const char* const* paths, uint64_t n_paths, uint32_t k, int32_t add_revcomp,
                                       int32_t n_threads, ntc_index** out) {
    if (device < 0) { set_error("device ordinal out of range"); return NTC_ERR_INVALID_ARG; }
    return build_from_files(paths, n_paths, k, add_revcomp, n_threads, device, out);
}
static int32_t build_from_files(const char* const* paths, uint64_t n_paths, uint32_t k, int32_t add_revcomp, int32_t n_threads, int32_t device,
                                ntc_index** out) {
    if (!out || (n_paths && !paths)) { set_error("null argument"); return NTC_ERR_INVALID_ARG; }
    std::vector<uint8_t*> fb(n_paths, nullptr);
    std::vector<uint64_t*> fo(n_paths, nullptr);
    std::vector<uint64_t> fn(n_paths, 0);
    int32_t rc = NTC_OK;
    for (uint64_t f = 0; f < n_paths && rc == NTC_OK; f++) rc = ntc_fastx_load(paths[f], n_threads, &fb[f], &fo[f], &fn[f], nullptr);
    if (rc == NTC_OK) {
        std::vector<const uint8_t*> ptrs;
        std::vector<uint64_t> lens;
        for (uint64_t f = 0; f < n_paths; f++)
            for (uint64_t i = 0; i < fn[f]; i++) { ptrs.push_back(fb[f] + fo[f][i]); lens.push_back(fo[f][i + 1] - fo[f][i]); }
        rc = device >= 0 ? ntc_index_build_gpu(device, ptrs.data(), lens.data(), ptrs.size(), k, add_revcomp, out, nullptr)
                         : ntc_index_build(ptrs.data(), lens.data(), ptrs.size(), k, add_revcomp, n_threads, out);
    }
    for (uint64_t f = 0; f < n_paths; f++) { free(fb[f]); free(fo[f]); }
    return rc;
}

}  // extern "C"
