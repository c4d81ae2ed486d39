// api.cu -- C ABI of the hot path: context creation (index re-layout + upload), the stream-ordered
// device pipelines, and the host-buffer entry points a Rust/cgo/ctypes caller binds.
// There is no CPU fallback here: without a GPU every compute entry point returns NTC_ERR_CUDA.
#include <sched.h>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <exception>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/ntcomp_b200.h"
#include "column_decode_kernels.cuh"
#include "column_kernels.cuh"
#include "column_plan.h"
#include "container.h"
#include "ctx.h"
#include "decode_kernels.cuh"
#include "encode_kernels.cuh"
#include "host_index.h"
#include "layout_kernels.cuh"
#include "synth_common.h"

using namespace ntc;

namespace {

struct CountLoad {
    const uint32_t* c;
    NTC_HD Pair64 operator()(uint64_t i) const { return Pair64{c[i], 0}; }
};

inline unsigned blocks_for(uint64_t n, unsigned per) { return (unsigned)((n + per - 1) / per); }

// Brackets the launches of one kernel class with CUDA events on the launching stream when
// profiling is on; always counts launches.
struct Timed {
    ntc_ctx* ctx; int cls; cudaStream_t st; cudaEvent_t e0 = nullptr, e1 = nullptr;
    Timed(ntc_ctx* c, int k, cudaStream_t s, int n_launches = 1) : ctx(c), cls(k), st(s) {
        ctx->launches += n_launches;
        ctx->class_launches[cls] += n_launches;
        if (!ctx->profiling) return;
        auto get = [&]() { cudaEvent_t e; if (ctx->event_pool.empty()) cudaEventCreate(&e); else { e = ctx->event_pool.back(); ctx->event_pool.pop_back(); } return e; };
        e0 = get(); e1 = get();
        cudaEventRecord(e0, st);
    }
    ~Timed() {
        if (!e0) return;
        cudaEventRecord(e1, st);
        ctx->spans.push_back(ntc_ctx::Span{cls, e0, e1});
    }
};

// Results the host waits for (scan totals, status word) are written straight into the lane's pinned landing
// zone (UVA: cudaHostAlloc memory is device-accessible) instead of going through a cudaMemcpyAsync: a small
// D2H copy queues behind the other lane's bulk D2H on the copy engine and would serialise the two lanes.
__global__ void publish_kernel(const uint64_t* __restrict__ pair, const uint32_t* __restrict__ status, volatile uint64_t* land) {
    land[0] = pair[0];
    land[1] = pair[1];
    land[2] = *status;
    __threadfence_system();
}

__global__ void policy_kernel(uint64_t* out) {
    const CachePolicy p = make_policies();
    out[0] = p.keep;
    out[1] = p.stream;
}
__global__ void copy_totals_kernel(const uint64_t* totals, uint64_t* counts) {
    counts[0] = totals[0];
    counts[1] = totals[0] * 8;
}

__global__ void __launch_bounds__(256) synth_reads_kernel(const uint8_t* __restrict__ genome, uint64_t genome_len, uint64_t seed,
                                                          uint64_t r0, uint64_t n_reads, uint32_t read_len, uint32_t err_ppm,
                                                          uint8_t* __restrict__ out) {
    uint64_t i = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    uint64_t total = n_reads * read_len;
    if (i >= total) return;
    uint64_t r = i / read_len;
    uint32_t t = (uint32_t)(i - r * read_len);
    out[i] = read_base(genome, genome_len, seed, r0 + r, read_len, err_ppm, t);
}

int32_t status_to_code(uint32_t st) {
    if (st & ST_READ_TOO_LONG) { set_error("a read of 2^32 bases or more is not supported"); return NTC_ERR_UNSUPPORTED; }
    if (st & ST_BAD_BASE) { set_error("read holds a non-ACGT byte or a base absent from the index (reference underflows at src/lib.rs:207)"); return NTC_ERR_BAD_BASE; }
    if (st & ST_EMPTY_READ) { set_error("empty read (EncodeError, src/lib.rs:224)"); return NTC_ERR_EMPTY_READ; }
    if (st & ST_LEN24) { set_error("record length >= 2^24 (src/lib.rs:226)"); return NTC_ERR_LEN24; }
    if (st & ST_BAD_RECORD) { set_error("record holds a colex rank >= n_sets"); return NTC_ERR_BAD_RECORD; }
    if (st & ST_CAPACITY) { set_error("output buffer too small"); return NTC_ERR_CAPACITY; }
    return NTC_OK;
}

// Tile length for a batch: long reads are cut into tiles (encode_kernels.cuh) so that a batch holds
// enough independent chains to fill the GPU and balance its tail (>= ~1.5 M work items when the batch
// allows); a tile costs k-1 warm-up bases, so tiles are never shorter than 512 positions by default.
// Reads no longer than the tile are one item each -- short-read batches are unaffected.
uint32_t tile_len_for(const ntc_ctx* ctx, uint64_t n_bases) {
    if (ctx->tile_len) return ctx->tile_len;
    uint64_t t = std::max<uint64_t>(512, n_bases / 1500000);
    t = std::min<uint64_t>((t + 63) & ~63ull, 1u << 22);
    return (uint32_t)t;
}

// matching statistics of a packed batch: tiles per read -> scan -> item descriptors -> ms_kernel over
// the items -> record counts of multi-tile reads.  counts may be null (diagnostic entry point).
int32_t ms_pipeline(ntc_ctx* ctx, Lane& lane, const uint64_t* packed, const uint64_t* d_offsets, uint64_t n_reads, uint64_t n_bases,
                    uint32_t T, uint64_t* scratch, uint32_t* counts, uint32_t* status, cudaStream_t st, bool keep_records = false) {
    if (!(ctx->uses & NTC_USE_ENCODE)) { set_error("this context was created without NTC_USE_ENCODE (no rank blocks / LCS words in HBM)"); return NTC_ERR_UNSUPPORTED; }
    const uint64_t max_items = n_reads + n_bases / T;                 // sum over reads of max(1, ceil(len / T)) is at most this
    NTC_CUDA_TRY(ensure(&lane.tile_off, (n_reads + 2) * 8 + 16));
    NTC_CUDA_TRY(ensure(&lane.desc, (max_items + 1) * 8));
    NTC_CUDA_TRY(ensure(&lane.tiles, scan_num_tiles(n_reads) * sizeof(Pair64)));
    uint64_t* tile_off = (uint64_t*)lane.tile_off.p;
    uint64_t* tile_tot = tile_off + n_reads + 2;                      // 2 x u64 scan totals
    {
        Timed t(ctx, KC_SCAN, st, 3);
        exclusive_scan_pairs(TileLoad{d_offsets, T}, n_reads, (Pair64*)lane.tiles.p, nullptr, tile_tot, tile_off, nullptr, st,
                             TileDescPost{(uint64_t*)lane.desc.p});
    }
    {
        Timed t(ctx, KC_MS, st);
        // stage buffer for the block's packed reads: 128 reads of the batch's mean length (+ alignment slack), at most 8 KB
        const uint64_t mean_len = n_reads ? (n_bases + n_reads - 1) / n_reads : 0;
        uint32_t stage_words = (uint32_t)std::min<uint64_t>(kMsStageWordsMax, ((kMsThreads * mean_len + 31) / 32 + 8 + 15) & ~15ull);
        static const long stage_env = [] { const char* e = getenv("NTC_MS_STAGE_WORDS"); return e ? atol(e) : -1L; }();   // tests/test_stress_determinism.py
        if (stage_env >= 0) stage_words = (uint32_t)std::min<long>(stage_env, (long)kMsStageWordsMax) & ~1u;
        static const int force_blocks = [] { const char* e = getenv("NTC_MS_BLOCKS"); return e ? atoi(e) : 0; }();
        const size_t smem = (size_t)stage_words * 8;
        const bool twelve = force_blocks ? force_blocks == 12 : stage_words <= 640;
        auto launch = [&](auto kernel) {
            kernel<<<blocks_for(max_items, kMsThreads), kMsThreads, smem, st>>>(ctx->dix, packed, d_offsets, (const uint64_t*)lane.desc.p,
                                                                                tile_off + n_reads, T, scratch, counts, status, keep_records ? 1 : 0,
                                                                                stage_words);
        };
        if (ctx->dix.rank_nodes == 64) { if (twelve) launch(ms_kernel<12, 64>); else launch(ms_kernel<10, 64>); }
        else { if (twelve) launch(ms_kernel<12, 32>); else launch(ms_kernel<10, 32>); }
    }
    lane.seg_len = 0;
    if (counts && n_bases > (uint64_t)T) {                            // no read can exceed T otherwise
        // One thread per read (count_kernel) walks a read's cover as one chain of dependent loads; the segment path cuts it into
        // 512-position pieces.  Measured on B200 (profiles/r2/ab_seg_parallel.txt): with enough reads to fill the GPU the two cost
        // the same (C5, 250 k x 10 kbp: 7.09 vs 7.43 ms for count + emit; both walk every record's entry once or twice out of DRAM),
        // with few long sequences -- contigs handed in as reads, as the reference's own test does -- the chain's latency is all
        // there is and the segments win by the number of pieces.  So: segments when the reads cannot fill the machine
        // (fewer than 148 SMs x 1024 threads).  NTC_SEG_SERIAL=1 / 0 forces either.
        static const int force_serial = [] { const char* e = getenv("NTC_SEG_SERIAL"); return e ? (atoi(e) != 0 ? 1 : 0) : -1; }();
        const bool serial = force_serial >= 0 ? force_serial == 1 : n_reads >= 148u * 1024u;
        if (keep_records && !serial) {
            // records of reads longer than the tile: speculative walk per 512-position segment, joined per read; seg_emit_kernel
            // (encode_pipeline) writes the words once the offsets are known
            static const uint32_t force_S = [] { const char* e = getenv("NTC_SEG_LEN"); return e ? (uint32_t)atoi(e) & ~63u : 0u; }();
            const uint32_t S = force_S >= 64 ? force_S : seg_len_for(n_bases, n_reads);
            const uint64_t max_segs = n_bases / S + n_reads;
            NTC_CUDA_TRY(ensure(&lane.seg_off, (n_reads + 2) * 8 + 16));
            NTC_CUDA_TRY(ensure(&lane.seg_desc, (max_segs + 1) * 8));
            NTC_CUDA_TRY(ensure(&lane.seg_bits, (max_segs + 1) * (S / 64) * 8));
            NTC_CUDA_TRY(ensure(&lane.seg_info, (max_segs + 1) * 16));
            uint64_t* seg_off = (uint64_t*)lane.seg_off.p;
            Timed t(ctx, KC_SCAN, st, 5);
            exclusive_scan_pairs(SegLoad{d_offsets, T, S}, n_reads, (Pair64*)lane.tiles.p, nullptr, seg_off + n_reads + 2, seg_off, nullptr, st,
                                 TileDescPost{(uint64_t*)lane.seg_desc.p});
            seg_spec_kernel<<<blocks_for(max_segs, 256), 256, 0, st>>>(ctx->dix, d_offsets, T, S, scratch, (const uint64_t*)lane.seg_desc.p,
                                                                       seg_off + n_reads, (uint64_t*)lane.seg_bits.p, (uint32_t*)lane.seg_info.p, status);
            seg_join_kernel<<<blocks_for(n_reads, 256), 256, 0, st>>>(ctx->dix, d_offsets, n_reads, T, S, scratch, seg_off, (const uint64_t*)lane.seg_bits.p,
                                                                      (uint32_t*)lane.seg_info.p, counts, status);
            lane.seg_len = S;
        } else {
            Timed t(ctx, KC_SCAN, st);
            count_kernel<<<blocks_for(n_reads, 256), 256, 0, st>>>(ctx->dix, packed, d_offsets, n_reads, T, scratch, counts, status,
                                                                   keep_records ? 1 : 0);
        }
    }
    NTC_CUDA_TRY(cudaGetLastError());
    return NTC_OK;
}

// ------------------------------------------------------------------------------------------
// encode pipeline on device buffers, using the workspaces of `lane` on stream `st`.
// MODE 0 records, 1 triples.  carry (device, 2 x u64, may be null) is added to every record
// offset; the grand total lands in totals_out (device, 2 x u64).  `rel` (device u64, may be null):
// records are written at out[offset - *rel] (chunked host path: out is the chunk's staging buffer).
// wait_scan: an event the scan must wait for (the previous chunk's scan on the other lane).
// ------------------------------------------------------------------------------------------
// d_bases == nullptr: lane.packed already holds the batch's 2-bit stream (packed on the host, api_host.inc).
int32_t encode_pipeline(ntc_ctx* ctx, Lane& lane, int mode, const uint8_t* d_bases, const uint64_t* d_offsets, uint64_t n_reads,
                        uint64_t n_bases, uint64_t* d_out, uint64_t cap, uint64_t* d_rec_offsets, const uint64_t* carry,
                        uint64_t* totals_out, const uint64_t* rel, cudaEvent_t wait_scan, cudaStream_t st) {
    if (n_reads == 0) {
        if (wait_scan) NTC_CUDA_TRY(cudaStreamWaitEvent(st, wait_scan, 0));
        NTC_CUDA_TRY(cudaMemsetAsync(totals_out, 0, 16, st));
        if (carry) NTC_CUDA_TRY(cudaMemcpyAsync(totals_out, carry, 16, cudaMemcpyDeviceToDevice, st));
        if (d_rec_offsets) NTC_CUDA_TRY(cudaMemcpyAsync(d_rec_offsets, totals_out, 8, cudaMemcpyDeviceToDevice, st));
        return NTC_OK;
    }
    const uint64_t n_words = n_bases / 32 + 2;
    NTC_CUDA_TRY(ensure(&lane.packed, (n_words + 4) * 8));
    NTC_CUDA_TRY(ensure(&lane.scratch, scratch_entries(n_bases, n_reads) * 8));
    NTC_CUDA_TRY(ensure(&lane.counts, n_reads * 4));
    NTC_CUDA_TRY(ensure(&lane.tiles, scan_num_tiles(n_reads) * sizeof(Pair64)));
    uint64_t* rec_off = d_rec_offsets;
    if (!rec_off) {
        NTC_CUDA_TRY(ensure(&lane.rec_off, (n_reads + 1) * 8));
        rec_off = (uint64_t*)lane.rec_off.p;
    }
    uint64_t* packed = (uint64_t*)lane.packed.p;
    uint64_t* scratch = (uint64_t*)lane.scratch.p;
    uint32_t* counts = (uint32_t*)lane.counts.p;

    if (n_bases && d_bases) {
        Timed t(ctx, KC_PACK, st);
        pack2bit_kernel<<<blocks_for((n_bases + 31) / 32, kPackThreads), kPackThreads, 0, st>>>(d_bases, d_offsets, n_bases, packed,
                                                                                                ctx->d_status);
    }
    const uint32_t T = tile_len_for(ctx, n_bases);
    int32_t rc = ms_pipeline(ctx, lane, packed, d_offsets, n_reads, n_bases, T, scratch, counts, ctx->d_status, st, mode == 0);
    if (rc) return rc;
    if (wait_scan) NTC_CUDA_TRY(cudaStreamWaitEvent(st, wait_scan, 0));
    {
        Timed t(ctx, KC_SCAN, st, 3);
        exclusive_scan_pairs(CountLoad{counts}, n_reads, (Pair64*)lane.tiles.p, carry, totals_out, rec_off, nullptr, st);
    }
    if (lane.ev_scan) NTC_CUDA_TRY(cudaEventRecord(lane.ev_scan, st));
    {
        Timed t(ctx, KC_EMIT, st);
        if (mode == 0) {
            emit_kernel<0><<<blocks_for(n_reads, kMsThreads), kMsThreads, 0, st>>>(ctx->dix, packed, d_offsets, n_reads, scratch, counts,
                                                                                   rec_off, rel, d_out, cap, ctx->d_status, T, lane.seg_len ? 1 : 0);
            if (lane.seg_len) {
                ctx->launches++; ctx->class_launches[KC_EMIT]++;
                seg_emit_kernel<<<blocks_for(n_bases / lane.seg_len + n_reads, 256), 256, 0, st>>>(
                    ctx->dix, packed, d_offsets, T, lane.seg_len, scratch, (const uint64_t*)lane.seg_desc.p, (const uint64_t*)lane.seg_off.p + n_reads,
                    (const uint32_t*)lane.seg_info.p, counts, rec_off, rel, d_out, cap, ctx->d_status);
            }
        } else {
            emit_kernel<1><<<blocks_for(n_reads, kMsThreads), kMsThreads, 0, st>>>(ctx->dix, packed, d_offsets, n_reads, scratch, counts,
                                                                                   rec_off, rel, d_out, cap, ctx->d_status, T, 0);
        }
    }
    NTC_CUDA_TRY(cudaGetLastError());
    return NTC_OK;
}

// ------------------------------------------------------------------------------------------
// decode on device buffers with the workspaces of `lane`: scans + per-sequence layout (prepare), then
// FASTA headers + the walk over all records.  Counts land in d_counts ([0] sequences, [1] bytes).
// off_carry is added to every entry of d_seq_offsets (chunked host path).
// ------------------------------------------------------------------------------------------
int32_t decode_prepare(ntc_ctx* ctx, Lane& lane, const uint64_t* d_records, uint64_t n_rec, uint64_t first_no, int fasta, uint64_t cap,
                       uint64_t* d_seq_offsets, uint64_t cap_seqs, uint64_t off_carry, uint64_t* d_counts, cudaStream_t st) {
    NTC_CUDA_TRY(ensure(&lane.base_off, (n_rec + 1) * 8));
    NTC_CUDA_TRY(ensure(&lane.seq_idx, (n_rec + 1) * 8));
    NTC_CUDA_TRY(ensure(&lane.seq_first, (n_rec + 1) * 8));
    NTC_CUDA_TRY(ensure(&lane.seq_endkey, (n_rec + 1) * 8));
    NTC_CUDA_TRY(ensure(&lane.dec_tiles, scan_num_tiles(n_rec) * sizeof(Pair64)));
    NTC_CUDA_TRY(ensure(&lane.dec_small, 64));
    uint64_t* base_off = (uint64_t*)lane.base_off.p;
    uint64_t* seq_idx = (uint64_t*)lane.seq_idx.p;
    uint64_t* seq_first = (uint64_t*)lane.seq_first.p;
    uint64_t* totals = (uint64_t*)lane.dec_small.p;
    {
        Timed t(ctx, KC_DSCAN, st, 3);
        exclusive_scan_pairs(RecordLoad{d_records}, n_rec, (Pair64*)lane.dec_tiles.p, nullptr, totals, base_off, seq_idx, st,
                             SeqFirstPost{seq_first});
    }
    {
        Timed t(ctx, KC_SEQ, st, 1);
        seq_layout_kernel<<<blocks_for(n_rec + 1, 256), 256, 0, st>>>(base_off, seq_first, totals, n_rec, (uint64_t*)lane.seq_endkey.p,
                                                                      d_seq_offsets, cap_seqs, off_carry, first_no, fasta, cap, d_counts,
                                                                      ctx->d_status);
    }
    NTC_CUDA_TRY(cudaGetLastError());
    return NTC_OK;
}

int32_t decode_walk(ntc_ctx* ctx, Lane& lane, const uint64_t* d_records, uint64_t n_rec, uint64_t first_no, int fasta, uint8_t* d_out,
                    uint64_t cap, const uint64_t* d_counts, cudaStream_t st) {
    if (n_rec == 0) return NTC_OK;
    if (!(ctx->uses & NTC_USE_DECODE)) { set_error("this context was created without NTC_USE_DECODE (no jump table in HBM)"); return NTC_ERR_UNSUPPORTED; }
    {
        // (a two-pass variant -- walk to 2-bit slots per record, then one thread per sequence formatting header + bases
        // left to right -- measured slower on B200: 3.31 vs 2.56 ms on C3, 9.6 vs 6.0 ms on C5; r1 A/B, profiles/r1)
        Timed t(ctx, KC_WALK, st);
        walk_kernel<<<blocks_for(n_rec, kWalkThreads), kWalkThreads, 0, st>>>(ctx->dix, d_records, n_rec, (const uint64_t*)lane.base_off.p,
                                                            (const uint64_t*)lane.seq_idx.p, (const uint64_t*)lane.seq_endkey.p,
                                                            (const uint64_t*)lane.dec_small.p, first_no, fasta, d_out, cap, d_counts,
                                                            ctx->d_status);
    }
    NTC_CUDA_TRY(cudaGetLastError());
    return NTC_OK;
}

int32_t decode_pipeline(ntc_ctx* ctx, const uint64_t* d_records, uint64_t n_rec, uint64_t first_no, int fasta, uint8_t* d_out,
                        uint64_t cap, uint64_t* d_seq_offsets, uint64_t cap_seqs, uint64_t* d_counts, cudaStream_t st) {
    if (n_rec == 0) {
        NTC_CUDA_TRY(cudaMemsetAsync(d_counts, 0, 16, st));
        if (d_seq_offsets) NTC_CUDA_TRY(cudaMemsetAsync(d_seq_offsets, 0, 8, st));
        return NTC_OK;
    }
    Lane& lane = ctx->lanes[0];
    int32_t rc = decode_prepare(ctx, lane, d_records, n_rec, first_no, fasta, cap, d_seq_offsets, cap_seqs, 0, d_counts, st);
    if (rc) return rc;
    return decode_walk(ctx, lane, d_records, n_rec, first_no, fasta, d_out, cap, d_counts, st);
}

// ---- page-locked staging for pageable caller buffers (ctx.h) ---------------------------------------------------------------
bool is_pageable(const void* p) {
    if (!p) return false;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return true; }
    return a.type == cudaMemoryTypeUnregistered;
}
// the pool, if host threads may be used on this context (created on first use)
HostPool* host_pool(ntc_ctx* ctx) {
    const int n = ctx->host_threads > 0 ? ctx->host_threads : (ctx->host_threads < 0 ? ctx->auto_threads : 0);
    if (n <= 0) return nullptr;
    if (!ctx->pool) ctx->pool = new HostPool(n);
    return ctx->pool;
}
cudaError_t ensure_pinned(ntc_ctx::Pinned* h, size_t bytes) {
    if (bytes <= h->cap) return cudaSuccess;
    if (h->p) cudaFreeHost(h->p);
    h->p = nullptr; h->cap = 0;
    const size_t want = bytes + bytes / 8 + 4096;
    cudaError_t e = cudaHostAlloc(&h->p, want, cudaHostAllocPortable);
    if (e == cudaSuccess) h->cap = want;
    return e;
}
void pool_memcpy(HostPool& pool, void* dst, const void* src, size_t bytes) {
    const size_t piece = 2u << 20;
    pool.run((bytes + piece - 1) / piece, [&](uint64_t i) {
        const size_t a = (size_t)i * piece, n = std::min(piece, bytes - a);
        memcpy((uint8_t*)dst + a, (const uint8_t*)src + a, n);
    });
}

int32_t read_status(ntc_ctx* ctx, cudaStream_t st) {
    uint32_t h = 0;
    NTC_CUDA_TRY(cudaMemcpyAsync(&h, ctx->d_status, 4, cudaMemcpyDeviceToHost, st));
    NTC_CUDA_TRY(cudaStreamSynchronize(st));
    if (h) NTC_CUDA_TRY(cudaMemsetAsync(ctx->d_status, 0, 4, st));
    return status_to_code(h);
}

}  // namespace

void free_column_work(ntc_ctx* ctx);

extern "C" {

int32_t ntc_ctx_create(int32_t device, const ntc_index* idx, ntc_ctx** out) {
    return ntc_ctx_create_for(device, idx, NTC_USE_ENCODE | NTC_USE_DECODE, out);
}

int32_t ntc_ctx_create_for(int32_t device, const ntc_index* idx, uint32_t uses, ntc_ctx** out) {
    if (!idx || !out) { set_error("null argument"); return NTC_ERR_INVALID_ARG; }
    *out = nullptr;
    if ((uses & (NTC_USE_ENCODE | NTC_USE_DECODE)) == 0 || (uses & ~(uint32_t)(NTC_USE_ENCODE | NTC_USE_DECODE))) {
        set_error("uses must be NTC_USE_ENCODE, NTC_USE_DECODE or both");
        return NTC_ERR_INVALID_ARG;
    }
    const bool trace = getenv("NTC_TRACE") != nullptr;
    const auto t0 = std::chrono::steady_clock::now();
    auto since = [&]() { return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); };
    // The derived HBM layouts (rank blocks, LCS words, jump and prefix tables) are built by kernels from the canonical arrays
    // (layout_kernels.cuh): 1.5 B per node go up instead of 14.  NTC_HOST_LAYOUTS=1 (and builds with 16-bit LCS words) take the
    // first version's route: host threads build them WHILE the CUDA context comes up on another thread, then they are uploaded.
    // An encode-only / decode-only context (the CLI) holds only the arrays its path reads.
    const bool host_layouts = getenv("NTC_HOST_LAYOUTS") && atoi(getenv("NTC_HOST_LAYOUTS")) != 0;
    int n_dev = 0;
    cudaError_t e_init = cudaSuccess;
    double t_cuda = 0;
    std::thread cuda_up([&]() {
        e_init = cudaGetDeviceCount(&n_dev);
        if (e_init == cudaSuccess && device >= 0 && device < n_dev) {
            e_init = cudaSetDevice(device);
            if (e_init == cudaSuccess) e_init = cudaFree(0);
        }
        t_cuda = since();
    });
    const int threads = (int)std::max(1u, std::thread::hardware_concurrency());
    std::vector<uint32_t> rank, pfx;
    std::vector<ntc_lcsx_t> lcs;
    std::vector<uint64_t> jump;
    uint32_t c_start[4] = {0, 0, 0, 0};
    const uint32_t rank_nodes = rank_nodes_for(idx->n);
    const bool use_pfx = idx->k >= 10 && !getenv("NTC_NO_PREFIX_TABLE");   // a tile's warm-up (k-1 bases) must cover the 8 bases
    double t_rank = 0, t_lcs = 0, t_jump = 0;
    std::exception_ptr build_err;
    if (host_layouts) {
        std::thread lcs_thread;
        std::exception_ptr lcs_err;
        try {
            if (uses & NTC_USE_ENCODE)                                                                                          // serial stack pass
                lcs_thread = std::thread([&]() {
                    try { build_lcs_words(*idx, &lcs); t_lcs = since(); }
                    catch (...) { lcs_err = std::current_exception(); }   // an exception leaving a thread is std::terminate
                });
            if (uses & NTC_USE_ENCODE) {
                build_rank_blocks(*idx, &rank, std::max(1, threads - 1), rank_nodes);
                if (use_pfx) build_prefix_table(rank, idx->n, &pfx, c_start, rank_nodes);
                t_rank = since();
            }
            if (uses & NTC_USE_DECODE) { build_jump_table(*idx, &jump, std::max(1, threads - 1)); t_jump = since(); }
        } catch (...) { build_err = std::current_exception(); }
        if (lcs_thread.joinable()) lcs_thread.join();
        if (lcs_err && !build_err) build_err = lcs_err;
    } else {
        for (int c = 0; c < 4; c++) c_start[c] = (uint32_t)idx->C[c];   // interval start after the single base c = rank_c(0) + C[c]
    }
    cuda_up.join();
    if (build_err) { set_error("out of host memory while deriving the index layouts"); return NTC_ERR_IO; }   // as container.cpp reports bad_alloc
    cudaError_t e = e_init;
    if (e != cudaSuccess || n_dev == 0) {
        set_error(std::string("no CUDA device available (this library has no CPU fallback): ") + cudaGetErrorString(e));
        return NTC_ERR_CUDA;
    }
    if (device < 0 || device >= n_dev) { set_error("device ordinal out of range"); return NTC_ERR_INVALID_ARG; }
    NTC_CUDA_TRY(cudaSetDevice(device));
    ntc_ctx* ctx = new ntc_ctx();
    ctx->device = device;
    ctx->k = idx->k;
    ctx->n = idx->n;
    ctx->uses = uses;
    memcpy(ctx->dix.c_start, c_start, sizeof c_start);
    if (const char* env = getenv("NTC_MAX_CHUNK_BASES")) {
        unsigned long long v = strtoull(env, nullptr, 10);
        if (v >= 1024) ctx->max_chunk_bases = v;
    }
    if (const char* env = getenv("NTC_COPY_DIRECTIONS")) ctx->copy_serial = strcmp(env, "serial") == 0;
    if (const char* env = getenv("NTC_MAX_CHUNK_RECORDS")) {
        unsigned long long v = strtoull(env, nullptr, 10);
        if (v >= 16) ctx->max_chunk_records = v;
    }
    auto fail = [&](const char* what, cudaError_t err) {
        set_error(std::string(what) + ": " + cudaGetErrorString(err));
        ntc_ctx_destroy(ctx);
        return NTC_ERR_CUDA;
    };
    const uint64_t n = idx->n;
    const uint64_t n_rank_blocks = n / rank_nodes + 1;                 // position n (interval end) must be addressable
    ctx->index_bytes[0] = (uses & NTC_USE_ENCODE) ? n_rank_blocks * 64 : 0;
    ctx->index_bytes[1] = (uses & NTC_USE_ENCODE) ? (n + 1) * sizeof(ntc_lcsx_t) : 0;
    ctx->index_bytes[2] = (uses & NTC_USE_DECODE) ? n * 8 : 0;
    const size_t pfx_bytes = (size_t)(1u << (2 * kPrefixBases)) * 8 * 4;
    if (uses & NTC_USE_ENCODE) {
        if ((e = cudaMalloc(&ctx->d_rank, ctx->index_bytes[0])) != cudaSuccess) return fail("cudaMalloc rank", e);
        if ((e = cudaMalloc(&ctx->d_lcs, ctx->index_bytes[1])) != cudaSuccess) return fail("cudaMalloc lcs", e);
        if (use_pfx && (e = cudaMalloc(&ctx->d_pfx, pfx_bytes)) != cudaSuccess) return fail("cudaMalloc prefix table", e);
    }
    if ((uses & NTC_USE_DECODE) && (e = cudaMalloc(&ctx->d_jump, ctx->index_bytes[2])) != cudaSuccess) return fail("cudaMalloc jump", e);
    if (host_layouts) {
        if (uses & NTC_USE_ENCODE) {
            if ((e = cudaMemcpy(ctx->d_rank, rank.data(), ctx->index_bytes[0], cudaMemcpyHostToDevice)) != cudaSuccess) return fail("upload rank", e);
            if ((e = cudaMemcpy(ctx->d_lcs, lcs.data(), ctx->index_bytes[1], cudaMemcpyHostToDevice)) != cudaSuccess) return fail("upload lcs", e);
            if (use_pfx && (e = cudaMemcpy(ctx->d_pfx, pfx.data(), pfx_bytes, cudaMemcpyHostToDevice)) != cudaSuccess) return fail("upload prefix table", e);
        }
        if ((uses & NTC_USE_DECODE) && (e = cudaMemcpy(ctx->d_jump, jump.data(), ctx->index_bytes[2], cudaMemcpyHostToDevice)) != cudaSuccess)
            return fail("upload jump", e);
    } else {
        // canonical arrays up (4 rows + LCS bytes), ones-before-word of every row by two scans, then one kernel per layout
        const uint64_t row_words = (n + 63) / 64;
        uint64_t *d_rows = nullptr, *d_pre = nullptr, *d_tmp = nullptr, *d_tot = nullptr;
        uint8_t* d_lcs8 = nullptr;
        Pair64* d_tiles = nullptr;
        auto cleanup = [&]() { cudaFree(d_rows); cudaFree(d_pre); cudaFree(d_tmp); cudaFree(d_tot); cudaFree(d_lcs8); cudaFree(d_tiles); };
        auto fail2 = [&](const char* what, cudaError_t err) { cleanup(); return fail(what, err); };
        if ((e = cudaMalloc(&d_rows, 4 * row_words * 8)) != cudaSuccess) return fail2("cudaMalloc rows", e);
        if ((e = cudaMalloc(&d_pre, 4 * (row_words + 1) * 8)) != cudaSuccess) return fail2("cudaMalloc row prefixes", e);
        if ((e = cudaMalloc(&d_lcs8, n + 1)) != cudaSuccess) return fail2("cudaMalloc lcs bytes", e);
        if ((e = cudaMalloc(&d_tiles, scan_num_tiles(row_words) * sizeof(Pair64))) != cudaSuccess) return fail2("cudaMalloc scan tiles", e);
        if ((e = cudaMalloc(&d_tot, 16)) != cudaSuccess) return fail2("cudaMalloc scan totals", e);
        for (int c = 0; c < 4; c++)
            if ((e = cudaMemcpy(d_rows + c * row_words, idx->rows[c].data(), row_words * 8, cudaMemcpyHostToDevice)) != cudaSuccess) return fail2("upload rows", e);
        if ((e = cudaMemcpy(d_lcs8, idx->lcs.data(), n, cudaMemcpyHostToDevice)) != cudaSuccess) return fail2("upload lcs bytes", e);
        lay::Rows rows;
        rows.row_words = row_words;
        for (int c = 0; c < 4; c++) { rows.row[c] = d_rows + c * row_words; rows.pre[c] = d_pre + c * (row_words + 1); }
        for (int c = 0; c < 5; c++) rows.C[c] = idx->C[c];
        for (int c = 0; c < 4; c += 2)
            ctx->launches += exclusive_scan_pairs(lay::RowPopLoad{rows.row[c], rows.row[c + 1]}, row_words, d_tiles, nullptr, d_tot,
                                                  d_pre + c * (row_words + 1), d_pre + (c + 1) * (row_words + 1), 0);
        if (uses & NTC_USE_ENCODE) {
            if (rank_nodes == 64) lay::rank_blocks_kernel<64><<<blocks_for(n_rank_blocks, 256), 256>>>(rows, d_lcs8, n, idx->k, (uint32_t*)ctx->d_rank, n_rank_blocks);
            else lay::rank_blocks_kernel<32><<<blocks_for(n_rank_blocks, 256), 256>>>(rows, d_lcs8, n, idx->k, (uint32_t*)ctx->d_rank, n_rank_blocks);
            lay::lcs_words_kernel<<<blocks_for(n + 1, 256), 256>>>(d_lcs8, n, (uint32_t*)ctx->d_lcs);
            if (use_pfx) {
                if (rank_nodes == 64) lay::prefix_table_kernel<64><<<256, 256>>>((const uint32_t*)ctx->d_rank, n, (uint32_t*)ctx->d_pfx);
                else lay::prefix_table_kernel<32><<<256, 256>>>((const uint32_t*)ctx->d_rank, n, (uint32_t*)ctx->d_pfx);
            }
            ctx->launches += use_pfx ? 3 : 2;
        }
        if (uses & NTC_USE_DECODE) {
            if ((e = cudaMalloc(&d_tmp, n * 8)) != cudaSuccess) return fail2("cudaMalloc jump scratch", e);
            // four doubling rounds ping-pong between the scratch and the table: start in the table so that the last round lands in it
            uint64_t* a = (uint64_t*)ctx->d_jump;
            uint64_t* b = d_tmp;
            lay::pred_init_kernel<<<blocks_for(row_words * 4, 256), 256>>>(rows, a);
            for (uint32_t t = 1; t < 16; t *= 2) {
                lay::pred_double_kernel<<<blocks_for(n, 256), 256>>>(a, b, t, n);
                std::swap(a, b);
            }
            ctx->launches += 5;                                        // a == d_jump again after four swaps
        }
        e = cudaDeviceSynchronize();
        if (e == cudaSuccess) e = cudaGetLastError();
        cleanup();
        if (e != cudaSuccess) return fail("deriving the index layouts on the device", e);
        t_rank = t_lcs = t_jump = since();
    }
    if ((e = cudaMalloc((void**)&ctx->d_status, 16)) != cudaSuccess) return fail("cudaMalloc status", e);
    if ((e = cudaMalloc((void**)&ctx->d_totals, 64)) != cudaSuccess) return fail("cudaMalloc totals", e);
    cudaMemset(ctx->d_status, 0, 16);
    cudaMemset(ctx->d_totals, 0, 64);
    for (Lane& ln : ctx->lanes) {
        if ((e = cudaStreamCreateWithFlags(&ln.stream, cudaStreamNonBlocking)) != cudaSuccess) return fail("cudaStreamCreate", e);
        if ((e = cudaEventCreateWithFlags(&ln.ev_scan, cudaEventDisableTiming)) != cudaSuccess) return fail("cudaEventCreate", e);
        if ((e = cudaEventCreateWithFlags(&ln.ev_done, cudaEventDisableTiming)) != cudaSuccess) return fail("cudaEventCreate", e);
        if ((e = cudaHostAlloc((void**)&ln.h_land, 64, cudaHostAllocMapped)) != cudaSuccess) return fail("cudaHostAlloc", e);
    }
    for (cudaEvent_t& ev : ctx->ev_out)
        if ((e = cudaEventCreateWithFlags(&ev, cudaEventDisableTiming)) != cudaSuccess) return fail("cudaEventCreate", e);
    ctx->stream = ctx->lanes[0].stream;
    ctx->dix.rank = (const uint4*)ctx->d_rank;
    ctx->dix.lcsx = (const ntc_lcsx_t*)ctx->d_lcs;
    ctx->dix.jump = (const uint64_t*)ctx->d_jump;
    ctx->dix.pfx = (const uint4*)ctx->d_pfx;
    ctx->dix.n = (uint32_t)idx->n;
    ctx->dix.k = idx->k;
    ctx->dix.rank_nodes = rank_nodes;
    spec_window(idx->n, &ctx->dix.spec_lo, &ctx->dix.spec_span);
    {
        // L2 cache-hint descriptors, made once on the device and handed to every kernel as (uniform) parameters
        policy_kernel<<<1, 1>>>((uint64_t*)ctx->d_totals);
        uint64_t pol[2] = {0, 0};
        if ((e = cudaMemcpy(pol, ctx->d_totals, 16, cudaMemcpyDeviceToHost)) != cudaSuccess) return fail("policy_kernel", e);
        cudaMemset(ctx->d_totals, 0, 64);
        ctx->dix.pol_keep = pol[0];
        ctx->dix.pol_stream = pol[1];
        // A persisting carve-out (cudaLimitPersistingL2CacheSize) is NOT set by default: measured on B200 (C3,
        // profiles/r1/ab_l2_persist.txt) a 64 / 79 MB set-aside slows ms_kernel from 15.3 to 20.1 / 26.4 ms, 32 MB changes
        // nothing -- the carve-out shrinks the L2 left for everything else.  (The evict_last / evict_first hints by
        // themselves measure neutral on v12: 12.01 ms without them, profiles/r1/ab_v12_variants.txt.)
    }
    ntc_ctx_set_host_threads(ctx, getenv("NTC_HOST_THREADS") ? atoi(getenv("NTC_HOST_THREADS")) : -1);   // default: staging for pageable caller buffers only
    if (const char* env = getenv("NTC_TILE_LEN")) ntc_ctx_set_tile_len(ctx, (uint32_t)strtoul(env, nullptr, 10));
    if (const char* env = getenv("NTC_SPEC_LO")) ctx->dix.spec_lo = (uint32_t)strtoul(env, nullptr, 10);      // tuning knobs
    if (const char* env = getenv("NTC_SPEC_SPAN")) ctx->dix.spec_span = (uint32_t)strtoul(env, nullptr, 10);
    if (trace && host_layouts)
        fprintf(stderr, "[ntc trace] ctx_create %.3f s: CUDA context up at %.3f (own thread) | host layouts: rank+prefix done at %.3f, lcs at %.3f, "
                "jump at %.3f | uploads + policies %.3f\n", since(), t_cuda, t_rank, t_lcs, t_jump, since() - std::max(std::max(t_cuda, t_jump), std::max(t_rank, t_lcs)));
    else if (trace)
        fprintf(stderr, "[ntc trace] ctx_create %.3f s: CUDA context up at %.3f | canonical arrays up + layouts by kernels done at %.3f | policies %.3f\n",
                since(), t_cuda, t_rank, since() - t_rank);
    *out = ctx;
    return NTC_OK;
}

void ntc_ctx_destroy(ntc_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    for (Lane& ln : ctx->lanes) {
        if (ln.stream) cudaStreamDestroy(ln.stream);
        if (ln.ev_scan) cudaEventDestroy(ln.ev_scan);
        if (ln.ev_done) cudaEventDestroy(ln.ev_done);
        if (ln.h_land) cudaFreeHost(ln.h_land);
        DevBuf* lb[] = {&ln.packed, &ln.scratch, &ln.counts, &ln.tiles, &ln.rec_off, &ln.tile_off, &ln.desc, &ln.seg_off, &ln.seg_desc, &ln.seg_bits, &ln.seg_info, &ln.base_off, &ln.seq_idx, &ln.seq_first, &ln.seq_endkey, &ln.dec_tiles, &ln.dec_small, &ln.st_in, &ln.st_off, &ln.st_out, &ln.st_out2};
        for (DevBuf* b : lb) release(b);
    }
    free_column_work(ctx);
    release(&ctx->hold); release(&ctx->hold2);
    delete ctx->pool;
    for (auto& h : ctx->h_in) if (h.p) cudaFreeHost(h.p);
    for (auto& h : ctx->h_out) if (h.p) cudaFreeHost(h.p);
    for (cudaEvent_t ev : ctx->ev_out) if (ev) cudaEventDestroy(ev);
    for (auto& sp : ctx->spans) { cudaEventDestroy(sp.e0); cudaEventDestroy(sp.e1); }
    for (auto ev : ctx->event_pool) cudaEventDestroy(ev);
    cudaFree(ctx->d_rank); cudaFree(ctx->d_lcs); cudaFree(ctx->d_jump); cudaFree(ctx->d_pfx);
    cudaFree(ctx->d_status); cudaFree(ctx->d_totals);
    delete ctx;
}

int32_t ntc_pinned_alloc(uint64_t bytes, void** out) {
    if (!out) { set_error("null argument"); return NTC_ERR_INVALID_ARG; }
    *out = nullptr;
    cudaError_t e = cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocPortable);
    if (e != cudaSuccess) { *out = nullptr; set_error(std::string("cudaHostAlloc: ") + cudaGetErrorString(e)); return NTC_ERR_CUDA; }
    return NTC_OK;
}
void ntc_pinned_free(void* p) { if (p) cudaFreeHost(p); }

int32_t ntc_ctx_bind_thread(ntc_ctx* ctx) {
    if (!ctx) { set_error("null ctx"); return NTC_ERR_INVALID_ARG; }
    NTC_CUDA_TRY(cudaSetDevice(ctx->device));
    return NTC_OK;
}

void ntc_ctx_index_bytes(const ntc_ctx* ctx, uint64_t out[3]) { memcpy(out, ctx->index_bytes, 24); }

int32_t ntc_ctx_layout_read(ntc_ctx* ctx, int32_t which, void* out, uint64_t cap, uint64_t* n_bytes) {
    if (!ctx || !n_bytes || which < 0 || which > 3) { set_error("invalid argument"); return NTC_ERR_INVALID_ARG; }
    const void* src[4] = {ctx->d_rank, ctx->d_lcs, ctx->d_jump, ctx->d_pfx};
    const uint64_t size[4] = {ctx->index_bytes[0], ctx->index_bytes[1], ctx->index_bytes[2], ctx->d_pfx ? (uint64_t)(1u << (2 * kPrefixBases)) * 32 : 0};
    *n_bytes = src[which] ? size[which] : 0;
    if (!out || *n_bytes == 0) return NTC_OK;
    if (cap < *n_bytes) { set_error("output buffer too small"); return NTC_ERR_CAPACITY; }
    NTC_CUDA_TRY(cudaSetDevice(ctx->device));
    NTC_CUDA_TRY(cudaMemcpy(out, src[which], *n_bytes, cudaMemcpyDeviceToHost));
    return NTC_OK;
}
uint64_t ntc_ctx_launch_count(const ntc_ctx* ctx) { return ctx ? ctx->launches : 0; }

int32_t ntc_ctx_set_tile_len(ntc_ctx* ctx, uint32_t tile_len) {
    if (!ctx) { set_error("null ctx"); return NTC_ERR_INVALID_ARG; }
    if (tile_len && tile_len < kTileMin) { set_error("tile length must be 0 (automatic) or >= 256"); return NTC_ERR_INVALID_ARG; }
    if (tile_len > (1u << 22)) tile_len = 1u << 22;
    ctx->tile_len = (tile_len + 3u) & ~3u;
    return NTC_OK;
}

int32_t ntc_ctx_set_host_threads(ntc_ctx* ctx, int32_t n_threads) {
    if (!ctx) { set_error("null argument"); return NTC_ERR_INVALID_ARG; }
    int n_dev = 1;
    cudaGetDeviceCount(&n_dev);
    int cpus = (int)std::thread::hardware_concurrency();
#if defined(__linux__)
    { cpu_set_t set; if (sched_getaffinity(0, sizeof set, &set) == 0 && CPU_COUNT(&set) > 0) cpus = CPU_COUNT(&set); }   // the CPUs this process may run on
#endif
    const int share = cpus / std::max(1, n_dev);   // this GPU's share of them (processes that see one GPU each: set NTC_HOST_THREADS)
    ctx->auto_threads = share >= 2 ? std::min(share, 16) : 0;
    ctx->host_threads = n_threads < 0 ? -1 : std::min(n_threads, 64);
    const int want = ctx->host_threads > 0 ? ctx->host_threads : ctx->auto_threads;
    if (ctx->pool && ctx->pool->threads() != want) { delete ctx->pool; ctx->pool = nullptr; }
    return NTC_OK;
}

int32_t ntc_ctx_set_copy_directions(ntc_ctx* ctx, int32_t serial) {
    if (!ctx) { set_error("null argument"); return NTC_ERR_INVALID_ARG; }
    ctx->copy_serial = serial != 0;
    return NTC_OK;
}

int32_t ntc_ctx_status(ntc_ctx* ctx, void* stream) {
    if (!ctx) { set_error("null ctx"); return NTC_ERR_INVALID_ARG; }
    NTC_CUDA_TRY(cudaSetDevice(ctx->device));
    return read_status(ctx, (cudaStream_t)stream);
}

// ---------------- device-resident entry points ----------------

int32_t ntc_encode_device(ntc_ctx* ctx, const uint8_t* d_bases, const uint64_t* d_offsets, uint64_t n_reads, uint64_t n_bases,
                          uint64_t* d_records, uint64_t cap, uint64_t* d_rec_offsets, uint64_t* d_counts, void* stream) {
    if (!ctx || !d_offsets || (!d_records && cap)) { set_error("null argument"); return NTC_ERR_INVALID_ARG; }
    NTC_CUDA_TRY(cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    int32_t rc = encode_pipeline(ctx, ctx->lanes[0], 0, d_bases, d_offsets, n_reads, n_bases, d_records, cap, d_rec_offsets, nullptr,
                                 ctx->d_totals, nullptr, nullptr, st);
    if (rc) return rc;
    if (d_counts) { Timed t(ctx, KC_MISC, st); copy_totals_kernel<<<1, 1, 0, st>>>(ctx->d_totals, d_counts); }
    return NTC_OK;
}

int32_t ntc_decode_device(ntc_ctx* ctx, const uint64_t* d_records, uint64_t n_records, uint64_t first_seq_number, int32_t fasta,
                          uint8_t* d_out, uint64_t cap, uint64_t* d_seq_offsets, uint64_t cap_seqs, uint64_t* d_counts, void* stream) {
    if (!ctx || (!d_records && n_records) || !d_counts) { set_error("null argument"); return NTC_ERR_INVALID_ARG; }
    if (fasta && first_seq_number == 0) { set_error("first_seq_number is 1-based"); return NTC_ERR_INVALID_ARG; }
    NTC_CUDA_TRY(cudaSetDevice(ctx->device));
    return decode_pipeline(ctx, d_records, n_records, first_seq_number, fasta, d_out, cap, d_seq_offsets, cap_seqs, d_counts,
                           (cudaStream_t)stream);
}

int32_t ntc_synth_reads_device(ntc_ctx* ctx, const uint8_t* d_genome, uint64_t genome_len, uint64_t seed, uint64_t r0, uint64_t n_reads,
                               uint32_t read_len, uint32_t err_ppm, uint8_t* d_out, void* stream) {
    if (!ctx || !d_genome || !d_out || read_len == 0 || genome_len < read_len) { set_error("bad argument"); return NTC_ERR_INVALID_ARG; }
    NTC_CUDA_TRY(cudaSetDevice(ctx->device));
    uint64_t total = n_reads * read_len;
    if (total) synth_reads_kernel<<<blocks_for(total, 256), 256, 0, (cudaStream_t)stream>>>(d_genome, genome_len, seed, r0, n_reads, read_len,
                                                                                             err_ppm, d_out);
    NTC_CUDA_TRY(cudaGetLastError());
    return NTC_OK;
}

// ---------------- measurement support ----------------

int32_t ntc_ctx_set_profiling(ntc_ctx* ctx, int32_t on) {
    if (!ctx) { set_error("null ctx"); return NTC_ERR_INVALID_ARG; }
    ctx->profiling = on != 0;
    return NTC_OK;
}

// Synchronises the device, folds the recorded event spans into per-class totals and returns them:
// ms[KC_COUNT] accumulated milliseconds, launches[KC_COUNT] kernel launches; reset != 0 clears.
int32_t ntc_ctx_kernel_times(ntc_ctx* ctx, double* ms, uint64_t* launches, int32_t reset) {
    if (!ctx) { set_error("null ctx"); return NTC_ERR_INVALID_ARG; }
    NTC_CUDA_TRY(cudaSetDevice(ctx->device));
    NTC_CUDA_TRY(cudaDeviceSynchronize());
    for (auto& sp : ctx->spans) {
        float t = 0.f;
        if (cudaEventElapsedTime(&t, sp.e0, sp.e1) == cudaSuccess) ctx->class_ms[sp.cls] += t;
        ctx->event_pool.push_back(sp.e0);
        ctx->event_pool.push_back(sp.e1);
    }
    ctx->spans.clear();
    for (int i = 0; i < KC_COUNT; i++) {
        if (ms) ms[i] = ctx->class_ms[i];
        if (launches) launches[i] = ctx->class_launches[i];
        if (reset) { ctx->class_ms[i] = 0; ctx->class_launches[i] = 0; }
    }
    return NTC_OK;
}

}  // extern "C"

// L2 random-sector micro-benchmark: every thread issues `loads` aligned 16-byte loads
// (one 32-byte sector each) at pseudo-random offsets of an L2-resident buffer.  dependent == 0:
// addresses come from a counter hash (independent, 8 in flight per thread); dependent != 0: each
// address depends on the previous load (pointer chase, the access pattern of a rank chain).
namespace {
template <int DEP>
__global__ void __launch_bounds__(256) l2_probe_kernel(const uint4* __restrict__ buf, uint32_t mask, uint32_t loads, uint32_t* sink) {
    uint32_t tid = blockIdx.x * 256 + threadIdx.x;
    uint32_t acc = 0;
    if (DEP) {
        uint32_t x = tid * 2654435761u;
        for (uint32_t i = 0; i < loads; i++) {
            uint4 v = __ldg(buf + (x & mask));
            x = x * 1664525u + 1013904223u + v.x;
            acc += v.y;
        }
        acc += x;
    } else {
        for (uint32_t i = 0; i < loads; i += 8) {
            uint4 v[8];
#pragma unroll
            for (int j = 0; j < 8; j++) {
                uint64_t h = ntc::splitmix64(((uint64_t)tid << 32) | (i + j));
                v[j] = __ldg(buf + ((uint32_t)h & mask));
            }
#pragma unroll
            for (int j = 0; j < 8; j++) acc += v[j].x;
        }
    }
    if (acc == 0x12345678u) *sink = acc;
}
}  // namespace

extern "C" int32_t ntc_l2_random_bench(ntc_ctx* ctx, uint64_t buffer_bytes, uint32_t loads_per_thread, int32_t dependent,
                                       uint32_t n_blocks, double* gsectors_per_s) {
    if (!ctx || !gsectors_per_s || buffer_bytes < 4096) { set_error("bad argument"); return NTC_ERR_INVALID_ARG; }
    NTC_CUDA_TRY(cudaSetDevice(ctx->device));
    uint64_t entries = 1;
    while (entries * 2 * 16 <= buffer_bytes) entries *= 2;
    DevBuf buf;
    NTC_CUDA_TRY(ensure(&buf, entries * 16));
    NTC_CUDA_TRY(cudaMemset(buf.p, 1, entries * 16));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    loads_per_thread = (loads_per_thread + 7) / 8 * 8;
    float best = 1e30f;
    for (int rep = 0; rep < 4; rep++) {
        cudaEventRecord(e0, ctx->stream);
        if (dependent) l2_probe_kernel<1><<<n_blocks, 256, 0, ctx->stream>>>((const uint4*)buf.p, (uint32_t)(entries - 1), loads_per_thread, ctx->d_status + 2);
        else l2_probe_kernel<0><<<n_blocks, 256, 0, ctx->stream>>>((const uint4*)buf.p, (uint32_t)(entries - 1), loads_per_thread, ctx->d_status + 2);
        cudaEventRecord(e1, ctx->stream);
        cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep > 0) best = std::min(best, ms);
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    release(&buf);
    NTC_CUDA_TRY(cudaGetLastError());
    *gsectors_per_s = (double)n_blocks * 256.0 * loads_per_thread / (best * 1e-3) / 1e9;
    return NTC_OK;
}

// ---------------- decode without any allocation: ntc_decode_device_static ------------------------------------------------
// ntc_decode_device keeps its scan workspaces in grow-only cudaMalloc'ed buffers, so a batch larger than any seen before
// makes it call cudaFree + cudaMalloc.  Some hosts of this library cannot have that happen inside a call:
//   * a host that handed the whole of HBM to its own pool (a database or an inference server with a caching allocator):
//     a cudaMalloc behind its back fails or fragments the pool;
//   * a latency-bound service: cudaMalloc / cudaFree synchronise the device and take a driver lock;
//   * a sandboxed plugin whose policy allows launches but not the allocation calls of the runtime.
// For them the same kernels run here on scratch that lives in the library image: one file-scope __device__ array, sized
// for kStaticMaxRecords records.  Every GPU that loads the library carries the array (about 16 MiB) until the process ends.
//
// In host code `g_static_scratch` is only a handle for the symbol -- NOT a device address; handing it to a kernel would
// fault (or scribble over host memory on a system where the GPU can reach it).  The address is looked up per device with
// cudaGetSymbolAddress while that device is current, and cached.
namespace {
constexpr uint64_t kStaticMaxRecords = NTC_STATIC_MAX_RECORDS;
constexpr size_t static_round(size_t b) { return (b + 255) / 256 * 256; }
constexpr size_t kStaticArrayBytes = static_round((kStaticMaxRecords + 1) * 8);     // base_off, seq_idx, seq_first, seq_endkey
constexpr size_t kStaticTileBytes = static_round(((kStaticMaxRecords + 1 + kScanTile - 1) / kScanTile) * sizeof(Pair64));
constexpr size_t kStaticSmallBytes = 256;                                           // scan totals (2 x u64)
constexpr size_t kStaticScratchBytes = 4 * kStaticArrayBytes + kStaticTileBytes + kStaticSmallBytes;
constexpr int kStaticMaxDevices = 64;

static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

std::mutex g_static_mu;
unsigned char* g_static_addr[kStaticMaxDevices] = {nullptr};

// device address of the array on `device`, which must be the current device; looked up once per device
int32_t static_scratch_address(int device, unsigned char** out) {
    if (device < 0 || device >= kStaticMaxDevices) { set_error("ntc_decode_device_static: device ordinal beyond 63"); return NTC_ERR_UNSUPPORTED; }
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_addr[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || !p) {
            set_error(std::string("cudaGetSymbolAddress(static decode scratch): ") + cudaGetErrorString(e));
            return NTC_ERR_CUDA;
        }
        g_static_addr[device] = (unsigned char*)p;
    }
    *out = g_static_addr[device];
    return NTC_OK;
}
}  // namespace

extern "C" uint64_t ntc_decode_static_max_records(void) { return kStaticMaxRecords; }

extern "C" int32_t ntc_decode_device_static(ntc_ctx* ctx, const uint64_t* d_records, uint64_t n_records, uint64_t first_seq_number,
                                            int32_t fasta, uint8_t* d_out, uint64_t cap, uint64_t* d_seq_offsets, uint64_t cap_seqs,
                                            uint64_t* d_counts, void* stream) {
    if (!ctx || (!d_records && n_records) || !d_counts) { set_error("null argument"); return NTC_ERR_INVALID_ARG; }
    if (fasta && first_seq_number == 0) { set_error("first_seq_number is 1-based"); return NTC_ERR_INVALID_ARG; }
    if (n_records > kStaticMaxRecords) {
        set_error("ntc_decode_device_static: " + std::to_string(n_records) + " records exceed the static scratch's limit of " +
                  std::to_string(kStaticMaxRecords) + " per call (split the batch at a `first` record, or use ntc_decode_device)");
        return NTC_ERR_CAPACITY;
    }
    if (n_records && !(ctx->uses & NTC_USE_DECODE)) { set_error("this context was created without NTC_USE_DECODE (no jump table in HBM)"); return NTC_ERR_UNSUPPORTED; }
    NTC_CUDA_TRY(cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    unsigned char* base = nullptr;
    if (int32_t rc = static_scratch_address(ctx->device, &base)) return rc;
    if (n_records == 0) {
        NTC_CUDA_TRY(cudaMemsetAsync(d_counts, 0, 16, st));
        if (d_seq_offsets) NTC_CUDA_TRY(cudaMemsetAsync(d_seq_offsets, 0, 8, st));
        return NTC_OK;
    }
    uint64_t* base_off = (uint64_t*)base;
    uint64_t* seq_idx = (uint64_t*)(base + kStaticArrayBytes);
    uint64_t* seq_first = (uint64_t*)(base + 2 * kStaticArrayBytes);
    uint64_t* seq_endkey = (uint64_t*)(base + 3 * kStaticArrayBytes);
    Pair64* tiles = (Pair64*)(base + 4 * kStaticArrayBytes);
    uint64_t* totals = (uint64_t*)(base + 4 * kStaticArrayBytes + kStaticTileBytes);
    // The array still holds the previous call's values.  Every word the kernels read is written earlier in the same call
    // (scan outputs for 0..n, tile sums for this call's tiles, seq_first / seq_endkey for this call's sequences); the grand
    // totals are the one place where a stale value could be mistaken for a result, so they are cleared.
    NTC_CUDA_TRY(cudaMemsetAsync(totals, 0, kStaticSmallBytes, st));
    const int n_scan = exclusive_scan_pairs(RecordLoad{d_records}, n_records, tiles, nullptr, totals, base_off, seq_idx, st,
                                            SeqFirstPost{seq_first});
    seq_layout_kernel<<<blocks_for(n_records + 1, 256), 256, 0, st>>>(base_off, seq_first, totals, n_records, seq_endkey, d_seq_offsets,
                                                                      cap_seqs, 0, first_seq_number, fasta, cap, d_counts, ctx->d_status);
    walk_kernel<<<blocks_for(n_records, kWalkThreads), kWalkThreads, 0, st>>>(ctx->dix, d_records, n_records, base_off, seq_idx, seq_endkey,
                                                                              totals, first_seq_number, fasta, d_out, cap, d_counts,
                                                                              ctx->d_status);
    NTC_CUDA_TRY(cudaGetLastError());
    ctx->launches += n_scan + 2;
    return NTC_OK;
}

#include "api_columns.inc"
#include "api_host.inc"
