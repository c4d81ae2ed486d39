// ctx.h -- per-GPU context: re-laid-out index in HBM, grow-only workspaces, status words.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <string>
#include <vector>

#include "device_common.cuh"
#include "host_pack.h"
#include "scan.cuh"

namespace ntc {
void set_error(const std::string& msg);

// kernel classes for launch accounting / profiling
enum KernelClass { KC_PACK = 0, KC_MS, KC_SCAN, KC_EMIT, KC_DSCAN, KC_SEQ, KC_WALK, KC_MISC, KC_COLENC, KC_COLDEC, KC_COUNT };

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
};

// One pipeline lane of the host-buffer API: its own stream, encode workspaces and staging buffers, so
// chunk i+1's H2D copy and kernels overlap chunk i's D2H copy (two lanes, ping-pong).
struct Lane {
    cudaStream_t stream = nullptr;
    DevBuf packed, scratch, counts, tiles, rec_off;   // encode workspaces
    DevBuf tile_off, desc;                            // tiles per read (scanned) and work-item descriptors of ms_kernel
    DevBuf seg_off, seg_desc, seg_bits, seg_info;     // long reads: segmentation segments (encode_kernels.cuh, seg_*_thread)
    uint32_t seg_len = 0;                             // != 0: the last ms_pipeline on this lane took the segment path with this S
    DevBuf st_in, st_off, st_out, st_out2;            // staging for host buffers
    // decode workspaces: base offsets / sequence ids per record, first record and layout word per sequence,
    // scan tiles, and dec_small = [0..1] scan totals (bases, sequences), [2..3] counts (sequences, bytes)
    DevBuf base_off, seq_idx, seq_first, seq_endkey, dec_tiles, dec_small;
    cudaEvent_t ev_scan = nullptr, ev_done = nullptr;
    uint64_t* h_land = nullptr;                       // pinned landing zone: [0..1] totals, [2] status
};
}  // namespace ntc

struct ntc_ctx {
    int device = 0;
    ntc::DevIndex dix{};
    void* d_rank = nullptr;
    void* d_lcs = nullptr;
    void* d_jump = nullptr;
    void* d_pfx = nullptr;   // prefix table (null when k < 10)
    uint64_t index_bytes[3] = {0, 0, 0};
    uint32_t k = 0;
    uint64_t n = 0;
    uint32_t uses = 3;       // NTC_USE_* mask: which derived arrays are resident

    uint32_t* d_status = nullptr;   // [4]
    uint64_t* d_totals = nullptr;   // [8]: ring of 3 encode scan totals (Pair64 each) + decode totals at [6]
    uint64_t launches = 0;
    uint64_t max_chunk_bases = 48ull << 20;   // host-buffer API chunk (overlap granularity; 16 M .. 192 M swept on B200, profiles/r1/sweep_chunks_v12.txt); NTC_MAX_CHUNK_BASES overrides
    uint32_t tile_len = 0;                    // matching-statistics tile length; 0 = chosen per batch (ntc_ctx_set_tile_len)

    ntc::Lane lanes[2];             // lanes[0] also serves the device-resident API (on the caller's stream)
    uint64_t max_chunk_records = 4ull << 20;  // host-buffer decode: records per chunk (NTC_MAX_CHUNK_RECORDS overrides)
    cudaStream_t stream = nullptr;  // == lanes[0].stream
    // Host-buffer API: 0 = both PCIe directions at once (results travel back while the next chunk's input goes up),
    // 1 = one direction at a time (all input first, results afterwards).  On a host whose copy engines share a ceiling
    // -- 8 GPUs copying on the round-2 box: 29.8 GB/s up alone, 14.9 down alone, 19.3 for BOTH together -- mixing the
    // directions looked costlier than the overlap gains (profiles/r2/e2e_probe_n8.json); measured, serial gains nothing there
    // (ranks out of phase: profiles/r2/ab_copy_directions.txt), so the default is 0.  NTC_COPY_DIRECTIONS=serial|duplex.
    int copy_serial = 0;
    ntc::DevBuf hold, hold2;        // copy_serial: records / offsets of a whole call parked in HBM
    // Host threads of the host-buffer entry points (ntc_ctx_set_host_threads; host_pack.cpp).
    //   PAGEABLE caller buffers (a Rust Vec, a numpy array) go through page-locked staging owned by the context, filled / emptied by the pool:
    //   a pageable cudaMemcpy runs at 7.5 GB/s up and 14 GB/s down on the B200 boxes (encode 200 ms, decode 119 ms per 1.5 G bases against
    //   32 + 31 ms from page-locked buffers; profiles/r2/pageable_probe.json).  Reads are packed to 2 bits per base on the way (the pool reads
    //   them once either way), records / text / offsets are copied.  Page-locked caller buffers are copied by the DMA engines directly, as before.
    //   host_threads: < 0 (default) = the above with auto_threads; 0 = never (the driver's pageable path); n > 0 = n threads, and the reads are
    //   packed on the host for page-locked buffers too (measured no faster there: profiles/r2/ab_host_pack.txt).
    int host_threads = -1;
    int auto_threads = 0;
    ntc::HostPool* pool = nullptr;  // created on first use
    struct Pinned { void* p = nullptr; size_t cap = 0; };
    Pinned h_in[2], h_out[2];       // per lane: staged input (packed reads + offsets, or records) and output (records / text + offsets)
    cudaEvent_t ev_out[2] = {nullptr, nullptr};   // the lane's D2H into h_out is done
    void* colwork = nullptr;        // workspaces + pinned landing buffers of the GPU column codec, both directions (api_columns.inc)

    // optional per-kernel-class timing with CUDA events on the launching stream (bench.py roofline)
    bool profiling = false;
    struct Span { int cls; cudaEvent_t e0, e1; };
    std::vector<Span> spans;
    std::vector<cudaEvent_t> event_pool;
    double class_ms[ntc::KC_COUNT] = {0};
    uint64_t class_launches[ntc::KC_COUNT] = {0};
};

namespace ntc {

#define NTC_CUDA_TRY(expr)                                                                   \
    do {                                                                                     \
        cudaError_t _e = (expr);                                                             \
        if (_e != cudaSuccess) {                                                             \
            ntc::set_error(std::string(#expr) + ": " + cudaGetErrorString(_e));              \
            return NTC_ERR_CUDA;                                                             \
        }                                                                                    \
    } while (0)

// grow-only device buffer
inline cudaError_t ensure(DevBuf* b, size_t bytes) {
    if (bytes <= b->cap) return cudaSuccess;
    if (b->p) { cudaFree(b->p); b->p = nullptr; b->cap = 0; }
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMalloc(&b->p, want);
    if (e != cudaSuccess) { want = bytes; e = cudaMalloc(&b->p, want); }
    if (e == cudaSuccess) b->cap = want;
    return e;
}
inline void release(DevBuf* b) {
    if (b->p) cudaFree(b->p);
    b->p = nullptr;
    b->cap = 0;
}

}  // namespace ntc
