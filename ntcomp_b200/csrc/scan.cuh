// scan.cuh -- device-wide exclusive prefix sums over pairs of u64 (reduce, scan of tile sums,
// downsweep).  Warp shuffles inside a tile; tile sums go through one small global array.
// Used for record offsets (encode) and base / sequence offsets (decode).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "device_common.cuh"

namespace ntc {

constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;
constexpr int kScanTile = kScanThreads * kScanItems;

struct Pair64 {
    uint64_t a, b;
};
NTC_HD Pair64 operator+(const Pair64& x, const Pair64& y) { return Pair64{x.a + y.a, x.b + y.b}; }

__device__ __forceinline__ Pair64 warp_inclusive(Pair64 v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint64_t ua = __shfl_up_sync(0xffffffffu, v.a, o);
        uint64_t ub = __shfl_up_sync(0xffffffffu, v.b, o);
        if (lane >= o) { v.a += ua; v.b += ub; }
    }
    return v;
}

// exclusive prefix of `v` over the block; returns block total in *total (valid in all threads)
__device__ __forceinline__ Pair64 block_exclusive(Pair64 v, Pair64* total) {
    __shared__ Pair64 warp_sums[kScanThreads / 32];
    __shared__ Pair64 block_total;
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    Pair64 inc = warp_inclusive(v, lane);
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        Pair64 w = lane < kScanThreads / 32 ? warp_sums[lane] : Pair64{0, 0};
        Pair64 winc = warp_inclusive(w, lane);
        if (lane < kScanThreads / 32) warp_sums[lane] = Pair64{winc.a - w.a, winc.b - w.b};
        if (lane == kScanThreads / 32 - 1) block_total = winc;
    }
    __syncthreads();
    Pair64 base = warp_sums[warp];
    *total = block_total;
    __syncthreads();
    return Pair64{base.a + inc.a - v.a, base.b + inc.b - v.b};
}

// pass 1: per-tile sums.  Load(i) -> Pair64 for i in [0, n)
template <typename Load>
__global__ void __launch_bounds__(kScanThreads) scan_reduce_kernel(Load load, uint64_t n, Pair64* tile_sums) {
    uint64_t base = (uint64_t)blockIdx.x * kScanTile + (uint64_t)threadIdx.x * kScanItems;
    Pair64 acc{0, 0};
#pragma unroll
    for (int i = 0; i < kScanItems; i++)
        if (base + i < n) acc = acc + load(base + i);
    Pair64 total;
    block_exclusive(acc, &total);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

// pass 2: one block scans the tile sums in place (exclusive) and adds the carried base offsets;
// the grand total (including carry) is written to totals[0..1] and becomes the next carry.
// Every thread owns one contiguous run of tiles (serial sum, ONE block-wide scan, serial write-back): the first
// version scanned 256 tiles per round with three barriers each -- 75 rounds and 82 us for the 19 k tiles of C3's
// record scan (ncu r1 launch list), as long as the reduce pass that reads all the records.
static __global__ void __launch_bounds__(kScanThreads) scan_tiles_kernel(Pair64* tile_sums, uint64_t n_tiles, const uint64_t* carry_in,
                                                                  uint64_t* totals) {
    const uint64_t per = (n_tiles + kScanThreads - 1) / kScanThreads;
    const uint64_t lo = tmin<uint64_t>(n_tiles, per * threadIdx.x), hi = tmin<uint64_t>(n_tiles, lo + per);
    const Pair64 carry = carry_in ? Pair64{carry_in[0], carry_in[1]} : Pair64{0, 0};   // may alias totals: read before the barriers
    Pair64 acc{0, 0};
    for (uint64_t i = lo; i < hi; i++) acc = acc + tile_sums[i];
    Pair64 total;
    Pair64 run = block_exclusive(acc, &total) + carry;
    total = total + carry;
    for (uint64_t i = lo; i < hi; i++) {
        const Pair64 v = tile_sums[i];
        tile_sums[i] = run;
        run = run + v;
    }
    if (threadIdx.x == 0) { totals[0] = total.a; totals[1] = total.b; }
}

// pass 3: exclusive prefixes.  out_a / out_b may be null; entry n (the total) is written by the
// thread that owns element n-1's successor slot.  post(i, prefix, value) runs for every element i < n (decode uses
// it to note the record that opens each sequence, which used to be a kernel of its own).
// A thread's 8 prefixes are 64 contiguous bytes per array: two 256-bit stores when the array is 32-byte aligned
// (first version: eight 8-byte stores per array, each warp instruction touching 32 different sectors --
// 369 us for the 0.94 GB of C3's record scan).
struct NoPost {
    NTC_HD void operator()(uint64_t, const Pair64&, const Pair64&) const {}
};
__device__ __forceinline__ void st32_plain(uint64_t* dst, uint64_t a, uint64_t b, uint64_t c, uint64_t d) {
    asm volatile("st.global.v4.b64 [%0], {%1,%2,%3,%4};" ::"l"(dst), "l"(a), "l"(b), "l"(c), "l"(d) : "memory");
}
template <typename Load, typename Post>
__global__ void __launch_bounds__(kScanThreads) scan_down_kernel(Load load, uint64_t n, const Pair64* tile_sums,
                                                                 uint64_t* out_a, uint64_t* out_b, Post post) {
    static_assert(kScanItems == 8, "vector store path writes 8 prefixes per thread");
    uint64_t base = (uint64_t)blockIdx.x * kScanTile + (uint64_t)threadIdx.x * kScanItems;
    Pair64 v[kScanItems];
    Pair64 acc{0, 0};
#pragma unroll
    for (int i = 0; i < kScanItems; i++) {
        v[i] = base + i < n ? load(base + i) : Pair64{0, 0};
        acc = acc + v[i];
    }
    Pair64 total;
    Pair64 ex = block_exclusive(acc, &total);
    Pair64 run = tile_sums[blockIdx.x] + ex;
    Pair64 pre[kScanItems];
#pragma unroll
    for (int i = 0; i < kScanItems; i++) {
        pre[i] = run;
        if (base + i < n) post(base + i, run, v[i]);
        run = run + v[i];
    }
    const bool full = base + kScanItems <= n + 1;  // slot n receives the grand total
    if (out_a) {
        if (full && ((uintptr_t)out_a & 31) == 0) {
            st32_plain(out_a + base, pre[0].a, pre[1].a, pre[2].a, pre[3].a);
            st32_plain(out_a + base + 4, pre[4].a, pre[5].a, pre[6].a, pre[7].a);
        } else {
#pragma unroll
            for (int i = 0; i < kScanItems; i++) if (base + i <= n) out_a[base + i] = pre[i].a;
        }
    }
    if (out_b) {
        if (full && ((uintptr_t)out_b & 31) == 0) {
            st32_plain(out_b + base, pre[0].b, pre[1].b, pre[2].b, pre[3].b);
            st32_plain(out_b + base + 4, pre[4].b, pre[5].b, pre[6].b, pre[7].b);
        } else {
#pragma unroll
            for (int i = 0; i < kScanItems; i++) if (base + i <= n) out_b[base + i] = pre[i].b;
        }
    }
}

inline uint64_t scan_num_tiles(uint64_t n) { return (n + 1 + kScanTile - 1) / kScanTile; }  // +1: slot n

// Exclusive scan of load(0..n) into out_a/out_b (n+1 entries each).  `tile_sums` must hold
// scan_num_tiles(n) Pair64; `totals` (device, 2 x u64) receives the grand totals; carry_in
// (device, 2 x u64, nullable) is added to every prefix.  Returns the number of kernels launched.
template <typename Load, typename Post = NoPost>
inline int exclusive_scan_pairs(Load load, uint64_t n, Pair64* tile_sums, const uint64_t* carry_in, uint64_t* totals,
                                uint64_t* out_a, uint64_t* out_b, cudaStream_t stream, Post post = Post()) {
    uint64_t tiles = scan_num_tiles(n);
    scan_reduce_kernel<Load><<<(unsigned)tiles, kScanThreads, 0, stream>>>(load, n, tile_sums);
    scan_tiles_kernel<<<1, kScanThreads, 0, stream>>>(tile_sums, tiles, carry_in, totals);
    scan_down_kernel<Load, Post><<<(unsigned)tiles, kScanThreads, 0, stream>>>(load, n, tile_sums, out_a, out_b, post);
    return 3;
}

}  // namespace ntc
