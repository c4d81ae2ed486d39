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
__global__ void __launch_bounds__(kScanThreads) scan_tiles_kernel(Pair64* tile_sums, uint64_t n_tiles, const uint64_t* carry_in,
                                                                  uint64_t* totals) {
    __shared__ Pair64 running;
    if (threadIdx.x == 0) running = carry_in ? Pair64{carry_in[0], carry_in[1]} : Pair64{0, 0};
    __syncthreads();
    for (uint64_t start = 0; start < n_tiles; start += kScanThreads) {
        uint64_t i = start + threadIdx.x;
        Pair64 v = i < n_tiles ? tile_sums[i] : Pair64{0, 0};
        Pair64 total;
        Pair64 ex = block_exclusive(v, &total);
        Pair64 r = running;
        if (i < n_tiles) tile_sums[i] = r + ex;
        __syncthreads();
        if (threadIdx.x == 0) running = r + total;
        __syncthreads();
    }
    if (threadIdx.x == 0) { totals[0] = running.a; totals[1] = running.b; }
}

// pass 3: exclusive prefixes.  out_a / out_b may be null; entry n (the total) is written by the
// thread that owns element n-1's successor slot.
template <typename Load>
__global__ void __launch_bounds__(kScanThreads) scan_down_kernel(Load load, uint64_t n, const Pair64* tile_sums,
                                                                 uint64_t* out_a, uint64_t* out_b) {
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
#pragma unroll
    for (int i = 0; i < kScanItems; i++) {
        if (base + i <= n) {  // slot n receives the grand total
            if (out_a) out_a[base + i] = run.a;
            if (out_b) out_b[base + i] = run.b;
        }
        run = run + v[i];
    }
}

inline uint64_t scan_num_tiles(uint64_t n) { return (n + 1 + kScanTile - 1) / kScanTile; }  // +1: slot n

// Exclusive scan of load(0..n) into out_a/out_b (n+1 entries each).  `tile_sums` must hold
// scan_num_tiles(n) Pair64; `totals` (device, 2 x u64) receives the grand totals; carry_in
// (device, 2 x u64, nullable) is added to every prefix.  Returns the number of kernels launched.
template <typename Load>
inline int exclusive_scan_pairs(Load load, uint64_t n, Pair64* tile_sums, const uint64_t* carry_in, uint64_t* totals,
                                uint64_t* out_a, uint64_t* out_b, cudaStream_t stream) {
    uint64_t tiles = scan_num_tiles(n);
    scan_reduce_kernel<Load><<<(unsigned)tiles, kScanThreads, 0, stream>>>(load, n, tile_sums);
    scan_tiles_kernel<<<1, kScanThreads, 0, stream>>>(tile_sums, tiles, carry_in, totals);
    scan_down_kernel<Load><<<(unsigned)tiles, kScanThreads, 0, stream>>>(load, n, tile_sums, out_a, out_b);
    return 3;
}

}  // namespace ntc
