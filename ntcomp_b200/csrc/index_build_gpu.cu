// index_build_gpu.cu -- ntc_index_build_gpu: SBWT + LCS construction on the GPU (replaces kbo::build as
// called at src/main.rs:111-134 / tests/fasta_data.rs:49-56).  The order of operations is build_pipeline.h, the element
// functors and the radix sort are build_kernels.cuh; this file is the CUDA executor (launches, scans, buffers) and the
// C entry point.  Same index, bit for bit, as the host builder (index_build.cpp) -- tests/test_index_build.py.
// No CPU fallback: without a device the call fails with NTC_ERR_CUDA.
#include <algorithm>
#include <chrono>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/ntcomp_b200.h"
#include "build_pipeline.h"
#include "ctx.h"
#include "host_index.h"
#include "scan.cuh"

using namespace ntc;

namespace {

// Device memory comes from one arena (a single cudaMalloc sized from the text length, a second chunk only if the dummy nodes
// outgrow it): the first version made ~25 cudaMalloc / cudaFree calls of 40-100 MB, more wall time than all kernels together.
struct GpuExec {
    cudaStream_t st = nullptr;
    struct Chunk { char* p; uint64_t cap, used; };
    std::vector<Chunk> chunks;
    uint64_t first_chunk = 256ull << 20;
    cudaError_t err = cudaSuccess;
    uint64_t launches = 0;
    Pair64* tile_sums = nullptr; uint64_t tile_cap = 0;
    uint64_t* d_totals = nullptr;
    uint32_t* block_hist = nullptr; uint64_t hist_cap = 0;
    uint32_t* digit_totals = nullptr;
    uint8_t* dig = nullptr; uint64_t dig_cap = 0;

    bool failed() const { return err != cudaSuccess; }
    void note(cudaError_t e) { if (err == cudaSuccess && e != cudaSuccess) err = e; }
    void* raw_alloc(uint64_t bytes) {
        if (failed()) return nullptr;
        bytes = (bytes + 255) & ~255ull;
        if (chunks.empty() || chunks.back().used + bytes > chunks.back().cap) {
            const uint64_t cap = std::max<uint64_t>(bytes, chunks.empty() ? first_chunk : 256ull << 20);
            void* p = nullptr;
            note(cudaMalloc(&p, cap));
            if (!p) return nullptr;
            chunks.push_back(Chunk{(char*)p, cap, 0});
        }
        Chunk& c = chunks.back();
        void* r = c.p + c.used;
        c.used += bytes;
        return r;
    }
    template <class T> T* alloc(uint64_t n) { return (T*)raw_alloc((n ? n : 1) * sizeof(T)); }
    void zero(void* p, uint64_t bytes) { if (!failed() && bytes) note(cudaMemsetAsync(p, 0, bytes, st)); }
    template <class F> void for_each(uint64_t n, F f) {
        if (failed() || n == 0) return;
        bld::for_each_kernel<F><<<(unsigned)((n + 255) / 256), 256, 0, st>>>(f, n);
        launches++;
    }
    template <class L, class P> uint64_t scan(L load, uint64_t n, P post) {
        if (failed() || n == 0) return 0;
        const uint64_t tiles = scan_num_tiles(n);
        if (tiles > tile_cap) { tile_sums = (Pair64*)raw_alloc(tiles * sizeof(Pair64)); tile_cap = tiles; }
        if (!d_totals) d_totals = (uint64_t*)raw_alloc(16);
        if (failed()) return 0;
        launches += exclusive_scan_pairs<L, P>(load, n, tile_sums, nullptr, d_totals, nullptr, nullptr, st, post);
        uint64_t h[2] = {0, 0};
        note(cudaMemcpyAsync(h, d_totals, 16, cudaMemcpyDeviceToHost, st));
        note(cudaStreamSynchronize(st));
        return failed() ? 0 : h[0];
    }
    template <class D> uint32_t* sort(uint32_t* a, uint32_t* b, uint64_t n, uint32_t t_lo, uint32_t t_hi, D digit) {
        if (failed() || n < 2) return a;
        const uint32_t n_blocks = (uint32_t)((n + bld::kSortTile - 1) / bld::kSortTile);
        if ((uint64_t)n_blocks * 256 > hist_cap) { block_hist = (uint32_t*)raw_alloc((uint64_t)n_blocks * 256 * 4); hist_cap = (uint64_t)n_blocks * 256; }
        if (n > dig_cap) { dig = (uint8_t*)raw_alloc(n); dig_cap = n; }
        if (!digit_totals) digit_totals = (uint32_t*)raw_alloc(256 * 4);
        if (failed()) return a;
        for (uint32_t t = t_lo; t < t_hi; t++) {
            bld::radix_hist_kernel<D><<<n_blocks, bld::kSortThreads, 0, st>>>(a, n, t, digit, dig, block_hist, n_blocks);
            bld::radix_scan_kernel<<<256, bld::kSortThreads, 0, st>>>(block_hist, n_blocks, digit_totals);
            bld::radix_scatter_kernel<<<n_blocks, bld::kSortThreads, 0, st>>>(a, dig, b, n, block_hist, n_blocks, digit_totals);
            launches += 3;
            uint32_t* s = a; a = b; b = s;
        }
        return a;
    }
    void download(void* host, const void* dev, uint64_t bytes) {
        if (failed() || !bytes) return;
        note(cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, st));
        note(cudaStreamSynchronize(st));
    }
    ~GpuExec() {
        for (Chunk& c : chunks) cudaFree(c.p);
        if (st) cudaStreamDestroy(st);
    }
};

}  // namespace

extern "C" int32_t ntc_index_build_gpu(int32_t device, const uint8_t* const* seqs, const uint64_t* lens, uint64_t n_seqs, uint32_t k,
                                       int32_t add_revcomp, ntc_index** out, double* stats) {
    if (!out || (n_seqs && (!seqs || !lens))) { set_error("null argument"); return NTC_ERR_INVALID_ARG; }
    *out = nullptr;
    if (k < kMinK || k > kMaxK) { set_error("k must be in [12,255]"); return NTC_ERR_UNSUPPORTED; }
    const auto t0 = std::chrono::steady_clock::now();
    auto ms_since = [&]() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(); };
    try {
        int n_dev = 0;
        cudaError_t e = cudaGetDeviceCount(&n_dev);
        if (e != cudaSuccess || n_dev == 0) {
            set_error(std::string("no CUDA device available (this library has no CPU fallback): ") + cudaGetErrorString(e));
            return NTC_ERR_CUDA;
        }
        if (device < 0 || device >= n_dev) { set_error("device ordinal out of range"); return NTC_ERR_INVALID_ARG; }
        NTC_CUDA_TRY(cudaSetDevice(device));
        // forward strand: every sequence followed by one separator, so no k-mer spans two sequences
        uint64_t fwd_len = 0;
        for (uint64_t s = 0; s < n_seqs; s++) fwd_len += lens[s] + 1;
        if (2 * fwd_len + bld::kPad + bld::kTail + 64 >= (1ull << 32)) { set_error("input too large: text positions are 32-bit"); return NTC_ERR_UNSUPPORTED; }
        std::vector<uint8_t> fwd(fwd_len ? fwd_len : 1);
        {
            uint64_t o = 0;
            for (uint64_t s = 0; s < n_seqs; s++) {
                if (lens[s]) memcpy(fwd.data() + o, seqs[s], lens[s]);
                o += lens[s];
                fwd[o++] = '\n';
            }
        }
        GpuExec x;
        // work arrays: ~45 bytes per text character (two strands + padding) when the dummy nodes are few, as in any genome
        x.first_chunk = ((uint64_t)(add_revcomp ? 2 : 1) * fwd_len + bld::kPad + bld::kTail + 64) * 48 + fwd_len + (32ull << 20);
        NTC_CUDA_TRY(cudaStreamCreateWithFlags(&x.st, cudaStreamNonBlocking));
        uint8_t* d_fwd = x.alloc<uint8_t>(fwd_len);
        if (x.failed()) { set_error(std::string("cudaMalloc: ") + cudaGetErrorString(x.err)); return NTC_ERR_CUDA; }
        NTC_CUDA_TRY(cudaMemcpyAsync(d_fwd, fwd.data(), fwd_len, cudaMemcpyHostToDevice, x.st));
        NTC_CUDA_TRY(cudaStreamSynchronize(x.st));
        const double t_up = ms_since();
        bld::BuildStats bs;
        const int32_t rc = bld::sbwt_build(x, d_fwd, fwd_len, k, add_revcomp != 0, out, &bs);
        if (rc == 2 || x.failed()) {
            cudaError_t le = x.failed() ? x.err : cudaGetLastError();
            set_error(std::string("GPU index construction failed: ") + cudaGetErrorString(le));
            if (*out) { ntc_index_free(*out); *out = nullptr; }
            return NTC_ERR_CUDA;
        }
        if (rc) return rc;
        cudaError_t le = cudaGetLastError();
        if (le != cudaSuccess) {
            set_error(std::string("GPU index construction failed: ") + cudaGetErrorString(le));
            ntc_index_free(*out); *out = nullptr;
            return NTC_ERR_CUDA;
        }
        if (stats) {
            stats[0] = t_up;                 // ms until the forward strand is in HBM (context start-up included on a cold process)
            stats[1] = ms_since() - t_up;    // ms of device work + download of rows / LCS + C
            stats[2] = (double)bs.occurrences;
            stats[3] = (double)bs.distinct;
            stats[4] = (double)bs.sources;
            stats[5] = (double)bs.dummies;
            stats[6] = (double)bs.sort_passes;   // radix passes run (over all occurrences, over the tied ones, over the dummy candidates)
            stats[7] = (double)x.launches;
        }
        return NTC_OK;
    } catch (const std::exception& ex) {
        set_error(std::string("index build failed: ") + ex.what());
        if (*out) { ntc_index_free(*out); *out = nullptr; }
        return NTC_ERR_INVALID_ARG;
    }
}
