// synth.cpp -- host side of the synthetic input generator.
#include <algorithm>
#include <thread>
#include <vector>

#include "../../include/ntcomp_b200.h"
#include "synth_common.h"

extern "C" {

void ntc_synth_genome(uint64_t seed, uint64_t n, uint8_t* out) {
    for (uint64_t j = 0; j < n; j++) out[j] = ntc::genome_base(seed, j);
}

void ntc_synth_reads(const uint8_t* genome, uint64_t genome_len, uint64_t seed, uint64_t r0, uint64_t n_reads,
                     uint32_t read_len, uint32_t err_ppm, uint8_t* out, int32_t n_threads) {
    if (n_threads < 1) n_threads = (int)std::max(1u, std::thread::hardware_concurrency());
    auto work = [&](uint64_t lo, uint64_t hi) {
        for (uint64_t i = lo; i < hi; i++)
            for (uint32_t t = 0; t < read_len; t++)
                out[i * read_len + t] = ntc::read_base(genome, genome_len, seed, r0 + i, read_len, err_ppm, t);
    };
    std::vector<std::thread> pool;
    uint64_t per = (n_reads + n_threads - 1) / n_threads;
    for (int t = 0; t < n_threads; t++) {
        uint64_t lo = std::min(n_reads, (uint64_t)t * per), hi = std::min(n_reads, lo + per);
        if (lo < hi) pool.emplace_back(work, lo, hi);
    }
    for (auto& th : pool) th.join();
}

}  // extern "C"
