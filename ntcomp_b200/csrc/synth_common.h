// synth_common.h -- counter-based synthetic genome/read generator, shared by the host
// generator (synth.cpp) and the device generator (kernels in api.cu).  Deterministic in (seed, index):
// any rank or the oracle can regenerate any read.
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
#define NTC_HD __host__ __device__ __forceinline__
#else
#define NTC_HD inline
#endif

namespace ntc {

NTC_HD uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    uint64_t z = x;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

NTC_HD uint8_t genome_base(uint64_t seed, uint64_t j) {
    const char a[4] = {'A', 'C', 'G', 'T'};
    return (uint8_t)a[splitmix64(seed + j) & 3];
}

NTC_HD uint64_t read_stream_key(uint64_t seed, uint64_t r) { return seed ^ (r * 0xD1B54A32D192ED03ull); }

NTC_HD int ascii_to_code(uint8_t b) { return b == 'A' ? 0 : b == 'C' ? 1 : b == 'G' ? 2 : 3; }

// base t of read r (length read_len) over `genome`
NTC_HD uint8_t read_base(const uint8_t* genome, uint64_t genome_len, uint64_t seed, uint64_t r, uint32_t read_len,
                         uint32_t err_ppm, uint32_t t) {
    const char a[4] = {'A', 'C', 'G', 'T'};
    uint64_t key = read_stream_key(seed, r);
    uint64_t start = splitmix64(key) % (genome_len - read_len + 1);
    uint64_t strand = splitmix64(key + 1) & 1;
    int code = strand ? 3 - ascii_to_code(genome[start + (read_len - 1 - t)]) : ascii_to_code(genome[start + t]);
    uint64_t u = splitmix64(key + 2 + t);
    if ((u >> 32) % 1000000ull < err_ppm) code = (code + 1 + (int)(u % 3)) & 3;
    return (uint8_t)a[code];
}

}  // namespace ntc
