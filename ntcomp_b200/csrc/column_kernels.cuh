// column_kernels.cuh -- GPU column codec ("next" row f1 of SURVEY §8): the record stream of one block is
// split into ntcomp's four columns (src/encode.rs:168-229) and each column is bit-packed with the
// reference's integer codes (Rice / minimal binary, src/encode.rs:59-94; dsi-bitstream conventions:
// MSB-first bit stream in big-endian u64 words, unary = zeros then a one).  Code lengths are
// prefix-summed (scan.cuh) and every value ORs its code into place, so the stream is identical to the
// sequential writer's.  DEFLATE stays on the host (container.cpp).
#pragma once
#include "device_common.cuh"
#include "scan.cuh"

namespace ntc {

// scan input for the split: a = 1 for reference records, b = literal bases of the record
struct SplitLoad {
    const uint64_t* rec;
    NTC_HD Pair64 operator()(uint64_t i) const {
        uint32_t flag = (uint32_t)(NTC_LDG(rec + i) >> 56);
        if (flag & 2u) return Pair64{0, tmin<uint32_t>(flag >> 2, 32u)};
        return Pair64{1, 0};
    }
};

// reductions gathered while scattering: [0] max colex, [1] sum len, [2] sum flags
NTC_HD void split_thread(const uint64_t* __restrict__ rec, uint64_t n, const uint64_t* __restrict__ ref_idx,
                         const uint64_t* __restrict__ lit_off, uint64_t* __restrict__ colex, uint64_t* __restrict__ len,
                         uint64_t* __restrict__ flags, unsigned long long* __restrict__ lits, unsigned long long* __restrict__ red,
                         uint64_t i) {
    if (i >= n) return;
    const uint64_t w = NTC_LDG(rec + i);
    const uint32_t flag = (uint32_t)(w >> 56);
    flags[i] = flag;
#if defined(__CUDA_ARCH__)
    atomicAdd(red + 2, (unsigned long long)flag);
#else
    red[2] += flag;
#endif
    if ((flag & 2u) == 0) {
        const uint64_t r = ref_idx[i], c = w & 0xFFFFFFFFull, l = (w >> 32) & 0xFFFFFFull;
        colex[r] = c;
        len[r] = l;
#if defined(__CUDA_ARCH__)
        atomicMax(red + 0, (unsigned long long)c);
        atomicAdd(red + 1, (unsigned long long)l);
#else
        red[0] = red[0] > c ? red[0] : c;
        red[1] += l;
#endif
    } else {
        const uint32_t L = tmin<uint32_t>(flag >> 2, 32u);
        const uint64_t payload = w & 0x00FFFFFFFFFFFFFFull;           // 7 bytes; bases 28..31 read as A (encode.rs:215-216)
        uint64_t p = lit_off[i];
        for (uint32_t j = 0; j < L; j++, p++) {                       // repack 31 bases per word (encode.rs:224)
            unsigned long long bits = ((payload >> (2 * j)) & 3ull) << (2 * (p % 31));
            if (bits) {
#if defined(__CUDA_ARCH__)
                atomicOr(lits + p / 31, bits);
#else
                lits[p / 31] |= bits;
#endif
            }
        }
    }
}
__global__ void __launch_bounds__(256) split_kernel(const uint64_t* __restrict__ rec, uint64_t n, const uint64_t* __restrict__ ref_idx,
                                                    const uint64_t* __restrict__ lit_off, uint64_t* __restrict__ colex,
                                                    uint64_t* __restrict__ len, uint64_t* __restrict__ flags,
                                                    unsigned long long* __restrict__ lits, unsigned long long* __restrict__ red) {
    split_thread(rec, n, ref_idx, lit_off, colex, len, flags, lits, red, (uint64_t)blockIdx.x * 256 + threadIdx.x);
}
__global__ void __launch_bounds__(256) max_kernel(const unsigned long long* __restrict__ v, uint64_t n, unsigned long long* out) {
    uint64_t i = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (i < n) atomicMax(out, v[i]);
}

// ---- integer codes -------------------------------------------------------------------------------
struct CodeParams {
    int rice;          // 1 Rice, 0 minimal binary
    uint32_t b;        // Rice: log2 of the modulus
    uint32_t l;        // minimal binary: floor(log2(upper))
    uint64_t limit;    // minimal binary: 2^(l+1) - upper
};
NTC_HD uint64_t code_bits(const CodeParams& p, uint64_t v) {
    if (p.rice) return (v >> p.b) + 1 + p.b;
    return (v + 1 < p.limit) ? p.l : p.l + 1;
}
struct CodeLenLoad {
    const uint64_t* v;
    CodeParams p;
    NTC_HD Pair64 operator()(uint64_t i) const { return Pair64{code_bits(p, v[i]), 0}; }
};

// ORs value i's code into the MSB-first stream (words are native-endian here; bswap_kernel finishes)
NTC_HD void pack_thread(const CodeParams& p, const uint64_t* __restrict__ v, uint64_t n, const uint64_t* __restrict__ bit_off,
                        unsigned long long* __restrict__ words, uint64_t i) {
    if (i >= n) return;
    uint64_t code, pos = bit_off[i];
    uint32_t clen;
    if (p.rice) {
        pos += v[i] >> p.b;                                           // unary part: zeros are already there
        code = (1ull << p.b) | (p.b ? (v[i] & ((1ull << p.b) - 1)) : 0);
        clen = p.b + 1;
    } else {
        uint64_t x = v[i] + 1;
        if (x < p.limit) { code = x; clen = p.l; } else { code = x + p.limit; clen = p.l + 1; }
    }
    if (clen == 0) return;
    const uint64_t w = pos >> 6;
    const uint32_t sh = (uint32_t)(pos & 63);
    unsigned long long first, second = 0;
    if (sh + clen <= 64) first = code << (64 - sh - clen);
    else { first = code >> (sh + clen - 64); second = code << (128 - sh - clen); }
#if defined(__CUDA_ARCH__)
    atomicOr(words + w, first);
    if (second) atomicOr(words + w + 1, second);
#else
    words[w] |= first;
    if (second) words[w + 1] |= second;
#endif
}
__global__ void __launch_bounds__(256) pack_kernel(CodeParams p, const uint64_t* __restrict__ v, uint64_t n,
                                                   const uint64_t* __restrict__ bit_off, unsigned long long* __restrict__ words) {
    pack_thread(p, v, n, bit_off, words, (uint64_t)blockIdx.x * 256 + threadIdx.x);
}
__global__ void __launch_bounds__(256) bswap_kernel(unsigned long long* __restrict__ words, uint64_t n) {
    uint64_t i = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    unsigned long long x = words[i];
    uint32_t lo = (uint32_t)x, hi = (uint32_t)(x >> 32);
    words[i] = ((unsigned long long)__byte_perm(lo, 0, 0x0123) << 32) | __byte_perm(hi, 0, 0x0123);
}

}  // namespace ntc
