// device_common.cuh -- HBM layout of the re-laid-out index and small device helpers (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "layout.h"

// Per-thread kernel bodies are __host__ __device__ so that tests/emul can run the very same code
// on the CPU (one loop iteration per CUDA thread) before any GPU time is spent.  The product
// library only ever launches them as kernels.
#define NTC_HD __host__ __device__ __forceinline__
#if defined(__CUDA_ARCH__)
#define NTC_LDG(p) __ldg(p)
#define NTC_POPCLL(x) __popcll(x)
#define NTC_ATOMIC_OR(p, v) atomicOr((p), (v))
#else
#define NTC_LDG(p) (*(p))
#define NTC_POPCLL(x) __builtin_popcountll(x)
#define NTC_ATOMIC_OR(p, v) (*(p) |= (v))
#endif

namespace ntc {

template <typename T> NTC_HD T tmin(T a, T b) { return a < b ? a : b; }
template <typename T> NTC_HD T tmax(T a, T b) { return a > b ? a : b; }

// status bits accumulated by kernels in ctx->d_status[0]
enum : uint32_t {
    ST_EMPTY_READ = 1u << 0,
    ST_BAD_BASE = 1u << 1,
    ST_LEN24 = 1u << 2,
    ST_CAPACITY = 1u << 3,
    ST_BAD_RECORD = 1u << 4,
    ST_READ_TOO_LONG = 1u << 5,   // a read of 2^32 bases or more: positions inside a read are 32-bit
};

// Index as the kernels see it.  All arrays are L2-resident at the BASELINE configs
// (5 Mbp / k=91: rank 13.4 MB, LCS 10 MB, jump 80 MB).
struct DevIndex {
    const uint4* rank;     // rank blocks, 64 bytes each (layout.h): rank_nodes == 32: [(n/32+1) * 4] entries, c = entry index & 3:
                           //   {C[c] + ones of row c before the block, row bits, Q bits, 0}; rank_nodes == 64: two sectors per block
    const ntc_lcsx_t* lcsx;  // [n + 1] LCS side array (layout.h)
    const uint64_t* jump;  // [n] pred^16 | chars16 << 32
    const uint4* pfx;      // [4^8 * 2] prefix table (host_index.h), null when k < 10
    uint32_t c_start[4];   // interval start after the single base c
    uint32_t n;
    uint32_t k;
    uint32_t spec_lo, spec_span;  // ms_kernel speculates LCS loads while spec_lo <= d < spec_lo + spec_span
    uint64_t pol_keep, pol_stream;  // L2 cache-hint descriptors (createpolicy, made once by policy_kernel at context
                                    // creation): kernel parameters are uniform, so a hinted load needs no R2UR pair
    uint32_t rank_nodes;   // 32 or 64: layout of `rank` (layout.h); read by the host side only (kernels are instantiated per layout)
};

// scratch entry written by the matching-statistics pass, one per base:
//   bits 0-31 interval start, 32-39 d, 40-63 length of the run of "unique predecessor" positions
//   ending here (replaces left_extend_kmer's 4 x search per base, src/lib.rs:94-128)
NTC_HD uint64_t make_entry(uint32_t s, uint32_t d, uint32_t urun) {
    return (uint64_t)s | ((uint64_t)d << 32) | ((uint64_t)urun << 40);
}
NTC_HD uint32_t entry_s(uint64_t e) { return (uint32_t)e; }
NTC_HD uint32_t entry_d(uint64_t e) { return (uint32_t)(e >> 32) & 0xFFu; }
NTC_HD uint32_t entry_urun(uint64_t e) { return (uint32_t)(e >> 40); }

// block / offset of node v, ones of the entry's row in block-relative positions [0, off), Q bit at off
#if defined(__CUDA_ARCH__)
#define NTC_POPC(x) __popc(x)
#else
#define NTC_POPC(x) __builtin_popcount(x)
#endif
// 2-bit base at chunk-relative position p of a packed stream (32 bases per u64, LSB first)
NTC_HD uint32_t packed_base(const uint64_t* __restrict__ packed, uint64_t p) {
    return (uint32_t)(NTC_LDG(packed + (p >> 5)) >> ((p & 31) * 2)) & 3u;
}
// up to 32 bases starting at position p, base j of the result at bits 2j..2j+1 (bitnuc as_2bit order)
NTC_HD uint64_t packed_slice(const uint64_t* __restrict__ packed, uint64_t p, uint32_t len) {
    uint64_t w0 = NTC_LDG(packed + (p >> 5));
    uint32_t sh = (uint32_t)(p & 31) * 2;
    uint64_t v = w0 >> sh;
    if (sh && (p & 31) + len > 32) v |= NTC_LDG(packed + (p >> 5) + 1) << (64 - sh);
    return len >= 32 ? v : (v & ((1ull << (2 * len)) - 1ull));
}

// ---- L2 residency control -------------------------------------------------------------------
// The index (rank blocks, LCS words, jump table) is re-read at random and must stay in the 126 MB L2;
// scratch / record / FASTA streams pass through once.  ncu r1 v4 showed 14.5 GB of DRAM reads in
// ms_kernel -- index lines evicted by the 12 GB scratch stream -- so index loads carry an evict_last
// policy and streaming stores an evict_first one (createpolicy + .L2::cache_hint).  Measured again on v12 the
// hints are neutral (12.01 vs 12.00 ms) and most of those DRAM reads turned out to be the record-count pass
// re-reading scratch entries; a persisting-L2 carve-out on top is harmful (api.cu, ntc_ctx_create).
// The descriptors are made once (policy_kernel) and travel as kernel parameters in DevIndex.
struct CachePolicy {
    uint64_t keep, stream;
};
NTC_HD CachePolicy make_policies() {
    CachePolicy p{0, 0};
#if defined(__CUDA_ARCH__)
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p.keep));
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p.stream));
#endif
    return p;
}
NTC_HD uint4 ld16_keep(const uint4* ptr, uint64_t pol) {
#if defined(__CUDA_ARCH__)
    uint4 v;
    asm volatile("ld.global.nc.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(ptr), "l"(pol));
    return v;
#else
    (void)pol;
    return *ptr;
#endif
}
NTC_HD uint32_t ld4_keep(const uint32_t* ptr, uint64_t pol) {
#if defined(__CUDA_ARCH__)
    uint32_t v;
    asm volatile("ld.global.nc.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(v) : "l"(ptr), "l"(pol));
    return v;
#else
    (void)pol;
    return *ptr;
#endif
}
// predicated forms: the load is issued only when `on` (result 0 otherwise) and stays a single predicated instruction,
// so several of them issue back to back with no branch and no use of an earlier load's result in between
// (ncu r1 v10/v11: `vb = other ? load : va` made ptxas copy va first, i.e. wait for it, before issuing the second load)
NTC_HD uint32_t ld4_keep_if(const uint32_t* ptr, uint64_t pol, bool on) {
#if defined(__CUDA_ARCH__)
    uint32_t v;
    asm volatile(
        "{\n .reg .pred p;\n setp.ne.u32 p, %3, 0;\n mov.u32 %0, 0;\n"
        " @p ld.global.nc.L2::cache_hint.u32 %0, [%1], %2;\n}"
        : "=&r"(v) : "l"(ptr), "l"(pol), "r"((uint32_t)on));
    return v;
#else
    (void)pol;
    return on ? *ptr : 0u;
#endif
}
NTC_HD uint32_t ld2_keep_if(const unsigned short* ptr, uint64_t pol, bool on) {
#if defined(__CUDA_ARCH__)
    uint32_t v;
    asm volatile(
        "{\n .reg .pred p;\n .reg .u16 h;\n setp.ne.u32 p, %3, 0;\n mov.u16 h, 0;\n"
        " @p ld.global.nc.L2::cache_hint.u16 h, [%1], %2;\n cvt.u32.u16 %0, h;\n}"
        : "=&r"(v) : "l"(ptr), "l"(pol), "r"((uint32_t)on));
    return v;
#else
    (void)pol;
    return on ? (uint32_t)*ptr : 0u;
#endif
}
NTC_HD uint2 ld8x2_keep_if(const uint4* ptr, uint64_t pol, bool on) {   // first two words of a rank entry
#if defined(__CUDA_ARCH__)
    uint2 v;
    asm volatile(
        "{\n .reg .pred p;\n setp.ne.u32 p, %4, 0;\n mov.u32 %0, 0;\n mov.u32 %1, 0;\n"
        " @p ld.global.nc.L2::cache_hint.v2.u32 {%0,%1}, [%2], %3;\n}"
        : "=&r"(v.x), "=&r"(v.y) : "l"(ptr), "l"(pol), "r"((uint32_t)on));
    return v;
#else
    (void)pol;
    return on ? uint2{ptr->x, ptr->y} : uint2{0u, 0u};
#endif
}
NTC_HD uint64_t ld8_keep_if(const uint64_t* ptr, uint64_t pol, bool on) {
#if defined(__CUDA_ARCH__)
    uint64_t v;
    asm volatile(
        "{\n .reg .pred p;\n setp.ne.u32 p, %3, 0;\n mov.u64 %0, 0;\n"
        " @p ld.global.nc.L2::cache_hint.u64 %0, [%1], %2;\n}"
        : "=&l"(v) : "l"(ptr), "l"(pol), "r"((uint32_t)on));
    return v;
#else
    (void)pol;
    return on ? *ptr : 0ull;
#endif
}
// Rank blocks (layout.h).  RankL<NODES>: block / offset of node v; ld_ent = everything a step reads of (block, base c) -- count, row
// bits, Q bits -- in ONE load of one sector; ld_part_if = the part a rank needs (count + row bits), loaded only when `on`;
// rank = ones of the entry's row in block-relative positions [0, off) + the count; rank_sel = the same in `vb` when `on`, else in `va`.
template <int NODES> struct RankL;
template <> struct RankL<32> {
    typedef uint4 Ent;                                                // {count, row bits, Q bits, spare}
    typedef uint2 Part;
    static NTC_HD uint32_t block(uint32_t v) { return v >> 5; }
    static NTC_HD uint32_t offset(uint32_t v) { return v & 31u; }
    static NTC_HD Ent ld_ent(const uint4* rank, uint32_t blk, uint32_t c, uint64_t pol) { return ld16_keep(rank + (blk * 4u + c), pol); }   // 32-bit entry index: n / 8 < 2^29
    static NTC_HD Part ld_part_if(const uint4* rank, uint32_t blk, uint32_t c, uint64_t pol, bool on) { return ld8x2_keep_if(rank + (blk * 4u + c), pol, on); }
    static NTC_HD uint32_t rank(const Ent& e, uint32_t off) { return e.x + (uint32_t)NTC_POPC(e.y & ((1u << off) - 1u)); }   // off in [0, 32)
    static NTC_HD uint32_t bit(const Ent& e, uint32_t off) { return (e.y >> off) & 1u; }
    static NTC_HD uint32_t qbit(const Ent& e, uint32_t off) { return (e.z >> off) & 1u; }
    static NTC_HD uint32_t rank_sel(const Ent& va, const Part& vb, bool on, uint32_t off) {
        const uint32_t cnt = on ? vb.x : va.x, bits = on ? vb.y : va.y;
        return cnt + (uint32_t)NTC_POPC(bits & ((1u << off) - 1u));
    }
};
template <> struct RankL<64> {
    struct Ent { uint32_t cnt; uint64_t bits, q; };
    struct Part { uint32_t cnt; uint64_t bits; };
    static NTC_HD uint32_t block(uint32_t v) { return v >> 6; }
    static NTC_HD uint32_t offset(uint32_t v) { return v & 63u; }
    static NTC_HD Ent ld_ent(const uint4* rank, uint32_t blk, uint32_t c, uint64_t pol) {
        const uint4* p = rank + (blk * 4u + (c >> 1) * 2u);          // the sector of bases {A,C} or {G,T}
        uint64_t w0, w1, w2, w3;
#if defined(__CUDA_ARCH__)
        asm volatile("ld.global.nc.L2::cache_hint.v4.u64 {%0,%1,%2,%3}, [%4], %5;" : "=l"(w0), "=l"(w1), "=l"(w2), "=l"(w3) : "l"(p), "l"(pol));
#else
        (void)pol;
        const uint64_t* q = (const uint64_t*)p;
        w0 = q[0]; w1 = q[1]; w2 = q[2]; w3 = q[3];
#endif
        Ent e;
        e.cnt = (c & 1u) ? (uint32_t)(w0 >> 32) : (uint32_t)w0;
        e.bits = (c & 1u) ? w2 : w1;
        e.q = w3;
        return e;
    }
    static NTC_HD Part ld_part_if(const uint4* rank, uint32_t blk, uint32_t c, uint64_t pol, bool on) {
        const uint4* p = rank + (blk * 4u + (c >> 1) * 2u);
        Part r;
        r.cnt = ld4_keep_if((const uint32_t*)p + (c & 1u), pol, on);
        r.bits = ld8_keep_if((const uint64_t*)p + 1u + (c & 1u), pol, on);
        return r;
    }
    static NTC_HD uint32_t rank(const Ent& e, uint32_t off) { return e.cnt + (uint32_t)NTC_POPCLL(e.bits & ((1ull << off) - 1ull)); }   // off in [0, 64)
    static NTC_HD uint32_t bit(const Ent& e, uint32_t off) { return (uint32_t)(e.bits >> off) & 1u; }
    static NTC_HD uint32_t qbit(const Ent& e, uint32_t off) { return (uint32_t)(e.q >> off) & 1u; }
    static NTC_HD uint32_t rank_sel(const Ent& va, const Part& vb, bool on, uint32_t off) {
        const uint32_t cnt = on ? vb.cnt : va.cnt;
        const uint64_t bits = on ? vb.bits : va.bits;
        return cnt + (uint32_t)NTC_POPCLL(bits & ((1ull << off) - 1ull));
    }
};
NTC_HD uint64_t ld8_keep(const uint64_t* ptr, uint64_t pol) {
#if defined(__CUDA_ARCH__)
    uint64_t v;
    asm volatile("ld.global.nc.L2::cache_hint.u64 %0, [%1], %2;" : "=l"(v) : "l"(ptr), "l"(pol));
    return v;
#else
    (void)pol;
    return *ptr;
#endif
}
// one full 32-byte sector in a single 256-bit store (sm_100: STG.E.256); dst 32-byte aligned
NTC_HD void st32_stream(uint64_t* dst, uint64_t a, uint64_t b, uint64_t c, uint64_t d, uint64_t pol) {
#if defined(__CUDA_ARCH__)
    asm volatile("st.global.L2::cache_hint.v4.b64 [%0], {%1,%2,%3,%4}, %5;" ::"l"(dst), "l"(a), "l"(b), "l"(c), "l"(d), "l"(pol) : "memory");
#else
    (void)pol;
    dst[0] = a; dst[1] = b; dst[2] = c; dst[3] = d;
#endif
}
NTC_HD void st8_stream(uint64_t* dst, uint64_t v, uint64_t pol) {
#if defined(__CUDA_ARCH__)
    asm volatile("st.global.L2::cache_hint.b64 [%0], %1, %2;" ::"l"(dst), "l"(v), "l"(pol) : "memory");
#else
    (void)pol;
    *dst = v;
#endif
}

// hinted stores of the widths the decode writer uses (text leaves once: evict_first keeps it from displacing the jump table)
NTC_HD void st16_stream(void* dst, uint64_t lo, uint64_t hi, uint64_t pol) {
#if defined(__CUDA_ARCH__)
    asm volatile("st.global.L2::cache_hint.v2.b64 [%0], {%1,%2}, %3;" ::"l"(dst), "l"(lo), "l"(hi), "l"(pol) : "memory");
#else
    (void)pol;
    ((uint64_t*)dst)[0] = lo; ((uint64_t*)dst)[1] = hi;
#endif
}
NTC_HD void st4_stream(void* dst, uint32_t v, uint64_t pol) {
#if defined(__CUDA_ARCH__)
    asm volatile("st.global.L2::cache_hint.b32 [%0], %1, %2;" ::"l"(dst), "r"(v), "l"(pol) : "memory");
#else
    (void)pol;
    *(uint32_t*)dst = v;
#endif
}
NTC_HD void st2_stream(void* dst, uint16_t v, uint64_t pol) {
#if defined(__CUDA_ARCH__)
    asm volatile("st.global.L2::cache_hint.b16 [%0], %1, %2;" ::"l"(dst), "h"(v), "l"(pol) : "memory");
#else
    (void)pol;
    *(uint16_t*)dst = v;
#endif
}
NTC_HD void st1_stream(void* dst, uint8_t v, uint64_t pol) {
#if defined(__CUDA_ARCH__)
    asm volatile("st.global.L2::cache_hint.b8 [%0], %1, %2;" ::"l"(dst), "r"((uint32_t)v), "l"(pol) : "memory");
#else
    (void)pol;
    *(uint8_t*)dst = v;
#endif
}

// ASCII of a 2-bit code
NTC_HD uint8_t code_ascii(uint32_t c) { return (uint8_t)(0x54474341u >> (8 * (c & 3u))); }

}  // namespace ntc
