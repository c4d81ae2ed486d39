// layout.h -- constants shared by the host-side layout builders (host_index.cpp) and the kernels.
#pragma once
// Nodes per 64-byte rank block: entry = {count, 32 row bits, 32 Q bits, spare} -- every rank is a 32-bit mask + popc
// and the block index a shift (2 B of index per node).  The first layout (v1-v11) packed 48 nodes per block
// (1.33 B per node, 64-bit masks, a division by 48 per probe); measured against it on B200 with otherwise identical v12
// code: 12.00 vs 12.73 ms on C3 (profiles/r1/ab_v12_variants.txt).
#define NTC_RANK_NODES 32
// LCS side array.  0: one u32 per node = LCS[i] | min(255, i - PSV(i)) << 8 | LCS[i+1] << 16 | min(255, NSV(i+1) - (i+1)) << 24 -- a
// singleton interval finds both of its sides in ONE word.  1: one u16 per node = LCS[i] | min(15, i - PSV(i)) << 8 |
// min(15, NSV(i) - i) << 12 -- half the bytes (rank + LCS of a 12 Mbp index: 96 instead of 144 MB, inside the L2), two loads per
// contraction instead of one.  A/B: profiles/r2/ab_lcs16.txt.
#ifndef NTC_LCS16
#define NTC_LCS16 0
#endif
#if NTC_LCS16
typedef unsigned short ntc_lcsx_t;
#define NTC_LCS_DIST_SAT 15u
#else
typedef unsigned int ntc_lcsx_t;
#define NTC_LCS_DIST_SAT 255u
#endif
