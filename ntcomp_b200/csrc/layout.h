// layout.h -- constants shared by the host-side layout builders (host_index.cpp) and the kernels.
#pragma once
// Rank blocks, two layouts, chosen per index by ntc_ctx_create (DevIndex::rank_nodes; NTC_RANK_NODES=32|64 forces one):
//   32 nodes per 64-byte block = 4 x 16-byte entries {count, 32 row bits, 32 Q bits, spare} -- every rank is a 32-bit mask + popc on one
//      128-bit load (2 B of index per node).  The first layout (v1-v11) packed 48 nodes per block (1.33 B per node, 64-bit masks, a
//      division by 48 per probe); measured against it on B200 with otherwise identical v12 code: 12.00 vs 12.73 ms on C3
//      (profiles/r1/ab_v12_variants.txt).
//   64 nodes per 64-byte block = two 32-byte sectors, sector 0 = { u32 count A, u32 count C, u64 row bits A, u64 row bits C, u64 Q bits },
//      sector 1 = the same for G, T (Q repeated): a probe is still one sector, now one 256-bit load, at 1 B of index per node.  The wider
//      load and the 64-bit masks cost 19 % on a 5 Mbp index (C3 11.5 -> 13.7 ms) and win once the 2 B / node array no longer stays in L2:
//      -10 % at 12 Mbp (C4 20.4 -> 18.4 ms, rank array 48 -> 24 MB), -28 % at 24 Mbp (32.3 -> 23.4 ms).  profiles/r2/ab_rank64.txt.
#define NTC_RANK64_MIN_NODES 20000000ull   // indexes with at least this many nodes take the 64-node layout
// LCS side array: one u32 per node = LCS[i] | min(255, i - PSV(i)) << 8 | LCS[i+1] << 16 | min(255, NSV(i+1) - (i+1)) << 24 -- a
// singleton interval finds both of its sides in ONE word.  (Rejected, profiles/r2/ab_lcs16.txt: a u16 per node with 4-bit
// distances -- half the bytes, but two loads per contraction; slower on C3 and on the 12 Mbp index.)
typedef unsigned int ntc_lcsx_t;
#define NTC_LCS_DIST_SAT 255u
