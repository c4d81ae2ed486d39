// layout.h -- constants shared by the host-side layout builders (host_index.cpp) and the kernels.
#pragma once
// Nodes per 64-byte rank block: entry = {count, 32 row bits, 32 Q bits, spare} -- every rank is a 32-bit mask + popc
// and the block index a shift (2 B of index per node).  The first layout (v1-v11) packed 48 nodes per block
// (1.33 B per node, 64-bit masks, a division by 48 per probe); measured against it on B200 with otherwise identical v12
// code: 12.00 vs 12.73 ms on C3 (profiles/r1/ab_v12_variants.txt).
#define NTC_RANK_NODES 32
