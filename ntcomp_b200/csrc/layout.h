// layout.h -- constants shared by the host-side layout builders (host_index.cpp) and the kernels.
#pragma once
// Nodes per 64-byte rank block.  32: entry = {count, 32 row bits, 32 Q bits, 0} -- every rank is a 32-bit
// mask + popc and the block index a shift (2 B of index per node).  48: entry = {count, 48 row bits, 48 Q
// bits} -- 1.33 B per node but 64-bit masks and a division by 48 in the inner loop (the round-1 layout up
// to v11; kept for A/B measurements: make EXTRA=-DNTC_RANK_NODES=48).
#ifndef NTC_RANK_NODES
#define NTC_RANK_NODES 32
#endif
#if NTC_RANK_NODES != 32 && NTC_RANK_NODES != 48
#error "NTC_RANK_NODES must be 32 or 48"
#endif
