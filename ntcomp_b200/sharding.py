"""Read sharding across the GPUs of one box: reads are independent, the index is
replicated, there is NO collective on the data path.  Shard boundaries are multiples of the
reference's block size (65,536 reads, src/main.rs:152) so every rank emits whole reference blocks and
the concatenation rank 0, 1, ... is byte-for-byte the single-GPU stream; the tail block goes to the
last shard that has work."""

BLOCK_READS = 65536  # src/main.rs:152


def shard_bounds(n_reads, world_size, block=BLOCK_READS):
    """[(lo, hi)] per rank, contiguous, covering [0, n_reads), boundaries on block multiples."""
    n_blocks = (n_reads + block - 1) // block
    base, extra = divmod(n_blocks, world_size)
    out, lo_b = [], 0
    for r in range(world_size):
        nb = base + (1 if r < extra else 0)
        lo, hi = min(n_reads, lo_b * block), min(n_reads, (lo_b + nb) * block)
        out.append((lo, hi))
        lo_b += nb
    return out


def block_bounds(lo, hi, block=BLOCK_READS):
    """reference blocks inside a shard: [(lo, hi)] with hi - lo == block except the tail"""
    return [(b, min(hi, b + block)) for b in range(lo, hi, block)]


def concat_in_rank_order(parts):
    """outputs are concatenated in input order: rank 0's blocks first (no reordering, no exchange)"""
    import numpy as np
    parts = [p for p in parts if len(p)]
    return np.concatenate(parts) if parts else np.zeros(0, dtype=np.uint64)
