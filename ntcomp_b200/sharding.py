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


def bind_to_gpu_numa(device):
    """Pin this process to the CPUs closest to `device` (NVML's ideal CPU affinity) before it allocates pinned
    host buffers, so that every rank's staging memory and copy-issuing thread sit on the NUMA node its GPU hangs
    off.  With 8 ranks on one box this is what keeps the host-buffer (end-to-end) path off the inter-socket
    link.  Returns the CPU list it bound to, or None when NVML / sched_setaffinity is unavailable."""
    import os
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(device)
        n_cpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (n_cpu + 63) // 64)
        cpus = [64 * w + b for w, mask in enumerate(words) for b in range(64) if (mask >> b) & 1]
        allowed = os.sched_getaffinity(0)
        cpus = [c for c in cpus if c in allowed]
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return cpus
    except Exception:
        return None
