"""Multi-GPU host logic on CPU: shard planning (65,536-read boundaries, src/main.rs:152) and a
world_size-2 gloo run where each rank handles its shard and rank 0 concatenates in rank order.
The per-shard compute stands in with the oracle here (test infrastructure): the product path needs a
GPU and is covered by the -m gpu suite; what is under test is partitioning + ordering + no exchange."""
import os
import socket
import subprocess
import sys

import numpy as np

from ntcomp_b200 import sharding
from conftest import ROOT


def test_shard_bounds_properties():
    for n_reads, world in [(0, 2), (1, 2), (65536, 2), (65537, 2), (10_000_000, 8), (100_000_000, 8), (3 * 65536 + 5, 4), (65536 * 8, 8)]:
        b = sharding.shard_bounds(n_reads, world)
        assert len(b) == world and b[0][0] == 0 and b[-1][1] == n_reads
        for (lo, hi), (lo2, _) in zip(b, b[1:]):
            assert hi == lo2 and lo <= hi
        for lo, hi in b[:-1]:
            assert lo % sharding.BLOCK_READS == 0 and (hi % sharding.BLOCK_READS == 0 or hi == n_reads)
        sizes = [(hi - lo + sharding.BLOCK_READS - 1) // sharding.BLOCK_READS for lo, hi in b]
        assert max(sizes) - min(sizes) <= 1
        # reference blocks of the whole job == union of the shards' blocks, in order
        whole = sharding.block_bounds(0, n_reads)
        parts = [blk for lo, hi in b for blk in sharding.block_bounds(lo, hi)]
        assert whole == parts


WORKER = r"""
import os, sys
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "tests"))
import numpy as np, torch, torch.distributed as dist
import helpers
from ntcomp_b200 import sharding
from oracle import binding as ob
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
rng = np.random.default_rng(1)
g = helpers.random_seq(rng, 8000)
o = ob.OracleIndex.build([g], 31, True)                     # index replicated: every rank builds the same one
reads = helpers.sample_reads(np.random.default_rng(2), [g], 50, [150], [0.01], 31)
block = 8                                                   # small block size so shards hold whole blocks
lo, hi = sharding.shard_bounds(len(reads), world, block)[rank]
bases, off = ob.concat_reads(reads[lo:hi])
rec, _ = o.encode_reads(bases, off, fast=True)             # this rank's shard, no exchange with the other rank
sizes = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
dist.all_gather(sizes, torch.tensor([len(rec)], dtype=torch.int64))
bufs = [torch.zeros(int(s.item()), dtype=torch.int64) for s in sizes]
mine = torch.from_numpy(rec.view(np.int64).copy())
if rank == 0:
    bufs[0] = mine
    for r in range(1, world):
        dist.recv(bufs[r], src=r)
    got = sharding.concat_in_rank_order([b.numpy().view(np.uint64) for b in bufs])
    allb, alloff = ob.concat_reads(reads)
    want, _ = o.encode_reads(allb, alloff, fast=True)
    assert (got == want).all(), "concatenated shards differ from the single-rank stream"
    db, doff = o.decode(got, fast=True)
    assert (db == allb).all() and (doff == alloff).all()
    print("GLOO_OK", len(got))
else:
    dist.send(mine, dst=0)
dist.barrier()
dist.destroy_process_group()
"""


def test_world_size_2_gloo_concatenation(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER % {"root": ROOT})
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    env = dict(os.environ, OMP_NUM_THREADS="1")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                          "--master-port", str(port), str(script)], capture_output=True, text=True, env=env, timeout=300)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "GLOO_OK" in out.stdout
