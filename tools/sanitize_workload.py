"""Small workload for `compute-sanitizer --tool memcheck|racecheck`: every kernel family of the
library on inputs small enough for the sanitizer's 10-100x slow-down, no torch in the process.  Checks results against the
original reads, so a sanitizer run that changes behaviour shows up too.

    compute-sanitizer --tool racecheck python tools/sanitize_workload.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ntcomp_b200 as nt  # noqa: E402
from ntcomp_b200 import synth  # noqa: E402


def main():
    g = synth.genome(40_000)
    index = nt.Index.build_gpu([g], 31)                       # bld::radix_* construction kernels, LCS
    ctx = nt.Context(index, 0)                                # lay::* layout kernels
    bases, off = synth.reads(g, 4242, 0, 700, 150, 10_000)
    long_bases, long_off = synth.reads(g, 99, 0, 3, 9000, 50_000)   # tiled long reads + the segment-parallel path
    for b, o in ((bases, off), (long_bases, long_off)):
        rec, roff = ctx.encode_batch(b, o)                    # pack, ms_kernel (TMA-staged reads), count, scans, emit
        db, doff = ctx.decode_batch(rec)                      # decode scans, seq_layout, walk_kernel (shared-memory sort)
        assert (db == b).all() and (doff == o).all()
        text, n_seqs = ctx.decode_fasta(rec, 1)
        assert n_seqs == len(o) - 1
    blocks, words, n_rec = ctx.encode_columns(bases, off, 256)   # fused column codec, both directions
    text2, n_seqs2, n_rec2 = ctx.decode_columns_fasta(blocks, words, 1, True)
    want, _ = ctx.decode_fasta(ctx.encode_batch(bases, off)[0], 1)
    assert bytes(text2) == bytes(want) and n_seqs2 == 700
    print("SANITIZE_WORKLOAD_OK launches=%d" % ctx.launch_count())


if __name__ == "__main__":
    main()
