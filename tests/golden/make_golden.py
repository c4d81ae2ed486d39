"""Generates tests/golden/*.npz with the CPU oracle (literal restatement).

NOT reference output: the reference (Rust) cannot be built or run in this environment and holds no
golden vectors of its own, so these files pin the ORACLE's answers (regression fixtures), and let the
-m gpu suite compare the CUDA path against bytes that are stored in the repository.
Run from the repo root:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import helpers  # noqa: E402
from oracle import binding as ob  # noqa: E402


def make(name, seed, k, contig_sizes, n_reads, lengths, rates):
    rng = np.random.default_rng(seed)
    contigs = [helpers.random_seq(rng, n) for n in contig_sizes]
    o = ob.OracleIndex.build(contigs, k, True)
    reads = helpers.sample_reads(rng, contigs, n_reads, lengths, rates, k) + contigs[:1]
    bases, off = ob.concat_reads(reads)
    rec, roff = o.encode_reads(bases, off, fast=False, threads=8)
    db, doff = o.decode(rec, fast=False)
    assert (db == bases).all() and (doff == off).all()
    cbases, coff = ob.concat_reads(contigs)
    d, s, e = o.matching_statistics(reads[0])
    np.savez_compressed(os.path.join(HERE, name + ".npz"), k=np.int64(k), contig_bases=cbases, contig_offsets=coff, n_sets=np.int64(o.n),
                        C=o.C(), lcs_head=o.lcs()[:4096], row_a_head=o.rows()[0][:64], read_bases=bases, read_offsets=off, records=rec,
                        rec_offsets=roff, ms0_d=d, ms0_s=s, ms0_e=e, fasta=np.frombuffer(helpers.fasta_text(reads, 1), dtype=np.uint8))
    print(name, "n_sets", o.n, "reads", len(reads), "records", len(rec))


if __name__ == "__main__":
    make("golden_k31", 31001, 31, [6000, 900], 60, [1, 30, 31, 32, 150, 150, 400], [0, 0.01, 0.05])
    make("golden_k91", 91001, 91, [9000], 40, [90, 91, 92, 93, 150, 150, 600], [0, 0.01, 0.05])
    make("golden_k255", 255001, 255, [1500, 700, 1999], 12, [254, 255, 256, 257, 600], [0, 0.01])
