#!/usr/bin/env python
"""tests/golden/*.npz -> tests/reference_dump/in/<name>.{k,contigs.fa,reads.fa}: the inputs of rust/ref_dump (the
unmodified reference run over this repository's golden fixtures).  Also writes the 7-contig fixture of the reference's
own test shape (tests/fasta_data.rs: k = 255, lengths < 2000) as `c1_shape`.
    python rust/export_fixtures.py [out_dir]
"""
import glob
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def write_fasta(path, seqs, prefix):
    with open(path, "wb") as f:
        for i, s in enumerate(seqs):
            f.write(b">%s.%d\n" % (prefix.encode(), i + 1) + bytes(s) + b"\n")


def split(bases, off):
    return [bases[int(off[i]): int(off[i + 1])] for i in range(len(off) - 1)]


def main():
    out = os.path.join(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "tests", "reference_dump"), "in")
    os.makedirs(out, exist_ok=True)
    for path in sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "*.npz"))):
        z = np.load(path)
        name = os.path.basename(path)[:-4]
        open(os.path.join(out, name + ".k"), "w").write("%d\n" % int(z["k"]))
        write_fasta(os.path.join(out, name + ".contigs.fa"), split(z["contig_bases"], z["contig_offsets"]), "contig")
        write_fasta(os.path.join(out, name + ".reads.fa"), split(z["read_bases"], z["read_offsets"]), "read")
    # the reference's own test shape: every contig encoded against the index that holds it (the C1 anomaly)
    rng = np.random.default_rng(20250731)
    lens = [1701, 1346, 266, 1647, 182, 933, 1285]          # two contigs whose remainder is a literal (266, 1285), one shorter than k
    contigs = [np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, size=n)] for n in lens]
    open(os.path.join(out, "c1_shape.k"), "w").write("255\n")
    write_fasta(os.path.join(out, "c1_shape.contigs.fa"), contigs, "contig")
    write_fasta(os.path.join(out, "c1_shape.reads.fa"), contigs, "contig")
    print("wrote fixtures to", out)


if __name__ == "__main__":
    main()
