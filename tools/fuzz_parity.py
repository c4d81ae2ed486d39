"""Randomised differential run of the whole GPU path against the CPU oracle (test infrastructure, like tests/):
random k, contigs, strands; index built on the GPU (compared with the host builder and the oracle's own construction); context with a random rank-block
layout and tile length; random reads (1 base .. contig-sized, 0-30 % substitutions); records against the oracle's, round trip, FASTA text.

    python tools/fuzz_parity.py <seconds> [seed]        -> one JSON line with the number of cases and mismatches
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers  # noqa: E402
import ntcomp_b200 as nt  # noqa: E402
from oracle import binding as ob  # noqa: E402


def one_case(rng, stats):
    k = int(rng.choice([12, 13, 16, 21, 31, 32, 33, 48, 64, 91, 127, 200, 255]))
    n_contigs = int(rng.integers(1, 6))
    alpha = helpers.ALPHA if rng.random() < 0.85 else helpers.ALPHA[:3]          # sometimes a skewed alphabet
    contigs = [alpha[rng.integers(0, len(alpha), size=int(rng.choice([k - 1, k, k + 1, 300, 2000, 20000, 60000])))] for _ in range(n_contigs)]
    if rng.random() < 0.3:
        contigs.append(contigs[0][5:].copy())                                      # a sequence whose start has a predecessor elsewhere
    if rng.random() < 0.2:
        c = contigs[-1].copy()
        if len(c) > 10:
            c[len(c) // 2] = ord("N")
            contigs.append(c)
    revcomp = bool(rng.random() < 0.8)
    gpu = nt.Index.build_gpu(contigs, k, revcomp)
    host = nt.Index.build(contigs, k, revcomp)
    same = gpu.n_sets == host.n_sets and (gpu.lcs() == host.lcs()).all() and all((a == b).all() for a, b in zip(gpu.rows(), host.rows()))
    stats["index"] += 1
    if not same:
        stats["mismatch"].append(("index", k, [len(c) for c in contigs]))
        return
    oracle = ob.OracleIndex.from_arrays(k, gpu.n_sets, gpu.rows(), gpu.lcs())
    os.environ["NTC_RANK_NODES"] = str(rng.choice([32, 64]))
    os.environ["NTC_MAX_CHUNK_BASES"] = str(int(rng.choice([4096, 100000, 50331648])))
    os.environ["NTC_MAX_CHUNK_RECORDS"] = str(int(rng.choice([64, 5000, 4194304])))
    ctx = nt.Context(gpu, 0)
    ctx.set_host_threads(int(rng.choice([-1, -1, 0, 3])))                    # auto: pageable numpy buffers are staged by host threads
    src = [c for c in contigs if len(c) > 0 and not (c == ord("N")).any()]
    if not src:
        return
    lengths = [1, 2, 5, k - 1, k, k + 1, 2 * k, 150, 150, 400, 1000, 5000, 30000]
    reads = helpers.sample_reads(rng, src, int(rng.integers(1, 200)), lengths, [0, 0, 0.002, 0.01, 0.05, 0.3], k)
    if rng.random() < 0.3:
        reads.append(np.ascontiguousarray(src[0]))                                 # a whole contig as one read
    bases, off = nt.concat_reads(reads)
    ctx.set_tile_len(int(rng.choice([0, 0, 256, 260, 512, 1024])))
    try:
        rec, roff = ctx.encode_batch(bases, off)
    except nt.NtcError as e:
        # a base that no node of the index carries (skewed alphabet, or the other strand of an index built without it): the
        # reference underflows there (Q3), the library reports NTC_ERR_BAD_BASE; the oracle must see d == 0 somewhere too
        if e.code != 5:
            raise
        d_zero = any((oracle.matching_statistics(r)[0] == 0).any() for r in reads)
        stats["rejected"] += 1
        if not d_zero:
            stats["mismatch"].append(("bad_base without d == 0", k, len(reads)))
        return
    orec, oroff = oracle.encode_reads(bases, off, fast=True, threads=8)
    stats["encode"] += 1
    stats["reads"] += len(reads)
    stats["bases"] += int(len(bases))
    if not ((roff == oroff).all() and len(rec) == len(orec) and (rec == orec).all()):
        stats["mismatch"].append(("records", k, len(reads)))
        return
    db, doff = ctx.decode_batch(rec)
    first = int(rng.choice([1, 9, 99995]))
    text, n = ctx.decode_fasta(rec, first)
    if not ((doff == off).all() and (db == bases).all() and n == len(reads) and text == helpers.fasta_text(reads, first)):
        stats["mismatch"].append(("decode", k, len(reads)))


def main():
    seconds = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    rng = np.random.default_rng(seed)
    stats = {"index": 0, "encode": 0, "rejected": 0, "reads": 0, "bases": 0, "mismatch": []}
    t0 = time.time()
    while time.time() - t0 < seconds:
        one_case(rng, stats)
    stats["seconds"] = round(time.time() - t0, 1)
    stats["seed"] = seed
    stats["mismatches"] = len(stats["mismatch"])
    print(json.dumps(stats))
    sys.exit(1 if stats["mismatch"] else 0)


if __name__ == "__main__":
    main()
