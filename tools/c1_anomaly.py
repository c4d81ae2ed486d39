#!/usr/bin/env python
"""The C1 anomaly, in numbers.

The reference's only test (tests/fasta_data.rs) indexes 7 random contigs (length = u64 % 2000, k = 255, both strands),
encodes every contig AGAINST ITS OWN INDEX in blocks of 3, decodes and asserts that all 7 contigs come back
(tests/fasta_data.rs:96-108).  `write_block_to` fails on a block without a literal record (minimal_binary_encode of an
empty column, src/encode.rs:79) and the test ignores that error (`let _ =`, tests/fasta_data.rs:75,81), so the test can
only pass when EVERY block -- including the last, which holds contig 7 alone -- has a literal record (<= 11 bases).

Under the oracle's reading of src/lib.rs:163-230 a contig of n > k + 1 bases that is in the index yields exactly two
records: a reference of h*k + 1 bases covering its end and a remainder of r = n - h*k - 1 bases (1 <= r <= k) covering
its start, h minimal with n - h*k <= k + 1; the remainder is a literal only when r <= 11, i.e. for 11 lengths in 255.
n = k gives [k], n = k + 1 gives [k, 1].  A contig SHORTER than k adds no k-mer to the index and is covered by chance
matches against the other contigs: many literal records, hardly ever a reference one.  This script

  1. checks that closed form against the oracle (literal restatement) on random self-indexed contigs;
  2. estimates by Monte Carlo how often a fixture of that shape would pass the test: 10^5 fixtures through the closed
     form, and a few thousand through the oracle end to end (index build, encode, blocks, decode) as a cross-check;
  3. rebuilds the fixture of seed 20250731 under candidate seedings of the `random` crate's default source
     (Xorshift128+; crate source absent, Cargo.lock pins random 0.14) and runs the oracle's encode -> blocks -> decode on
     each: which assertion the upstream test would trip on, if the oracle were the reference.  One passing candidate
     would be strong evidence for that seeding AND the oracle (chance: ~1 % each).

Usage: python tools/c1_anomaly.py [--fixtures 100000] [--check 60] [--full 5000] > profiles/r2/c1_anomaly.json
Needs only the oracle (gcc); no GPU.
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import binding as ob  # noqa: E402

K = 255
MASK = (1 << 64) - 1


# ---- the closed form -------------------------------------------------------------------------------------------------
def predicted_record_lengths(n, k=K):
    """record lengths in DICTIONARY order (the first record covers the END of the contig) of a contig of n >= k bases
    encoded against an index that holds it, src/lib.rs:177-216.  None for n < k: such a contig adds no k-mer to the index
    and is covered by chance matches against the other contigs (many records of ~log4(index) bases, nearly all literal)."""
    if n < k:
        return None
    if n == k:
        return [k]                      # one record: d = k at the last base, i > k + 1 is false, i -= d - 1 leaves nothing
    if n == k + 1:
        return [k, 1]                   # i > k + 1 still false: the k-mer, then the first base by itself
    h = -(-(n - (k + 1)) // k)          # left_extend_kmer reaches n - 1 bases; the hop loop lands on i = n - h*k <= k + 1
    return [h * k + 1, n - h * k - 1]   # remainder r in [1, k]: a literal only when r <= 11


def fixture_passes(lengths, block=3, k=K):
    """would tests/fasta_data.rs pass on contigs of these lengths, under the closed form?  Every block needs >= 1 literal
    and >= 1 reference record (src/encode.rs:79 via split_encoded_dictionary, error dropped at tests/fasta_data.rs:75,81);
    an empty contig panics in encode_sequence(..).unwrap().  Contigs shorter than k are taken to give literals only (a
    chance match of >= 12 bases against ~14 kbp of index is rare: the full-oracle run below counts those too)."""
    if any(n == 0 for n in lengths):
        return False
    for b in range(0, len(lengths), block):
        lit = ref = False
        for n in lengths[b: b + block]:
            recs = predicted_record_lengths(n, k)
            if recs is None:
                lit = True
            else:
                lit |= any(x <= 11 for x in recs)
                ref |= any(x > 11 for x in recs)
        if not (lit and ref):
            return False
    return True


# ---- Xorshift128+ (Vigna 2014), the generator behind random::default ------------------------------------------------
class Xorshift128Plus:
    def __init__(self, s0, s1, shifts=(23, 17, 26)):
        self.s = [s0 & MASK, s1 & MASK]
        self.a, self.b, self.c = shifts

    def read_u64(self):
        x, y = self.s
        self.s[0] = y
        x ^= (x << self.a) & MASK
        x ^= x >> self.b
        x ^= y ^ (y >> self.c)
        self.s[1] = x
        return (x + y) & MASK


def splitmix(x):
    x = (x + 0x9E3779B97F4A7C15) & MASK
    z = x
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & MASK
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & MASK
    return z ^ (z >> 31), x


def candidate_seedings(seed):
    sm1, st = splitmix(seed)
    sm2, _ = splitmix(st)
    out = {
        "[seed, seed]": (seed, seed),
        "[seed, 0]": (seed, 0),
        "[0, seed]": (0, seed),
        "[seed, !seed]": (seed, ~seed & MASK),
        "[seed, seed+1]": (seed, seed + 1),
        "[seed+1, seed]": (seed + 1, seed),
        "[42^seed, 69^seed]": (42 ^ seed, 69 ^ seed),
        "[42, seed]": (42, seed),
        "[seed, 69]": (seed, 69),
        "[splitmix64(seed), splitmix64 next]": (sm1, sm2),
        "[seed, splitmix64(seed)]": (seed, sm1),
        "[seed*0x9E3779B97F4A7C15, seed]": ((seed * 0x9E3779B97F4A7C15) & MASK, seed),
    }
    return out


def fixture(rng, n_contigs=7, max_len=2000):
    contigs = []
    for _ in range(n_contigs):
        n = rng.read_u64() % max_len                                     # tests/fasta_data.rs:45
        contigs.append(bytes(b"ACGT"[rng.read_u64() % 4] for _ in range(n)))   # tests/fasta_data.rs:19-26
    return contigs


def run_reference_test_on_oracle(contigs, block=3):
    """tests/fasta_data.rs:51-108 with the oracle in the reference's place; returns where the test would end"""
    if any(len(c) == 0 for c in contigs):
        return {"outcome": "panic: encode_sequence(empty).unwrap()", "decoded": 0}
    ix = ob.OracleIndex.build([np.frombuffer(c, dtype=np.uint8) for c in contigs], K, add_revcomp=True)
    blob = b""
    dropped = []
    lens = []
    for b in range(0, len(contigs), block):
        recs = []
        for c in contigs[b: b + block]:
            d = ix.encode_sequence(np.frombuffer(c, dtype=np.uint8), fast=False)
            lens.append([x[0] for x in d])
            recs.append(ix.encode_dictionary(d))
        rec = np.concatenate(recs)
        num = block if len(contigs[b: b + block]) == block else len(contigs) % block
        data = ob.write_block(rec, num)
        if data is None:
            dropped.append(b // block)                                   # `let _ = write_block_to(..)`: silently nothing written
        else:
            blob += data
    got = []
    pos = 0
    while True:
        try:
            r = ob.read_block(blob, pos)
        except RuntimeError:
            break
        if r is None:
            break
        rec, _, used = r
        bases, off = ix.decode(rec)
        for i in range(len(off) - 1):
            got.append(bytes(bases[int(off[i]): int(off[i + 1])]))
        pos += used
    outcome = "pass"
    for i, g in enumerate(got):
        if i >= len(contigs) or g != contigs[i]:
            outcome = "assert_eq!(*nucleotides, contigs[%d]) fails (tests/fasta_data.rs:99)" % i
            break
    if outcome == "pass" and len(got) != len(contigs):
        outcome = "assert_eq!(n_contigs, i) fails with i = %d (tests/fasta_data.rs:105)" % len(got)
    return {"outcome": outcome, "decoded": len(got), "blocks_dropped": dropped, "record_lengths": lens,
            "closed_form_agrees": all(predicted_record_lengths(len(c)) in (None, l) for l, c in zip(lens, contigs))}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--fixtures", type=int, default=100000)
    ap.add_argument("--check", type=int, default=60, help="random self-indexed contigs for closed form vs oracle")
    ap.add_argument("--full", type=int, default=2000, help="fixtures run through the oracle end to end")
    args = ap.parse_args()
    rng = np.random.default_rng(20250731)

    # 1. closed form vs oracle (a second, long contig keeps the index non-empty for the short ones)
    agree = 0
    lengths = [int(x) for x in rng.integers(255, 2000, size=args.check)] + [255, 256, 257, 258, 266, 267, 268, 510, 511, 512, 521, 522, 523, 1999]
    for n in lengths:
        seq = np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, size=n)]
        other = np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, size=900)]
        ix = ob.OracleIndex.build([seq, other], K, add_revcomp=True)
        got = [x[0] for x in ix.encode_sequence(seq, fast=False)]
        agree += got == predicted_record_lengths(n)
        if got != predicted_record_lengths(n):
            print("closed form differs at n = %d: oracle %s, predicted %s" % (n, got, predicted_record_lengths(n)), file=sys.stderr)

    # 2a. Monte Carlo over fixtures of the test's shape, closed form
    L = rng.integers(0, 2000, size=(args.fixtures, 7))
    passes = sum(fixture_passes([int(x) for x in row]) for row in L)
    # 2b. the same question asked of the oracle itself, on fewer fixtures (index build + encode + blocks + decode each)
    full_pass = 0
    full_agree = 0
    outcomes = {}
    for f in range(args.full):
        lens = [int(x) for x in rng.integers(0, 2000, size=7)]
        contigs = [bytes(np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, size=n)]) for n in lens]
        res = run_reference_test_on_oracle(contigs)
        full_pass += res["outcome"] == "pass"
        full_agree += (res["outcome"] == "pass") == fixture_passes(lens)
        key = res["outcome"].split(" with ")[0].split("[")[0]
        outcomes[key] = outcomes.get(key, 0) + 1

    # 3. the seed of the test under candidate seedings
    cands = {}
    for shifts in ((23, 17, 26), (23, 18, 5)):
        for name, (s0, s1) in candidate_seedings(20250731).items():
            contigs = fixture(Xorshift128Plus(s0, s1, shifts))
            res = run_reference_test_on_oracle(contigs)
            res["contig_lengths"] = [len(c) for c in contigs]
            res["closed_form_says"] = "pass" if fixture_passes([len(c) for c in contigs]) else "fail"
            res.pop("record_lengths")
            cands["shifts %s, state %s" % (shifts, name)] = res

    out = {
        "closed_form_vs_oracle": {"contigs": len(lengths), "agree": int(agree)},
        "monte_carlo": {"fixtures": args.fixtures, "would_pass": int(passes), "p_pass": passes / args.fixtures,
                        "note": "a fixture passes only if each of its 3 blocks (3 + 3 + 1 contigs) holds a literal AND a reference record"},
        "monte_carlo_full_oracle": {"fixtures": args.full, "would_pass": int(full_pass), "p_pass": full_pass / max(args.full, 1),
                                    "closed_form_verdict_agrees": int(full_agree), "outcomes": outcomes},
        "seed_20250731_under_candidate_seedings": cands,
        "any_candidate_passes": any(v["outcome"] == "pass" for v in cands.values()),
    }
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
