"""Seeded test-data helpers shared by the CPU and GPU suites."""
import numpy as np

ALPHA = np.frombuffer(b"ACGT", dtype=np.uint8)
COMP = np.zeros(256, dtype=np.uint8)
for a, b in zip(b"ACGT", b"TGCA"):
    COMP[a] = b


def random_seq(rng, n):
    return ALPHA[rng.integers(0, 4, size=n)]


def revcomp(a):
    return COMP[a[::-1]]


def mutate(rng, read, rate):
    """substitutions to a different base"""
    read = read.copy()
    if rate <= 0 or len(read) == 0:
        return read
    hit = rng.random(len(read)) < rate
    codes = np.searchsorted(ALPHA, read)
    codes[hit] = (codes[hit] + rng.integers(1, 4, size=int(hit.sum()))) & 3
    return ALPHA[codes]


def sample_reads(rng, contigs, n_reads, lengths, error_rates, k):
    """reads drawn from the contigs: both strands, ragged lengths, substitution errors, some starting
    at contig position 0 (dummy-node territory), some shorter than k, some of length 1"""
    out = []
    for _ in range(n_reads):
        src = contigs[rng.integers(0, len(contigs))]
        L = int(lengths[rng.integers(0, len(lengths))])
        L = max(1, min(L, len(src)))
        a = 0 if rng.random() < 0.1 else int(rng.integers(0, len(src) - L + 1))
        r = src[a:a + L]
        if rng.random() < 0.5:
            r = revcomp(r)
        r = mutate(rng, r, float(error_rates[rng.integers(0, len(error_rates))]))
        out.append(np.ascontiguousarray(r))
    return out


def fasta_text(reads, first=1):
    """the writer loop of src/main.rs:201-209"""
    parts = []
    for i, r in enumerate(reads):
        parts.append(b">seq.%d\n" % (first + i))
        parts.append(bytes(r))
        parts.append(b"\n")
    return b"".join(parts)


def fixture_c1(seed=20250731):
    """tests/fasta_data.rs shape: 7 contigs of length rng % 2000, k = 255, block size 3.  The Rust
    `random` crate's stream is not reproducible here, so any seeded ACGT generator of the same shape
    exercises the same property."""
    rng = np.random.default_rng(seed)
    contigs = []
    while len(contigs) < 7:
        n = int(rng.integers(0, 2000))
        if n == 0:
            continue  # a zero-length contig makes the reference test panic at fasta_data.rs:69
        contigs.append(random_seq(rng, n))
    return contigs, 255, 3
