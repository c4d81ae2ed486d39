"""numpy restatement of the synthetic genome / read generator of SURVEY section 8d (splitmix64, counter based).

TEST INFRASTRUCTURE, like everything under oracle/: it lets the reference arm of bench.py (`--impl reference`)
and the tests make their inputs WITHOUT loading the product library.  tests/test_index_build.py checks it
against the product's C++ / CUDA generators (ntcomp_b200/csrc/synth_common.h).
"""
import numpy as np

GENOME_SEED = 20250731
ALPHA = np.frombuffer(b"ACGT", dtype=np.uint8)
_M1 = np.uint64(0xBF58476D1CE4E5B9)
_M2 = np.uint64(0x94D049BB133111EB)
_GOLD = np.uint64(0x9E3779B97F4A7C15)
_READ_MUL = np.uint64(0xD1B54A32D192ED03)


def splitmix64(x):
    with np.errstate(over="ignore"):
        z = np.asarray(x, dtype=np.uint64) + _GOLD
        z = (z ^ (z >> np.uint64(30))) * _M1
        z = (z ^ (z >> np.uint64(27))) * _M2
        return z ^ (z >> np.uint64(31))


def genome(n, seed=GENOME_SEED):
    out = np.empty(n, dtype=np.uint8)
    step = 1 << 22
    for lo in range(0, n, step):
        hi = min(n, lo + step)
        with np.errstate(over="ignore"):
            j = np.uint64(seed) + np.arange(lo, hi, dtype=np.uint64)
        out[lo:hi] = ALPHA[(splitmix64(j) & np.uint64(3)).astype(np.intp)]
    return out


def reads(genome_arr, seed, r0, n_reads, read_len, err_ppm):
    """reads r0 .. r0+n_reads: (bases uint8[n_reads * read_len], offsets uint64[n_reads + 1])"""
    g = np.ascontiguousarray(genome_arr, dtype=np.uint8)
    G, L = len(g), int(read_len)
    codes = np.zeros(256, dtype=np.int8)
    for i, b in enumerate(b"ACGT"):
        codes[b] = i
    gc = codes[g]
    out = np.empty(n_reads * L, dtype=np.uint8)
    t = np.arange(L, dtype=np.uint64)
    step = max(1, (1 << 22) // L)
    with np.errstate(over="ignore"):
        for lo in range(0, n_reads, step):
            hi = min(n_reads, lo + step)
            r = np.arange(r0 + lo, r0 + hi, dtype=np.uint64)
            key = np.uint64(seed) ^ (r * _READ_MUL)
            start = (splitmix64(key) % np.uint64(G - L + 1)).astype(np.int64)
            strand = (splitmix64(key + np.uint64(1)) & np.uint64(1)).astype(bool)
            pos = np.where(strand[:, None], start[:, None] + (L - 1 - t.astype(np.int64))[None, :], start[:, None] + t.astype(np.int64)[None, :])
            c = gc[pos].astype(np.int64)
            c = np.where(strand[:, None], 3 - c, c)
            u = splitmix64(key[:, None] + np.uint64(2) + t[None, :])
            hit = ((u >> np.uint64(32)) % np.uint64(1_000_000)) < np.uint64(err_ppm)
            c = np.where(hit, (c + 1 + (u % np.uint64(3)).astype(np.int64)) & 3, c)
            out[lo * L: hi * L] = ALPHA[c.reshape(-1)]
    offsets = np.arange(n_reads + 1, dtype=np.uint64) * np.uint64(L)
    return out, offsets
