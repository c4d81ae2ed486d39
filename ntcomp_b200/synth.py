"""Synthetic genomes and reads of BASELINE.json's shapes: counter-based splitmix64, so
every rank (and the oracle) can regenerate any read from (seed, read index) alone."""
import ctypes as C

import numpy as np

from . import _lib

GENOME_SEED = 20250731  # homage to tests/fasta_data.rs:37

# BASELINE.json configs (C1 is the tests/fasta_data.rs shape and lives in tests/)
CONFIGS = {
    "C2": dict(genome=5_000_000, k=31, n_reads=1_000_000, read_len=150, err_ppm=10_000, seed=20250801 + 2000),
    "C3": dict(genome=5_000_000, k=91, n_reads=10_000_000, read_len=150, err_ppm=10_000, seed=20250801 + 3000),
    "C4": dict(genome=12_000_000, k=91, n_reads=100_000_000, read_len=150, err_ppm=10_000, seed=20250801 + 4000),
    "C5": dict(genome=5_000_000, k=91, n_reads=2_000_000, read_len=10_000, err_ppm=50_000, seed=20250801 + 5000),
}


def genome(n, seed=GENOME_SEED):
    out = np.empty(n, dtype=np.uint8)
    _lib.lib().ntc_synth_genome(seed, n, out.ctypes.data_as(C.c_void_p))
    return out


def reads(genome_arr, seed, r0, n_reads, read_len, err_ppm, n_threads=0, out=None):
    """reads r0 .. r0+n_reads as one uint8 array of n_reads*read_len bases plus uint64 offsets"""
    g = np.ascontiguousarray(genome_arr, dtype=np.uint8)
    if out is None:
        out = np.empty(n_reads * read_len, dtype=np.uint8)
    _lib.lib().ntc_synth_reads(g.ctypes.data_as(C.c_void_p), len(g), seed, r0, n_reads, read_len, err_ppm,
                               out.ctypes.data_as(C.c_void_p), n_threads)
    offsets = np.arange(n_reads + 1, dtype=np.uint64) * np.uint64(read_len)
    return out, offsets
