#!/usr/bin/env python
"""Fused column codec on 16 blocks of C3 reads (row f1): ntc_encode_columns + ntc_decode_columns_fasta, three repetitions.
The target of the ncu captures of the column kernels (profiles/r2/ncu_*_columns.txt); prints per-class kernel times.
    python tools/columns_probe.py [n_blocks]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ntcomp_b200 as nt  # noqa: E402
from ntcomp_b200 import synth  # noqa: E402

n_blocks = int(sys.argv[1]) if len(sys.argv) > 1 else 16
cfg = synth.CONFIGS["C3"]
g = synth.genome(cfg["genome"])
idx = nt.Index.build([g], cfg["k"])
ctx = nt.Context(idx, 0)
n = n_blocks * 65536
bases, off = synth.reads(g, cfg["seed"], 0, n, cfg["read_len"], cfg["err_ppm"])
ctx.set_profiling(True)
for rep in range(3):
    t0 = time.perf_counter()
    blocks, words, n_rec = ctx.encode_columns(bases, off, 65536)
    t1 = time.perf_counter()
    text, n_seqs, n_dec = ctx.decode_columns_fasta(blocks, words, 1, full_last_literal_word=True)
    t2 = time.perf_counter()
    assert n_seqs == n and n_dec == n_rec
kt = ctx.kernel_times(reset=True)
print("reads=%d records=%d coded_bytes=%d (%.2f B/record) encode_call=%.2f ms decode_call=%.2f ms" %
      (n, n_rec, len(words) * 8, len(words) * 8 / n_rec, (t1 - t0) * 1e3, (t2 - t1) * 1e3))
print({k: (round(v[0] / 3, 3), v[1] // 3) for k, v in kt.items() if v[1]})
