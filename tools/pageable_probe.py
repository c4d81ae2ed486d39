"""ntc_encode_batch / ntc_decode_fasta with PAGEABLE host buffers (what a Rust Vec or a numpy array is) against page-locked ones, with and without
host-side packing (ntc_ctx_set_host_threads): C3 reads, one GPU.  Prints one JSON line."""
import ctypes as C
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ntcomp_b200 as nt  # noqa: E402
from ntcomp_b200 import _lib, synth  # noqa: E402

n_reads = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
L = 150
g = synth.genome(5_000_000)
idx = nt.Index.build_gpu([g], 91)
ctx = nt.Context(idx, 0)
bases, off = synth.reads(g, 20250801 + 3000, 0, n_reads, L, 10_000)        # numpy arrays: pageable
nb = len(bases)
lib = _lib.lib()
rec = np.empty(nb // 6, dtype=np.uint64)
text = np.empty(nb + n_reads * 24 + 64, dtype=np.uint8)
p_bases = torch.from_numpy(bases).pin_memory()
p_off = torch.from_numpy(off.astype(np.int64)).pin_memory()
p_rec = torch.empty(nb // 6, dtype=torch.int64).pin_memory()
p_text = torch.empty(nb + n_reads * 24 + 64, dtype=torch.uint8).pin_memory()
out = {"reads": n_reads, "bases": nb, "host_threads_available": os.cpu_count()}


def run(label, b_ptr, o_ptr, r_ptr, t_ptr, threads):
    ctx.set_host_threads(threads)
    n_rec, n_bytes, n_seqs = C.c_uint64(0), C.c_uint64(0), C.c_uint64(0)
    best = [1e9, 1e9]
    for _ in range(3):
        t0 = time.perf_counter()
        _lib.check(lib.ntc_encode_batch(ctx._h, b_ptr, o_ptr, n_reads, r_ptr, nb // 6, C.byref(n_rec), None))
        t1 = time.perf_counter()
        _lib.check(lib.ntc_decode_fasta(ctx._h, r_ptr, n_rec.value, 1, t_ptr, nb + n_reads * 24 + 64, C.byref(n_bytes), C.byref(n_seqs)))
        t2 = time.perf_counter()
        best = [min(best[0], t1 - t0), min(best[1], t2 - t1)]
    out[label] = {"encode_ms": round(best[0] * 1e3, 2), "decode_ms": round(best[1] * 1e3, 2), "records": n_rec.value}


for threads in (0, 8, 16):
    run("pinned_threads%d" % threads, p_bases.data_ptr(), p_off.data_ptr(), p_rec.data_ptr(), p_text.data_ptr(), threads)
    run("pageable_threads%d" % threads, bases.ctypes.data, off.ctypes.data, rec.ctypes.data, text.ctypes.data, threads)
print(json.dumps(out))
