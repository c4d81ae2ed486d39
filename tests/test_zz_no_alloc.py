"""ntc_decode_device_static: the decode entry point for callers that are not allowed to allocate.

Who needs it: a host process that has handed all of the GPU's memory to its own pool (so a cudaMalloc made behind its back
fails or fragments the pool), a service that cannot take the device-wide synchronisation of cudaMalloc / cudaFree inside a
request, and a sandboxed plugin that may launch kernels but not allocate.  ntc_decode_device grows its scan workspaces with
cudaFree + cudaMalloc when a batch is larger than any before; the static variant runs the same kernels on an array that is
part of the library image and therefore has a fixed limit (NTC_STATIC_MAX_RECORDS records per call).

The file name sorts last on purpose: pytest -x stops at the first failure, so a problem here must not hide any other test.
The test builds its own index and context, so it passes alone in a fresh process as well as after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

import helpers
import ntcomp_b200 as nt
from ntcomp_b200 import _lib
from conftest import has_gpu

GUARD = 0xA5


def _decode(fn, ctx, torch, d_rec, n_rec, first_no, fasta, out_bytes, n_seqs, stream):
    """one call of `fn` (either decode entry point, raw C ABI) into fresh guarded buffers -> (text, seq offsets, counts)"""
    dev = d_rec.device
    d_out = torch.full((out_bytes + 128,), GUARD, dtype=torch.uint8, device=dev)
    d_soff = torch.full((n_seqs + 1 + 8,), -1, dtype=torch.int64, device=dev)
    d_cnt = torch.full((4,), -1, dtype=torch.int64, device=dev)
    rc = fn(ctx._h, d_rec.data_ptr(), n_rec, first_no, int(fasta), d_out.data_ptr() + 64, out_bytes,
            0 if fasta else d_soff.data_ptr(), 0 if fasta else n_seqs, d_cnt.data_ptr(), stream)
    assert rc == 0, (rc, _lib.lib().ntc_last_error())
    ctx.status(stream)                         # synchronises the stream and raises on a device-side error
    out = d_out.cpu().numpy()
    assert (out[:64] == GUARD).all() and (out[64 + out_bytes:] == GUARD).all()     # nothing outside [0, out_bytes)
    return bytes(out[64: 64 + out_bytes]), d_soff.cpu().numpy()[: n_seqs + 1].view(np.uint64), d_cnt.cpu().numpy()[:2]


@pytest.mark.gpu
def test_static_scratch_decode_matches_decode_device_twice_in_a_row():
    """Two calls in a row through the C ABI, each compared byte for byte (integer work: no tolerance) with what
    ntc_decode_device gives for the same records: first a whole batch as plain bases with sequence offsets, then a shorter
    batch as FASTA text with another first sequence number -- the second call finds the first call's values in the shared
    array and must not be affected by them.  Then the first batch once more, and the size limit."""
    if not has_gpu():
        pytest.skip("needs a GPU")
    import torch
    L = _lib.lib()
    rng = np.random.default_rng(20251106)
    contigs = [helpers.random_seq(rng, 40000), helpers.random_seq(rng, 3000)]
    k = 31
    index = nt.Index.build(contigs, k, True)
    ctx = nt.Context(index, 0)
    reads = helpers.sample_reads(rng, contigs, 600, lengths=[1, 5, k - 1, k, k + 1, 3 * k + 7, 150, 150, 150, 400, 2500],
                                 error_rates=[0, 0.01, 0.01, 0.05, 0.3], k=k)
    bases, off = nt.concat_reads(reads)
    rec, roff = ctx.encode_batch(bases, off)
    dev = torch.device("cuda:0")
    d_rec = torch.from_numpy(rec.view(np.int64)).to(dev)
    stream = torch.cuda.current_stream().cuda_stream
    assert nt.Context.static_max_records() == 524288 and len(rec) <= 524288

    # batch 1: everything, plain bases + offsets.  batch 2: the first 211 reads only, FASTA numbered from 1000.
    m = 211
    n_rec2 = int(roff[m])
    text2 = helpers.fasta_text(reads[:m], 1000)
    jobs = [(len(rec), 1, False, len(bases), len(reads)), (n_rec2, 1000, True, len(text2), m), (len(rec), 1, False, len(bases), len(reads))]
    want = [_decode(L.ntc_decode_device, ctx, torch, d_rec, *job, stream) for job in jobs]
    assert want[0][0] == bytes(bases) and (want[0][1] == off).all() and want[1][0] == text2          # the yardstick itself is right
    for job, (w_text, w_off, w_cnt) in zip(jobs, want):                                              # back to back on the static scratch
        g_text, g_off, g_cnt = _decode(L.ntc_decode_device_static, ctx, torch, d_rec, *job, stream)
        assert (g_cnt == w_cnt).all(), (job, g_cnt, w_cnt)
        assert g_text == w_text, job
        if not job[2]:
            assert (g_off == w_off).all(), job

    # an empty batch is fine; one record too many is refused before anything is enqueued
    d_cnt = torch.full((4,), -1, dtype=torch.int64, device=dev)
    assert L.ntc_decode_device_static(ctx._h, 0, 0, 1, 0, 0, 0, 0, 0, d_cnt.data_ptr(), stream) == 0
    ctx.status(stream)
    assert d_cnt[:2].tolist() == [0, 0]
    d_cnt.fill_(-1)
    rc = L.ntc_decode_device_static(ctx._h, d_rec.data_ptr(), 524288 + 1, 1, 0, 0, 0, 0, 0, d_cnt.data_ptr(), stream)
    assert rc == 3 and b"524288" in L.ntc_last_error()                                               # NTC_ERR_CAPACITY, says the limit
    ctx.status(stream)
    assert d_cnt[:2].tolist() == [-1, -1]
    ctx.close()
