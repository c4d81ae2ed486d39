"""Host block container of the product (ntcomp_b200/csrc/container.cpp) against the oracle's
restatement (P2: header fields and inflated payload words identical; both link system zlib so the
members are byte-identical here too)."""
import ctypes as C

import numpy as np
import pytest

import helpers
import ntcomp_b200 as nt
from ntcomp_b200 import _lib
from oracle import binding as ob


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def column_code(values, rice):
    v = np.ascontiguousarray(values, dtype=np.uint64)
    cap = 2 * len(v) + 1024
    words = np.zeros(cap, dtype=np.uint64)
    nw, param = C.c_uint64(0), C.c_uint64(0)
    rc = _lib.lib().ntc_column_code(_ptr(v), len(v), int(rice), _ptr(words), cap, C.byref(nw), C.byref(param))
    return rc, words[: nw.value].copy(), param.value


def block_write(rec, num_reads):
    rec = np.ascontiguousarray(rec, dtype=np.uint64)
    cap = int(_lib.lib().ntc_block_bound(len(rec)))
    out = np.zeros(cap, dtype=np.uint8)
    nb = C.c_uint64(0)
    rc = _lib.lib().ntc_block_write(_ptr(rec), len(rec), num_reads, _ptr(out), cap, C.byref(nb))
    return rc, out[: nb.value].tobytes()


def block_read(data, offset=0):
    buf = np.frombuffer(data, dtype=np.uint8)[offset:].copy()
    cap = 8 * len(buf) + 1024
    rec = np.zeros(cap, dtype=np.uint64)
    n, used = C.c_uint64(0), C.c_uint64(0)
    rc = _lib.lib().ntc_block_read(_ptr(buf) if len(buf) else None, len(buf), _ptr(rec), cap, C.byref(n), C.byref(used))
    return rc, rec[: n.value].copy(), used.value


def test_column_coders_match_oracle():
    rng = np.random.default_rng(1)
    for _ in range(30):
        n = int(rng.integers(1, 4000))
        vals = rng.integers(0, int(rng.choice([2, 48, 300, 1 << 12, 1 << 24, 1 << 32])), size=n).astype(np.uint64)
        for rice in (False, True):
            if rice and vals.max() > (1 << 16):
                continue
            rc, w, p = column_code(vals, rice)
            ow, op = ob.code_column(vals, rice)
            assert rc == 0 and p == op and (w == ow).all()
    rc, _, _ = column_code(np.zeros(0, dtype=np.uint64), False)
    assert rc == 10                                                  # NTC_ERR_EMPTY_COLUMN (Q4)
    rc, w, p = column_code(np.array([5, 0, 1 << 20], dtype=np.uint64), True)   # long unary run
    ow, op = ob.code_column(np.array([5, 0, 1 << 20], dtype=np.uint64), True)
    assert rc == 0 and p == op and (w == ow).all()


def test_block_write_read_matches_oracle():
    rng = np.random.default_rng(2)
    g = helpers.random_seq(rng, 6000)
    o = ob.OracleIndex.build([g], 20, True)
    for n_reads, rate in [(300, 0.03), (40, 0.0), (1000, 0.1)]:
        reads = helpers.sample_reads(rng, [g], n_reads, [150, 30, 400], [rate], 20)
        bases, off = ob.concat_reads(reads)
        rec, _ = o.encode_reads(bases, off, fast=True, threads=4)
        rc, blob = block_write(rec, n_reads)
        oblob = ob.write_block(rec, n_reads)
        if oblob is None:
            assert rc == 10
            continue
        assert rc == 0 and blob == oblob
        n_lit = int((((rec >> np.uint64(56)) & np.uint64(2)) >> np.uint64(1) * ((rec >> np.uint64(58)) & np.uint64(63))).sum())
        rc, rec2, used = block_read(blob + b"\0" * 7)
        orec = ob.read_block(blob) if True else None
        if n_lit % 31 == 0 and n_lit:                               # Q5: the reference decoder panics
            assert rc == 7
        else:
            assert rc == 0 and used == len(blob) and (rec2 == rec).all() and (orec[0] == rec).all()
    rc, rec2, used = block_read(b"")                                 # EOF
    assert rc == 0 and used == 0 and len(rec2) == 0
    rc, _, _ = block_read(blob[: len(blob) // 2])                    # truncated
    assert rc == 7


def test_q5_block_is_refused_like_the_reference_and_decoded_with_the_flag():
    """literal bases a non-zero multiple of 31: the reference's zip_block_contents reads the last literal word
    as empty and panics (src/decode.rs:114-118,138); ntc_block_read mirrors that, ntc_block_read_ex with
    NTC_BLOCK_FULL_LAST_LITERAL_WORD returns what the encoder wrote (the file pipeline uses it)."""
    import ntcomp_b200 as nt
    rng = np.random.default_rng(31)
    for n_lit_records, L in [(31, 1), (31, 2), (31, 11), (62, 3)]:
        recs = [7 | (40 << 32) | (1 << 56)]                                            # one reference record
        for i in range(n_lit_records):
            bits = int(rng.integers(0, 1 << (2 * L)))
            recs.append(bits | (((i % 2) + 2) | (L << 2)) << 56)
        rec = np.array(recs, dtype=np.uint64)
        rc, blob = block_write(rec, 5)
        assert rc == 0 and blob == ob.write_block(rec, 5)
        rc, _, _ = block_read(blob)
        assert rc == 7
        rec2, used = nt.block_read(blob, full_last_literal_word=True)
        assert used == len(blob) and (rec2 == rec).all()
    # the flag changes nothing for blocks the reference can read
    rec = np.array([7 | (40 << 32) | (1 << 56), 0b1101 | ((0 + 2) | (2 << 2)) << 56], dtype=np.uint64)
    rc, blob = block_write(rec, 1)
    assert (nt.block_read(blob, True)[0] == rec).all() and (nt.block_read(blob, False)[0] == rec).all()


def test_q4_blocks_are_refused_like_the_reference():
    rec = np.array([5 | (20 << 32) | (1 << 56), 9 | (30 << 32)], dtype=np.uint64)     # reference records only
    assert block_write(rec, 1)[0] == 10
    lit = np.array([(0b1101 | ((1 + 2) | (2 << 2)) << 56)], dtype=np.uint64)          # literal records only
    assert block_write(lit, 1)[0] == 10


def _columns_of(rec):
    """(words, param) of the four columns through the host coder, via the oracle's split"""
    cols = ob.split(rec)
    return [column_code(cols[q], rice)[1:] for q, rice in enumerate((False, True, True, False))]


def _blocks_of(rec, roff, block_reads):
    """(record slice, reads in block) per block of block_reads reads"""
    n = len(roff) - 1
    return [(rec[int(roff[lo]): int(roff[min(n, lo + block_reads)])], min(n, lo + block_reads) - lo) for lo in range(0, n, block_reads)]


def _q4_free_reads(rng, g, n, k):
    """error-free reads whose cover holds reference records only (no literal): with one such block in a batch the
    reference drops it (Q4) while its neighbours are written"""
    out = []
    while len(out) < n:
        a = int(rng.integers(0, len(g) - 2 * k - 40))
        out.append(np.ascontiguousarray(g[a: a + k + 30]))
    return out


def test_gpu_column_codec_bodies_on_cpu_emulation():
    """column_kernels.cuh (split to ranks, gathered literal words, one code-length scan over all (block, column) segments,
    one thread per output word) + the planning of column_plan.h, run on the CPU over MULTI-block batches == the oracle's
    sequential coders per block (src/encode.rs:59-94,168-229)."""
    from emul_binding import lib as emul_lib
    L = emul_lib()
    L.emul_batch_columns.restype = C.c_int
    L.emul_batch_columns.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]
    rng = np.random.default_rng(3)
    g = helpers.random_seq(rng, 8000)
    o = ob.OracleIndex.build([g], 20, True)
    for n_reads, rate, lens, block_reads in [(400, 0.03, [150], 64), (50, 0.2, [40, 400], 7), (800, 0.01, [150, 2000], 800), (90, 0.0, [35], 30)]:
        reads = helpers.sample_reads(rng, [g], n_reads, lens, [rate], 20)
        rec, roff = o.encode_reads(*ob.concat_reads(reads), fast=True, threads=4)
        blk_rec = np.array([int(roff[min(n_reads, b)]) for b in range(0, n_reads + block_reads, block_reads)][: (n_reads + block_reads - 1) // block_reads + 1], dtype=np.uint64)
        n_blocks = len(blk_rec) - 1
        words = np.zeros(4 * len(rec) + 4096, dtype=np.uint64)
        seg = np.zeros(16 * n_blocks, dtype=np.uint64)
        status = np.zeros(n_blocks, dtype=np.int32)
        total = C.c_uint64(0)
        assert L.emul_batch_columns(_ptr(rec), _ptr(blk_rec), n_blocks, _ptr(words), len(words), _ptr(seg), _ptr(status), C.byref(total)) == 0
        seg = seg.reshape(n_blocks, 4, 4)
        for b in range(n_blocks):
            brec = rec[int(blk_rec[b]): int(blk_rec[b + 1])]
            cols = ob.split(brec)
            if len(cols[0]) == 0 or len(cols[3]) == 0:
                assert status[b] == 10                                   # Q4
                continue
            assert status[b] == 0
            for q, rice in enumerate((False, True, True, False)):
                ow, op = ob.code_column(cols[q], rice)
                nv, nw, param, wb = (int(x) for x in seg[b, q])
                assert (nv, nw, param) == (len(cols[q]), len(ow), op), (b, q)
                assert (words[wb: wb + nw] == ow).all(), (b, q)


def test_gpu_literal_and_packing_bodies_edge_cases_on_cpu_emulation():
    """literal records of every length 0..32 (foreign streams; bases 28..31 of a record read as A, src/encode.rs:215-216),
    literal words that straddle many records, long unary runs, values needing 63-bit minimal-binary codes"""
    from emul_binding import lib as emul_lib
    L = emul_lib()
    L.emul_batch_columns.restype = C.c_int
    L.emul_batch_columns.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]
    rng = np.random.default_rng(33)
    recs = []
    for i in range(3000):
        if rng.random() < 0.4:
            ln = int(rng.choice([12, 92, 150, 1 << 20, (1 << 24) - 1])) if rng.random() < 0.2 else int(rng.integers(12, 200))
            recs.append(int(rng.integers(0, 1 << 32)) | (ln << 32) | (int(rng.integers(0, 2)) << 56))
        else:
            Lb = int(rng.integers(0, 33))
            bits = int(rng.integers(0, 1 << 56)) & ((1 << min(56, 2 * Lb)) - 1) if Lb else 0
            recs.append(bits | (((int(rng.integers(0, 2)) + 2) | (Lb << 2)) << 56))
    rec = np.array(recs, dtype=np.uint64)
    for cuts in ([0, 3000], [0, 1, 2, 700, 701, 3000], [0, 1500, 1500, 3000]):
        blk_rec = np.array(cuts, dtype=np.uint64)
        n_blocks = len(cuts) - 1
        words = np.zeros(40 * len(rec) + 4096, dtype=np.uint64)
        seg = np.zeros(16 * n_blocks, dtype=np.uint64)
        status = np.zeros(n_blocks, dtype=np.int32)
        total = C.c_uint64(0)
        assert L.emul_batch_columns(_ptr(rec), _ptr(blk_rec), n_blocks, _ptr(words), len(words), _ptr(seg), _ptr(status), C.byref(total)) == 0
        seg = seg.reshape(n_blocks, 4, 4)
        for b in range(n_blocks):
            brec = rec[cuts[b]: cuts[b + 1]]
            cols = ob.split(brec) if len(brec) else None
            if cols is None or len(cols[0]) == 0 or len(cols[3]) == 0:
                assert status[b] in (4, 10)
                continue
            for q, rice in enumerate((False, True, True, False)):
                ow, op = ob.code_column(cols[q], rice)
                nv, nw, param, wb = (int(x) for x in seg[b, q])
                assert (nv, nw, param) == (len(cols[q]), len(ow), op), (cuts, b, q)
                assert (words[wb: wb + nw] == ow).all(), (cuts, b, q)


def _emul_decode(words, n, rice, param):
    from emul_binding import lib as emul_lib
    L = emul_lib()
    L.emul_column_decode.restype = C.c_int64
    L.emul_column_decode.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_int, C.c_uint64, C.c_void_p]
    w = np.ascontiguousarray(words, dtype=np.uint64)
    out = np.zeros(max(n, 1), dtype=np.uint64)
    total = L.emul_column_decode(_ptr(w), len(w), n, int(rice), int(param), _ptr(out))
    return total, out[:n]


def test_gpu_column_decoder_bodies_on_cpu_emulation():
    """column_decode_kernels.cuh (speculative per-chunk tables, resolve, extract) run on the CPU == the values that
    were coded == the oracle's sequential decoder.  Long unary runs spanning several chunks (state M), remainder bits
    straddling chunk boundaries (P_k), minimal-binary codes of every length up to 63 bits, Rice parameter 0."""
    rng = np.random.default_rng(17)
    cases = []
    for hi in (2, 3, 48, 300, 1 << 12, 1 << 24, 1 << 32, 1 << 45, (1 << 62) - 7):
        for n in (1, 2, 63, 700, 5000):
            cases.append(rng.integers(0, hi, size=n).astype(np.uint64))
    spiky = rng.integers(0, 4, size=3000).astype(np.uint64)
    spiky[[5, 900, 901, 2999]] = [9000, 40000, 3000, 70000]           # unary runs of thousands of zeros at a small parameter
    cases.append(spiky)
    cases.append(np.zeros(10000, dtype=np.uint64))                     # Rice: sum 0 -> parameter 0, one bit per code
    cases.append(np.full(4097, 1, dtype=np.uint64))
    for vals in cases:
        for rice in (False, True):
            if rice and int(vals.max()) >= (1 << 40):
                continue                                               # a 2^40-zero unary run is not a test case
            rc, words, param = column_code(vals, rice)
            assert rc == 0
            total, got = _emul_decode(words, len(vals), rice, param)
            assert total >= len(vals), (rice, param, len(vals))
            assert (got == vals).all(), (rice, param, len(vals))
            assert (ob.decode_column(words, len(vals), rice, param) == vals).all()
    # a stream that ends early is reported: fewer codes than announced
    rc, words, param = column_code(np.arange(1, 2000, dtype=np.uint64) * 3, True)
    total, _ = _emul_decode(words[: len(words) // 2], 1999, True, param)
    assert total < 1999


def _cols_of_blocks(blobs):
    """disassemble written blocks into the (cols, words) arrays emul_columns_decode / ntc_decode_columns_fasta take"""
    cols, words, blocks = [], [], []
    at = 0
    for blob in blobs:
        blk, w, used = nt.block_disassemble(blob)
        assert used == len(blob)
        for q in range(4):
            blk.col[q].word_offset += at
            cols += [blk.col[q].n_values, blk.col[q].n_words, blk.col[q].param, blk.col[q].word_offset]
        at += len(w)
        words.append(w)
        blocks.append(blk)
    return np.array(cols, dtype=np.uint64), (np.concatenate(words) if words else np.zeros(0, dtype=np.uint64)), blocks


def _emul_columns_decode(cols, n_blocks, words, full, trim, cap):
    from emul_binding import lib as emul_lib
    L = emul_lib()
    L.emul_columns_decode.restype = C.c_int
    L.emul_columns_decode.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]
    out = np.zeros(max(cap, 1), dtype=np.uint64)
    n, detail = C.c_uint64(0), C.c_uint32(0)
    rc = L.emul_columns_decode(_ptr(cols), n_blocks, _ptr(words), len(words), full, trim, _ptr(out), cap, C.byref(n), C.byref(detail))
    return rc, out[: n.value], detail.value


def test_gpu_column_decode_and_zip_bodies_on_cpu_emulation():
    """multi-block batches through plan_decode + the decoder bodies + zip_thread (+ the per-block trim) == the records the
    oracle reads back block by block (src/decode.rs:51-149); Q5 and zip errors as the host reader reports them"""
    rng = np.random.default_rng(19)
    g = helpers.random_seq(rng, 8000)
    o = ob.OracleIndex.build([g], 20, True)
    for n_reads, rate, block_reads in [(300, 0.03, 64), (60, 0.2, 60), (500, 0.01, 137)]:
        reads = helpers.sample_reads(rng, [g], n_reads, [150, 35, 400], [rate], 20)
        rec, roff = o.encode_reads(*ob.concat_reads(reads), fast=True, threads=4)
        blobs, kept = [], []
        for brec, nr in _blocks_of(rec, roff, block_reads):
            blob = ob.write_block(brec, nr)
            if blob is not None:
                blobs.append(blob)
                kept.append(brec)
        cols, words, _ = _cols_of_blocks(blobs)
        q5 = any(int(np.where((r >> np.uint64(57)) & np.uint64(1), (r >> np.uint64(58)) & np.uint64(63), 0).sum()) % 31 == 0 for r in kept)
        for full in (0, 1):
            rc, got, detail = _emul_columns_decode(cols, len(blobs), words, full, 0, len(rec))
            if q5 and not full:
                assert rc == 7 and detail == (2 << 1)                    # Q5: ZIP_ERR_LIT
            else:
                assert rc == 0 and (got == np.concatenate(kept)).all()
                for blob, brec in zip(blobs, kept):
                    orec = ob.read_block(blob) if not q5 else None
                    assert orec is None or (orec[0] == brec).all()
    # per-block trim: records in front of a block's first `first` flag become empty literals, other blocks untouched
    a = np.array([5 | (20 << 32), 6 | (30 << 32), 7 | (40 << 32) | (1 << 56), (0b11 | ((2 | (1 << 2)) << 56))], dtype=np.uint64)
    b = np.array([9 | (25 << 32) | (1 << 56), (0b10 | ((2 | (1 << 2)) << 56))], dtype=np.uint64)
    cols, words, _ = _cols_of_blocks([ob.write_block(b, 1), ob.write_block(a, 1), ob.write_block(b, 1)])
    rc, got, _ = _emul_columns_decode(cols, 3, words, 0, 1, 64)
    want = np.concatenate([b, np.array([2 << 56, 2 << 56], dtype=np.uint64), a[2:], b])
    assert rc == 0 and (got == want).all()
    rc, got, _ = _emul_columns_decode(cols, 3, words, 0, 0, 64)
    assert rc == 0 and (got == np.concatenate([b, a, b])).all()
    # header fields from a hostile file: rejected by the planner before anything is sized by them
    bad = cols.copy(); bad[4 * 1 + 2] = 64                              # Rice parameter of the len column
    assert _emul_columns_decode(bad, 3, words, 0, 0, 64)[0] == 7
    bad = cols.copy(); bad[2] = 1                                       # minimal-binary bound below 2
    assert _emul_columns_decode(bad, 3, words, 0, 0, 64)[0] == 7
    bad = cols.copy(); bad[0] = 1 << 40                                 # more values than bits
    assert _emul_columns_decode(bad, 3, words, 0, 0, 64)[0] == 7
    bad = cols.copy(); bad[3] = len(words) + 5                          # words outside the buffer
    assert _emul_columns_decode(bad, 3, words, 0, 0, 64)[0] == 1


def test_malformed_headers_are_rejected_by_the_host_reader():
    """a hostile .dat must produce an error code, never an out-of-range shift, a 32 GiB reserve or an
    exception across the C ABI.  The GPU reader applies the same checks (test_block_read_gpu_matches_oracle)."""
    import struct
    rec = np.array([7 | (40 << 32) | (1 << 56), 0b1101 | ((0 + 2) | (2 << 2)) << 56], dtype=np.uint64)
    rc, blob = block_write(rec, 1)
    assert rc == 0
    for field_off, value in [(16, 64), (16, 1 << 40)]:                                     # Rice parameter of the len column (2nd sub-block)
        bs0 = struct.unpack_from("<I", blob, 0)[0]
        b = bytearray(blob)
        struct.pack_into("<Q", b, 32 + bs0 + field_off, value)
        assert block_read(bytes(b))[0] == 7
    for value in (0, 1):                                                                    # minimal-binary bound of the colex column
        b = bytearray(blob)
        struct.pack_into("<Q", b, 16, value)
        assert block_read(bytes(b))[0] == 7
    b = bytearray(blob)
    struct.pack_into("<I", b, 8, 0xFFFFFFFF)                                                # num_u64: 4 G values announced for a few bytes
    assert block_read(bytes(b))[0] == 7
    b = bytearray(blob)
    struct.pack_into("<I", b, 0, 0xFFFFFFF0)                                                # block_size beyond the data
    assert block_read(bytes(b))[0] == 7


def test_block_disassemble_and_assemble_round_trip_on_the_host():
    """the host halves of the fused codec (headers + gzip members) need no GPU: oracle block -> ntc_block_disassemble ->
    ntc_block_assemble / 4 x ntc_sub_block_assemble (what the file pipeline's gzip tasks call) -> the same bytes"""
    rng = np.random.default_rng(12)
    rec = []
    for i in range(3000):
        if rng.random() < 0.3:
            L = int(rng.integers(1, 12))
            rec.append(int(rng.integers(0, 1 << (2 * L))) | (((i % 7 == 0) + 2) | (L << 2)) << 56)
        else:
            rec.append(int(rng.integers(0, 1 << 32)) | (int(rng.integers(12, 1 << 24)) << 32) | (int(i % 7 == 0) << 56))
    rec = np.array(rec, dtype=np.uint64)
    want = ob.write_block(rec, 65536)
    blk, words, used = nt.block_disassemble(want)
    assert used == len(want) and blk.n_records == len(rec) and blk.num_reads == 65536
    assert nt.block_assemble(blk, words) == want
    assert b"".join(nt.sub_block_assemble(blk, q, words) for q in range(4)) == want
    with pytest.raises(nt.NtcError):
        nt.sub_block_assemble(blk, 4, words)
    # a missing word pointer for a non-empty column is an argument error, not a crash; a short buffer is a capacity error
    import ctypes as C
    from ntcomp_b200 import _lib
    out = np.zeros(len(want) + 1024, dtype=np.uint8)
    nb = C.c_uint64(7)
    L = _lib.lib()
    assert L.ntc_block_assemble(C.byref(blk), None, out.ctypes.data_as(C.c_void_p), len(out), C.byref(nb)) == 1 and nb.value == 0
    assert L.ntc_block_assemble(C.byref(blk), words.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p), len(want) - 1, C.byref(nb)) == 3
    assert nb.value == 0
    assert L.ntc_block_assemble(C.byref(blk), words.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p), len(out), C.byref(nb)) == 0
    assert bytes(out[: nb.value]) == want


@pytest.mark.gpu
def test_block_write_gpu_matches_oracle_bytes():
    """ntc_block_write_gpu (split + codes on the GPU) against the ORACLE's write_block directly (and the host container)"""
    rng = np.random.default_rng(4)
    g = helpers.random_seq(rng, 30000)
    idx = nt.Index.build([g], 31)
    ctx = nt.Context(idx, 0)
    o = ob.OracleIndex.from_arrays(idx.k, idx.n_sets, idx.rows(), idx.lcs())
    for n_reads, rate, lens in [(3000, 0.01, [150]), (65536, 0.01, [150]), (200, 0.3, [60, 1000]), (3, 0.05, [150])]:
        reads = helpers.sample_reads(rng, [g], n_reads, lens, [rate], 31)
        rec, _ = o.encode_reads(*ob.concat_reads(reads), fast=True, threads=8)
        want = ob.write_block(rec, n_reads)
        cap = int(_lib.lib().ntc_block_bound(len(rec)))
        out = np.zeros(cap, dtype=np.uint8)
        nb = C.c_uint64(0)
        rc = _lib.lib().ntc_block_write_gpu(ctx._h, _ptr(rec), len(rec), n_reads, _ptr(out), cap, C.byref(nb))
        if want is None:
            assert rc == 10
            continue
        assert rc == 0 and out[: nb.value].tobytes() == want
        assert block_write(rec, n_reads) == (0, want)
    only_ref = rec[((rec >> np.uint64(56)) & np.uint64(2)) == 0]
    nb = C.c_uint64(0)
    assert _lib.lib().ntc_block_write_gpu(ctx._h, _ptr(only_ref), len(only_ref), 1, _ptr(out), cap, C.byref(nb)) == 10


@pytest.mark.gpu
def test_block_read_gpu_matches_oracle():
    """ntc_block_read_gpu (integer decoders + zip on the GPU) against the ORACLE's read_block directly; errors as the host
    reader reports them"""
    import struct
    rng = np.random.default_rng(23)
    g = helpers.random_seq(rng, 60000)
    index = nt.Index.build([g], 31)
    ctx = nt.Context(index, 0)
    o = ob.OracleIndex.from_arrays(index.k, index.n_sets, index.rows(), index.lcs())

    def gpu_read(blob, flags=0):
        buf = np.frombuffer(blob, dtype=np.uint8)
        rec = np.zeros(8 * len(buf) + 1024, dtype=np.uint64)
        n, used = C.c_uint64(0), C.c_uint64(0)
        rc = _lib.lib().ntc_block_read_gpu(ctx._h, _ptr(buf) if len(buf) else None, len(buf), _ptr(rec), len(rec), C.byref(n), C.byref(used), flags)
        return rc, rec[: n.value].copy(), used.value

    for n_reads, rate, lens in [(65536, 0.01, [150]), (3000, 0.05, [150, 2000, 10000]), (500, 0.3, [60]), (7, 0.0, [400])]:
        reads = helpers.sample_reads(rng, [g], n_reads, lens, [rate], 31)
        rec, _ = o.encode_reads(*ob.concat_reads(reads), fast=True, threads=8)
        blob = ob.write_block(rec, n_reads)
        if blob is None:
            continue
        good = blob
        n_lit = int(np.where((rec >> np.uint64(57)) & np.uint64(1), (rec >> np.uint64(58)) & np.uint64(63), 0).sum())
        for flags in (0, 1):
            grc, grec, gused = gpu_read(blob, flags)
            hrc = block_read(blob)[0] if flags == 0 else 0
            assert grc == hrc, (n_reads, flags)
            if n_lit % 31 == 0 and not flags:
                assert grc == 7                                         # Q5: the reference panics
                continue
            orec, _, oused = ob.read_block(blob) if n_lit % 31 else (rec, None, len(blob))
            assert grc == 0 and gused == oused == len(blob) and (grec == orec).all() and (grec == rec).all()
    assert gpu_read(b"")[0] == 0 and gpu_read(b"")[2] == 0          # EOF
    assert gpu_read(good[: len(good) // 2])[0] == 7                    # truncated
    # Q5 block: refused like the reference without the flag, decoded with it
    recs = [7 | (40 << 32) | (1 << 56)] + [int(rng.integers(0, 4)) | ((2 | (1 << 2)) << 56) for _ in range(31)]
    q5 = np.array(recs, dtype=np.uint64)
    blob = ob.write_block(q5, 1)
    assert gpu_read(blob, 0)[0] == 7
    grc, grec, _ = gpu_read(blob, 1)
    assert grc == 0 and (grec == q5).all()
    # malformed header fields: same verdict as the host reader
    bs0 = struct.unpack_from("<I", good, 0)[0]
    for off, fmt, value in [(32 + bs0 + 16, "<Q", 64), (16, "<Q", 1), (8, "<I", 0xFFFFFFFF), (0, "<I", 0xFFFFFFF0)]:
        b = bytearray(good)
        struct.pack_into(fmt, b, off, value)
        assert gpu_read(bytes(b))[0] == 7 and block_read(bytes(b))[0] == 7


@pytest.mark.gpu
def test_fused_columns_encode_and_decode_match_oracle_blocks():
    """Row f1 as SURVEY 8(f)1 defines it: ntc_encode_columns (records stay in HBM, only coded words travel) + the host
    remainder ntc_block_assemble == the oracle's write_block per block, byte for byte (both link system zlib), including a
    block the reference drops (Q4); ntc_decode_columns_fasta over the same coded blocks == the FASTA text of the reads."""
    rng = np.random.default_rng(41)
    g = helpers.random_seq(rng, 50000)
    index = nt.Index.build([g], 31)
    ctx = nt.Context(index, 0)
    o = ob.OracleIndex.from_arrays(index.k, index.n_sets, index.rows(), index.lcs())
    for n_reads, rate, lens, block_reads in [(5000, 0.01, [150], 512), (70000, 0.01, [150], 65536), (300, 0.05, [150, 3000], 64), (10, 0.0, [100], 100)]:
        reads = helpers.sample_reads(rng, [g], n_reads, lens, [rate], 31)
        if n_reads == 5000:
            reads[1024: 1536] = _q4_free_reads(rng, g, 512, 31)          # block 2: reference records only
        bases, off = nt.concat_reads(reads)
        rec, roff = o.encode_reads(bases, off, fast=True, threads=8)
        blocks, words, n_rec = ctx.encode_columns(bases, off, block_reads)
        assert n_rec == len(rec) and len(blocks) == (n_reads + block_reads - 1) // block_reads
        kept_blocks, kept_reads, dropped = [], [], 0
        for b, (brec, nr) in enumerate(_blocks_of(rec, roff, block_reads)):
            want = ob.write_block(brec, nr)
            assert blocks[b].n_records == len(brec) and blocks[b].num_reads == nr
            if want is None:
                assert blocks[b].status == 10
                dropped += 1
                continue
            assert blocks[b].status == 0
            assert nt.block_assemble(blocks[b], words) == want, (n_reads, b)
            assert b"".join(nt.sub_block_assemble(blocks[b], q, words) for q in range(4)) == want
            kept_blocks.append(blocks[b])
            kept_reads += reads[b * block_reads: b * block_reads + nr]
        assert dropped == (1 if n_reads == 5000 else 0)
        n_lit_ok = all(int(sum(int(c.n_values) for c in [blk.col[3]])) >= 0 for blk in kept_blocks)
        text, n_seqs, n_dec = ctx.decode_columns_fasta(kept_blocks, words, 7, full_last_literal_word=True)
        assert n_lit_ok and n_seqs == len(kept_reads) and n_dec == sum(int(b.n_records) for b in kept_blocks)
        assert text == helpers.fasta_text(kept_reads, 7)
    # blocks read back from bytes (ntc_block_disassemble), one of them with orphan records in front of its first flag
    a = np.array([5 | (20 << 32), 7 | (40 << 32) | (1 << 56), (0b11 | ((2 | (1 << 2)) << 56))], dtype=np.uint64)
    b = np.array([9 | (25 << 32) | (1 << 56), (0b10 | ((2 | (1 << 2)) << 56))], dtype=np.uint64)
    _, words, blks = _cols_of_blocks([ob.write_block(b, 1), ob.write_block(a, 1), ob.write_block(b, 1)])
    text, n_seqs, _ = ctx.decode_columns_fasta(blks, words, 1)
    sb = b"".join(o.decode(b, fast=True)[0].tobytes() for _ in range(1))
    sa = o.decode(a, fast=True)[0].tobytes()
    assert n_seqs == 3 and text == b">seq.1\n" + sb + b"\n>seq.2\n" + sa + b"\n>seq.3\n" + sb + b"\n"


def _blobs_with_crafted_len_column(b=3):
    """One block whose match-length column (Rice, sub-block 1) is written by hand: (records, intact blob, [blobs whose LAST
    Rice code has its terminating one bit but not all of its b remainder bits])."""
    lens = [40, 13, 150, 91, 12, 77]
    rec = [(1000 + i) | (L << 32) | (int(i == 0) << 56) for i, L in enumerate(lens)]
    rec += [0b1101 | (((0 + 2) | (2 << 2)) << 56)]                                          # one literal: every column non-empty
    rec = np.array(rec, dtype=np.uint64)
    rc, blob = block_write(rec, 1)
    assert rc == 0
    blk, words, used = nt.block_disassemble(blob)
    assert used == len(blob) and blk.col[1].n_values == len(lens)
    code = lambda v: "0" * (v >> b) + "1" + format(v & ((1 << b) - 1), "0%db" % b)

    def with_bits(bits):
        assert len(bits) % 64 == 0
        col1 = np.packbits(np.array([int(c) for c in bits], dtype=np.uint8)).view(np.uint64)   # stream bytes in order = the file's words
        parts = [words[int(blk.col[q].word_offset): int(blk.col[q].word_offset + blk.col[q].n_words)] for q in range(4)]
        parts[1] = col1
        nb = nt.CodedBlock.from_buffer_copy(blk)
        pos = 0
        for q in range(4):
            nb.col[q].word_offset, nb.col[q].n_words = pos, len(parts[q])
            pos += len(parts[q])
        nb.col[1].param = b
        return nt.block_assemble(nb, np.concatenate(parts))

    whole = "".join(code(v) for v in lens)
    good = with_bits(whole + "0" * (-len(whole) % 64))
    head = "".join(code(v) for v in lens[:-1])
    bad = []
    for missing in range(1, b + 1):                                                         # 1 .. b remainder bits cut off
        tail_zeros = -(len(head) + 1 + (b - missing)) % 64
        if tail_zeros > (lens[-1] >> b) + 64:
            continue
        bits = head + "0" * tail_zeros + "1" + "0" * (b - missing)
        bad.append(with_bits(bits))
    return rec, good, bad


def test_truncated_final_rice_code_is_rejected_by_the_host_reader():
    """a length column whose last code ends in the middle of its remainder bits is not a column (dsi-bitstream's
    reader runs out of words, src/decode.rs:51-62).  The hand-written intact column decodes to the block's records."""
    rec, good, bad = _blobs_with_crafted_len_column()
    rc, got, used = block_read(good)
    assert rc == 0 and used == len(good) and (got == rec).all()
    assert len(bad) == 3
    for blob in bad:
        assert block_read(blob)[0] == 7


@pytest.mark.gpu
def test_truncated_final_rice_code_same_verdict_on_gpu_and_host():
    """The GPU column decoder used to count a Rice code whose remainder bits lie beyond the stream (read as zero padding), so
    ntc_block_read_gpu accepted what the host reader refuses.  Same verdict now, and the intact column gives the same records."""
    rng = np.random.default_rng(3)
    index = nt.Index.build([helpers.random_seq(rng, 5000)], 31)
    ctx = nt.Context(index, 0)

    def gpu_read(blob):
        buf = np.frombuffer(blob, dtype=np.uint8)
        out = np.zeros(8 * len(buf) + 1024, dtype=np.uint64)
        n, used = C.c_uint64(0), C.c_uint64(0)
        rc = _lib.lib().ntc_block_read_gpu(ctx._h, _ptr(buf), len(buf), _ptr(out), len(out), C.byref(n), C.byref(used), 0)
        return rc, out[: n.value].copy(), used.value

    rec, good, bad = _blobs_with_crafted_len_column()
    rc, got, used = gpu_read(good)
    assert rc == 0 and used == len(good) and (got == rec).all()
    for blob in bad:
        assert block_read(blob)[0] == 7 and gpu_read(blob)[0] == 7
