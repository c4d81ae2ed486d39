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


def test_gpu_column_codec_bodies_on_cpu_emulation():
    """column_kernels.cuh (split + prefix-summed bit packing) run on the CPU == sequential coders"""
    from emul_binding import lib as emul_lib
    L = emul_lib()
    L.emul_block_columns.restype = C.c_int
    L.emul_block_columns.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]
    rng = np.random.default_rng(3)
    g = helpers.random_seq(rng, 8000)
    o = ob.OracleIndex.build([g], 20, True)
    for n_reads, rate, lens in [(400, 0.03, [150]), (50, 0.2, [40, 400]), (800, 0.01, [150, 2000])]:
        reads = helpers.sample_reads(rng, [g], n_reads, lens, [rate], 20)
        rec, _ = o.encode_reads(*ob.concat_reads(reads), fast=True, threads=4)
        words = np.zeros(4 * len(rec) + 4096, dtype=np.uint64)
        nw = np.zeros(4, dtype=np.uint64)
        params = np.zeros(4, dtype=np.uint64)
        rc = L.emul_block_columns(_ptr(rec), len(rec), _ptr(words), len(words), _ptr(nw), _ptr(params))
        assert rc == 0
        pos = 0
        for q, (w, p) in enumerate(_columns_of(rec)):
            assert int(params[q]) == p and int(nw[q]) == len(w), q
            assert (words[pos: pos + len(w)] == w).all(), q
            pos += len(w)


@pytest.mark.gpu
def test_block_write_gpu_is_byte_identical_to_host():
    rng = np.random.default_rng(4)
    g = helpers.random_seq(rng, 30000)
    idx = nt.Index.build([g], 31)
    ctx = nt.Context(idx, 0)
    for n_reads, rate, lens in [(3000, 0.01, [150]), (65536, 0.01, [150]), (200, 0.3, [60, 1000])]:
        reads = helpers.sample_reads(rng, [g], n_reads, lens, [rate], 31)
        rec, _ = ctx.encode_batch(*nt.concat_reads(reads))
        rc, blob = block_write(rec, n_reads)
        cap = int(_lib.lib().ntc_block_bound(len(rec)))
        out = np.zeros(cap, dtype=np.uint8)
        nb = C.c_uint64(0)
        rc2 = _lib.lib().ntc_block_write_gpu(ctx._h, _ptr(rec), len(rec), n_reads, _ptr(out), cap, C.byref(nb))
        assert rc2 == rc
        if rc == 0:
            assert out[: nb.value].tobytes() == blob
            rc3, rec2, used = block_read(blob)
            n_lit = int((np.where(((rec >> np.uint64(56)) & np.uint64(2)) != 0, (rec >> np.uint64(58)) & np.uint64(63), 0)).sum())
            if n_lit % 31:
                assert rc3 == 0 and (rec2 == rec).all()
    only_ref = rec[((rec >> np.uint64(56)) & np.uint64(2)) == 0]
    nb = C.c_uint64(0)
    assert _lib.lib().ntc_block_write_gpu(ctx._h, _ptr(only_ref), len(only_ref), 1, _ptr(out), cap, C.byref(nb)) == 10


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


def test_gpu_zip_body_on_cpu_emulation():
    from emul_binding import lib as emul_lib
    L = emul_lib()
    L.emul_zip.restype = C.c_uint32
    L.emul_zip.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_int, C.c_void_p]
    rng = np.random.default_rng(19)
    g = helpers.random_seq(rng, 8000)
    o = ob.OracleIndex.build([g], 20, True)
    for n_reads, rate in [(300, 0.03), (60, 0.2), (500, 0.01)]:
        reads = helpers.sample_reads(rng, [g], n_reads, [150, 35, 400], [rate], 20)
        rec, _ = o.encode_reads(*ob.concat_reads(reads), fast=True, threads=4)
        cols = [np.ascontiguousarray(c, dtype=np.uint64) for c in ob.split(rec)]
        n_lit = int(sum((int(f) >> 2) for f in cols[2] if int(f) & 2))
        out = np.zeros(len(rec), dtype=np.uint64)
        for full in (0, 1):
            err = L.emul_zip(_ptr(cols[0]), _ptr(cols[1]), len(cols[0]), _ptr(cols[2]), len(cols[2]), _ptr(cols[3]), len(cols[3]), full, _ptr(out))
            if n_lit % 31 == 0 and n_lit and not full:
                assert err == 2                                         # Q5
            else:
                assert err == 0 and (out == rec).all()
    # more reference records than colex values / a literal longer than 32 bases
    flags = np.array([1, 0, 0], dtype=np.uint64)
    one = np.array([5], dtype=np.uint64)
    assert L.emul_zip(_ptr(one), _ptr(one), 1, _ptr(flags), 3, _ptr(one), 1, 0, _ptr(np.zeros(3, dtype=np.uint64))) & 1
    flags = np.array([(33 << 2) | 3], dtype=np.uint64)
    assert L.emul_zip(_ptr(one), _ptr(one), 1, _ptr(flags), 1, _ptr(one), 1, 0, _ptr(np.zeros(1, dtype=np.uint64))) & 4


@pytest.mark.gpu
def test_block_read_gpu_matches_host_reader():
    """ntc_block_read_gpu (integer decoders + zip on the GPU) == ntc_block_read_ex on the same bytes, errors included"""
    import ntcomp_b200 as nt
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
        rc, blob = block_write(rec, n_reads)
        if rc == 10:
            continue
        good = blob
        for flags in (0, 1):
            hrc, hrec, hused = block_read(blob) if flags == 0 else (0,) + nt.block_read(blob, True)
            grc, grec, gused = gpu_read(blob, flags)
            assert grc == hrc, (n_reads, flags)
            if hrc == 0:
                assert gused == hused == len(blob) and (grec == hrec).all() and (grec == rec).all()
    assert gpu_read(b"")[0] == 0 and gpu_read(b"")[2] == 0          # EOF
    assert gpu_read(good[: len(good) // 2])[0] == 7                    # truncated
    # Q5 block: refused like the reference without the flag, decoded with it
    recs = [7 | (40 << 32) | (1 << 56)] + [int(rng.integers(0, 4)) | ((2 | (1 << 2)) << 56) for _ in range(31)]
    q5 = np.array(recs, dtype=np.uint64)
    rc, blob = block_write(q5, 1)
    assert rc == 0 and gpu_read(blob, 0)[0] == 7
    grc, grec, _ = gpu_read(blob, 1)
    assert grc == 0 and (grec == q5).all()
