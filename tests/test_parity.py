"""Parity tests against the CPU oracle on the same inputs.  Bar: bit-exact (integer / byte work).
P1 record stream, P3 decoded bytes + FASTA, P4 cross decode.

Every case runs on two backends:
  gpu   -- the CUDA path through the C ABI (ntcomp_b200.Context), marked `gpu`: the parity gate proper
  emul  -- the same per-thread kernel bodies executed on the CPU (tests/emul), so kernel logic is
           checked in the "not gpu" suite too.  The emulation is test infrastructure, not a fallback.
"""
import numpy as np
import pytest

import helpers
import ntcomp_b200 as nt
from ntcomp_b200 import synth
from oracle import binding as ob

BACKENDS = [pytest.param("emul"), pytest.param("gpu", marks=pytest.mark.gpu)]


def make_ctx(backend, index):
    if backend == "gpu":
        return nt.Context(index, 0)
    from emul_binding import EmulContext
    return EmulContext(index)


class Case:
    def __init__(self, backend, contigs, k, add_revcomp=True):
        self.backend = backend
        self.contigs = contigs
        self.k = k
        self.index = nt.Index.build(contigs, k, add_revcomp)
        self.oracle = ob.OracleIndex.from_arrays(k, self.index.n_sets, self.index.rows(), self.index.lcs())
        self.ctx = make_ctx(backend, self.index)


@pytest.fixture(scope="module", params=BACKENDS)
def backend(request):
    return request.param


@pytest.fixture(scope="module")
def small_cases(backend):
    rng = np.random.default_rng(11)
    cases = []
    for k, sizes in [(12, [400, 90]), (16, [3000]), (31, [20000, 500, 31, 30]), (91, [30000, 2000]), (255, [1500, 900, 300])]:
        cases.append(Case(backend, [helpers.random_seq(rng, n) for n in sizes], k))
    # skewed index: almost no T, so a T in a read forces contract_left down to very short contexts
    # (LCS runs longer than 255 nodes: the saturated-distance fallback of ms_kernel, and t == 0)
    g = helpers.ALPHA[rng.integers(0, 3, size=30000)]
    g[[5000, 12000, 12001, 20000, 29000]] = ord("T")
    cases.append(Case(backend, [g], 16, add_revcomp=False))
    return cases


def _reads_for(rng, case, n):
    k = case.k
    reads = helpers.sample_reads(rng, [c for c in case.contigs if len(c) > 0], n,
                                 lengths=[1, 2, 5, k - 1, k, k + 1, k + 2, 2 * k, 3 * k + 7, 150, 150, 150, 400, 1000],
                                 error_rates=[0, 0, 0.01, 0.01, 0.05, 0.3], k=k)
    reads += [np.ascontiguousarray(c) for c in case.contigs if len(c) > 0]            # the contigs themselves (C1 style)
    reads += [helpers.random_seq(rng, 200), helpers.random_seq(rng, 7)]                # unrelated sequence
    return reads


def test_matching_statistics_parity(small_cases):
    rng = np.random.default_rng(5)
    for case in small_cases:
        reads = _reads_for(rng, case, 200)
        bases, off = nt.concat_reads(reads)
        d, s = case.ctx.matching_statistics(bases, off)
        for i, r in enumerate(reads):
            od, os_, _ = case.oracle.matching_statistics(r)
            lo, hi = int(off[i]), int(off[i + 1])
            assert (d[lo:hi] == od).all(), (case.k, i)
            assert (s[lo:hi] == os_).all(), (case.k, i)


def test_encode_records_parity_literal_oracle(small_cases):
    """P1 against the literal restatement (4 x search left extension, access_kmer literals)."""
    rng = np.random.default_rng(6)
    for case in small_cases:
        reads = _reads_for(rng, case, 150)
        bases, off = nt.concat_reads(reads)
        rec, roff = case.ctx.encode_batch(bases, off)
        orec, oroff = case.oracle.encode_reads(bases, off, fast=False, threads=8)
        assert (roff == oroff).all(), case.k
        assert (rec == orec).all(), case.k


def test_rank_blocks_of_64_nodes(backend, monkeypatch):
    """Indexes of 20 M nodes or more keep their rank blocks as 64 nodes per 64 bytes (layout.h: one 256-bit load per probe, half the
    bytes); NTC_RANK_NODES=64 selects that layout for small indexes.  Same matching statistics, records and round trips as the
    oracle: reads across block borders, intervals whose ends lie in different blocks, the prefix table built from these blocks."""
    monkeypatch.setenv("NTC_RANK_NODES", "64")
    rng = np.random.default_rng(64)
    for k, sizes in [(12, [700, 90]), (31, [20000, 500]), (91, [30000]), (255, [1500, 900])]:
        case = Case(backend, [helpers.random_seq(rng, n) for n in sizes], k)
        reads = _reads_for(rng, case, 60)
        bases, off = nt.concat_reads(reads)
        d, s = case.ctx.matching_statistics(bases, off)
        for i, r in enumerate(reads[:25]):
            od, os_, _ = case.oracle.matching_statistics(r)
            lo, hi = int(off[i]), int(off[i + 1])
            assert (d[lo:hi] == od).all() and (s[lo:hi] == os_).all(), (k, i)
        for T in (0, 256):
            case.ctx.set_tile_len(T)
            rec, roff = case.ctx.encode_batch(bases, off)
            orec, oroff = case.oracle.encode_reads(bases, off, fast=True, threads=8)
            assert (roff == oroff).all() and (rec == orec).all(), (k, T)
        case.ctx.set_tile_len(0)
        db, doff = case.ctx.decode_batch(rec)
        assert (doff == off).all() and (db == bases).all()


def test_long_read_segment_path(backend):
    """Reads longer than the tile length are segmented in 512-position segments (seg_spec / seg_join / seg_emit in
    encode_kernels.cuh): speculative walks per segment joined by the true walk.  Noisy long reads (walks merge at every
    mismatch), error-free copies of a 30 kbp contig (one record spanning many segments: the speculative walks give up at
    kSpecExtCap = 4096 and the true walk takes the long step), a long match broken once, read lengths around multiples of 512,
    short reads in the same batch (the one-item path), capacity errors."""
    rng = np.random.default_rng(77)
    g = helpers.random_seq(rng, 30000)
    case = Case(backend, [g, helpers.random_seq(rng, 700)], 31)
    reads = helpers.sample_reads(rng, [g], 40, lengths=[511, 512, 513, 1023, 1024, 1025, 1536, 4096, 4097, 4608, 9000, 150, 40],
                                 error_rates=[0, 0.002, 0.01, 0.05, 0.2], k=31)
    broken = g[2000:22000].copy()
    broken[10000] = ord("A") if broken[10000] != ord("A") else ord("C")
    reads += [g.copy(), g[100:9000].copy(), g[5:4700].copy(), broken, helpers.random_seq(rng, 3000), g[:600].copy()]
    bases, off = nt.concat_reads(reads)
    for T in (256, 512, 1000, 0):
        case.ctx.set_tile_len(T)
        rec, roff = case.ctx.encode_batch(bases, off)
        orec, oroff = case.oracle.encode_reads(bases, off, fast=True, threads=8)
        assert (roff == oroff).all() and (rec == orec).all(), T
        db, doff = case.ctx.decode_batch(rec)
        assert (doff == off).all() and (db == bases).all()
    case.ctx.set_tile_len(256)
    lrec, _ = case.oracle.encode_reads(bases, off, fast=False, threads=8)
    assert (lrec == rec).all()
    with pytest.raises(nt.NtcError) as e:
        case.ctx.encode_batch(bases, off, cap=len(rec) - 1)
    assert e.value.code == 3
    case.ctx.set_tile_len(0)


def test_tile_length_invariance(small_cases):
    """Long reads are cut into tiles with a k-1 base warm-up (ms_thread); matching statistics, record
    streams and round trips must not depend on the tile length.  Reads of exactly T, T+1, 2T-1, 2T, ...
    bases, error-free contigs (unique-predecessor runs spanning many tiles) and noisy long reads."""
    rng = np.random.default_rng(61)
    for case in small_cases:
        src = [c for c in case.contigs if len(c) > 0]
        for T in (256, 260, 1024):
            case.ctx.set_tile_len(T)
            reads = helpers.sample_reads(rng, src, 30, lengths=[T - 1, T, T + 1, 2 * T - 1, 2 * T, 2 * T + 1, 3 * T + 5, 5000, 150, 1],
                                         error_rates=[0, 0.01, 0.05, 0.2], k=case.k)
            reads += [np.ascontiguousarray(c) for c in src] + [helpers.random_seq(rng, 3 * T + 2)]
            bases, off = nt.concat_reads(reads)
            d, s = case.ctx.matching_statistics(bases, off)
            for i, r in enumerate(reads):
                od, os_, _ = case.oracle.matching_statistics(r)
                lo, hi = int(off[i]), int(off[i + 1])
                assert (d[lo:hi] == od).all() and (s[lo:hi] == os_).all(), (case.k, T, i, len(r))
            rec, roff = case.ctx.encode_batch(bases, off)
            orec, oroff = case.oracle.encode_reads(bases, off, fast=True, threads=8)
            assert (roff == oroff).all() and (rec == orec).all(), (case.k, T)
            db, doff = case.ctx.decode_batch(rec)
            assert (doff == off).all() and (db == bases).all()
        case.ctx.set_tile_len(0)
        lrec, _ = case.oracle.encode_reads(bases, off, fast=False, threads=8)   # literal restatement on the last batch
        assert (lrec == rec).all()


def test_encode_sequence_triples_and_dictionary(small_cases):
    rng = np.random.default_rng(7)
    for case in small_cases:
        reads = _reads_for(rng, case, 40)
        for r in reads:
            got = nt.encode_sequence(r, case.ctx)
            want = case.oracle.encode_sequence(r, fast=False)
            assert [(ln, (rg.start, rg.stop)) for ln, rg in got] == want, case.k
            words = nt.encode_dictionary(got, case.ctx)
            assert (words == case.oracle.encode_dictionary(want)).all()


def test_decode_parity_and_roundtrip(small_cases):
    rng = np.random.default_rng(8)
    for case in small_cases:
        reads = _reads_for(rng, case, 150)
        bases, off = nt.concat_reads(reads)
        rec, _ = case.ctx.encode_batch(bases, off)
        db, doff = case.ctx.decode_batch(rec)
        ob_, ooff = case.oracle.decode(rec, fast=False)          # literal decode_sequence
        assert (doff == ooff).all() and (db == ob_).all(), case.k
        assert (doff == off).all() and (db == bases).all(), case.k
        seqs = nt.decode_sequence(rec, case.ctx)
        assert seqs == [bytes(r) for r in reads]
        # P4: oracle-encoded stream through the GPU decoder
        orec, _ = case.oracle.encode_reads(bases, off, fast=True, threads=8)
        db2, doff2 = case.ctx.decode_batch(orec)
        assert (db2 == bases).all() and (doff2 == off).all()


def test_decode_fasta_text(small_cases):
    rng = np.random.default_rng(9)
    case = small_cases[2]
    reads = _reads_for(rng, case, 120)
    bases, off = nt.concat_reads(reads)
    rec, _ = case.ctx.encode_batch(bases, off)
    for first in (1, 9, 98, 65530, 99999999):
        text, n_seqs = case.ctx.decode_fasta(rec, first)
        assert n_seqs == len(reads)
        assert text == helpers.fasta_text(reads, first)


def test_decode_drops_records_before_first_flag(small_cases):
    case = small_cases[2]
    rng = np.random.default_rng(10)
    reads = _reads_for(rng, case, 20)
    bases, off = nt.concat_reads(reads)
    rec, roff = case.ctx.encode_batch(bases, off)
    cut = rec[int(roff[3]) + 1:]                     # starts in the middle of read 3's records
    db, doff = case.ctx.decode_batch(cut)
    ob_, ooff = case.oracle.decode(cut, fast=False)
    assert (doff == ooff).all() and (db == ob_).all()
    assert len(doff) - 1 == len(reads) - 4
    empty_b, empty_off = case.ctx.decode_batch(np.zeros(0, dtype=np.uint64))
    assert len(empty_b) == 0 and list(empty_off) == [0]


def test_decode_handcrafted_literals_up_to_32_bases(small_cases):
    """literal records longer than the encoder's 11 (flag length field is 6 bits; from_2bit takes <= 32)"""
    case = small_cases[2]
    rng = np.random.default_rng(21)
    recs = []
    for L in list(range(1, 29)) + [33, 40, 63]:          # 7 payload bytes hold 28 bases; > 32 decodes to nothing (lib.rs:303)
        bits = int(rng.integers(0, 1 << 56))
        bits &= (1 << (2 * min(L, 28))) - 1
        first = int(rng.integers(0, 2)) if recs else 1
        recs.append(bits | (((first + 2) | (L << 2)) << 56))
    good = bytes(case.contigs[0][100:180])
    ref_rec, _ = case.oracle.encode_reads(*nt.concat_reads([good]), fast=True)
    stream = np.array(recs + list(ref_rec) + recs[::-1], dtype=np.uint64)
    stream[len(recs)] |= np.uint64(1 << 56)
    db, doff = case.ctx.decode_batch(stream)
    ob_, ooff = case.oracle.decode(stream, fast=False)
    assert (doff == ooff).all() and (db == ob_).all()
    text, n = case.ctx.decode_fasta(stream, 7)
    seqs = [bytes(ob_[int(ooff[i]): int(ooff[i + 1])]) for i in range(len(ooff) - 1)]
    assert text == helpers.fasta_text(seqs, 7) and n == len(seqs)


def test_fixture_c1_shape_roundtrip(backend):
    """BASELINE config #1: tests/fasta_data.rs shape (7 contigs < 2000 bp, k = 255, block size 3)."""
    contigs, k, block = helpers.fixture_c1()
    case = Case(backend, contigs, k)
    decoded = []
    for b in range(0, len(contigs), block):
        bases, off = nt.concat_reads(contigs[b:b + block])
        rec, _ = case.ctx.encode_batch(bases, off)
        orec, _ = case.oracle.encode_reads(bases, off, fast=False)
        assert (rec == orec).all()
        blob = ob.write_block(rec, min(block, len(contigs) - b))   # host container stays the reference's (oracle restatement here)
        if blob is None:
            decoded += nt.decode_sequence(rec, case.ctx)            # Q4: block the reference would drop
            continue
        rec2, _, _ = ob.read_block(blob)
        assert (rec2 == rec).all()
        decoded += nt.decode_sequence(rec2, case.ctx)
    assert decoded == [bytes(c) for c in contigs]                    # fasta_data.rs:93,99
    assert sum(map(len, decoded)) == sum(map(len, contigs))          # fasta_data.rs:100


def test_error_codes(small_cases):
    case = small_cases[2]
    good = bytes(case.contigs[0][:100])
    with pytest.raises(nt.EncodeError):
        nt.encode_sequence(b"", case.ctx)                            # lib.rs:224
    with pytest.raises(nt.NtcError) as e:
        case.ctx.encode_batch(*nt.concat_reads([good, b"ACGTNACGT"]))
    assert e.value.code == 5                                         # NTC_ERR_BAD_BASE (Q3)
    with pytest.raises(nt.NtcError) as e:
        case.ctx.encode_batch(*nt.concat_reads([good, b"", good]))
    assert e.value.code == 4
    bases, off = nt.concat_reads([good] * 8)
    with pytest.raises(nt.NtcError) as e:
        case.ctx.encode_batch(bases, off, cap=3)
    assert e.value.code == 3
    # the context stays usable after an error
    rec, _ = case.ctx.encode_batch(bases, off)
    assert nt.decode_sequence(rec, case.ctx) == [good] * 8
    bad = np.array([(case.index.n_sets + 5) | (20 << 32) | (1 << 56)], dtype=np.uint64)
    with pytest.raises(nt.NtcError) as e:
        case.ctx.decode_batch(bad)
    assert e.value.code == 8


def test_index_with_missing_base_flags_absent_base(backend):
    """d == 0: the reference underflows at lib.rs:207; the ABI reports NTC_ERR_BAD_BASE."""
    contig = np.frombuffer(b"ACACACCACAACCCACACAAACCACACCA" * 4, dtype=np.uint8)
    case = Case(backend, [contig], 12, add_revcomp=False)                    # index holds only A and C
    with pytest.raises(nt.NtcError) as e:
        case.ctx.encode_batch(*nt.concat_reads([b"ACACACCACAACCCAGACAAACC"]))
    assert e.value.code == 5
    with pytest.raises(ob.OracleError) as oe:
        case.oracle.encode_sequence(b"ACACACCACAACCCAGACAAACC")
    assert oe.value.code == -3


def _decode_with_cap(ctx, rec, cap):
    import ctypes as C
    from ntcomp_b200 import _lib
    rec = np.ascontiguousarray(rec, dtype=np.uint64)
    out = np.zeros(cap + 64, dtype=np.uint8)
    soff = np.zeros(len(rec) + 1, dtype=np.uint64)
    n = C.c_uint64(0)
    _lib.check(_lib.lib().ntc_decode_batch(ctx._h, rec.ctypes.data_as(C.c_void_p), len(rec), out.ctypes.data_as(C.c_void_p), cap,
                                           soff.ctypes.data_as(C.c_void_p), len(rec), C.byref(n)))
    return out, soff, n.value


@pytest.mark.gpu
def test_chunked_host_path_matches_unchunked(monkeypatch, small_cases):
    if small_cases[0].backend != "gpu":
        pytest.skip("host chunking lives in the C ABI")
    case0 = small_cases[3]
    rng = np.random.default_rng(12)
    reads = _reads_for(rng, case0, 300)
    bases, off = nt.concat_reads(reads)
    rec, roff = case0.ctx.encode_batch(bases, off)
    monkeypatch.setenv("NTC_MAX_CHUNK_BASES", "4096")
    monkeypatch.setenv("NTC_MAX_CHUNK_RECORDS", "64")           # decode: many chunks, alternating lanes
    ctx2 = nt.Context(case0.index, 0)
    rec2, roff2 = ctx2.encode_batch(bases, off)
    assert (rec == rec2).all() and (roff == roff2).all()
    tri, _ = case0.ctx.encode_sequence_batch(bases, off)
    tri2, _ = ctx2.encode_sequence_batch(bases, off)
    assert (tri == tri2).all()
    # chunked decode (chunks start on records that open a sequence; numbering / offsets carried across) == input
    db, doff = ctx2.decode_batch(rec2)
    assert (db == bases).all() and (doff == off).all()
    for first in (1, 9995, 12345):                               # digit-count changes inside a chunk and across chunks
        text, n = ctx2.decode_fasta(rec2, first)
        assert text == helpers.fasta_text(reads, first) and n == len(reads)
    # leading records without a `first` flag are dropped (src/lib.rs:308-317), also when they fill whole chunks
    tail = rec[int(roff[len(reads) - 3]) + 1:]                   # starts inside the last contig-sized read: > 64 orphan records
    db2, doff2 = ctx2.decode_batch(tail)
    db3, doff3 = case0.ctx.decode_batch(tail)
    assert (db2 == db3).all() and (doff2 == doff3).all()
    with pytest.raises(nt.NtcError) as e:                        # capacity error in a later chunk
        _decode_with_cap(ctx2, rec2, len(bases) - 5)
    assert e.value.code == 3
    with pytest.raises(nt.NtcError) as e:                        # an error in a middle chunk drains both lanes cleanly
        ctx2.encode_batch(*nt.concat_reads(reads[:200] + [b"ACGTNNNN"] + reads[200:]))
    assert e.value.code == 5
    rec3, roff3 = ctx2.encode_batch(bases, off)
    assert (rec3 == rec).all() and (roff3 == roff).all()
    # one PCIe direction at a time (ntc_ctx_set_copy_directions): records parked in HBM until the input is up, all records
    # up before any text comes down -- same results, chunked and unchunked, including the capacity error
    for c in (ctx2, case0.ctx):
        c.set_copy_directions(True)
        try:
            rec4, roff4 = c.encode_batch(bases, off)
            assert (rec4 == rec).all() and (roff4 == roff).all()
            tri4, _ = c.encode_sequence_batch(bases, off)
            assert (tri4 == tri).all()
            db4, doff4 = c.decode_batch(rec4)
            assert (db4 == bases).all() and (doff4 == off).all()
            text, n = c.decode_fasta(rec4, 9995)
            assert text == helpers.fasta_text(reads, 9995) and n == len(reads)
            with pytest.raises(nt.NtcError) as e:
                _decode_with_cap(c, rec4, len(bases) - 5)
            assert e.value.code == 3
        finally:
            c.set_copy_directions(False)
    # who packs the reads to 2 bits (ntc_ctx_set_host_threads): host threads before the copy (1, 3 and 16 of them; partial last words,
    # chunks of a few thousand bases) or pack2bit_kernel after it -- same records, the same error for a non-ACGT byte
    for c in (ctx2, case0.ctx):
        for threads in (0, 1, 3, 16):
            c.set_host_threads(threads)
            rec5, roff5 = c.encode_batch(bases, off)
            assert (rec5 == rec).all() and (roff5 == roff).all(), threads
            tri5, _ = c.encode_sequence_batch(bases, off)
            assert (tri5 == tri).all()
            for bad in (b"ACGTNNNN", b"ACGTacgt", b"ACGU" * 20):
                with pytest.raises(nt.NtcError) as e:
                    c.encode_batch(*nt.concat_reads(reads[:200] + [bad] + reads[200:]))
                assert e.value.code == 5, (threads, bad)
            rec5, roff5 = c.encode_batch(bases, off)
            assert (rec5 == rec).all() and (roff5 == roff).all(), threads
        c.set_host_threads(-1)


@pytest.mark.parametrize("cfg", ["C2-small", "C3-small", "C5-small"])
def test_synthetic_config_shapes(cfg, backend):
    """BASELINE configs at reduced genome/read counts: oracle parity on a sample + full round trip."""
    shape = {"C2-small": dict(genome=200_000, k=31, n_reads=20_000, read_len=150, err_ppm=10_000),
             "C3-small": dict(genome=200_000, k=91, n_reads=20_000, read_len=150, err_ppm=10_000),
             "C5-small": dict(genome=200_000, k=91, n_reads=60, read_len=10_000, err_ppm=50_000)}[cfg]
    if backend == "emul":                      # the CPU emulation is ~1000x slower: shrink the batch
        shape = dict(shape, n_reads=max(8, shape["n_reads"] // 40))
    g = synth.genome(shape["genome"])
    case = Case(backend, [g], shape["k"])
    bases, off = synth.reads(g, 777, 0, shape["n_reads"], shape["read_len"], shape["err_ppm"])
    rec, roff = case.ctx.encode_batch(bases, off)
    n_check = min(shape["n_reads"], 3000)
    orec, oroff = case.oracle.encode_reads(bases[: int(off[n_check])], off[: n_check + 1], fast=True, threads=8)
    assert (roff[: n_check + 1] == oroff).all()
    assert (rec[: int(roff[n_check])] == orec).all()
    n_lit = min(shape["n_reads"], 40)
    lrec, _ = case.oracle.encode_reads(bases[: int(off[n_lit])], off[: n_lit + 1], fast=False, threads=8)
    assert (rec[: int(roff[n_lit])] == lrec).all()
    db, doff = case.ctx.decode_batch(rec)
    assert (doff == off).all() and (db == bases).all()
    text, n_seqs = case.ctx.decode_fasta(rec, 1)
    assert n_seqs == shape["n_reads"]
    head = helpers.fasta_text([bases[int(off[i]): int(off[i + 1])] for i in range(min(50, shape["n_reads"]))], 1)
    assert text[: len(head)] == head


@pytest.mark.gpu
def test_device_resident_api_with_torch(small_cases):
    if small_cases[0].backend != "gpu":
        pytest.skip("device pointers need a GPU")
    import torch
    case = small_cases[3]
    rng = np.random.default_rng(13)
    reads = _reads_for(rng, case, 200)
    bases, off = nt.concat_reads(reads)
    dev = torch.device("cuda:0")
    d_bases = torch.from_numpy(bases).to(dev)
    d_off = torch.from_numpy(off.astype(np.int64)).to(dev)
    cap = len(bases)
    d_rec = torch.zeros(cap, dtype=torch.int64, device=dev)
    d_roff = torch.zeros(len(reads) + 1, dtype=torch.int64, device=dev)
    d_cnt = torch.zeros(4, dtype=torch.int64, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    case.ctx.encode_device(d_bases.data_ptr(), d_off.data_ptr(), len(reads), len(bases), d_rec.data_ptr(), cap, d_roff.data_ptr(),
                           d_cnt.data_ptr(), stream)
    case.ctx.status(stream)
    n_rec = int(d_cnt[0].item())
    rec = d_rec[:n_rec].cpu().numpy().view(np.uint64)
    orec, oroff = case.oracle.encode_reads(bases, off, fast=True, threads=8)
    assert (rec == orec).all() and (d_roff.cpu().numpy().view(np.uint64) == oroff).all()
    d_out = torch.zeros(len(bases) + 64, dtype=torch.uint8, device=dev)
    d_soff = torch.zeros(len(reads) + 1, dtype=torch.int64, device=dev)
    case.ctx.decode_device(d_rec.data_ptr(), n_rec, 1, False, d_out.data_ptr(), len(bases), d_soff.data_ptr(), len(reads), d_cnt.data_ptr(), stream)
    case.ctx.status(stream)
    assert int(d_cnt[0].item()) == len(reads) and int(d_cnt[1].item()) == len(bases)
    assert (d_out[: len(bases)].cpu().numpy() == bases).all()
    assert (d_soff.cpu().numpy().view(np.uint64) == off).all()
    want = helpers.fasta_text(reads, 99)
    d_text = torch.zeros(len(want) + 64, dtype=torch.uint8, device=dev)
    for mis in (1, 3, 5, 6, 8, 9, 12, 15):                # caller-misaligned output buffer (the writer pairs words into 16-byte stores)
        d_text.zero_()
        case.ctx.decode_device(d_rec.data_ptr(), n_rec, 99, True, d_text.data_ptr() + mis, len(want), 0, 0, d_cnt.data_ptr(), stream)
        case.ctx.status(stream)
        assert int(d_cnt[1].item()) == len(want)
        assert bytes(d_text[mis: mis + len(want)].cpu().numpy()) == want
        assert int(d_text[:mis].sum().item()) == 0 and int(d_text[mis + len(want):].sum().item()) == 0   # nothing outside the range


@pytest.mark.gpu
def test_config_c2_full_size_round_trip_and_sampled_parity():
    """BASELINE config C2 at full size (5 Mbp genome, k = 31, 1 M x 150 bp, 1 % substitutions): the
    whole batch must round-trip bit-exactly (size-independent property), the record stream of a sample
    must equal the oracle's, and shard-wise encoding must concatenate to the single-batch stream."""
    from ntcomp_b200 import sharding
    cfg = synth.CONFIGS["C2"]
    g = synth.genome(cfg["genome"])
    index = nt.Index.build([g], cfg["k"])
    ctx = nt.Context(index, 0)
    bases, off = synth.reads(g, cfg["seed"], 0, cfg["n_reads"], cfg["read_len"], cfg["err_ppm"])
    rec, roff = ctx.encode_batch(bases, off)
    assert int(roff[-1]) == len(rec)
    db, doff = ctx.decode_batch(rec)
    assert (doff == off).all() and (db == bases).all()
    text, n_seqs = ctx.decode_fasta(rec, 1)
    header_bytes = sum((6 + d + 1) * c for d, c in [(1, 9), (2, 90), (3, 900), (4, 9000), (5, 90000), (6, 900000), (7, 1)])
    assert n_seqs == cfg["n_reads"] == 1_000_000 and len(text) == len(bases) + header_bytes
    assert text.endswith(bytes(bases[-150:]) + b"\n") and text.startswith(b">seq.1\n" + bytes(bases[:150]) + b"\n>seq.2\n")
    oracle = ob.OracleIndex.from_arrays(index.k, index.n_sets, index.rows(), index.lcs())
    for lo in (0, 500_000, 999_000):
        hi = lo + 1000
        orec, _ = oracle.encode_reads(bases[int(off[lo]): int(off[hi])], off[lo: hi + 1] - off[lo], fast=True, threads=8)
        assert (rec[int(roff[lo]): int(roff[hi])] == orec).all()
    # two shards on 65,536-read boundaries, encoded separately, concatenate to the same stream
    parts = []
    for lo, hi in sharding.shard_bounds(cfg["n_reads"], 2):
        r, _ = ctx.encode_batch(bases[int(off[lo]): int(off[hi])], off[lo: hi + 1] - off[lo])
        parts.append(r)
    assert (sharding.concat_in_rank_order(parts) == rec).all()


def test_record_length_limit_2_pow_24(backend):
    """Maximum sizes: a read that is one unbroken unique match.  Self-encoding a contig of n bases gives a first record
    of hops * k + 1 bases (src/lib.rs:186-203); at >= 2^24 the reference asserts (src/lib.rs:226) -> NTC_ERR_LEN24, just
    below it the record is legal and must survive the round trip.  Also the longest chain of matching-statistics tiles
    the default tile length produces (unique-predecessor runs continuing across every tile border)."""
    rng = np.random.default_rng(2024)
    k = 31
    contig = helpers.random_seq(rng, (1 << 24) + 4 * k)
    index = nt.Index.build([contig], k, False)
    ctx = make_ctx(backend, index)
    with pytest.raises(nt.NtcError) as e:
        ctx.encode_batch(*nt.concat_reads([contig]))
    assert e.value.code == 6                                          # NTC_ERR_LEN24
    n_ok = ((1 << 24) - 2) // k * k - k + 5                           # first record well below the limit
    read = np.ascontiguousarray(contig[100: 100 + n_ok])
    bases, off = nt.concat_reads([read, contig[:500], read[: 3 * k]])
    rec, roff = ctx.encode_batch(bases, off)
    lens = np.where((rec >> np.uint64(57)) & np.uint64(1), rec >> np.uint64(58), (rec >> np.uint64(32)) & np.uint64(0xFFFFFF))
    assert int(lens.max()) > (1 << 23) and int(lens.max()) < (1 << 24)
    assert int(lens[int(roff[0]): int(roff[1])].sum()) == n_ok        # src/lib.rs:227
    db, doff = ctx.decode_batch(rec)
    assert (doff == off).all() and (db == bases).all()


def test_prefix_table_boundaries(small_cases):
    """ms_kernel takes the first 8 bases of a read (and of a tile's warm-up) from the prefix table when k >= 10:
    read lengths around 8 and around the 16-base word boundary, reads at every alignment of the packed stream,
    reads whose first 8-mer is not in the index (table entry empty -> ordinary path), with and without tiles."""
    rng = np.random.default_rng(88)
    for case in small_cases:
        src = [c for c in case.contigs if len(c) > 0]
        reads = helpers.sample_reads(rng, src, 120, lengths=[1, 2, 7, 8, 9, 10, 15, 16, 17, 23, 24, 25, 31, 32, 33, 40, 300],
                                     error_rates=[0, 0.02, 0.1], k=case.k)
        # foreign starts: random 8-mers (mostly absent from a small index) in front of genuine sequence
        for _ in range(40):
            r = helpers.sample_reads(rng, src, 1, lengths=[12, 60, 200], error_rates=[0], k=case.k)[0]
            reads.append(np.concatenate([helpers.random_seq(rng, int(rng.integers(1, 12))), r]))
        for T in (0, 256):
            case.ctx.set_tile_len(T)
            bases, off = nt.concat_reads(reads)
            d, s = case.ctx.matching_statistics(bases, off)
            for i, r in enumerate(reads):
                od, os_, _ = case.oracle.matching_statistics(r)
                lo, hi = int(off[i]), int(off[i + 1])
                assert (d[lo:hi] == od).all() and (s[lo:hi] == os_).all(), (case.k, T, i, len(r))
            rec, roff = case.ctx.encode_batch(bases, off)
            orec, oroff = case.oracle.encode_reads(bases, off, fast=False, threads=8)
            assert (roff == oroff).all() and (rec == orec).all(), (case.k, T)
            for r in reads[:30]:
                got = nt.encode_sequence(r, case.ctx)
                want = case.oracle.encode_sequence(r, fast=False)
                assert [(ln, (rg.start, rg.stop)) for ln, rg in got] == want, case.k
        case.ctx.set_tile_len(0)
