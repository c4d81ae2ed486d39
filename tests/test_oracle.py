"""CPU tests of the oracle (oracle/*.cpp): against the definition-level Python model, literal vs fast
restatements, the reference's own round-trip property (tests/fasta_data.rs), its in-code
invariants and quirks Q1-Q6.  PARITY UNPINNED: the reference holds no golden vectors."""
import numpy as np
import pytest

import helpers
from bruteforce import BruteSBWT
from oracle import binding as ob


def _bits(row, n):
    return [(int(row[i >> 6]) >> (i & 63)) & 1 for i in range(n)]


@pytest.mark.parametrize("seed", range(6))
def test_index_matches_definition_model(seed):
    rng = np.random.default_rng(seed)
    k = int(rng.choice([3, 4, 5, 7, 12, 16, 31, 33]))
    seqs = [bytes(helpers.random_seq(rng, int(rng.integers(0, 120)))).decode() for _ in range(int(rng.integers(1, 4)))]
    if seed % 2 == 0:
        seqs.append("ACGTNNACGTACGTTTGACN")
    rc = bool(rng.random() < 0.7)
    b = BruteSBWT(seqs, k, rc)
    o = ob.OracleIndex.build([s.encode() for s in seqs], k, rc)
    assert o.n == b.n and list(o.C()) == b.C and list(o.lcs()) == b.lcs
    for c, row in enumerate(o.rows()):
        assert _bits(row, b.n) == b.rows[c]
    for v in range(b.n):
        assert o.access_kmer(v).decode() == b.nodes[v]
    for s in seqs:
        s = "".join(ch if ch in "ACGT" else "A" for ch in s)
        if len(s) < 2:
            continue
        read = bytes(helpers.mutate(rng, np.frombuffer(s.encode(), dtype=np.uint8), 0.05)).decode()
        d, st, en = o.matching_statistics(read.encode())
        assert [(int(d[i]), (int(st[i]), int(en[i]))) for i in range(len(read))] == b.matching_statistics(read)
        iv = o.search(read[:k].encode())
        assert iv == b.interval_of_suffix(read[:k]) or (iv is None and b.interval_of_suffix(read[:k]) is None)


def test_handmade_known_answer():
    """Worked by hand from the SBWT definition: one sequence ACGT, k = 3, no reverse complements.
    k-mers ACG, CGT; ACG has no predecessor -> dummies $$$, $$A, $AC.  Colex order:
    $$$ < $$A < $AC < ACG < CGT."""
    o = ob.OracleIndex.build([b"ACGT"], 3, add_revcomp=False)
    assert o.n == 5
    assert [o.access_kmer(v) for v in range(5)] == [b"$$$", b"$$A", b"$AC", b"ACG", b"CGT"]
    rows = [_bits(r, 5) for r in o.rows()]
    assert rows[0] == [1, 0, 0, 0, 0]   # $$$ -A-> $$A
    assert rows[1] == [0, 1, 0, 0, 0]   # $$A -C-> $AC
    assert rows[2] == [0, 0, 1, 0, 0]   # $AC -G-> ACG
    assert rows[3] == [0, 0, 0, 1, 0]   # ACG -T-> CGT
    assert list(o.C()) == [1, 2, 3, 4, 5]
    assert list(o.lcs()) == [0, 0, 0, 0, 0]
    d, s, e = o.matching_statistics(b"ACGTA")
    assert list(d) == [1, 2, 3, 3, 1] and list(s) == [1, 2, 3, 4, 1] and list(e) == [2, 3, 4, 5, 2]


@pytest.mark.parametrize("k", [12, 15, 31, 91])
def test_literal_equals_fast_and_round_trips(k):
    rng = np.random.default_rng(100 + k)
    contigs = [helpers.random_seq(rng, n) for n in (6000, 700)]
    o = ob.OracleIndex.build(contigs, k, True)
    reads = helpers.sample_reads(rng, contigs, 120, [1, 3, k - 1, k, k + 1, k + 2, 2 * k + 1, 150, 500], [0, 0.01, 0.05, 0.3], k)
    reads += contigs
    for r in reads:
        lit = o.encode_sequence(r, fast=False)
        assert lit == o.encode_sequence(r, fast=True)
        assert sum(x[0] for x in lit) == len(r)                          # lib.rs:227
        assert max(x[0] for x in lit) < (1 << 24)                        # lib.rs:226
    bases, off = ob.concat_reads(reads)
    rec, roff = o.encode_reads(bases, off, fast=False, threads=4)
    rec_f, roff_f = o.encode_reads(bases, off, fast=True, threads=4)
    assert (rec == rec_f).all() and (roff == roff_f).all()
    for fast in (False, True):
        db, doff = o.decode(rec, fast=fast)
        assert (db == bases).all() and (doff == off).all()
    # format invariants of the record word (encode.rs:145-158)
    flags = (rec >> np.uint64(56)).astype(np.uint8)
    first = flags & 1
    assert int(first.sum()) == len(reads) and all(first[int(roff[i])] == 1 for i in range(len(reads)))
    lit_mask = (flags & 2) != 0
    assert ((flags[lit_mask] >> 2) <= 11).all() and ((flags[lit_mask] >> 2) >= 1).all()
    assert (((rec[~lit_mask] >> np.uint64(32)) & np.uint64(0xFFFFFF)) > 11).all()
    assert ((flags[~lit_mask] >> 1) == 0).all()


def test_q1_full_match_without_extension_is_one_base_literal():
    """Q1: ext == 0 turns a full-k match into a 1-base literal (lib.rs:199)."""
    rng = np.random.default_rng(3)
    g = helpers.random_seq(rng, 3000)
    k = 16
    o = ob.OracleIndex.build([g], k, True)
    seen = False
    for a in range(0, 2000, 37):
        read = g[a:a + 60].copy()
        read[60 - k - 1] = helpers.ALPHA[(int(np.searchsorted(helpers.ALPHA, read[60 - k - 1])) + 1) & 3]  # error just left of the last k-mer
        d = o.encode_sequence(read)
        if d[0][0] == 1:
            seen = True
            words = o.encode_dictionary(d)
            assert (int(words[0]) >> 56) == (1 + 2) | (1 << 2)
    assert seen


def test_appendix_b_worked_example():
    """error-free 150-bp read, k = 91, unique region -> records (92, first) + (58)."""
    rng = np.random.default_rng(4)
    g = helpers.random_seq(rng, 5000)
    o = ob.OracleIndex.build([g], 91, True)
    d = o.encode_sequence(g[1000:1150])
    assert [x[0] for x in d] == [92, 58]
    words = o.encode_dictionary(d)
    assert (int(words[0]) >> 56) == 1 and (int(words[1]) >> 56) == 0
    assert ((int(words[0]) >> 32) & 0xFFFFFF) == 92


def test_reference_roundtrip_fixture_c1():
    """tests/fasta_data.rs restated: 7 contigs, k = 255, block size 3, through the block container."""
    contigs, k, block = helpers.fixture_c1()
    o = ob.OracleIndex.build(contigs, k, True)
    stream = bytearray(32)                                                # encode_file_header(0,0,0,0), lib.rs:52
    written = []
    for b in range(0, len(contigs), block):
        blk = contigs[b:b + block]
        bases, off = ob.concat_reads(blk)
        rec, _ = o.encode_reads(bases, off, fast=False)
        db, doff = o.decode(rec, fast=False)                              # the property itself, per block
        assert (db == bases).all() and (doff == off).all()
        blob = ob.write_block(rec, len(blk))
        flags = (rec >> np.uint64(56)) & np.uint64(2)
        if blob is None:
            # Q4: a block without a literal (or without a reference) record cannot be coded
            # (minimal_binary_encode on an empty column, encode.rs:80) and main.rs:170 drops it.
            # Self-encoding an error-free contig longer than k+1 yields two reference records
            # unless (len-1) % k <= 11, so this is the COMMON case for this fixture shape under the
            # restated semantics -- recorded in DESIGN.md as the main "parity unpinned" observation.
            assert flags.all() or not flags.any()
            continue
        written += blk
        stream += blob
    contigs = written
    pos, decoded = 32, []
    while True:
        got = ob.read_block(bytes(stream), pos)
        if got is None:
            break
        rec, hdr, used = got
        pos += used
        db, doff = o.decode(rec, fast=False)
        decoded += [bytes(db[int(doff[i]): int(doff[i + 1])]) for i in range(len(doff) - 1)]
        for q in range(4):
            assert hdr[q][5] == 8                                         # bitpacker_exponent, encode.rs:118
    assert decoded == [bytes(c) for c in contigs]                         # fasta_data.rs:93
    assert sum(map(len, decoded)) == sum(map(len, contigs))


def test_errors_empty_and_absent_base():
    rng = np.random.default_rng(5)
    o = ob.OracleIndex.build([helpers.random_seq(rng, 500)], 12, True)
    with pytest.raises(ob.OracleError) as e:
        o.encode_sequence(b"")
    assert e.value.code == -1                                             # lib.rs:224
    with pytest.raises(ob.OracleError) as e:
        o.encode_sequence(b"ACGTNACGT")
    assert e.value.code == -3                                             # Q3


# ---- container ----

def test_column_codecs_round_trip_and_parameters():
    rng = np.random.default_rng(6)
    for _ in range(20):
        n = int(rng.integers(1, 3000))
        vals = rng.integers(0, int(rng.choice([2, 50, 1000, 1 << 20, 1 << 32])), size=n).astype(np.uint64)
        w, p = ob.code_column(vals, rice=False)
        assert p == int(vals.max()) + 2                                   # encode.rs:82
        assert (ob.decode_column(w, n, False, p) == vals).all()
        if vals.sum() > 0:
            small = (vals % np.uint64(4096)).astype(np.uint64)
            if small.sum() == 0:
                continue
            w, p = ob.code_column(small, rice=True)
            assert (ob.decode_column(w, n, True, p) == small).all()
    # Rice parameter = ceil(log2(-ln(phi)/ln(1-p))), p = n/sum (encode.rs:62-63)
    w, p = ob.code_column(np.full(100, 35, dtype=np.uint64), rice=True)
    assert p == 5
    w, p = ob.code_column(np.array([0, 1, 0, 1, 1], dtype=np.uint64), rice=True)   # mean < 1: NaN/-inf saturate to 0
    assert p == 0
    # bit order: MSB-first stream, words stored big-endian (first byte in memory holds the first bits)
    w, p = ob.code_column(np.array([0], dtype=np.uint64), rice=False)     # max = 2: l = 1, limit = 2; n+1 = 1 -> bit '1'
    assert p == 2 and w.tobytes()[0] == 0x80


def test_split_zip_and_quirks():
    rng = np.random.default_rng(7)
    g = helpers.random_seq(rng, 4000)
    o = ob.OracleIndex.build([g], 20, True)
    reads = helpers.sample_reads(rng, [g], 300, [150], [0.03], 20)
    bases, off = ob.concat_reads(reads)
    rec, _ = o.encode_reads(bases, off, fast=True, threads=4)
    colex, ln, flags, lits = ob.split(rec)
    assert len(colex) == len(ln) and len(flags) == len(rec)               # lib.rs:239-240
    n_lit_bases = int(((flags >> np.uint64(2)) * ((flags >> np.uint64(1)) & np.uint64(1))).sum())
    assert len(lits) == (n_lit_bases + 30) // 31                          # encode.rs:224
    blob = ob.write_block(rec, 300)
    if n_lit_bases % 31 != 0:                                             # Q5: multiple of 31 panics in the reference decoder
        rec2, hdr, used = ob.read_block(blob)
        assert used == len(blob) and (rec2 == rec).all()
        assert all(int(hdr[q][1]) == 300 for q in range(4))               # Q6: num_records is the block size argument
        assert int(hdr[0][2]) == len(colex) and int(hdr[2][2]) == len(rec)
    # Q4: a block with no literal record (or no reference record) cannot be written
    only_ref = rec[((rec >> np.uint64(56)) & np.uint64(2)) == 0]
    assert ob.write_block(only_ref, 1) is None
    only_lit = rec[((rec >> np.uint64(56)) & np.uint64(2)) != 0]
    assert ob.write_block(only_lit, 1) is None
