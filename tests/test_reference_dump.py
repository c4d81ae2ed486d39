"""The one-command parity pin: `rust/ref_dump` runs the UNMODIFIED reference over the
golden fixtures on a machine with cargo and writes tests/reference_dump/*; these tests compare every dumped artefact
with the oracle, the product's host code and -- on a GPU -- the CUDA path.  The directory is absent in this repository
(no Rust toolchain in its image): the comparisons then SKIP, and the oracle's parity stays "unpinned" (DESIGN.md
section 2).  `test_dump_comparator_self_check` feeds the comparator a dump written by the oracle in the very same
file formats, so the reader / comparator code is exercised either way.

    python rust/export_fixtures.py && cargo run --release --manifest-path rust/ref_dump/Cargo.toml -- tests/reference_dump
"""
import glob
import os

import numpy as np
import pytest

import ntcomp_b200 as nt
from oracle import binding as ob

HERE = os.path.dirname(os.path.abspath(__file__))
DUMP = os.path.join(HERE, "reference_dump")
SKIP_MSG = ("tests/reference_dump/ holds no reference dump: run rust/ref_dump where cargo exists "
            "(python rust/export_fixtures.py; cargo run --release --manifest-path rust/ref_dump/Cargo.toml -- tests/reference_dump)")


def _names(d):
    return sorted(os.path.basename(p)[: -len(".index.bin")] for p in glob.glob(os.path.join(d, "*.index.bin")))


def _read_fasta(path):
    seqs = []
    for line in open(path, "rb").read().split(b"\n"):
        if line.startswith(b">"):
            seqs.append(b"")
        elif line:
            seqs[-1] += line
    return [np.frombuffer(s, dtype=np.uint8) for s in seqs]


def load_dump(d, name):
    """the artefacts rust/ref_dump/src/main.rs writes for one fixture"""
    raw = open(os.path.join(d, name + ".index.bin"), "rb").read()
    k, n = (int(x) for x in np.frombuffer(raw[:16], dtype="<u8"))
    words = (n + 63) // 64
    rows = [np.frombuffer(raw[16 + c * words * 8: 16 + (c + 1) * words * 8], dtype="<u8") for c in range(4)]
    lcs = np.frombuffer(raw[16 + 4 * words * 8: 16 + 4 * words * 8 + n], dtype=np.uint8)
    assert len(raw) == 16 + 4 * words * 8 + n, "index.bin has trailing or missing bytes"
    ms0 = np.frombuffer(open(os.path.join(d, name + ".ms0.bin"), "rb").read(), dtype="<u8").reshape(-1, 3)
    reads = _read_fasta(os.path.join(d, "in", name + ".reads.fa"))
    contigs = _read_fasta(os.path.join(d, "in", name + ".contigs.fa"))
    rec_raw = np.frombuffer(open(os.path.join(d, name + ".records.bin"), "rb").read(), dtype="<u8")
    counts, records = rec_raw[: len(reads)], rec_raw[len(reads):]
    assert int(counts.sum()) == len(records)
    return {"k": k, "n": n, "rows": rows, "lcs": lcs, "ms0": ms0, "reads": reads, "contigs": contigs, "counts": counts, "records": records,
            "encoded": open(os.path.join(d, name + ".encoded.dat"), "rb").read(),
            "decoded": open(os.path.join(d, name + ".decoded.fa"), "rb").read()}


def _fasta(seqs, first=1):
    return b"".join(b">seq.%d\n%s\n" % (first + i, bytes(s)) for i, s in enumerate(seqs))


def _blocks(blob):
    """record streams of the blocks of an encoded.dat (oracle container reader; DEFLATE bytes are not compared: zlib vs zlib-rs)"""
    out, pos = [], 32
    while True:
        r = ob.read_block(blob, pos)
        if r is None:
            return out
        out.append(r[0])
        pos += r[2]


def compare_with_host(z):
    """reference dump vs the oracle and the product's host code (index builder, container)"""
    k = z["k"]
    # index: product builder and oracle builder against kbo::build
    idx = nt.Index.build(z["contigs"], k)
    assert idx.n_sets == z["n"], "n_sets differs from kbo::build (dummy-node handling?)"
    for c in range(4):
        assert (idx.rows()[c] == z["rows"][c]).all(), "subset row %d differs" % c
    assert (idx.lcs() == z["lcs"]).all(), "LCS array differs"
    o = ob.OracleIndex.from_arrays(k, z["n"], z["rows"], z["lcs"])
    # matching statistics of read 0 (StreamingIndex::matching_statistics)
    d, s, e = o.matching_statistics(z["reads"][0])
    assert (d == z["ms0"][:, 0]).all() and (s == z["ms0"][:, 1]).all() and (e == z["ms0"][:, 2]).all(), "matching statistics differ"
    # record stream: literal restatement of encode_sequence + encode_dictionary
    bases, off = nt.concat_reads(z["reads"])
    rec, roff = o.encode_reads(bases, off, fast=False, threads=4)
    assert (np.diff(roff.astype(np.int64)) == z["counts"].astype(np.int64)).all(), "records per read differ"
    assert (rec == z["records"]).all(), "record words differ"
    # container: the blocks of the reference's file hold the same records (or were dropped: Q4)
    got = _blocks(z["encoded"])
    n_reads = len(z["reads"])
    want = []
    for r0 in range(0, n_reads, 65536):
        r1 = min(r0 + 65536, n_reads)
        blk = z["records"][int(roff[r0]): int(roff[r1])]
        if ob.write_block(blk, r1 - r0) is not None:
            want.append(blk)
    assert len(got) == len(want) and all((a == b).all() for a, b in zip(got, want)), "encoded.dat blocks differ"
    # decode: FASTA text of the surviving blocks
    text = b""
    first = 1
    for blk in got:
        b, o_ = o.decode(blk, fast=False)
        seqs = [b[int(o_[i]): int(o_[i + 1])] for i in range(len(o_) - 1)]
        text += _fasta(seqs, first)
        first += len(seqs)
    assert text == z["decoded"], "decoded FASTA differs"
    return got


def compare_with_gpu(z, blocks):
    """reference dump vs the CUDA path through the C ABI"""
    idx = nt.Index.from_arrays(z["k"], z["n"], z["rows"], z["lcs"])
    ctx = nt.Context(idx, 0)
    bases, off = nt.concat_reads(z["reads"])
    rec, roff = ctx.encode_batch(bases, off)
    assert (np.diff(roff.astype(np.int64)) == z["counts"].astype(np.int64)).all() and (rec == z["records"]).all()
    d, s = ctx.matching_statistics(*nt.concat_reads([z["reads"][0]]))
    assert (d == z["ms0"][:, 0]).all() and (s == z["ms0"][:, 1]).all()
    text, first = b"", 1
    for blk in blocks:
        t, n = ctx.decode_fasta(blk, first)
        text += t
        first += n
    assert text == z["decoded"]


def write_dump_with_oracle(d, name, k, contigs, reads):
    """the file formats of rust/ref_dump/src/main.rs, written by the ORACLE (comparator self-check only)"""
    os.makedirs(os.path.join(d, "in"), exist_ok=True)
    open(os.path.join(d, "in", name + ".k"), "w").write("%d\n" % k)
    for tag, seqs in (("contigs", contigs), ("reads", reads)):
        with open(os.path.join(d, "in", "%s.%s.fa" % (name, tag)), "wb") as f:
            for i, sq in enumerate(seqs):
                f.write(b">x.%d\n%s\n" % (i + 1, bytes(sq)))
    o = ob.OracleIndex.build(contigs, k)
    with open(os.path.join(d, name + ".index.bin"), "wb") as f:
        f.write(np.array([k, o.n], dtype="<u8").tobytes())
        for r in o.rows():
            f.write(r.astype("<u8").tobytes())
        f.write(o.lcs().tobytes())
    dd, s, e = o.matching_statistics(reads[0])
    np.stack([dd.astype("<u8"), s.astype("<u8"), e.astype("<u8")], axis=1).tofile(os.path.join(d, name + ".ms0.bin"))
    bases, off = nt.concat_reads(reads)
    rec, roff = o.encode_reads(bases, off, fast=False)
    np.concatenate([np.diff(roff.astype(np.int64)).astype("<u8"), rec.astype("<u8")]).tofile(os.path.join(d, name + ".records.bin"))
    blob = bytes(32)
    text, first = b"", 1
    blk = ob.write_block(rec, len(reads))
    if blk is not None:
        blob += blk
        text = _fasta(reads, first)
    open(os.path.join(d, name + ".encoded.dat"), "wb").write(blob)
    open(os.path.join(d, name + ".decoded.fa"), "wb").write(text)


def _self_dump(tmp_path):
    rng = np.random.default_rng(5)
    acgt = np.frombuffer(b"ACGT", dtype=np.uint8)
    g = acgt[rng.integers(0, 4, size=3000)]
    reads = []
    for _ in range(30):
        p = int(rng.integers(0, 3000 - 120))
        r = g[p: p + 120].copy()
        r[int(rng.integers(0, 120))] = acgt[int(rng.integers(0, 4))]
        reads.append(r)
    write_dump_with_oracle(str(tmp_path), "self", 31, [g], reads)
    assert _names(str(tmp_path)) == ["self"]
    return load_dump(str(tmp_path), "self")


def test_dump_comparator_self_check(tmp_path):
    z = _self_dump(tmp_path)
    blocks = compare_with_host(z)
    assert len(blocks) == 1
    z["records"] = z["records"].copy()
    z["records"][3] ^= np.uint64(1)                                       # the comparator must notice a one-bit difference
    with pytest.raises(AssertionError):
        compare_with_host(z)


@pytest.mark.gpu
def test_dump_comparator_self_check_gpu(tmp_path):
    z = _self_dump(tmp_path)
    compare_with_gpu(z, _blocks(z["encoded"]))


@pytest.mark.parametrize("name", _names(DUMP) or ["<absent>"])
def test_reference_dump_matches_oracle_and_host_code(name):
    if name == "<absent>":
        pytest.skip(SKIP_MSG)
    compare_with_host(load_dump(DUMP, name))


@pytest.mark.gpu
@pytest.mark.parametrize("name", _names(DUMP) or ["<absent>"])
def test_reference_dump_matches_gpu_path(name):
    if name == "<absent>":
        pytest.skip(SKIP_MSG)
    z = load_dump(DUMP, name)
    compare_with_gpu(z, _blocks(z["encoded"]))
