"""File-level host pipeline and command-line front end: `ntcomp build / encode / decode`
(src/main.rs:97-211) over the C ABI.  CPU part: fastX parsing and the build subcommand; the encode /
decode subcommands must fail loudly without a GPU.  GPU part: files round-trip, block streams equal the
oracle's record stream block by block, FASTA text is the writer loop's."""
import gzip
import os
import subprocess
import sys

import numpy as np
import pytest

import helpers
import ntcomp_b200 as nt
from ntcomp_b200 import _lib
from oracle import binding as ob

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "ntcomp_b200", "bin", "ntcomp-b200")


def _write_fasta(path, seqs, width=70, lower=False, opener=open):
    with opener(path, "wb") as f:
        for i, s in enumerate(seqs):
            f.write(b">contig%d some description\n" % i)
            b = bytes(s).lower() if lower else bytes(s)
            for j in range(0, len(b), width):
                f.write(b[j:j + width] + b"\r\n" if i % 2 else b[j:j + width] + b"\n")


def _write_fastq(path, seqs, opener=open):
    with opener(path, "wb") as f:
        for i, s in enumerate(seqs):
            q = b"@" * len(s)                      # quality lines that start with '@' must not confuse the parser
            f.write(b"@read%d\n%s\n+\n%s\n" % (i, bytes(s), q))


@pytest.fixture(scope="module")
def genome_and_reads():
    rng = np.random.default_rng(77)
    contigs = [helpers.random_seq(rng, n) for n in (40000, 2500, 90)]
    reads = helpers.sample_reads(rng, contigs, 700, lengths=[150, 150, 150, 91, 40, 400, 1000, 3], error_rates=[0, 0.01, 0.05], k=31)
    return contigs, reads


def test_build_from_files_matches_in_memory_build(tmp_path, genome_and_reads):
    contigs, _ = genome_and_reads
    fa, fq = tmp_path / "a.fa", tmp_path / "b.fq.gz"
    _write_fasta(fa, contigs[:2], lower=True)
    _write_fastq(fq, contigs[2:], opener=gzip.open)
    a = nt.Index.build_from_files([fa, fq], 31)
    b = nt.Index.build(contigs, 31)
    assert a.n_sets == b.n_sets and (a.lcs() == b.lcs()).all()
    for x, y in zip(a.rows(), b.rows()):
        assert (x == y).all()
    with pytest.raises(nt.NtcError) as e:
        nt.Index.build_from_files([tmp_path / "missing.fa"], 31)
    assert e.value.code == 7
    bad = tmp_path / "bad.fq"
    bad.write_bytes(b"@r1\nACGT\n+\nII\n")          # truncated quality
    with pytest.raises(nt.NtcError):
        nt.Index.build_from_files([bad], 31)


def _fastx_load(path, n_threads=3):
    import ctypes as C
    L = _lib.lib()
    b, o = C.c_void_p(), C.c_void_p()
    n, nb = C.c_uint64(0), C.c_uint64(0)
    rc = L.ntc_fastx_load(str(path).encode(), n_threads, C.byref(b), C.byref(o), C.byref(n), C.byref(nb))
    if rc:
        return rc, None
    bases = np.ctypeslib.as_array(C.cast(b, C.POINTER(C.c_uint8)), shape=(max(nb.value, 1),))[: nb.value].copy()
    off = np.ctypeslib.as_array(C.cast(o, C.POINTER(C.c_uint64)), shape=(n.value + 1,)).copy()
    L.ntc_free(b)
    L.ntc_free(o)
    return 0, [bytes(bases[int(off[i]): int(off[i + 1])]) for i in range(n.value)]


def test_parallel_fastx_parser(tmp_path):
    """The encode pipeline's front end (reader thread cutting at record boundaries + parse pool, in-order hand-over):
    files of several pieces, FASTA with wrapped / CRLF / lower-case lines, 4-line FASTQ whose quality lines start with
    '@', multi-line FASTQ (no safe cut: parsed serially), gzip input, malformed input."""
    rng = np.random.default_rng(123)
    seqs = [helpers.random_seq(rng, int(n)) for n in rng.integers(1, 400, size=60000)]       # ~12 MB of sequence: several 4 MB pieces
    seqs[17] = helpers.random_seq(rng, 5_000_000)                                             # one record longer than a piece
    want = [bytes(s) for s in seqs]
    fa = tmp_path / "a.fa"
    _write_fasta(fa, seqs, width=61, lower=True)
    assert _fastx_load(fa) == (0, want)
    assert _fastx_load(fa, 1) == (0, want)
    fq = tmp_path / "a.fq"
    _write_fastq(fq, seqs)
    assert _fastx_load(fq) == (0, want)
    fqz = tmp_path / "a.fq.gz"
    _write_fastq(fqz, seqs[:3000], opener=lambda p, m: gzip.open(p, m, compresslevel=1))
    assert _fastx_load(fqz) == (0, want[:3000])
    ml = tmp_path / "multi.fq"                                                                 # wrapped FASTQ: sequence and quality over several lines
    with open(ml, "wb") as f:
        for i, s in enumerate(seqs[:40000]):
            b = bytes(s)
            f.write(b"@r%d\n" % i + b"\n".join(b[j:j + 50] for j in range(0, len(b), 50)) + b"\n+\n" +
                    b"\n".join(b"@" * len(b[j:j + 50]) for j in range(0, len(b), 50)) + b"\n")
    assert _fastx_load(ml) == (0, want[:40000])
    u = tmp_path / "u.fa"
    u.write_bytes(b">x\nacgu ACGU\tnn\n\n>y\n\n>z\nA")
    assert _fastx_load(u) == (0, [b"ACGTACGTNN", b"", b"A"])
    for bad in (b"@r1\nACGT\n+\nII\n", b"@r1\nACGT\n", b"hello\n"):
        p = tmp_path / "bad.fq"
        p.write_bytes(bad)
        assert _fastx_load(p)[0] == 7, bad
    assert _fastx_load(tmp_path / "missing")[0] == 7
    e = tmp_path / "empty.fa"
    e.write_bytes(b"")
    assert _fastx_load(e) == (0, [])


def test_cli_build_and_no_gpu_failure(tmp_path, genome_and_reads):
    contigs, reads = genome_and_reads
    assert os.path.exists(CLI), "ntcomp-b200 is built by make -C ntcomp_b200/csrc"
    fa = tmp_path / "ref.fa"
    _write_fasta(fa, contigs)
    prefix = str(tmp_path / "idx")
    lst = tmp_path / "list.tsv"
    lst.write_text("ref\t%s\n" % fa)
    subprocess.run([CLI, "build", "-k", "31", "-o", prefix, "-l", str(lst), "-t", "2"], check=True, capture_output=True)
    idx = nt.Index.load(prefix + ".ntcidx")
    assert idx.k == 31 and idx.n_sets == nt.Index.build(contigs, 31).n_sets
    assert subprocess.run([CLI], capture_output=True).returncode == 2
    import torch
    if not torch.cuda.is_available():                # product path must fail loudly without a GPU
        fq = tmp_path / "r.fq"
        _write_fastq(fq, reads[:5])
        r = subprocess.run([CLI, "encode", "-i", prefix, str(fq)], capture_output=True)
        assert r.returncode == 1 and b"no CUDA device" in r.stderr and r.stdout == b""


@pytest.mark.gpu
def test_cli_round_trip_and_block_parity(tmp_path, genome_and_reads, monkeypatch):
    contigs, reads = genome_and_reads
    # 3 blocks: shrink nothing -- 65,536-read blocks need > 131,072 reads; repeat the read set
    reads = (reads * 190)[: 2 * 65536 + 777]
    fa, fq = tmp_path / "ref.fa", tmp_path / "reads.fq.gz"
    _write_fasta(fa, contigs)
    _write_fastq(fq, reads, opener=lambda p, m: gzip.open(p, m, compresslevel=1))
    prefix = str(tmp_path / "idx")
    subprocess.run([CLI, "build", "-k", "31", "-o", prefix, str(fa)], check=True, capture_output=True)
    # `build --gpu` (ntc_index_build_gpu behind the same front end) writes the same file, byte for byte
    subprocess.run([CLI, "build", "--gpu", "-k", "31", "-o", prefix + "_gpu", str(fa)], check=True, capture_output=True)
    with open(prefix + ".ntcidx", "rb") as fh, open(prefix + "_gpu.ntcidx", "rb") as fg:
        assert fh.read() == fg.read()
    assert (nt.Index.build_from_files([fa], 31, gpu=0).lcs() == nt.Index.load(prefix + ".ntcidx").lcs()).all()
    enc, dec = tmp_path / "encoded.dat", tmp_path / "decoded.fasta"
    with open(enc, "wb") as f:
        r = subprocess.run([CLI, "encode", "-i", prefix, str(fq)], stdout=f, stderr=subprocess.PIPE)
    assert r.returncode == 0, r.stderr
    with open(dec, "wb") as f:
        r = subprocess.run([CLI, "decode", "-i", prefix, str(enc)], stdout=f, stderr=subprocess.PIPE)
    assert r.returncode == 0, r.stderr
    assert dec.read_bytes() == helpers.fasta_text(reads, 1)                      # README.md:48-53 diff procedure
    # block stream: 32 zero bytes, then per 65,536 reads the oracle's records (P1) in the reference's container
    data = enc.read_bytes()
    assert data[:32] == bytes(32)
    index = nt.Index.load(prefix + ".ntcidx")
    oracle = ob.OracleIndex.from_arrays(index.k, index.n_sets, index.rows(), index.lcs())
    bases, off = nt.concat_reads(reads)
    orec, oroff = oracle.encode_reads(bases, off, fast=True, threads=8)
    pos, lo, n_blocks = 32, 0, 0
    while pos < len(data):
        rec, used = nt.block_read(data[pos:])
        assert used > 0
        hi = min(len(reads), lo + 65536)
        assert (rec == orec[int(oroff[lo]): int(oroff[hi])]).all()
        num_records = int(np.frombuffer(data[pos + 4: pos + 8], dtype=np.uint32)[0])
        assert num_records == hi - lo                                             # src/main.rs:170,176
        pos += used
        lo = hi
        n_blocks += 1
    assert n_blocks == 3 and lo == len(reads)
    # same through the Python mirror of the file API
    ctx = nt.Context(index, 0)
    st = ctx.encode_file(fq, tmp_path / "e2.dat", n_threads=3)
    assert st["reads"] == len(reads) and st["blocks"] == 3 and st["dropped_blocks"] == 0 and st["bytes"] == len(data)
    assert (tmp_path / "e2.dat").read_bytes() == data
    st = ctx.decode_file(tmp_path / "e2.dat", tmp_path / "d2.fa", n_threads=3)
    assert st["sequences"] == len(reads) and (tmp_path / "d2.fa").read_bytes() == dec.read_bytes()
    # several GPU batches per file (one block per batch): same bytes, same text
    monkeypatch.setenv("NTC_PIPELINE_BATCH_BLOCKS", "1")
    ctx.encode_file(fq, tmp_path / "e3.dat", n_threads=2)
    assert (tmp_path / "e3.dat").read_bytes() == data
    ctx.decode_file(tmp_path / "e3.dat", tmp_path / "d3.fa", n_threads=2)
    assert (tmp_path / "d3.fa").read_bytes() == dec.read_bytes()
    monkeypatch.delenv("NTC_PIPELINE_BATCH_BLOCKS")
    # one record per base (reads of 1-3 bases): the pipeline's first record buffer is too small and it retries with the bound
    tiny = [helpers.random_seq(np.random.default_rng(5), n) for n in ([1, 2, 3, 1] * 700)]
    _write_fasta(tmp_path / "tiny.fa", tiny)
    st = ctx.encode_file(tmp_path / "tiny.fa", tmp_path / "tiny.dat")
    assert st["reads"] == len(tiny) and st["records"] >= len(tiny)
    if st["dropped_blocks"] == 0:
        ctx.decode_file(tmp_path / "tiny.dat", tmp_path / "tiny.out")
        assert (tmp_path / "tiny.out").read_bytes() == helpers.fasta_text(tiny, 1)
    # a read with an N makes encode fail with NTC_ERR_BAD_BASE instead of the reference's underflow (Q3)
    bad = tmp_path / "bad.fa"
    bad.write_bytes(b">x\nACGTNACGT\n")
    with pytest.raises(nt.NtcError) as e:
        ctx.encode_file(bad, tmp_path / "bad.dat")
    assert e.value.code == 5


_SECOND_GPU_SCRIPT = r"""
import ctypes as C, sys
import numpy as np
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, sys.argv[1] + "/tests")
import helpers, ntcomp_b200 as nt
cu = C.CDLL("libcuda.so.1")
def active(dev):
    d, flags, act = C.c_int(), C.c_uint(), C.c_int()
    assert cu.cuInit(0) == 0 and cu.cuDeviceGet(C.byref(d), dev) == 0
    assert cu.cuDevicePrimaryCtxGetState(d, C.byref(flags), C.byref(act)) == 0
    return act.value
rng = np.random.default_rng(5)
contigs = [helpers.random_seq(rng, 30000)]
reads = helpers.sample_reads(rng, contigs, 3000, lengths=[150, 150, 400], error_rates=[0, 0.01], k=31)
tmp = sys.argv[2]
with open(tmp + "/r.fa", "wb") as f:
    f.write(helpers.fasta_text(reads, 1))
ctx = nt.Context(nt.Index.build(contigs, 31, True), 1)
assert active(0) == 0 and active(1) == 1
ctx.encode_file(tmp + "/r.fa", tmp + "/enc.dat", 4)
assert active(0) == 0, "ntc_encode_file on GPU 1 created a CUDA context on GPU 0"
ctx.decode_file(tmp + "/enc.dat", tmp + "/dec.fa", 4)
assert active(0) == 0, "ntc_decode_file on GPU 1 created a CUDA context on GPU 0"
assert open(tmp + "/dec.fa", "rb").read() == helpers.fasta_text(reads, 1)
print("SECOND_GPU_OK")
"""


@pytest.mark.gpu
def test_file_pipeline_on_second_gpu_leaves_gpu0_alone(tmp_path):
    """A context on GPU 1 must not bring up a CUDA context on GPU 0: the pipeline's batch-assembler thread allocates
    page-locked buffers and a new thread's current device is 0 (ntc_ctx_bind_thread).  Checked from a fresh process (this
    one may already hold GPU 0 through torch) with the driver's own cuDevicePrimaryCtxGetState.  Needs two GPUs."""
    import torch
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    r = subprocess.run([sys.executable, "-c", _SECOND_GPU_SCRIPT, ROOT, str(tmp_path)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "SECOND_GPU_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


@pytest.mark.gpu
def test_decode_file_writes_the_blocks_before_a_damaged_one(tmp_path, genome_and_reads, monkeypatch):
    """src/main.rs:203 (`while let Ok(..)`) writes every block decoded before the first bad one.  ntc_decode_file
    must do the same whatever its internal grouping of blocks is (and then report the error, a deliberate difference).
    Third block of three damaged (a) in its gzip bytes -- caught while unpacking -- and (b) in a header count that only the
    column decoder on the GPU can see."""
    import struct
    contigs, reads = genome_and_reads
    reads = (reads * 190)[: 2 * 65536 + 777]
    fq = tmp_path / "reads.fa"
    fq.write_bytes(helpers.fasta_text(reads, 1))
    ctx = nt.Context(nt.Index.build(contigs, 31, True), 0)
    enc = tmp_path / "enc.dat"
    st = ctx.encode_file(fq, enc, 4)
    assert st["blocks"] == 3 and st["dropped_blocks"] == 0
    data = enc.read_bytes()
    starts, pos = [], 32
    while pos < len(data):
        starts.append(pos)
        for _ in range(4):
            pos += 32 + struct.unpack_from("<I", data, pos)[0]
    assert len(starts) == 3 and pos == len(data)
    want = helpers.fasta_text(reads[: 2 * 65536], 1)
    b3 = starts[2]
    bs0 = struct.unpack_from("<I", data, b3)[0]
    gz_damage = bytearray(data)
    for i in range(40, 60):                                            # inside the first column's deflate stream
        gz_damage[b3 + 32 + i] ^= 0xFF
    count_damage = bytearray(data)                                     # one more match length announced than the stream holds
    n_len = struct.unpack_from("<I", data, b3 + 32 + bs0 + 8)[0]
    struct.pack_into("<I", count_damage, b3 + 32 + bs0 + 8, n_len + 1)
    for name, blob in (("gzip", gz_damage), ("count", count_damage)):
        for group in ("1", "2", None):
            if group is None:
                monkeypatch.delenv("NTC_PIPELINE_BATCH_BLOCKS", raising=False)
            else:
                monkeypatch.setenv("NTC_PIPELINE_BATCH_BLOCKS", group)
            bad, out = tmp_path / "bad.dat", tmp_path / "out.fa"
            bad.write_bytes(bytes(blob))
            with pytest.raises(nt.NtcError):
                ctx.decode_file(bad, out, 4)
            assert out.read_bytes() == want, (name, group)
    monkeypatch.delenv("NTC_PIPELINE_BATCH_BLOCKS", raising=False)
    ctx.decode_file(enc, tmp_path / "ok.fa", 4)                        # the context is still usable after the failures
    assert (tmp_path / "ok.fa").read_bytes() == helpers.fasta_text(reads, 1)
    ctx.close()
