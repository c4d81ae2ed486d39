"""The C-ABI library loads, exports every symbol include/ntcomp_b200.h declares, and refuses to
compute without a GPU (no CPU fallback)."""
import os
import re

import numpy as np
import pytest

import ntcomp_b200 as nt
from ntcomp_b200 import _lib
from conftest import ROOT, has_gpu


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "ntcomp_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ntc_[A-Za-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported_and_bound():
    L = _lib.lib()
    names = _declared_symbols()
    assert len(names) >= 25
    for n in names:
        assert hasattr(L, n), "libntcomp_b200.so does not export %s" % n
        assert n in _lib.SIGNATURES, "%s has no ctypes prototype" % n
    assert sorted(_lib.SIGNATURES) == names
    assert b"sm_100a" in L.ntc_version()


def test_library_holds_sm100a_code_only():
    import subprocess
    out = subprocess.run(["cuobjdump", "-lelf", _lib.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_product_does_not_link_or_import_the_oracle():
    import subprocess
    out = subprocess.run(["ldd", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in out
    for dirpath, _, files in os.walk(os.path.join(ROOT, "ntcomp_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".inc")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.replace("the oracle", "").replace("or the oracle", "") or f in ("synth.py", "synth_common.h"), f


@pytest.mark.skipif(has_gpu(), reason="CPU-only behaviour")
def test_fails_loudly_without_gpu():
    g = np.frombuffer(b"ACGTTGCATGCATGCAGTCAGTCATGCATGCATGCAGTACGTAGCTAGCTAGCTAGCATG" * 4, dtype=np.uint8)
    idx = nt.Index.build([g], 16)
    with pytest.raises(nt.NtcError) as e:
        nt.Context(idx, 0)
    assert e.value.code == 2 and "no CPU fallback" in str(e.value)


def test_bounds_helpers():
    L = _lib.lib()
    assert L.ntc_encode_bound(1234) == 1234
    rec = np.array([(5 | (20 << 32) | (1 << 56)), (3 << 2 | 2) << 56, (7 | (12 << 32) | (1 << 56))], dtype=np.uint64)
    ctxless = nt.Context.__new__(nt.Context)
    nb, ns, fb = nt.Context.decode_bound(ctxless, rec, 9)
    assert (nb, ns) == (35, 2)
    assert fb == 35 + len(b">seq.9\n\n") + len(b">seq.10\n\n")


def test_kernel_bodies_under_address_sanitizer():
    """compute-sanitizer is closed on the GPU pool, so the per-thread kernel bodies are run on the CPU
    under AddressSanitizer with exactly-sized buffers (tests/emul/asan_driver.cu)."""
    import subprocess
    emul = os.path.join(ROOT, "tests", "emul")
    _lib.lib()
    subprocess.check_call(["make", "-C", emul, "-s", "asan_driver"], stderr=subprocess.DEVNULL)
    env = dict(os.environ, ASAN_OPTIONS="detect_leaks=0:protect_shadow_gap=0")
    out = subprocess.run([os.path.join(emul, "asan_driver")], capture_output=True, text=True, env=env, timeout=600)
    assert out.returncode == 0 and "ASAN_DRIVER_OK" in out.stdout, out.stdout[-1500:] + out.stderr[-3000:]


def test_rust_sys_crate_declares_every_header_symbol():
    """rust/ntcomp-b200-sys/src/lib.rs (the FFI crate a Rust host links) is written by hand against the
    header -- there is no cargo in this image -- so its `extern "C"` block is held to the header's symbol list and argument
    counts here."""
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    header = open(os.path.join(root, "include", "ntcomp_b200.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    rust = open(os.path.join(root, "rust", "ntcomp-b200-sys", "src", "lib.rs")).read()
    rust = re.sub(r"//[^\n]*", "", rust)
    c_decl = {m.group(1): m.group(2) for m in re.finditer(r"\b(ntc_[a-zA-Z_0-9]+)\s*\(([^;{]*?)\)\s*;", header)}
    r_decl = {m.group(1): m.group(2) for m in re.finditer(r"pub fn (ntc_[a-zA-Z_0-9]+)\s*\(([^;{]*?)\)\s*(?:->\s*[^;]+)?;", rust, flags=re.S)}
    assert set(c_decl) == set(r_decl), (sorted(set(c_decl) - set(r_decl)), sorted(set(r_decl) - set(c_decl)))

    def n_args(s):
        s = s.strip()
        return 0 if s in ("", "void") else s.count(",") + 1
    bad = [n for n in c_decl if n_args(c_decl[n]) != n_args(r_decl[n])]
    assert not bad, bad
    for const in ("NTC_USE_ENCODE", "NTC_USE_DECODE", "NTC_BLOCK_FULL_LAST_LITERAL_WORD", "NTC_ERR_EMPTY_COLUMN"):
        assert const in rust


@pytest.mark.gpu
def test_one_direction_contexts_and_pinned_buffers():
    """ntc_ctx_create_for: an encode-only context holds no jump table and refuses to decode (and the reverse) with
    NTC_ERR_UNSUPPORTED instead of touching memory it does not have; both produce what a full context produces.
    ntc_pinned_alloc hands out page-locked memory the host-buffer entry points accept like any other pointer."""
    import ctypes as C
    rng = np.random.default_rng(21)
    g = np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, size=20000)]
    idx = nt.Index.build([g], 31)
    reads = [g[p: p + 150].copy() for p in rng.integers(0, 20000 - 150, size=300)]
    bases, off = nt.concat_reads(reads)
    full = nt.Context(idx, 0)
    enc = nt.Context(idx, 0, uses=nt.Context.USE_ENCODE)
    dec = nt.Context(idx, 0, uses=nt.Context.USE_DECODE)
    rec, roff = full.encode_batch(bases, off)
    rec_e, roff_e = enc.encode_batch(bases, off)
    assert (rec == rec_e).all() and (roff == roff_e).all()
    assert enc.index_bytes()["jump"] == 0 and enc.index_bytes()["rank"] > 0      # only the arrays of one direction are resident
    assert dec.index_bytes()["rank"] == 0 and dec.index_bytes()["lcs"] == 0 and dec.index_bytes()["jump"] > 0
    text, n = dec.decode_fasta(rec, 1)
    assert (text, n) == full.decode_fasta(rec, 1)
    with pytest.raises(nt.NtcError) as e:
        enc.decode_fasta(rec, 1)
    assert e.value.code == 9
    with pytest.raises(nt.NtcError) as e:
        dec.encode_batch(bases, off)
    assert e.value.code == 9
    with pytest.raises(nt.NtcError):
        nt.Context(idx, 0, uses=0)
    # pinned buffers through the ABI
    L = _lib.lib()
    p = C.c_void_p()
    _lib.check(L.ntc_pinned_alloc(len(bases), C.byref(p)))
    C.memmove(p, bases.ctypes.data, len(bases))
    out = np.zeros(len(bases), dtype=np.uint64)
    n_rec = C.c_uint64(0)
    _lib.check(L.ntc_encode_batch(full._h, p, off.ctypes.data_as(C.c_void_p), len(reads), out.ctypes.data_as(C.c_void_p), len(out), C.byref(n_rec), None))
    assert n_rec.value == len(rec) and (out[: len(rec)] == rec).all()
    L.ntc_pinned_free(p)
    # every buffer page-locked (the DMA engines copy directly, many chunks) against every buffer pageable (staged by host threads)
    import torch
    os.environ["NTC_MAX_CHUNK_BASES"], os.environ["NTC_MAX_CHUNK_RECORDS"] = "4096", "64"
    try:
        small = nt.Context(idx, 0)
    finally:
        del os.environ["NTC_MAX_CHUNK_BASES"], os.environ["NTC_MAX_CHUNK_RECORDS"]
    pb, po = torch.from_numpy(bases.copy()).pin_memory(), torch.from_numpy(off.astype(np.int64)).pin_memory()
    pr, pro = torch.zeros(len(bases), dtype=torch.int64).pin_memory(), torch.zeros(len(off), dtype=torch.int64).pin_memory()
    pt = torch.zeros(len(bases) + 24 * len(reads) + 64, dtype=torch.uint8).pin_memory()
    n_bytes, n_seqs = C.c_uint64(0), C.c_uint64(0)
    for threads in (-1, 0, 5):
        small.set_host_threads(threads)
        _lib.check(L.ntc_encode_batch(small._h, pb.data_ptr(), po.data_ptr(), len(reads), pr.data_ptr(), pr.numel(), C.byref(n_rec), pro.data_ptr()))
        assert n_rec.value == len(rec) and (pr.numpy()[: len(rec)].view(np.uint64) == rec).all() and (pro.numpy().view(np.uint64) == roff).all()
        _lib.check(L.ntc_decode_fasta(small._h, pr.data_ptr(), n_rec.value, 7, pt.data_ptr(), pt.numel(), C.byref(n_bytes), C.byref(n_seqs)))
        assert bytes(pt.numpy()[: n_bytes.value]) == full.decode_fasta(rec, 7)[0] and n_seqs.value == len(reads)
        rec_s, roff_s = small.encode_batch(bases, off)                  # pageable numpy buffers, same context
        assert (rec_s == rec).all() and (roff_s == roff).all()
        assert small.decode_fasta(rec_s, 7) == full.decode_fasta(rec, 7)
        db, doff = small.decode_batch(rec_s)
        assert (db == bases).all() and (doff == off).all()
