"""ctypes binding of tests/emul/libntc_emul.so: the product's per-thread kernel bodies run on the CPU.
TEST INFRASTRUCTURE ONLY -- exposes the same methods as ntcomp_b200.Context so the parity cases can
run against either backend."""
import ctypes as C
import os
import subprocess

import numpy as np

import ntcomp_b200 as nt
from ntcomp_b200 import _lib

_HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "emul")
_LIB = None

# status bits (device_common.cuh) -> C-ABI codes, same priority as api.cu::status_to_code
_ST = [(1 << 1, 5), (1 << 0, 4), (1 << 2, 6), (1 << 4, 8), (1 << 3, 3)]


def lib():
    global _LIB
    if _LIB is None:
        _lib.lib()
        subprocess.check_call(["make", "-C", _HERE, "-s"], stderr=subprocess.DEVNULL)
        L = C.CDLL(os.path.join(_HERE, "libntc_emul.so"))
        vp, u64, u32, i32 = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int32
        L.emul_index_create.restype = vp
        L.emul_index_create.argtypes = [vp]
        L.emul_index_free.argtypes = [vp]
        L.emul_set_tile_len.argtypes = [vp, u32]
        L.emul_encode.restype = u32
        L.emul_encode.argtypes = [vp, vp, vp, u64, i32, vp, u64, vp, vp]
        L.emul_ms.restype = u32
        L.emul_ms.argtypes = [vp, vp, vp, u64, vp, vp]
        L.emul_dict_pack.restype = u32
        L.emul_dict_pack.argtypes = [vp, vp, u64, vp]
        L.emul_decode.restype = u32
        L.emul_decode.argtypes = [vp, vp, u64, u64, i32, vp, u64, vp, u64, vp]
        L.emul_layouts_check.restype = i32
        L.emul_layouts_check.argtypes = [vp]
        L.emul_index_build.restype = i32
        L.emul_index_build.argtypes = [vp, vp, u64, u32, i32, C.POINTER(vp), vp]
        _LIB = L
    return _LIB


def layouts_check(index):
    """layout_kernels.cuh bodies in loops against the host builders: 0 = rank blocks (both layouts), prefix table, LCS words and jump
    table identical"""
    return int(lib().emul_layouts_check(index._h))


def index_build(seqs, k, add_revcomp=True, stats=None):
    """build_pipeline.h (the GPU index construction) run with loops on the CPU -> nt.Index"""
    arrs = [np.ascontiguousarray(np.frombuffer(bytes(s), dtype=np.uint8) if isinstance(s, (bytes, bytearray)) else s, dtype=np.uint8) for s in seqs]
    n = len(arrs)
    ptrs = (C.c_void_p * max(n, 1))(*[a.ctypes.data for a in arrs])
    lens = np.array([len(a) for a in arrs], dtype=np.uint64)
    h = C.c_void_p()
    st = np.zeros(7, dtype=np.uint64)
    rc = lib().emul_index_build(C.cast(ptrs, C.c_void_p), _ptr(lens), n, k, int(add_revcomp), C.byref(h), _ptr(st))
    if rc:
        raise nt.NtcError(rc, "emulated index construction failed")
    if stats is not None:
        stats.update(zip(("text_chars", "occurrences", "distinct", "sources", "dummies", "nodes", "sort_passes"), (int(v) for v in st)))
    return nt.Index(h)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _raise(status):
    for bit, code in _ST:
        if status & bit:
            raise nt.NtcError(code, "emulated kernel status 0x%x" % status)


class EmulContext:
    def __init__(self, index, device=0):
        self._h = lib().emul_index_create(index._h)
        self.k, self.n_sets = index.k, index.n_sets
        self._index = index

    def __del__(self):
        try:
            lib().emul_index_free(self._h)
        except Exception:
            pass

    def set_tile_len(self, tile_len=0):
        lib().emul_set_tile_len(self._h, int(tile_len))

    def _encode(self, bases, offsets, mode, cap=None):
        bases = np.ascontiguousarray(bases, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        n = len(offsets) - 1
        if (np.diff(offsets.astype(np.int64)) == 0).any():
            raise nt.NtcError(4, "empty read")
        full = max(int(offsets[-1] - offsets[0]), 1)
        out = np.zeros((3 if mode else 1) * full, dtype=np.uint64)
        roff = np.zeros(n + 1, dtype=np.uint64)
        n_rec = C.c_uint64(0)
        st = lib().emul_encode(self._h, _ptr(bases), _ptr(offsets), n, mode, _ptr(out), cap if cap is not None else full, C.byref(n_rec), _ptr(roff))
        _raise(st)
        return out, roff, n_rec.value

    def encode_batch(self, bases, offsets, cap=None, want_offsets=True):
        out, roff, n = self._encode(bases, offsets, 0, cap)
        return out[:n], roff

    def encode_sequence_batch(self, bases, offsets):
        out, roff, n = self._encode(bases, offsets, 1)
        return out[: 3 * n].reshape(-1, 3), roff

    def encode_dictionary(self, triples):
        tri = np.ascontiguousarray(triples, dtype=np.uint64).reshape(-1, 3)
        out = np.zeros(max(len(tri), 1), dtype=np.uint64)
        _raise(lib().emul_dict_pack(self._h, _ptr(tri), len(tri), _ptr(out)))
        return out[: len(tri)]

    decode_bound = nt.Context.decode_bound

    def decode_batch(self, records):
        rec = np.ascontiguousarray(records, dtype=np.uint64)
        nb, ns, _ = self.decode_bound(rec)
        bases = np.zeros(max(nb, 1) + 72, dtype=np.uint8)[3:]       # deliberately misaligned output buffer
        soff = np.zeros(ns + 1, dtype=np.uint64)
        cnt = np.zeros(2, dtype=np.uint64)
        _raise(lib().emul_decode(self._h, _ptr(rec), len(rec), 1, 0, _ptr(bases), nb, _ptr(soff), ns, _ptr(cnt)))
        return bases[: int(cnt[1])], soff[: int(cnt[0]) + 1]

    def decode_fasta(self, records, first_seq_number=1):
        rec = np.ascontiguousarray(records, dtype=np.uint64)
        _, _, fb = self.decode_bound(rec, first_seq_number)
        text = np.zeros(max(fb, 1) + 64, dtype=np.uint8)
        cnt = np.zeros(2, dtype=np.uint64)
        _raise(lib().emul_decode(self._h, _ptr(rec), len(rec), first_seq_number, 1, _ptr(text), fb, None, 0, _ptr(cnt)))
        return text[: int(cnt[1])].tobytes(), int(cnt[0])

    def matching_statistics(self, bases, offsets):
        bases = np.ascontiguousarray(bases, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        nb = int(offsets[-1] - offsets[0])
        d = np.zeros(max(nb, 1), dtype=np.uint8)
        s = np.zeros(max(nb, 1), dtype=np.uint32)
        lib().emul_ms(self._h, _ptr(bases), _ptr(offsets), len(offsets) - 1, _ptr(d), _ptr(s))
        return d[:nb], s[:nb]
