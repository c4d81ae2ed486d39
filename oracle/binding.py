"""ctypes binding of the CPU oracle (oracle/liboracle.so).

TEST INFRASTRUCTURE ONLY -- PARITY UNPINNED (see oracle/oracle_index.h).  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this module;
nothing under ntcomp_b200/ does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

ERRORS = {-1: "EncodeError (empty)", -2: "capacity", -3: "absent base (d == 0)", -4: "reference assert", -5: "len >= 2^24"}


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".cpp", ".h"))]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return so


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(so):
            build()
        L = C.CDLL(so)
        vp, u64, i64, i32 = C.c_void_p, C.c_uint64, C.c_int64, C.c_int
        L.orc_last_error.restype = C.c_char_p
        L.orc_build.restype = vp
        L.orc_build.argtypes = [vp, vp, u64, i32, i32]
        L.orc_from_arrays.restype = vp
        L.orc_from_arrays.argtypes = [i32, u64, vp, vp, vp, vp, vp]
        L.orc_free.argtypes = [vp]
        L.orc_n.restype = u64
        L.orc_n.argtypes = [vp]
        L.orc_k.restype = i32
        L.orc_k.argtypes = [vp]
        L.orc_get_C.argtypes = [vp, vp]
        L.orc_get_row.argtypes = [vp, i32, vp]
        L.orc_get_lcs.argtypes = [vp, vp]
        L.orc_ms.argtypes = [vp, vp, u64, vp, vp, vp]
        L.orc_search.restype = i32
        L.orc_search.argtypes = [vp, vp, u64, vp]
        L.orc_access_kmer.argtypes = [vp, u64, vp]
        L.orc_encode_sequence.restype = i64
        L.orc_encode_sequence.argtypes = [vp, vp, u64, i32, vp, u64]
        L.orc_encode_dictionary.restype = i64
        L.orc_encode_dictionary.argtypes = [vp, vp, u64, vp]
        L.orc_encode_reads.restype = i64
        L.orc_encode_reads.argtypes = [vp, vp, vp, u64, i32, vp, u64, vp, i32]
        L.orc_decode.restype = i64
        L.orc_decode.argtypes = [vp, vp, u64, i32, vp, u64, vp, u64]
        L.orc_code_column.restype = i64
        L.orc_code_column.argtypes = [vp, u64, i32, vp, u64, vp]
        L.orc_decode_column.restype = i64
        L.orc_decode_column.argtypes = [vp, u64, u64, i32, u64, vp]
        L.orc_split.restype = i64
        L.orc_split.argtypes = [vp, u64, vp, vp, vp, vp, vp]
        L.orc_write_block.restype = i64
        L.orc_write_block.argtypes = [vp, u64, u64, vp, u64]
        L.orc_read_block.restype = i64
        L.orc_read_block.argtypes = [vp, u64, vp, u64, vp, vp]
        _LIB = L
    return _LIB


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def _as_u8(x):
    if isinstance(x, (bytes, bytearray)):
        return np.frombuffer(bytes(x), dtype=np.uint8)
    return np.ascontiguousarray(x, dtype=np.uint8)


def concat_reads(reads):
    """list of bytes / uint8 arrays -> (bases uint8[], offsets uint64[n+1])"""
    arrs = [_as_u8(r) for r in reads]
    offsets = np.zeros(len(arrs) + 1, dtype=np.uint64)
    if arrs:
        offsets[1:] = np.cumsum([len(a) for a in arrs], dtype=np.uint64)
    bases = np.concatenate(arrs) if arrs else np.zeros(0, dtype=np.uint8)
    return np.ascontiguousarray(bases), offsets


class OracleError(RuntimeError):
    def __init__(self, code):
        super().__init__("oracle error %d: %s" % (code, ERRORS.get(code, "?")))
        self.code = code


class OracleIndex:
    def __init__(self, handle):
        if not handle:
            raise RuntimeError("oracle index build failed: %s" % lib().orc_last_error().decode())
        self.h = handle
        L = lib()
        self.n = int(L.orc_n(self.h))
        self.k = int(L.orc_k(self.h))

    @classmethod
    def build(cls, seqs, k, add_revcomp=True):
        bases, offsets = concat_reads(seqs)
        if len(bases) == 0:
            bases = np.zeros(1, dtype=np.uint8)
        return cls(lib().orc_build(_ptr(bases), _ptr(offsets), len(offsets) - 1, k, int(add_revcomp)))

    @classmethod
    def from_arrays(cls, k, n, rows, lcs):
        rows = [np.ascontiguousarray(r, dtype=np.uint64) for r in rows]
        lcs = np.ascontiguousarray(lcs, dtype=np.uint8)
        return cls(lib().orc_from_arrays(k, n, _ptr(rows[0]), _ptr(rows[1]), _ptr(rows[2]), _ptr(rows[3]), _ptr(lcs)))

    def __del__(self):
        try:
            if self.h:
                lib().orc_free(self.h)
                self.h = None
        except Exception:
            pass

    # ---- raw arrays ----
    def C(self):
        out = np.zeros(5, dtype=np.uint64)
        lib().orc_get_C(self.h, _ptr(out))
        return out

    def rows(self):
        out = []
        for c in range(4):
            a = np.zeros((self.n + 63) // 64, dtype=np.uint64)
            lib().orc_get_row(self.h, c, _ptr(a))
            out.append(a)
        return out

    def lcs(self):
        a = np.zeros(self.n, dtype=np.uint8)
        lib().orc_get_lcs(self.h, _ptr(a))
        return a

    # ---- primitives ----
    def matching_statistics(self, read):
        r = _as_u8(read)
        d = np.zeros(len(r), dtype=np.uint32)
        s = np.zeros(len(r), dtype=np.uint64)
        e = np.zeros(len(r), dtype=np.uint64)
        lib().orc_ms(self.h, _ptr(r), len(r), _ptr(d), _ptr(s), _ptr(e))
        return d, s, e

    def search(self, pat):
        p = _as_u8(pat)
        out = np.zeros(2, dtype=np.uint64)
        if not lib().orc_search(self.h, _ptr(p), len(p), _ptr(out)):
            return None
        return int(out[0]), int(out[1])

    def access_kmer(self, v):
        out = np.zeros(self.k, dtype=np.uint8)
        lib().orc_access_kmer(self.h, int(v), _ptr(out))
        return out.tobytes()

    # ---- codec ----
    def encode_sequence(self, read, fast=False):
        r = _as_u8(read)
        cap = max(len(r), 1)
        out = np.zeros(3 * cap, dtype=np.uint64)
        rc = lib().orc_encode_sequence(self.h, _ptr(r), len(r), int(fast), _ptr(out), cap)
        if rc < 0:
            raise OracleError(rc)
        return [(int(out[3 * i]), (int(out[3 * i + 1]), int(out[3 * i + 2]))) for i in range(rc)]

    def encode_dictionary(self, dictionary):
        n = len(dictionary)
        tri = np.zeros(3 * max(n, 1), dtype=np.uint64)
        for i, (ln, (s, e)) in enumerate(dictionary):
            tri[3 * i], tri[3 * i + 1], tri[3 * i + 2] = ln, s, e
        out = np.zeros(max(n, 1), dtype=np.uint64)
        rc = lib().orc_encode_dictionary(self.h, _ptr(tri), n, _ptr(out))
        if rc < 0:
            raise OracleError(rc)
        return out[:rc].copy()

    def encode_reads(self, bases, offsets, fast=False, threads=1):
        bases = _as_u8(bases)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        n = len(offsets) - 1
        cap = max(int(offsets[-1]) - int(offsets[0]), 1)
        rec = np.zeros(cap, dtype=np.uint64)
        roff = np.zeros(n + 1, dtype=np.uint64)
        rc = lib().orc_encode_reads(self.h, _ptr(bases), _ptr(offsets), n, int(fast), _ptr(rec), cap, _ptr(roff), threads)
        if rc < 0:
            raise OracleError(rc)
        return rec[:rc].copy(), roff

    def decode(self, records, fast=False):
        rec = np.ascontiguousarray(records, dtype=np.uint64)
        lens = np.where(((rec >> np.uint64(56)) & np.uint64(2)) == 0, (rec >> np.uint64(32)) & np.uint64(0xFFFFFF),
                        (rec >> np.uint64(58)) & np.uint64(0x3F))
        cap = int(lens.sum()) + 1
        bases = np.zeros(cap, dtype=np.uint8)
        off = np.zeros(len(rec) + 2, dtype=np.uint64)
        rc = lib().orc_decode(self.h, _ptr(rec), len(rec), int(fast), _ptr(bases), cap, _ptr(off), len(rec) + 1)
        if rc < 0:
            raise OracleError(rc)
        off = off[: rc + 1].copy()
        return bases[: int(off[-1])].copy(), off


# ---- container helpers ----
def code_column(ints, rice):
    a = np.ascontiguousarray(ints, dtype=np.uint64)
    cap = 2 * len(a) + 1024
    words = np.zeros(cap, dtype=np.uint64)
    param = C.c_uint64(0)
    rc = lib().orc_code_column(_ptr(a), len(a), int(rice), _ptr(words), cap, C.byref(param))
    if rc < 0:
        return None, None
    return words[:rc].copy(), int(param.value)


def decode_column(words, n, rice, param):
    w = np.ascontiguousarray(words, dtype=np.uint64)
    out = np.zeros(max(n, 1), dtype=np.uint64)
    rc = lib().orc_decode_column(_ptr(w), len(w), n, int(rice), param, _ptr(out))
    if rc < 0:
        return None
    return out[:rc].copy()


def split(records):
    rec = np.ascontiguousarray(records, dtype=np.uint64)
    n = len(rec)
    cols = [np.zeros(max(n, 1), dtype=np.uint64) for _ in range(4)]
    sizes = np.zeros(4, dtype=np.uint64)
    rc = lib().orc_split(_ptr(rec), n, _ptr(cols[0]), _ptr(cols[1]), _ptr(cols[2]), _ptr(cols[3]), _ptr(sizes))
    if rc < 0:
        return None
    return [cols[i][: int(sizes[i])].copy() for i in range(4)]


def write_block(records, num_records):
    rec = np.ascontiguousarray(records, dtype=np.uint64)
    cap = 16 * len(rec) + 4096
    out = np.zeros(cap, dtype=np.uint8)
    rc = lib().orc_write_block(_ptr(rec), len(rec), num_records, _ptr(out), cap)
    if rc < 0:
        return None
    return out[:rc].tobytes()


def read_block(data, offset=0):
    """returns (records, headers[4][6], bytes_consumed) or None at EOF"""
    buf = np.frombuffer(data, dtype=np.uint8)[offset:]
    buf = np.ascontiguousarray(buf)
    cap = 8 * len(buf) + 1024
    rec = np.zeros(cap, dtype=np.uint64)
    n_rec = C.c_uint64(0)
    hdr = np.zeros(24, dtype=np.uint64)
    rc = lib().orc_read_block(_ptr(buf), len(buf), _ptr(rec), cap, C.byref(n_rec), _ptr(hdr))
    if rc == 0:
        return None
    if rc < 0:
        raise RuntimeError("oracle read_block error %d" % rc)
    return rec[: n_rec.value].copy(), hdr.reshape(4, 6), int(rc)
