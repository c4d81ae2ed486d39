"""Host-side mirror of ntcomp's lib.rs encode/decode API over the C ABI (include/ntcomp_b200.h).

Names, argument meaning and error behaviour follow the reference:

    encode_sequence(nucleotides, sbwt, lcs)   src/lib.rs:163     -> [(len, range(start, end)), ...]
    encode_dictionary(dictionary, sbwt)       src/encode.rs:129  -> uint64 record words
    decode_sequence(encoding, sbwt)           src/lib.rs:254     -> [bytes, ...]

In the reference `sbwt` / `lcs` are `SbwtIndexVariant` / `LcsArray`; here both live in one
`Context` (the index re-laid-out in one GPU's HBM), which is passed for either argument.
The batched calls (`Context.encode_batch`, `.decode_batch`, `.decode_fasta`) are what a host should
use; the per-read functions above are thin shims over a batch of one, as in INTEGRATION.md.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import NtcError, check


class CodedColumn(C.Structure):
    """ntc_coded_column (include/ntcomp_b200.h)"""
    _fields_ = [("n_values", C.c_uint64), ("n_words", C.c_uint64), ("param", C.c_uint64), ("word_offset", C.c_uint64)]


class CodedBlock(C.Structure):
    """ntc_coded_block: one 65,536-read block as four coded columns (write_block_to minus gzip, src/lib.rs:232-252)"""
    _fields_ = [("n_records", C.c_uint64), ("num_reads", C.c_uint64), ("status", C.c_int32), ("reserved", C.c_int32), ("col", CodedColumn * 4)]


class EncodeError(ValueError):
    """src/encode.rs:38-47: 'invalid input to encode'."""

    def __str__(self):
        return "invalid input to encode"


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _u8(x):
    if isinstance(x, (bytes, bytearray, memoryview)):
        return np.frombuffer(bytes(x), dtype=np.uint8)
    return np.ascontiguousarray(x, dtype=np.uint8)


def concat_reads(reads):
    """list of bytes / uint8 arrays -> (bases uint8[total], offsets uint64[n+1])"""
    arrs = [_u8(r) for r in reads]
    offsets = np.zeros(len(arrs) + 1, dtype=np.uint64)
    if arrs:
        offsets[1:] = np.cumsum([len(a) for a in arrs], dtype=np.uint64)
    bases = np.concatenate(arrs) if arrs else np.zeros(0, dtype=np.uint8)
    return np.ascontiguousarray(bases, dtype=np.uint8), offsets


class Index:
    """Host flat SBWT + LCS (stands in for kbo::build / kbo::index::load_sbwt, src/main.rs:134,149)."""

    def __init__(self, handle):
        self._h = handle
        L = _lib.lib()
        self.k = int(L.ntc_index_k(handle))
        self.n_sets = int(L.ntc_index_n_sets(handle))

    @classmethod
    def build(cls, seqs, k, add_revcomp=True, n_threads=0):
        L = _lib.lib()
        arrs = [_u8(s) for s in seqs]
        n = len(arrs)
        ptrs = (C.c_void_p * max(n, 1))(*[a.ctypes.data for a in arrs])
        lens = np.array([len(a) for a in arrs], dtype=np.uint64)
        h = C.c_void_p()
        check(L.ntc_index_build(C.cast(ptrs, C.c_void_p), _ptr(lens), n, k, int(add_revcomp), n_threads, C.byref(h)))
        return cls(h)

    @classmethod
    def build_gpu(cls, seqs, k, add_revcomp=True, device=0, stats=None):
        """The same construction on the GPU (ntc_index_build_gpu; `ntcomp build`, src/main.rs:97-140).  `stats`, when a
        dict, receives the phase times and counts the call reports."""
        L = _lib.lib()
        arrs = [_u8(s) for s in seqs]
        n = len(arrs)
        ptrs = (C.c_void_p * max(n, 1))(*[a.ctypes.data for a in arrs])
        lens = np.array([len(a) for a in arrs], dtype=np.uint64)
        h = C.c_void_p()
        st = np.zeros(8, dtype=np.float64)
        check(L.ntc_index_build_gpu(device, C.cast(ptrs, C.c_void_p), _ptr(lens), n, k, int(add_revcomp), C.byref(h), _ptr(st)))
        if stats is not None:
            stats.update(upload_ms=float(st[0]), device_ms=float(st[1]), occurrences=int(st[2]), distinct=int(st[3]),
                         sources=int(st[4]), dummies=int(st[5]), radix_passes=int(st[6]), launches=int(st[7]))
        return cls(h)

    @classmethod
    def from_arrays(cls, k, n_sets, rows, lcs):
        L = _lib.lib()
        rows = [np.ascontiguousarray(r, dtype=np.uint64) for r in rows]
        lcs = np.ascontiguousarray(lcs, dtype=np.uint8)
        h = C.c_void_p()
        check(L.ntc_index_from_arrays(k, n_sets, _ptr(rows[0]), _ptr(rows[1]), _ptr(rows[2]), _ptr(rows[3]), _ptr(lcs), C.byref(h)))
        return cls(h)

    @classmethod
    def build_from_files(cls, paths, k, add_revcomp=True, n_threads=0, gpu=None):
        """`ntcomp build` front end (src/main.rs:97-134): every fastX record of every file is one input sequence; gpu = a device
        ordinal runs the construction there (ntc_index_build_gpu)"""
        arr = (C.c_char_p * len(paths))(*[str(p).encode() for p in paths])
        h = C.c_void_p()
        if gpu is None:
            check(_lib.lib().ntc_index_build_from_files(arr, len(paths), k, int(add_revcomp), n_threads, C.byref(h)))
        else:
            check(_lib.lib().ntc_index_build_from_files_gpu(int(gpu), arr, len(paths), k, int(add_revcomp), n_threads, C.byref(h)))
        return cls(h)

    @classmethod
    def load(cls, path):
        h = C.c_void_p()
        check(_lib.lib().ntc_index_load(str(path).encode(), C.byref(h)))
        return cls(h)

    def save(self, path):
        check(_lib.lib().ntc_index_save(self._h, str(path).encode()))

    def C(self):
        out = np.zeros(5, dtype=np.uint64)
        _lib.lib().ntc_index_get_C(self._h, _ptr(out))
        return out

    def rows(self):
        out = []
        for c in range(4):
            a = np.zeros((self.n_sets + 63) // 64, dtype=np.uint64)
            _lib.lib().ntc_index_get_row(self._h, c, _ptr(a))
            out.append(a)
        return out

    def lcs(self):
        a = np.zeros(self.n_sets, dtype=np.uint8)
        _lib.lib().ntc_index_get_lcs(self._h, _ptr(a))
        return a

    def close(self):
        if self._h:
            _lib.lib().ntc_index_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Context:
    """One GPU: index in HBM + workspaces.  Raises NtcError(NTC_ERR_CUDA) when no GPU is present."""

    USE_ENCODE, USE_DECODE = 1, 2

    def __init__(self, index, device=0, uses=3):
        h = C.c_void_p()
        check(_lib.lib().ntc_ctx_create_for(device, index._h, uses, C.byref(h)))
        self._h = h
        self.device = device
        self.k = index.k
        self.n_sets = index.n_sets

    def close(self):
        if getattr(self, "_h", None):
            _lib.lib().ntc_ctx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def index_bytes(self):
        out = np.zeros(3, dtype=np.uint64)
        _lib.lib().ntc_ctx_index_bytes(self._h, _ptr(out))
        return {"rank": int(out[0]), "lcs": int(out[1]), "jump": int(out[2])}

    def layout(self, which):
        """derived device array `which` (0 rank blocks, 1 LCS words, 2 jump table, 3 prefix table) as bytes -- diagnostic"""
        n = C.c_uint64(0)
        check(_lib.lib().ntc_ctx_layout_read(self._h, which, None, 0, C.byref(n)))
        out = np.zeros(n.value, dtype=np.uint8)
        if n.value:
            check(_lib.lib().ntc_ctx_layout_read(self._h, which, _ptr(out), n.value, C.byref(n)))
        return out

    KERNEL_CLASSES = ["pack2bit", "ms", "scan", "emit", "decode_scan", "seq_offsets", "walk", "misc", "column_encode", "column_decode"]

    def set_tile_len(self, tile_len=0):
        """matching-statistics tile length (0 = automatic); results do not depend on it"""
        check(_lib.lib().ntc_ctx_set_tile_len(self._h, int(tile_len)))

    def set_copy_directions(self, serial):
        """host-buffer API: one PCIe direction at a time (True) or results travelling back while input goes up (False)"""
        check(_lib.lib().ntc_ctx_set_copy_directions(self._h, int(bool(serial))))

    def set_host_threads(self, n_threads=-1):
        """host-buffer encode: host threads that pack the reads to 2 bits per base before they cross PCIe (0 = off, < 0 = auto)"""
        check(_lib.lib().ntc_ctx_set_host_threads(self._h, int(n_threads)))

    def encode_file(self, fastx_path, out_path, n_threads=0):
        """`ntcomp encode` (src/main.rs:141-181): fastX file -> encoded.dat; returns the stats dict"""
        st = np.zeros(6, dtype=np.uint64)
        check(_lib.lib().ntc_encode_file(self._h, str(fastx_path).encode(), str(out_path).encode(), n_threads, _ptr(st)))
        return dict(zip(["reads", "bases", "records", "bytes", "blocks", "dropped_blocks"], (int(x) for x in st)))

    def decode_file(self, in_path, out_path, n_threads=0):
        """`ntcomp decode` (src/main.rs:183-211): encoded.dat -> FASTA text; returns the stats dict"""
        st = np.zeros(6, dtype=np.uint64)
        check(_lib.lib().ntc_decode_file(self._h, str(in_path).encode(), str(out_path).encode(), n_threads, _ptr(st)))
        return dict(zip(["sequences", "fasta_bytes", "records", "bytes_read", "blocks"], (int(x) for x in st)))

    def set_profiling(self, on=True):
        check(_lib.lib().ntc_ctx_set_profiling(self._h, int(on)))

    def kernel_times(self, reset=True):
        """{class: (ms, launches)} from CUDA events on the launching stream (synchronises)"""
        ms = np.zeros(len(self.KERNEL_CLASSES), dtype=np.float64)
        ln = np.zeros(len(self.KERNEL_CLASSES), dtype=np.uint64)
        check(_lib.lib().ntc_ctx_kernel_times(self._h, _ptr(ms), _ptr(ln), int(reset)))
        return {n: (float(ms[i]), int(ln[i])) for i, n in enumerate(self.KERNEL_CLASSES)}

    def l2_random_bench(self, buffer_bytes, loads_per_thread=256, dependent=False, n_blocks=148 * 8 * 4):
        out = C.c_double(0)
        check(_lib.lib().ntc_l2_random_bench(self._h, buffer_bytes, loads_per_thread, int(dependent), n_blocks, C.byref(out)))
        return out.value

    def launch_count(self):
        return int(_lib.lib().ntc_ctx_launch_count(self._h))

    # ---- host-buffer API ----
    def encode_batch(self, bases, offsets, cap=None, want_offsets=True):
        """flat_map(encode_dictionary . encode_sequence): returns (records uint64[], rec_offsets uint64[n+1])."""
        bases = _u8(bases)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        n = len(offsets) - 1
        if cap is None:
            cap = max(int(offsets[-1] - offsets[0]), 1)
        rec = np.empty(cap, dtype=np.uint64)
        roff = np.zeros(n + 1, dtype=np.uint64) if want_offsets else None
        n_rec = C.c_uint64(0)
        check(_lib.lib().ntc_encode_batch(self._h, _ptr(bases), _ptr(offsets), n, _ptr(rec), cap, C.byref(n_rec), _ptr(roff)))
        return rec[: n_rec.value], roff

    def encode_sequence_batch(self, bases, offsets):
        """encode_sequence over a batch: returns (triples uint64[n_rec,3], rec_offsets)."""
        bases = _u8(bases)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        n = len(offsets) - 1
        cap = max(int(offsets[-1] - offsets[0]), 1)
        tri = np.empty(3 * cap, dtype=np.uint64)
        roff = np.zeros(n + 1, dtype=np.uint64)
        n_rec = C.c_uint64(0)
        check(_lib.lib().ntc_encode_sequence_batch(self._h, _ptr(bases), _ptr(offsets), n, _ptr(tri), cap, C.byref(n_rec), _ptr(roff)))
        return tri[: 3 * n_rec.value].reshape(-1, 3), roff

    def encode_dictionary(self, triples):
        tri = np.ascontiguousarray(triples, dtype=np.uint64).reshape(-1, 3)
        out = np.empty(max(len(tri), 1), dtype=np.uint64)
        check(_lib.lib().ntc_encode_dictionary(self._h, _ptr(tri), len(tri), _ptr(out)))
        return out[: len(tri)]

    def decode_bound(self, records, first_seq_number=1):
        rec = np.ascontiguousarray(records, dtype=np.uint64)
        nb, ns, fb = C.c_uint64(0), C.c_uint64(0), C.c_uint64(0)
        _lib.lib().ntc_decode_bound(_ptr(rec), len(rec), first_seq_number, C.byref(nb), C.byref(ns), C.byref(fb))
        return nb.value, ns.value, fb.value

    def decode_batch(self, records):
        """decode_sequence: returns (bases uint8[], seq_offsets uint64[n_seqs+1])."""
        rec = np.ascontiguousarray(records, dtype=np.uint64)
        nb, ns, _ = self.decode_bound(rec)
        bases = np.empty(max(nb, 1), dtype=np.uint8)
        soff = np.zeros(ns + 1, dtype=np.uint64)
        n_seqs = C.c_uint64(0)
        check(_lib.lib().ntc_decode_batch(self._h, _ptr(rec), len(rec), _ptr(bases), nb, _ptr(soff), ns, C.byref(n_seqs)))
        return bases[:nb], soff[: n_seqs.value + 1]

    def decode_fasta(self, records, first_seq_number=1):
        """decode_sequence + the writer loop of src/main.rs:201-209: returns (text bytes, n_seqs)."""
        rec = np.ascontiguousarray(records, dtype=np.uint64)
        _, _, fb = self.decode_bound(rec, first_seq_number)
        text = np.empty(max(fb, 1), dtype=np.uint8)
        n_bytes, n_seqs = C.c_uint64(0), C.c_uint64(0)
        check(_lib.lib().ntc_decode_fasta(self._h, _ptr(rec), len(rec), first_seq_number, _ptr(text), fb, C.byref(n_bytes), C.byref(n_seqs)))
        return text[: n_bytes.value].tobytes(), n_seqs.value

    def matching_statistics(self, bases, offsets):
        bases = _u8(bases)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        nb = int(offsets[-1] - offsets[0])
        d = np.zeros(max(nb, 1), dtype=np.uint8)
        s = np.zeros(max(nb, 1), dtype=np.uint32)
        check(_lib.lib().ntc_matching_statistics(self._h, _ptr(bases), _ptr(offsets), len(offsets) - 1, _ptr(d), _ptr(s)))
        return d[:nb], s[:nb]

    # ---- fused column codec (records stay in HBM; row f1) ----
    def encode_columns(self, bases, offsets, block_reads=65536):
        """encode + column codec per block of block_reads reads: returns (blocks: CodedBlock array, words uint64[] copy, n_records)"""
        bases = _u8(bases)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        n = len(offsets) - 1
        cap = max((n + block_reads - 1) // block_reads, 1)
        blocks = (CodedBlock * cap)()
        nb, nw, nr = C.c_uint64(0), C.c_uint64(0), C.c_uint64(0)
        wp = C.c_void_p()
        check(_lib.lib().ntc_encode_columns(self._h, _ptr(bases), _ptr(offsets), n, block_reads, C.cast(blocks, C.c_void_p), cap, C.byref(nb),
                                            C.byref(wp), C.byref(nw), C.byref(nr)))
        words = np.ctypeslib.as_array(C.cast(wp, C.POINTER(C.c_uint64)), shape=(nw.value,)).copy() if nw.value else np.zeros(0, dtype=np.uint64)
        return [blocks[i] for i in range(nb.value)], words, nr.value

    def decode_columns_fasta(self, blocks, words, first_seq_number=1, full_last_literal_word=False):
        """coded columns of a batch of blocks -> FASTA text (decode_block minus gunzip + decode_sequence + writer loop)"""
        arr = (CodedBlock * max(len(blocks), 1))(*blocks)
        words = np.ascontiguousarray(words, dtype=np.uint64)
        tp = C.c_void_p()
        nb, ns, nr = C.c_uint64(0), C.c_uint64(0), C.c_uint64(0)
        check(_lib.lib().ntc_decode_columns_fasta(self._h, C.cast(arr, C.c_void_p), len(blocks), _ptr(words) if len(words) else None, len(words),
                                                  first_seq_number, 1 if full_last_literal_word else 0, C.byref(tp), C.byref(nb), C.byref(ns),
                                                  C.byref(nr)))
        # (C.string_at takes a C int: texts of 2 GiB and more need the array route)
        text = memoryview((C.c_ubyte * nb.value).from_address(tp.value)).tobytes() if nb.value else b""
        return text, ns.value, nr.value

    # ---- device-resident API (raw device pointers as ints, e.g. torch.Tensor.data_ptr()) ----
    def encode_device(self, d_bases, d_offsets, n_reads, n_bases, d_records, cap, d_rec_offsets, d_counts, stream=0):
        check(_lib.lib().ntc_encode_device(self._h, d_bases, d_offsets, n_reads, n_bases, d_records, cap, d_rec_offsets, d_counts, stream))

    def decode_device(self, d_records, n_records, first_seq_number, fasta, d_out, cap, d_seq_offsets, cap_seqs, d_counts, stream=0):
        check(_lib.lib().ntc_decode_device(self._h, d_records, n_records, first_seq_number, int(fasta), d_out, cap, d_seq_offsets, cap_seqs,
                                           d_counts, stream))

    def decode_device_static(self, d_records, n_records, first_seq_number, fasta, d_out, cap, d_seq_offsets, cap_seqs, d_counts, stream=0):
        """decode_device on the library's built-in scratch array: no cudaMalloc / cudaFree inside the call, at most
        static_max_records() records per call, calls on one GPU must not overlap (include/ntcomp_b200.h)."""
        check(_lib.lib().ntc_decode_device_static(self._h, d_records, n_records, first_seq_number, int(fasta), d_out, cap, d_seq_offsets,
                                                  cap_seqs, d_counts, stream))

    @staticmethod
    def static_max_records():
        return int(_lib.lib().ntc_decode_static_max_records())

    def status(self, stream=0):
        check(_lib.lib().ntc_ctx_status(self._h, stream))


# ---- reference-named shims (batch of one) ----

def encode_sequence(nucleotides, sbwt, lcs=None):
    """src/lib.rs:163.  `sbwt` is a Context (it also holds the LCS array)."""
    seq = _u8(nucleotides)
    try:
        tri, _ = sbwt.encode_sequence_batch(seq, np.array([0, len(seq)], dtype=np.uint64))
    except NtcError as e:
        if e.code == 4:  # NTC_ERR_EMPTY_READ
            raise EncodeError() from e
        raise
    return [(int(t[0]), range(int(t[1]), int(t[2]))) for t in tri]


def encode_dictionary(dictionary, sbwt):
    """src/encode.rs:129."""
    if len(dictionary) == 0:
        raise EncodeError()
    tri = np.array([[ln, rg.start, rg.stop] for ln, rg in dictionary], dtype=np.uint64)
    return sbwt.encode_dictionary(tri)


def decode_sequence(encoding, sbwt):
    """src/lib.rs:254."""
    bases, off = sbwt.decode_batch(encoding)
    return [bases[int(off[i]): int(off[i + 1])].tobytes() for i in range(len(off) - 1)]


def block_write(records, num_reads):
    """write_block_to (src/lib.rs:232-252) on the host: one block of encoded.dat (four coded + gzipped columns)"""
    rec = np.ascontiguousarray(records, dtype=np.uint64)
    cap = int(_lib.lib().ntc_block_bound(len(rec)))
    out = np.zeros(cap, dtype=np.uint8)
    nb = C.c_uint64(0)
    check(_lib.lib().ntc_block_write(_ptr(rec), len(rec), int(num_reads), _ptr(out), cap, C.byref(nb)))
    return out[: nb.value].tobytes()


def block_read(data, full_last_literal_word=False):
    """decode_block up to decode_sequence (src/lib.rs:320-363): (records, bytes consumed); consumed == 0 at EOF.
    full_last_literal_word: decode blocks the reference panics on (literal bases a multiple of 31)"""
    buf = np.frombuffer(data, dtype=np.uint8)
    n, used = C.c_uint64(0), C.c_uint64(0)
    cap = 1024
    if len(buf) >= 32 * 3:   # capacity from the flags column's value count (third sub-block header)
        pos = 0
        for q in range(3):
            if pos + 32 > len(buf):
                break
            bs = int(np.frombuffer(buf[pos: pos + 4].tobytes(), dtype=np.uint32)[0])
            if q == 2:
                cap = int(np.frombuffer(buf[pos + 8: pos + 12].tobytes(), dtype=np.uint32)[0]) + 16
            pos += 32 + bs
    rec = np.zeros(cap, dtype=np.uint64)
    check(_lib.lib().ntc_block_read_ex(_ptr(buf) if len(buf) else None, len(buf), _ptr(rec), cap, C.byref(n), C.byref(used),
                                       1 if full_last_literal_word else 0))
    return rec[: n.value].copy(), used.value


def block_assemble(block, words):
    """the host remainder of write_block_to: four gzip members + 32-byte headers of one CodedBlock"""
    words = np.ascontiguousarray(words, dtype=np.uint64)
    n_w = sum(int(c.n_words) for c in block.col)
    cap = 4 * 32 + 4 * 64 + n_w * 9 + 1024
    out = np.zeros(cap, dtype=np.uint8)
    nb = C.c_uint64(0)
    check(_lib.lib().ntc_block_assemble(C.byref(block), _ptr(words) if len(words) else None, _ptr(out), cap, C.byref(nb)))
    return out[: nb.value].tobytes()


def sub_block_assemble(block, column, words):
    """one of the four sub-blocks of block_assemble (0 colex, 1 lengths, 2 flags, 3 literal words): header + gzip member"""
    words = np.ascontiguousarray(words, dtype=np.uint64)
    cap = 32 + 64 + (int(block.col[column].n_words) if 0 <= column < 4 else 0) * 9 + 1024
    out = np.zeros(cap, dtype=np.uint8)
    nb = C.c_uint64(0)
    check(_lib.lib().ntc_sub_block_assemble(C.byref(block), column, _ptr(words) if len(words) else None, _ptr(out), cap, C.byref(nb)))
    return out[: nb.value].tobytes()


def block_disassemble(data, offset=0):
    """the host head of decode_block: (CodedBlock, words, bytes consumed); consumed == 0 at EOF"""
    buf = np.frombuffer(data, dtype=np.uint8)[offset:]
    buf = np.ascontiguousarray(buf)
    cap = 0
    pos = 0
    for q in range(4):                       # encoded_size of the four headers
        if pos + 32 > len(buf):
            break
        bs, _, _, es = (int(x) for x in np.frombuffer(buf[pos: pos + 16].tobytes(), dtype=np.uint32))
        cap += es
        pos += 32 + bs
    words = np.zeros(cap + 8, dtype=np.uint64)
    blk = CodedBlock()
    nw, used = C.c_uint64(0), C.c_uint64(0)
    check(_lib.lib().ntc_block_disassemble(_ptr(buf) if len(buf) else None, len(buf), C.byref(blk), _ptr(words), len(words), C.byref(nw), C.byref(used)))
    return blk, words[: nw.value].copy(), used.value
