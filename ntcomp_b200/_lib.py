"""ctypes loader for libntcomp_b200.so (the C ABI declared in include/ntcomp_b200.h).

The library is built in-tree (ntcomp_b200/libntcomp_b200.so) by `__graft_entry__.build()` /
`make -C ntcomp_b200/csrc`.  There is no fallback: a missing library raises ImportError-like
RuntimeError, and compute calls fail with NTC_ERR_CUDA when no GPU is present.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("NTC_LIB_PATH") or os.path.join(_HERE, "libntcomp_b200.so")   # override: kernel-variant experiments

NTC_OK = 0
ERROR_NAMES = {
    1: "NTC_ERR_INVALID_ARG", 2: "NTC_ERR_CUDA", 3: "NTC_ERR_CAPACITY", 4: "NTC_ERR_EMPTY_READ", 5: "NTC_ERR_BAD_BASE",
    6: "NTC_ERR_LEN24", 7: "NTC_ERR_IO", 8: "NTC_ERR_BAD_RECORD", 9: "NTC_ERR_UNSUPPORTED", 10: "NTC_ERR_EMPTY_COLUMN",
}

# every symbol include/ntcomp_b200.h declares: (restype, argtypes)
_vp, _u64, _u32, _i32 = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int32
_pp = C.POINTER(C.c_void_p)
SIGNATURES = {
    "ntc_last_error": (C.c_char_p, []),
    "ntc_version": (C.c_char_p, []),
    "ntc_index_build": (_i32, [_vp, _vp, _u64, _u32, _i32, _i32, _pp]),
    "ntc_index_build_gpu": (_i32, [_i32, _vp, _vp, _u64, _u32, _i32, _pp, _vp]),
    "ntc_index_from_arrays": (_i32, [_u32, _u64, _vp, _vp, _vp, _vp, _vp, _pp]),
    "ntc_index_save": (_i32, [_vp, C.c_char_p]),
    "ntc_index_load": (_i32, [C.c_char_p, _pp]),
    "ntc_index_k": (_u32, [_vp]),
    "ntc_index_n_sets": (_u64, [_vp]),
    "ntc_index_get_C": (None, [_vp, _vp]),
    "ntc_index_get_row": (None, [_vp, _i32, _vp]),
    "ntc_index_get_lcs": (None, [_vp, _vp]),
    "ntc_index_free": (None, [_vp]),
    "ntc_ctx_create": (_i32, [_i32, _vp, _pp]),
    "ntc_ctx_create_for": (_i32, [_i32, _vp, _u32, _pp]),
    "ntc_pinned_alloc": (_i32, [_u64, _pp]),
    "ntc_pinned_free": (None, [_vp]),
    "ntc_ctx_bind_thread": (_i32, [_vp]),
    "ntc_ctx_destroy": (None, [_vp]),
    "ntc_ctx_index_bytes": (None, [_vp, _vp]),
    "ntc_ctx_layout_read": (_i32, [_vp, _i32, _vp, _u64, _vp]),
    "ntc_encode_batch": (_i32, [_vp, _vp, _vp, _u64, _vp, _u64, _vp, _vp]),
    "ntc_encode_sequence_batch": (_i32, [_vp, _vp, _vp, _u64, _vp, _u64, _vp, _vp]),
    "ntc_encode_dictionary": (_i32, [_vp, _vp, _u64, _vp]),
    "ntc_encode_bound": (_u64, [_u64]),
    "ntc_decode_batch": (_i32, [_vp, _vp, _u64, _vp, _u64, _vp, _u64, _vp]),
    "ntc_decode_fasta": (_i32, [_vp, _vp, _u64, _u64, _vp, _u64, _vp, _vp]),
    "ntc_decode_bound": (None, [_vp, _u64, _u64, _vp, _vp, _vp]),
    "ntc_block_bound": (_u64, [_u64]),
    "ntc_block_write": (_i32, [_vp, _u64, _u64, _vp, _u64, _vp]),
    "ntc_block_read": (_i32, [_vp, _u64, _vp, _u64, _vp, _vp]),
    "ntc_block_read_ex": (_i32, [_vp, _u64, _vp, _u64, _vp, _vp, _i32]),
    "ntc_block_read_gpu": (_i32, [_vp, _vp, _u64, _vp, _u64, _vp, _vp, _i32]),
    "ntc_block_write_gpu": (_i32, [_vp, _vp, _u64, _u64, _vp, _u64, _vp]),
    "ntc_column_code": (_i32, [_vp, _u64, _i32, _vp, _u64, _vp, _vp]),
    "ntc_encode_columns": (_i32, [_vp, _vp, _vp, _u64, _u64, _vp, _u64, _vp, _vp, _vp, _vp]),
    "ntc_block_assemble": (_i32, [_vp, _vp, _vp, _u64, _vp]),
    "ntc_sub_block_assemble": (_i32, [_vp, _i32, _vp, _vp, _u64, _vp]),
    "ntc_block_disassemble": (_i32, [_vp, _u64, _vp, _vp, _u64, _vp, _vp]),
    "ntc_decode_columns_fasta": (_i32, [_vp, _vp, _u64, _vp, _u64, _u64, _i32, _vp, _vp, _vp, _vp]),
    "ntc_encode_device": (_i32, [_vp, _vp, _vp, _u64, _u64, _vp, _u64, _vp, _vp, _vp]),
    "ntc_decode_device": (_i32, [_vp, _vp, _u64, _u64, _i32, _vp, _u64, _vp, _u64, _vp, _vp]),
    "ntc_decode_device_static": (_i32, [_vp, _vp, _u64, _u64, _i32, _vp, _u64, _vp, _u64, _vp, _vp]),
    "ntc_decode_static_max_records": (_u64, []),
    "ntc_ctx_status": (_i32, [_vp, _vp]),
    "ntc_ctx_launch_count": (_u64, [_vp]),
    "ntc_ctx_set_profiling": (_i32, [_vp, _i32]),
    "ntc_encode_file": (_i32, [_vp, C.c_char_p, C.c_char_p, _i32, _vp]),
    "ntc_decode_file": (_i32, [_vp, C.c_char_p, C.c_char_p, _i32, _vp]),
    "ntc_fastx_load": (_i32, [C.c_char_p, _i32, _vp, _vp, _vp, _vp]),
    "ntc_free": (None, [_vp]),
    "ntc_index_build_from_files": (_i32, [_vp, _u64, _u32, _i32, _i32, _pp]),
    "ntc_index_build_from_files_gpu": (_i32, [_i32, _vp, _u64, _u32, _i32, _i32, _pp]),
    "ntc_ctx_set_tile_len": (_i32, [_vp, _u32]),
    "ntc_ctx_set_copy_directions": (_i32, [_vp, _i32]),
    "ntc_ctx_set_host_threads": (_i32, [_vp, _i32]),
    "ntc_ctx_kernel_times": (_i32, [_vp, _vp, _vp, _i32]),
    "ntc_l2_random_bench": (_i32, [_vp, _u64, _u32, _i32, _u32, _vp]),
    "ntc_matching_statistics": (_i32, [_vp, _vp, _vp, _u64, _vp, _vp]),
    "ntc_synth_genome": (None, [_u64, _u64, _vp]),
    "ntc_synth_reads": (None, [_vp, _u64, _u64, _u64, _u64, _u32, _u32, _vp, _i32]),
    "ntc_synth_reads_device": (_i32, [_vp, _vp, _u64, _u64, _u64, _u64, _u32, _u32, _vp, _vp]),
}

_LIB = None


class NtcError(RuntimeError):
    def __init__(self, code, message):
        super().__init__("%s (%d): %s" % (ERROR_NAMES.get(code, "NTC_ERR"), code, message))
        self.code = code


def lib():
    """Loads the shared library (once) and applies the prototypes."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "ntcomp_b200: %s is missing -- build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "or `make -C ntcomp_b200/csrc`; there is no CPU fallback" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError here means the .so does not export a declared symbol
            fn.restype = res
            fn.argtypes = args
        _LIB = L
    return _LIB


def check(rc):
    if rc != NTC_OK:
        raise NtcError(rc, lib().ntc_last_error().decode("utf-8", "replace"))
