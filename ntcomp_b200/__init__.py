"""ntcomp_b200 -- B200-native (sm_100a) encode/decode hot path of tmaklin/ntcomp behind its lib.rs API.

Only what the path needs: `csrc/` (CUDA kernels + the C ABI of include/ntcomp_b200.h), the ctypes
loader (`_lib`), the host-side mirror of the reference interface (`api`), the synthetic inputs of
SURVEY §8d (`synth`) and read sharding across GPUs (`sharding`).
"""
from .api import (CodedBlock, CodedColumn, Context, EncodeError, Index, block_assemble, block_disassemble, block_read, block_write,  # noqa: F401
                  concat_reads, decode_sequence, encode_dictionary, encode_sequence, sub_block_assemble)
from ._lib import NtcError  # noqa: F401
