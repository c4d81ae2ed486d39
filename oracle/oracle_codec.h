// oracle/oracle_codec.h -- TEST INFRASTRUCTURE ONLY.  PARITY UNPINNED (see oracle_index.h).
// CPU restatement of ntcomp's sequence codec (SURVEY §8a A2,A3,A5,A6,A7):
//   literal_*  follow src/lib.rs and src/encode.rs statement by statement (same cost profile:
//              4 x k-character search per extended base, k selects per access_kmer);
//   fast_*     use the O(1) equivalences of SURVEY Appendix C and must return identical results.
#pragma once
#include "oracle_index.h"

namespace orc {

// error codes (negative returns)
enum : int64_t {
    ERR_EMPTY = -1,        // EncodeError: empty read / empty dictionary (lib.rs:224, encode.rs:133)
    ERR_CAPACITY = -2,     // caller buffer too small
    ERR_ABSENT_BASE = -3,  // a position has d == 0: reference underflows at lib.rs:207
    ERR_ASSERT = -4,       // a reference assert!/slice bound would panic
    ERR_LEN24 = -5,        // record length >= 2^24 (lib.rs:226)
};

struct DictEntry { uint64_t len, start, end; };  // (usize, Range<usize>) of lib.rs:167

// encode_sequence, src/lib.rs:163-230
int64_t literal_encode_sequence(const Index& ix, const uint8_t* read, uint64_t n, std::vector<DictEntry>* out);
int64_t fast_encode_sequence(const Index& ix, const uint8_t* read, uint64_t n, std::vector<DictEntry>* out);

// encode_dictionary, src/encode.rs:129-166 (literal bases come from access_kmer)
int64_t literal_encode_dictionary(const Index& ix, const DictEntry* dict, uint64_t n, std::vector<uint64_t>* out);
// same words, literal bases taken from the read slice ending at the record's end position
int64_t fast_encode_read(const Index& ix, const uint8_t* read, uint64_t n, std::vector<uint64_t>* out);

// decode_sequence, src/lib.rs:254-318.  Appends sequences to `bases`, boundaries to `offsets`
// (offsets must start with one entry: the current size of bases).
int64_t literal_decode_sequence(const Index& ix, const uint64_t* rec, uint64_t n, std::vector<uint8_t>* bases,
                                std::vector<uint64_t>* offsets);
int64_t fast_decode_sequence(const Index& ix, const uint64_t* rec, uint64_t n, std::vector<uint8_t>* bases,
                             std::vector<uint64_t>* offsets);

// bitnuc 0.2.11 semantics (encode.rs:154,220,225; lib.rs:303; decode.rs:120,140):
// base j of the slice at bits 2j..2j+1, A0 C1 G2 T3; at most 32 bases.
bool as_2bit(const uint8_t* seq, uint64_t len, uint64_t* out);
bool from_2bit(uint64_t word, uint64_t len, std::vector<uint8_t>* out);

}  // namespace orc
