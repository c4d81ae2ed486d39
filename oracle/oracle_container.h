// oracle/oracle_container.h -- TEST INFRASTRUCTURE ONLY.  PARITY UNPINNED (see oracle_index.h).
#pragma once
#include <cstdint>
#include <vector>

namespace orc {

struct Columns {                       // src/encode.rs:175-226
    std::vector<uint64_t> colex;       // u32 of every reference record
    std::vector<uint64_t> len;         // u24 of every reference record
    std::vector<uint64_t> flags;       // byte 7 of every record
    std::vector<uint64_t> literals;    // literal bases re-packed 31 per word
};

struct SubHeader {                     // BlockHeader, src/lib.rs:37-47
    uint32_t block_size = 0, num_records = 0, num_u64 = 0, encoded_size = 0;
    uint64_t rice_param = 0;
    uint8_t bitpacker_exponent = 0;
};

uint64_t rice_log2_b(double p);
bool rice_encode(const std::vector<uint64_t>& ints, std::vector<uint64_t>* words, uint64_t* param);
bool minimal_binary_encode(const std::vector<uint64_t>& ints, std::vector<uint64_t>* words, uint64_t* param);
bool rice_decode(const uint64_t* words, uint64_t n_words, uint64_t n, uint64_t b, std::vector<uint64_t>* out);
bool minimal_binary_decode(const uint64_t* words, uint64_t n_words, uint64_t n, uint64_t max, std::vector<uint64_t>* out);
bool code_column(const std::vector<uint64_t>& ints, bool rice, std::vector<uint64_t>* words, uint64_t* param);

bool split_encoded_dictionary(const uint64_t* rec, uint64_t n, Columns* c);
int zip_block_contents(const Columns& c, std::vector<uint64_t>* rec);

bool write_block(const uint64_t* rec, uint64_t n, uint64_t num_records, std::vector<uint8_t>* sink);
int64_t read_block(const uint8_t* data, uint64_t size, std::vector<uint64_t>* rec, SubHeader hdrs[4],
                   std::vector<uint64_t> inflated[4]);

}  // namespace orc
