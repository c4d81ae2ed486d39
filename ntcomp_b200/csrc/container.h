// container.h -- host-side block container helpers shared by container.cpp and the GPU column codec.
#pragma once
#include <cstdint>
#include <vector>

namespace ntc {

struct BlockColumns {                // src/encode.rs:175-226
    std::vector<uint64_t> colex, len, flags, literals;
};
struct CodedColumn {
    std::vector<uint64_t> words;     // big-endian u64 words of the MSB-first bit stream
    uint64_t n_values = 0;
    uint64_t param = 0;              // Rice log2(b) or minimal-binary upper bound
};

uint64_t rice_parameter(uint64_t n_values, uint64_t sum);
void minimal_binary_shape(uint64_t upper, unsigned* l, uint64_t* limit);
bool code_column_host(const uint64_t* v, uint64_t n, int rice, std::vector<uint64_t>* words, uint64_t* param);
bool decode_column_host(const uint64_t* words, uint64_t n_words, uint64_t n, int rice, uint64_t param, std::vector<uint64_t>* out);
void split_records(const uint64_t* rec, uint64_t n, BlockColumns* c);
int zip_columns(const BlockColumns& c, std::vector<uint64_t>* rec, bool full_last_word = false);
void put_sub_block(const CodedColumn& col, uint64_t num_records, std::vector<uint8_t>* sink, bool* ok);
// 0 ok, -1 deflate failed, -2 out too small
int32_t put_sub_block_raw(const uint64_t* words, uint64_t n_words, uint64_t n_values, uint64_t param, uint64_t num_records, uint8_t* out,
                          uint64_t cap, uint64_t* n_bytes);
int64_t inflate_sub_block(const uint8_t* data, uint64_t size, std::vector<uint64_t>* words, uint64_t* n_values, uint64_t* param);
int64_t read_sub_block(const uint8_t* data, uint64_t size, int rice, std::vector<uint64_t>* values);

}  // namespace ntc
