// host_pack.h -- ASCII <-> 2-bit conversion on host threads (host_pack.cpp) for the host-buffer entry points.
#pragma once
#include <cstdint>
#include <functional>

namespace ntc {

// persistent worker pool: run(n, f) calls f(0) .. f(n-1) on the pool's threads and the caller, returns when all are done
class HostPool {
public:
    explicit HostPool(int n_threads);
    ~HostPool();
    HostPool(const HostPool&) = delete;
    HostPool& operator=(const HostPool&) = delete;
    void run(uint64_t n_items, const std::function<void(uint64_t)>& f);
    int threads() const { return n_; }
private:
    struct Impl;
    Impl* impl_;
    int n_;
};

// n_bases ASCII bases -> ceil(n_bases / 32) words in pack2bit_kernel's layout (base j at bits 2(j % 32) of word j / 32, A0 C1 G2 T3; the
// last word's unused bits are 0).  Returns true when some byte is not A, C, G or T (its code is then unspecified).
bool host_pack2bit(HostPool& pool, const uint8_t* src, uint64_t n_bases, uint64_t* dst);
// the inverse: n_bases bases of a packed stream -> ASCII
void host_unpack2bit(HostPool& pool, const uint64_t* src, uint64_t n_bases, uint8_t* dst);

}  // namespace ntc
