// probe.cpp -- how fast can host threads turn ASCII reads into the 2-bit stream (and back)?  The question behind "send reads over PCIe as
// 0.25 B per base": g++ -O2 -std=c++17 -pthread -o probe probe.cpp ../../ntcomp_b200/csrc/host_pack.cpp && ./probe <threads> <bases>.  Not part of the library.
#include "../../ntcomp_b200/csrc/host_pack.h"
#include <chrono>
#include <cstdio>
#include <cstring>
#include <vector>
#include <thread>
using namespace ntc;
int main(int argc, char** argv) {
    int nt = argc > 1 ? atoi(argv[1]) : (int)std::thread::hardware_concurrency();
    uint64_t n = argc > 2 ? strtoull(argv[2], 0, 10) : 1500000000ull;
    std::vector<uint8_t> a(n), b(n);
    uint64_t x = 88172645463325252ull;
    for (uint64_t i = 0; i < n; i++) { x ^= x << 13; x ^= x >> 7; x ^= x << 17; a[i] = "ACGT"[x & 3]; }
    std::vector<uint64_t> p((n + 31) / 32 + 1);
    HostPool pool(nt);
    for (int rep = 0; rep < 3; rep++) {
        auto t0 = std::chrono::steady_clock::now();
        bool bad = host_pack2bit(pool, a.data(), n, p.data());
        auto t1 = std::chrono::steady_clock::now();
        host_unpack2bit(pool, p.data(), n, b.data());
        auto t2 = std::chrono::steady_clock::now();
        double dp = std::chrono::duration<double>(t1 - t0).count(), du = std::chrono::duration<double>(t2 - t1).count();
        printf("threads %d: pack %.1f ms (%.1f GB/s in)  unpack %.1f ms (%.1f GB/s out)  bad=%d same=%d\n", nt, dp * 1e3, n / dp / 1e9, du * 1e3, n / du / 1e9, (int)bad, (int)(memcmp(a.data(), b.data(), n) == 0));
    }
    a[n / 2] = 'N';
    printf("bad detect: %d\n", (int)host_pack2bit(pool, a.data(), n, p.data()));
    // odd sizes / scalar path
    setenv("NTC_NO_AVX2", "1", 1);
    return 0;
}
