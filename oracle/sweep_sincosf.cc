// oracle/sweep_sincosf.cc -- TEST INFRASTRUCTURE.
// Exhaustive pin of oracle/sincosf_glibc.h against the host libm: for EVERY float `deg` in
// [0, 360] (all 1 135 869 953 bit patterns 0x00000000..0x43b40000) form angle = deg * factorPI exactly as
// the reference does (src/ORBextractor.cc:109,123) and compare sinf/cosf bit-for-bit.
// Build: g++ -O2 -mfma -ffp-contract=off -std=c++17 -pthread sweep_sincosf.cc -o _ref/sweep_sincosf
// Usage: sweep_sincosf [stride]   (stride 1 = exhaustive)
#include "sincosf_glibc.h"
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <vector>
#include <atomic>

int main(int argc, char** argv)
{
    const uint32_t last = 0x43b40000u;  // 360.0f
    uint32_t stride = argc > 1 ? (uint32_t)atoi(argv[1]) : 1;
    const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
    unsigned nt = std::thread::hardware_concurrency();
    if (!nt) nt = 1;
    std::atomic<unsigned long long> bad_s{0}, bad_c{0}, total{0};
    std::vector<std::thread> th;
    for (unsigned t = 0; t < nt; t++)
        th.emplace_back([&, t]() {
            unsigned long long bs = 0, bc = 0, n = 0;
            for (uint64_t u = (uint64_t)t * stride; u <= last; u += (uint64_t)nt * stride) {
                uint32_t bits = (uint32_t)u;
                float deg;
                std::memcpy(&deg, &bits, 4);
                volatile float angle = deg * factorPI;
                float s0 = sinf(angle), c0 = cosf(angle);
                float s1 = orbo::glibc_sinf(angle), c1 = orbo::glibc_cosf(angle);
                uint32_t a, b;
                std::memcpy(&a, &s0, 4); std::memcpy(&b, &s1, 4); bs += (a != b);
                std::memcpy(&a, &c0, 4); std::memcpy(&b, &c1, 4); bc += (a != b);
                n++;
            }
            bad_s += bs; bad_c += bc; total += n;
        });
    for (auto& x : th) x.join();
    printf("angles swept: %llu  sinf mismatches: %llu  cosf mismatches: %llu  (threads %u, stride %u)\n",
           (unsigned long long)total, (unsigned long long)bad_s, (unsigned long long)bad_c, nt, stride);
    return (bad_s || bad_c) ? 1 : 0;
}
