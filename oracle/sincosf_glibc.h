// oracle/sincosf_glibc.h -- TEST INFRASTRUCTURE.
//
// The reference computes `float a = (float)cos(angle), b = (float)sin(angle)` with `angle` a float
// and `using namespace std` active (src/ORBextractor.cc:33,123-125), i.e. libm cosf/sinf.  The
// result therefore depends on the host libm.  This header restates glibc >= 2.28's sinf/cosf
// (sysdeps/ieee754/flt-32/s_sincosf.h, the ARM optimized-routines algorithm) in the exact
// operation order of glibc 2.39's x86-64 FMA variant (__sinf_fma/__cosf_fma, read from the
// disassembly of /lib/x86_64-linux-gnu/libm.so.6 in this image): every `fma` below is one fused
// operation there, every `*` one rounded multiply.  Constants were read from the binary's
// __sincosf_table.  Pinned exhaustively against the host libm over every float angle in
// [0, 360] degrees by oracle/sweep_sincosf.cc (result recorded in DESIGN.md).
//
// Only |x| < 120 is handled (the path's domain is [0, 2*pi]); larger inputs fall back to libm.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

namespace orbo {

struct SinCosTab {
    double c0, c1, s1, c2, s2, c3, s3, c4;
};
// table[0]: +cos / +sin,  table[1]: -cos (sin coefficients unchanged; sign comes from sign[])
static const SinCosTab kSinCosTab[2] = {
    {0x1p0, -0x1.ffffffd0c621cp-2, -0x1.555545995a603p-3, 0x1.55553e1068f19p-5, 0x1.1107605230bc4p-7,
     -0x1.6c087e89a359dp-10, -0x1.994eb3774cf24p-13, 0x1.99343027bf8c3p-16},
    {-0x1p0, 0x1.ffffffd0c621cp-2, -0x1.555545995a603p-3, -0x1.55553e1068f19p-5, 0x1.1107605230bc4p-7,
     0x1.6c087e89a359dp-10, -0x1.994eb3774cf24p-13, -0x1.99343027bf8c3p-16}};
static const double kSinCosSign[4] = {1.0, -1.0, -1.0, 1.0};
static const double kHpiInv = 0x1.45f306dc9c883p+23;  // 2/pi * 2^24
static const double kHpi = 0x1.921fb54442d18p0;

static inline uint32_t f32_abstop12(float x)
{
    uint32_t u;
    std::memcpy(&u, &x, 4);
    return (u >> 20) & 0x7ff;
}

static inline float sincos_poly(double x, double x2, const SinCosTab& p, int n)
{
    if ((n & 1) == 0) {
        double s1 = std::fma(x2, p.s3, p.s2);
        double x3 = x2 * x;
        double x7 = x3 * x2;
        double s = std::fma(x3, p.s1, x);
        return (float)std::fma(s1, x7, s);
    } else {
        double x4 = x2 * x2;
        double c1 = std::fma(x2, p.c1, p.c0);
        double c2 = std::fma(x2, p.c4, p.c3);
        double x6 = x4 * x2;
        double c = std::fma(x4, p.c2, c1);
        return (float)std::fma(c2, x6, c);
    }
}

static inline float glibc_sinf(float y)
{
    double x = y;
    uint32_t top = f32_abstop12(y);
    if (top < 0x3f4) {  // |y| < pi/4
        if (top < 0x398) return y;  // |y| < 2^-12
        return sincos_poly(x, x * x, kSinCosTab[0], 0);
    }
    if (top < 0x42f) {  // |y| < 120
        double r = x * kHpiInv;
        int n = ((int32_t)r + 0x800000) >> 24;
        x = std::fma(-(double)n, kHpi, x);
        const SinCosTab& p = kSinCosTab[(n & 2) ? 1 : 0];
        double s = kSinCosSign[n & 3];
        return sincos_poly(x * s, x * x, p, n);
    }
    return sinf(y);
}

static inline float glibc_cosf(float y)
{
    double x = y;
    uint32_t top = f32_abstop12(y);
    if (top < 0x3f4) {
        if (top < 0x398) return 1.0f;
        return sincos_poly(x, x * x, kSinCosTab[0], 1);
    }
    if (top < 0x42f) {
        double r = x * kHpiInv;
        int n = ((int32_t)r + 0x800000) >> 24;
        x = std::fma(-(double)n, kHpi, x);
        const SinCosTab& p = kSinCosTab[(n & 2) ? 1 : 0];
        double s = kSinCosSign[n & 3];
        return sincos_poly(x * s, x * x, p, n ^ 1);
    }
    return cosf(y);
}

}  // namespace orbo
