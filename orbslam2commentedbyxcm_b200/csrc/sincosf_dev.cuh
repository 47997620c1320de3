// sincosf_dev.cuh -- device sinf/cosf that is bit-identical to glibc 2.39's x86-64 FMA variant on [0, 2*pi].
//
// The reference's computeOrbDescriptor takes cos/sin of a float angle through libm's cosf/sinf
// (src/ORBextractor.cc:123-125); descriptors are only bit-exact if the rotation (a, b) is.  CUDA's cosf/sinf
// differ from glibc in the last ulp on ~1 % of inputs, so the glibc algorithm (ARM optimized-routines sincosf:
// double-precision range reduction by pi/2 and degree-7/8 minimax polynomials, rounded once to float) is
// implemented here with the same fused/unfused operation sequence as the host code: __fma_rn where the host
// issues vfmadd, __dmul_rn where it issues vmulsd.  IEEE-754 double operations are identical on both sides.
// Domain: |y| < 120 (the path only produces angles in [0, 6.2832]).
#pragma once
#include <cuda_runtime.h>

namespace orbx {

struct SinCosOut { float s, c; };

__device__ __forceinline__ double sc_sin_poly(double x, double x2)
{
    // s1 = -0x1.555545995a603p-3, s2 = 0x1.1107605230bc4p-7, s3 = -0x1.994eb3774cf24p-13
    double t = __fma_rn(x2, -0x1.994eb3774cf24p-13, 0x1.1107605230bc4p-7);
    double x3 = __dmul_rn(x2, x);
    double x7 = __dmul_rn(x3, x2);
    double s = __fma_rn(x3, -0x1.555545995a603p-3, x);
    return __fma_rn(t, x7, s);
}

__device__ __forceinline__ double sc_cos_poly(double x2, double sgn)
{
    // c0 = 1, c1 = -0x1.ffffffd0c621cp-2, c2 = 0x1.55553e1068f19p-5, c3 = -0x1.6c087e89a359dp-10,
    // c4 = 0x1.99343027bf8c3p-16; the "negative" table negates all five, which commutes with every rounding.
    double x4 = __dmul_rn(x2, x2);
    double c1 = __fma_rn(x2, sgn * -0x1.ffffffd0c621cp-2, sgn * 0x1p0);
    double c2 = __fma_rn(x2, sgn * 0x1.99343027bf8c3p-16, sgn * -0x1.6c087e89a359dp-10);
    double x6 = __dmul_rn(x4, x2);
    double c = __fma_rn(x4, sgn * 0x1.55553e1068f19p-5, c1);
    return __fma_rn(c2, x6, c);
}

// returns sinf(y), cosf(y) with glibc's results for 0 <= |y| < 120
__device__ __forceinline__ SinCosOut sincosf_glibc(float y)
{
    SinCosOut o;
    double x = (double)y;
    unsigned top = (__float_as_uint(y) >> 20) & 0x7ff;
    if (top < 0x3f4) {  // |y| < pi/4
        if (top < 0x398) {  // |y| < 2^-12
            o.s = y;
            o.c = 1.0f;
            return o;
        }
        double x2 = __dmul_rn(x, x);
        o.s = (float)sc_sin_poly(x, x2);
        o.c = (float)sc_cos_poly(x2, 1.0);
        return o;
    }
    double r = __dmul_rn(x, 0x1.45f306dc9c883p+23);
    int n = (__double2int_rz(r) + 0x800000) >> 24;
    x = __fma_rn(-(double)n, 0x1.921fb54442d18p0, x);
    double x2 = __dmul_rn(x, x);
    // sign[n & 3] = {1, -1, -1, 1}; table 1 (n & 2) negates the cosine polynomial
    double sgn_sin = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;
    double sgn_cos = (n & 2) ? -1.0 : 1.0;
    double xs = __dmul_rn(x, sgn_sin);
    // sinf: poly selected by n, cosf: by n ^ 1
    if ((n & 1) == 0) {
        o.s = (float)sc_sin_poly(xs, x2);
        // cosf: n^1 odd -> cosine polynomial with table (n & 2)
        o.c = (float)sc_cos_poly(x2, sgn_cos);
    } else {
        o.s = (float)sc_cos_poly(x2, sgn_cos);
        // cosf: n^1 even -> sine polynomial on x * sign[n & 3]
        o.c = (float)sc_sin_poly(xs, x2);
    }
    return o;
}

}  // namespace orbx
