// er_libm.cuh — the six libm functions the path calls, with the bits of the reference's host libm.
//
// The reference reaches libm through sycl::{sin,cos,acos,atan2,pow,log} (src/sycl.h:8; call sites
// src/Texture.cpp:262-292, src/HDRI.cpp:100-117, src/Sampling.h:20-53, src/Disney.cpp:160-170,
// src/kernel.cpp:152-153). Compiled for the host (the oracle) those are glibc's float functions.
// CUDA's own sinf/acosf/... differ from them by an ulp in 1-20 % of calls; a one-ulp change of a bounce
// direction is enough to flip a self-intersection test on flat geometry (t = +-1e-9), so a pixel could
// leave the 1e-3 radiance tolerance through libm alone. These restatements remove that source of
// divergence: they follow the PUBLISHED algorithms glibc 2.39 (the image's libm, third-party to the
// reference) uses for binary32 —
//   sinf / cosf   ARM Optimized Routines' sincosf: double-precision reduction by pi/2 (|x| < 120) and a
//                 double polynomial per quadrant parity, rounded once
//   logf / powf   ARM Optimized Routines: 16-entry (1/c, log c) tables, double polynomials, exp2 by a
//                 32-entry table; the log tables are log(1/invc) rounded to double, the exp2 table is
//                 2^(i/32) rounded to double (tools/gen in tests/cpp/libm_check.cpp re-derives both)
//   acosf / atan2f / atanf   the fdlibm binary32 routines (rational approximation / 4-interval reduction
//                 + odd-even polynomial), all in float arithmetic
// — and are PINNED against the host's libm itself: tests/cpp/libm_check.cpp compiles this very file for
// the host and compares it with glibc over 10^8-10^9 arguments per function (bit-identical for acosf,
// atanf, atan2f, logf; sinf/cosf/powf differ in < 2e-8 of arguments, where glibc's FMA build rounds a
// double intermediate the other way); tests/test_gpu_units.py::test_libm_bits does the same for the
// device code through er_test_libm. Everything is compiled without FMA contraction (-fmad=false on the
// device, -ffp-contract=off on the host).
//
// Outside the argument ranges the path produces (|x| >= 120 for sin/cos, x <= 0 or non-finite arguments
// for pow/log) the functions fall back to the toolchain's own libm.
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#ifdef __CUDACC__
#ifdef ERLM_INLINE
#define ERLM_FN __device__ __forceinline__
#else
#define ERLM_FN __device__ __noinline__
#endif
#define ERLM_INL __device__ __forceinline__
#define ERLM_TAB __device__ __constant__ const
#else
#define ERLM_FN static inline
#define ERLM_INL static inline
#define ERLM_TAB static const
#endif

ERLM_INL float erlm_f(uint32_t u) {
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float f;
    memcpy(&f, &u, 4);
    return f;
#endif
}
ERLM_INL uint32_t erlm_u(float f) {
#ifdef __CUDA_ARCH__
    return __float_as_uint(f);
#else
    uint32_t u;
    memcpy(&u, &f, 4);
    return u;
#endif
}
ERLM_INL double erlm_d(uint64_t u) {
#ifdef __CUDA_ARCH__
    return __longlong_as_double((long long)u);
#else
    double f;
    memcpy(&f, &u, 8);
    return f;
#endif
}
ERLM_INL uint64_t erlm_u64(double f) {
#ifdef __CUDA_ARCH__
    return (uint64_t)__double_as_longlong(f);
#else
    uint64_t u;
    memcpy(&u, &f, 8);
    return u;
#endif
}

// ---------------------------------------------------------------------------------------------------
// sinf / cosf
// ---------------------------------------------------------------------------------------------------
// n even: sine polynomial of the reduced argument; n odd: cosine polynomial, negated when `neg`. (The published
// routine keeps a second coefficient table with every cosine coefficient negated; negating the result instead is the
// same number: rounding to nearest is symmetric in sign for every product and sum of the evaluation.)
ERLM_INL double erlm_sincos_poly(double x, double x2, bool neg, int n) {
    if ((n & 1) == 0) {
        const double s1 = -0x1.555545995a603p-3, s2 = 0x1.1107605230bc4p-7, s3 = -0x1.994eb3774cf24p-13;
        const double x3 = x * x2;
        const double t = s2 + x2 * s3;
        const double x7 = x3 * x2;
        const double s = x + x3 * s1;
        return s + x7 * t;
    }
    const double c0 = 0x1p0, c1 = -0x1.ffffffd0c621cp-2, c2 = 0x1.55553e1068f19p-5, c3 = -0x1.6c087e89a359dp-10,
                 c4 = 0x1.99343027bf8c3p-16;
    const double x4 = x2 * x2;
    const double d2 = c3 + x2 * c4;
    const double d1 = c0 + x2 * c1;
    const double x6 = x4 * x2;
    const double c = d1 + x4 * c2;
    const double r = c + x6 * d2;
    return neg ? -r : r;
}
ERLM_INL uint32_t erlm_top12(float f) { return (erlm_u(f) >> 20) & 0x7ffu; }

// which = 0: sine, 1: cosine
ERLM_INL float erlm_sincosf(float y, int which) {
    double x = (double)y;
    if (erlm_top12(y) < erlm_top12(0x1.921FB6p-1f)) {  // |y| < pi/4
        const double x2 = x * x;
        if (erlm_top12(y) < erlm_top12(0x1p-12f)) return which ? 1.0f : y;
        return (float)erlm_sincos_poly(x, x2, false, which);
    }
    if (erlm_top12(y) < erlm_top12(120.0f)) {
        const double r = x * 0x1.45F306DC9C883p+23;  // 2/pi * 2^24: the quadrant lands in bits 24..31
        const int n = ((int32_t)r + 0x800000) >> 24;
        x = x - (double)n * 0x1.921FB54442D18p0;
        const double xs = ((n + 1) & 2) ? -x : x;  // sign of the sine in quadrants 0..3: + - - +
        return (float)erlm_sincos_poly(xs, x * x, (n & 2) != 0, n ^ which);
    }
    return which ? cosf(y) : sinf(y);  // never on the path (arguments are angles within [-pi, 2 pi])
}
ERLM_FN float erlm_sinf(float x) { return erlm_sincosf(x, 0); }
ERLM_FN float erlm_cosf(float x) { return erlm_sincosf(x, 1); }

// ---------------------------------------------------------------------------------------------------
// acosf (fdlibm binary32: R(z) = P(z)/Q(z) on |x| < 0.5, 2 asin(sqrt((1-|x|)/2)) outside)
// ---------------------------------------------------------------------------------------------------
ERLM_INL float erlm_acos_r(float z) {
    const float pS0 = 1.6666667163e-01f, pS1 = -3.2556581497e-01f, pS2 = 2.0121252537e-01f, pS3 = -4.0055535734e-02f,
                pS4 = 7.9153501429e-04f, pS5 = 3.4793309169e-05f, qS1 = -2.4033949375e+00f, qS2 = 2.0209457874e+00f,
                qS3 = -6.8828397989e-01f, qS4 = 7.7038154006e-02f;
    const float p = z * (pS0 + z * (pS1 + z * (pS2 + z * (pS3 + z * (pS4 + z * pS5)))));
    const float q = 1.0f + z * (qS1 + z * (qS2 + z * (qS3 + z * qS4)));
    return p / q;
}
ERLM_FN float erlm_acosf(float x) {
    const float pi = 3.1415925026e+00f, pio2_hi = 1.5707962513e+00f, pio2_lo = 7.5497894159e-08f;
    const int32_t hx = (int32_t)erlm_u(x), ix = hx & 0x7fffffff;
    if (ix == 0x3f800000) return hx > 0 ? 0.0f : pi + 2.0f * pio2_lo;
    if (ix > 0x3f800000) return (x - x) / (x - x);
    if (ix < 0x3f000000) {
        if (ix <= 0x32800000) return pio2_hi + pio2_lo;
        const float r = erlm_acos_r(x * x);
        return pio2_hi - (x - (pio2_lo - x * r));
    }
    if (hx < 0) {
        const float z = (1.0f + x) * 0.5f;
        const float r = erlm_acos_r(z);
        const float s = sqrtf(z);
        const float w = r * s - pio2_lo;
        return pi - 2.0f * (s + w);
    }
    const float z = (1.0f - x) * 0.5f;
    const float s = sqrtf(z);
    const float df = erlm_f(erlm_u(s) & 0xfffff000u);
    const float c = (z - df * df) / (s + df);
    const float r = erlm_acos_r(z);
    const float w = r * s + c;
    return 2.0f * (df + w);
}

// ---------------------------------------------------------------------------------------------------
// atanf / atan2f (fdlibm binary32)
// ---------------------------------------------------------------------------------------------------
ERLM_INL float erlm_atanf(float x) {
    const float hi0 = 4.6364760399e-01f, hi1 = 7.8539812565e-01f, hi2 = 9.8279368877e-01f, hi3 = 1.5707962513e+00f;
    const float lo0 = 5.0121582440e-09f, lo1 = 3.7748947079e-08f, lo2 = 3.4473217170e-08f, lo3 = 7.5497894159e-08f;
    const float a0 = 3.3333334327e-01f, a1 = -2.0000000298e-01f, a2 = 1.4285714924e-01f, a3 = -1.1111110449e-01f,
                a4 = 9.0908870101e-02f, a5 = -7.6918758452e-02f, a6 = 6.6610731184e-02f, a7 = -5.8335702866e-02f,
                a8 = 4.9768779427e-02f, a9 = -3.6531571299e-02f, a10 = 1.6285819933e-02f;
    const int32_t hx = (int32_t)erlm_u(x), ix = hx & 0x7fffffff;
    int id;
    float hi = 0.f, lo = 0.f;
    if (ix >= 0x4c000000) {  // |x| >= 2^25
        if (ix > 0x7f800000) return x + x;
        return hx > 0 ? hi3 + lo3 : -hi3 - lo3;
    }
    if (ix < 0x3ee00000) {  // |x| < 7/16
        if (ix < 0x31000000) return x;
        id = -1;
    } else {
        x = fabsf(x);
        if (ix < 0x3f980000) {
            if (ix < 0x3f300000) {
                id = 0, hi = hi0, lo = lo0;
                x = (2.0f * x - 1.0f) / (2.0f + x);
            } else {
                id = 1, hi = hi1, lo = lo1;
                x = (x - 1.0f) / (x + 1.0f);
            }
        } else if (ix < 0x401c0000) {
            id = 2, hi = hi2, lo = lo2;
            x = (x - 1.5f) / (1.0f + 1.5f * x);
        } else {
            id = 3, hi = hi3, lo = lo3;
            x = -1.0f / x;
        }
    }
    const float z = x * x;
    const float w = z * z;
    const float s1 = z * (a0 + w * (a2 + w * (a4 + w * (a6 + w * (a8 + w * a10)))));
    const float s2 = w * (a1 + w * (a3 + w * (a5 + w * (a7 + w * a9))));
    if (id < 0) return x - x * (s1 + s2);
    const float r = hi - ((x * (s1 + s2) - lo) - x);
    return hx < 0 ? -r : r;
}
ERLM_FN float erlm_atan2f(float y, float x) {
    const float tiny = 1.0e-30f, pi_o_4 = 7.8539818525e-01f, pi_o_2 = 1.5707963705e+00f, pi = 3.1415927410e+00f,
                pi_lo = -8.7422776573e-08f;
    const int32_t hx = (int32_t)erlm_u(x), hy = (int32_t)erlm_u(y);
    const int32_t ix = hx & 0x7fffffff, iy = hy & 0x7fffffff;
    if (ix > 0x7f800000 || iy > 0x7f800000) return x + y;
    if (hx == 0x3f800000) return erlm_atanf(y);
    const int m = ((hy >> 31) & 1) | ((hx >> 30) & 2);  // 2 sign(x) + sign(y)
    if (iy == 0) return m < 2 ? y : (m == 2 ? pi + tiny : -pi - tiny);
    if (ix == 0) return hy < 0 ? -pi_o_2 - tiny : pi_o_2 + tiny;
    if (ix == 0x7f800000) {
        if (iy == 0x7f800000)
            return m == 0 ? pi_o_4 + tiny : m == 1 ? -pi_o_4 - tiny : m == 2 ? 3.0f * pi_o_4 + tiny : -3.0f * pi_o_4 - tiny;
        return m == 0 ? 0.0f : m == 1 ? -0.0f : m == 2 ? pi + tiny : -pi - tiny;
    }
    if (iy == 0x7f800000) return hy < 0 ? -pi_o_2 - tiny : pi_o_2 + tiny;
    const int k = (iy - ix) >> 23;
    float z;
    if (k > 60)
        z = pi_o_2 + 0.5f * pi_lo;
    else if (hx < 0 && k < -60)
        z = 0.0f;
    else
        z = erlm_atanf(fabsf(y / x));
    switch (m) {
        case 0: return z;
        case 1: return erlm_f(erlm_u(z) ^ 0x80000000u);
        case 2: return pi - (z - pi_lo);
        default: return (z - pi_lo) - pi;
    }
}

// ---------------------------------------------------------------------------------------------------
// logf / powf
// ---------------------------------------------------------------------------------------------------
// 1/c for the 16 sub-intervals of [0x1.66p-1, 0x1.66p0), shared by logf and powf
ERLM_TAB double erlm_invc[16] = {
    0x1.661ec79f8f3bep+0, 0x1.571ed4aaf883dp+0, 0x1.49539f0f010bp+0,  0x1.3c995b0b80385p+0,
    0x1.30d190c8864a5p+0, 0x1.25e227b0b8eap+0,  0x1.1bb4a4a1a343fp+0, 0x1.12358f08ae5bap+0,
    0x1.0953f419900a7p+0, 0x1p+0,               0x1.e608cfd9a47acp-1, 0x1.ca4b31f026aap-1,
    0x1.b2036576afce6p-1, 0x1.9c2d163a1aa2dp-1, 0x1.886e6037841edp-1, 0x1.767dcf5534862p-1};
// ln(c) = -ln(invc), rounded to double
ERLM_TAB double erlm_logc[16] = {
    -0x1.57bf7808caadep-2, -0x1.2bef0a7c06ddbp-2, -0x1.01eae7f513a67p-2, -0x1.b31d8a68224e9p-3,
    -0x1.6574f0ac07758p-3, -0x1.1aa2bc79c81p-3,   -0x1.a4e76ce8c0e5ep-4, -0x1.1973c5a611cccp-4,
    -0x1.252f438e10c1ep-5, 0x0p+0,                0x1.aa5aa5df25984p-5,  0x1.c5e53aa362eb4p-4,
    0x1.526e57720db08p-3,  0x1.bc2860d22477p-3,   0x1.1058bc8a07ee1p-2,  0x1.4043057b6ee09p-2};
// log2(c) = -log2(invc), rounded to double (powf scales it by 32)
ERLM_TAB double erlm_log2c[16] = {
    -0x1.efec65b963019p-2, -0x1.b0b6832d4fca4p-2, -0x1.7418b0a1fb77bp-2, -0x1.39de91a6dcf7bp-2,
    -0x1.01d9bf3f2b631p-2, -0x1.97c1d1b3b7afp-3,  -0x1.2f9e393af3c9fp-3, -0x1.960cbbf788d5cp-4,
    -0x1.a6f9db6475fcep-5, 0x0p+0,                0x1.338ca9f24f53dp-4,  0x1.476a9543891bap-3,
    0x1.e840b4ac4e4d2p-3,  0x1.40645f0c6651cp-2,  0x1.88e9c2c1b9ff8p-2,  0x1.ce0a44eb17bccp-2};
// bits of 2^(i/32) minus i << 47
ERLM_TAB uint64_t erlm_exp2t[32] = {
    0x3ff0000000000000ull, 0x3fefd9b0d3158574ull, 0x3fefb5586cf9890full, 0x3fef9301d0125b51ull, 0x3fef72b83c7d517bull,
    0x3fef54873168b9aaull, 0x3fef387a6e756238ull, 0x3fef1e9df51fdee1ull, 0x3fef06fe0a31b715ull, 0x3feef1a7373aa9cbull,
    0x3feedea64c123422ull, 0x3feece086061892dull, 0x3feebfdad5362a27ull, 0x3feeb42b569d4f82ull, 0x3feeab07dd485429ull,
    0x3feea47eb03a5585ull, 0x3feea09e667f3bcdull, 0x3fee9f75e8ec5f74ull, 0x3feea11473eb0187ull, 0x3feea589994cce13ull,
    0x3feeace5422aa0dbull, 0x3feeb737b0cdc5e5ull, 0x3feec49182a3f090ull, 0x3feed503b23e255dull, 0x3feee89f995ad3adull,
    0x3feeff76f2fb5e47ull, 0x3fef199bdd85529cull, 0x3fef3720dcef9069ull, 0x3fef5818dcfba487ull, 0x3fef7c97337b9b5full,
    0x3fefa4afa2a490daull, 0x3fefd0765b6e4540ull};

ERLM_FN float erlm_logf(float x) {
    uint32_t ix = erlm_u(x);
    if (ix == 0x3f800000u) return 0.0f;
    if (ix - 0x00800000u >= 0x7f800000u - 0x00800000u) {  // x < 2^-126, inf or nan
        if (ix * 2u == 0u) return -INFINITY;
        if (ix == 0x7f800000u) return x;
        if ((ix & 0x80000000u) || ix * 2u >= 0xff000000u) return (x - x) / (x - x);
        ix = erlm_u(x * 0x1p23f);  // subnormal: normalise
        ix -= 23u << 23;
    }
    // x = 2^k z, z in [0x1.66p-1, 0x1.66p0)
    const uint32_t tmp = ix - 0x3f330000u;
    const int i = (int)((tmp >> 19) & 15u);
    const int k = (int32_t)tmp >> 23;
    const uint32_t iz = ix - (tmp & 0xff800000u);
    const double z = (double)erlm_f(iz);
    const double r = z * erlm_invc[i] - 1.0;
    const double y0 = erlm_logc[i] + (double)k * 0x1.62e42fefa39efp-1;
    const double r2 = r * r;
    double y = 0x1.5575b0be00b6ap-2 * r + -0x1.ffffef20a4123p-2;
    y = -0x1.00ea348b88334p-2 * r2 + y;
    y = y * r2 + (y0 + r);
    return (float)y;
}

ERLM_FN float erlm_powf(float x, float y) {
    uint32_t ix = erlm_u(x);
    const uint32_t iy = erlm_u(y);
    const bool y_special = 2u * iy - 1u >= 2u * 0x7f800000u - 1u;  // y is 0, inf or nan
    if (ix - 0x00800000u >= 0x7f800000u - 0x00800000u || y_special) {
        if (x == 0.0f && !y_special && !(ix & 0x80000000u)) return (iy & 0x80000000u) ? INFINITY : 0.0f;
        if (y_special || (ix & 0x80000000u) || ix >= 0x7f800000u) return powf(x, y);  // never on the path
        ix = erlm_u(x * 0x1p23f);  // positive subnormal: normalise
        ix &= 0x7fffffffu;
        ix -= 23u << 23;
    }
    // 32 log2(x)
    const uint32_t tmp = ix - 0x3f330000u;
    const int i = (int)((tmp >> 19) & 15u);
    const uint32_t top = tmp & 0xff800000u;
    const uint32_t iz = ix - top;
    const int k = (int32_t)top >> (23 - 5);
    const double z = (double)erlm_f(iz);
    const double r = z * erlm_invc[i] - 1.0;
    const double y0 = erlm_log2c[i] * 32.0 + (double)k;
    const double r2 = r * r;
    double l = (0x1.27616c9496e0bp-2 * 32.0) * r + (-0x1.71969a075c67ap-2 * 32.0);
    const double p = (0x1.ec70a6ca7baddp-2 * 32.0) * r + (-0x1.7154748bef6c8p-1 * 32.0);
    const double r4 = r2 * r2;
    double q = (0x1.71547652ab82bp0 * 32.0) * r + y0;
    q = p * r2 + q;
    l = l * r4 + q;
    const double ylogx = (double)y * l;
    if (((erlm_u64(ylogx) >> 47) & 0xffffu) >= (erlm_u64(126.0 * 32.0) >> 47)) {
        if (ylogx > 0x1.fffffffd1d571p+6 * 32.0) return INFINITY;
        if (ylogx <= -150.0 * 32.0) return 0.0f;
    }
    // 2^(ylogx / 32)
    const double shift = 0x1.8p+52;
    double kd = ylogx + shift;
    const uint64_t ki = erlm_u64(kd);
    kd -= shift;
    const double rr = ylogx - kd;
    const uint64_t t = erlm_exp2t[ki & 31u] + (ki << 47);
    const double s = erlm_d(t);
    const double c0 = 0x1.c6af84b912394p-5 / 32 / 32 / 32, c1 = 0x1.ebfce50fac4f3p-3 / 32 / 32, c2 = 0x1.62e42ff0c52d6p-1 / 32;
    const double zz = c0 * rr + c1;
    const double rr2 = rr * rr;
    double e = c2 * rr + 1.0;
    e = zz * rr2 + e;
    e = e * s;
    return (float)e;
}
