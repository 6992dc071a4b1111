// er_math.cuh — FP32 vector primitives with the reference's exact operation order.
//
// Every function here restates the arithmetic of the reference's Vector3 / Math.hpp
// (src/Vector.h:97-256, src/Math.hpp:22-67) so that, compiled with -fmad=false and IEEE
// div/sqrt, results are bit-identical to the reference built without FMA contraction.
// NaN ordering matters: er_minf/er_maxf are `a<b?a:b` / `a>b?a:b` (Math.hpp:61-67), NOT fminf.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "er_libm.cuh"

#define ER_PIF 3.14159265358979323846f
#define ER_HD __host__ __device__ __forceinline__
#define ER_D __device__ __forceinline__
// big shading helpers that are called from several sites. -DER_NOINLINE_SHADE keeps one out-of-line copy
// of each (render kernel 183 KB -> 117 KB of SASS); measured 7 % SLOWER on B200 (profiles/README.md), so
// the default inlines them
#ifdef ER_NOINLINE_SHADE
#define ER_SHADE_FN __device__ __noinline__
#else
#define ER_SHADE_FN __device__ __forceinline__
#endif
// scalar-in / scalar-out helpers (libm calls, the texel fetch): out of line they cost a call and no
// local memory, and take ~25 KB of straight-line code out of the render kernel's instruction stream
// (the kernel is instruction-fetch bound in its shading code, profiles/README.md). -DER_INLINE_SCALAR
// inlines them again.
#ifdef ER_INLINE_SCALAR
#define ER_SCALAR_FN __device__ __forceinline__
#else
#define ER_SCALAR_FN __device__ __noinline__
#endif

struct V3 {
    float x, y, z;
};

ER_HD V3 v3(float x, float y, float z) {
    V3 r;
    r.x = x;
    r.y = y;
    r.z = z;
    return r;
}
ER_HD V3 v3s(float s) { return v3(s, s, s); }
ER_HD V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
ER_HD V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
ER_HD V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
ER_HD V3 operator*(float s, V3 a) { return v3(a.x * s, a.y * s, a.z * s); }
ER_HD V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
ER_HD V3 operator/(V3 a, float s) { return v3(a.x / s, a.y / s, a.z / s); }
ER_HD V3 operator/(V3 a, V3 b) { return v3(a.x / b.x, a.y / b.y, a.z / b.z); }
ER_HD V3 operator+(V3 a, float s) { return v3(a.x + s, a.y + s, a.z + s); }
ER_HD V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }  // Vector.h:111 (x * -1 == -x bitwise)

// Vector.h:163-165 — left-to-right sum of products
ER_HD float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
// Vector.h:167-169 — the middle component is the negated difference
ER_HD V3 cross(V3 a, V3 b) {
    return v3((a.y * b.z - a.z * b.y), -(a.x * b.z - a.z * b.x), (a.x * b.y - a.y * b.x));
}
ER_HD float length(V3 a) { return sqrtf(a.x * a.x + a.y * a.y + a.z * a.z); }
// Vector.h:180-183
#if !defined(ER_INLINE_SCALAR)
// one out-of-line copy on the device (sqrt + three IEEE divisions, ~50 instructions per inlined site);
// Ray's constructor (Ray.h:13-17) is the same arithmetic, see er_ray_dir
__device__ __noinline__ float3 er_normalize3(float x, float y, float z) {
    float l = sqrtf(x * x + y * y + z * z);
    if (l == 0) return make_float3(x, y, z);
    return make_float3(x / l, y / l, z / l);
}
ER_D V3 normalized(V3 a) {
    const float3 r = er_normalize3(a.x, a.y, a.z);
    return v3(r.x, r.y, r.z);
}
#else
ER_HD V3 normalized(V3 a) {
    float l = length(a);
    if (l == 0) return a;
    return a / l;
}
#endif
// Vector.h:248-250
ER_HD V3 reflect(V3 v1, V3 v2) { return v1 - 2 * (dot(v1, v2)) * v2; }

ER_HD float er_minf(float a, float b) { return a < b ? a : b; }
ER_HD float er_maxf(float a, float b) { return a > b ? a : b; }
ER_HD float er_clamp(float a, float b, float c) { return a < b ? b : a > c ? c : a; }  // Math.hpp:26-28
ER_HD V3 er_clamp(V3 v, float b, float c) { return v3(er_clamp(v.x, b, c), er_clamp(v.y, b, c), er_clamp(v.z, b, c)); }
// Math.hpp:34-41 with FAST_LERP
ER_HD float er_lerp(float a, float b, float c) { return a + c * (b - a); }
ER_HD V3 er_lerp(V3 a, V3 b, float c) { return v3(er_lerp(a.x, b.x, c), er_lerp(a.y, b.y, c), er_lerp(a.z, b.z, c)); }
// Math.hpp:44-47: u += -(u > 1) + -(u < 0) in int arithmetic, then int -> float
ER_HD void er_limit_uv(float& u, float& v) {
    u += (float)(-(int)(u > 1) + -(int)(u < 0));
    v += (float)(-(int)(v > 1) + -(int)(v < 0));
}
// C++ float -> int conversion of the reference's host build (x86 cvttss2si): truncation toward
// zero, INT_MIN for NaN / out of range.  CUDA's cvt.rzi saturates instead, so restate it.
ER_HD int er_f2i(float f) {
    if (!(f > -2147483904.0f && f < 2147483648.0f)) return (int)0x80000000;
    return (int)f;
}

// libm calls of the shading code, one out-of-line copy each (the inlined range reductions are 100+
// instructions per call site)
// libm with the bits of the reference's host libm (er_libm.cuh); -DER_CUDA_LIBM = CUDA's own functions
#ifdef ER_CUDA_LIBM
ER_SCALAR_FN float er_sinf(float x) { return sinf(x); }
ER_SCALAR_FN float er_cosf(float x) { return cosf(x); }
ER_SCALAR_FN float er_acosf(float x) { return acosf(x); }
ER_SCALAR_FN float er_atan2f(float y, float x) { return atan2f(y, x); }
ER_SCALAR_FN float er_powf(float x, float y) { return powf(x, y); }
ER_SCALAR_FN float er_logf(float x) { return logf(x); }
#else
#define er_sinf erlm_sinf
#define er_cosf erlm_cosf
#define er_acosf erlm_acosf
#define er_atan2f erlm_atan2f
#define er_powf erlm_powf
#define er_logf erlm_logf
#endif
