// er_device.cuh — device functions of the path-tracing hot path (sm_100a).
//
// Each function cites the reference code whose RESULTS it must reproduce; the structure is new:
// implicit-index stackless traversal over 64-B child-pair records, candidates evaluated from split
// 48-B vertex / normal blocks, full hit attributes rebuilt only for the winner.
// Compile with -fmad=false (no contraction on the decision path), default IEEE div/sqrt.
#pragma once
#include "er_math.cuh"
#include "er_scene.h"

// ---------------------------------------------------------------------------------------------
// RNG (src/kernel.cpp:25-47)
// ---------------------------------------------------------------------------------------------
ER_HD uint32_t er_jenkins(uint32_t seed) {
    uint32_t hash = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        hash += (seed >> (i * 8)) & 0xFF;
        hash += (hash << 10);
        hash ^= (hash >> 6);
    }
    hash += (hash << 3);
    hash ^= (hash >> 11);
    hash += (hash << 15);
    return hash;
}
ER_HD uint32_t er_rng_seed(uint32_t idx) { return er_jenkins(idx + 1); }
// xorshift32; float(state)/float(UINT_MAX): the divisor rounds to 2^32, so [0,1] inclusive
ER_HD float er_rng_next(uint32_t& state) {
    state ^= state << 13;
    state ^= state >> 17;
    state ^= state << 5;
    return (float)state / 4294967296.0f;
}

// ---------------------------------------------------------------------------------------------
// Textures (src/Texture.cpp:172-292)
// ---------------------------------------------------------------------------------------------
// Texture::getValueFromCoordinates (Texture.cpp:172-205) on scalars: wrap, mirror negatives, decode 1/2/>=3
// channels. One out-of-line copy: the two signed remainders are ~70 instructions per inlined site and
// the shading code has ~40 sites.
ER_SCALAR_FN float3 er_texel(const float* __restrict__ data, int width, int height, int ch, int x, int y) {
    x %= width;
    y %= height;
    if (x < 0) x *= -1;
    if (y < 0) y *= -1;
    const int p = y * width + x;
    if (ch == 0) return make_float3(0.f, 0.f, 0.f);
    if (ch == 1) {
        const float a = __ldg(data + p);
        return make_float3(a, a, a);
    }
    if (ch == 2) return make_float3(__ldg(data + 2 * p), __ldg(data + 2 * p + 1), 0.f);
    return make_float3(__ldg(data + ch * p), __ldg(data + ch * p + 1), __ldg(data + ch * p + 2));
}
ER_D V3 er_tex_coords(const DTexture& T, int x, int y) {
    const float3 c = er_texel(T.data, T.width, T.height, T.channels, x, y);
    return v3(c.x, c.y, c.z);
}
ER_D V3 er_tex_uv(const DTexture& T, float u, float v) {
    return er_tex_coords(T, er_f2i(u * (float)T.width), er_f2i(v * (float)T.height));
}
ER_D V3 er_tex_bilinear(const DTexture& T, float u, float v) {
    float x = u * (float)T.width;
    float y = v * (float)T.height;
    float t1x = floorf(x), t1y = floorf(y);
    float t2x = t1x + 1, t2y = t1y + 1;
    float a = (x - t1x) / (t2x - t1x);
    float b = (y - t1y) / (t2y - t1y);
    V3 p1 = er_tex_coords(T, er_f2i(t1x), er_f2i(t1y));
    V3 p2 = er_tex_coords(T, er_f2i(t2x), er_f2i(t1y));
    V3 p3 = er_tex_coords(T, er_f2i(t1x), er_f2i(t2y));
    V3 p4 = er_tex_coords(T, er_f2i(t2x), er_f2i(t2y));
    return er_lerp(er_lerp(p1, p2, a), er_lerp(p3, p4, a), b);
}
ER_D V3 er_tex_filtered(const DTexture& T, float u, float v) {
    return T.filter == 1 ? er_tex_bilinear(T, u, v) : er_tex_uv(T, u, v);
}
// Texture::sphericalMapping(Vector3(), point, 1, u, v)  (Texture.cpp:229-242)
ER_D void er_spherical_mapping(V3 point, float& u, float& v) {
    V3 p = (point - v3s(0.f)) / 1.0f;
    float theta = er_acosf(-p.y);
    float phi = er_atan2f(-p.z, p.x) + ER_PIF;
    u = phi / (2 * ER_PIF);
    v = theta / ER_PIF;
    er_limit_uv(u, v);
}
// Texture::inverseTransformUV (Texture.cpp:267-279); returns (nu, nv)
ER_D void er_inverse_transform_uv(const DTexture& T, float u, float v, float& nu, float& nv) {
    int x = er_f2i(u * (float)T.width);
    int y = er_f2i(v * (float)T.height);
    nu = (float)x / (float)T.width;
    nv = (float)y / (float)T.height;
    er_limit_uv(nu, nv);
}
// Texture::reverseSphericalMapping (Texture.cpp:281-292)
ER_D V3 er_reverse_spherical(float u, float v) {
    float phi = u * 2 * ER_PIF;
    float theta = v * ER_PIF;
    float px = er_cosf(phi - ER_PIF);
    float py = -er_cosf(theta);
    float pz = -er_sinf(phi - ER_PIF);
    float a = sqrtf(1 - py * py);
    return v3(a * px, py, a * pz);
}

// ---------------------------------------------------------------------------------------------
// HDRI (src/HDRI.cpp:85-117)
// ---------------------------------------------------------------------------------------------
ER_D int er_hdri_search(const float* __restrict__ arr, float value, int length) {
    int from = 0;
    int to = length - 1;
    while (to - from > 0) {
        int m = from + (to - from) / 2;
        if (m >= length || m < 0) return 0;
        float a = __ldg(arr + m);
        if (value == a) return m;
        if (value < a) to = m - 1;
        if (value > a) from = m + 1;
    }
    return to;
}
ER_D float er_hdri_pdf(const DScene& S, int x, int y) {
    V3 dv = er_tex_coords(S.hdri, x, y);
    float theta = (((float)y / (float)S.hdri.height)) * ER_PIF;
    return ((dv.x + dv.y + dv.z) / S.radianceSum) * (float)S.hdri.width * (float)S.hdri.height /
           (2.0f * ER_PIF * er_sinf(theta));
}
