// er_shade.cuh — HitData generation, Disney BRDF, samplers and the camera ray.
// Restates src/kernel.cpp:76-172,371-473, src/Disney.cpp:34-230, src/Sampling.h:20-53,
// src/shader.cpp:6-10 + src/shader.h:11-12 with the reference's operation order.
#pragma once
#include "er_trace.cuh"

struct HitData {  // src/kernel.h:92-121
    float metallic, roughness, clearcoatGloss, clearcoat, anisotropic, eta, transmission, specular,
        specularTint, sheenTint, subsurface, sheen, opacity, ax, ay;
    V3 emission, albedo;
    int triIdx;
    V3 position, normal, gnormal, tangent, bitangent;
};

ER_D V3 a3(const float* p) { return v3(p[0], p[1], p[2]); }

// generateHitData (kernel.cpp:76-172)
ER_D void er_generate_hitdata(const DScene& S, const DMaterial& m, const HitFull& hit, HitData& hd) {
    hd.albedo = m.albedoTextureID < 0 ? a3(m.albedo) : er_tex_filtered(S.textures[m.albedoTextureID], hit.tu, hit.tv);
    hd.opacity = m.opacityTextureID < 0 ? m.opacity : er_tex_filtered(S.textures[m.opacityTextureID], hit.tu, hit.tv).x;
    hd.emission =
        m.emissionTextureID < 0 ? a3(m.emission) : er_tex_filtered(S.textures[m.emissionTextureID], hit.tu, hit.tv);
    // roughness/metallic are raised to 2.2 (kernel.cpp:152-153); constants were raised on the host
    if (m.roughnessTextureID < 0)
        hd.roughness = m.roughness_pow;
    else
        hd.roughness = er_powf(er_tex_filtered(S.textures[m.roughnessTextureID], hit.tu, hit.tv).x, 2.2f);
    if (m.metallicTextureID < 0)
        hd.metallic = m.metallic_pow;
    else
        hd.metallic = er_powf(er_tex_filtered(S.textures[m.metallicTextureID], hit.tu, hit.tv).x, 2.2f);
    hd.transmission =
        m.transmissionTextureID < 0 ? m.transmission : er_tex_filtered(S.textures[m.transmissionTextureID], hit.tu, hit.tv).x;
    if (m.normalTextureID < 0) {
        hd.normal = hit.normal;
    } else {
        V3 ncolor = er_tex_uv(S.textures[m.normalTextureID], hit.tu, hit.tv);
        V3 localNormal = (ncolor * 2.0f) - v3s(1.0f);
        hd.normal = normalized(localNormal.x * hit.tangent - localNormal.y * hit.bitangent + localNormal.z * hit.normal);
    }
    hd.clearcoatGloss = m.clearcoatGloss;
    hd.clearcoat = m.clearcoat;
    hd.anisotropic = m.anisotropic;
    hd.eta = m.eta;
    hd.specular = m.specular;
    hd.specularTint = m.specularTint;
    hd.sheenTint = m.sheenTint;
    hd.subsurface = m.subsurface;
    hd.sheen = m.sheen;
    hd.ax = m.ax;
    hd.ay = m.ay;
    hd.gnormal = hit.gnormal;
    hd.tangent = hit.tangent;
    hd.bitangent = hit.bitangent;
    hd.position = hit.position;
    hd.triIdx = hit.triIdx;
}

// asl_shade with the placeholder bodies (shader.cpp:6-10, shader.h:11-12): slots 0..3 write (1,1,0)
ER_D void er_asl_shade(int id, V3& albedo) {
    albedo = v3s(0.f);  // kernel.cpp:531-533
    if (id >= 0 && id < 4) albedo = v3(1.f, 1.f, 0.f);
}

// ---- Disney helpers (Disney.cpp:34-89) -------------------------------------------------------
ER_D float er_schlick(float u) {
    float m = er_clamp(1.0f - u, 0.0f, 1.0f);
    float m2 = m * m;
    return m2 * m2 * m;
}
ER_D float er_gtr1(float NDotH, float a) {
    if (a >= 1.0f) return (1.0f / ER_PIF);
    float a2 = a * a;
    float t = 1.0f + (a2 - 1.0f) * NDotH * NDotH;
    return (a2 - 1.0f) / (ER_PIF * er_logf(a2) * t);
}
ER_D float er_gtr2_aniso(float NDotH, float HDotX, float HDotY, float ax, float ay) {
    float a = HDotX / ax;
    float b = HDotY / ay;
    float c = a * a + b * b + NDotH * NDotH;
    return 1.0f / (ER_PIF * ax * ay * c * c);
}
ER_D float er_smithg_ggx(float NDotV, float alphaG) {
    float a = alphaG * alphaG;
    float b = NDotV * NDotV;
    return 1.0f / (NDotV + sqrtf(a + b - a * b));
}
ER_D float er_smithg_ggx_aniso(float NDotV, float VDotX, float VDotY, float ax, float ay) {
    float a = VDotX * ax;
    float b = VDotY * ay;
    float c = NDotV;
    return 1.0f / (NDotV + sqrtf(a * a + b * b + c * c));
}

// DisneyPdf (Disney.cpp:97-133)
ER_SHADE_FN float er_disney_pdf(const HitData& hd, V3 V, V3 N, V3 L) {
    V3 H = normalized(L + V);
    float NDotH = fabsf(dot(N, H));
    if (dot(N, L) <= 0.0f) return 1.0f;
    float clearcoatAlpha = er_lerp(0.1f, 0.001f, hd.clearcoatGloss);
    float diffuseRatio = 0.5f * (1.0f - hd.metallic);
    float specularRatio = 1.0f - diffuseRatio;
    float aspect = sqrtf(1.0f - hd.anisotropic * 0.9f);
    float ax = er_maxf(0.001f, hd.roughness / aspect);
    float ay = er_maxf(0.001f, hd.roughness * aspect);
    float pdfGTR2_aniso = er_gtr2_aniso(NDotH, dot(H, hd.tangent), dot(H, hd.bitangent), ax, ay) * NDotH;
    float pdfGTR1 = er_gtr1(NDotH, clearcoatAlpha) * NDotH;
    float ratio = 1.0f / (1.0f + hd.clearcoat);
    float pdfSpec = er_lerp(pdfGTR1, pdfGTR2_aniso, ratio) / (4.0f * fabsf(dot(L, H)));
    float pdfDiff = fabsf(dot(L, N)) * (1.0f / ER_PIF);
    return diffuseRatio * pdfDiff + specularRatio * pdfSpec;
}

// Sampling.h:30-53
ER_D V3 er_cosine_hemisphere(float u1, float u2) {
    V3 dir;
    float r = sqrtf(u1);
    float phi = 2.0f * ER_PIF * u2;
    dir.x = r * er_cosf(phi);
    dir.y = r * er_sinf(phi);
    dir.z = sqrtf(er_maxf(0.0f, 1.0f - dir.x * dir.x - dir.y * dir.y));
    return dir;
}
ER_D V3 er_sample_ggx(float rgh, float r1, float r2) {
    float a = er_maxf(0.001f, rgh);
    float phi = r1 * ER_PIF * 2;
    float cosTheta = sqrtf((1.0f - r2) / (1.0f + (a * a - 1.0f) * r2));
    float sinTheta = er_clamp(sqrtf(1.0f - (cosTheta * cosTheta)), 0.0f, 1.0f);
    float sinPhi = er_sinf(phi);
    float cosPhi = er_cosf(phi);
    return v3(sinTheta * cosPhi, sinTheta * sinPhi, cosTheta);
}

// DisneySample (Disney.cpp:136-158)
ER_D V3 er_disney_sample(const HitData& hd, V3 V, V3 N, float r1, float r2, float r3) {
    V3 T = hd.tangent, B = hd.bitangent;
    float diffuseRatio = 0.5f * (1.0f - hd.metallic);
    if (r3 < diffuseRatio) {
        V3 H = er_cosine_hemisphere(r1, r2);
        H = T * H.x + B * H.y + N * H.z;
        return H;
    }
    V3 H = er_sample_ggx(hd.roughness, r1, r2);
    H = T * H.x + B * H.y + N * H.z;
    return reflect(-1.0f * V, H);
}

// DisneyEval (Disney.cpp:160-230)
ER_SHADE_FN V3 er_disney_eval(const HitData& hd, V3 V, V3 N, V3 L) {
    V3 T = hd.tangent, B = hd.bitangent;
    V3 H = normalized(L + V);
    float NDotL = fabsf(dot(N, L));
    float NDotV = fabsf(dot(N, V));
    float NDotH = fabsf(dot(N, H));
    float LDotH = fabsf(dot(L, H));
    V3 brdf = v3s(0.0f);
    if (hd.transmission < 1.0f && dot(N, L) > 0.0f && dot(N, V) > 0.0f) {
        V3 Cdlin = hd.albedo;
        float Cdlum = 0.3f * Cdlin.x + 0.6f * Cdlin.y + 0.1f * Cdlin.z;
        V3 Ctint = Cdlum > 0.0f ? Cdlin / Cdlum : v3s(1.0f);
        V3 Cspec0 = er_lerp(hd.specular * 0.08f * er_lerp(v3s(1.0f), Ctint, hd.specularTint), Cdlin, hd.metallic);
        V3 Csheen = er_lerp(v3s(1.0f), Ctint, hd.sheenTint);
        float FL = er_schlick(NDotL);
        float FV = er_schlick(NDotV);
        float Fd90 = 0.5f + 2.0f * LDotH * LDotH * hd.roughness;
        float Fd = er_lerp(1.0f, Fd90, FL) * er_lerp(1.0f, Fd90, FV);
        float Fss90 = LDotH * LDotH * hd.roughness;
        float Fss = er_lerp(1.0f, Fss90, FL) * er_lerp(1.0f, Fss90, FV);
        float ss = 1.25f * (Fss * (1.0f / (NDotL + NDotV) - 0.5f) + 0.5f);
        float aspect = sqrtf(1.0f - hd.anisotropic * 0.9f);
        float ax = er_maxf(0.001f, hd.roughness / aspect);
        float ay = er_maxf(0.001f, hd.roughness * aspect);
        float Ds = er_gtr2_aniso(NDotH, dot(H, T), dot(H, B), ax, ay);
        float FH = er_schlick(LDotH);
        V3 Fs = er_lerp(Cspec0, v3s(1.0f), FH);
        float Gs = er_smithg_ggx_aniso(NDotL, dot(L, T), dot(L, B), ax, ay);
        Gs *= er_smithg_ggx_aniso(NDotV, dot(V, T), dot(V, B), ax, ay);
        V3 Fsheen = FH * hd.sheen * Csheen;
        float Dr = er_gtr1(NDotH, er_lerp(0.1f, 0.001f, hd.clearcoatGloss));
        float Fr = er_lerp(0.04f, 1.0f, FH);
        float Gr = er_smithg_ggx(NDotL, 0.25f) * er_smithg_ggx(NDotV, 0.25f);
        brdf = ((1.0f / ER_PIF) * er_lerp(Fd, ss, hd.subsurface) * Cdlin + Fsheen) * (1.0f - hd.metallic) +
               Gs * Fs * Ds + 0.25f * hd.clearcoat * Gr * Fr * Dr;
    }
    return brdf;
}

// calculateCameraRay (kernel.cpp:371-473); sin/cos of the Euler angles come from the host.
ER_D void er_camera_ray(const DScene& S, int x, int y, float r1, float r2, float r3, float r4, float r5, V3& ro,
                        V3& rd) {
    const DCamera& c = S.cam;
    float dx = c.px + ((float)x) / ((float)S.x_res) * c.sensorWidth;
    float dy = c.py + ((float)y) / ((float)S.y_res) * c.sensorHeight;
    float odx = (-c.sensorWidth / 2.0f) + dx;
    float ody = (-c.sensorHeight / 2.0f) + dy;
    float rx = (1.0f / (float)S.x_res) * (r1 - 0.5f) * c.sensorWidth;
    float ry = (1.0f / (float)S.y_res) * (r2 - 0.5f) * c.sensorHeight;
    float SPx = odx + rx;
    float SPy = ody + ry;
    float SPz = c.pz + c.focalLength;
    V3 pos = v3(c.px, c.py, c.pz);
    V3 dir = v3(SPx, SPy, SPz) - pos;
    V3 dX = v3(dir.x, dir.y * c.cx - dir.z * c.sx, dir.y * c.sx + dir.z * c.cx);
    V3 dY = v3(dX.x * c.cy + dX.z * c.sy, dX.y, dX.z * c.cy - dX.x * c.sy);
    V3 dZ = v3(dY.x * c.cz - dY.y * c.sz, dY.x * c.sz + dY.y * c.cz, dY.z);
    ro = pos;
    rd = er_ray_dir(dZ);
    if (c.bokeh) {
        float diameter = c.focalLength / c.aperture;
        float l = c.focusDistance + c.focalLength;
        V3 focusPoint = ro + rd * l;
        // uniformCircleSampling (Sampling.h:20-28)
        float t = 2 * ER_PIF * r3;
        float uu = r4 + r5;
        float r = uu > 1 ? 2 - uu : uu;
        float rIPx = r * er_cosf(t);
        float rIPy = r * er_sinf(t);
        rIPx *= diameter * 0.5f;
        rIPy *= diameter * 0.5f;
        V3 rIP = v3(rIPx, rIPy, 0.f);
        V3 bX = v3(rIP.x, rIP.y * c.cx - rIP.z * c.sx, rIP.y * c.sx + rIP.z * c.cx);
        V3 bY = v3(bX.x * c.cy + bX.z * c.sy, bX.y, bX.z * c.cy - bX.x * c.sy);
        V3 bZ = v3(bY.x * c.cz - bY.y * c.sz, bY.x * c.sz + bY.y * c.cz, bY.z);
        V3 orig = pos + bZ;
        ro = orig;
        rd = er_ray_dir(focusPoint - orig);
    }
}
