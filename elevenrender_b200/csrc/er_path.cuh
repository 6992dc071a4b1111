// er_path.cuh — one path of renderingKernel (src/kernel.cpp:477-646): camera ray, bounce loop with
// environment NEE + Disney sampling, and the running-mean accumulation into the passes.
//
// The bounce body is split at the shadow ray so that the persistent kernel (er_render.cuh) can send
// nearest-hit and shadow queries through ONE traversal site:
//   er_shade_hit()   everything of kernel.cpp:524-587 that does not depend on the shadow ray's answer
//   er_shade_finish()  light/reduction update once `blocked` is known
#pragma once
#include "er_shade.cuh"

struct PathOpts {
    int max_depth;  // the reference hard-codes 5 (kernel.cpp:508)
    int rtl;        // 1 = draw argument-list uniforms right-to-left (g++ build of the reference)
    int ext;        // ErExtension bits (er_cuda.h): 1 MIS weights, 2 point-light NEE, 4 Russian roulette; 0 = the reference
};
#define ER_EXT_MIS_BIT 1
#define ER_EXT_POINT_BIT 2
#define ER_EXT_RR_BIT 4

struct PathStats {
    TraceStats tr;
    unsigned rays, shaded_hits, nee_samples;
};

// What a shaded hit leaves behind for the rest of the bounce.
struct Bounce {
    V3 add_open;     // emission + hdriInt with the environment sample visible      (kernel.cpp:567,575)
    V3 add_blocked;  // emission + hdriInt with hdriValue = Vector3()               (kernel.cpp:560-562)
    V3 factor;       // DisneyEval(wibrdf) * |wibrdf.n| / brdfpdf                   (kernel.cpp:577)
    V3 next_o, next_d;  // the continued ray, direction already normalised (kernel.cpp:587 / :590)
    V3 so, sd;       // shadow ray (kernel.cpp:555)
    V3 normal, tangent, bitangent;  // hitdata AOVs (kernel.cpp:581-585)
    int own_slot;
    bool opaque;       // rnd.next() <= opacity
    bool need_shadow;  // false when the shadow ray's answer cannot change the result
    bool shadow_any;   // true: the first candidate decides (own triangle cannot be hit)
    // extensions (opt.ext != 0, k_render_pixel only)
    float bw;          // ER_EXT_MIS: weight of the environment seen by the BRDF-sampled ray (kernel.cpp:572)
    bool has_pl;       // ER_EXT_POINT_LIGHTS: a point light was sampled
    V3 pl_o, pl_d;     //   shadow ray towards it (kernel.cpp:283)
    V3 pl_point;       //   the shaded point distances are measured from (kernel.cpp:280,285)
    float pl_dist;     //   distance to the light
    V3 pl_value;       //   radiance / dist^2 * DisneyEval * |cos| / pdf (the commented-out return of kernel.cpp:296-299)
};

ER_D bool er_finite3(V3 v) { return isfinite(v.x) && isfinite(v.y) && isfinite(v.z); }

// kernel.cpp:524-587 for a valid nearest hit `rec` of ray (ro, rd). EXT = false is the reference's estimator (what
// the production kernels instantiate: the extension code costs them nothing); EXT = true reads opt.ext.
template <bool EXT>
ER_D void er_shade_hit(const DScene& S, const PathOpts& opt, const HitRec& rec, V3 ro, V3 rd, uint32_t& rng, Bounce& b) {
    HitFull hit;
    er_make_hit(S, rec, ro, rd, hit);
    const DMaterial& m = S.materials[hit.material_id];
    HitData hd;
    er_generate_hitdata(S, m, hit, hd);
    if (m.albedoShaderID != -1) er_asl_shade(m.albedoShaderID, hd.albedo);

    b.bw = 1.f;
    b.has_pl = false;
    b.opaque = er_rng_next(rng) <= hd.opacity;
    if (!b.opaque) {
        b.next_o = hit.position + rd * 0.001f;
        b.next_d = er_ray_dir(rd);
        b.need_shadow = false;
        return;
    }
    V3 wo = -rd;
    // HDRI::sample (HDRI.cpp:109-117)
    int count = er_hdri_search(S.cdf, er_rng_next(rng), S.hdri.width * S.hdri.height);
    float tcx = (float)(count % S.hdri.width);
    float tcy = (float)(count / S.hdri.width);
    float d0 = er_rng_next(rng), d1 = er_rng_next(rng), d2 = er_rng_next(rng);
    V3 wibrdf = opt.rtl ? er_disney_sample(hd, wo, hd.normal, d2, d1, d0) : er_disney_sample(hd, wo, hd.normal, d0, d1, d2);
    float nu = tcx / (float)S.hdri.width;
    float nv = tcy / (float)S.hdri.height;
    float iu, iv;
    er_inverse_transform_uv(S.hdri, nu, nv, iu, iv);
    V3 wihdri = -normalized(er_reverse_spherical(iu, iv));
    b.so = hd.position + hd.normal * 0.001f;
    b.sd = er_ray_dir(wihdri);
    b.own_slot = hd.triIdx;

    V3 hdriValue = er_tex_uv(S.hdri, iu, iv);
    float hdripdf = er_hdri_pdf(S, er_f2i(iu * (float)S.hdri.width), er_f2i(iv * (float)S.hdri.height));
    // DisneyEval is called for the environment direction and for the sampled direction
    // (kernel.cpp:567,577): ONE copy of its code, run twice (the render kernel is instruction-fetch
    // bound in this function, profiles/README.md); -DER_NO_DEDUP inlines it twice as before
#ifndef ER_NO_DEDUP
    V3 evalh = v3s(0.f), evalb = v3s(0.f);
#pragma unroll 1
    for (int k = 0; k < 2; k++) {
        const V3 e = er_disney_eval(hd, wo, hd.normal, k ? wibrdf : wihdri);
        if (k)
            evalb = e;
        else
            evalh = e;
    }
#else
    V3 evalh = er_disney_eval(hd, wo, hd.normal, wihdri);
    V3 evalb = er_disney_eval(hd, wo, hd.normal, wibrdf);
#endif
    float cosh = fabsf(dot(wihdri, hd.normal));
    V3 hdriInt = hdriValue * evalh * cosh / hdripdf;
    V3 hdriIntBlocked = v3s(0.f) * evalh * cosh / hdripdf;
    b.add_open = hd.emission + hdriInt;
    b.add_blocked = hd.emission + hdriIntBlocked;
    // With a zero BRDF and finite texel both variants are +-0 or the same NaN: the visibility query
    // cannot change the sample (at most the sign of a zero), so it is not traced.
    b.need_shadow = !(evalh.x == 0.f && evalh.y == 0.f && evalh.z == 0.f && er_finite3(hdriValue));
    if (b.need_shadow) {
        // kernel.cpp:555-562: blocked iff the NEAREST hit exists and is not the shaded triangle. If the
        // shadow ray cannot hit the shaded triangle at all, any candidate blocks it.
        const float4* pv = S.tri_v + 3 * (size_t)b.own_slot;
        float u, v, t;
        b.shadow_any = !er_mt(ld3(pv), ld3(pv + 1), ld3(pv + 2), b.so, b.sd, u, v, t);
    } else {
        b.shadow_any = true;
    }
    float brdfpdf = er_disney_pdf(hd, wo, hd.normal, wibrdf);
    if (EXT && (opt.ext & ER_EXT_MIS_BIT)) {
        // the balance heuristic of kernel.cpp:571-572, applied instead of dropped: the environment is then counted
        // once (NEE term weighted hw, the environment the BRDF-sampled ray sees weighted bw) instead of twice
        float hw = hdripdf / (brdfpdf + hdripdf);
        float bw = brdfpdf / (brdfpdf + hdripdf);
        if (!(hw >= 0.f && hw <= 1.f)) hw = 0.f, bw = 1.f;  // infinite / NaN pdf (texel row 0): the NEE term is zero there
        b.add_open = hd.emission + hdriInt * hw;
        b.bw = bw;
    }
    if (EXT && (opt.ext & ER_EXT_POINT_BIT) && S.point_light_count > 0) {
        // pointLight() (kernel.cpp:269-301): one light chosen by an extra uniform drawn AFTER the bounce's other
        // draws, pdf = count / (2 pi), quadratic attenuation, the BRDF and cosine of its commented-out return
        const float r1 = er_rng_next(rng);
        int li = (int)((float)S.point_light_count * r1);
        if (li >= S.point_light_count) li = S.point_light_count - 1;
        const float* pl = S.point_lights + 6 * li;
        const V3 lp = v3(pl[0], pl[1], pl[2]), lr = v3(pl[3], pl[4], pl[5]);
        const V3 point = hd.position;
        const V3 newDir = normalized(lp - point);
        const float dist = length(lp - point);
        const float pdf = ((float)S.point_light_count) / (2.0f * ER_PIF);
        const V3 value = lr / (dist * dist);
        b.has_pl = true;
        b.pl_point = point;
        b.pl_dist = dist;
        b.pl_o = point + newDir * 0.001f;
        b.pl_d = er_ray_dir(newDir);
        b.pl_value = value * er_disney_eval(hd, wo, hd.normal, newDir) * fabsf(dot(newDir, hd.normal)) / pdf;
    }
    b.factor = evalb * fabsf(dot(wibrdf, hd.normal)) / brdfpdf;
    b.normal = hd.normal;
    b.tangent = hd.tangent;
    b.bitangent = hd.bitangent;
    b.next_o = hit.position + wibrdf * 0.001f;
    b.next_d = er_ray_dir(wibrdf);
}

// One sample of one pixel, bounce by bounce (the one-thread-per-pixel form).
template <bool COUNT, bool ORDERED>
ER_D void er_trace_path(const DScene& S, const PathOpts& opt, int x, int y, uint32_t& rng, V3& light, V3& normal,
                        V3& tangent, V3& bitangent, PathStats& ps) {
    float c0 = er_rng_next(rng), c1 = er_rng_next(rng), c2 = er_rng_next(rng), c3 = er_rng_next(rng),
          c4 = er_rng_next(rng);
    V3 ro, rd;
    if (opt.rtl)
        er_camera_ray(S, x, y, c4, c3, c2, c1, c0, ro, rd);
    else
        er_camera_ray(S, x, y, c0, c1, c2, c3, c4, ro, rd);

    light = v3s(0.f);
    normal = v3s(0.f);
    tangent = v3s(0.f);
    bitangent = v3s(0.f);
    V3 reduction = v3s(1.f);
    float env_w = 1.f;  // ER_EXT_MIS: weight of the environment on a miss (1 for camera rays)

    for (int i = 0; i < opt.max_depth; i++) {
        if ((opt.ext & ER_EXT_RR_BIT) && i >= 2) {
            // Russian roulette: unbiased termination of dim paths from the third bounce on
            float p = fmaxf(reduction.x, fmaxf(reduction.y, reduction.z));
            p = p < 0.05f ? 0.05f : p > 1.f ? 1.f : p;
            if (er_rng_next(rng) > p) break;
            reduction = reduction / p;
        }
        HitRec rec;
        er_nearest<COUNT, ORDERED>(S, ro, rd, rec, ps.tr);
        if (COUNT) ps.rays++;
        if (rec.slot < 0) {
            float u, v;
            er_spherical_mapping(-1.0f * rd, u, v);
            light = light + reduction * er_tex_filtered(S.hdri, u, v) * env_w;
            break;
        }
        Bounce b;
        if (opt.ext)
            er_shade_hit<true>(S, opt, rec, ro, rd, rng, b);
        else
            er_shade_hit<false>(S, opt, rec, ro, rd, rng, b);
        if (COUNT) ps.shaded_hits++;
        if (b.opaque) {
            bool blocked = false;
            if (COUNT) ps.nee_samples++;
            if (b.need_shadow || !ORDERED) {
                HitRec sr;
                sr.slot = -1;
                sr.dist = 0.f;
                sr.rank = 0;
                sr.u = sr.v = sr.t = 0.f;
                const bool any = ORDERED && b.shadow_any;
                er_traverse<COUNT, ORDERED>(S, any, b.so, b.sd, sr, ps.tr);
                if (COUNT) ps.rays++;
                blocked = sr.slot >= 0 && (any || sr.slot != b.own_slot);
            }
            light = light + reduction * (blocked ? b.add_blocked : b.add_open);
            if (b.has_pl) {
                // kernel.cpp:283-289: visible unless the nearest hit of the shadow ray is closer than the light
                HitRec pr;
                er_nearest<COUNT, ORDERED>(S, b.pl_o, b.pl_d, pr, ps.tr);
                if (COUNT) ps.rays++;
                bool lit = true;
                if (pr.slot >= 0) {
                    HitFull ph;
                    er_make_hit(S, pr, b.pl_o, b.pl_d, ph);
                    lit = !(length(ph.position - b.pl_point) < b.pl_dist);
                }
                if (lit) light = light + reduction * b.pl_value;
            }
            reduction = reduction * b.factor;
            env_w = b.bw;
            if (i == 0) {
                normal = b.normal;
                tangent = b.tangent;
                bitangent = b.bitangent;
            }
        }
        ro = b.next_o;
        rd = b.next_d;
    }
    light = er_clamp(light, 0.f, 10.f);
}

#ifndef ER_PASS_LD
#define ER_PASS_LD(p) __ldcg(p)
#endif
// The pass update of kernel.cpp:600-644 for one finished sample. Returns the new sample counter.
ER_D uint32_t er_accumulate(float4* __restrict__ passes, size_t P, size_t idx, uint32_t sa, V3 light, V3 normal,
                            V3 tangent, V3 bitangent) {
    if (isnan(light.x) || isnan(light.y) || isnan(light.z)) return sa;
    const float k = ((float)sa) / ((float)(sa + 1));
    const float den = (float)(sa + 1);
    const V3 add[5] = {light, v3s(0.f), normal, tangent, bitangent};
#pragma unroll
    for (int pass = 0; pass < 5; pass++) {
        if (pass == 1) continue;  // DENOISE is never written
        float4 p = ER_PASS_LD(passes + pass * P + idx);  // L2: the previous sample may have been written by another SM
        if (sa > 0) {
            p.x *= k;
            p.y *= k;
            p.z *= k;
        }
        p.x += add[pass].x / den;
        p.y += add[pass].y / den;
        p.z += add[pass].z / den;
        passes[pass * P + idx] = p;
    }
    return sa + 1;
}
