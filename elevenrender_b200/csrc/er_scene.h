// er_scene.h — device-resident scene layout (what er_upload_scene flattens the reference's
// AoS Scene / BVH / Tri / Material / Texture / HDRI into; replaces src/SYCLCopy.cpp:3-104).
//
// All records are 16-byte aligned and read with 128-bit loads.
//
//  pairs   one 64-B record per INTERNAL BVH node (depth 0..17), addressed by the node's pre-order
//          rank among internal nodes:  left(i) = i+1, right(i, depth) = i + (1 << (17-depth)).
//          A record holds the boxes of BOTH children (the reference tests both every step,
//          src/BVH.cpp:73-77) plus what is needed to skip/prune them:
//            q0 = (L.b1.x, L.b1.y, L.b1.z, L.b2.x)
//            q1 = (L.b2.y, L.b2.z, R.b1.x, R.b1.y)
//            q2 = (R.b1.z, R.b2.x, R.b2.y, R.b2.z)
//            q3 = depth < 17 : (int triCount(L), int triCount(R), float slack(L), float slack(R))
//                 depth == 17: (int fromL, int toL (== fromR), int toR, float slack(L u R))
//          slack(subtree) bounds |shadingPosition - geomPosition| for every triangle below
//          (src/Tri.h:100-117): distance pruning must keep any node whose entry distance is
//          within slack of the best hit so far.
//  tri_v   per BVH slot (triangles pre-permuted by triIndices): 3 x float4 = v0,v1,v2 (.w unused)
//  tri_n   per slot: 3 x float4 = vertex normals n0,n1,n2
//  tri_a   per slot: 5 x float4 = (uv0.x,uv0.y,uv1.x,uv1.y) (uv2.x,uv2.y,tangentsSign,objectID bits)
//                                 (t0.xyz, materialID bits) (t1.xyz,0) (t2.xyz,0)
#pragma once
#include <stdint.h>

#define ER_DEPTH 18
#define ER_INTERNAL_COUNT ((1 << ER_DEPTH) - 1) /* 262143 pair records */
#define ER_LEAF_COUNT (1 << ER_DEPTH)

struct DTexture {
    int width, height, channels, filter;
    const float* data;
};

struct DMaterial {  // 32 words, 128 B
    int albedoTextureID, emissionTextureID, roughnessTextureID, metallicTextureID, normalTextureID,
        opacityTextureID, transmissionTextureID, albedoShaderID;
    float albedo[3];
    float emission[3];
    float opacity, roughness, metallic, clearcoatGloss, clearcoat, anisotropic, eta, transmission, specular,
        specularTint, sheenTint, subsurface, sheen, ax, ay;
    // powf(roughness, 2.2f) / powf(metallic, 2.2f) of the constants (src/kernel.cpp:152-153),
    // evaluated once on the host; used when the channel has no texture
    float roughness_pow, metallic_pow;
    float pad;
};

struct DCamera {
    float focalLength, sensorWidth, sensorHeight, aperture, focusDistance;
    float px, py, pz;
    // sin/cos of rotation * (PIF/180) (src/kernel.cpp:412-427), evaluated once on the host
    float cx, sx, cy, sy, cz, sz;
    int bokeh;
    int pad;
};

struct DScene {
    const float4* pairs;
    const float4* tri_v;
    const float4* tri_n;
    const float4* tri_a;
    const float4* tri_g;  // 2 per group of 4 consecutive slots: (lo.xyz, hi.x) (hi.yz, max|coord|, 0): conservative box
    const DMaterial* materials;
    const DTexture* textures;
    DTexture hdri;
    const float* cdf;
    float radianceSum;
    int tri_count;
    int material_count, texture_count;
    int x_res, y_res;
    DCamera cam;
    const float* point_lights;  // 6 floats each: position, radiance (src/PointLight.h:8-9); read by ER_EXT_POINT_LIGHTS only
    int point_light_count;
    // render state (src/kernel.h:44-46)
    float4* passes;      // 5 * P float4
    // per pixel {rng state, sample counter (dev_samples), sample chunks completed, 0}: one 16-B record so
    // that k_render_wavefront hands a pixel from one sample chunk to the next with ONE 128-bit store / load
    uint4* pix;          // P
};

struct DCounters {
    unsigned long long rays, box_tests, tri_tests, candidates, shaded_hits, nee_samples, samples;
};
