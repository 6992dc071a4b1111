// TEST INFRASTRUCTURE — not product code.  Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may load the library this file is built into.
//
// C-ABI harness around the UNMODIFIED reference sources (compiled by path from
// /root/reference/src by oracle/Makefile together with the stand-in headers under oracle/shim/).
// It builds a reference `Scene` from flat arrays, lets the reference build its own BVH / HDRI CDF,
// deep-copies it with the reference's copy_scene (SYCLCopy.cpp:3-104, through the host shim) and
// drives setupKernel / renderingKernel once per pixel index (kernel.cpp:176, :477) — NOT through
// kernel_render_enqueue, whose padded launch samples some pixels twice.
//
// Built with g++, so rnd.next() calls inside an argument list are evaluated right-to-left
// (SURVEY F5): the CUDA path is compared with rng_order = RTL.
#include <atomic>
#include <cstddef>
#include <sstream>
#include <thread>
#include <vector>

#include "BVH.h"
#include "Camera.h"
#include "Disney.h"
#include "HDRI.h"
#include "Hit.h"
#include "Material.h"
#include "ObjLoader.h"
#include "Ray.h"
#include "SYCLCopy.h"
#include "Scene.h"
#include "Texture.h"
#include "kernel.h"
#include "shader.h"

// external-linkage functions of kernel.cpp that no reference header declares
void setupKernel(dev_Scene* dev_scene_g, int idx);                       // kernel.cpp:176
Hit throwRay(Ray ray, dev_Scene* scene);                                 // kernel.cpp:218
void generateHitData(dev_Scene*, Material*, HitData&, Hit);              // kernel.cpp:76
void calculateCameraRay(int x, int y, dev_Scene& scene, Camera& camera, Ray& ray, float r1, float r2,
                        float r3, float r4, float r5);                   // kernel.cpp:371

namespace {

// ---- flat exchange records (same words as include/er_cuda.h) -------------------------------
constexpr int TRI_WORDS = 39;   // vertices9 uv9 normals9 tangents9 sign objectID materialID
constexpr int MAT_WORDS = 29;   // 8 ids, albedo3, emission3, 15 scalars
constexpr int CAM_WORDS = 12;   // focal sw sh aperture focusDist rot3 bokeh pos3
constexpr int NODE_WORDS = 11;  // b1(3) b2(3) from to idx depth valid
constexpr int HIT_WORDS = 21;   // pos3 n3 t3 b3 gn3 valid objectID material_id tu tv triIdx

static_assert(offsetof(Tri, tangentsSign) == 144, "Tri layout");
static_assert(offsetof(Tri, materialID) == 152, "Tri layout");
static_assert(sizeof(Node) == 44, "Node layout");
static_assert(sizeof(Hit) == 84, "Hit layout");
static_assert(sizeof(Camera) == 48, "Camera layout");

struct RefCtx {
    Scene scene;
    dev_Scene* host_view = nullptr;  // `temp` in renderSetup (kernel.cpp:655)
    dev_Scene* ds = nullptr;         // deep copy made by copy_scene
    bool finalized = false;
};

void fill_tri(Tri& t, const float* w) {
    std::memcpy(&t.vertices[0], w + 0, 36);
    std::memcpy(&t.uv[0], w + 9, 36);
    std::memcpy(&t.normals[0], w + 18, 36);
    std::memcpy(&t.tangents[0], w + 27, 36);
    t.tangentsSign = w[36];
    std::memcpy(&t.objectID, w + 37, 4);
    std::memcpy(&t.materialID, w + 38, 4);
}

void dump_tri(const Tri& t, float* w) {
    std::memcpy(w + 0, &t.vertices[0], 36);
    std::memcpy(w + 9, &t.uv[0], 36);
    std::memcpy(w + 18, &t.normals[0], 36);
    std::memcpy(w + 27, &t.tangents[0], 36);
    w[36] = t.tangentsSign;
    std::memcpy(w + 37, &t.objectID, 4);
    std::memcpy(w + 38, &t.materialID, 4);
}

void fill_material(Material& m, const float* w) {
    int ids[8];
    std::memcpy(ids, w, 32);
    m.albedoTextureID = ids[0];
    m.emissionTextureID = ids[1];
    m.roughnessTextureID = ids[2];
    m.metallicTextureID = ids[3];
    m.normalTextureID = ids[4];
    m.opacityTextureID = ids[5];
    m.transmissionTextureID = ids[6];
    m.albedoShaderID = ids[7];
    m.albedo = Vector3(w[8], w[9], w[10]);
    m.emission = Vector3(w[11], w[12], w[13]);
    m.opacity = w[14];
    m.roughness = w[15];
    m.metallic = w[16];
    m.clearcoatGloss = w[17];
    m.clearcoat = w[18];
    m.anisotropic = w[19];
    m.eta = w[20];
    m.transmission = w[21];
    m.specular = w[22];
    m.specularTint = w[23];
    m.sheenTint = w[24];
    m.subsurface = w[25];
    m.sheen = w[26];
    m.ax = w[27];
    m.ay = w[28];
}

void fill_camera(Camera& c, const float* w) {
    c.focalLength = w[0];
    c.sensorWidth = w[1];
    c.sensorHeight = w[2];
    c.aperture = w[3];
    c.focusDistance = w[4];
    c.rotation = Vector3(w[5], w[6], w[7]);
    int b;
    std::memcpy(&b, w + 8, 4);
    c.bokeh = b != 0;
    c.position = Vector3(w[9], w[10], w[11]);
}

void dump_hit(const Hit& h, float* w) {
    std::memcpy(w + 0, &h.position, 12);
    std::memcpy(w + 3, &h.normal, 12);
    std::memcpy(w + 6, &h.tangent, 12);
    std::memcpy(w + 9, &h.bitangent, 12);
    std::memcpy(w + 12, &h.gnormal, 12);
    int v = h.valid ? 1 : 0;
    std::memcpy(w + 15, &v, 4);
    // objectID/material_id/tu/tv are uninitialised in the reference on a miss (Hit.h:9-11)
    unsigned int oid = h.valid ? h.objectID : 0u, mid = h.valid ? h.material_id : 0u;
    float tu = h.valid ? h.tu : 0.f, tv = h.valid ? h.tv : 0.f;
    std::memcpy(w + 16, &oid, 4);
    std::memcpy(w + 17, &mid, 4);
    w[18] = tu;
    w[19] = tv;
    std::memcpy(w + 20, &h.triIdx, 4);
}

template <typename F>
void parallel_for_idx(long begin, long end, int nthreads, F f) {
    if (nthreads <= 1) {
        for (long i = begin; i < end; i++) f(i);
        return;
    }
    std::atomic<long> next(begin);
    const long chunk = 256;
    std::vector<std::thread> pool;
    for (int t = 0; t < nthreads; t++) {
        pool.emplace_back([&]() {
            for (;;) {
                long s = next.fetch_add(chunk);
                if (s >= end) break;
                long e = s + chunk < end ? s + chunk : end;
                for (long i = s; i < e; i++) f(i);
            }
        });
    }
    for (auto& th : pool) th.join();
}

}  // namespace

extern "C" {

// ---- size table ------------------------------------------------------------
void ref_sizeof_table(int* out /*[11]*/) {
    out[0] = sizeof(Tri);
    out[1] = sizeof(Node);
    out[2] = (int)(sizeof(BVH) & 0x7fffffff);
    out[3] = sizeof(Hit);
    out[4] = sizeof(HitData);
    out[5] = sizeof(Material);
    out[6] = sizeof(Texture);
    out[7] = sizeof(HDRI);
    out[8] = sizeof(Camera);
    out[9] = sizeof(dev_Scene);
    out[10] = sizeof(RngGenerator);
}

void ref_flush_stdout() {
    fflush(stdout);
    std::cout.flush();
}

int ref_hardware_threads() { return (int)std::thread::hardware_concurrency(); }

// ---- scene construction -------------------------------------------------------------------
void* ref_scene_create(int x_res, int y_res) {
    RefCtx* c = new RefCtx();
    c->scene.x_res = x_res;
    c->scene.y_res = y_res;
    return c;
}

void ref_scene_destroy(void* h) {
    // the reference never frees its deep copy; the harness leaks it deliberately as well except
    // for the big blocks
    RefCtx* c = (RefCtx*)h;
    if (c->ds) {
        std::free(c->ds->tris);
        if (c->ds->bvh) {
            std::free(c->ds->bvh->triIndices);
            std::free(c->ds->bvh);
        }
        std::free(c->ds->dev_passes);
        std::free(c->ds->dev_samples);
        std::free(c->ds->dev_randstate);
    }
    if (c->host_view && c->host_view->bvh) {
        delete[] c->host_view->bvh->triIndices;
        delete c->host_view->bvh;
    }
    delete c;
}

void ref_scene_set_camera(void* h, const float* cam) { fill_camera(((RefCtx*)h)->scene.camera, cam); }

// material 0 ("default") exists already (Scene.h:43); returns the new material's index
int ref_scene_add_material(void* h, const float* mat_words, const char* name) {
    RefCtx* c = (RefCtx*)h;
    Material m;
    fill_material(m, mat_words);
    m.name = name ? name : "";
    c->scene.addMaterial(m);
    return c->scene.materialCount() - 1;
}

void ref_scene_set_material(void* h, int i, const float* mat_words) {
    fill_material(((RefCtx*)h)->scene.materials[i], mat_words);
}

int ref_scene_add_texture(void* h, int w, int ht, int ch, int bilinear, const float* data, const char* name) {
    RefCtx* c = (RefCtx*)h;
    float* copy = new float[(size_t)w * ht * ch];
    std::memcpy(copy, data, sizeof(float) * (size_t)w * ht * ch);
    Texture t(name, w, ht, ch, copy, bilinear ? Texture::Filter::BILINEAR : Texture::Filter::NO_FILTER,
              Texture::CS::LINEAR);
    c->scene.addTexture(t);
    return c->scene.textureCount() - 1;
}

// HDRI(Texture) builds the CDF (HDRI.cpp:6-10,62-83); pixel_shift as LoadHDRI does is the caller's
// business (CommandManager.cpp) — the harness takes the final texels.
void ref_scene_set_hdri(void* h, int w, int ht, int ch, int bilinear, const float* data) {
    RefCtx* c = (RefCtx*)h;
    float* copy = new float[(size_t)w * ht * ch];
    std::memcpy(copy, data, sizeof(float) * (size_t)w * ht * ch);
    Texture t("hdri", w, ht, ch, copy, bilinear ? Texture::Filter::BILINEAR : Texture::Filter::NO_FILTER,
              Texture::CS::LINEAR);
    c->scene.addHDRI(HDRI(t));
}

void ref_scene_set_hdri_color(void* h, float r, float g, float b) {
    ((RefCtx*)h)->scene.addHDRI(HDRI(Vector3(r, g, b)));
}

void ref_scene_add_point_light(void* h, const float* pos, const float* rad) {
    ((RefCtx*)h)->scene.addPointLight(
        PointLight(Vector3(pos[0], pos[1], pos[2]), Vector3(rad[0], rad[1], rad[2])));
}

// one MeshObject made of n flat triangles (Scene::addMeshObject, Scene.cpp:48-66)
void ref_scene_add_tris(void* h, const float* tri_words, long n) {
    RefCtx* c = (RefCtx*)h;
    std::vector<Tri>* tris = new std::vector<Tri>((size_t)n);
    for (long i = 0; i < n; i++) fill_tri((*tris)[i], tri_words + i * TRI_WORDS);
    MeshObject mo;
    mo.name = "mesh";
    mo.tris = tris->data();
    mo.triCount = (unsigned)n;
    c->scene.tris.reserve(c->scene.tris.size() + (size_t)n);
    c->scene.addMeshObject(mo);
    // the reference overwrites objectID in addMeshObject; restore what the caller gave
    size_t base = c->scene.tris.size() - (size_t)n;
    for (long i = 0; i < n; i++) std::memcpy(&c->scene.tris[base + i].objectID, tri_words + i * TRI_WORDS + 37, 4);
    delete tris;
}

// OBJ/MTL text through the reference ingest path (ObjLoader.cpp:154-164 -> :69-147, MikkTSpace),
// then Scene::addMeshObject + pair_materials. Returns number of triangles now in the scene.
long ref_scene_load_obj(void* h, const char* obj_text, const char* mtl_text, int recompute_normals) {
    RefCtx* c = (RefCtx*)h;
    std::istringstream obj_stream{std::string(obj_text)};
    std::vector<MeshObject> objs;
    ObjLoader loader;
    loader.loadObjsRapid(obj_stream, std::string_view(mtl_text ? mtl_text : ""), objs, recompute_normals != 0);
    for (auto& mo : objs) c->scene.addMeshObject(mo);
    c->scene.pair_materials();
    return (long)c->scene.tris.size();
}

long ref_scene_tri_count(void* h) { return (long)((RefCtx*)h)->scene.tris.size(); }

void ref_scene_get_tris(void* h, float* out_words) {
    RefCtx* c = (RefCtx*)h;
    for (size_t i = 0; i < c->scene.tris.size(); i++) dump_tri(c->scene.tris[i], out_words + i * TRI_WORDS);
}

// new dev_Scene(scene) builds the BVH with the reference builder (kernel.cpp:244-266 ->
// Scene.cpp:122-143 -> BVH.cpp:132-281); copy_scene deep-copies; setupKernel per idx.
// If nodes/tri_indices are given (both non-null), the reference build is skipped and the given
// tree is installed instead (used for very large scenes once the builder under test has been
// shown to be identical on smaller ones).
void ref_scene_finalize(void* h, const float* nodes_words, const int* tri_indices) {
    RefCtx* c = (RefCtx*)h;
    Scene* scene = &c->scene;
    dev_Scene* temp;
    if (nodes_words && tri_indices) {
        temp = (dev_Scene*)std::calloc(1, sizeof(dev_Scene));
        temp->camera = &scene->camera;
        temp->meshObjectCount = scene->meshObjectCount();
        temp->materialCount = scene->materialCount();
        temp->textureCount = scene->textureCount();
        temp->triCount = scene->triCount();
        temp->pointLightCount = scene->pointLightCount();
        temp->pointLights = scene->getPointLights();
        temp->meshObjects = scene->getMeshObjects();
        temp->materials = scene->getMaterials();
        temp->textures = scene->getTextures();
        temp->hdri = &scene->hdri;
        temp->tris = scene->getTris();
        BVH* bvh = new BVH();
        bvh->triIndices = new int[scene->triCount() > 0 ? scene->triCount() : 1];
        std::memcpy(bvh->triIndices, tri_indices, sizeof(int) * (size_t)scene->triCount());
        static_assert(sizeof(Node) == NODE_WORDS * 4, "node words");
        std::memcpy(bvh->nodes, nodes_words, sizeof(Node) * ((2 << BVH_DEPTH) - 1));
        temp->bvh = bvh;
        temp->x_res = scene->x_res;
        temp->y_res = scene->y_res;
    } else {
        temp = new dev_Scene(scene);
    }
    c->host_view = temp;
    sycl::queue q;
    c->ds = sycl::malloc_device<dev_Scene>(1, q);
    copy_scene(temp, c->ds, q);
    c->finalized = true;
}

void ref_setup(void* h) {
    RefCtx* c = (RefCtx*)h;
    long P = (long)c->ds->x_res * c->ds->y_res;
    for (long i = 0; i < P; i++) setupKernel(c->ds, (int)i);
}

// ---- exports of host-built data -----------------------------------------------------------
long ref_bvh_node_count() { return (2 << BVH_DEPTH) - 1; }

void ref_get_bvh(void* h, float* nodes_words, int* tri_indices) {
    RefCtx* c = (RefCtx*)h;
    std::memcpy(nodes_words, c->host_view->bvh->nodes, sizeof(Node) * ((2 << BVH_DEPTH) - 1));
    std::memcpy(tri_indices, c->host_view->bvh->triIndices, sizeof(int) * (size_t)c->scene.triCount());
}

void ref_get_hdri(void* h, int* whc, float* radiance_sum, float* cdf /*w*h+1 or null*/, float* texels /*or null*/) {
    RefCtx* c = (RefCtx*)h;
    HDRI& hd = c->scene.hdri;
    whc[0] = hd.texture.width;
    whc[1] = hd.texture.height;
    whc[2] = (int)hd.texture.channels;
    *radiance_sum = hd.radianceSum;
    size_t n = (size_t)hd.texture.width * hd.texture.height;
    if (cdf) std::memcpy(cdf, hd.cdf, sizeof(float) * (n + 1));
    if (texels) std::memcpy(texels, hd.texture.data, sizeof(float) * n * hd.texture.channels);
}

// ---- the hot path -------------------------------------------------------------------------
// `passes` renderingKernel calls per pixel index in [idx_begin, idx_end) (kernel.cpp:477-646)
void ref_render(void* h, int passes, int nthreads, long idx_begin, long idx_end) {
    RefCtx* c = (RefCtx*)h;
    dev_Scene* ds = c->ds;
    parallel_for_idx(idx_begin, idx_end, nthreads, [=](long i) {
        for (int s = 0; s < passes; s++) renderingKernel(ds, (int)i, passes);
    });
}

void ref_get_passes(void* h, float* out /*5*P*4*/) {
    RefCtx* c = (RefCtx*)h;
    size_t P = (size_t)c->ds->x_res * c->ds->y_res;
    std::memcpy(out, c->ds->dev_passes, sizeof(float) * PASSES_COUNT * P * 4);
}

void ref_get_samples(void* h, unsigned* out) {
    RefCtx* c = (RefCtx*)h;
    size_t P = (size_t)c->ds->x_res * c->ds->y_res;
    std::memcpy(out, c->ds->dev_samples, sizeof(unsigned) * P);
}

void ref_get_rng(void* h, unsigned* out) {
    RefCtx* c = (RefCtx*)h;
    size_t P = (size_t)c->ds->x_res * c->ds->y_res;
    static_assert(sizeof(RngGenerator) == 4, "rng");
    std::memcpy(out, c->ds->dev_randstate, 4 * P);
}

// nearest-hit queries for host-supplied rays (o.xyz, d.xyz); Ray ctor normalises (Ray.h:13-17)
void ref_trace(void* h, const float* rays, long n, float* hits_words, int nthreads) {
    RefCtx* c = (RefCtx*)h;
    dev_Scene* ds = c->ds;
    parallel_for_idx(0, n, nthreads, [=](long i) {
        const float* r = rays + 6 * i;
        Ray ray(Vector3(r[0], r[1], r[2]), Vector3(r[3], r[4], r[5]));
        Hit hit = throwRay(ray, ds);
        dump_hit(hit, hits_words + i * HIT_WORDS);
    });
}

// ---- unit entry points (T1) ---------------------------------------------------------------
void ref_rng_stream(int idx, int n, unsigned* states, float* values) {
    RngGenerator r((uint32_t)idx);
    for (int i = 0; i < n; i++) {
        values[i] = r.next();
        states[i] = r.state;
    }
}

void ref_camera_ray(void* h, int x, int y, const float* r5, float* out_ray6) {
    RefCtx* c = (RefCtx*)h;
    Ray ray;
    calculateCameraRay(x, y, *c->ds, *c->ds->camera, ray, r5[0], r5[1], r5[2], r5[3], r5[4]);
    std::memcpy(out_ray6, &ray.origin, 12);
    std::memcpy(out_ray6 + 3, &ray.direction, 12);
}

int ref_box_intersect(const float* ray6_normalised, const float* b1, const float* b2) {
    static BVH* bvh = new BVH();
    Ray ray;
    ray.origin = Vector3(ray6_normalised[0], ray6_normalised[1], ray6_normalised[2]);
    ray.direction = Vector3(ray6_normalised[3], ray6_normalised[4], ray6_normalised[5]);
    return bvh->intersect(ray, Vector3(b1[0], b1[1], b1[2]), Vector3(b2[0], b2[1], b2[2])) ? 1 : 0;
}

int ref_tri_hit(const float* tri_words, const float* ray6, float* hit_words) {
    Tri t;
    fill_tri(t, tri_words);
    Ray ray(Vector3(ray6[0], ray6[1], ray6[2]), Vector3(ray6[3], ray6[4], ray6[5]));
    Hit hit;
    bool ok = t.hit(ray, hit);
    dump_hit(hit, hit_words);
    return ok ? 1 : 0;
}

// mode 0: getValueFromCoordinates(int(a),int(b)); 1: getValueFromUV; 2: getValueBilinear; 3: Filtered
void ref_texture_lookup(void* h, int tex /* -1 = hdri */, int mode, const float* ab, long n, float* out3) {
    RefCtx* c = (RefCtx*)h;
    Texture* t = tex < 0 ? &c->ds->hdri->texture : &c->ds->textures[tex];
    for (long i = 0; i < n; i++) {
        Vector3 v;
        float a = ab[2 * i], b = ab[2 * i + 1];
        if (mode == 0) v = t->getValueFromCoordinates((int)a, (int)b);
        else if (mode == 1) v = t->getValueFromUV(a, b);
        else if (mode == 2) v = t->getValueBilinear(a, b);
        else v = t->getValueFromUVFiltered(a, b);
        std::memcpy(out3 + 3 * i, &v, 12);
    }
}

// the libm the reference's device code reaches: sycl::{sin,cos,acos,atan2,pow,log} (src/sycl.h:8), here the
// host's glibc through the stand-in header. fn: 0 sin, 1 cos, 2 acos, 3 atan2(a,b), 4 pow(a,b), 5 log
void ref_libm(int fn, const float* a, const float* b, long n, float* out) {
    for (long i = 0; i < n; i++) {
        const float x = a[i], y = b ? b[i] : 0.f;
        float r;
        switch (fn) {
            case 0: r = sycl::sin(x); break;
            case 1: r = sycl::cos(x); break;
            case 2: r = sycl::acos(x); break;
            case 3: r = sycl::atan2(x, y); break;
            case 4: r = sycl::pow(x, y); break;
            default: r = sycl::log(x); break;
        }
        out[i] = r;
    }
}

void ref_spherical_mapping(const float* dirs3, long n, float* uv2) {
    for (long i = 0; i < n; i++) {
        float u, v;
        Texture::sphericalMapping(Vector3(), Vector3(dirs3[3 * i], dirs3[3 * i + 1], dirs3[3 * i + 2]), 1, u, v);
        uv2[2 * i] = u;
        uv2[2 * i + 1] = v;
    }
}

// env NEE direction chain of kernel.cpp:543-553,558,565: r -> texel -> (iu,iv) -> wihdri, value, pdf
void ref_hdri_sample(void* h, const float* r, long n, float* out /* n*10: x y iu iv wi3 val3 */, float* pdf) {
    RefCtx* c = (RefCtx*)h;
    HDRI* hd = c->ds->hdri;
    for (long i = 0; i < n; i++) {
        Vector3 tc = hd->sample(r[i]);
        float nu = tc.x / static_cast<float>(hd->texture.width);
        float nv = tc.y / static_cast<float>(hd->texture.height);
        float iu = hd->texture.inverseTransformUV(nu, nv).x;
        float iv = hd->texture.inverseTransformUV(nu, nv).y;
        Vector3 wi = -hd->texture.reverseSphericalMapping(iu, iv).normalized();
        Vector3 val = hd->texture.getValueFromUV(iu, iv);
        float* o = out + 10 * i;
        o[0] = tc.x; o[1] = tc.y; o[2] = iu; o[3] = iv;
        o[4] = wi.x; o[5] = wi.y; o[6] = wi.z;
        o[7] = val.x; o[8] = val.y; o[9] = val.z;
        pdf[i] = hd->pdf(iu * hd->texture.width, iv * hd->texture.height);
    }
}

// hitdata_words: 37 floats in HitData member order (kernel.h:92-121) with triIdx as int bits
static void fill_hitdata(HitData& hd, const float* w) { std::memcpy(&hd, w, sizeof(HitData)); }

void ref_disney(const float* hitdata_words, const float* V3, const float* L3, const float* r3,
                float* out_sample3, float* out_eval3, float* out_pdf) {
    static_assert(sizeof(HitData) == 37 * 4, "HitData words");
    HitData hd;
    fill_hitdata(hd, hitdata_words);
    Vector3 V(V3[0], V3[1], V3[2]), L(L3[0], L3[1], L3[2]);
    Vector3 s = DisneySample(hd, V, hd.normal, r3[0], r3[1], r3[2]);
    Vector3 e = DisneyEval(hd, V, hd.normal, L);
    *out_pdf = DisneyPdf(hd, V, hd.normal, L);
    std::memcpy(out_sample3, &s, 12);
    std::memcpy(out_eval3, &e, 12);
}

// generateHitData for a hit record + material index (kernel.cpp:76-172) -> 37 HitData words
void ref_generate_hitdata(void* h, const float* hit_words, float* hitdata_words) {
    RefCtx* c = (RefCtx*)h;
    Hit hit;
    std::memcpy(&hit.position, hit_words + 0, 12);
    std::memcpy(&hit.normal, hit_words + 3, 12);
    std::memcpy(&hit.tangent, hit_words + 6, 12);
    std::memcpy(&hit.bitangent, hit_words + 9, 12);
    std::memcpy(&hit.gnormal, hit_words + 12, 12);
    hit.valid = true;
    std::memcpy(&hit.objectID, hit_words + 16, 4);
    std::memcpy(&hit.material_id, hit_words + 17, 4);
    hit.tu = hit_words[18];
    hit.tv = hit_words[19];
    std::memcpy(&hit.triIdx, hit_words + 20, 4);
    HitData hd;
    std::memset(&hd, 0, sizeof(hd));
    generateHitData(c->ds, &c->ds->materials[hit.material_id], hd, hit);
    std::memcpy(hitdata_words, &hd, sizeof(HitData));
}

void ref_asl_shade(int id, float* rgb) {
    asl_shade(id, 0, 0, 0, 0, 0, 1, 0, 1, 0, 0, 1, 0, 0.5f, 0.5f, rgb[0], rgb[1], rgb[2]);
}

}  // extern "C"
