// er_hostscene.hpp — the host-side Scene of the reference (src/Scene.h:17-52, src/Scene.cpp:3-120) with its
// name-based pairing of textures and materials, and the host texture operations of src/Texture.cpp:60-153,
// kept layout-compatible with the C ABI records (ErTri / ErMaterial / ErTexture / ErCamera).
#pragma once
#include <math.h>

#include <map>
#include <string>
#include <vector>

#include "../../include/er_manager.hpp"

namespace erh {

// src/Math.hpp:12-20. The union is initialised through its float member and the routine then reads and
// writes x[1], which lies beyond it: the result is assembled from x[0] = 0, i.e. +0.0f for every input
// (SURVEY F11). Kept: an sRGB texel above 0.04045 becomes 0 exactly as in the reference.
inline float fast_pow(float a, float b) {
    union {
        float d;
        int x[2];
    } u = {a};
    u.x[1] = static_cast<int>(b * (u.x[1] - 1072632447) + 1072632447);
    u.x[0] = 0;
    return u.d;
}
inline float sRGBToLinear(float s) {  // src/Texture.cpp:114-121
    if (s <= 0.04045f) return s / 12.92f;
    return fast_pow((s + 0.055f) / 1.055f, 2.4f);
}

struct Texture {  // src/Texture.h:10-28
    enum class Filter { NO_FILTER, BILINEAR };
    enum class CS { LINEAR, sRGB };
    std::string name;
    int width = 1, height = 1, channels = 1;
    Filter filter = Filter::NO_FILTER;
    std::vector<float> data = std::vector<float>(1, 0.f);

    Texture() = default;
    Texture(const std::string& n, int w, int h, int ch, const float* d, size_t count, Filter f, CS cs)  // src/Texture.cpp:123-131
        : name(n), width(w), height(h), channels(ch), filter(f) {
        const size_t need = (size_t)w * h * ch;
        if (w <= 0 || h <= 0 || ch <= 0 || count < need) throw std::runtime_error("texture payload smaller than width*height*channels");
        data.assign(d, d + need);
        if (cs == CS::sRGB)
            for (float& v : data) v = sRGBToLinear(v);
    }
    static Texture color(float r, float g, float b) {  // src/Texture.cpp:133-141
        Texture t;
        t.channels = 3;
        t.data = {r, g, b};
        return t;
    }
    void mirror_x() {  // src/Texture.cpp:60-71
        std::vector<float> n(data.size());
        for (int x = 0; x < width; x++)
            for (int y = 0; y < height; y++)
                for (int c = 0; c < channels; c++) n[(size_t)channels * (y * width + x) + c] = data[(size_t)channels * (y * width + (width - x - 1)) + c];
        data.swap(n);
    }
    void mirror_y() {  // src/Texture.cpp:73-84
        std::vector<float> n(data.size());
        for (int x = 0; x < width; x++)
            for (int y = 0; y < height; y++)
                for (int c = 0; c < channels; c++) n[(size_t)channels * (y * width + x) + c] = data[(size_t)channels * ((height - y - 1) * width + x) + c];
        data.swap(n);
    }
    void pixel_shift(const float x_amount, const float y_amount) {  // src/Texture.cpp:113-127
        std::vector<float> n(data.size());
        for (int x = 0; x < width; x++) {
            const int sx = static_cast<int>(x + width * x_amount) % width;
            for (int y = 0; y < height; y++) {
                const int sy = static_cast<int>(y + height * y_amount) % height;
                for (int c = 0; c < channels; c++) n[(size_t)channels * (sy * width + sx) + c] = data[(size_t)channels * (y * width + x) + c];
            }
        }
        data.swap(n);
    }
    ErTexture view() const {
        ErTexture t{};
        t.width = width;
        t.height = height;
        t.channels = channels;
        t.filter = filter == Filter::BILINEAR ? ER_FILTER_BILINEAR : ER_FILTER_NONE;
        t.data = data.data();
        return t;
    }
};

struct Material {  // src/Material.h:6-56 (defaults :30-47)
    std::string name, albedo_map, emission_map, roughness_map, metallic_map, normal_map, opacity_map, transmission_map;
    ErMaterial m;
    Material() {
        m.albedoTextureID = m.emissionTextureID = m.roughnessTextureID = m.metallicTextureID = m.normalTextureID =
            m.opacityTextureID = m.transmissionTextureID = m.albedoShaderID = -1;
        m.albedo[0] = m.albedo[1] = m.albedo[2] = 0.5f;
        m.emission[0] = m.emission[1] = m.emission[2] = 0.f;
        m.opacity = 1, m.roughness = 1, m.metallic = 0, m.clearcoatGloss = 0, m.clearcoat = 0, m.anisotropic = 0, m.eta = 0,
        m.transmission = 0, m.specular = 0.5f, m.specularTint = 0, m.sheenTint = 0.5f, m.subsurface = 0, m.sheen = 0, m.ax = 0, m.ay = 0;
    }
    static Material DefaultMaterial() {
        Material mat;
        mat.name = "default";
        return mat;
    }
};

struct MeshObject {  // src/MeshObject.hpp:16-27 (the triangles live in Scene::tris)
    std::string name;
    size_t first = 0, triCount = 0;
    int objectID = 0;
};

class Scene {
   public:
    std::vector<Material> materials;
    std::vector<MeshObject> meshObjects;
    std::vector<Texture> textures;
    std::vector<ErTri> tris;
    std::vector<std::string> triMatNames;  // Tri::matName (src/Tri.h:20)
    std::vector<ErPointLight> pointLights;
    std::map<std::string, unsigned> textureIDs;
    unsigned x_res = 0, y_res = 0;
    Texture hdri = Texture::color(0.5f, 0.5f, 0.5f);  // HDRI() : HDRI(Vector3(0.5)) (src/HDRI.cpp:19)
    ErCamera camera;

    Scene() {  // src/Scene.h:41 + Camera defaults (src/Camera.h:9-21)
        materials.push_back(Material::DefaultMaterial());
        camera = ErCamera{};
        camera.focalLength = 35 * 0.001f;
        camera.sensorWidth = 36 * 0.001f;
        camera.sensorHeight = 24 * 0.001f;
        camera.aperture = 2.8f;
        camera.focusDistance = 1000000;
    }
    void addTexture(const Texture& t) {  // src/Scene.cpp:41-48: a name already present is ignored
        if (textureIDs.count(t.name) == 0) {
            textureIDs[t.name] = (unsigned)textures.size();
            textures.push_back(t);
        }
    }
    void addMaterial(const Material& m) { materials.push_back(m); }
    void addHDRI(const Texture& t) { hdri = t; }
    // src/Scene.cpp:51-68: the object id is its index in the scene
    void addMeshObject(const std::string& name, const std::vector<ErTri>& t, const std::vector<std::string>& matNames) {
        MeshObject mo;
        mo.name = name;
        mo.objectID = (int)meshObjects.size();
        mo.first = tris.size();
        mo.triCount = t.size();
        for (size_t i = 0; i < t.size(); i++) {
            ErTri tri = t[i];
            tri.objectID = mo.objectID;
            tris.push_back(tri);
            triMatNames.push_back(matNames[i]);
        }
        meshObjects.push_back(mo);
    }
    void pair_textures() {  // src/Scene.cpp:72-102 (the transmission map is not paired there either)
        for (Material& mt : materials)
            for (size_t j = 0; j < textures.size(); j++) {
                const std::string& n = textures[j].name;
                if (n.empty()) continue;
                if (n == mt.albedo_map) mt.m.albedoTextureID = (int)j;
                if (n == mt.emission_map) mt.m.emissionTextureID = (int)j;
                if (n == mt.roughness_map) mt.m.roughnessTextureID = (int)j;
                if (n == mt.metallic_map) mt.m.metallicTextureID = (int)j;
                if (n == mt.opacity_map) mt.m.opacityTextureID = (int)j;
                if (n == mt.normal_map) mt.m.normalTextureID = (int)j;
            }
    }
    void pair_materials() {  // src/Scene.cpp:104-120: material 0 unless a name matches; the LAST match wins
        for (size_t i = 0; i < tris.size(); i++) {
            tris[i].materialID = 0;
            for (size_t j = 0; j < materials.size(); j++)
                if (materials[j].name == triMatNames[i]) tris[i].materialID = (int)j;
        }
    }
    // the plain view er::RenderingManager uploads (texel pointers stay owned by this Scene)
    void fill(er::HostScene& hs) const {
        hs.x_res = x_res;
        hs.y_res = y_res;
        hs.tris = tris;
        hs.materials.clear();
        for (const Material& m : materials) hs.materials.push_back(m.m);
        hs.textures.clear();
        for (const Texture& t : textures) hs.textures.push_back(t.view());
        hs.hdri = hdri.view();
        hs.camera = camera;
        hs.pointLights = pointLights;
        hs.nodes.clear();
        hs.triIndices.clear();
        hs.cdf.clear();
    }
};

}  // namespace erh
