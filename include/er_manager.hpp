// er_manager.hpp — the reference's RenderingManager (src/Managers.h:41-67, src/Managers.cpp:211-302) as
// a C++ class over the C ABI of er_cuda.h. Header-only; links against liber_cuda.so only.
//
// Same method names, argument meaning and error behaviour as the reference class:
//   start_rendering(scene)   Managers.cpp:234-275  device selection by "name|platform", BVH build, upload,
//                                                  asynchronous render of `sample_target` samples
//   get_render_info()        Managers.cpp:211-232  {samples}: dev_samples[0], i.e. completed samples + 1
//   get_pass(name)           Managers.cpp:287-302  new float[w*h*4] (caller owns it, as in the reference)
// Errors are C++ exceptions (std::runtime_error with er_last_error()), which the reference's command loop
// catches and logs (src/main.cpp:229-236). There is no CPU fallback: without a B200 every call throws.
//
// The scene is handed over through HostScene, a plain view of what the reference's Scene holds
// (src/Scene.h:17-52): the maintainer fills it from Scene's own arrays; the test
// program tests/cpp/manager_main.cpp fills it procedurally.
#pragma once
#include <stdexcept>
#include <string>
#include <vector>

#include "er_cuda.h"

namespace er {

struct RenderParameters {  // src/kernel.h:56-75
    unsigned width = 1280, height = 720, sampleTarget = 100;
    bool denoise = false;
    std::string device;       // "name|platform" as sent by --load_config; empty = first compatible device
    unsigned block_size = 8;  // accepted and ignored: the launch shape is the library's
};

struct RenderInfo {  // src/Managers.h:36-38
    unsigned samples = 0;
};

struct HostScene {  // what copy_scene deep-copies (src/SYCLCopy.cpp:3-104)
    unsigned x_res = 0, y_res = 0;
    std::vector<ErTri> tris;
    std::vector<ErMaterial> materials;
    std::vector<ErTexture> textures;  // texel pointers stay owned by the caller
    ErTexture hdri{};
    ErCamera camera{};
    std::vector<ErPointLight> pointLights;
    // filled by buildBVH() / buildCDF()
    std::vector<ErNode> nodes;
    std::vector<int32_t> triIndices;
    std::vector<float> cdf;
    float radianceSum = 0.f;

    void buildBVH(int nthreads = 0) {  // Scene::buildBVH, src/Scene.cpp:122-143 (same arrays, multi-threaded)
        nodes.resize(ER_BVH_NODE_COUNT);
        triIndices.resize(tris.size());
        if (er_bvh_build(tris.data(), tris.size(), sizeof(ErTri), nodes.data(), triIndices.data(), nthreads))
            throw std::runtime_error(er_last_error());
    }
    void buildCDF() {  // HDRI::generateCDF, src/HDRI.cpp:62-83
        cdf.resize((size_t)hdri.width * hdri.height + 1);
        if (er_hdri_build_cdf(&hdri, cdf.data(), &radianceSum)) throw std::runtime_error(er_last_error());
    }
};

inline int parsePass(const std::string& s) {  // src/kernel.cpp:50-73
    if (s == "beauty") return ER_PASS_BEAUTY;
    if (s == "denoise") return ER_PASS_DENOISE;
    if (s == "normal") return ER_PASS_NORMAL;
    if (s == "tangent") return ER_PASS_TANGENT;
    if (s == "bitangent") return ER_PASS_BITANGENT;
    throw std::runtime_error("unknown pass '" + s + "'");
}

class RenderingManager {
   public:
    RenderParameters rend_pars;

    explicit RenderingManager(RenderParameters p = RenderParameters()) : rend_pars(std::move(p)) {}
    RenderingManager(const RenderingManager&) = delete;
    RenderingManager& operator=(const RenderingManager&) = delete;
    ~RenderingManager() {
        if (ctx_) er_destroy(ctx_);  // joins the render thread (the reference detaches it)
    }

    // get_sycl_info's device list (src/CommandManager.cpp:303-362): same keys, CUDA devices
    static std::vector<ErDeviceInfo> devices() {
        int n = 0;
        must(er_device_count(&n));
        std::vector<ErDeviceInfo> out((size_t)n);
        for (int i = 0; i < n; i++) must(er_device_info(i, &out[(size_t)i]));
        return out;
    }

    void start_rendering(HostScene* scene, const ErRenderOpts* opts = nullptr) {
        if (!ctx_) must(er_create(select_device(), &ctx_));
        scene->x_res = rend_pars.width;
        scene->y_res = rend_pars.height;
        if (scene->nodes.empty()) scene->buildBVH();
        if (scene->cdf.empty()) scene->buildCDF();
        ErSceneDesc d{};
        d.x_res = scene->x_res;
        d.y_res = scene->y_res;
        d.tris = scene->tris.data();
        d.tri_count = scene->tris.size();
        d.tri_stride = sizeof(ErTri);
        d.nodes = scene->nodes.data();
        d.node_stride = sizeof(ErNode);
        d.tri_indices = scene->triIndices.data();
        d.materials = scene->materials.data();
        d.material_count = (int32_t)scene->materials.size();
        d.textures = scene->textures.data();
        d.texture_count = (int32_t)scene->textures.size();
        d.hdri = scene->hdri;
        d.hdri_cdf = scene->cdf.data();
        d.hdri_radiance_sum = scene->radianceSum;
        d.camera = &scene->camera;
        d.point_lights = scene->pointLights.data();
        d.point_light_count = (int32_t)scene->pointLights.size();
        must(er_upload_scene(ctx_, &d));  // renderSetup: copy_scene + setupKernel (src/kernel.cpp:651-677)
        ErRenderOpts o{};
        if (opts) o = *opts;
        must(er_render_start(ctx_, rend_pars.sampleTarget, &o));  // kernel_render_enqueue on the library's thread
    }

    RenderInfo get_render_info() {
        RenderInfo r;
        must(er_get_render_info(ctx_, &r.samples));
        return r;
    }

    float* get_pass(const std::string& pass) {
        float* out = new float[(size_t)rend_pars.width * rend_pars.height * 4];
        const int rc = er_get_pass(ctx_, parsePass(pass), out);
        if (rc) {
            delete[] out;
            must(rc);
        }
        return out;
    }

    void wait() { must(er_render_wait(ctx_)); }
    void abort() { must(er_render_abort(ctx_)); }
    ErCtx* context() { return ctx_; }

   private:
    ErCtx* ctx_ = nullptr;

    static void must(int rc) {
        if (rc) throw std::runtime_error(std::string("er_cuda: ") + er_last_error());
    }
    int select_device() const {  // NameSelector, src/Managers.cpp:191-208
        const std::vector<ErDeviceInfo> ds = devices();
        int first = -1;
        for (size_t i = 0; i < ds.size(); i++) {
            if (!ds[i].is_compatible) continue;
            if (first < 0) first = (int)i;
            if (std::string(ds[i].name) + "|" + ds[i].platform == rend_pars.device) return (int)i;
        }
        if (first < 0) throw std::runtime_error("er_cuda: no compute-capability-10 device");
        return first;
    }
};

}  // namespace er
