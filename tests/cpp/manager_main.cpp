// Drives er::RenderingManager (include/er_manager.hpp) the way the reference's command loop drives its
// RenderingManager (src/main.cpp:36-187): load scene, start, poll get_render_info, get_pass.
//   manager_main <scene.bin> <out.bin>
// scene.bin (written by tests/test_cpp_manager.py): int32 x_res, y_res, spp, rng_order, ntris, nmats, hw, hh, hch,
// then ErTri[ntris], ErMaterial[nmats], ErCamera, float hdri[hw*hh*hch].
// out.bin: uint32 samples (get_render_info), then the beauty, normal, tangent, bitangent passes (w*h*4 floats each).
// Exit codes: 0 ok, 2 usage / io, 3 the library reported an error (e.g. no B200: there is no CPU fallback).
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <thread>

#include "er_manager.hpp"

template <class T>
static bool rd(FILE* f, T* p, size_t n) {
    return fread(p, sizeof(T), n, f) == n;
}

int main(int argc, char** argv) {
    if (argc != 3) {
        fprintf(stderr, "usage: %s scene.bin out.bin\n", argv[0]);
        return 2;
    }
    FILE* f = fopen(argv[1], "rb");
    if (!f) return 2;
    int32_t h[9];
    if (!rd(f, h, 9)) return 2;
    er::HostScene scene;
    scene.tris.resize((size_t)h[4]);
    scene.materials.resize((size_t)h[5]);
    std::vector<float> texels((size_t)h[6] * h[7] * h[8]);
    if (!rd(f, scene.tris.data(), scene.tris.size()) || !rd(f, scene.materials.data(), scene.materials.size()) ||
        !rd(f, &scene.camera, 1) || !rd(f, texels.data(), texels.size()))
        return 2;
    fclose(f);
    scene.hdri = ErTexture{h[6], h[7], h[8], ER_FILTER_NONE, texels.data()};
    try {
        er::RenderParameters pars;
        pars.width = (unsigned)h[0];
        pars.height = (unsigned)h[1];
        pars.sampleTarget = (unsigned)h[2];
        er::RenderingManager rm(pars);
        for (const ErDeviceInfo& d : er::RenderingManager::devices())
            fprintf(stderr, "device %s|%s sm=%d compatible=%d\n", d.name, d.platform, d.max_compute_units, d.is_compatible);
        ErRenderOpts o{};
        o.rng_order = h[3];
        rm.start_rendering(&scene, &o);
        // the client polls --get_info until the target is reached (N+1 semantics)
        unsigned seen = 0;
        for (int i = 0; i < 60000; i++) {
            seen = rm.get_render_info().samples;
            if (seen >= pars.sampleTarget + 1) break;
            std::this_thread::sleep_for(std::chrono::milliseconds(1));
        }
        rm.wait();
        const unsigned samples = rm.get_render_info().samples;
        FILE* g = fopen(argv[2], "wb");
        if (!g) return 2;
        fwrite(&samples, 4, 1, g);
        for (const char* p : {"beauty", "normal", "tangent", "bitangent"}) {
            std::unique_ptr<float[]> buf(rm.get_pass(p));
            fwrite(buf.get(), sizeof(float), (size_t)pars.width * pars.height * 4, g);
        }
        fclose(g);
        fprintf(stderr, "rendered %u samples (get_render_info = %u)\n", pars.sampleTarget, samples);
    } catch (const std::exception& e) {
        fprintf(stderr, "error: %s\n", e.what());
        return 3;
    }
    return 0;
}
