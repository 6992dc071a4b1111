// er_host.h — host-side prerequisites of the device path (plain C++, no CUDA types):
//   * the BVH builder whose nodes[] / triIndices[] are bit-identical to BVH::build
//     (src/BVH.cpp:132-156,239-281,327-415,447-534)
//   * HDRI::generateCDF (src/HDRI.cpp:62-83)
//   * flattening of the reference's AoS scene into the device SoA layout of er_scene.h
//     (takes the place of copy_scene, src/SYCLCopy.cpp:3-104)
#pragma once
#include <stddef.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/er_cuda.h"

namespace erh {

struct F4 {
    float x, y, z, w;
};

// host image of the device geometry buffers (er_scene.h)
struct FlatScene {
    std::vector<F4> pairs;  // 4 per internal node, internal pre-order rank
    std::vector<F4> tri_v;  // 3 per BVH slot
    std::vector<F4> tri_n;  // 3 per BVH slot
    std::vector<F4> tri_a;  // 5 per BVH slot
    std::vector<F4> tri_g;  // 2 per GROUP of 4 consecutive slots: conservative box of the group's triangles
    std::vector<float> tri_slack;  // per slot: bound of |shadingPosition - geomPosition| (scratch of the flattening)
    float max_slack = 0.f;
};

// Returns 0 or a negative ErStatus; `err` receives the reason.
int bvh_build(const void* tris, size_t tri_count, size_t tri_stride, ErNode* nodes_out, int32_t* tri_indices_out,
              int nthreads, std::string& err);
int hdri_build_cdf(const ErTexture* tex, float* cdf_out, float* radiance_sum_out, std::string& err);
// device_slack: per-slot slack computed on the device (the per-slot records are then not built here, only the tree)
int flatten_geometry(const ErSceneDesc* d, FlatScene& out, std::string& err, const float* device_slack = nullptr);

// inverse of the pair-record packing, used by the T2 round-trip test: child boxes + leaf ranges back
// into reference pre-order nodes (only b1,b2,from,to of depth>=1 nodes are recoverable)
void unflatten_nodes(const FlatScene& fs, ErNode* nodes_out);

}  // namespace erh
