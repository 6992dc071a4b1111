// er_ingest.cpp — C-ABI entry of the scene ingest (host only; ../host/er_objloader.hpp, er_mikktspace.hpp).
#include <string.h>

#include <string>
#include <vector>

#include "../../include/er_cuda.h"
#include "../host/er_objloader.hpp"

namespace erh_api {
int fail_invalid(const char* msg);  // er_api.cu: sets er_last_error()
}

extern "C" ER_API int er_load_obj(const char* obj_text, size_t obj_len, const char* mtl_text, size_t mtl_len, int recompute_normals,
                                  ErTri* tris_out, size_t tri_capacity, size_t* tri_count, int32_t* object_of_tri,
                                  char* mat_names_out, size_t mat_names_capacity, size_t* mat_names_len) {
    if (!obj_text || !tri_count) return erh_api::fail_invalid("er_load_obj: null argument");
    try {
        const std::vector<erh::LoadedObject> objs =
            erh::load_obj_text(std::string(obj_text, obj_len), std::string(mtl_text ? mtl_text : "", mtl_text ? mtl_len : 0), recompute_normals != 0);
        size_t n = 0;
        std::string names;
        for (size_t o = 0; o < objs.size(); o++)
            for (size_t i = 0; i < objs[o].tris.size(); i++, n++) {
                if (tris_out && n < tri_capacity) {
                    tris_out[n] = objs[o].tris[i];
                    tris_out[n].objectID = (int32_t)o;  // Scene::addMeshObject numbers the objects in load order
                    if (object_of_tri) object_of_tri[n] = (int32_t)o;
                }
                names += objs[o].matNames[i];
                names += '\n';
            }
        *tri_count = n;
        if (mat_names_len) *mat_names_len = names.size();
        if (mat_names_out && mat_names_capacity) {
            const size_t k = names.size() < mat_names_capacity - 1 ? names.size() : mat_names_capacity - 1;
            memcpy(mat_names_out, names.data(), k);
            mat_names_out[k] = '\0';
        }
        return ER_OK;
    } catch (const std::exception& e) {
        return erh_api::fail_invalid(e.what());
    }
}
