// er_objloader.hpp — scene ingest for `--load_object`: Wavefront OBJ text + MTL text -> ErTri records, the way
// ObjLoader::loadObjsRapid's stream path does it (src/ObjLoader.cpp:66-164):
//   * rapidobj's parse (third-party, pinned submodule of the reference): `v`, `vt` (2 components kept), `vn`,
//     `f` with v, v/vt, v//vn, v/vt/vn and negative indices, a new shape at every `g` / `o`, `usemtl` by name
//     from the `newmtl` lines of the material text (every other MTL line is dropped: ObjLoader.cpp:156-157);
//   * rapidobj::Triangulate: triangles kept; a quad is split along its SHORTER diagonal — (0,1,2)(0,2,3) when
//     |p0-p2|^2 < |p1-p3|^2, else (0,1,3)(1,2,3); larger polygons are fanned here (rapidobj runs earcut on
//     them: a deviation, Blender exports triangles and quads);
//   * per corner: vertex (x, y, -z), normal (nx, ny, -nz).normalized(), uv (u, v, 0)  (ObjLoader.cpp:106-108);
//   * --recompute_normals: face-area-weighted normals over all triangles sharing a position, keyed by exact
//     position and accumulated across the shapes seen so far, cross(edge2, edge1) (ObjLoader.cpp:53-64,112-116);
//   * shapes that hold line elements are skipped (ObjLoader.cpp:83);
//   * tangents: MikkTSpace's genTangSpaceDefault per object (er_mikktspace.hpp, src/mikktspaceCallback.cpp).
#pragma once
#include <math.h>

#include <map>
#include <sstream>
#include <string>
#include <tuple>
#include <vector>

#include "../../include/er_cuda.h"
#include "er_mikktspace.hpp"

namespace erh {

struct LoadedObject {
    std::string name;
    std::vector<ErTri> tris;
    std::vector<std::string> matNames;
};

namespace objdetail {
struct V3f { float x, y, z; };
inline V3f sub(V3f a, V3f b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3f cross(V3f a, V3f b) { return {(a.y * b.z - a.z * b.y), -(a.x * b.z - a.z * b.x), (a.x * b.y - a.y * b.x)}; }  // src/Vector.h:173-175
inline float length(V3f v) { return sqrtf(v.x * v.x + v.y * v.y + v.z * v.z); }
inline V3f normalized(V3f v) {  // src/Vector.h:186-189
    if (length(v) == 0) return v;
    const float l = length(v);
    return {v.x / l, v.y / l, v.z / l};
}
struct Corner { int p, t, n; };
inline bool parse_index(const char*& s, int count, int& out) {  // 1-based, negative = relative to the end
    char* e = nullptr;
    const long v = strtol(s, &e, 10);
    if (e == s) return false;
    s = e;
    out = v > 0 ? (int)v - 1 : v < 0 ? count + (int)v : -1;
    return true;
}
}  // namespace objdetail

inline std::vector<LoadedObject> load_obj_text(const std::string& obj_text, const std::string& mtl_text, bool recompute_normals) {
    using namespace objdetail;
    // material names: the `newmtl` lines only
    std::vector<std::string> mtl_names;
    {
        std::istringstream ms(mtl_text);
        std::string line;
        while (std::getline(ms, line)) {
            if (!line.empty() && line.back() == '\r') line.pop_back();
            if (line.compare(0, 7, "newmtl ") == 0 || line.compare(0, 7, "newmtl\t") == 0) {
                std::string n = line.substr(7);
                while (!n.empty() && (n.back() == ' ' || n.back() == '\t')) n.pop_back();
                size_t b = 0;
                while (b < n.size() && (n[b] == ' ' || n[b] == '\t')) b++;
                mtl_names.push_back(n.substr(b));
            }
        }
    }
    std::vector<float> pos, tex, nor;
    struct Shape {
        std::string name;
        std::vector<Corner> corners;
        std::vector<int> face_sizes;
        std::vector<int> face_mat;  // index into mtl_names or -1
        bool has_lines = false;
    };
    std::vector<Shape> shapes;
    int cur_mat = -1;
    auto shape = [&]() -> Shape& {
        if (shapes.empty()) shapes.push_back(Shape());
        return shapes.back();
    };
    std::istringstream os(obj_text);
    std::string line;
    while (std::getline(os, line)) {
        if (!line.empty() && line.back() == '\r') line.pop_back();
        size_t b = 0;
        while (b < line.size() && (line[b] == ' ' || line[b] == '\t')) b++;
        const char* s = line.c_str() + b;
        if (s[0] == 'v' && (s[1] == ' ' || s[1] == '\t')) {
            char* e;
            const char* q = s + 2;
            for (int k = 0; k < 3; k++) {
                pos.push_back(strtof(q, &e));
                q = e;
            }
        } else if (s[0] == 'v' && s[1] == 't' && (s[2] == ' ' || s[2] == '\t')) {
            char* e;
            const char* q = s + 3;
            const float u = strtof(q, &e);
            q = e;
            const float v = strtof(q, &e);
            tex.push_back(u);
            tex.push_back(e == q ? 0.f : v);
        } else if (s[0] == 'v' && s[1] == 'n' && (s[2] == ' ' || s[2] == '\t')) {
            char* e;
            const char* q = s + 3;
            for (int k = 0; k < 3; k++) {
                nor.push_back(strtof(q, &e));
                q = e;
            }
        } else if (s[0] == 'f' && (s[1] == ' ' || s[1] == '\t')) {
            Shape& sh = shape();
            const char* q = s + 2;
            int n = 0;
            for (;;) {
                while (*q == ' ' || *q == '\t') q++;
                if (!*q) break;
                Corner c{-1, -1, -1};
                if (!parse_index(q, (int)(pos.size() / 3), c.p)) break;
                if (*q == '/') {
                    q++;
                    if (*q != '/') parse_index(q, (int)(tex.size() / 2), c.t);
                    if (*q == '/') {
                        q++;
                        parse_index(q, (int)(nor.size() / 3), c.n);
                    }
                }
                sh.corners.push_back(c);
                n++;
            }
            if (n >= 3) {
                sh.face_sizes.push_back(n);
                sh.face_mat.push_back(cur_mat);
            } else {
                sh.corners.resize(sh.corners.size() - (size_t)n);
            }
        } else if ((s[0] == 'g' || s[0] == 'o') && (s[1] == ' ' || s[1] == '\t')) {
            Shape sh;
            sh.name = s + 2;
            while (!sh.name.empty() && (sh.name.back() == ' ' || sh.name.back() == '\t')) sh.name.pop_back();
            size_t k = 0;
            while (k < sh.name.size() && (sh.name[k] == ' ' || sh.name[k] == '\t')) k++;
            sh.name = sh.name.substr(k);
            shapes.push_back(sh);
        } else if (strncmp(s, "usemtl", 6) == 0 && (s[6] == ' ' || s[6] == '\t')) {
            std::string n = s + 7;
            while (!n.empty() && (n.back() == ' ' || n.back() == '\t')) n.pop_back();
            size_t k = 0;
            while (k < n.size() && (n[k] == ' ' || n[k] == '\t')) k++;
            n = n.substr(k);
            cur_mat = -1;
            for (size_t i = 0; i < mtl_names.size(); i++)
                if (mtl_names[i] == n) cur_mat = (int)i;
        } else if (s[0] == 'l' && (s[1] == ' ' || s[1] == '\t')) {
            shape().has_lines = true;
        }
    }

    auto P = [&](int i) -> V3f { return i >= 0 && (size_t)(3 * i + 2) < pos.size() ? V3f{pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]} : V3f{0, 0, 0}; };
    std::map<std::tuple<float, float, float>, std::vector<ErTri>> faces;  // shared by all shapes (ObjLoader.cpp:77)
    std::vector<LoadedObject> out;
    for (const Shape& sh : shapes) {
        if (sh.has_lines) continue;
        LoadedObject mo;
        mo.name = sh.name;
        size_t base = 0;
        auto emit = [&](const Corner& c0, const Corner& c1, const Corner& c2, int mat) {
            ErTri tri{};
            const Corner cs[3] = {c0, c1, c2};
            for (int j = 0; j < 3; j++) {
                const V3f p = P(cs[j].p);
                V3f n{0, 0, 0};
                if (cs[j].n >= 0 && (size_t)(3 * cs[j].n + 2) < nor.size()) n = {nor[3 * cs[j].n], nor[3 * cs[j].n + 1], nor[3 * cs[j].n + 2]};
                float tu = 0.f, tv = 0.f;
                if (cs[j].t >= 0 && (size_t)(2 * cs[j].t + 1) < tex.size()) tu = tex[2 * cs[j].t], tv = tex[2 * cs[j].t + 1];
                tri.vertices[j][0] = p.x, tri.vertices[j][1] = p.y, tri.vertices[j][2] = -p.z;
                const V3f nn = normalized(V3f{n.x, n.y, -n.z});
                tri.normals[j][0] = nn.x, tri.normals[j][1] = nn.y, tri.normals[j][2] = nn.z;
                tri.uv[j][0] = tu, tri.uv[j][1] = tv, tri.uv[j][2] = 0.f;
            }
            if (recompute_normals)
                for (int j = 0; j < 3; j++) faces[std::make_tuple(tri.vertices[j][0], tri.vertices[j][1], tri.vertices[j][2])].push_back(tri);
            mo.tris.push_back(tri);
            mo.matNames.push_back(mat >= 0 ? mtl_names[(size_t)mat] : std::string());
        };
        for (size_t f = 0; f < sh.face_sizes.size(); f++) {
            const int n = sh.face_sizes[f];
            const Corner* c = sh.corners.data() + base;
            base += (size_t)n;
            if (n == 3) {
                emit(c[0], c[1], c[2], sh.face_mat[f]);
            } else if (n == 4) {
                const V3f e02 = sub(P(c[0].p), P(c[2].p)), e13 = sub(P(c[1].p), P(c[3].p));
                const float d02 = e02.x * e02.x + e02.y * e02.y + e02.z * e02.z;
                const float d13 = e13.x * e13.x + e13.y * e13.y + e13.z * e13.z;
                const bool less = d02 < d13;
                emit(c[0], c[1], less ? c[2] : c[3], sh.face_mat[f]);
                emit(less ? c[0] : c[1], c[2], c[3], sh.face_mat[f]);
            } else {
                for (int k = 1; k + 1 < n; k++) emit(c[0], c[k], c[k + 1], sh.face_mat[f]);
            }
        }
        if (recompute_normals) {  // recompute_normals_face_weight (ObjLoader.cpp:53-64)
            for (ErTri& t : mo.tris)
                for (int j = 0; j < 3; j++) {
                    const std::vector<ErTri>& fs = faces[std::make_tuple(t.vertices[j][0], t.vertices[j][1], t.vertices[j][2])];
                    V3f n{0, 0, 0};
                    for (const ErTri& ft : fs) {
                        const V3f v0{ft.vertices[0][0], ft.vertices[0][1], ft.vertices[0][2]};
                        const V3f e2 = sub(V3f{ft.vertices[2][0], ft.vertices[2][1], ft.vertices[2][2]}, v0);
                        const V3f e1 = sub(V3f{ft.vertices[1][0], ft.vertices[1][1], ft.vertices[1][2]}, v0);
                        const V3f cr = cross(e2, e1);
                        n.x += cr.x, n.y += cr.y, n.z += cr.z;
                    }
                    const V3f nn = normalized(n);
                    t.normals[j][0] = nn.x, t.normals[j][1] = nn.y, t.normals[j][2] = nn.z;
                }
        }
        mikk::gen_tangents(mo.tris);
        out.push_back(std::move(mo));
    }
    return out;
}

}  // namespace erh
