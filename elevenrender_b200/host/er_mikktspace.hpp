// er_mikktspace.hpp — tangents for loaded objects: a restatement of MikkTSpace's genTangSpaceDefault
// (Morten S. Mikkelsen's published tangent-space algorithm; the reference vendors its C implementation as a
// pinned submodule and calls it through src/mikktspaceCallback.cpp:24-29,96-103) for triangle meshes, with
// the reference's callbacks: position / vertex normal / uv of Tri, SMOOTH_SHADING on (src/Definitions.h:3),
// m_setTSpaceBasic -> tangents[corner] and tangentsSign (the LAST corner written wins the sign).
//
// Steps (float arithmetic and summation order as published, so the results match bit for bit — checked
// against the reference's loader on the test meshes, tests/test_ingest_cpu.py):
//   1. weld corners with identical (position, normal, uv); mark triangles with two equal positions degenerate
//      and keep them out of the topology (they copy a tangent from a good triangle at the same vertex);
//   2. per triangle: first-order derivatives vOs/vOt (eq. 18/19), orientation = sign of the uv area,
//      magnitudes; a triangle with a zero uv area or zero magnitude "groups with any";
//   3. neighbours across shared edges with opposite winding;
//   4. groups: the fan of triangles around a welded vertex with equal orientation, flood-filled over the
//      neighbours (left, then right);
//   5. per group one tangent: sum over its triangles IN ASCENDING TRIANGLE ORDER of the derivative projected
//      into the vertex's normal plane and normalised, weighted by the angle of the triangle at the vertex;
//      normalise (default angular threshold 180 degrees: the whole group joins unless two projected derivatives
//      are exactly opposite).
#pragma once
#include <float.h>
#include <math.h>

#include <algorithm>
#include <map>
#include <tuple>
#include <vector>

#include "../../include/er_cuda.h"

namespace erh {
namespace mikk {

struct V { float x, y, z; };
inline V vsub(V a, V b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V vadd(V a, V b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V vscale(float f, V v) { return {f * v.x, f * v.y, f * v.z}; }
inline float vdot(V a, V b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline float vlen(V v) { return sqrtf(v.x * v.x + v.y * v.y + v.z * v.z); }
inline bool not_zero(float f) { return fabsf(f) > FLT_MIN; }
inline bool vnot_zero(V v) { return not_zero(v.x) || not_zero(v.y) || not_zero(v.z); }
inline V vnormalize(V v) { return vscale(1.0f / vlen(v), v); }
inline bool veq(V a, V b) { return a.x == b.x && a.y == b.y && a.z == b.z; }
inline V project(V n, V v) {  // v minus its component along n, normalised unless zero
    V r = vsub(v, vscale(vdot(n, v), n));
    if (vnot_zero(r)) r = vnormalize(r);
    return r;
}

struct TriInfo {
    int neighbor[3] = {-1, -1, -1};
    int group[3] = {-1, -1, -1};
    V vOs{0, 0, 0}, vOt{0, 0, 0};
    float fMagS = 0, fMagT = 0;
    bool orient_preserving = false, group_with_any = true;
    int org = 0;  // triangle index in the caller's array
};
struct Group {
    int vertex = -1;
    bool orient = false;
    std::vector<int> faces;
};
struct TSpace {
    V vOs{1, 0, 0};
    bool orient = false;
};

inline void gen_tangents(std::vector<ErTri>& tris) {
    const int total = (int)tris.size();
    if (!total) return;
    auto pos = [&](int t, int c) { return V{tris[t].vertices[c][0], tris[t].vertices[c][1], tris[t].vertices[c][2]}; };
    auto nor = [&](int t, int c) { return V{tris[t].normals[c][0], tris[t].normals[c][1], tris[t].normals[c][2]}; };
    // 1. weld: corner id = first corner (in triangle order) with the same position, normal and uv
    std::vector<int> corner_id((size_t)total * 3);
    {
        typedef std::tuple<float, float, float, float, float, float, float, float> Key;
        std::map<Key, int> first;
        for (int t = 0; t < total; t++)
            for (int c = 0; c < 3; c++) {
                // (-0 == +0 under the float compares MikkTSpace welds with; normalise the key accordingly)
                auto z = [](float f) { return f == 0.f ? 0.f : f; };
                const Key k(z(tris[t].vertices[c][0]), z(tris[t].vertices[c][1]), z(tris[t].vertices[c][2]), z(tris[t].normals[c][0]),
                            z(tris[t].normals[c][1]), z(tris[t].normals[c][2]), z(tris[t].uv[c][0]), z(tris[t].uv[c][1]));
                auto it = first.find(k);
                if (it == first.end()) it = first.emplace(k, t * 3 + c).first;
                corner_id[(size_t)t * 3 + c] = it->second;
            }
    }
    // good triangles first (stable), degenerate ones behind
    std::vector<int> order;
    std::vector<int> degenerate;
    for (int t = 0; t < total; t++) {
        const V p0 = pos(t, 0), p1 = pos(t, 1), p2 = pos(t, 2);
        (veq(p0, p1) || veq(p0, p2) || veq(p1, p2) ? degenerate : order).push_back(t);
    }
    const int n = (int)order.size();
    std::vector<TriInfo> info((size_t)n);
    std::vector<int> vid((size_t)n * 3);  // welded vertex of every corner of the good triangles
    for (int f = 0; f < n; f++) {
        info[f].org = order[f];
        for (int c = 0; c < 3; c++) vid[(size_t)f * 3 + c] = corner_id[(size_t)order[f] * 3 + c];
    }
    // 2. derivatives
    for (int f = 0; f < n; f++) {
        const int t = order[f];
        const V v1 = pos(t, 0), v2 = pos(t, 1), v3 = pos(t, 2);
        const float t21x = tris[t].uv[1][0] - tris[t].uv[0][0], t21y = tris[t].uv[1][1] - tris[t].uv[0][1];
        const float t31x = tris[t].uv[2][0] - tris[t].uv[0][0], t31y = tris[t].uv[2][1] - tris[t].uv[0][1];
        const V d1 = vsub(v2, v1), d2 = vsub(v3, v1);
        const float area2 = t21x * t31y - t21y * t31x;
        const V vOs = vsub(vscale(t31y, d1), vscale(t21y, d2));
        const V vOt = vadd(vscale(-t31x, d1), vscale(t21x, d2));
        TriInfo& ti = info[f];
        ti.orient_preserving = area2 > 0;
        if (not_zero(area2)) {
            const float fAbsArea = fabsf(area2);
            const float fLenOs = vlen(vOs), fLenOt = vlen(vOt);
            const float fS = ti.orient_preserving ? 1.0f : -1.0f;
            if (not_zero(fLenOs)) ti.vOs = vscale(fS / fLenOs, vOs);
            if (not_zero(fLenOt)) ti.vOt = vscale(fS / fLenOt, vOt);
            ti.fMagS = fLenOs / fAbsArea;
            ti.fMagT = fLenOt / fAbsArea;
            if (not_zero(ti.fMagS) && not_zero(ti.fMagT)) ti.group_with_any = false;
        }
    }
    // 3. neighbours: edges keyed by the sorted vertex pair; a pair matches when the windings are opposite,
    // candidates in ascending triangle order
    {
        struct Edge { int i0, i1, f; };
        std::vector<Edge> edges;
        edges.reserve((size_t)n * 3);
        for (int f = 0; f < n; f++)
            for (int c = 0; c < 3; c++) {
                const int a = vid[(size_t)f * 3 + c], b = vid[(size_t)f * 3 + (c < 2 ? c + 1 : 0)];
                edges.push_back({a < b ? a : b, a < b ? b : a, f});
            }
        std::sort(edges.begin(), edges.end(), [](const Edge& x, const Edge& y) { return std::tie(x.i0, x.i1, x.f) < std::tie(y.i0, y.i1, y.f); });
        // edge number and direction of (i0,i1) inside a triangle
        auto get_edge = [&](int f, int i0, int i1, int& a, int& b, int& edgenum) {
            const int* v = &vid[(size_t)f * 3];
            if (v[0] == i0 || v[0] == i1) {
                if (v[1] == i0 || v[1] == i1) {
                    edgenum = 0, a = v[0], b = v[1];
                } else {
                    edgenum = 2, a = v[2], b = v[0];
                }
            } else {
                edgenum = 1, a = v[1], b = v[2];
            }
        };
        for (size_t i = 0; i < edges.size(); i++) {
            const int i0 = edges[i].i0, i1 = edges[i].i1, f = edges[i].f;
            int a0, a1, ea;
            get_edge(f, i0, i1, a0, a1, ea);
            if (info[f].neighbor[ea] != -1) continue;
            for (size_t j = i + 1; j < edges.size() && edges[j].i0 == i0 && edges[j].i1 == i1; j++) {
                const int t = edges[j].f;
                int b0, b1, eb;
                get_edge(t, i0, i1, b1, b0, eb);  // flipped
                if (a0 == b0 && a1 == b1 && info[t].neighbor[eb] == -1) {
                    info[f].neighbor[ea] = t;
                    info[t].neighbor[eb] = f;
                    break;
                }
            }
        }
    }
    // 4. groups
    std::vector<Group> groups;
    {
        struct Item { int tri; };
        for (int f = 0; f < n; f++)
            for (int i = 0; i < 3; i++) {
                if (info[f].group_with_any || info[f].group[i] != -1) continue;
                const int g = (int)groups.size();
                groups.push_back(Group());
                groups[g].vertex = vid[(size_t)f * 3 + i];
                groups[g].orient = info[f].orient_preserving;
                info[f].group[i] = g;
                groups[g].faces.push_back(f);
                // depth-first over the neighbours: left edge first, then right (explicit stack, same visit order)
                std::vector<int> stack;
                const int nl = info[f].neighbor[i], nr = info[f].neighbor[i > 0 ? i - 1 : 2];
                if (nr >= 0) stack.push_back(nr);
                if (nl >= 0) stack.push_back(nl);
                while (!stack.empty()) {
                    const int t = stack.back();
                    stack.pop_back();
                    TriInfo& ti = info[t];
                    int k = -1;
                    for (int c = 0; c < 3; c++)
                        if (k < 0 && vid[(size_t)t * 3 + c] == groups[g].vertex) k = c;
                    if (k < 0) continue;
                    if (ti.group[k] != -1) continue;  // in this group already, or in another one
                    if (ti.group_with_any && ti.group[0] == -1 && ti.group[1] == -1 && ti.group[2] == -1)
                        ti.orient_preserving = groups[g].orient;  // the first group to reach it decides
                    if (ti.orient_preserving != groups[g].orient) continue;
                    groups[g].faces.push_back(t);
                    ti.group[k] = g;
                    const int l = ti.neighbor[k], r = ti.neighbor[k > 0 ? k - 1 : 2];
                    if (r >= 0) stack.push_back(r);
                    if (l >= 0) stack.push_back(l);
                }
            }
    }
    // 5. one tangent per (group, set of joining triangles)
    std::vector<TSpace> ts((size_t)total * 3);
    const float thres_cos = (float)cos((180.0f * (float)3.14159265358979323846) / 180.0f);
    auto vertex_normal = [&](int corner) { return nor(corner / 3, corner % 3); };
    auto vertex_pos = [&](int corner) { return pos(corner / 3, corner % 3); };
    auto eval = [&](const std::vector<int>& members, int vertex) {
        V sum{0, 0, 0};
        for (int f : members) {
            if (info[f].group_with_any) continue;
            int i = -1;
            for (int c = 0; c < 3; c++)
                if (i < 0 && vid[(size_t)f * 3 + c] == vertex) i = c;
            if (i < 0) continue;
            const V nrm = vertex_normal(vid[(size_t)f * 3 + i]);
            const V vOs = project(nrm, info[f].vOs);
            const int i2 = vid[(size_t)f * 3 + (i < 2 ? i + 1 : 0)], i1 = vid[(size_t)f * 3 + i], i0 = vid[(size_t)f * 3 + (i > 0 ? i - 1 : 2)];
            const V p0 = vertex_pos(i0), p1 = vertex_pos(i1), p2 = vertex_pos(i2);
            const V v1 = project(nrm, vsub(p0, p1)), v2 = project(nrm, vsub(p2, p1));
            float fCos = vdot(v1, v2);
            fCos = fCos > 1 ? 1 : (fCos < (-1) ? (-1) : fCos);
            const float fAngle = (float)acos((double)fCos);
            sum = vadd(sum, vscale(fAngle, vOs));
        }
        if (vnot_zero(sum)) sum = vnormalize(sum);
        return sum;
    };
    for (const Group& g : groups) {
        std::vector<std::vector<int>> uni_members;
        std::vector<V> uni_tangent;
        for (int f : g.faces) {
            int index = -1;
            for (int c = 0; c < 3; c++)
                if (index < 0 && info[f].group[c] == (int)(&g - groups.data())) index = c;
            if (index < 0) continue;
            const V nrm = vertex_normal(vid[(size_t)f * 3 + index]);
            const V vOs = project(nrm, info[f].vOs), vOt = project(nrm, info[f].vOt);
            std::vector<int> members;
            for (int t : g.faces) {
                const V vOs2 = project(nrm, info[t].vOs), vOt2 = project(nrm, info[t].vOt);
                const bool any = info[f].group_with_any || info[t].group_with_any;
                const bool same = info[f].org == info[t].org;
                const float fCosS = vdot(vOs, vOs2), fCosT = vdot(vOt, vOt2);
                if (any || same || (fCosS > thres_cos && fCosT > thres_cos)) members.push_back(t);
            }
            std::sort(members.begin(), members.end());
            size_t l = 0;
            while (l < uni_members.size() && uni_members[l] != members) l++;
            if (l == uni_members.size()) {
                uni_tangent.push_back(eval(members, g.vertex));
                uni_members.push_back(members);
            }
            TSpace& out = ts[(size_t)info[f].org * 3 + index];
            out.vOs = uni_tangent[l];
            out.orient = g.orient;
        }
    }
    // degenerate triangles copy the tangent space of the first good corner at the same welded vertex
    if (!degenerate.empty()) {
        std::map<int, int> first_good;  // welded vertex -> f * 3 + c, first in good-triangle order
        for (int f = 0; f < n; f++)
            for (int c = 0; c < 3; c++) first_good.emplace(vid[(size_t)f * 3 + c], f * 3 + c);
        for (int t : degenerate)
            for (int c = 0; c < 3; c++) {
                auto it = first_good.find(corner_id[(size_t)t * 3 + c]);
                if (it != first_good.end()) ts[(size_t)t * 3 + c] = ts[(size_t)info[it->second / 3].org * 3 + it->second % 3];
            }
    }
    for (int t = 0; t < total; t++)
        for (int c = 0; c < 3; c++) {
            const TSpace& s = ts[(size_t)t * 3 + c];
            tris[t].tangents[c][0] = s.vOs.x, tris[t].tangents[c][1] = s.vOs.y, tris[t].tangents[c][2] = s.vOs.z;
            tris[t].tangentsSign = s.orient ? 1.0f : -1.0f;
        }
}

}  // namespace mikk
}  // namespace erh
