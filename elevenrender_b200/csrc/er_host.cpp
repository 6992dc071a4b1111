// er_host.cpp — host prerequisites: result-identical BVH build, HDRI CDF, SoA flattening.
// Plain C++ (g++ -O2 -ffp-contract=off); see er_host.h.
#include "er_host.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <limits>
#include <atomic>
#include <chrono>
#include <thread>

namespace erh {
namespace {

// Math.hpp:61-67 — NaN/-0 order sensitive, so NOT fminf/fmaxf
inline float minf(float a, float b) { return a < b ? a : b; }
inline float maxf(float a, float b) { return a > b ? a : b; }
// float -> int as the reference's x86-64 host build does it (cvttss2si): truncation, INT_MIN when
// NaN or out of range
inline int f2i(float f) {
    if (!(f >= -2147483648.0f && f < 2147483648.0f)) return std::numeric_limits<int>::min();
    return (int)f;
}
// Math.hpp:22-24
inline float mapf(float a, float b, float c, float d, float e) { return d + ((a - b) / (c - b)) * (e - d); }

struct B3 {
    float x, y, z;
};
inline float comp(const B3& v, int a) { return a == 0 ? v.x : a == 1 ? v.y : a == 2 ? v.z : v.x; }  // Vector.h:113-118

// BVH::boundsArea (BVH.cpp:474-482)
inline float area(const B3& b1, const B3& b2) {
    float x = b2.x - b1.x;
    float y = b2.y - b1.y;
    float z = b2.z - b1.z;
    return 2 * (x * y + x * z + y * z);
}

// BVH::boundsUnion (BVH.cpp:447-472); arguments by value as in the reference (outputs may alias inputs)
inline void bounds_union(B3 b1, B3 b2, B3 b3, B3 b4, B3& b5, B3& b6) {
    if (area(b1, b2) <= 0 || area(b3, b4) <= 0) {
        if (area(b1, b2) <= 0) {
            b5 = b3;
            b6 = b4;
        }
        if (area(b3, b4) <= 0) {
            b5 = b1;
            b6 = b2;
        }
    } else {
        b5.x = minf(b1.x, minf(b2.x, minf(b3.x, b4.x)));
        b5.y = minf(b1.y, minf(b2.y, minf(b3.y, b4.y)));
        b5.z = minf(b1.z, minf(b2.z, minf(b3.z, b4.z)));
        b6.x = maxf(b1.x, maxf(b2.x, maxf(b3.x, b4.x)));
        b6.y = maxf(b1.y, maxf(b2.y, maxf(b3.y, b4.y)));
        b6.z = maxf(b1.z, maxf(b2.z, maxf(b3.z, b4.z)));
    }
}

constexpr int DEPTH = ER_BVH_DEPTH;
constexpr int BINS = ER_BVH_SAHBINS;

struct Builder {
    const char* tris;
    size_t stride;
    size_t n;
    std::vector<B3> cen, tb1, tb2;  // per source triangle: centroid, box
    ErNode* nodes;
    int32_t* order;  // slot -> triangle; partitioned in place, final content = triIndices
    std::vector<int32_t> scratch;
    int par_depth;

    const float* vert(int32_t t) const { return (const float*)(tris + (size_t)t * stride); }

    void put_node(int idx, B3 b1, B3 b2, int from, int to, int depth) {
        ErNode& nd = nodes[idx];
        nd.b1[0] = b1.x, nd.b1[1] = b1.y, nd.b1[2] = b1.z;
        nd.b2[0] = b2.x, nd.b2[1] = b2.y, nd.b2[2] = b2.z;
        nd.from = from;
        nd.to = to;
        nd.idx = idx;
        nd.depth = depth;
        nd.valid = 1;
    }

    // BVH::bounds(std::vector<BVHTri>*) (BVH.cpp:497-534)
    void set_bounds(int from, int count, B3& b1, B3& b2) const {
        if (count <= 0) return;
        const float* v = vert(order[from]);
        b1 = B3{v[0], v[1], v[2]};
        b2 = b1;
        for (int i = 0; i < count; i++) {
            v = vert(order[from + i]);
            for (int k = 0; k < 3; k++) {
                b1.x = minf(v[3 * k + 0], b1.x);
                b1.y = minf(v[3 * k + 1], b1.y);
                b1.z = minf(v[3 * k + 2], b1.z);
            }
            for (int k = 0; k < 3; k++) {
                b2.x = maxf(v[3 * k + 0], b2.x);
                b2.y = maxf(v[3 * k + 1], b2.y);
                b2.z = maxf(v[3 * k + 2], b2.z);
            }
        }
    }

    // BVH::divideSAH (BVH.cpp:327-415). Stable in-place partition of order[from, from+count);
    // returns the number of triangles that go left.
    int divide_sah(int from, int count, const B3& totalB1, const B3& totalB2) {
        int bestBin = 0, bestAxis = 0;
        float bestHeuristic = std::numeric_limits<float>::max();
        for (int axis = 0; axis < 3; axis++) {
            B3 b1s[BINS], b2s[BINS];
            int cnt[BINS];
            for (int i = 0; i < BINS; i++) {
                cnt[i] = 0;
                b1s[i] = B3{0, 0, 0};
                b2s[i] = B3{0, 0, 0};
            }
            const float lo = comp(totalB1, axis), hi = comp(totalB2, axis);
            for (int i = 0; i < count; i++) {
                const int32_t t = order[from + i];
                int bin = 0;
                if (lo != hi) bin = f2i(mapf(comp(cen[t], axis), lo, hi, 0, BINS - 1));
                if (bin < 0) bin = 0;  // out of range is undefined behaviour in the reference
                if (bin > BINS - 1) bin = BINS - 1;
                cnt[bin]++;
                bounds_union(b1s[bin], b2s[bin], tb1[t], tb2[t], b1s[bin], b2s[bin]);
            }
            for (int i = 0; i < BINS; i++) {
                int count1 = 0, count2 = 0;
                B3 b1{0, 0, 0}, b2{0, 0, 0}, b3{0, 0, 0}, b4{0, 0, 0};
                for (int j = 0; j < i; j++) {
                    count1 += cnt[j];
                    bounds_union(b1, b2, b1s[j], b2s[j], b1, b2);
                }
                for (int k = i; k < BINS; k++) {
                    count2 += cnt[k];
                    bounds_union(b3, b4, b1s[k], b2s[k], b3, b4);
                }
                float heuristic = area(b1, b2) * static_cast<float>(count1) + area(b3, b4) * static_cast<float>(count2);
                if (heuristic < bestHeuristic) {
                    bestHeuristic = heuristic;
                    bestBin = i;
                    bestAxis = axis;
                }
            }
        }
        // division: no degenerate-axis guard here in the reference (x/0 -> INT_MIN -> left)
        const float lo = comp(totalB1, bestAxis), hi = comp(totalB2, bestAxis);
        int nl = 0, nr = 0;
        int32_t* tmp = scratch.data() + from;
        for (int i = 0; i < count; i++) {
            const int32_t t = order[from + i];
            int bin = f2i(mapf(comp(cen[t], bestAxis), lo, hi, 0, BINS - 1));
            if (bin < bestBin)
                order[from + nl++] = t;
            else
                tmp[nr++] = t;
        }
        if (nr) memcpy(order + from + nl, tmp, sizeof(int32_t) * (size_t)nr);
        return nl;
    }

    // BVH::buildAux (BVH.cpp:239-281); node index and slot range are known up front because the
    // tree is complete (left = idx+1, right = idx + 2^(18-depth)) and the partition is stable.
    void build(int depth, int idx, int from, int count) {
        B3 b1{0, 0, 0}, b2{0, 0, 0};
        set_bounds(from, count, b1, b2);
        if (depth == DEPTH) {
            put_node(idx, b1, b2, from, from + count, depth);
            return;
        }
        put_node(idx, b1, b2, 0, 0, depth);
        int nl = 0;
        if (count > 0) nl = divide_sah(from, count, b1, b2);
        const int li = idx + 1, ri = idx + (1 << (DEPTH - depth));
        if (depth < par_depth && count > 8192) {
            std::thread th([=]() { build(depth + 1, ri, from + nl, count - nl); });
            build(depth + 1, li, from, nl);
            th.join();
        } else {
            build(depth + 1, li, from, nl);
            build(depth + 1, ri, from + nl, count - nl);
        }
    }
};

}  // namespace

int bvh_build(const void* tris, size_t tri_count, size_t tri_stride, ErNode* nodes_out, int32_t* tri_indices_out,
              int nthreads, std::string& err) {
    if ((!tris && tri_count) || !nodes_out || (!tri_indices_out && tri_count)) {
        err = "er_bvh_build: null argument";
        return ER_ERR_INVALID;
    }
    if (tri_stride < 36 || tri_count > 0x7fffffffull) {
        err = "er_bvh_build: tri_stride must cover the 9 vertex floats and tri_count must fit int32";
        return ER_ERR_INVALID;
    }
    Builder b;
    b.tris = (const char*)tris;
    b.stride = tri_stride;
    b.n = tri_count;
    b.nodes = nodes_out;
    b.order = tri_indices_out;
    b.scratch.resize(tri_count ? tri_count : 1);
    b.cen.resize(tri_count);
    b.tb1.resize(tri_count);
    b.tb2.resize(tri_count);
    if (nthreads <= 0) nthreads = (int)std::thread::hardware_concurrency();
    if (nthreads < 1) nthreads = 1;
    b.par_depth = 0;
    while ((1 << b.par_depth) < nthreads) b.par_depth++;
    if (nthreads > 1) b.par_depth += 2;  // more tasks than threads: SAH splits are uneven
    for (size_t i = 0; i < tri_count; i++) {
        const float* v = b.vert((int32_t)i);
        // Tri::centroid (Tri.h:30-35): the double 3.0 narrows to float in operator/(float)
        b.cen[i] = B3{(v[0] + v[3] + v[6]) / 3.0f, (v[1] + v[4] + v[7]) / 3.0f, (v[2] + v[5] + v[8]) / 3.0f};
        // BVH::bounds(Tri) (BVH.cpp:484-493)
        b.tb1[i] = B3{minf(v[0], minf(v[3], v[6])), minf(v[1], minf(v[4], v[7])), minf(v[2], minf(v[5], v[8]))};
        b.tb2[i] = B3{maxf(v[0], maxf(v[3], v[6])), maxf(v[1], maxf(v[4], v[7])), maxf(v[2], maxf(v[5], v[8]))};
        tri_indices_out[i] = (int32_t)i;
    }
    memset(nodes_out, 0, sizeof(ErNode) * (size_t)ER_BVH_NODE_COUNT);
    b.build(0, 0, 0, (int)tri_count);
    return ER_OK;
}

// HDRI::generateCDF (HDRI.cpp:62-83) with Texture::getValueFromCoordinates' channel decode
int hdri_build_cdf(const ErTexture* tex, float* cdf_out, float* radiance_sum_out, std::string& err) {
    if (!tex || !tex->data || !cdf_out || !radiance_sum_out || tex->width <= 0 || tex->height <= 0 || tex->channels < 0) {
        err = "er_hdri_build_cdf: bad texture";
        return ER_ERR_INVALID;
    }
    const int w = tex->width, h = tex->height, ch = tex->channels;
    auto lum = [&](int x, int y) -> float {
        const size_t p = (size_t)y * w + x;
        float px = 0, py = 0, pz = 0;
        if (ch == 1)
            px = py = pz = tex->data[p];
        else if (ch == 2)
            px = tex->data[2 * p], py = tex->data[2 * p + 1];
        else if (ch >= 3)
            px = tex->data[ch * p], py = tex->data[ch * p + 1], pz = tex->data[ch * p + 2];
        return px + py + pz;
    };
    float radianceSum = 0;
    cdf_out[0] = 0;
    for (int j = 0; j < h; j++)
        for (int i = 0; i < w; i++) radianceSum += lum(i, j);
    size_t c = 0;
    for (int j = 0; j < h; j++)
        for (int i = 0; i < w; i++) {
            cdf_out[c + 1] = cdf_out[c] + lum(i, j) / radianceSum;
            c++;
        }
    *radiance_sum_out = radianceSum;
    return ER_OK;
}

namespace {
struct SubInfo {
    int count = 0;
    float slack = 0.f;
    int first_from = 0, last_to = 0;  // leaf range covered, in pre-order
    bool ok = true;                   // the leaves below tile [first_from, last_to)
};
struct Flattener {
    const ErSceneDesc* d;
    FlatScene* out;
    const float* tri_slack;  // per slot
    std::vector<int> leaf_from, leaf_to;  // per leaf, in pre-order (= slot) order
    const ErNode& node(int i) const { return *(const ErNode*)((const char*)d->nodes + (size_t)i * d->node_stride); }

    // returns count/slack of the subtree rooted at reference node `idx` (depth `depth`), and whether its
    // leaves tile [first_from, last_to) in pre-order; `rank` = internal pre-order rank of that node when
    // depth < 18. The top `par_depth` levels fork a thread per left subtree (disjoint pair records).
    SubInfo walk(int idx, int depth, int rank, int par_depth, size_t T, int ord = 0) {
        if (depth == DEPTH) {
            const ErNode& nd = node(idx);
            SubInfo s{nd.to - nd.from, 0.f, nd.from, nd.to, true};
            if (nd.from < 0 || nd.to < nd.from || (size_t)nd.to > T) {
                s.ok = false;
                return s;
            }
            leaf_from[(size_t)ord] = nd.from;
            leaf_to[(size_t)ord] = nd.to;
            for (int i = nd.from; i < nd.to; i++) s.slack = std::max(s.slack, tri_slack[i]);
            return s;
        }
        const int li = idx + 1, ri = idx + (1 << (DEPTH - depth));
        const ErNode& L = node(li);
        const ErNode& R = node(ri);
        SubInfo sl, sr;
        if (depth < par_depth) {
            std::thread th([&]() { sl = walk(li, depth + 1, rank + 1, par_depth, T, ord); });
            sr = walk(ri, depth + 1, rank + (1 << (DEPTH - 1 - depth)), par_depth, T, ord + (1 << (DEPTH - 1 - depth)));
            th.join();
        } else {
            sl = walk(li, depth + 1, rank + 1, par_depth, T, ord);
            sr = walk(ri, depth + 1, rank + (1 << (DEPTH - 1 - depth)), par_depth, T, ord + (1 << (DEPTH - 1 - depth)));
        }
        F4* q = out->pairs.data() + 4 * (size_t)rank;
        q[0] = F4{L.b1[0], L.b1[1], L.b1[2], L.b2[0]};
        q[1] = F4{L.b2[1], L.b2[2], R.b1[0], R.b1[1]};
        q[2] = F4{R.b1[2], R.b2[0], R.b2[1], R.b2[2]};
        if (depth < DEPTH - 1) {
            float cl, cr;
            memcpy(&cl, &sl.count, 4);
            memcpy(&cr, &sr.count, 4);
            q[3] = F4{cl, cr, sl.slack, sr.slack};
        } else {
            float a, b, c;
            memcpy(&a, &L.from, 4);
            memcpy(&b, &L.to, 4);
            memcpy(&c, &R.to, 4);
            q[3] = F4{a, b, c, std::max(sl.slack, sr.slack)};
        }
        return SubInfo{sl.count + sr.count, std::max(sl.slack, sr.slack), sl.first_from, sr.last_to,
                       sl.ok && sr.ok && sl.last_to == sr.first_from};
    }
};
}  // namespace

int flatten_geometry(const ErSceneDesc* d, FlatScene& out, std::string& err, const float* device_slack) {
    const size_t T = d->tri_count;
    if (T > 0x7fffffffull) {
        err = "tri_count must fit int32";
        return ER_ERR_INVALID;
    }
    if (!d->nodes || d->node_stride < sizeof(ErNode)) {
        err = "nodes missing or node_stride < 44";
        return ER_ERR_INVALID;
    }
    if (T && (!d->tris || !d->tri_indices || d->tri_stride < sizeof(ErTri))) {
        err = "tris/tri_indices missing or tri_stride < 156";
        return ER_ERR_INVALID;
    }
    const bool timing = getenv("ER_TIMING") != nullptr;
    auto t_mark = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) {
        if (!timing) return;
        const auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "  flatten %-10s %8.2f ms\n", what, std::chrono::duration<double, std::milli>(now - t_mark).count());
        t_mark = now;
    };
    lap("tiling");
    out.pairs.assign(4 * (size_t)((1 << DEPTH) - 1), F4{0, 0, 0, 0});
    // device_slack != nullptr: the per-slot records were built on the device (er_api.cu k_flatten_tris) and only
    // their slack came back; what is left for the host is the tree
    if (!device_slack) {
        out.tri_v.resize(3 * T);
        out.tri_n.resize(3 * T);
        out.tri_a.resize(5 * T);
        out.tri_slack.resize(T);  // lives with the staging: no fresh pages per upload
    }
    Flattener f;
    f.d = d;
    f.out = &out;
    f.tri_slack = device_slack ? device_slack : out.tri_slack.data();
    float* const slack_out = device_slack ? nullptr : out.tri_slack.data();
    lap("resize");
    const char* slack_env = getenv("ER_SLACK");
    const bool legacy_slack = slack_env && !strcmp(slack_env, "legacy");
    // per-slot records: independent of each other, so the slots are split over the host threads (10 M
    // triangles: 1 s single-threaded, most of er_upload_scene)
    std::atomic<bool> bad_index{false}, bad_material{false};
    auto fill = [&](size_t s0, size_t s1) {
    for (size_t s = s0; s < s1; s++) {
        const int32_t t = d->tri_indices[s];
        if (t < 0 || (size_t)t >= T) {
            bad_index = true;
            return;
        }
        const ErTri& tr = *(const ErTri*)((const char*)d->tris + (size_t)t * d->tri_stride);
        if (tr.materialID < 0 || tr.materialID >= d->material_count) bad_material = true;  // (every triangle is in exactly one slot)
        for (int k = 0; k < 3; k++) {
            out.tri_v[3 * s + k] = F4{tr.vertices[k][0], tr.vertices[k][1], tr.vertices[k][2], 0.f};
            out.tri_n[3 * s + k] = F4{tr.normals[k][0], tr.normals[k][1], tr.normals[k][2], 0.f};
        }
        float oid, mid;
        memcpy(&oid, &tr.objectID, 4);
        memcpy(&mid, &tr.materialID, 4);
        out.tri_a[5 * s + 0] = F4{tr.uv[0][0], tr.uv[0][1], tr.uv[1][0], tr.uv[1][1]};
        out.tri_a[5 * s + 1] = F4{tr.uv[2][0], tr.uv[2][1], tr.tangentsSign, oid};
        out.tri_a[5 * s + 2] = F4{tr.tangents[0][0], tr.tangents[0][1], tr.tangents[0][2], mid};
        out.tri_a[5 * s + 3] = F4{tr.tangents[1][0], tr.tangents[1][1], tr.tangents[1][2], 0.f};
        out.tri_a[5 * s + 4] = F4{tr.tangents[2][0], tr.tangents[2][1], tr.tangents[2][2], 0.f};
        // slack >= |shadingPosition - geomPosition| (Tri.h:100-108). With P the hit point, p_i = P - ((P - v_i).n_i) n_i
        // and the shading position a convex combination of the p_i:
        //   |shadingPosition - P| <= max_i |n_i| |(P - v_i).n_i| <= max_i |n_i| max_j |(v_j - v_i).n_i|
        // (P lies in the triangle, a linear function peaks at a vertex): O(edge^2 * curvature) on a smooth
        // mesh. The rounding of P itself is covered by the ray-side margins of the traversal, which assume
        // |n_i| ~ 1; other normals (or ER_SLACK=legacy) get the cruder bound maxEdge * max|n_i|^2.
        // Non-finite input -> inf -> never pruned.
        double e2 = 0, n2 = 0, tight = 0;
        for (int k = 0; k < 3; k++) {
            const float* a = tr.vertices[k];
            const float* b = tr.vertices[(k + 1) % 3];
            double dx = (double)a[0] - b[0], dy = (double)a[1] - b[1], dz = (double)a[2] - b[2];
            e2 = std::max(e2, dx * dx + dy * dy + dz * dz);
            const float* nn = tr.normals[k];
            const double nk2 = (double)nn[0] * nn[0] + (double)nn[1] * nn[1] + (double)nn[2] * nn[2];
            n2 = std::max(n2, nk2);
            for (int j = 0; j < 3; j++) {
                const float* c = tr.vertices[j];
                const double dt = ((double)c[0] - a[0]) * nn[0] + ((double)c[1] - a[1]) * nn[1] + ((double)c[2] - a[2]) * nn[2];
                tight = std::max(tight, fabs(dt) * sqrt(nk2));
            }
        }
        double sl = sqrt(e2) * n2 * 1.001 + 1e-30;
        if (!legacy_slack && n2 <= 1.21) sl = std::min(sl, tight * 1.001 + 1e-30);
        if (!(sl == sl) || !(e2 == e2) || !(n2 == n2) || !(tight == tight)) sl = INFINITY;
        slack_out[s] = (float)std::min(sl, (double)INFINITY);
        if ((double)slack_out[s] < sl) slack_out[s] = nextafterf(slack_out[s], INFINITY);
    }
    };
    if (!device_slack) {
        size_t nt = std::thread::hardware_concurrency();
        if (nt < 1) nt = 1;
        if (nt > 64) nt = 64;
        if (T < 65536) nt = 1;
        std::vector<std::thread> pool;
        for (size_t k = 1; k < nt; k++) pool.emplace_back(fill, T * k / nt, T * (k + 1) / nt);
        fill(0, T / nt);
        for (auto& th : pool) th.join();
    }
    if (bad_index) {
        err = "tri_indices entry out of range";
        return ER_ERR_INVALID;
    }
    if (bad_material) {
        err = "triangle materialID out of range";
        return ER_ERR_INVALID;
    }
    lap("triangles");
    // the leaves (pre-order, depth 18) must tile [0, T) in order: that is what BVH::build produces
    f.leaf_from.assign((size_t)1 << DEPTH, 0);
    f.leaf_to.assign((size_t)1 << DEPTH, 0);
    SubInfo root = f.walk(0, 0, 0, 4, T);
    if (!root.ok || root.first_from != 0 || (size_t)root.last_to != T) {
        err = "BVH leaves do not tile [0, tri_count) in pre-order";
        return ER_ERR_INVALID;
    }
    out.max_slack = root.slack;
    lap("walk");
    // Group boxes (round 2). The warp-cooperative leaf test (wf_leaf_coop_g) tests one conservative box per group of
    // up to 4 triangles of a leaf first and runs Moeller-Trumbore only on the triangles of the boxes the ray meets.
    // The triangles of a leaf are grouped by the Morton order of their centroids (slots keep their place: a group
    // names its triangles by their offset from the leaf's first slot). Leaf `ord` (pre-order = slot order) owns the
    // groups starting at (from >> 2) + ord — disjoint for consecutive leaves, computable on the device from the
    // leaf-pair record and the path. Record: (lo.xyz, hi.x) (hi.yz, max|coordinate|, 4 offsets as bytes, 255 = none).
    // The reference tests EVERY triangle of a visited leaf, so a box must never reject a ray for which Tri::hit passes
    // in float arithmetic: it is the exact bounding box widened by 4e-5 x (largest |coordinate|) + 1e-6 (hundreds of
    // ulps; the ray-origin part of the rounding is added at test time). Non-finite coordinates: an infinite box.
    // Leaves with more than 255 triangles get no groups (the device falls back to the full loop).
    // Measured slower than the plain cooperative loop (er_wavefront.cuh): built only when ER_GROUP_BOXES=1.
    out.tri_g.clear();
    if (const char* gbx = getenv("ER_GROUP_BOXES"); !device_slack && gbx && atoi(gbx) != 0) {
        const size_t LEAVES = (size_t)1 << DEPTH;
        out.tri_g.assign(2 * ((T >> 2) + LEAVES + 2), F4{0, 0, 0, 0});
        auto boxes = [&](size_t l0, size_t l1) {
            std::vector<std::pair<uint32_t, int>> order;
            for (size_t l = l0; l < l1; l++) {
                const int from = f.leaf_from[l], to = f.leaf_to[l], n = to - from;
                if (n <= 0 || n > 255) continue;
                float clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
                auto centroid = [&](int s, int a) {
                    const F4* v = &out.tri_v[3 * (size_t)s];
                    const float c[3][3] = {{v[0].x, v[0].y, v[0].z}, {v[1].x, v[1].y, v[1].z}, {v[2].x, v[2].y, v[2].z}};
                    return (c[0][a] + c[1][a] + c[2][a]) * (1.0f / 3.0f);
                };
                for (int s = from; s < to; s++)
                    for (int a = 0; a < 3; a++) {
                        const float c = centroid(s, a);
                        if (c < clo[a]) clo[a] = c;
                        if (c > chi[a]) chi[a] = c;
                    }
                order.clear();
                for (int s = from; s < to; s++) {
                    uint32_t code = 0;
                    uint32_t q[3];
                    for (int a = 0; a < 3; a++) {
                        const float w = chi[a] - clo[a];
                        float u = w > 0 ? (centroid(s, a) - clo[a]) / w : 0.f;
                        if (!(u >= 0.f)) u = 0.f;
                        if (u > 1.f) u = 1.f;
                        q[a] = (uint32_t)(u * 1023.0f);
                    }
                    for (int bit = 9; bit >= 0; bit--)
                        for (int a = 0; a < 3; a++) code = (code << 1) | ((q[a] >> bit) & 1u);
                    order.push_back({code, s});
                }
                std::sort(order.begin(), order.end());
                const size_t gb = ((size_t)from >> 2) + l;
                for (int g = 0; 4 * g < n; g++) {
                    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY}, mx = 0.f;
                    bool finite = true;
                    uint32_t offs = 0xffffffffu;
                    for (int j = 0; j < 4 && 4 * g + j < n; j++) {
                        const int s = order[(size_t)(4 * g + j)].second;
                        offs = (offs & ~(0xffu << (8 * j))) | ((uint32_t)(s - from) << (8 * j));
                        for (int k = 0; k < 3; k++) {
                            const F4& v = out.tri_v[3 * (size_t)s + k];
                            const float c[3] = {v.x, v.y, v.z};
                            for (int a = 0; a < 3; a++) {
                                if (!(c[a] - c[a] == 0.f)) finite = false;
                                lo[a] = std::min(lo[a], c[a]);
                                hi[a] = std::max(hi[a], c[a]);
                                mx = std::max(mx, fabsf(c[a]));
                            }
                        }
                    }
                    const float m = 4e-5f * mx + 1e-6f;
                    float ow;
                    memcpy(&ow, &offs, 4);
                    if (!finite) {
                        out.tri_g[2 * (gb + g)] = F4{-INFINITY, -INFINITY, -INFINITY, INFINITY};
                        out.tri_g[2 * (gb + g) + 1] = F4{INFINITY, INFINITY, 0.f, ow};
                    } else {
                        out.tri_g[2 * (gb + g)] = F4{lo[0] - m, lo[1] - m, lo[2] - m, hi[0] + m};
                        out.tri_g[2 * (gb + g) + 1] = F4{hi[1] + m, hi[2] + m, mx, ow};
                    }
                }
            }
        };
        size_t nt = std::thread::hardware_concurrency();
        if (nt < 1) nt = 1;
        if (nt > 64) nt = 64;
        if (T < 65536) nt = 1;
        std::vector<std::thread> pool;
        for (size_t k = 1; k < nt; k++) pool.emplace_back(boxes, LEAVES * k / nt, LEAVES * (k + 1) / nt);
        boxes(0, LEAVES / nt);
        for (auto& th : pool) th.join();
    }
    lap("groups");
    return ER_OK;
}

void unflatten_nodes(const FlatScene& fs, ErNode* nodes_out) {
    memset(nodes_out, 0, sizeof(ErNode) * (size_t)ER_BVH_NODE_COUNT);
    struct It {
        int idx, depth, rank;
    };
    std::vector<It> stack;
    stack.push_back({0, 0, 0});
    while (!stack.empty()) {
        It it = stack.back();
        stack.pop_back();
        if (it.depth == DEPTH) continue;
        const F4* q = fs.pairs.data() + 4 * (size_t)it.rank;
        const int li = it.idx + 1, ri = it.idx + (1 << (DEPTH - it.depth));
        ErNode& L = nodes_out[li];
        ErNode& R = nodes_out[ri];
        L.b1[0] = q[0].x, L.b1[1] = q[0].y, L.b1[2] = q[0].z, L.b2[0] = q[0].w, L.b2[1] = q[1].x, L.b2[2] = q[1].y;
        R.b1[0] = q[1].z, R.b1[1] = q[1].w, R.b1[2] = q[2].x, R.b2[0] = q[2].y, R.b2[1] = q[2].z, R.b2[2] = q[2].w;
        L.idx = li, R.idx = ri, L.depth = R.depth = it.depth + 1, L.valid = R.valid = 1;
        if (it.depth == DEPTH - 1) {
            memcpy(&L.from, &q[3].x, 4);
            memcpy(&L.to, &q[3].y, 4);
            R.from = L.to;
            memcpy(&R.to, &q[3].z, 4);
        }
        stack.push_back({ri, it.depth + 1, it.rank + (1 << (DEPTH - 1 - it.depth))});
        stack.push_back({li, it.depth + 1, it.rank + 1});
    }
}

}  // namespace erh
