// er_trace.cuh — nearest-hit query: BVH::transverse + BVH::intersect + BVH::intersectNode +
// Tri::hit (src/BVH.cpp:27-120, src/Tri.h:41-144) restated for the pair-record layout.
//
// Result contract: the winner is the lexicographic minimum of (dist, rank)
// over all Möller–Trumbore-passing triangles in the reference's visit set, where
//   dist = |hit.position - origin| with the shadow-terminator position (Tri.h:100-117), and
//   rank = the reference's visit order (DFS left-first; sibling leaves right-before-left,
//          BVH.cpp:81-87; ascending slot inside a leaf).
// The reference keeps the FIRST strict minimum in visit order, which is that minimum; so the
// traversal order here is free (near child first) and distance pruning is allowed as long as a
// node is only dropped when nothing below it can beat the current best (slack, er_scene.h).
#pragma once
#include "er_device.cuh"

struct HitRec {
    int slot;  // -1 = miss
    float u, v, t;
    float dist;
    uint32_t rank;
};

struct TraceStats {
    unsigned box_tests, tri_tests, candidates;
};

// BVH::intersect (BVH.cpp:27-61) with 1/d hoisted (same IEEE op, same bits). Returns overlap and
// the entry distance tmin used only for ordering/pruning.
ER_D bool er_box(V3 b1, V3 b2, V3 o, V3 inv, float& tmin_out) {
    float t1 = (b1.x - o.x) * inv.x;
    float t2 = (b2.x - o.x) * inv.x;
    float t3 = (b1.y - o.y) * inv.y;
    float t4 = (b2.y - o.y) * inv.y;
    float t5 = (b1.z - o.z) * inv.z;
    float t6 = (b2.z - o.z) * inv.z;
    float tmin = er_maxf(er_maxf(er_minf(t1, t2), er_minf(t3, t4)), er_minf(t5, t6));
    float tmax = er_minf(er_minf(er_maxf(t1, t2), er_maxf(t3, t4)), er_maxf(t5, t6));
    tmin_out = tmin;
    if (tmax < 0) return false;
    if (tmin > tmax) return false;
    return true;
}

// Möller–Trumbore exactly as Tri.h:43-73.
ER_D bool er_mt(V3 v0, V3 v1, V3 v2, V3 o, V3 d, float& u, float& v, float& t) {
    const float EPSILON = 0.0000001f;
    V3 edge1 = v1 - v0;
    V3 edge2 = v2 - v0;
    V3 pvec = cross(d, edge2);
    float det = dot(edge1, pvec);
    float inv_det = 1.0f / det;
    if (det > -EPSILON && det < EPSILON) return false;
    V3 tvec = o - v0;
    u = dot(tvec, pvec) * inv_det;
    if (u < 0 || u > 1) return false;
    V3 qvec = cross(tvec, edge1);
    v = dot(d, qvec) * inv_det;
    if (v < 0 || (u + v) > 1) return false;
    t = dot(edge2, qvec) * inv_det;
    if (t < 0) return false;
    return true;
}

ER_D V3 er_project_on_plane(V3 position, V3 origin, V3 normal) {  // Tri.h:37-39
    return position - dot(position - origin, normal) * normal;
}

// hit.position of Tri.h:76-112 for a candidate (u,v,t); also returns the shading normal.
ER_D V3 er_hit_position(V3 v0, V3 v1, V3 v2, V3 n0, V3 n1, V3 n2, V3 o, V3 d, float u, float v, float t,
                        V3& shadingNormal) {
    V3 geomPosition = o + d * t;
    shadingNormal = normalized(n0 + (n1 - n0) * u + (n2 - n0) * v);
    V3 p0 = er_project_on_plane(geomPosition, v0, n0);
    V3 p1 = er_project_on_plane(geomPosition, v1, n1);
    V3 p2 = er_project_on_plane(geomPosition, v2, n2);
    V3 shadingPosition = p0 + (p1 - p0) * u + (p2 - p0) * v;
    bool convex = dot(shadingPosition - geomPosition, shadingNormal) > 0;
    return convex ? shadingPosition : geomPosition;
}

ER_D V3 ld3(const float4* __restrict__ p) {
    float4 q = __ldg(p);
    return v3(q.x, q.y, q.z);
}

// One leaf: slots [from,to); rank_base = rank of slot `from`.
template <bool COUNT>
ER_D bool er_leaf(const DScene& S, bool any, int from, int to, uint32_t rank_base, V3 o, V3 d, HitRec& best,
                  TraceStats& st) {
    for (int i = from; i < to; i++) {
        const float4* pv = S.tri_v + 3 * (size_t)i;
        V3 v0 = ld3(pv), v1 = ld3(pv + 1), v2 = ld3(pv + 2);
        float u, v, t;
        if (COUNT) st.tri_tests++;
        if (!er_mt(v0, v1, v2, o, d, u, v, t)) continue;
        if (COUNT) st.candidates++;
        if (any) {
            best.slot = i;
            return true;
        }
        const float4* pn = S.tri_n + 3 * (size_t)i;
        V3 n0 = ld3(pn), n1 = ld3(pn + 1), n2 = ld3(pn + 2);
        V3 sn;
        V3 pos = er_hit_position(v0, v1, v2, n0, n1, n2, o, d, u, v, t, sn);
        float dist = length(pos - o);
        uint32_t rank = rank_base + (uint32_t)(i - from);
        // reference: replace iff !valid || dist < best (strict), first in visit order wins ties
        if (best.slot < 0 || dist < best.dist || (dist == best.dist && rank < best.rank)) {
            best.slot = i;
            best.u = u;
            best.v = v;
            best.t = t;
            best.dist = dist;
            best.rank = rank;
        }
    }
    return false;
}

#define ER_PRUNE_REL 1.0001f
#define ER_PRUNE_ABS 1e-6f
#define ER_PRUNE_ORIGIN 2e-5f  // x largest |origin coordinate|: the exact dist |o + d t - o| carries ~ulp(|o|) of noise

// any = true: stop at the first Möller–Trumbore-passing triangle of the reference visit set
// (occlusion query). ORDERED = false keeps the reference's left-first order and visits the
// reference's full visit set (no pruning) — used to cross-check the pruned traversal.
template <bool COUNT, bool ORDERED>
ER_D void er_traverse(const DScene& S, bool any, V3 o, V3 d, HitRec& best, TraceStats& st) {
    // BVH::intersect computes dirfrac per box; hoisted here
    V3 inv = v3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    uint32_t path = 0, todo = 0;
    int depth = 0;
    // prune margin: relative in the distance, plus a term that follows the magnitude of the origin (far from the
    // world origin a near-tie in another subtree differs by rounding of o + d t, not of t)
    const float pabs = ER_PRUNE_ORIGIN * fmaxf(fmaxf(fabsf(o.x), fabsf(o.y)), fabsf(o.z)) + ER_PRUNE_ABS;
    for (;;) {
        const int idx = (int)path + depth - __popc(path);
        const float4* rec = S.pairs + 4 * (size_t)idx;
        const float4 q0 = __ldg(rec), q1 = __ldg(rec + 1), q2 = __ldg(rec + 2), q3 = __ldg(rec + 3);
        float tminL, tminR;
        bool hitL = er_box(v3(q0.x, q0.y, q0.z), v3(q0.w, q1.x, q1.y), o, inv, tminL);
        bool hitR = er_box(v3(q1.z, q1.w, q2.x), v3(q2.y, q2.z, q2.w), o, inv, tminR);
        if (COUNT) st.box_tests += 2;
        const uint32_t b = 1u << (ER_DEPTH - 1 - depth);
        bool descended = false;
        if (depth < ER_DEPTH - 1) {
            bool goL = hitL && __float_as_int(q3.x) > 0;
            bool goR = hitR && __float_as_int(q3.y) > 0;
            if (ORDERED && best.slot >= 0) {
                // keep a node unless its entry distance is beyond best + slack (with margin)
                if (tminL > (best.dist + q3.z) * ER_PRUNE_REL + pabs) goL = false;
                if (tminR > (best.dist + q3.w) * ER_PRUNE_REL + pabs) goR = false;
            }
            if (goL || goR) {
                bool right_first = !goL || (ORDERED && goR && tminR < tminL);
                if (goL && goR) todo |= b;
                if (right_first) path |= b;
                depth++;
                descended = true;
            }
        } else {
            const int fromL = __float_as_int(q3.x), toL = __float_as_int(q3.y), toR = __float_as_int(q3.z);
            const uint32_t nR = (uint32_t)(toR - toL);
            bool doR = hitR && toR > toL;
            bool doL = hitL && toL > fromL;
            if (ORDERED && best.slot >= 0) {
                if (tminL > (best.dist + q3.w) * ER_PRUNE_REL + pabs) doL = false;
                if (tminR > (best.dist + q3.w) * ER_PRUNE_REL + pabs) doR = false;
            }
            // reference rank inside the pair: right leaf's slots first, then the left leaf's
            if (doR && er_leaf<COUNT>(S, any, toL, toR, (uint32_t)fromL, o, d, best, st)) return;
            if (doL && er_leaf<COUNT>(S, any, fromL, toL, (uint32_t)fromL + nR, o, d, best, st)) return;
        }
        if (!descended) {
            if (todo == 0) return;
            const uint32_t pb = todo & (0u - todo);  // deepest pending sibling
            path = (path & ~(pb - 1)) ^ pb;
            depth = ER_DEPTH - (__ffs((int)pb) - 1);
            todo ^= pb;
        }
    }
}

// The reference `Hit` for the winner (Tri.h:76-141 + BVH.cpp:115-116).
struct HitFull {
    V3 position, normal, tangent, bitangent, gnormal;
    float tu, tv;
    uint32_t objectID, material_id;
    int triIdx;
    bool valid;
};

ER_D void er_make_hit(const DScene& S, const HitRec& r, V3 o, V3 d, HitFull& h) {
    if (r.slot < 0) {
        h.valid = false;
        h.triIdx = -2;
        h.position = h.normal = h.tangent = h.bitangent = h.gnormal = v3s(0.f);
        h.tu = h.tv = 0.f;
        h.objectID = h.material_id = 0;
        return;
    }
    const float4* pv = S.tri_v + 3 * (size_t)r.slot;
    const float4* pn = S.tri_n + 3 * (size_t)r.slot;
    const float4* pa = S.tri_a + 5 * (size_t)r.slot;
    V3 v0 = ld3(pv), v1 = ld3(pv + 1), v2 = ld3(pv + 2);
    V3 n0 = ld3(pn), n1 = ld3(pn + 1), n2 = ld3(pn + 2);
    float4 a0 = __ldg(pa), a1 = __ldg(pa + 1), a2 = __ldg(pa + 2), a3 = __ldg(pa + 3), a4 = __ldg(pa + 4);
    const float u = r.u, v = r.v;
    // tUV = uv0 + (uv1-uv0)*u + (uv2-uv0)*v
    h.tu = a0.x + (a0.z - a0.x) * u + (a1.x - a0.x) * v;
    h.tv = a0.y + (a0.w - a0.y) * u + (a1.y - a0.y) * v;
    V3 sn;
    h.position = er_hit_position(v0, v1, v2, n0, n1, n2, o, d, u, v, r.t, sn);
    V3 edge1 = v1 - v0, edge2 = v2 - v0;
    V3 compNormal = normalized(cross(edge1, edge2));
    if (dot(compNormal, d) > 0) compNormal = compNormal * -1.0f;
    V3 t0 = v3(a2.x, a2.y, a2.z), t1 = v3(a3.x, a3.y, a3.z), t2 = v3(a4.x, a4.y, a4.z);
    h.tangent = t0 + (t1 - t0) * u + (t2 - t0) * v;
    h.normal = sn;
    h.gnormal = compNormal;
    h.bitangent = a1.z * cross(h.normal, h.tangent);
    h.valid = true;
    h.objectID = (uint32_t)__float_as_int(a1.w);
    h.material_id = (uint32_t)__float_as_int(a2.w);
    h.triIdx = r.slot;
}

// throwRay (src/kernel.cpp:218-240): ray direction must already be normalised (Ray ctor).
template <bool COUNT, bool ORDERED>
ER_D void er_nearest(const DScene& S, V3 o, V3 d, HitRec& best, TraceStats& st) {
    best.slot = -1;
    best.dist = 0.f;
    best.rank = 0;
    best.u = best.v = best.t = 0.f;
    er_traverse<COUNT, ORDERED>(S, false, o, d, best, st);
}

// Ray ctor (Ray.h:13-17): normalise unless the length is zero
ER_D V3 er_ray_dir(V3 d) { return normalized(d); }
