// er_wavefront.cuh — k_render_wavefront: the persistent render kernel (takes the place of the
// per-sample parallel_for of renderingKernel, src/kernel.cpp:689-700, body :477-646).
//
// Design (B200: 148 SMs, 228 KB shared memory / SM, 32-wide warps):
//   * one persistent CTA per SM; the CTA owns WF_SLOTS path slots whose whole state lives in SHARED
//     memory (SoA, one 32-bit word array per field), so any lane can work on any path;
//   * a path alternates between three CTA-local ring queues:
//        QT  rays to trace (nearest-hit and shadow queries share the one traversal site)
//        QH  nearest hits to shade  (generateHitData + HDRI NEE + Disney sample/eval/pdf)
//        QE  everything cheap: shadow answers, misses (environment lookup), finished samples
//            (clamp + running mean into the passes), new camera rays, new pixel tickets
//     warps pick their role dynamically: a full batch of 32 hits or 32 epilogues is always executed
//     by 32 dense lanes (wavefront-style compaction by ballot + popc); otherwise the warp traces;
//   * tracing is a while-while loop: all lanes descend child-pair records together, lanes that reach
//     a leaf pair park until enough lanes wait, then the parked lanes run their Möller–Trumbore tests
//     together; a lane that finishes its ray is refilled from QT at once. Rays of a warp that is
//     needed for shading are suspended (7 words) back into QT and resumed by whichever warp is free;
//   * a pixel's samples run back to back in one slot (its xorshift stream is sequential across
//     samples, kernel.cpp:483,645); slots fetch new pixels through a warp-aggregated global ticket.
//
// Nearest-hit contract is that of er_trace.cuh: winner = min (dist, reference visit rank) over the
// Möller–Trumbore-passing triangles of the reference visit set. Here the exact `dist` of a candidate
// (Tri.h:76-117 shading position, 3 normal fetches, 2 sqrt) is evaluated only when two candidates are
// closer to each other than the interval bounds  t -/+ (1e-5 (t + |o|max) + slack(leaf pair))  can
// separate; a sole surviving candidate wins without ever being evaluated.
#pragma once
#include "er_render.cuh"

// One CTA of 20 warps per SM owning 1024 path slots (92 KB of shared memory; the rest of the SM's 256 KB
// stays L1, which the BVH records of the upper tree levels live in). Measured on B200, C2
// (profiles/README.md): 2 CTAs x 384 threads x 512 slots reaches 641 Msamples/s at its best batch
// threshold, 768 x 1024 slots 718, 640 x 1024 761; 2048 slots (L1 cut to 28 KB) 660.
#ifndef WF_NO_LEAN
#define WF_LEAN 1  // round 2: the lean traversal step (st_step / st_leaf_open); -DWF_NO_LEAN = round 1's wf_descend
#endif
#ifndef WF_THREADS
#define WF_THREADS 640
#endif
#ifndef WF_SLOTS
#define WF_SLOTS 1024
#endif
#ifndef WF_MIN_CTAS
#define WF_MIN_CTAS 1
#endif
// ring size of the queues: the power of two >= WF_SLOTS
#define WF_RING (WF_SLOTS <= 512 ? 512 : WF_SLOTS <= 1024 ? 1024 : 2048)
#define WF_MASK (WF_RING - 1)
static_assert(WF_SLOTS <= WF_RING && (WF_RING & WF_MASK) == 0, "the ring holds every slot and its size is a power of two");
#ifndef WF_SMEM_PAD
#define WF_SMEM_PAD 0  // extra dynamic shared memory (experiments on the L1 / shared memory split)
#endif
#define WF_FULLMASK 0xffffffffu

// per-slot fields (32-bit words), s_state[field * WF_SLOTS + slot]
enum {
    F_OX = 0, F_OY, F_OZ, F_DX, F_DY, F_DZ,        // ray to trace (direction normalised)
    F_LAX, F_LAY, F_LAZ,                           // light so far; while a shadow ray is out: light if it is NOT blocked
    F_RX, F_RY, F_RZ,                              // reduction (already multiplied by the bounce factor while a shadow ray is out)
    F_RNG, F_IDX, F_META, F_RES,                   // rng state, pixel index, packed meta, traversal answer
    F_TPATH, F_TTODO, F_TDEPTH, F_THI, F_TSLOT, F_TLO, F_TDIST,  // suspended traversal
    F_COUNT
};
// The records of the top WF_TOP_LEVELS levels of the tree (depth 0 .. WF_TOP_LEVELS-1: every ray walks them) are copied
// into shared memory when the kernel starts, in heap order (depth d, decisions p: (1 << d) - 1 + p). The lanes of a
// warp sit at unrelated records, so a record load from global memory costs the L1 one wavefront per lane and 32-byte
// half (the L1/TEX data path is the busiest unit of the traversal, 72 % in the round-2 profile); the same 64 bytes from
// shared memory cost a quarter of that. 9 levels = 511 records = 32 KB. 0 = off.
// MEASURED (round 2, profiles/README.md): 7 / 8 / 9 / 10 levels are all 1-3 % SLOWER than none on C2 (647-651 vs 668
// Msamples/s) and C3 (213-216 vs 219): the top levels are L1 hits already and the shared memory they take is L1 lost
// to the deeper ones. Off by default.
#ifndef WF_TOP_LEVELS
#define WF_TOP_LEVELS 0
#endif
#define WF_TOP_RECORDS ((1 << WF_TOP_LEVELS) - 1)
#define WF_TOP_BYTES (WF_TOP_LEVELS > 0 ? WF_TOP_RECORDS * 64 : 0)
#define WF_SMEM_BYTES (F_COUNT * WF_SLOTS * 4 + WF_TOP_BYTES + WF_SMEM_PAD)

// What a slot needs only across a shadow query or at the end of a sample is parked in a per-slot record
// in GLOBAL memory (L2-resident, 96 B, accessed with 128-bit loads/stores by dense batches) so that the
// shared memory of an SM holds twice as many slots:
//   r0 = (LB.xyz, NO.x)  r1 = (NO.yz, ND.xy)  r2 = (ND.z, OWN, SA, PG)     LB: light if the shadow ray IS
//   r3 = (N.xyz, T.x)    r4 = (T.yz, B.xy)    r5 = (B.z, -, -, -)          blocked; NO/ND: the continued ray;
// OWN: shaded triangle; SA: the pixel's sample counter; N/T/B: first-hit AOVs (valid iff WF_AOV in meta).
#define WF_SCRATCH_F4 6
#define WF_SCRATCH_BYTES_PER_CTA ((size_t)WF_SLOTS * WF_SCRATCH_F4 * 16)

// meta: bits 0-15 samples left for this pixel in this launch, 16-23 bounce, 24-26 phase, 27 any-hit, 28 suspended
enum { WPH_FETCH = 0, WPH_NEAREST = 1, WPH_SHADOW = 2, WPH_COMPLETE = 3, WPH_DEAD = 4, WPH_RETRY = 5 };
#define WF_META(s_left, bounce, phase) ((uint32_t)(s_left) | ((uint32_t)(bounce) << 16) | ((uint32_t)(phase) << 24))
#define WF_ANY (1u << 27)
#define WF_SUSP (1u << 28)
#define WF_AOV (1u << 29)

// QW: slots that hold a ticket for a pixel that is not ready yet (CHUNKED launches with fewer pixels than
// slots). They are looked at again only now and then (polling them like QE would starve the very rays
// they wait for).
enum { WQ_T = 0, WQ_H = 1, WQ_E = 2, WQ_W = 3 };

// Ring of WF_RING 32-bit entries. Entry of queue index i (position i & WF_MASK, lap i / WF_RING):
//   (lap + 1) << WF_QBITS | slot       written by the producer that reserved index i
//   (lap + 1) << WF_QBITS | WF_QSLOT   written back by the consumer that claimed index i
// A producer waits for the previous lap's consumed mark (0 on lap 0), a consumer for its own lap's
// entry, so neither can act on a value that belongs to another lap of the ring.
#define WF_QBITS 12
#define WF_QSLOT ((1u << WF_QBITS) - 1u)
struct WfQueue {
    unsigned head, tail;
    unsigned buf[WF_RING];
};
static_assert(WF_SLOTS < WF_QSLOT, "slot ids must fit below the consumed mark");

#define WF_SPIN_LIMIT (1u << 22)
__device__ int* g_wf_watchdog = nullptr;  // host-mapped watchdog record (set by the launch code)

// -DWF_CHECK: bounds checks on every index the scheduler hands around; a violation is written to the
// watchdog record ([2] code, [3] value, [4] CTA) and trips the watchdog instead of faulting
#ifdef WF_CHECK
#define WF_ASSERT(cond, code, val)                                           \
    do {                                                                     \
        if (!(cond) && g_wf_watchdog) {                                      \
            if (atomicCAS_system(g_wf_watchdog + 2, 0, (code)) == 0) {       \
                g_wf_watchdog[3] = (int)(val);                               \
                g_wf_watchdog[4] = (int)blockIdx.x;                          \
            }                                                                \
            atomicAdd_system(g_wf_watchdog, 1);                              \
        }                                                                    \
    } while (0)
#else
#define WF_ASSERT(cond, code, val) \
    do {                           \
    } while (0)
#endif

ER_D unsigned wf_lane() { return threadIdx.x & 31u; }
ER_D unsigned wf_vload(const unsigned* p) { return *(const volatile unsigned*)p; }
ER_D void wf_spin_fail(int what, unsigned idx, const WfQueue& q) {
    if (!g_wf_watchdog) return;
    if (atomicCAS_system(g_wf_watchdog + 5, 0, what) == 0) {
        g_wf_watchdog[6] = (int)idx;
        g_wf_watchdog[7] = (int)wf_vload(&q.head);
        g_wf_watchdog[8] = (int)wf_vload(&q.tail);
        g_wf_watchdog[9] = (int)wf_vload(&q.buf[idx & WF_MASK]);
        g_wf_watchdog[10] = (int)blockIdx.x;
        g_wf_watchdog[11] = (int)threadIdx.x;
    }
    atomicAdd_system(g_wf_watchdog, 0x10000);
}

// All 32 lanes call; lanes with pred append `slot`. State written before is made visible first.
ER_D void wf_push(WfQueue& q, bool pred, int slot) {
    const unsigned m = __ballot_sync(WF_FULLMASK, pred);
    if (!m) return;
    __threadfence_block();
    const int leader = __ffs((int)m) - 1;
    unsigned base = 0;
    if ((int)wf_lane() == leader) base = atomicAdd(&q.tail, (unsigned)__popc(m));
    base = __shfl_sync(WF_FULLMASK, base, leader);
    if (pred) {
        const unsigned idx = base + (unsigned)__popc(m & ((1u << wf_lane()) - 1u));
        const unsigned lap = idx / WF_RING;
        const unsigned expect = lap ? ((lap << WF_QBITS) | WF_QSLOT) : 0u;
        volatile unsigned* e = q.buf + (idx & WF_MASK);
        for (unsigned spin = 0; *e != expect; spin++)
            if (spin > WF_SPIN_LIMIT) {  // cannot happen: at most WF_SLOTS entries are ever outstanding
                wf_spin_fail(1, idx, q);
                break;
            }
        *e = ((lap + 1u) << WF_QBITS) | (unsigned)slot;
    }
}

// All 32 lanes call; returns how many entries the warp claimed (warp-uniform); lane i < n gets its slot.
ER_D int wf_claim(WfQueue& q, int want, int& slot) {
    unsigned h = 0;
    int n = 0;
    if (wf_lane() == 0) {
        for (;;) {
            h = wf_vload(&q.head);
            const unsigned t = wf_vload(&q.tail);
            n = (int)(t - h);
            if (n > want) n = want;
            if (n <= 0) {
                n = 0;
                break;
            }
            if (atomicCAS(&q.head, h, h + (unsigned)n) == h) break;
        }
    }
    n = __shfl_sync(WF_FULLMASK, n, 0);
    h = __shfl_sync(WF_FULLMASK, h, 0);
    slot = -1;
    if ((int)wf_lane() < n) {
        const unsigned idx = h + wf_lane();
        const unsigned tag = (idx / WF_RING + 1u) << WF_QBITS;
        volatile unsigned* e = q.buf + (idx & WF_MASK);
        unsigned v;
        bool stuck = false;
        for (unsigned spin = 0; ((v = *e) & ~WF_QSLOT) != tag || (v & WF_QSLOT) == WF_QSLOT; spin++)
            if (spin > WF_SPIN_LIMIT) {
                wf_spin_fail(2, idx, q);
                stuck = true;
                break;
            }
        if (!stuck) {  // a stuck lane (cannot happen) stays empty-handed and the launch is reported as failed
            *e = tag | WF_QSLOT;
            slot = (int)(v & WF_QSLOT);
        }
        WF_ASSERT(stuck || (slot >= 0 && slot < WF_SLOTS), 1, slot);
    }
    if (n) __threadfence_block();
    return n;
}

ER_D int wf_size(const WfQueue& q) { return (int)(wf_vload(&q.tail) - wf_vload(&q.head)); }

// ---------------------------------------------------------------------------------------------
// per-lane traversal state (registers)
// ---------------------------------------------------------------------------------------------
enum { T_IDLE = 0, T_DESCEND = 1, T_LEAF = 2, T_DONE = 3 };

struct Trav {
    int state;
    int slot;  // path slot this lane is tracing for
    V3 o, d, inv;
    float omax;
    uint32_t path, todo;
    int depth;
    bool any, slow, shadow;
    float best_hi;  // upper bound of the winner's dist (prunes nodes, culls candidates)
    int st_slot;    // surviving candidate (BVH slot) or -1
    float st_lo;    // lower bound of its dist while it has not been evaluated
    float st_dist;  // its exact dist once evaluated, else -1
    int result;     // any-hit answer
    // leaf pair being tested: reference ranks [lf_cur, lf_end) of the pair; ranks below lf_mid are the
    // RIGHT leaf's slots (tested first, BVH.cpp:81-87), the rest the left leaf's
    int lf_cur, lf_end, lf_mid, lf_nL, lf_nR;
    float lf_slack;
    unsigned work;  // COUNT build: traversal work of this ray (er_row_cost), in units of about one warp instruction
};

// BVH::intersect for rays without a zero/denormal direction component: no NaN can arise, so the
// NaN-order-sensitive a<b?a:b selects (Math.hpp:61-67) are the plain min/max (one FMNMX each); the
// sign of a zero result never reaches a decision (tmax < 0, tmin > tmax).
ER_D bool er_box_fast(V3 b1, V3 b2, V3 o, V3 inv, float& tmin_out) {
    float t1 = (b1.x - o.x) * inv.x;
    float t2 = (b2.x - o.x) * inv.x;
    float t3 = (b1.y - o.y) * inv.y;
    float t4 = (b2.y - o.y) * inv.y;
    float t5 = (b1.z - o.z) * inv.z;
    float t6 = (b2.z - o.z) * inv.z;
    float tmin = fmaxf(fmaxf(fminf(t1, t2), fminf(t3, t4)), fminf(t5, t6));
    float tmax = fminf(fminf(fmaxf(t1, t2), fmaxf(t3, t4)), fmaxf(t5, t6));
    tmin_out = tmin;
    return !(tmax < 0) && !(tmin > tmax);
}

// 32 bytes from a 32-byte-aligned read-only address in one instruction (sm_100: ld.global.nc.v8)
ER_D void wf_ldg256(const float4* p, float4& a, float4& b) {
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
                 : "l"(p));
}

ER_D void wf_pop(Trav& t) {
    if (t.todo == 0) {
        t.state = T_DONE;
        return;
    }
    const uint32_t pb = t.todo & (0u - t.todo);  // deepest pending sibling
    t.path = (t.path & ~(pb - 1)) ^ pb;
    t.depth = ER_DEPTH - (__ffs((int)pb) - 1);
    t.todo ^= pb;
    t.state = T_DESCEND;
}

// one child-pair record (BVH::transverse step, BVH.cpp:69-101, both child boxes tested)
template <bool COUNT>
ER_D void wf_descend(const DScene& S, Trav& t, TraceStats& st, const int pf_min) {
    int idx = (int)t.path + t.depth - __popc(t.path);
    WF_ASSERT(idx >= 0 && idx < ER_INTERNAL_COUNT && t.depth >= 0 && t.depth < ER_DEPTH, 3, (t.depth << 24) ^ idx);
#ifdef WF_CHECK
    if (!(idx >= 0 && idx < ER_INTERNAL_COUNT)) idx = 0;
#endif
    const float4* rec = S.pairs + 4 * (size_t)idx;
#ifdef WF_PREFETCH
    // the children's records are at implicit addresses: start fetching both into L1 while this one is in flight
    if (t.depth < ER_DEPTH - 1) {
        asm volatile("prefetch.global.L1 [%0];" ::"l"(rec + 4));
        asm volatile("prefetch.global.L1 [%0];" ::"l"(rec + 4 * ((size_t)1 << (ER_DEPTH - 1 - t.depth))));
    }
#endif
    // a 64-B record = two 256-bit loads (LDG.E.256): half the L1 tag lookups of four LDG.128 — the lanes
    // of a warp are at unrelated records, so every load instruction costs one lookup per active lane
    float4 q0, q1, q2, q3;
    wf_ldg256(rec, q0, q1);
    wf_ldg256(rec + 2, q2, q3);
    float tminL, tminR;
    bool hitL, hitR;
    if (t.slow) {
        hitL = er_box(v3(q0.x, q0.y, q0.z), v3(q0.w, q1.x, q1.y), t.o, t.inv, tminL);
        hitR = er_box(v3(q1.z, q1.w, q2.x), v3(q2.y, q2.z, q2.w), t.o, t.inv, tminR);
    } else {
        hitL = er_box_fast(v3(q0.x, q0.y, q0.z), v3(q0.w, q1.x, q1.y), t.o, t.inv, tminL);
        hitR = er_box_fast(v3(q1.z, q1.w, q2.x), v3(q2.y, q2.z, q2.w), t.o, t.inv, tminR);
    }
    if (COUNT) st.box_tests += 2;
    if (COUNT) t.work += 4;
    if (t.depth < ER_DEPTH - 1) {
        bool goL = hitL && __float_as_int(q3.x) > 0;
        bool goR = hitR && __float_as_int(q3.y) > 0;
        if (tminL > (t.best_hi + q3.z) * ER_PRUNE_REL + ER_PRUNE_ABS) goL = false;
        if (tminR > (t.best_hi + q3.w) * ER_PRUNE_REL + ER_PRUNE_ABS) goR = false;
        if (goL || goR) {
            const uint32_t b = 1u << (ER_DEPTH - 1 - t.depth);
            const bool right_first = !goL || (goR && tminR < tminL);
            if (goL && goR) t.todo |= b;
            if (right_first) t.path |= b;
            t.depth++;
            return;
        }
    } else {
        const int fromL = __float_as_int(q3.x), toL = __float_as_int(q3.y), toR = __float_as_int(q3.z);
        bool doR = hitR && toR > toL;
        bool doL = hitL && toL > fromL;
        if (tminL > (t.best_hi + q3.w) * ER_PRUNE_REL + ER_PRUNE_ABS) doL = false;
        if (tminR > (t.best_hi + q3.w) * ER_PRUNE_REL + ER_PRUNE_ABS) doR = false;
        if (doL || doR) {
            t.lf_nL = toL - fromL;
            t.lf_nR = toR - toL;
            t.lf_mid = fromL + t.lf_nR;
            t.lf_cur = doR ? fromL : t.lf_mid;
            t.lf_end = doL ? toR : t.lf_mid;
            t.lf_slack = q3.w;
            t.state = T_LEAF;
            // A big leaf pair is tested later, by the whole warp, once few lanes still descend: ask L2 for its
            // triangles now (one bulk prefetch of the contiguous slot range, no register or shared memory
            // cost), so that the coalesced loads of wf_leaf_coop find them there instead of in HBM
            if (pf_min > 0 && t.lf_end - t.lf_cur >= pf_min) {
                const int s0 = doL ? fromL : toL, s1 = doR ? toR : toL;
                asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(S.tri_v + 3 * (size_t)s0), "r"(48 * (s1 - s0)) : "memory");
            }
            return;
        }
    }
    wf_pop(t);
}

// exact |hit.position - origin| of BVH slot `slot` for ray (o,d): Tri.h:43-117 (the slot passed MT)
__device__ __noinline__ float wf_exact_dist(const float4* __restrict__ tri_v, const float4* __restrict__ tri_n, int slot,
                                            float ox, float oy, float oz, float dx, float dy, float dz) {
    const V3 o = v3(ox, oy, oz), d = v3(dx, dy, dz);
    const float4* pv = tri_v + 3 * (size_t)slot;
    const float4* pn = tri_n + 3 * (size_t)slot;
    const V3 v0 = ld3(pv), v1 = ld3(pv + 1), v2 = ld3(pv + 2);
    float u = 0.f, v = 0.f, tt = 0.f;
    er_mt(v0, v1, v2, o, d, u, v, tt);
    V3 sn;
    const V3 pos = er_hit_position(v0, v1, v2, ld3(pn), ld3(pn + 1), ld3(pn + 2), o, d, u, v, tt, sn);
    return length(pos - o);
}

// a Möller–Trumbore-passing triangle at parameter tt
ER_D void wf_candidate(const DScene& S, Trav& t, int slot, float tt) {
    if (t.any) {
        t.result = slot;
        t.state = T_DONE;
        return;
    }
    const float m = 1e-5f * (tt + t.omax);
    const float lo = tt - m - t.lf_slack, hi = tt + m + t.lf_slack;
    if (lo > t.best_hi) return;  // cannot beat the survivor
    if (t.st_slot >= 0) {
        const float s_lo = t.st_dist >= 0.f ? t.st_dist : t.st_lo;
        if (!(s_lo > hi)) {
            // the two cannot be separated by bounds: compare (dist, reference rank) exactly
            if (!(t.st_dist >= 0.f))
                t.st_dist = wf_exact_dist(S.tri_v, S.tri_n, t.st_slot, t.o.x, t.o.y, t.o.z, t.d.x, t.d.y, t.d.z);
            const float nd = wf_exact_dist(S.tri_v, S.tri_n, slot, t.o.x, t.o.y, t.o.z, t.d.x, t.d.y, t.d.z);
            // rank order: inside one leaf pair = test order (the survivor came first); across pairs = slot order
            const bool same_pair = t.st_slot >= t.lf_mid - t.lf_nR && t.st_slot < t.lf_mid + t.lf_nL;
            if (nd < t.st_dist || (nd == t.st_dist && !same_pair && slot < t.st_slot)) {
                t.st_slot = slot;
                t.st_dist = nd;
            }
            t.best_hi = fminf(t.best_hi, t.st_dist);
            return;
        }
    }
    t.st_slot = slot;
    t.st_lo = lo;
    t.st_dist = -1.f;
    t.best_hi = fminf(t.best_hi, hi);
}

// one triangle of the parked leaf pair (BVH::intersectNode body, BVH.cpp:109-118)
template <bool COUNT>
ER_D void wf_tri(const DScene& S, Trav& t, TraceStats& st) {
    const int r = t.lf_cur++;
    int slot = r < t.lf_mid ? r + t.lf_nL : r - t.lf_nR;
    WF_ASSERT(slot >= 0 && slot < S.tri_count, 4, slot);
#ifdef WF_CHECK
    if (!(slot >= 0 && slot < S.tri_count)) slot = 0;
#endif
    const float4* pv = S.tri_v + 3 * (size_t)slot;
    float u, v, tt;
    if (COUNT) st.tri_tests++;
    if (COUNT) t.work += 8;
    if (er_mt(ld3(pv), ld3(pv + 1), ld3(pv + 2), t.o, t.d, u, v, tt)) {
        if (COUNT) st.candidates++;
        wf_candidate(S, t, slot, tt);
    }
    if (t.state == T_LEAF && t.lf_cur >= t.lf_end) wf_pop(t);
}

// The parked leaf pair of lane `src`, tested by all 32 lanes (BVH::intersectNode, BVH.cpp:109-118, for
// both leaves of the pair). Candidates reach the owner in the reference's test order, so wf_candidate
// sees exactly the sequence the per-lane loop would give it.
template <bool COUNT>
ER_D void wf_leaf_coop(const DScene& S, Trav& t, int src, TraceStats& st) {
    const V3 o = v3(__shfl_sync(WF_FULLMASK, t.o.x, src), __shfl_sync(WF_FULLMASK, t.o.y, src),
                    __shfl_sync(WF_FULLMASK, t.o.z, src));
    const V3 d = v3(__shfl_sync(WF_FULLMASK, t.d.x, src), __shfl_sync(WF_FULLMASK, t.d.y, src),
                    __shfl_sync(WF_FULLMASK, t.d.z, src));
    const int cur = __shfl_sync(WF_FULLMASK, t.lf_cur, src), end = __shfl_sync(WF_FULLMASK, t.lf_end, src);
    const int mid = __shfl_sync(WF_FULLMASK, t.lf_mid, src);
    const int nL = __shfl_sync(WF_FULLMASK, t.lf_nL, src), nR = __shfl_sync(WF_FULLMASK, t.lf_nR, src);
    const bool owner = (int)wf_lane() == src;
#ifdef WF_COOP_PIPE
    // Software pipeline (round 2, -DWF_COOP_PIPE): the vertices of the NEXT 32 triangles are in flight while these 32
    // are tested (the loop body ends in votes and a conditional break, which keeps the compiler from hoisting the
    // loads itself; the long-scoreboard wait on these loads was 56-67 % of the stall samples of the leaf test on C3).
    // MEASURED SLOWER (profiles/r04c_pipe.txt): C3 219.8 -> 209.4, C2 674.6 -> 650.1 Msamples/s. The kernel sits at its
    // register cap (96 at 640 threads); the 12 extra live registers turn into spills (172 -> 200 bytes) on paths C2
    // runs too. Off by default.
    float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0, a2 = a0;
    {
        const int r = cur + (int)wf_lane();
        if (r < end) {
            int slot = r < mid ? r + nL : r - nR;
#ifdef WF_CHECK
            if (!(slot >= 0 && slot < S.tri_count)) slot = 0;
#endif
            const float4* pv = S.tri_v + 3 * (size_t)slot;
            a0 = __ldg(pv), a1 = __ldg(pv + 1), a2 = __ldg(pv + 2);
        }
    }
    for (int base = cur; base < end; base += 32) {
        float4 b0 = a0, b1 = a1, b2 = a2;
        {
            const int rn = base + 32 + (int)wf_lane();
            if (rn < end) {
                int sn = rn < mid ? rn + nL : rn - nR;
#ifdef WF_CHECK
                if (!(sn >= 0 && sn < S.tri_count)) sn = 0;
#endif
                const float4* pn = S.tri_v + 3 * (size_t)sn;
                b0 = __ldg(pn), b1 = __ldg(pn + 1), b2 = __ldg(pn + 2);
            }
        }
        const int r = base + (int)wf_lane();
        int slot = -1;
        float u, v, tt = 0.f;
        bool pass = false;
        if (r < end) {
            slot = r < mid ? r + nL : r - nR;
            WF_ASSERT(slot >= 0 && slot < S.tri_count, 4, slot);
            if (COUNT) st.tri_tests++;
            pass = er_mt(v3(a0.x, a0.y, a0.z), v3(a1.x, a1.y, a1.z), v3(a2.x, a2.y, a2.z), o, d, u, v, tt);
            if (COUNT && pass) st.candidates++;
        }
        unsigned cm = __ballot_sync(WF_FULLMASK, pass);
        while (cm) {
            const int j = __ffs((int)cm) - 1;
            cm &= cm - 1;
            const int cs = __shfl_sync(WF_FULLMASK, slot, j);
            const float ct = __shfl_sync(WF_FULLMASK, tt, j);
            if (owner && t.state == T_LEAF) wf_candidate(S, t, cs, ct);
        }
        if (__shfl_sync(WF_FULLMASK, t.state, src) != T_LEAF) break;  // an any-hit query is answered
        a0 = b0, a1 = b1, a2 = b2;
    }
#else
    for (int base = cur; base < end; base += 32) {
        const int r = base + (int)wf_lane();
        int slot = -1;
        float u, v, tt = 0.f;
        bool pass = false;
        if (r < end) {
            slot = r < mid ? r + nL : r - nR;
            WF_ASSERT(slot >= 0 && slot < S.tri_count, 4, slot);
#ifdef WF_CHECK
            if (!(slot >= 0 && slot < S.tri_count)) slot = 0;
#endif
            const float4* pv = S.tri_v + 3 * (size_t)slot;
            if (COUNT) st.tri_tests++;
            pass = er_mt(ld3(pv), ld3(pv + 1), ld3(pv + 2), o, d, u, v, tt);
            if (COUNT && pass) st.candidates++;
        }
        unsigned cm = __ballot_sync(WF_FULLMASK, pass);
        while (cm) {
            const int j = __ffs((int)cm) - 1;
            cm &= cm - 1;
            const int cs = __shfl_sync(WF_FULLMASK, slot, j);
            const float ct = __shfl_sync(WF_FULLMASK, tt, j);
            if (owner && t.state == T_LEAF) wf_candidate(S, t, cs, ct);
        }
        if (__shfl_sync(WF_FULLMASK, t.state, src) != T_LEAF) break;  // an any-hit query is answered
    }
#endif
    if (COUNT && owner) t.work += 3u * (unsigned)(end - cur);
    if (owner && t.state == T_LEAF) {
        t.lf_cur = t.lf_end;
        wf_pop(t);
    }
}

// ---------------------------------------------------------------------------------------------
// Warp-cooperative leaf test with group boxes (round 2). The reference tests EVERY triangle of a visited leaf
// (BVH::intersectNode, BVH.cpp:109-118); with ~38 triangles per leaf (10 M-triangle scenes) almost all of those tests
// fail. DScene::tri_g holds one conservative box per group of up to 4 triangles of a leaf, grouped by the Morton order
// of their centroids (er_host.cpp; leaf `ord` owns the groups from (from >> 2) + ord): the warp first tests the
// boxes of the whole leaf pair (one lane per group: up to 128 triangles in one step), then runs Moeller-Trumbore only
// on the triangles of the groups the ray meets (4 lanes per group, up to 8 groups in one step), and hands the
// passing ones to the owning lane in the reference's test order (rank: the right leaf's slots, then the left leaf's),
// so wf_candidate sees the sequence the full loop would give it minus triangles that cannot pass. A box never
// rejects a ray for which Tri::hit passes: it is the exact box widened by hundreds of ulps of the coordinates
// (host) and of the ray origin (here). More than 32 groups or more than 8 groups hit: the full loop (wf_leaf_coop).
// MEASURED (C3, 8 spp): triangle tests per ray 253 -> 112, but 219 -> 149 Msamples/s: the box pass, the second
// dependent load and the hand-over cost more than the skipped Moeller-Trumbore tests saved. OFF by default: compiled
// only with -DWF_GROUP_BOXES, and the host builds the boxes only when ER_GROUP_BOXES=1 is in the environment.
// ---------------------------------------------------------------------------------------------

ER_D bool wf_group_box(float4 g0, float4 g1, V3 o, V3 inv, float omax) {
    const float e = 4e-5f * (omax + g1.z);  // g1.z = largest |coordinate| of the group
    const float t1 = ((g0.x - e) - o.x) * inv.x, t2 = ((g0.w + e) - o.x) * inv.x;
    const float t3 = ((g0.y - e) - o.y) * inv.y, t4 = ((g1.x + e) - o.y) * inv.y;
    const float t5 = ((g0.z - e) - o.z) * inv.z, t6 = ((g1.y + e) - o.z) * inv.z;
    const float tmin = fmaxf(fmaxf(fminf(t1, t2), fminf(t3, t4)), fminf(t5, t6));
    const float tmax = fminf(fminf(fmaxf(t1, t2), fmaxf(t3, t4)), fmaxf(t5, t6));
    // widened once more in t: the products above round too
    return !(tmax < -1e-6f - 1e-5f * fabsf(tmax)) && !(tmin > tmax + 1e-6f + 1e-5f * (fabsf(tmin) + fabsf(tmax)));
}

template <bool COUNT>
ER_D void wf_leaf_coop_g(const DScene& S, Trav& t, int src, TraceStats& st) {
    const int lane = (int)wf_lane();
    const int cur = __shfl_sync(WF_FULLMASK, t.lf_cur, src), end = __shfl_sync(WF_FULLMASK, t.lf_end, src);
    const int mid = __shfl_sync(WF_FULLMASK, t.lf_mid, src);
    const int nL = __shfl_sync(WF_FULLMASK, t.lf_nL, src), nR = __shfl_sync(WF_FULLMASK, t.lf_nR, src);
    const unsigned pair = __shfl_sync(WF_FULLMASK, t.path, src) >> 1;  // ordinal of the leaf pair: leaves 2 pair, 2 pair + 1
    // ranks [cur, end): those below mid are the RIGHT leaf's slots (slot = rank + nL), the rest the left leaf's (rank - nR).
    // A pair comes here whole (its first vote): cur is the start of an enabled leaf, end the end of one.
    const int fromL = mid - nR, toL = fromL + nL;
    const bool hasR = cur < mid && nR > 0, hasL = end > mid && nL > 0;
    const int gR = hasR ? (nR + 3) >> 2 : 0, gL = hasL ? (nL + 3) >> 2 : 0;  // groups: the right leaf's first (rank order)
    if (gR + gL > 32 || nR > 255 || nL > 255 || cur != (hasR ? fromL : mid)) {
        wf_leaf_coop<COUNT>(S, t, src, st);
        return;
    }
    const V3 o = v3(__shfl_sync(WF_FULLMASK, t.o.x, src), __shfl_sync(WF_FULLMASK, t.o.y, src), __shfl_sync(WF_FULLMASK, t.o.z, src));
    const V3 inv = v3(__shfl_sync(WF_FULLMASK, t.inv.x, src), __shfl_sync(WF_FULLMASK, t.inv.y, src),
                      __shfl_sync(WF_FULLMASK, t.inv.z, src));
    const float omax = __shfl_sync(WF_FULLMASK, t.omax, src);
    bool hitg = false;
    unsigned offs = 0xffffffffu;
    int leaf_from = 0;
    bool right = false;
    if (lane < gR + gL) {
        right = lane < gR;
        leaf_from = right ? toL : fromL;
        const size_t g = right ? ((size_t)toL >> 2) + 2u * pair + 1u + (unsigned)lane
                               : ((size_t)fromL >> 2) + 2u * pair + (unsigned)(lane - gR);
        const float4* pg = S.tri_g + 2 * g;
        const float4 g0 = __ldg(pg), g1 = __ldg(pg + 1);
        hitg = wf_group_box(g0, g1, o, inv, omax);
        offs = __float_as_uint(g1.w);
        if (COUNT) st.box_tests++;
    }
    const unsigned gm = __ballot_sync(WF_FULLMASK, hitg);
    const int ng = __popc(gm);
    if (ng > 8) {  // (rare) the passing triangles of more than one step would reach the owner out of rank order
        wf_leaf_coop<COUNT>(S, t, src, st);
        return;
    }
    const bool owner = lane == src;
    if (ng > 0) {
        const V3 d = v3(__shfl_sync(WF_FULLMASK, t.d.x, src), __shfl_sync(WF_FULLMASK, t.d.y, src), __shfl_sync(WF_FULLMASK, t.d.z, src));
        // lane l tests triangle (l & 3) of the (l >> 2)-th group that was hit
        const int gi = (lane >> 2) < ng ? (int)__fns(gm, 0, (lane >> 2) + 1) : 0;
        const unsigned g_offs = __shfl_sync(WF_FULLMASK, offs, gi);
        const int g_from = __shfl_sync(WF_FULLMASK, leaf_from, gi);
        const bool g_right = __shfl_sync(WF_FULLMASK, (int)right, gi) != 0;
        int slot = -1, rank = 0x7fffffff;
        float tt = 0.f;
        bool pass = false;
        const unsigned off = (g_offs >> (8 * (lane & 3))) & 0xffu;
        if ((lane >> 2) < ng && off != 0xffu) {
            slot = g_from + (int)off;
            WF_ASSERT(slot >= 0 && slot < S.tri_count, 4, slot);
#ifdef WF_CHECK
            if (!(slot >= 0 && slot < S.tri_count)) slot = 0;
#endif
            const float4* pv = S.tri_v + 3 * (size_t)slot;
            float u, v;
            if (COUNT) st.tri_tests++;
            pass = er_mt(ld3(pv), ld3(pv + 1), ld3(pv + 2), o, d, u, v, tt);
            if (COUNT && pass) st.candidates++;
            rank = g_right ? slot - nL : slot + nR;
        }
        unsigned pm = __ballot_sync(WF_FULLMASK, pass);
        while (pm) {
            const int best = __reduce_min_sync(WF_FULLMASK, pass ? rank : 0x7fffffff);
            const int j = __ffs((int)__ballot_sync(WF_FULLMASK, pass && rank == best)) - 1;
            const int cs = __shfl_sync(WF_FULLMASK, slot, j);
            const float ct = __shfl_sync(WF_FULLMASK, tt, j);
            if (owner && t.state == T_LEAF) wf_candidate(S, t, cs, ct);
            if (lane == j) pass = false;
            pm &= ~(1u << j);
            if (__shfl_sync(WF_FULLMASK, t.state, src) != T_LEAF) break;  // an any-hit query is answered
        }
    }
    if (COUNT && owner) t.work += 3u + 3u * (unsigned)ng;
    if (owner && t.state == T_LEAF) {
        t.lf_cur = t.lf_end;
        wf_pop(t);
    }
}

// ---------------------------------------------------------------------------------------------
// The lean traversal step (round 2): what a lane does between two records, cut down to what the profile of
// round 1 showed it needs (profiles/README.md: 84 warp instructions of bookkeeping around 74 of box tests):
//   * st_step handles INTERIOR records only (depth 0..16); a lane that arrives at depth 17 waits for the
//     leaf phase, where st_leaf_open reads the leaf-pair record and the triangles are tested — no divergent
//     leaf branch inside the descend loop;
//   * rays with a zero or denormal direction component (BVH::intersect's NaN-ordered min/max matters for
//     them, BVH.cpp:27-61, Math.hpp:61-67) are rare: they are traced by the scalar er_traverse when they
//     are loaded (st_trace_slow), so the loop carries no `slow` variant of the box test;
//   * the prune threshold of the best hit so far (distance + margins) is one register, refreshed after
//     every leaf phase; per child it costs one FADD with the subtree's slack.
// Used by k_render_wavefront below and by k_st_trace (er_staged.cuh).
// ---------------------------------------------------------------------------------------------
#define ST_LEAF_DEPTH (ER_DEPTH - 1)  // records at this depth hold two leaves

ER_D void st_pop(Trav& t) {
    if (t.todo == 0) {
        t.state = T_DONE;
        return;
    }
    const uint32_t pb = t.todo & (0u - t.todo);  // deepest pending sibling
    t.path = (t.path & ~(pb - 1)) ^ pb;
    t.depth = ER_DEPTH - (__ffs((int)pb) - 1);
    t.todo ^= pb;
    t.state = T_DESCEND;
}

// the margin of distance pruning: a node whose entry distance exceeds best_hi by more than this (plus its
// subtree's slack) cannot hold the winner. Relative part for the rounding of tmin and dist at distance
// best_hi, origin-scaled part for the rounding of |position - origin| (ulp(|o|) each).
// (prune = best_hi * 1.0001 + (2e-5 |o|max + 1e-6), refreshed after every leaf phase)
ER_D float st_omax(V3 o) { return fmaxf(fmaxf(fabsf(o.x), fabsf(o.y)), fabsf(o.z)); }
#define ST_SLOT_MASK 0x0fffffff
#define ST_ANY_BIT 0x40000000
#define ST_SHADOW_BIT 0x20000000

// one interior child-pair record (BVH::transverse step, BVH.cpp:69-101, both child boxes tested)
template <bool COUNT>
ER_D void st_step(const DScene& S, Trav& t, const float prune, TraceStats& st, const float4* top = nullptr) {
    float4 q0, q1, q2, q3;
    if (WF_TOP_LEVELS > 0 && top != nullptr && t.depth < WF_TOP_LEVELS) {
        const float4* rec = top + 4 * (((1u << t.depth) - 1u) + (t.depth ? t.path >> (ER_DEPTH - t.depth) : 0u));
        q0 = rec[0], q1 = rec[1], q2 = rec[2], q3 = rec[3];
    } else {
        const int idx = (int)t.path + t.depth - __popc(t.path);
        const float4* rec = S.pairs + 4 * (size_t)idx;
        wf_ldg256(rec, q0, q1);
        wf_ldg256(rec + 2, q2, q3);
    }
    float tminL, tminR;
    bool goL = er_box_fast(v3(q0.x, q0.y, q0.z), v3(q0.w, q1.x, q1.y), t.o, t.inv, tminL);
    bool goR = er_box_fast(v3(q1.z, q1.w, q2.x), v3(q2.y, q2.z, q2.w), t.o, t.inv, tminR);
    if (COUNT) st.box_tests += 2;
    goL = goL && __float_as_int(q3.x) > 0 && !(tminL > prune + q3.z);
    goR = goR && __float_as_int(q3.y) > 0 && !(tminR > prune + q3.w);
    if (goL || goR) {
        const uint32_t b = 1u << (ER_DEPTH - 1 - t.depth);
        if (goL && goR) t.todo |= b;
        if (!goL || (goR && tminR < tminL)) t.path |= b;
        t.depth++;
        return;
    }
    st_pop(t);
}

// the leaf-pair record of a lane waiting at depth 17 (BVH.cpp:79-87: both leaf boxes tested, R leaf first)
template <bool COUNT>
ER_D void st_leaf_open(const DScene& S, Trav& t, const float prune, TraceStats& st) {
    const int idx = (int)t.path + t.depth - __popc(t.path);
    const float4* rec = S.pairs + 4 * (size_t)idx;
    float4 q0, q1, q2, q3;
    wf_ldg256(rec, q0, q1);
    wf_ldg256(rec + 2, q2, q3);
    float tminL, tminR;
    const bool hitL = er_box_fast(v3(q0.x, q0.y, q0.z), v3(q0.w, q1.x, q1.y), t.o, t.inv, tminL);
    const bool hitR = er_box_fast(v3(q1.z, q1.w, q2.x), v3(q2.y, q2.z, q2.w), t.o, t.inv, tminR);
    if (COUNT) st.box_tests += 2;
    const int fromL = __float_as_int(q3.x), toL = __float_as_int(q3.y), toR = __float_as_int(q3.z);
    const bool doR = hitR && toR > toL && !(tminR > prune + q3.w);
    const bool doL = hitL && toL > fromL && !(tminL > prune + q3.w);
    if (doL || doR) {
        t.lf_nL = toL - fromL;
        t.lf_nR = toR - toL;
        t.lf_mid = fromL + t.lf_nR;
        t.lf_cur = doR ? fromL : t.lf_mid;
        t.lf_end = doL ? toR : t.lf_mid;
        t.lf_slack = q3.w;
        t.state = T_LEAF;
        return;
    }
    st_pop(t);
}

// a ray with a zero / denormal direction component or a non-finite origin: the literal box test
// (er_box), one ray at a time, out of line so that it costs the hot loop no registers
template <bool COUNT>
__device__ __noinline__ int st_trace_slow(const DScene& S, bool any, float ox, float oy, float oz, float dx, float dy, float dz,
                                          TraceStats& st) {
    HitRec best;
    best.slot = -1;
    best.dist = 0.f;
    best.rank = 0;
    best.u = best.v = best.t = 0.f;
    er_traverse<COUNT, true>(S, any, v3(ox, oy, oz), v3(dx, dy, dz), best, st);
    return best.slot;
}

// ---------------------------------------------------------------------------------------------
// slot state access
// ---------------------------------------------------------------------------------------------
#ifdef WF_CHECK
ER_D int wf_index(int f, int s) {
    WF_ASSERT(s >= 0 && s < WF_SLOTS && f >= 0 && f < F_COUNT, 8, (f << 16) ^ (s & 0xffff));
    return (s >= 0 && s < WF_SLOTS) ? f * WF_SLOTS + s : f * WF_SLOTS;
}
#else
#define wf_index(f, s) ((f) * WF_SLOTS + (s))
#endif
#define WF_U(f, s) (st_[wf_index((f), (s))])
#define WF_F(f, s) (__uint_as_float(st_[wf_index((f), (s))]))
#define WF_SETF(f, s, v) (st_[wf_index((f), (s))] = __float_as_uint(v))
ER_D V3 wf_ld3(const uint32_t* st_, int f, int s) { return v3(WF_F(f, s), WF_F(f + 1, s), WF_F(f + 2, s)); }
ER_D void wf_st3(uint32_t* st_, int f, int s, V3 v) {
    WF_SETF(f, s, v.x);
    WF_SETF(f + 1, s, v.y);
    WF_SETF(f + 2, s, v.z);
}

// QH handler: kernel.cpp:524-587 for one nearest hit (dense: every calling lane with slot >= 0 has one).
// Returns the queue the slot goes to next (WQ_T or WQ_E).
template <bool COUNT>
ER_D int wf_shade(const DScene& S, const RenderParams& P, uint32_t* st_, float4* sc_, int slot, PathStats& ps) {
    const V3 ro = wf_ld3(st_, F_OX, slot), rd = wf_ld3(st_, F_DX, slot);
    uint32_t rng = WF_U(F_RNG, slot);
    const uint32_t meta = WF_U(F_META, slot);
    int bounce = (int)((meta >> 16) & 0xffu);
    const uint32_t s_left = meta & 0xffffu;
    HitRec rec;
    rec.slot = (int)WF_U(F_RES, slot);
    WF_ASSERT(rec.slot >= 0 && rec.slot < S.tri_count, 2, rec.slot);
    WF_ASSERT(((meta >> 24) & 7u) == WPH_NEAREST && !(meta & WF_SUSP), 6, meta);
#ifdef WF_CHECK
    if (!(rec.slot >= 0 && rec.slot < S.tri_count)) rec.slot = 0;
#endif
    rec.dist = 0.f;
    rec.rank = 0;
    {
        // u, v, t of the winner: the same Möller–Trumbore evaluation the traversal did (same code, same bits)
        const float4* pv = S.tri_v + 3 * (size_t)rec.slot;
        er_mt(ld3(pv), ld3(pv + 1), ld3(pv + 2), ro, rd, rec.u, rec.v, rec.t);
    }
    Bounce b;
    er_shade_hit<false>(S, P.opt, rec, ro, rd, rng, b);
    if (COUNT) ps.shaded_hits++;
    if (COUNT && P.row_cost) atomicAdd(P.row_cost + WF_U(F_IDX, slot) / (uint32_t)S.x_res, 80u);
    WF_U(F_RNG, slot) = rng;
    uint32_t aov = meta & WF_AOV;
    float4* sc = sc_ + (size_t)slot * WF_SCRATCH_F4;
    if (b.opaque) {
        if (COUNT) ps.nee_samples++;
        const V3 light = wf_ld3(st_, F_LAX, slot);
        const V3 reduction = wf_ld3(st_, F_RX, slot);
        if (bounce == 0) {
            __stcg(sc + 3, make_float4(b.normal.x, b.normal.y, b.normal.z, b.tangent.x));
            __stcg(sc + 4, make_float4(b.tangent.y, b.tangent.z, b.bitangent.x, b.bitangent.y));
            __stcg((float*)(sc + 5), b.bitangent.z);
            aov = WF_AOV;
        }
        wf_st3(st_, F_RX, slot, reduction * b.factor);
        if (b.need_shadow) {
            const V3 lb = light + reduction * b.add_blocked;
            wf_st3(st_, F_LAX, slot, light + reduction * b.add_open);
            __stcg(sc + 0, make_float4(lb.x, lb.y, lb.z, b.next_o.x));
            __stcg(sc + 1, make_float4(b.next_o.y, b.next_o.z, b.next_d.x, b.next_d.y));
            __stcg((float2*)(sc + 2), make_float2(b.next_d.z, __int_as_float(b.own_slot)));
            wf_st3(st_, F_OX, slot, b.so);
            wf_st3(st_, F_DX, slot, b.sd);
            WF_U(F_META, slot) = WF_META(s_left, bounce, WPH_SHADOW) | (b.shadow_any ? WF_ANY : 0u) | aov;
            return WQ_T;
        }
        wf_st3(st_, F_LAX, slot, light + reduction * b.add_open);
    }
    wf_st3(st_, F_OX, slot, b.next_o);
    wf_st3(st_, F_DX, slot, b.next_d);
    bounce++;
    if (bounce >= P.opt.max_depth) {
        WF_U(F_META, slot) = WF_META(s_left, bounce, WPH_COMPLETE) | aov;
        return WQ_E;
    }
    WF_U(F_META, slot) = WF_META(s_left, bounce, WPH_NEAREST) | aov;
    return WQ_T;
}

// QE handler. All 32 lanes call (converged); lanes without a slot pass slot = -1.
// shadow answers (kernel.cpp:555-587), misses (:517-522), finished samples (:597-645), new samples
// (:483-505) and new pixels. Returns true when the slot has a ray for QT.
// Pixel tickets are handed to a CTA in chunks of WF_CHUNK consecutive tickets (a compact 32x8-pixel
// region, er_ticket_pixel), so that the paths one SM works on at a time are neighbours in the image and
// their rays meet in the same BVH records and triangles (L1 hits). Local ticket l of the CTA belongs to
// local chunk l / WF_CHUNK; the lane that draws a chunk's first ticket reserves the chunk from the global
// counter and publishes its base in a small ring of tagged entries, the others wait for that entry.
#ifndef WF_CHUNK
#define WF_CHUNK 256
#endif
#define WF_CHUNK_RING 64
struct WfTickets {
    unsigned next;                   // local tickets handed out so far
    unsigned tag[WF_CHUNK_RING];     // local chunk id + 1 once base[] is valid
    unsigned base[WF_CHUNK_RING];    // first global ticket of the chunk
};

// All 32 lanes call (converged); lanes with pred hold local ticket `local`. Returns the global ticket.
ER_D unsigned wf_ticket(WfTickets& tk, const RenderParams& P, bool pred, unsigned local) {
    const unsigned k = local / WF_CHUNK, e = k % WF_CHUNK_RING;
    if (pred && local % WF_CHUNK == 0) {
        const unsigned g = atomicAdd(P.ticket, (unsigned)WF_CHUNK);
        ((volatile unsigned*)tk.base)[e] = g;
        __threadfence_block();
        ((volatile unsigned*)tk.tag)[e] = k + 1u;
    }
    __syncwarp();
    unsigned w = 0xffffffffu;
    if (pred) {
        bool ok = true;
        for (unsigned spin = 0; ((volatile unsigned*)tk.tag)[e] != k + 1u; spin++)
            if (spin > WF_SPIN_LIMIT) {  // cannot happen: the ticket is dropped and the launch reported as failed
                if (g_wf_watchdog) atomicAdd_system(g_wf_watchdog, 0x10000);
                ok = false;
                break;
            }
        __threadfence_block();
        if (ok) w = ((volatile unsigned*)tk.base)[e] + local % WF_CHUNK;
    }
    return w;
}

// CHUNKED launches hand a pixel out in chunks of chunk_spp samples. The first work_total tickets are the
// pixels themselves; a slot that finishes a chunk which is not the pixel's last appends the pixel to a
// READY RING in global memory, and ticket work_total + j is served by ring entry j: whichever pixel became
// ready j-th. A slot therefore waits only when no pixel at all is ready (fewer pixels than slots, or the
// very end of the launch), never for one particular pixel. The ring has one position per hand-over of the
// launch (er_api.cu), so no position is written twice. Entry = (lap tag << 24) | pixel index, written
// with a release store after the pixel's record {rng, samples, chunks done} and its passes, and read with
// a RELAXED gpu-scope load: an acquire load makes the compiler invalidate the SM's whole L1 (CCTL.IVALL),
// which is where the upper BVH levels live (measured: -6 %); what the reader loads afterwards (record,
// passes) is loaded past L1 (ld.cg) behind a branch on the entry.
ER_D unsigned wf_ld_relaxed(const unsigned* p) {
    unsigned v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
ER_D void wf_st_release(unsigned* p, unsigned v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Returns where the slot goes next: 0 nowhere (no work left), 1 QT (it has a ray), 2 QW (it holds a ticket
// for a ready-ring entry that has not been written yet, and asks again later).
// CHUNKED = false: one ticket = all of a pixel's samples of this launch (no hand-over, no QW).
template <bool COUNT, bool CHUNKED>
ER_D int wf_epilogue(const DScene& S, const RenderParams& P, uint32_t* st_, float4* sc_, int* s_live, WfTickets& tk,
                     int slot, unsigned& done) {
    const bool have = slot >= 0;
    const size_t NP = (size_t)S.x_res * S.y_res;
    bool sample_done = false, to_trace = false, need_fetch = false, start_sample = false, dead = false;
    bool check = false, retry = false, again = false;
    bool from_ring = false;
    uint32_t meta = 0, rng = 0, sa = 0, s_left = 0, ticket = 0;
    size_t idx = 0;
    int bounce = 0;
    V3 light = v3s(0.f);
    float4* sc = sc_ + (size_t)(have ? slot : 0) * WF_SCRATCH_F4;
    if (have) {
        meta = WF_U(F_META, slot);
        const int phase = (int)((meta >> 24) & 7u);
        bounce = (int)((meta >> 16) & 0xffu);
        s_left = meta & 0xffffu;
        if (phase == WPH_SHADOW) {
            const int res = (int)WF_U(F_RES, slot);
            const float4 r0 = __ldcg(sc), r1 = __ldcg(sc + 1);
            const float2 r2 = __ldcg((const float2*)(sc + 2));
            const bool blocked = res >= 0 && ((meta & WF_ANY) || res != __float_as_int(r2.y));
            light = blocked ? v3(r0.x, r0.y, r0.z) : wf_ld3(st_, F_LAX, slot);
            bounce++;
            if (bounce >= P.opt.max_depth) {
                sample_done = true;
            } else {
                wf_st3(st_, F_LAX, slot, light);
                wf_st3(st_, F_OX, slot, v3(r0.w, r1.x, r1.y));
                wf_st3(st_, F_DX, slot, v3(r1.z, r1.w, r2.x));
                WF_U(F_META, slot) = WF_META(s_left, bounce, WPH_NEAREST) | (meta & WF_AOV);
                to_trace = true;
            }
        } else if (phase == WPH_NEAREST) {  // miss: environment (kernel.cpp:517-522)
            const V3 rd = wf_ld3(st_, F_DX, slot);
            float u, v;
            er_spherical_mapping(-1.0f * rd, u, v);
            light = wf_ld3(st_, F_LAX, slot) + wf_ld3(st_, F_RX, slot) * er_tex_filtered(S.hdri, u, v);
            sample_done = true;
        } else if (phase == WPH_COMPLETE) {
            light = wf_ld3(st_, F_LAX, slot);
            sample_done = true;
        } else if (CHUNKED && phase == WPH_RETRY) {
            ticket = WF_U(F_RES, slot);
            check = from_ring = true;
        } else {
            need_fetch = true;
        }
    }
    if (sample_done) {
        idx = (size_t)WF_U(F_IDX, slot);
        WF_ASSERT(idx < NP, 5, idx);
#ifdef WF_CHECK
        if (!(idx < NP)) idx = 0;
#endif
        const float2 sp = __ldcg((const float2*)(sc + 2) + 1);  // (sample counter, chunks done) in one load
        sa = __float_as_uint(sp.x);
        light = er_clamp(light, 0.f, 10.f);
        V3 an = v3s(0.f), at = v3s(0.f), ab = v3s(0.f);
        if (meta & WF_AOV) {
            const float4 r3 = __ldcg(sc + 3), r4 = __ldcg(sc + 4);
            const float r5 = __ldcg((const float*)(sc + 5));
            an = v3(r3.x, r3.y, r3.z);
            at = v3(r3.w, r4.x, r4.y);
            ab = v3(r4.z, r4.w, r5);
        }
        sa = er_accumulate(S.passes, NP, idx, sa, light, an, at, ab);
        done++;
        s_left--;
        if (s_left > 0) {
            __stcg((float*)(sc + 2) + 2, __uint_as_float(sa));
            rng = WF_U(F_RNG, slot);
            start_sample = true;
        } else {
            // the chunk (or, not CHUNKED, the pixel's share of this launch) is done
            const uint32_t cdone = __float_as_uint(sp.y) + (CHUNKED ? 1u : 0u);
            __stcg(S.pix + idx, make_uint4(WF_U(F_RNG, slot), sa, cdone, 0u));
            if (CHUNKED) again = cdone - P.progress_base < P.chunk_count;
#ifdef WF_TIMELINE
            if (P.pixel_done != nullptr && !again) {
                unsigned long long ns;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns));
                P.pixel_done[idx] = (unsigned)(ns >> 10);
            }
#endif
            need_fetch = true;
        }
    }
    if (CHUNKED) {
        // pixels with chunks left go to the ready ring (warp-aggregated tail, one release store each)
        const unsigned am = __ballot_sync(WF_FULLMASK, again);
        if (am) {
            const int leader = __ffs((int)am) - 1;
            unsigned j = 0;
            if ((int)wf_lane() == leader) j = atomicAdd(P.ring_tail, (unsigned)__popc(am));
            j = __shfl_sync(WF_FULLMASK, j, leader) + (unsigned)__popc(am & ((1u << wf_lane()) - 1u));
            if (again) wf_st_release(P.ring + (j & P.ring_mask), ((((j >> P.ring_shift) + 1u) & 0xffu) << 24) | (unsigned)idx);
        }
    }
    // new pixel: warp-aggregated ticket
    for (;;) {
        const unsigned m = __ballot_sync(WF_FULLMASK, need_fetch);
        if (!m) break;
        const int leader = __ffs((int)m) - 1;
        unsigned base = 0;
        if ((int)wf_lane() == leader) base = atomicAdd(&tk.next, (unsigned)__popc(m));
        base = __shfl_sync(WF_FULLMASK, base, leader);
        const unsigned g = wf_ticket(tk, P, need_fetch, base + (unsigned)__popc(m & ((1u << wf_lane()) - 1u)));
        if (need_fetch) {
            int x, y;
            if (g >= (CHUNKED ? P.ticket_total : P.work_total)) {  // also the 0xffffffff of a dropped ticket
                need_fetch = false;
                dead = true;
            } else if (CHUNKED && g >= P.work_total) {
                ticket = g;
                need_fetch = false;
                check = from_ring = true;
            } else if (er_ticket_pixel(S, P, g, x, y)) {
                idx = (size_t)y * S.x_res + x;
                WF_U(F_IDX, slot) = (uint32_t)idx;
                need_fetch = false;
                check = true;
            }
        }
    }
    if (check) {
        bool ready = true;
        if (CHUNKED && from_ring) {
            const unsigned j = ticket - P.work_total;
            const unsigned e = wf_ld_relaxed(P.ring + (j & P.ring_mask));
            ready = (e >> 24) == (((j >> P.ring_shift) + 1u) & 0xffu);
            if (ready) {
                idx = (size_t)(e & 0xffffffu);
                WF_ASSERT(idx < NP, 9, idx);
#ifdef WF_CHECK
                if (!(idx < NP)) idx = 0;
#endif
                WF_U(F_IDX, slot) = (uint32_t)idx;
            }
        }
        if (ready) {
            // the pixel's RNG stream and running means continue where its previous chunk / launch left them
            const uint4 px = __ldcg(S.pix + idx);
            rng = px.x;
            __stcg((float2*)(sc + 2) + 1, make_float2(__uint_as_float(px.y), __uint_as_float(px.z)));
            if (CHUNKED) {
                // chunk c of the pixel: its size by level, and the samples the earlier chunks held
                const uint32_t c = px.z - P.progress_base;
                const uint32_t c1 = c < P.chunk_n1 ? c : P.chunk_n1, r1 = c - c1;
                const uint32_t c2 = r1 < P.chunk_n2 ? r1 : P.chunk_n2, c3 = r1 - c2;
                const uint32_t before = c1 * (uint32_t)P.chunk_spp + c2 * (uint32_t)P.chunk_k2 + c3 * (uint32_t)P.chunk_k3;
                const uint32_t size = c < P.chunk_n1 ? (uint32_t)P.chunk_spp : r1 < P.chunk_n2 ? (uint32_t)P.chunk_k2 : (uint32_t)P.chunk_k3;
                const uint32_t left = (uint32_t)P.spp - before;
                s_left = left < size ? left : size;
            } else {
                s_left = (uint32_t)P.spp;
            }
            start_sample = true;
        } else {
            WF_U(F_RES, slot) = ticket;
            WF_U(F_META, slot) = WF_META(0, 0, WPH_RETRY);
            retry = true;
        }
    }
    {
        const unsigned md = __ballot_sync(WF_FULLMASK, dead);
        if (dead) WF_U(F_META, slot) = WF_META(0, 0, WPH_DEAD);
        if (md && wf_lane() == 0) atomicSub(s_live, __popc(md));
    }
    if (start_sample) {  // kernel.cpp:492-505
        if (!sample_done) idx = (size_t)WF_U(F_IDX, slot);
        const int x = (int)(idx % (size_t)S.x_res), y = (int)(idx / (size_t)S.x_res);
        float c0 = er_rng_next(rng), c1 = er_rng_next(rng), c2 = er_rng_next(rng), c3 = er_rng_next(rng),
              c4 = er_rng_next(rng);
        V3 ro, rd;
        if (P.opt.rtl)
            er_camera_ray(S, x, y, c4, c3, c2, c1, c0, ro, rd);
        else
            er_camera_ray(S, x, y, c0, c1, c2, c3, c4, ro, rd);
        wf_st3(st_, F_OX, slot, ro);
        wf_st3(st_, F_DX, slot, rd);
        wf_st3(st_, F_LAX, slot, v3s(0.f));
        wf_st3(st_, F_RX, slot, v3s(1.f));
        WF_U(F_RNG, slot) = rng;
        WF_U(F_META, slot) = WF_META(s_left, 0, WPH_NEAREST);
        to_trace = true;
    }
    return to_trace ? 1 : retry ? 2 : 0;
}

// ---------------------------------------------------------------------------------------------
// the kernel
// ---------------------------------------------------------------------------------------------
ER_D void wf_trav_idle(Trav& t) {
    t.state = T_IDLE;
    t.slot = -1;
    t.o = t.d = t.inv = v3s(0.f);
    t.omax = 0.f;
    t.path = t.todo = 0;
    t.depth = 0;
    t.any = t.slow = t.shadow = false;
    t.best_hi = 0.f;
    t.st_slot = -1;
    t.st_lo = 0.f;
    t.st_dist = -1.f;
    t.result = -1;
    t.lf_cur = t.lf_end = t.lf_mid = t.lf_nL = t.lf_nR = 0;
    t.lf_slack = 0.f;
    t.work = 0;
}

// load the ray of `slot` into this lane (fresh query, or one suspended by another warp)
ER_D void wf_trav_load(const uint32_t* st_, int slot, Trav& t) {
    const uint32_t meta = WF_U(F_META, slot);
    WF_ASSERT(((meta >> 24) & 7u) == WPH_NEAREST || ((meta >> 24) & 7u) == WPH_SHADOW, 7, meta);
    t.slot = slot;
    t.o = wf_ld3(st_, F_OX, slot);
    t.d = wf_ld3(st_, F_DX, slot);
    // BVH::intersect computes dirfrac per box (BVH.cpp:29-33); hoisted: same IEEE division, same bits
    t.inv = v3(1.0f / t.d.x, 1.0f / t.d.y, 1.0f / t.d.z);
    t.omax = fmaxf(fmaxf(fabsf(t.o.x), fabsf(t.o.y)), fabsf(t.o.z));
    t.slow = !(isfinite(t.inv.x) && isfinite(t.inv.y) && isfinite(t.inv.z) && isfinite(t.omax));
    t.any = (meta & WF_ANY) != 0;
    t.shadow = ((meta >> 24) & 7u) == WPH_SHADOW;
    t.result = -1;
    t.work = 0;
    if (meta & WF_SUSP) {
        t.path = WF_U(F_TPATH, slot);
        t.todo = WF_U(F_TTODO, slot);
        t.depth = (int)WF_U(F_TDEPTH, slot);
        t.best_hi = WF_F(F_THI, slot);
        t.st_slot = (int)WF_U(F_TSLOT, slot);
        t.st_lo = WF_F(F_TLO, slot);
        t.st_dist = WF_F(F_TDIST, slot);
    } else {
        t.path = t.todo = 0;
        t.depth = 0;
        t.best_hi = __int_as_float(0x7f800000);
        t.st_slot = -1;
        t.st_lo = 0.f;
        t.st_dist = -1.f;
    }
    t.state = T_DESCEND;
}

__device__ __noinline__ void wf_watchdog_report(int* watchdog, int* flag, int act, int nt, int nh, int ne, int live, int w5, int w6, int w7) {
    atomicAdd(flag, 1);
    atomicAdd_system(watchdog, 1);
    if (atomicCAS_system(watchdog + 1, 0, (int)blockIdx.x + 1) == 0 || watchdog[1] == (int)blockIdx.x + 1) {
        int* w = watchdog + 16 + 8 * (threadIdx.x >> 5);
        w[0] = act, w[1] = nt, w[2] = nh, w[3] = ne, w[4] = live;
        w[5] = w5;  // chunked: ready-ring entries pushed so far, else this warp's scheduler iterations
        w[6] = w6;  // the global ticket counter
        w[7] = w7;
    }
}

template <bool COUNT, bool CHUNKED>
__global__ void __launch_bounds__(WF_THREADS, WF_MIN_CTAS) k_render_wavefront(const __grid_constant__ DScene S,
                                                                   const __grid_constant__ RenderParams P) {
    extern __shared__ __align__(16) uint32_t st_[];
    __shared__ WfQueue s_q[CHUNKED ? 4 : 3];
    __shared__ int s_live;
    __shared__ int s_mode;  // bit 0: QH is being drained, bit 1: QE is being drained
    __shared__ WfTickets s_tk;

    for (int i = threadIdx.x; i < WF_RING; i += blockDim.x) {
        s_q[WQ_T].buf[i] = 0u;
        s_q[WQ_H].buf[i] = 0u;
        if (CHUNKED) s_q[CHUNKED ? WQ_W : 0].buf[i] = 0u;
        s_q[WQ_E].buf[i] = i < P.slot_limit ? (1u << WF_QBITS) | (unsigned)i : 0u;  // every slot starts by fetching a pixel (lap 0 of QE)
        if (i < WF_SLOTS) WF_U(F_META, i) = WF_META(0, 0, WPH_FETCH);
        if (i < WF_CHUNK_RING) s_tk.tag[i] = 0u;
    }
    // the top levels of the tree, heap order (see WF_TOP_LEVELS)
    float4* const top_ = (float4*)(st_ + F_COUNT * WF_SLOTS);
    for (int i = threadIdx.x; i < (WF_TOP_LEVELS > 0 ? WF_TOP_RECORDS : 0) * 4; i += blockDim.x) {
        const unsigned h = (unsigned)i >> 2, d = 31u - (unsigned)__clz((int)(h + 1u)), p = h + 1u - (1u << d);
        const unsigned pathbits = d ? p << (ER_DEPTH - d) : 0u;
        top_[i] = S.pairs[4 * (size_t)(pathbits + d - (unsigned)__popc(pathbits)) + (i & 3)];
    }
    if (threadIdx.x == 0) {
        s_q[WQ_T].head = s_q[WQ_T].tail = 0;
        s_q[WQ_H].head = s_q[WQ_H].tail = 0;
        if (CHUNKED) s_q[CHUNKED ? WQ_W : 0].head = s_q[CHUNKED ? WQ_W : 0].tail = 0;
        s_q[WQ_E].head = 0;
        s_q[WQ_E].tail = (unsigned)P.slot_limit;
        s_live = P.slot_limit;
        s_mode = 2;  // every slot starts in QE
        s_tk.next = 0;
    }
    __syncthreads();

    float4* const sc_ = P.scratch + (size_t)blockIdx.x * WF_SLOTS * WF_SCRATCH_F4;
    PathStats ps = {};
    unsigned done = 0;
    Trav t;
    wf_trav_idle(t);
    const int thr_descend = P.thr_descend, coop_min = P.coop_min, pf_min = P.prefetch_min;
    float prune = 0.f;  // WF_LEAN: prune threshold of this lane's best hit so far (st_step)
    (void)pf_min;
    (void)prune;
    const long long t_start = clock64();
    unsigned iter = 0;

    for (;;) {
        // watchdog: a launch that runs longer than P.watchdog_cycles gives up and reports its scheduler
        // state through the host-mapped record (er_api.cu turns that into an error, never into a hang)
        ++iter;
#ifdef WF_TIMELINE  // debugging aid, compiled in by `make EXTRA=-DWF_TIMELINE` only (it costs C2 3 % through code layout)
        if ((iter & 255u) == 0 && P.timeline != nullptr && threadIdx.x == 0) {
            const unsigned k = (unsigned)((clock64() - t_start) >> P.timeline_shift);
            if (k < 512u) {
                unsigned* w = P.timeline + ((size_t)blockIdx.x * 512u + k) * 8u;
                if (w[0] == 0u) {
                    w[1] = (unsigned)wf_size(s_q[WQ_T]), w[2] = (unsigned)wf_size(s_q[WQ_H]), w[3] = (unsigned)wf_size(s_q[WQ_E]);
                    w[4] = *(volatile unsigned*)P.ticket;
                    w[0] = 1u + (unsigned)*(volatile int*)&s_live;
                }
            }
        }
#endif
        if ((iter & 255u) == 0 && P.watchdog != nullptr) {
            bool trip = false;
            if (wf_lane() == 0) trip = clock64() - t_start > P.watchdog_cycles || *(volatile int*)P.watchdog_flag != 0;
            if (__shfl_sync(WF_FULLMASK, trip, 0)) {
                const unsigned act = __ballot_sync(WF_FULLMASK, t.state != T_IDLE);
                if (wf_lane() == 0)  // the report is out of line: it runs once in a launch that has already failed
                    wf_watchdog_report(P.watchdog, P.watchdog_flag, (int)act, wf_size(s_q[WQ_T]), wf_size(s_q[WQ_H]), wf_size(s_q[WQ_E]),
                                       *(volatile int*)&s_live, CHUNKED ? (int)*(volatile unsigned*)P.ring_tail : (int)iter,
                                       (int)*(volatile unsigned*)P.ticket,
                                       (t.state & 0xf) | ((t.depth & 0xf) << 4) | ((CHUNKED ? wf_size(s_q[WQ_W]) : 0) << 8));
                break;
            }
        }
        // ---- 1. finished rays: publish the answer, hand the slot to the shading queues ------------
        {
            const bool fin = t.state == T_DONE;
            int res = -1;
            if (fin) {
                res = t.any ? t.result : t.st_slot;
                WF_U(F_RES, t.slot) = (uint32_t)res;
                if (COUNT) ps.rays++;
                if (COUNT && P.row_cost) atomicAdd(P.row_cost + WF_U(F_IDX, t.slot) / (uint32_t)S.x_res, t.work + 24u);
            }
            const bool to_h = fin && !t.shadow && res >= 0;
            wf_push(s_q[WQ_H], to_h, t.slot);
            wf_push(s_q[WQ_E], fin && !to_h, t.slot);
            if (fin) t.state = T_IDLE;
        }
        // ---- 2. role: dense shading batches come first ----------------------------------------------
        int nact = __popc(__ballot_sync(WF_FULLMASK, t.state != T_IDLE));
        // lanes 0..3 read one queue's head and tail each (this iteration is on every warp's critical path)
        int qn = 0;
        if (wf_lane() < (CHUNKED ? 4 : 3)) qn = wf_size(s_q[wf_lane()]);
        else if (wf_lane() == 4) qn = *(volatile int*)&s_live;
        const int nt = __shfl_sync(WF_FULLMASK, qn, WQ_T), nh = __shfl_sync(WF_FULLMASK, qn, WQ_H);
        const int ne = __shfl_sync(WF_FULLMASK, qn, WQ_E), nw = CHUNKED ? __shfl_sync(WF_FULLMASK, qn, WQ_W) : 0;
        // paths of this CTA that still have work (in lanes or in the three work queues); slots that only wait for a
        // chunk to become ready (QW) have nothing to run: near the end of a chunked launch most slots sit there
        const int live_now = __shfl_sync(WF_FULLMASK, qn, 4) - nw;
        // When few paths are left, a warp beyond live / park_lanes stops taking work: fewer warps, fuller lanes and
        // fewer instruction streams in the instruction cache (with 32 live paths per SM the 20 warps of the CTA
        // spent 60-90 % of their stall samples waiting for instructions, profiles/r04g).
        const bool parked = P.park_lanes > 0 && (int)(threadIdx.x >> 5) >= 2 && (int)(threadIdx.x >> 5) * P.park_lanes >= live_now;
        if (parked && nact == 0) {
            if (live_now + nw == 0) break;
            __nanosleep((unsigned)P.park_ns);
            continue;
        }
        // the batch threshold follows the live paths: what is left once every lane of the CTA has a ray
        int thr_b = live_now - P.batch_reserve;
        thr_b = thr_b < P.batch_min ? P.batch_min : thr_b;
        thr_b = thr_b > P.thr_batch ? P.thr_batch : thr_b;
        // Shading is batched in TIME as well: a queue is drained only once it holds thr_batch entries
        // (hysteresis: until it falls below one warp's worth), so that several warps run through the long
        // straight-line shading code close behind each other and share its instruction fetches.
        int mode = 0;
        if (wf_lane() == 0) {
            mode = *(volatile int*)&s_mode;
            int m2 = mode;
            if (nh >= thr_b) m2 |= 1; else if (nh < 32) m2 &= ~1;
            if (ne >= thr_b) m2 |= 2; else if (ne < 32) m2 &= ~2;
            if (m2 != mode) {
                if (m2 & ~mode) atomicOr(&s_mode, m2 & ~mode);
                if (mode & ~m2) atomicAnd(&s_mode, ~(mode & ~m2));
                mode = m2;
            }
        }
        mode = __shfl_sync(WF_FULLMASK, mode, 0);
        int role = -1;
        if (nh >= 32 && (mode & 1))
            role = WQ_H;
        else if (ne >= 32 && (mode & 2))
            role = WQ_E;
        else if (nt == 0 && nact < P.thr_switch && (nh > 0 || ne > 0))
            role = nh >= ne ? WQ_H : WQ_E;
        else if (CHUNKED && nw > 0 && ((iter & 63u) == 0 || (nt == 0 && nact == 0)))
            role = WQ_W;
        if (role >= 0) {
            int slot;
            const int n = wf_claim(s_q[role], 32, slot);
            if (n > 0) {
                // suspend this warp's rays (all are between two records) so that the batch runs dense
                const bool susp = t.state == T_DESCEND;
                if (susp) {
                    WF_U(F_TPATH, t.slot) = t.path;
                    WF_U(F_TTODO, t.slot) = t.todo;
                    WF_U(F_TDEPTH, t.slot) = (uint32_t)t.depth;
                    WF_SETF(F_THI, t.slot, t.best_hi);
                    WF_U(F_TSLOT, t.slot) = (uint32_t)t.st_slot;
                    WF_SETF(F_TLO, t.slot, t.st_lo);
                    WF_SETF(F_TDIST, t.slot, t.st_dist);
                    WF_U(F_META, t.slot) |= WF_SUSP;
                }
                wf_push(s_q[WQ_T], susp, t.slot);
                wf_trav_idle(t);
                if (role == WQ_H) {
                    int next = -1;
                    if (slot >= 0) next = wf_shade<COUNT>(S, P, st_, sc_, slot, ps);
                    wf_push(s_q[WQ_T], next == WQ_T, slot);
                    wf_push(s_q[WQ_E], next == WQ_E, slot);
                } else {
                    const int next = wf_epilogue<COUNT, CHUNKED>(S, P, st_, sc_, &s_live, s_tk, slot, done);
                    wf_push(s_q[WQ_T], next == 1, slot);
                    if (CHUNKED) wf_push(s_q[CHUNKED ? WQ_W : 0], next == 2, slot);
                }
                continue;
            }
        }
        // ---- 3. refill idle lanes from QT -----------------------------------------------------------
        if (nact < 32 && nt > 0 && !parked) {
            int slot;
            const int n = wf_claim(s_q[WQ_T], 32 - nact, slot);
            // the i-th idle lane takes the i-th claimed slot
            const unsigned idle = __ballot_sync(WF_FULLMASK, t.state == T_IDLE);
            const int my = __popc(idle & ((1u << wf_lane()) - 1u));
            const int got = __shfl_sync(WF_FULLMASK, slot, my & 31);
            if (t.state == T_IDLE && my < n && got >= 0) {
                wf_trav_load(st_, got, t);
                WF_U(F_META, got) &= ~WF_SUSP;
#ifdef WF_LEAN
                prune = t.best_hi * 1.0001f + (2e-5f * t.omax + 1e-6f);
                if (t.slow) {  // the literal box test, one ray at a time, out of line (never suspended: it finishes here)
                    t.st_slot = t.result = st_trace_slow<COUNT>(S, t.any, t.o.x, t.o.y, t.o.z, t.d.x, t.d.y, t.d.z, ps.tr);
                    t.state = T_DONE;
                }
#endif
            }
            nact += n;
        }
        if (nact == 0) {
            int live = 0;
            if (wf_lane() == 0) live = *(volatile int*)&s_live;
            if (__shfl_sync(WF_FULLMASK, live, 0) == 0) break;
            __nanosleep(200);
            continue;
        }
#ifdef WF_LEAN
        // ---- 4. descend together (interior records only) until too few lanes still descend -----------
        const int nidle = 32 - nact;
        for (;;) {
            const bool want = t.state == T_DESCEND && t.depth < ST_LEAF_DEPTH;
            const int nd = __popc(__ballot_sync(WF_FULLMASK, want));
            if (nd == 0) break;
            // lanes waiting for something else: at a leaf pair, finished, or idle with rays queued
            const int waiting = (nact - nd) + (nt > 0 ? nidle : 0);
            if (nd < thr_descend && waiting > 0) break;
            if (want) {
                st_step<COUNT>(S, t, prune, ps.tr, WF_TOP_LEVELS > 0 ? top_ : nullptr);
                if (COUNT) t.work += 4;
            }
        }
        // ---- 5. leaf phase: lanes at depth 17 open their leaf pair and test its triangles (small pairs per
        // lane, big ones by the whole warp), until no lane of the warp is at a leaf
        for (;;) {
            if (t.state == T_DESCEND && t.depth == ST_LEAF_DEPTH) {
                st_leaf_open<COUNT>(S, t, prune, ps.tr);
                if (COUNT) t.work += 4;
            }
            const bool want = t.state == T_LEAF;
            const bool again = t.state == T_DESCEND && t.depth == ST_LEAF_DEPTH;  // popped onto the sibling pair
            const unsigned code = want ? (t.lf_end - t.lf_cur >= coop_min ? 95u - wf_lane() : 1u) : again ? 1u : 0u;
            const unsigned top = __reduce_max_sync(WF_FULLMASK, code);
            if (top == 0u) break;
            if (top >= 64u) {
#ifdef WF_GROUP_BOXES
                wf_leaf_coop_g<COUNT>(S, t, (int)(95u - top), ps.tr);
#else
                wf_leaf_coop<COUNT>(S, t, (int)(95u - top), ps.tr);
#endif
                continue;
            }
            if (want) wf_tri<COUNT>(S, t, ps.tr);
        }
        prune = t.best_hi * 1.0001f + (2e-5f * t.omax + 1e-6f);
    }
#else
        // ---- 4. descend together until too few lanes still descend ----------------------------------
        const int nidle = 32 - nact;
        for (;;) {
            const bool want = t.state == T_DESCEND;
            const int nd = __popc(__ballot_sync(WF_FULLMASK, want));
            if (nd == 0) break;
            // lanes waiting for something else: parked at a leaf pair, finished, or idle with rays queued
            const int waiting = (nact - nd) + (nt > 0 ? nidle : 0);
            if (nd < thr_descend && waiting > 0) break;
            if (want) wf_descend<COUNT>(S, t, ps.tr, pf_min);
        }
        // ---- 5. parked lanes test their leaf triangles ----------------------------------------------
        // A leaf pair with few triangles is tested by its own lane, all parked lanes side by side. A big
        // one (10M-triangle scenes: ~38 per leaf) is tested by the whole warp: 32 consecutive triangles
        // per step, coalesced loads, and the few that pass are handed to the owning lane in rank order.
        // (prefetching the triangles of all parked pairs with prefetch.global.L1 before testing them one
        // after the other was measured SLOWER on B200: C3 227 -> 217, C2 749 -> 731 Msamples/s)
        for (;;) {
            const bool want = t.state == T_LEAF;
            // one vote: 0 nobody parked, 1 small pairs only, 64 + (31 - lane) the first lane with a big pair
            const unsigned code = want ? (t.lf_end - t.lf_cur >= coop_min ? 95u - wf_lane() : 1u) : 0u;
            const unsigned top = __reduce_max_sync(WF_FULLMASK, code);
            if (top == 0u) break;
            if (top >= 64u) {
                wf_leaf_coop<COUNT>(S, t, (int)(95u - top), ps.tr);
                continue;
            }
            if (want) wf_tri<COUNT>(S, t, ps.tr);
        }
    }
#endif
    if (COUNT) er_flush_stats(P.counters, ps, done);
}
