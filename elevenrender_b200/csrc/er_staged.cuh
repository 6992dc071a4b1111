// er_staged.cuh — the staged render path: renderingKernel (src/kernel.cpp:477-646) as three small kernels
// per wave over per-pixel path state in HBM, instead of one persistent kernel with every role in it.
//
//   k_st_logic   kernel.cpp:483-505 (new sample: camera ray), :517-522 (miss: environment), :555-587 (shadow
//                answer), :597-645 (clamp + running means), per-pixel sample loop
//   k_st_trace   throwRay for every path that has a ray out (nearest-hit and shadow queries share the one
//                traversal site): BVH::transverse + Tri::hit, er_wavefront.cuh's warp traversal functions
//   k_st_shade   kernel.cpp:524-587 for every nearest hit (generateHitData, HDRI NEE, DisneySample/Eval/Pdf)
//
// Why (profiles/README.md, round 1): in the persistent kernel 38 % of the executed warp instructions were
// queue and role bookkeeping, the traversal ran at 20 warps per SM because it shared its register budget
// (96) with the shading code, and the 122 KB of SASS thrashed the 32 KB instruction cache. Here
//   * every pixel owns ONE path slot for the whole launch (slot = pixel ticket, 8x4-tile order), so there
//     are no queues at all: a slot's `phase` byte says which kernel has work for it, each kernel walks
//     the phase bytes in chunks claimed from one counter and compacts the slots it wants (ballot/popc) —
//     dense lanes in the shading code, spatially coherent rays in a warp;
//   * the traversal kernel holds only traversal state: 64 registers, 32 warps per SM, 6 KB of code;
//   * path state moves through HBM/L2 as 16-B vectors, ~300 B per ray: at 2-3 G rays/s that is 10-15 % of
//     the 6.5 TB/s the HBM delivers, and it overlaps the latency-bound traversal.
// A pixel's samples stay sequential (its xorshift stream persists across samples, kernel.cpp:483,645), so
// a launch of `spp` samples takes as many waves as the busiest pixel traces rays; late waves are thin.
// The host enqueues waves without synchronising; every kernel leaves at once when no slot is alive.
//
// Bit-identity: the per-path arithmetic is the same device functions the persistent kernel runs
// (er_shade_hit, er_camera_ray, er_accumulate, wf_descend/wf_tri/wf_leaf_coop/wf_candidate), every pixel's
// operations happen in the same order, so the passes equal k_render_wavefront's bit for bit
// (tests/test_gpu_staged.py) — and the oracle's within the radiance tolerance.
#pragma once
#include "er_wavefront.cuh"

enum {
    SP_DEAD = 0,         // no work left in this launch (or a ticket outside the image)
    SP_START = 1,        // launch begins: load the pixel record, start the first sample
    SP_NEAREST = 2,      // ray out: nearest-hit query                       -> k_st_trace
    SP_SHADOW = 3,       // ray out: shadow query, nearest hit wanted        -> k_st_trace
    SP_SHADOW_ANY = 4,   // ray out: shadow query, any hit answers it        -> k_st_trace
    SP_HIT = 5,          // nearest hit found                                -> k_st_shade
    SP_MISS = 6,         // nearest-hit query missed: environment            -> k_st_logic
    SP_SHADOW_DONE = 7,  // shadow query answered                            -> k_st_logic
    SP_COMPLETE = 8      // max depth reached                                -> k_st_logic
};

// per-slot state, SoA of 16-B vectors
//   ray[2s]   = (o.xyz, d.x)        ray[2s+1] = (d.y, d.z, -, sample counter bits)
//   lr[2s]    = (light.xyz, red.x)  lr[2s+1]  = (red.y, red.z, rng bits, meta bits)
//       meta: bits 0-15 samples left in this launch, 16-23 bounce, 24 first-hit AOVs valid
//   park[3s..]= (LB.xyz, NO.x) (NO.yz, ND.xy) (ND.z, own slot bits, -, -)   across a shadow query: light if it IS
//                                                                           blocked, the continued ray, the shaded triangle
//   aov[3s..] = (N.xyz, T.x) (T.yz, B.xy) (B.z, -, -, -)                    first-hit AOVs of the running sample
struct StagedState {
    float4* ray;
    float4* lr;
    float4* park;
    float4* aov;
    int* res;         // traversal answer: BVH slot of the hit, -1 = none
    uint8_t* phase;
    unsigned* ctr;    // [0] trace chunks claimed, [1] shade chunks, [2] logic chunks, [3] live slots
    unsigned slots;   // multiple of ST_CHUNK
};
enum { SC_TRACE = 0, SC_SHADE = 1, SC_LOGIC = 2, SC_LIVE = 3 };

#define ST_CHUNK 32           // slots per claimed chunk of the trace kernel: one 8x4-pixel tile, one per lane
#ifndef ST_TRACE_THREADS
#define ST_TRACE_THREADS 256
#endif
#define ST_TRACE_WARPS (ST_TRACE_THREADS / 32)
#define ST_LIST 64            // pending slots a warp can hold: < 32 left over + one chunk
#ifndef ST_TRACE_MIN_CTAS
#define ST_TRACE_MIN_CTAS 3
#endif
#define ST_BLOCK 256          // threads (= slots per chunk) of the shade / logic kernels
#define ST_META(s_left, bounce, aov) ((uint32_t)(s_left) | ((uint32_t)(bounce) << 16) | ((aov) ? 1u << 24 : 0u))

__global__ void k_st_init(const __grid_constant__ DScene S, const __grid_constant__ RenderParams P, StagedState W) {
    const unsigned s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s == 0) {
        W.ctr[SC_TRACE] = W.ctr[SC_SHADE] = W.ctr[SC_LOGIC] = 0u;
    }
    if (s >= W.slots) return;
    int x, y;
    const bool ok = s < P.work_total && er_ticket_pixel(S, P, s, x, y);
    W.phase[s] = ok ? SP_START : SP_DEAD;
    const unsigned m = __ballot_sync(__activemask(), ok);
    if (m && (threadIdx.x & 31u) == (unsigned)(__ffs((int)m) - 1)) atomicAdd(W.ctr + SC_LIVE, (unsigned)__popc(m));
}

// ---------------------------------------------------------------------------------------------------
// CTA-wide walk over the phase bytes: chunks of ST_BLOCK slots claimed from one counter, the wanted slots
// compacted into shared memory and handed to `body` by dense threads.
// ---------------------------------------------------------------------------------------------------
template <typename Want, typename Body>
ER_D void st_for_each(const StagedState& W, unsigned* counter, Want want, Body body) {
    __shared__ unsigned s_list[2 * ST_BLOCK];
    __shared__ unsigned s_warp_n[ST_BLOCK / 32];
    __shared__ unsigned s_chunk;
    const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const unsigned chunks = W.slots / ST_BLOCK;
    unsigned n = 0;  // entries in s_list (CTA-uniform)
    bool more = true;
    for (;;) {
        while (n < ST_BLOCK && more) {
            if (tid == 0) s_chunk = atomicAdd(counter, 1u);
            __syncthreads();
            const unsigned c = s_chunk;
            if (c >= chunks) {
                more = false;
                break;
            }
            const unsigned slot = c * ST_BLOCK + tid;
            const bool w = want(W.phase[slot]);
            const unsigned m = __ballot_sync(WF_FULLMASK, w);
            if (lane == 0) s_warp_n[warp] = (unsigned)__popc(m);
            __syncthreads();
            unsigned base = n, total = 0;
#pragma unroll
            for (unsigned k = 0; k < ST_BLOCK / 32; k++) {
                const unsigned v = s_warp_n[k];
                if (k < warp) base += v;
                total += v;
            }
            if (w) s_list[base + (unsigned)__popc(m & ((1u << lane) - 1u))] = slot;
            n += total;
            __syncthreads();
        }
        if (n == 0) break;
        const unsigned take = n < ST_BLOCK ? n : ST_BLOCK;
        body(tid < take ? (int)s_list[n - take + tid] : -1);
        n -= take;
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------------------
// k_st_logic: everything cheap between two rays of a path. All 32 lanes of a warp call the body (lanes
// without a slot pass -1): it holds warp votes.
// ---------------------------------------------------------------------------------------------------
template <bool COUNT>
__global__ void __launch_bounds__(ST_BLOCK) k_st_logic(const __grid_constant__ DScene S, const __grid_constant__ RenderParams P,
                                                        StagedState W) {
    if (*(volatile unsigned*)(W.ctr + SC_LIVE) == 0u) return;
    if (blockIdx.x == 0 && threadIdx.x == 0) W.ctr[SC_TRACE] = 0u;
    const size_t NP = (size_t)S.x_res * S.y_res;
    unsigned done = 0;
    st_for_each(
        W, W.ctr + SC_LOGIC, [](uint8_t ph) { return ph == SP_START || ph == SP_MISS || ph == SP_SHADOW_DONE || ph == SP_COMPLETE; },
        [&](int slot) {
            const bool have = slot >= 0;
            bool sample_done = false, start_sample = false, dead = false;
            V3 light = v3s(0.f), red = v3s(1.f);
            uint32_t rng = 0, meta = 0, sa = 0, s_left = 0;
            size_t idx = 0;
            int x = 0, y = 0;
            if (have) {
                er_ticket_pixel(S, P, (unsigned)slot, x, y);
                idx = (size_t)y * S.x_res + x;
                const int ph = W.phase[slot];
                if (ph == SP_START) {
                    // the pixel's RNG stream and running means continue where the previous launch left them
                    const uint4 px = __ldcg(S.pix + idx);
                    rng = px.x;
                    sa = px.y;
                    s_left = (uint32_t)P.spp;
                    start_sample = true;
                } else {
                    const float4 l0 = __ldcg(W.lr + 2 * (size_t)slot), l1 = __ldcg(W.lr + 2 * (size_t)slot + 1);
                    const float4 r1 = __ldcg(W.ray + 2 * (size_t)slot + 1);
                    light = v3(l0.x, l0.y, l0.z);
                    red = v3(l0.w, l1.x, l1.y);
                    rng = __float_as_uint(l1.z);
                    meta = __float_as_uint(l1.w);
                    sa = __float_as_uint(r1.w);
                    s_left = meta & 0xffffu;
                    int bounce = (int)((meta >> 16) & 0xffu);
                    if (ph == SP_SHADOW_DONE) {  // kernel.cpp:555-587
                        const int res = W.res[slot];
                        const float4 p0 = __ldcg(W.park + 3 * (size_t)slot), p1 = __ldcg(W.park + 3 * (size_t)slot + 1);
                        const float4 p2 = __ldcg(W.park + 3 * (size_t)slot + 2);
                        // p2.z: 1 = an any-hit query (the shaded triangle cannot be hit), any answer blocks
                        const bool blocked = res >= 0 && (__float_as_int(p2.z) != 0 || res != __float_as_int(p2.y));
                        if (blocked) light = v3(p0.x, p0.y, p0.z);
                        bounce++;
                        if (bounce >= P.opt.max_depth) {
                            sample_done = true;
                        } else {
                            __stcg(W.ray + 2 * (size_t)slot, make_float4(p0.w, p1.x, p1.y, p1.z));
                            __stcg(W.ray + 2 * (size_t)slot + 1, make_float4(p1.w, p2.x, 0.f, r1.w));
                            __stcg(W.lr + 2 * (size_t)slot, make_float4(light.x, light.y, light.z, red.x));
                            __stcg(W.lr + 2 * (size_t)slot + 1,
                                   make_float4(red.y, red.z, l1.z, __uint_as_float(ST_META(s_left, bounce, meta & (1u << 24)))));
                            W.phase[slot] = SP_NEAREST;
                        }
                    } else if (ph == SP_MISS) {  // environment (kernel.cpp:517-522)
                        const float4 r0 = __ldcg(W.ray + 2 * (size_t)slot);
                        const V3 rd = v3(r0.w, r1.x, r1.y);
                        float u, v;
                        er_spherical_mapping(-1.0f * rd, u, v);
                        light = light + red * er_tex_filtered(S.hdri, u, v);
                        sample_done = true;
                    } else {  // SP_COMPLETE
                        sample_done = true;
                    }
                }
            }
            if (sample_done) {  // kernel.cpp:597-645
                light = er_clamp(light, 0.f, 10.f);
                V3 an = v3s(0.f), at = v3s(0.f), ab = v3s(0.f);
                if (meta & (1u << 24)) {
                    const float4 a0 = __ldcg(W.aov + 3 * (size_t)slot), a1 = __ldcg(W.aov + 3 * (size_t)slot + 1);
                    const float a2 = __ldcg((const float*)(W.aov + 3 * (size_t)slot + 2));
                    an = v3(a0.x, a0.y, a0.z);
                    at = v3(a0.w, a1.x, a1.y);
                    ab = v3(a1.z, a1.w, a2);
                }
                sa = er_accumulate(S.passes, NP, idx, sa, light, an, at, ab);
                done++;
                s_left--;
                if (s_left > 0) {
                    start_sample = true;
                } else {
                    __stcg(S.pix + idx, make_uint4(rng, sa, 0u, 0u));
                    W.phase[slot] = SP_DEAD;
                    dead = true;
                }
            }
            if (start_sample) {  // kernel.cpp:492-505
                float c0 = er_rng_next(rng), c1 = er_rng_next(rng), c2 = er_rng_next(rng), c3 = er_rng_next(rng),
                      c4 = er_rng_next(rng);
                V3 ro, rd;
                if (P.opt.rtl)
                    er_camera_ray(S, x, y, c4, c3, c2, c1, c0, ro, rd);
                else
                    er_camera_ray(S, x, y, c0, c1, c2, c3, c4, ro, rd);
                __stcg(W.ray + 2 * (size_t)slot, make_float4(ro.x, ro.y, ro.z, rd.x));
                __stcg(W.ray + 2 * (size_t)slot + 1, make_float4(rd.y, rd.z, 0.f, __uint_as_float(sa)));
                __stcg(W.lr + 2 * (size_t)slot, make_float4(0.f, 0.f, 0.f, 1.f));
                __stcg(W.lr + 2 * (size_t)slot + 1, make_float4(1.f, 1.f, __uint_as_float(rng), __uint_as_float(ST_META(s_left, 0, 0))));
                W.phase[slot] = SP_NEAREST;
            }
            const unsigned md = __ballot_sync(WF_FULLMASK, dead);
            if (md && (threadIdx.x & 31u) == 0) atomicSub(W.ctr + SC_LIVE, (unsigned)__popc(md));
        });
    if (COUNT) {
        PathStats ps = {};
        er_flush_stats(P.counters, ps, done);
    }
}

// ---------------------------------------------------------------------------------------------------
// k_st_shade: kernel.cpp:524-587 for every nearest hit
// ---------------------------------------------------------------------------------------------------
template <bool COUNT>
__global__ void __launch_bounds__(ST_BLOCK) k_st_shade(const __grid_constant__ DScene S, const __grid_constant__ RenderParams P,
                                                        StagedState W) {
    if (*(volatile unsigned*)(W.ctr + SC_LIVE) == 0u) return;
    if (blockIdx.x == 0 && threadIdx.x == 0) W.ctr[SC_LOGIC] = 0u;
    PathStats ps = {};
    st_for_each(
        W, W.ctr + SC_SHADE, [](uint8_t ph) { return ph == SP_HIT; },
        [&](int slot) {
            if (slot < 0) return;
            const float4 r0 = __ldcg(W.ray + 2 * (size_t)slot), r1 = __ldcg(W.ray + 2 * (size_t)slot + 1);
            const float4 l0 = __ldcg(W.lr + 2 * (size_t)slot), l1 = __ldcg(W.lr + 2 * (size_t)slot + 1);
            const V3 ro = v3(r0.x, r0.y, r0.z), rd = v3(r0.w, r1.x, r1.y);
            V3 light = v3(l0.x, l0.y, l0.z);
            const V3 red = v3(l0.w, l1.x, l1.y);
            uint32_t rng = __float_as_uint(l1.z);
            const uint32_t meta = __float_as_uint(l1.w);
            const uint32_t s_left = meta & 0xffffu;
            int bounce = (int)((meta >> 16) & 0xffu);
            bool aov = (meta & (1u << 24)) != 0;
            HitRec rec;
            rec.slot = W.res[slot];
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
            V3 red_new = red;
            if (b.opaque) {
                if (COUNT) ps.nee_samples++;
                if (bounce == 0) {
                    __stcg(W.aov + 3 * (size_t)slot, make_float4(b.normal.x, b.normal.y, b.normal.z, b.tangent.x));
                    __stcg(W.aov + 3 * (size_t)slot + 1, make_float4(b.tangent.y, b.tangent.z, b.bitangent.x, b.bitangent.y));
                    __stcg((float*)(W.aov + 3 * (size_t)slot + 2), b.bitangent.z);
                    aov = true;
                }
                red_new = red * b.factor;
                if (b.need_shadow) {
                    const V3 lb = light + red * b.add_blocked;
                    light = light + red * b.add_open;
                    __stcg(W.park + 3 * (size_t)slot, make_float4(lb.x, lb.y, lb.z, b.next_o.x));
                    __stcg(W.park + 3 * (size_t)slot + 1, make_float4(b.next_o.y, b.next_o.z, b.next_d.x, b.next_d.y));
                    __stcg(W.park + 3 * (size_t)slot + 2,
                           make_float4(b.next_d.z, __int_as_float(b.own_slot), __int_as_float(b.shadow_any ? 1 : 0), 0.f));
                    __stcg(W.ray + 2 * (size_t)slot, make_float4(b.so.x, b.so.y, b.so.z, b.sd.x));
                    __stcg(W.ray + 2 * (size_t)slot + 1, make_float4(b.sd.y, b.sd.z, 0.f, r1.w));
                    __stcg(W.lr + 2 * (size_t)slot, make_float4(light.x, light.y, light.z, red_new.x));
                    __stcg(W.lr + 2 * (size_t)slot + 1,
                           make_float4(red_new.y, red_new.z, __uint_as_float(rng), __uint_as_float(ST_META(s_left, bounce, aov))));
                    W.phase[slot] = b.shadow_any ? SP_SHADOW_ANY : SP_SHADOW;
                    return;
                }
                light = light + red * b.add_open;
            }
            bounce++;
            __stcg(W.ray + 2 * (size_t)slot, make_float4(b.next_o.x, b.next_o.y, b.next_o.z, b.next_d.x));
            __stcg(W.ray + 2 * (size_t)slot + 1, make_float4(b.next_d.y, b.next_d.z, 0.f, r1.w));
            __stcg(W.lr + 2 * (size_t)slot, make_float4(light.x, light.y, light.z, red_new.x));
            __stcg(W.lr + 2 * (size_t)slot + 1,
                   make_float4(red_new.y, red_new.z, __uint_as_float(rng), __uint_as_float(ST_META(s_left, bounce, aov))));
            W.phase[slot] = bounce >= P.opt.max_depth ? SP_COMPLETE : SP_NEAREST;
        });
    if (COUNT) er_flush_stats(P.counters, ps, 0);
}

// ---------------------------------------------------------------------------------------------------
// k_st_trace: persistent warps; every warp claims chunks of ST_CHUNK consecutive slots (one 8x4-pixel tile),
// keeps the ones with a ray out in its own pending list and runs a while-while traversal over them,
// refilling lanes as their rays finish.
//
// The traversal is er_wavefront.cuh's (same visit set, same candidate selection: wf_candidate, wf_tri,
// wf_leaf_coop), with the hot loop cut down to what a lane needs between two records:
//   * st_step handles INTERIOR records only (depth 0..16); a lane that arrives at depth 17 waits for the
//     leaf phase, where the leaf-pair record is opened and its triangles are tested — no divergent
//     leaf branch inside the descend loop;
//   * rays with a zero or denormal direction component (BVH::intersect's NaN-ordered min/max matters for
//     them, BVH.cpp:27-61, Math.hpp:61-67) are rare: they are traced by the scalar er_traverse at refill
//     time, so the loop carries no `slow` variant of the box test;
//   * the prune threshold of the best hit so far (distance + margins) is one register, refreshed after
//     every leaf phase; per child it costs one FADD with the subtree's slack.
// ---------------------------------------------------------------------------------------------------
// Chunks are handed out in two levels so that the warps of one CTA work on neighbouring tiles at the same
// time (their rays meet in the same BVH records: L1 hits): a CTA claims a SUPER-chunk of ST_SUPER chunks
// (a 32x8-pixel region) from the global counter, its warps draw the chunks of it from one shared word
//   s_word = (super-chunk << 5) | draws from it so far (0 .. ST_SUPER - 1 are its chunks; ST_SUPER = being replaced)
// with one shared-memory atomicAdd; the warp that draws ST_SUPER exactly claims the next super-chunk and
// publishes it, warps that draw more wait for that (a few hundred cycles, once per 8 chunks).
#define ST_SUPER 8
#define ST_DRAW_BITS 5
#define ST_DRAW_MASK ((1u << ST_DRAW_BITS) - 1u)
// the draw count runs up to ST_SUPER (the draw that fetches) + 1 + one draw of every other warp that waits
static_assert(ST_SUPER + ST_TRACE_WARPS < (1 << ST_DRAW_BITS), "the draw count of a super-chunk must not carry into its id");

ER_D unsigned st_claim_chunk(unsigned* s_word, unsigned* g_counter) {  // lane 0 only
    for (;;) {
        const unsigned old = atomicAdd(s_word, 1u);
        const unsigned k = old & ST_DRAW_MASK;
        if (k < ST_SUPER) return (old >> ST_DRAW_BITS) * ST_SUPER + k;
        if (k == ST_SUPER) {
            const unsigned sup = atomicAdd(g_counter, 1u);
            atomicExch(s_word, (sup << ST_DRAW_BITS) | 1u);
            return sup * ST_SUPER;
        }
        while ((*(volatile unsigned*)s_word & ST_DRAW_MASK) >= ST_SUPER) __nanosleep(32);
    }
}

template <bool COUNT>
__global__ void __launch_bounds__(ST_TRACE_THREADS, ST_TRACE_MIN_CTAS) k_st_trace(const __grid_constant__ DScene S,
                                                                                 const __grid_constant__ RenderParams P,
                                                                                 StagedState W) {
    if (*(volatile unsigned*)(W.ctr + SC_LIVE) == 0u) return;
    if (blockIdx.x == 0 && threadIdx.x == 0) W.ctr[SC_SHADE] = 0u;
    __shared__ unsigned s_pending[ST_TRACE_WARPS][ST_LIST];
    __shared__ unsigned s_word;
    if (threadIdx.x == 0) s_word = ST_SUPER;
    __syncthreads();
    const unsigned lane = wf_lane();
    unsigned* const list = s_pending[threadIdx.x >> 5];
    const unsigned chunks = W.slots / ST_CHUNK;
    int list_n = 0;    // pending slots of this warp (warp-uniform)
    bool more = true;  // chunks left to claim (warp-uniform)
    PathStats ps = {};
    Trav t;
    wf_trav_idle(t);
    float prune = 0.f;
    // a warp steps its descending lanes until thr_leaf lanes wait at a leaf pair or thr_refill lanes are free
    const int thr_leaf = P.thr_descend, thr_refill = P.thr_switch, coop_min = P.coop_min;

    for (;;) {
        const bool at_leaf = t.state == T_DESCEND && t.depth == ST_LEAF_DEPTH;
        const bool stepping = t.state == T_DESCEND && t.depth < ST_LEAF_DEPTH;
        const int n_step = __popc(__ballot_sync(WF_FULLMASK, stepping));
        const int n_leaf = __popc(__ballot_sync(WF_FULLMASK, at_leaf));
        const int n_free = 32 - n_step - n_leaf;  // idle or finished
        const bool queued = list_n > 0 || more;
        if ((n_free >= thr_refill && queued) || n_step + n_leaf == 0) {
            // ---- finished rays: publish the answer, hand the slot to the next kernel ---------------------
            if (t.state == T_DONE) {
                const int res = t.st_slot, slot = t.slot & ST_SLOT_MASK;
                W.res[slot] = res;
                W.phase[slot] = (uint8_t)((t.slot & ST_SHADOW_BIT) ? SP_SHADOW_DONE : res >= 0 ? SP_HIT : SP_MISS);
                if (COUNT) ps.rays++;
                t.state = T_IDLE;
            }
            // ---- refill the free lanes ------------------------------------------------------------------
            while (list_n < n_free && more) {
                unsigned c = 0;
#ifdef ST_SIMPLE_CLAIM
                if (lane == 0) c = atomicAdd(W.ctr + SC_TRACE, 1u);
#else
                if (lane == 0) c = st_claim_chunk(&s_word, W.ctr + SC_TRACE);
#endif
                c = __shfl_sync(WF_FULLMASK, c, 0);
                if (c >= chunks) {
                    more = false;
                    break;
                }
                const unsigned ph = W.phase[(size_t)c * ST_CHUNK + lane];
                const bool w = ph == SP_NEAREST || ph == SP_SHADOW || ph == SP_SHADOW_ANY;
                const unsigned m = __ballot_sync(WF_FULLMASK, w);
                if (w) list[list_n + __popc(m & ((1u << lane) - 1u))] = c * ST_CHUNK + lane;
                list_n += __popc(m);
                __syncwarp();
            }
            const unsigned idle = __ballot_sync(WF_FULLMASK, t.state == T_IDLE);
            const int take = list_n < n_free ? list_n : n_free;
            const int my = __popc(idle & ((1u << lane) - 1u));
            if (t.state == T_IDLE && my < take) {
                const int slot = (int)list[list_n - 1 - my];
                const float4 r0 = __ldcg(W.ray + 2 * (size_t)slot), r1 = __ldcg(W.ray + 2 * (size_t)slot + 1);
                const int ph = W.phase[slot];
                t.slot = slot | (ph == SP_SHADOW_ANY ? ST_ANY_BIT : 0) | (ph != SP_NEAREST ? ST_SHADOW_BIT : 0);
                t.o = v3(r0.x, r0.y, r0.z);
                const V3 d = v3(r0.w, r1.x, r1.y);
                // BVH::intersect computes dirfrac per box (BVH.cpp:29-33); hoisted: same IEEE division, same bits
                t.inv = v3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
                const float omax = st_omax(t.o);
                t.path = t.todo = 0;
                t.depth = 0;
                t.best_hi = __int_as_float(0x7f800000);
                t.st_slot = -1;
                t.st_lo = 0.f;
                t.st_dist = -1.f;
                t.state = T_DESCEND;
                prune = t.best_hi;
                if (!(isfinite(t.inv.x) && isfinite(t.inv.y) && isfinite(t.inv.z) && isfinite(omax))) {
                    // a zero / denormal direction component or a non-finite origin: the literal box test, one ray at a time
                    t.st_slot = st_trace_slow<COUNT>(S, ph == SP_SHADOW_ANY, t.o.x, t.o.y, t.o.z, d.x, d.y, d.z, ps.tr);
                    t.state = T_DONE;
                }
            }
            __syncwarp();
            list_n -= take;
            if (take == 0 && n_step + n_leaf == 0) {
                // nothing in flight and nothing to take: the slow-path lanes of the last refill may still have to
                // publish; otherwise this warp is done
                if (__ballot_sync(WF_FULLMASK, t.state == T_DONE) == 0u) break;
            }
            continue;
        }
        if (n_leaf >= thr_leaf || n_step == 0) {
            // ---- leaf phase: lanes at depth 17 open their leaf pair and test its triangles (small pairs per
            // lane, big ones by the whole warp), until no lane of the warp is at a leaf.
            // What only this phase needs is set up here and dead outside it (the step loop then carries origin,
            // 1/direction, the path words and the best hit only): the direction comes back from the slot's ray record.
            t.lf_cur = t.lf_end = t.lf_mid = t.lf_nL = t.lf_nR = 0;
            t.lf_slack = 0.f;
            t.result = -1;
            t.any = (t.slot & ST_ANY_BIT) != 0;
            t.omax = st_omax(t.o);
            t.d = v3s(0.f);
            if (at_leaf) {
                const int slot = t.slot & ST_SLOT_MASK;
                const float4 r0 = __ldcg(W.ray + 2 * (size_t)slot), r1 = __ldcg(W.ray + 2 * (size_t)slot + 1);
                t.d = v3(r0.w, r1.x, r1.y);
            }
            for (;;) {
                if (t.state == T_DESCEND && t.depth == ST_LEAF_DEPTH) st_leaf_open<COUNT>(S, t, prune, ps.tr);
                const bool want = t.state == T_LEAF;
                const bool again = t.state == T_DESCEND && t.depth == ST_LEAF_DEPTH && at_leaf;  // popped onto the sibling pair
                const unsigned code = want ? (t.lf_end - t.lf_cur >= coop_min ? 95u - lane : 1u) : again ? 1u : 0u;
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
            if (t.any && t.result >= 0) t.st_slot = t.result;  // an any-hit query: the first candidate is the answer
            if (at_leaf) prune = t.best_hi * 1.0001f + (2e-5f * t.omax + 1e-6f);
            continue;
        }
        // ---- one record for every descending lane ---------------------------------------------------
        if (stepping) st_step<COUNT>(S, t, prune, ps.tr);
    }
    if (COUNT) er_flush_stats(P.counters, ps, 0);
}
