// er_render.cuh — the render kernels (take the place of the two parallel_for launches,
// src/kernel.cpp:665-671 and :693-699).
//
//  k_setup            setupKernel (kernel.cpp:176-213) for the pixels this shard owns
//  k_render_pixel     one thread per pixel, `spp` samples, bounce by bounce (simple form; cross-check)
//  k_render_persist   persistent threads: every lane owns one pixel at a time and runs its samples
//                     back to back (the per-pixel xorshift stream is sequential across samples,
//                     kernel.cpp:483,645, so a pixel can never be split over lanes). Lanes are
//                     refilled with the next pixel through a warp-aggregated ticket (ballot +
//                     popc + one atomicAdd per warp); nearest-hit and shadow queries of all lanes
//                     go through ONE traversal site per loop iteration so a warp keeps traversing
//                     together whatever bounce or sample each lane is in.
#pragma once
#include "er_path.cuh"

struct RenderParams {
    PathOpts opt;
    int spp;                 // samples per pixel in this launch
    unsigned work_total;     // work items (tile-ordered pixel tickets, some fall outside the image)
    unsigned* ticket;        // global ticket counter (zeroed before the launch)
    int shard_index, shard_count, rows_per_block;
    int tiles_x;             // ceil(x_res / 8)
    int owned_blocks;        // row blocks owned by this shard
    DCounters* counters;
};

// ticket -> pixel. Tickets enumerate the owned row blocks; inside a block 8x4-pixel tiles (one warp's
// worth) left to right, top to bottom. Returns false for tickets that fall outside the image.
ER_D bool er_ticket_pixel(const DScene& S, const RenderParams& P, unsigned w, int& x, int& y) {
    const unsigned per_block = (unsigned)P.tiles_x * (unsigned)(P.rows_per_block / 4) * 32u;
    const unsigned ob = w / per_block;
    const unsigned r = w - ob * per_block;
    const unsigned tile = r >> 5, lane = r & 31u;
    const unsigned tx = tile % (unsigned)P.tiles_x, ty = tile / (unsigned)P.tiles_x;
    const int block = (int)ob * P.shard_count + P.shard_index;
    x = (int)(tx * 8u + (lane & 7u));
    y = block * P.rows_per_block + (int)(ty * 4u + (lane >> 3));
    return x < S.x_res && y < S.y_res;
}

ER_D bool er_owns_row(int y, int shard_index, int shard_count, int rows_per_block) {
    return ((y / rows_per_block) % shard_count) == shard_index;
}

__global__ void k_setup(const __grid_constant__ DScene S, int shard_index, int shard_count, int rows_per_block) {
    const size_t P = (size_t)S.x_res * S.y_res;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < P; idx += (size_t)gridDim.x * blockDim.x) {
        const int y = (int)(idx / S.x_res);
        const bool own = er_owns_row(y, shard_index, shard_count, rows_per_block);
        S.rng[idx] = er_rng_seed((uint32_t)idx);
        S.samples[idx] = 1;
        // non-owned pixels hold exact zeros so that a sum over shards equals the single-GPU buffer
        const float4 v = own ? make_float4(0.f, 0.f, 0.f, 1.f) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int p = 0; p < 5; p++) S.passes[p * P + idx] = v;
    }
}

ER_D void er_flush_stats(DCounters* c, const PathStats& ps, unsigned samples) {
    // warp-reduce then one atomic per warp and counter
    unsigned long long v[7] = {ps.rays, ps.tr.box_tests, ps.tr.tri_tests, ps.tr.candidates, ps.shaded_hits, ps.nee_samples, samples};
    const unsigned m = __activemask();
#pragma unroll
    for (int k = 0; k < 7; k++) {
        unsigned long long x = v[k];
        for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(m, x, o);
        v[k] = x;
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&c->rays, v[0]);
        atomicAdd(&c->box_tests, v[1]);
        atomicAdd(&c->tri_tests, v[2]);
        atomicAdd(&c->candidates, v[3]);
        atomicAdd(&c->shaded_hits, v[4]);
        atomicAdd(&c->nee_samples, v[5]);
        atomicAdd(&c->samples, v[6]);
    }
}

template <bool COUNT, bool ORDERED>
__global__ void __launch_bounds__(128) k_render_pixel(const __grid_constant__ DScene S, const __grid_constant__ RenderParams P) {
    PathStats ps = {};
    unsigned done = 0;
    const unsigned w = blockIdx.x * blockDim.x + threadIdx.x;
    int x, y;
    if (w < P.work_total && er_ticket_pixel(S, P, w, x, y)) {
        const size_t idx = (size_t)y * S.x_res + x;
        const size_t NP = (size_t)S.x_res * S.y_res;
        uint32_t rng = S.rng[idx];
        uint32_t sa = S.samples[idx];
        for (int s = 0; s < P.spp; s++) {
            V3 light, n, t, b;
            er_trace_path<COUNT, ORDERED>(S, P.opt, x, y, rng, light, n, t, b, ps);
            sa = er_accumulate(S.passes, NP, idx, sa, light, n, t, b);
            done++;
        }
        S.rng[idx] = rng;
        S.samples[idx] = sa;
    }
    if (COUNT) {
        __syncwarp();
        er_flush_stats(P.counters, ps, done);
    }
}

enum { PH_FETCH = 0, PH_CAMERA = 1, PH_NEAREST = 2, PH_SHADOW = 3 };

#define ER_PERSIST_THREADS 128

template <bool COUNT>
__global__ void __launch_bounds__(ER_PERSIST_THREADS) k_render_persist(const __grid_constant__ DScene S,
                                                                        const __grid_constant__ RenderParams P) {
    const unsigned FULL = 0xffffffffu;
    const unsigned lane = threadIdx.x & 31u;
    const size_t NP = (size_t)S.x_res * S.y_res;
    // first-hit AOVs of the sample in flight: parked in shared memory, touched twice per sample
    __shared__ float s_aov[9][ER_PERSIST_THREADS];

    PathStats ps = {};
    unsigned done = 0;

    bool alive = true;
    int phase = PH_FETCH;
    int x = 0, y = 0, s_left = 0, bounce = 0;
    size_t idx = 0;
    uint32_t rng = 0, sa = 0;
    V3 ro = v3s(0.f), rd = v3s(0.f);
    V3 light = v3s(0.f), reduction = v3s(1.f);
    // bounce state parked across the shadow query
    V3 add_open = v3s(0.f), add_blocked = v3s(0.f), factor = v3s(0.f), next_o = v3s(0.f), next_d = v3s(0.f);
    int own_slot = -1;
    bool shadow_any = false;

    for (;;) {
        // ---- refill idle lanes: warp-aggregated ticket ------------------------------------------
        const bool need = alive && phase == PH_FETCH;
        const unsigned m = __ballot_sync(FULL, need);
        if (m) {
            const int leader = __ffs((int)m) - 1;
            unsigned base = 0;
            if ((int)lane == leader) base = atomicAdd(P.ticket, (unsigned)__popc(m));
            base = __shfl_sync(FULL, base, leader);
            if (need) {
                const unsigned w = base + (unsigned)__popc(m & ((1u << lane) - 1u));
                if (w >= P.work_total) {
                    alive = false;
                } else if (er_ticket_pixel(S, P, w, x, y)) {
                    idx = (size_t)y * S.x_res + x;
                    rng = S.rng[idx];
                    sa = S.samples[idx];
                    s_left = P.spp;
                    phase = PH_CAMERA;
                }
            }
        }
        if (!__any_sync(FULL, alive)) break;

        // ---- new sample: camera ray (kernel.cpp:492-505) -------------------------------------------
        if (alive && phase == PH_CAMERA) {
            float c0 = er_rng_next(rng), c1 = er_rng_next(rng), c2 = er_rng_next(rng), c3 = er_rng_next(rng),
                  c4 = er_rng_next(rng);
            if (P.opt.rtl)
                er_camera_ray(S, x, y, c4, c3, c2, c1, c0, ro, rd);
            else
                er_camera_ray(S, x, y, c0, c1, c2, c3, c4, ro, rd);
            light = v3s(0.f);
            reduction = v3s(1.f);
#pragma unroll
            for (int k = 0; k < 9; k++) s_aov[k][threadIdx.x] = 0.f;
            bounce = 0;
            phase = PH_NEAREST;
        }

        // ---- the one traversal site ---------------------------------------------------------------
        const bool tracing = alive && (phase == PH_NEAREST || phase == PH_SHADOW);
        HitRec rec;
        rec.slot = -1;
        rec.dist = 0.f;
        rec.rank = 0;
        rec.u = rec.v = rec.t = 0.f;
        if (tracing) {
            er_traverse<COUNT, true>(S, phase == PH_SHADOW && shadow_any, ro, rd, rec, ps.tr);
            if (COUNT) ps.rays++;
        }

        // ---- consume the answer -------------------------------------------------------------------
        bool sample_done = false;
        if (tracing && phase == PH_SHADOW) {
            const bool blocked = rec.slot >= 0 && (shadow_any || rec.slot != own_slot);
            light = light + reduction * (blocked ? add_blocked : add_open);
            reduction = reduction * factor;
            ro = next_o;
            rd = next_d;
            bounce++;
            phase = PH_NEAREST;
            if (bounce >= P.opt.max_depth) sample_done = true;
        } else if (tracing) {
            if (rec.slot < 0) {
                float u, v;
                er_spherical_mapping(-1.0f * rd, u, v);
                light = light + reduction * er_tex_filtered(S.hdri, u, v);
                sample_done = true;
            } else {
                Bounce b;
                er_shade_hit(S, P.opt, rec, ro, rd, rng, b);
                if (COUNT) ps.shaded_hits++;
                if (b.opaque) {
                    if (COUNT) ps.nee_samples++;
                    if (bounce == 0) {
                        s_aov[0][threadIdx.x] = b.normal.x, s_aov[1][threadIdx.x] = b.normal.y, s_aov[2][threadIdx.x] = b.normal.z;
                        s_aov[3][threadIdx.x] = b.tangent.x, s_aov[4][threadIdx.x] = b.tangent.y, s_aov[5][threadIdx.x] = b.tangent.z;
                        s_aov[6][threadIdx.x] = b.bitangent.x, s_aov[7][threadIdx.x] = b.bitangent.y, s_aov[8][threadIdx.x] = b.bitangent.z;
                    }
                    if (b.need_shadow) {
                        add_open = b.add_open;
                        add_blocked = b.add_blocked;
                        factor = b.factor;
                        next_o = b.next_o;
                        next_d = b.next_d;
                        own_slot = b.own_slot;
                        shadow_any = b.shadow_any;
                        ro = b.so;
                        rd = b.sd;
                        phase = PH_SHADOW;
                    } else {
                        light = light + reduction * b.add_open;
                        reduction = reduction * b.factor;
                        ro = b.next_o;
                        rd = b.next_d;
                        bounce++;
                        if (bounce >= P.opt.max_depth) sample_done = true;
                    }
                } else {
                    ro = b.next_o;
                    rd = b.next_d;
                    bounce++;
                    if (bounce >= P.opt.max_depth) sample_done = true;
                }
            }
        }

        // ---- sample end: clamp + running mean (kernel.cpp:597-645) ----------------------------------
        if (sample_done) {
            light = er_clamp(light, 0.f, 10.f);
            const V3 n = v3(s_aov[0][threadIdx.x], s_aov[1][threadIdx.x], s_aov[2][threadIdx.x]);
            const V3 t = v3(s_aov[3][threadIdx.x], s_aov[4][threadIdx.x], s_aov[5][threadIdx.x]);
            const V3 bt = v3(s_aov[6][threadIdx.x], s_aov[7][threadIdx.x], s_aov[8][threadIdx.x]);
            sa = er_accumulate(S.passes, NP, idx, sa, light, n, t, bt);
            done++;
            if (--s_left > 0) {
                phase = PH_CAMERA;
            } else {
                S.rng[idx] = rng;
                S.samples[idx] = sa;
                phase = PH_FETCH;
            }
        }
    }
    if (COUNT) er_flush_stats(P.counters, ps, done);
}

// nearest-hit queries for caller-supplied rays (throwRay, kernel.cpp:218-240)
struct DHit {  // == ErHit
    float position[3], normal[3], tangent[3], bitangent[3], gnormal[3];
    int valid;
    uint32_t objectID, material_id;
    float tu, tv;
    int triIdx;
};

template <bool COUNT, bool ORDERED>
__global__ void __launch_bounds__(128) k_trace(const __grid_constant__ DScene S, const float* __restrict__ rays, size_t n,
                                               DHit* __restrict__ hits, DCounters* counters) {
    PathStats ps = {};
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float* r = rays + 6 * i;
        V3 o = v3(r[0], r[1], r[2]);
        V3 d = er_ray_dir(v3(r[3], r[4], r[5]));
        HitRec rec;
        er_nearest<COUNT, ORDERED>(S, o, d, rec, ps.tr);
        if (COUNT) ps.rays++;
        HitFull h;
        er_make_hit(S, rec, o, d, h);
        DHit out;
        out.position[0] = h.position.x, out.position[1] = h.position.y, out.position[2] = h.position.z;
        out.normal[0] = h.normal.x, out.normal[1] = h.normal.y, out.normal[2] = h.normal.z;
        out.tangent[0] = h.tangent.x, out.tangent[1] = h.tangent.y, out.tangent[2] = h.tangent.z;
        out.bitangent[0] = h.bitangent.x, out.bitangent[1] = h.bitangent.y, out.bitangent[2] = h.bitangent.z;
        out.gnormal[0] = h.gnormal.x, out.gnormal[1] = h.gnormal.y, out.gnormal[2] = h.gnormal.z;
        out.valid = h.valid ? 1 : 0;
        out.objectID = h.objectID;
        out.material_id = h.material_id;
        out.tu = h.tu;
        out.tv = h.tv;
        out.triIdx = h.triIdx;
        hits[i] = out;
    }
    if (COUNT) {
        __syncwarp();
        er_flush_stats(counters, ps, 0);
    }
}
