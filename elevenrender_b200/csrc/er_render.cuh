// er_render.cuh — the render kernels (take the place of the two parallel_for launches,
// src/kernel.cpp:665-671 and :693-699).
//
//  k_setup            setupKernel (kernel.cpp:176-213) for the pixels this shard owns
//  k_render_pixel     one thread per pixel, `spp` samples, bounce by bounce (simple form; cross-check)
//  (the production kernel, k_render_wavefront, is in er_wavefront.cuh)
#pragma once
#include "er_path.cuh"

struct RenderParams {
    PathOpts opt;
    int spp;                 // samples per pixel in this launch
    // CHUNKED k_render_wavefront launches (er_wavefront.cuh): a pixel is handed out in chunk_count chunks of
    // chunk_spp samples; tickets >= work_total are served by the ready ring; pix[].z == progress_base when
    // the launch starts
    int chunk_spp;
    unsigned chunk_count, progress_base;
    // chunk sizes shrink towards the end of a pixel (the launch ends when the slowest pixels finish their LAST chunk):
    // the first chunk_n1 chunks hold chunk_spp samples, the next chunk_n2 hold chunk_k2, the rest chunk_k3
    unsigned chunk_n1, chunk_n2;
    int chunk_k2, chunk_k3;
    unsigned ticket_total;   // work_total + owned pixels * (chunk_count - 1)
    unsigned* ring;          // ready ring: (lap tag << 24) | pixel index
    unsigned* ring_tail;
    unsigned ring_mask, ring_shift;
    unsigned work_total;     // pixel tickets (tile-ordered, some fall outside the image)
    unsigned* ticket;        // global ticket counter (zeroed before the launch)
    // pixel ownership (er_owns_row): row blocks first_block, first_block + block_step, ... below row y_limit.
    // Interleaved shards: first_block = shard index, block_step = shard count, y_limit = y_res; a contiguous
    // stripe [y0, y1): first_block = y0 / rows_per_block, block_step = 1, y_limit = y1.
    int first_block, block_step, rows_per_block, y_limit;
    // optional (er_set_block_order): the i-th owned row block to hand out is owned block block_order[i] — the costly
    // blocks first, so that the pixels with the longest sample chains start at once (a launch ends with its slowest pixel)
    const int* block_order;
    int tiles_x;             // ceil(x_res / 8)
    int owned_blocks;        // row blocks owned by this shard
    unsigned* row_cost;      // COUNT build only, may be null: per image row, traversal work of the rays of its pixels
    DCounters* counters;
    // k_render_wavefront scheduling thresholds (er_wavefront.cuh)
    int thr_descend;         // keep descending while at least this many lanes of the warp want to
    int thr_switch;          // a tracing warp with fewer active lanes and an empty ray queue goes shading
    int thr_batch;           // a shading queue is drained once it holds this many entries (>= 32)
    int coop_min;            // a parked leaf pair with at least this many triangles left is tested by the whole warp
    // low-occupancy behaviour (the drain of a launch, small launches; profiles/README.md round 2): the batch threshold
    // follows the live paths of the CTA, thr = clamp(live - batch_reserve, batch_min, thr_batch), and warps beyond
    // live / park_lanes (at least 2 stay) stop taking work and sleep park_ns at a time
    int batch_reserve, batch_min, park_lanes, park_ns;
    // debugging aid (ER_WF_TIMELINE=<file>, else null): every CTA writes, once per 2^timeline_shift cycles, a record
    // (live paths, QT, QH, QE sizes, global ticket counter, 0, 0, 0) at timeline[(cta * 512 + k) * 8]
    unsigned* timeline;
    int timeline_shift;
    unsigned* pixel_done;    // with the timeline: per pixel, %globaltimer >> 10 (~us) when its share of the launch was complete
    int slot_limit;          // path slots a CTA uses, 32..WF_SLOTS (ER_WF_SLOT_LIMIT: occupancy experiments; default all)
    int prefetch_min;        // 0 = off; a leaf pair with at least this many triangles is prefetched into L2 when its lane parks
    float4* scratch;         // k_render_wavefront: per-CTA parked path state (er_wavefront.cuh), grid * WF_SCRATCH_BYTES_PER_CTA
    int* watchdog;           // host-mapped: [0] trip count, [1] reporting CTA + 1, [16 + 8 * warp ..] its scheduler state
    int* watchdog_flag;      // device word: non-zero once any warp of the launch has given up
    long long watchdog_cycles;
};

// ticket -> pixel. Tickets enumerate the owned row blocks; inside a block 8x4-pixel tiles (one warp's
// worth) column by column (top to bottom, then to the right), so that a run of consecutive tickets is a
// compact region: 256 tickets = 32x8 pixels with 8-row blocks. Returns false for tickets that fall
// outside the image.
ER_D bool er_ticket_pixel(const DScene& S, const RenderParams& P, unsigned w, int& x, int& y) {
    const unsigned tiles_y = (unsigned)(P.rows_per_block / 4);
    const unsigned per_block = (unsigned)P.tiles_x * tiles_y * 32u;
    unsigned ob = w / per_block;
    const unsigned r = w - ob * per_block;
    if (P.block_order != nullptr) ob = (unsigned)__ldg(P.block_order + ob);
    const unsigned tile = r >> 5, lane = r & 31u;
    const unsigned tx = tile / tiles_y, ty = tile - tx * tiles_y;
    const int block = (int)ob * P.block_step + P.first_block;
    x = (int)(tx * 8u + (lane & 7u));
    y = block * P.rows_per_block + (int)(ty * 4u + (lane >> 3));
    return x < S.x_res && y < P.y_limit;
}

ER_HD bool er_owns_row(int y, int first_block, int block_step, int rows_per_block, int y_limit) {
    const int b = y / rows_per_block;
    return y < y_limit && b >= first_block && ((b - first_block) % block_step) == 0;
}

__global__ void k_setup(const __grid_constant__ DScene S, int first_block, int block_step, int rows_per_block, int y_limit) {
    const size_t P = (size_t)S.x_res * S.y_res;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < P; idx += (size_t)gridDim.x * blockDim.x) {
        const int y = (int)(idx / S.x_res);
        const bool own = er_owns_row(y, first_block, block_step, rows_per_block, y_limit);
        S.pix[idx] = make_uint4(er_rng_seed((uint32_t)idx), 1u, 0u, 0u);
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
        const uint4 px = S.pix[idx];
        uint32_t rng = px.x;
        uint32_t sa = px.y;
        for (int s = 0; s < P.spp; s++) {
            V3 light, n, t, b;
            er_trace_path<COUNT, ORDERED>(S, P.opt, x, y, rng, light, n, t, b, ps);
            sa = er_accumulate(S.passes, NP, idx, sa, light, n, t, b);
            done++;
        }
        S.pix[idx] = make_uint4(rng, sa, px.z, 0u);
    }
    if (COUNT) {
        __syncwarp();
        er_flush_stats(P.counters, ps, done);
    }
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
