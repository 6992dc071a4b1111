// er_api.cu — the C ABI of include/er_cuda.h: context, scene upload (copy_scene's replacement),
// kernel launches (renderSetup / kernel_render_enqueue's replacement), readout and test hooks.
// sm_100a only; there is no CPU fallback anywhere in this file.
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <mutex>
#include <string>
#include <chrono>
#include <thread>
#include <vector>

#include <nvtx3/nvToolsExt.h>  // header-only NVTX 3: ranges show up in nsys / ncu timelines, no-ops when no tool is attached

#include "../../include/er_cuda.h"
#include "er_host.h"
#include "er_wavefront.cuh"
#include "er_staged.cuh"

static_assert(sizeof(DHit) == sizeof(ErHit), "ErHit layout");
static_assert(sizeof(HitData) == 37 * 4, "HitData words");
static_assert(sizeof(ErTri) == 156 && sizeof(ErNode) == 44 && sizeof(ErCamera) == 48, "record layouts");
static_assert(sizeof(DCounters) == sizeof(ErCounters), "counters");

namespace {

thread_local std::string g_err;

int fail(int code, const std::string& msg) {
    g_err = msg;
    return code;
}

#define ER_CK(call)                                                                                   \
    do {                                                                                              \
        cudaError_t e__ = (call);                                                                     \
        if (e__ != cudaSuccess) {                                                                     \
            char buf__[512];                                                                          \
            snprintf(buf__, sizeof buf__, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
            return fail(e__ == cudaErrorMemoryAllocation ? ER_ERR_OOM : ER_ERR_CUDA, buf__);          \
        }                                                                                             \
    } while (0)

}  // namespace

namespace erh_api {
int fail_invalid(const char* msg) { return fail(ER_ERR_INVALID, msg); }
}  // namespace erh_api

struct ErCtx {
    int device = 0;
    int sm_count = 0;
    cudaStream_t own_stream = nullptr;
    cudaStream_t stream = nullptr;      // launch stream (own or caller's)
    cudaStream_t read_stream = nullptr;  // tiny concurrent readbacks (get_render_info)
    DScene S{};
    bool uploaded = false;
    std::vector<std::pair<void*, size_t>> scene_allocs, alloc_pool;
    erh::FlatScene flat;
    std::vector<std::pair<void*, size_t>> pinned;  // page-locked staging ranges of `flat` (stage_flat)
    size_t pairs_bytes = 0;
    bool l2_window = false;
    std::vector<DTexture> h_textures;
    // pixel ownership: row blocks first_block, first_block + block_step, ... below row y_limit (0 = y_res);
    // er_set_shard installs interleaved blocks, er_set_shard_stripe a contiguous stripe of rows
    int first_block = 0, block_step = 1, rows_per_block = 8, y_limit = 0;
    unsigned* d_row_cost = nullptr;  // er_row_cost: per-row traversal work of a 1-spp COUNT launch
    size_t row_cost_rows = 0;
    int* d_block_order = nullptr;  // er_set_block_order
    float* h_slack = nullptr;      // page-locked: per-slot slack back from k_flatten_tris
    size_t h_slack_n = 0;
    int* d_flag = nullptr;         // k_flatten_tris: bit 0 bad index, bit 1 bad material
    int block_order_n = 0, block_order_cap = 0;
    unsigned* d_ticket = nullptr;
    DCounters* d_counters = nullptr;
    int* h_watchdog = nullptr;  // host-mapped watchdog record of k_render_wavefront (RenderParams::watchdog)
    int* d_watchdog = nullptr;
    int* d_wdflag = nullptr;
    unsigned* d_ring = nullptr;   // ready ring of CHUNKED launches
    size_t ring_bytes = 0;
    unsigned progress_base = 0;   // value of S.pix[].z of every owned pixel between launches
    float4* d_scratch = nullptr;  // k_render_wavefront's parked path state, one record block per CTA
    size_t scratch_bytes = 0;
    void* d_staged = nullptr;     // the staged path's per-slot state (er_staged.cuh), one allocation
    size_t staged_bytes = 0;
    int* h_live = nullptr;        // host-mapped: (live slots, launch id) of the staged launch the GPU is working on
    int staged_launch_id = 0;
    // kernel timing: event pairs recorded around render launches
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> ev_pool, ev_live;
    float kernel_ms = 0.f;
    int launches = 0;
    // render thread (RenderingManager::start_rendering's t_rend, Managers.cpp:271 — joined here)
    std::thread worker;
    std::mutex batch_mutex;  // held for one launch+sync; readout takes it to see whole batches only
    std::atomic<bool> abort{false};
    std::atomic<bool> running{false};
    std::atomic<unsigned> completed{0};  // samples finished since the last setup
    std::atomic<int> worker_status{0};
    std::string worker_error;
};

namespace {

// NVTX range over the calls a profile of the host side wants to see (tracing)
struct NvtxRange {
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
    NvtxRange(const NvtxRange&) = delete;
    NvtxRange& operator=(const NvtxRange&) = delete;
};

int use(ErCtx* c) {
    if (!c) return fail(ER_ERR_INVALID, "null context");
    ER_CK(cudaSetDevice(c->device));
    return ER_OK;
}

// Device buffers of a replaced scene are kept in a pool until the next upload has taken what it can reuse
// (a re-upload of a scene of the same shape — new camera, new materials, the e2e bench — allocates nothing).
void free_scene(ErCtx* c) {
    for (auto& a : c->scene_allocs) c->alloc_pool.push_back(a);
    c->scene_allocs.clear();
    c->h_textures.clear();
    c->S = DScene{};
    c->uploaded = false;
}

int dev_alloc(ErCtx* c, size_t bytes, void** out) {
    for (size_t i = 0; i < c->alloc_pool.size(); i++)
        if (c->alloc_pool[i].second == bytes) {
            *out = c->alloc_pool[i].first;
            c->scene_allocs.push_back(c->alloc_pool[i]);
            c->alloc_pool.erase(c->alloc_pool.begin() + (long)i);
            return ER_OK;
        }
    void* p = nullptr;
    ER_CK(cudaMalloc(&p, bytes));
    c->scene_allocs.push_back({p, bytes});
    *out = p;
    return ER_OK;
}

void drain_pool(ErCtx* c) {
    for (auto& a : c->alloc_pool) cudaFree(a.first);
    c->alloc_pool.clear();
}

template <typename T>
int dev_upload(ErCtx* c, const T* host, size_t n, const T** out) {
    void* p = nullptr;
    const size_t bytes = sizeof(T) * (n ? n : 1);
    int rc = dev_alloc(c, bytes, &p);
    if (rc) return rc;
    if (n) ER_CK(cudaMemcpyAsync(p, host, sizeof(T) * n, cudaMemcpyHostToDevice, c->stream));
    *out = (const T*)p;
    return ER_OK;
}

int upload_texture(ErCtx* c, const ErTexture& t, DTexture& d, const char* what) {
    if (t.width <= 0 || t.height <= 0 || t.channels < 0 || (!t.data && t.channels > 0))
        return fail(ER_ERR_INVALID, std::string("bad texture record: ") + what);
    d.width = t.width;
    d.height = t.height;
    d.channels = t.channels;
    d.filter = t.filter;
    const float* p = nullptr;
    int rc = dev_upload<float>(c, t.data, (size_t)t.width * t.height * t.channels, &p);
    d.data = p;
    return rc;
}

DCamera make_camera(const ErCamera& cam) {
    DCamera d{};
    d.focalLength = cam.focalLength;
    d.sensorWidth = cam.sensorWidth;
    d.sensorHeight = cam.sensorHeight;
    d.aperture = cam.aperture;
    d.focusDistance = cam.focusDistance;
    d.px = cam.position[0];
    d.py = cam.position[1];
    d.pz = cam.position[2];
    // kernel.cpp:412-414: rotation *= (PIF / 180.0f); the sin/cos of the three angles are the same for
    // every pixel and sample, so they are evaluated once here
    const float k = ER_PIF / 180.0f;
    const float rx = cam.rotation[0] * k, ry = cam.rotation[1] * k, rz = cam.rotation[2] * k;
    d.cx = cosf(rx), d.sx = sinf(rx);
    d.cy = cosf(ry), d.sy = sinf(ry);
    d.cz = cosf(rz), d.sz = sinf(rz);
    d.bokeh = cam.bokeh ? 1 : 0;
    return d;
}

// A render launch that exceeded its time limit gave up (er_wavefront.cuh): report it as an error.
int check_watchdog(ErCtx* c) {
    if (!c->h_watchdog || ((volatile int*)c->h_watchdog)[0] == 0) return ER_OK;
    volatile int* w = c->h_watchdog;
    std::string msg = "render kernel watchdog: check code " + std::to_string(w[2]) + " value " + std::to_string(w[3]) + " CTA " + std::to_string(w[4]) +
                      "; launch exceeded its time limit or failed a check; CTA " + std::to_string(w[1] - 1) + " warps:";
    for (int i = 0; i < WF_THREADS / 32; i++) {
        char buf[200];
        snprintf(buf, sizeof buf, " [w%d act=%08x qt=%d qh=%d qe=%d qw=%d live=%d iter|pushed=%d ticket=%d st=%x]", i, (unsigned)w[16 + 8 * i],
                 w[17 + 8 * i], w[18 + 8 * i], w[19 + 8 * i], w[23 + 8 * i] >> 8, w[20 + 8 * i], w[21 + 8 * i], w[22 + 8 * i], (unsigned)w[23 + 8 * i] & 0xffu);
        msg += buf;
    }
    if (w[5]) {
        char buf[200];
        snprintf(buf, sizeof buf, " QUEUE SPIN LIMIT kind=%d (1 push, 2 claim) idx=%d head=%d tail=%d entry=%x cta=%d thread=%d flags=%x", w[5], w[6], w[7], w[8], (unsigned)w[9], w[10], w[11], (unsigned)w[0]);
        msg += buf;
    }
    fprintf(stderr, "%s\n", msg.c_str());
    if (const char* dump = getenv("ER_WF_DEBUG_DUMP")) {
        // debugging aid for CHUNKED launches: ticket counter, ring tail, the ready ring and the pixel records
        if (c->d_ring && c->S.pix) {
            const size_t np = (size_t)c->S.x_res * c->S.y_res;
            std::vector<unsigned> ring(c->ring_bytes / 4), pix(np * 4);
            unsigned hdr[4] = {0, 0, (unsigned)ring.size(), (unsigned)np};
            cudaMemcpy(hdr, c->d_ticket, 8, cudaMemcpyDeviceToHost);
            cudaMemcpy(ring.data(), c->d_ring, c->ring_bytes, cudaMemcpyDeviceToHost);
            cudaMemcpy(pix.data(), c->S.pix, np * 16, cudaMemcpyDeviceToHost);
            if (FILE* f = fopen(dump, "wb")) {
                fwrite(hdr, 4, 4, f);
                fwrite(ring.data(), 4, ring.size(), f);
                fwrite(pix.data(), 4, pix.size(), f);
                fclose(f);
            }
        }
    }
    memset(c->h_watchdog, 0, 4096);
    return fail(ER_ERR_CUDA, msg);
}

int grid_for(ErCtx* c, size_t n, int block) {
    size_t g = (n + block - 1) / block;
    size_t cap = (size_t)c->sm_count * 16;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return (int)g;
}

int y_lim(const ErCtx* c) { return c->y_limit > 0 && c->y_limit < c->S.y_res ? c->y_limit : c->S.y_res; }

// image rows [y0, y1) of every owned row block, in order
template <typename F>
void for_owned_ranges(const ErCtx* c, F f) {
    const int yl = y_lim(c), rpb = c->rows_per_block;
    for (long long b = c->first_block; b * rpb < yl; b += c->block_step) {
        const int y0 = (int)(b * rpb), y1 = y0 + rpb < yl ? y0 + rpb : yl;
        f(y0, y1);
    }
}

int launch_setup(ErCtx* c) {
    const size_t P = (size_t)c->S.x_res * c->S.y_res;
    k_setup<<<grid_for(c, P, 256), 256, 0, c->stream>>>(c->S, c->first_block, c->block_step, c->rows_per_block, y_lim(c));
    ER_CK(cudaGetLastError());
    c->completed = 0;
    c->progress_base = 0;
    return ER_OK;
}

int fold_events(ErCtx* c) {
    for (auto& pr : c->ev_live) {
        ER_CK(cudaEventSynchronize(pr.second));
        float ms = 0.f;
        ER_CK(cudaEventElapsedTime(&ms, pr.first, pr.second));
        c->kernel_ms += ms;
        c->ev_pool.push_back(pr);
    }
    c->ev_live.clear();
    return ER_OK;
}

int get_event_pair(ErCtx* c, std::pair<cudaEvent_t, cudaEvent_t>& pr) {
    if (c->ev_live.size() >= 256) {
        int rc = fold_events(c);
        if (rc) return rc;
    }
    if (!c->ev_pool.empty()) {
        pr = c->ev_pool.back();
        c->ev_pool.pop_back();
        return ER_OK;
    }
    ER_CK(cudaEventCreate(&pr.first));
    ER_CK(cudaEventCreate(&pr.second));
    return ER_OK;
}

ErRenderOpts norm_opts(const ErRenderOpts* o) {
    ErRenderOpts r{};
    if (o) r = *o;
    if (r.max_depth <= 0) r.max_depth = 5;
    if (r.spp_per_launch <= 0) r.spp_per_launch = 16;
    return r;
}

// The staged path (er_staged.cuh): per-slot state for every pixel ticket, then waves of logic / trace /
// shade kernels. A sample takes at most 2 * max_depth waves (one per ray), so the worst case is known; the
// kernels of a wave that finds no slot alive return at once, and the host stops enqueueing when the
// host-mapped live count (written by the GPU every wave) reads zero. No host synchronisation.
__global__ void k_st_publish_live(StagedState W, int* live_out, int launch_id) {
    ((volatile int*)live_out)[0] = (int)W.ctr[SC_LIVE];
    __threadfence_system();
    ((volatile int*)live_out)[1] = launch_id;  // launches are enqueued without host sync: an older launch's count must not end this one
}

int launch_staged(ErCtx* c, RenderParams& P, int spp, bool count) {
    if (spp > 0xffff) return fail(ER_ERR_INVALID, "spp_per_launch must be <= 65535");
    const size_t M = ((size_t)P.work_total + 255) / 256 * 256;
    // ray 32 + lr 32 + park 48 + aov 48 + res 4 + phase 1 (padded to 4) bytes per slot, counters behind
    const size_t bytes = M * (32 + 32 + 48 + 48 + 4 + 4) + 256;
    if (c->staged_bytes < bytes) {
        if (c->d_staged) ER_CK(cudaFree(c->d_staged));
        c->d_staged = nullptr;
        c->staged_bytes = 0;
        ER_CK(cudaMalloc(&c->d_staged, bytes));
        c->staged_bytes = bytes;
    }
    if (!c->h_live) {
        ER_CK(cudaHostAlloc((void**)&c->h_live, 64, cudaHostAllocMapped));
        memset(c->h_live, 0, 64);
    }
    int* d_live = nullptr;
    ER_CK(cudaHostGetDevicePointer((void**)&d_live, c->h_live, 0));
    StagedState W{};
    char* base = (char*)c->d_staged;
    W.ray = (float4*)base;
    W.lr = (float4*)(base + M * 32);
    W.park = (float4*)(base + M * 64);
    W.aov = (float4*)(base + M * 112);
    W.res = (int*)(base + M * 160);
    W.phase = (uint8_t*)(base + M * 164);
    W.ctr = (unsigned*)(base + M * 168);
    W.slots = (unsigned)M;
    const char* e1 = getenv("ER_WF_THR_DESCEND");
    const char* e5 = getenv("ER_WF_COOP_MIN");
    const char* e2 = getenv("ER_ST_THR_REFILL");
    e1 = getenv("ER_ST_THR_LEAF");
    P.thr_descend = e1 ? atoi(e1) : 12;  // k_st_trace: leaf phase once this many lanes wait at a leaf pair
    if (P.thr_descend < 1) P.thr_descend = 1;
    P.thr_switch = e2 ? atoi(e2) : 12;   // k_st_trace: refill once this many lanes are free
    if (P.thr_switch < 1) P.thr_switch = 1;
    P.coop_min = e5 ? atoi(e5) : 8;
    if (P.coop_min < 1) P.coop_min = 1;
    ER_CK(cudaMemsetAsync(W.ctr, 0, 64, c->stream));
    const int launch_id = ++c->staged_launch_id;
    k_st_init<<<(unsigned)(M / 256), 256, 0, c->stream>>>(c->S, P, W);
    int per_t = 0, per_s = 0, per_l = 0;
    if (count) {
        ER_CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_t, k_st_trace<true>, ST_TRACE_THREADS, 0));
        ER_CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_s, k_st_shade<true>, ST_BLOCK, 0));
        ER_CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_l, k_st_logic<true>, ST_BLOCK, 0));
    } else {
        ER_CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_t, k_st_trace<false>, ST_TRACE_THREADS, 0));
        ER_CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_s, k_st_shade<false>, ST_BLOCK, 0));
        ER_CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_l, k_st_logic<false>, ST_BLOCK, 0));
    }
    auto grid_of = [&](int per_sm, size_t units) {
        size_t g = (size_t)c->sm_count * (size_t)(per_sm < 1 ? 1 : per_sm);
        if (g > units) g = units;
        return (unsigned)(g < 1 ? 1 : g);
    };
    const unsigned g_t = grid_of(per_t, (M / ST_CHUNK + ST_TRACE_WARPS - 1) / ST_TRACE_WARPS);
    const unsigned g_s = grid_of(per_s, M / ST_BLOCK), g_l = grid_of(per_l, M / ST_BLOCK);
    const long waves = (long)spp * 2 * P.opt.max_depth + 2;
    for (long w = 0; w < waves; w++) {
        // the GPU has finished this launch already (it publishes (live, launch id) every 8 waves)
        if (w > 8 && ((volatile int*)c->h_live)[1] == launch_id && ((volatile int*)c->h_live)[0] == 0) break;
        if (count) {
            k_st_logic<true><<<g_l, ST_BLOCK, 0, c->stream>>>(c->S, P, W);
            k_st_trace<true><<<g_t, ST_TRACE_THREADS, 0, c->stream>>>(c->S, P, W);
            k_st_shade<true><<<g_s, ST_BLOCK, 0, c->stream>>>(c->S, P, W);
        } else {
            k_st_logic<false><<<g_l, ST_BLOCK, 0, c->stream>>>(c->S, P, W);
            k_st_trace<false><<<g_t, ST_TRACE_THREADS, 0, c->stream>>>(c->S, P, W);
            k_st_shade<false><<<g_s, ST_BLOCK, 0, c->stream>>>(c->S, P, W);
        }
        if ((w & 7) == 7) k_st_publish_live<<<1, 1, 0, c->stream>>>(W, d_live, launch_id);
    }
    ER_CK(cudaGetLastError());
    return ER_OK;
}

// one kernel launch tracing `spp` samples for every owned pixel; no host sync
int launch_render(ErCtx* c, int spp, const ErRenderOpts& o) {
    NvtxRange nvtx_("er_cuda: render launch");
    if (!c->uploaded) return fail(ER_ERR_INVALID, "no scene uploaded");
    RenderParams P{};
    P.opt.max_depth = o.max_depth;
    P.opt.rtl = o.rng_order == ER_RNG_RTL ? 1 : 0;
    P.opt.ext = o.extensions;
    P.spp = spp;
    P.first_block = c->first_block;
    P.block_step = c->block_step;
    P.rows_per_block = c->rows_per_block;
    P.y_limit = y_lim(c);
    P.tiles_x = (c->S.x_res + 7) / 8;
    const int blocks_total = (P.y_limit + c->rows_per_block - 1) / c->rows_per_block;
    P.owned_blocks = blocks_total > c->first_block ? (blocks_total - c->first_block + c->block_step - 1) / c->block_step : 0;
    P.row_cost = o.count_stats == 2 ? c->d_row_cost : nullptr;
    P.block_order = c->block_order_n > 0 && c->block_order_n == P.owned_blocks ? c->d_block_order : nullptr;
    const unsigned long long work =
        (unsigned long long)P.owned_blocks * (unsigned long long)P.tiles_x * (unsigned long long)(c->rows_per_block / 4) * 32ull;
    if (work > 0xf0000000ull) return fail(ER_ERR_INVALID, "image too large for 32-bit tickets");  // headroom: every CTA overshoots by up to one chunk
    P.work_total = (unsigned)work;
    P.ticket = c->d_ticket;
    P.counters = c->d_counters;
    if (P.work_total == 0) return ER_OK;
    if (o.kernel == 0 && !o.extensions && spp > 0xffff) return fail(ER_ERR_INVALID, "spp_per_launch must be <= 65535");
    std::pair<cudaEvent_t, cudaEvent_t> ev;
    int rc = get_event_pair(c, ev);
    if (rc) return rc;
    ER_CK(cudaEventRecord(ev.first, c->stream));
    const bool count = o.count_stats != 0;
    const int kernel = o.extensions ? 1 : o.kernel;  // the extensions live in the one-thread-per-pixel kernel only
    if (kernel == 0) {
        ER_CK(cudaMemsetAsync(c->d_ticket, 0, 2 * sizeof(unsigned), c->stream));  // ticket counter, ring tail
        int per_sm = 0;
        if (count)
            ER_CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_render_wavefront<true, true>, WF_THREADS, WF_SMEM_BYTES));
        else
            ER_CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_render_wavefront<false, true>, WF_THREADS, WF_SMEM_BYTES));
        if (per_sm < 1) per_sm = 1;
        // persistent: one CTA per residency slot; fewer when the frame has less work than path slots
        unsigned grid = (unsigned)(c->sm_count * per_sm);
        const char* e7 = getenv("ER_WF_SLOT_LIMIT");
        P.slot_limit = e7 ? atoi(e7) : WF_SLOTS;
        if (P.slot_limit < 32) P.slot_limit = 32;
        if (P.slot_limit > WF_SLOTS) P.slot_limit = WF_SLOTS;
        const unsigned need = (P.work_total + (unsigned)P.slot_limit - 1) / (unsigned)P.slot_limit;
        if (grid > need) grid = need;
        bool chunked;
        double rounds = 0.0;  // pixels per path slot
        {
            // A pixel's samples are sequential, so the end of a launch is unbalanced by up to one pixel's
            // worth of samples per slot. When the launch has few pixels per slot (multi-GPU shards, small
            // frames) the pixels are handed out in chunks of chunk_spp samples instead (the CHUNKED kernel:
            // perfectly balanced, ~3 % slower per sample); a big frame takes the plain kernel.
            const char* ec = getenv("ER_WF_CHUNK_SPP");
            const char* em = getenv("ER_WF_CHUNKED");
            // chunk size: 8 samples, more for long launches (measured on a 1/8 shard of C2 at 512 spp:
            // 487 / 496 / 506 Msamples/s for chunks of 4 / 8 / 16)
            // round 2, 1/8 stripes of C3 at 256 spp on one GPU (profiles/README.md): 4.75x / 5.15x / 5.25x / 5.13x the
            // full frame's rate summed over the stripes for chunks of 8 / 16 / 32 / 64 samples (with thr_batch 256)
            int k = ec ? atoi(ec) : (spp / 8 < 8 ? 8 : spp / 8 > 64 ? 64 : spp / 8);
            if (k < 1) k = 1;
            // tickets fit 32 bits, and the ready ring (4 B per hand-over) stays below 256 MiB
            while (((unsigned long long)(spp + k - 1) / k) * work > 0x04000000ull && k < spp) k *= 2;
            rounds = (double)P.work_total / ((double)grid * WF_SLOTS);
            // measured on 1/K shards of C2 (profiles/README.md, round 1): chunked wins below ~6 pixels per slot
            // (K = 3, 4, 8: +6 %, +7 %, +6 %) and loses above (K = 2: -2 %, full frame: -7 %). Round 2, C3 at 256 spp
            // with the adaptive batch threshold and parking (profiles/r04l_shards.txt), plain / chunked, slowest
            // stripe and sum: 4 stripes 743-790 / 734 ms and 2819-2877 / 2718 ms, 8 stripes 495-515 / 458-479 ms and
            // 3412-3452 / 3407-3473 ms; 2 stripes (7 pixels per slot) 1235 / 1303 ms: the rule stands.
            chunked = em ? atoi(em) != 0 : (rounds < 6.0 && spp > k);
            const size_t NPIX = (size_t)c->S.x_res * c->S.y_res;
            if (NPIX > (1u << 24)) chunked = false;  // a ready-ring entry holds a 24-bit pixel index
            P.chunk_spp = chunked ? k : spp;
            P.chunk_count = chunked ? (unsigned)((spp + k - 1) / k) : 1u;
            P.chunk_n1 = P.chunk_count, P.chunk_n2 = 0u, P.chunk_k2 = P.chunk_k3 = k;
            if (chunked) {
                // Shrinking chunks (ER_WF_CHUNK_TAIL=f1:d2:f2:d3: f1 of a pixel's samples in chunks of k, f2 in chunks of
                // k / d2, the rest in chunks of k / d3). The timeline of a 1/8 stripe of C3 (profiles/r04p) ends with 150 of
                // 430 ms in which no chunk is left to hand out, but smaller last chunks do NOT shorten it (0.625:4:0.25:8:
                // 436 -> 447 ms on the stripe, 3510 -> 3595 ms over the 8 stripes, profiles/r04q, r04s): what is left at
                // the end are the costliest pixels, fewer than slots, each waiting on its own sequential samples.
                // Default: every chunk k samples.
                double f1 = 1.0, f2 = 0.0;
                int d2 = 1, d3 = 1;
                if (const char* et = getenv("ER_WF_CHUNK_TAIL")) sscanf(et, "%lf:%d:%lf:%d", &f1, &d2, &f2, &d3);
                if (d2 < 1) d2 = 1;
                if (d3 < 1) d3 = 1;
                const int k2 = k / d2 > 0 ? k / d2 : 1, k3 = k / d3 > 0 ? k / d3 : 1;
                unsigned n1 = (unsigned)(f1 * spp / k);
                if ((long long)n1 * k > spp) n1 = (unsigned)(spp / k);
                const int rest1 = spp - (int)n1 * k;
                unsigned n2 = (unsigned)(f2 * spp / k2);
                if ((long long)n2 * k2 > rest1) n2 = (unsigned)(rest1 / k2);
                const int rest2 = rest1 - (int)n2 * k2;
                const unsigned n3 = (unsigned)((rest2 + k3 - 1) / k3);
                if ((unsigned long long)(n1 + n2 + n3) * work <= 0x04000000ull) {  // tickets and ring stay in range
                    P.chunk_n1 = n1, P.chunk_n2 = n2, P.chunk_k2 = k2, P.chunk_k3 = k3;
                    P.chunk_count = n1 + n2 + n3;
                }
            }
            P.progress_base = c->progress_base;
            P.ticket_total = P.work_total;
            if (chunked) {
                c->progress_base += P.chunk_count;
                size_t owned = 0;  // image pixels of the owned row blocks
                for_owned_ranges(c, [&](int y0, int y1) { owned += (size_t)(y1 - y0) * c->S.x_res; });
                P.ticket_total = (unsigned)(work + owned * (P.chunk_count - 1));
                // One ring position per hand-over of the launch: a position is never written twice. (A ring of
                // 2 x owned entries with lap tags was not enough: a pixel waits for ONE consumer while the other
                // pixels push up to chunk_count - 1 entries each, so a slow consumer's entry was overwritten by
                // the next lap and its pixel never continued: the stall of test_sample_chunks with 1-sample chunks.)
                unsigned shift = 10;
                while (((size_t)1 << shift) < owned * (P.chunk_count - 1)) shift++;
                const size_t bytes = sizeof(unsigned) << shift;
                if (c->ring_bytes < bytes) {
                    if (c->d_ring) ER_CK(cudaFree(c->d_ring));
                    c->d_ring = nullptr;
                    c->ring_bytes = 0;
                    ER_CK(cudaMalloc((void**)&c->d_ring, bytes));
                    c->ring_bytes = bytes;
                }
                ER_CK(cudaMemsetAsync(c->d_ring, 0, bytes, c->stream));
                P.ring = c->d_ring;
                P.ring_tail = c->d_ticket + 1;
                P.ring_shift = shift;
                P.ring_mask = (1u << shift) - 1u;
            }
        }
        if (c->scratch_bytes < (size_t)grid * WF_SCRATCH_BYTES_PER_CTA) {
            if (c->d_scratch) ER_CK(cudaFree(c->d_scratch));
            c->d_scratch = nullptr;
            c->scratch_bytes = (size_t)c->sm_count * per_sm * WF_SCRATCH_BYTES_PER_CTA;
            ER_CK(cudaMalloc((void**)&c->d_scratch, c->scratch_bytes));
        }
        P.scratch = c->d_scratch;
        const char* e1 = getenv("ER_WF_THR_DESCEND");
        const char* e2 = getenv("ER_WF_THR_SWITCH");
        // measured with the lean step (profiles/README.md, round 2): scenes whose leaf pairs are tested by the whole
        // warp (10 M triangles: ~76 per pair) want the descend loop to run on (C3: 198 / 216 / 220 Msamples/s for
        // 8 / 12 / 16), scenes with a triangle or two per leaf want the parked lanes served early (C2: 670 / 661 / 641)
        P.thr_descend = e1 ? atoi(e1) : (c->S.tri_count > 2 * ER_LEAF_COUNT * 4 ? 16 : 8);
        P.thr_switch = e2 ? atoi(e2) : 1;
        if (P.thr_switch < 1) P.thr_switch = 1;  // a warp with nothing to trace must always take shading work
        if (P.thr_descend < 1) P.thr_descend = 1;
        const char* e5 = getenv("ER_WF_COOP_MIN");
        P.coop_min = e5 ? atoi(e5) : 8;
        if (P.coop_min < 1) P.coop_min = 1;
        const char* e6 = getenv("ER_WF_PREFETCH_MIN");
        P.prefetch_min = e6 ? atoi(e6) : 0;
        const char* e4 = getenv("ER_WF_THR_BATCH");
        // shading batches: 384 entries on big frames; a chunked (few pixels per slot) launch does better with 256
        // (a launch with few pixels per slot is bound by its longest pixels, and smaller batches shorten a sample:
        // 1/8 stripe of C3 426 / 409 / 406 / 433 ms for 384 / 320 / 256 / 192; the full frame wants 320-384)
        P.thr_batch = e4 ? atoi(e4) : (chunked ? WF_SLOTS / 4 : rounds < 3.0 ? 5 * WF_SLOTS / 16 : 3 * WF_SLOTS / 8);
        if (P.thr_batch < 32) P.thr_batch = 32;
        if (P.thr_batch > WF_SLOTS / 2) P.thr_batch = WF_SLOTS / 2;
        {
            const char* b1 = getenv("ER_WF_BATCH_RESERVE");
            const char* b2 = getenv("ER_WF_BATCH_MIN");
            const char* b3 = getenv("ER_WF_PARK_LANES");
            const char* b4 = getenv("ER_WF_PARK_NS");
            P.batch_reserve = b1 ? atoi(b1) : WF_THREADS;
            // measured (C3, 256 spp; profiles/r04h-r04j): full frame 217 -> 225 Msamples/s, 1/2 stripe 1310 -> 1235 ms,
            // 1/8 stripe 462 -> 406 ms, 1/16 stripe 360 -> 298 ms; steady state with 256 / 64 live paths per SM 66 -> 110
            // / 24 -> 47 Msamples/s. ER_WF_BATCH_MIN=<thr_batch> ER_WF_PARK_LANES=0 is round 1's behaviour.
            P.batch_min = b2 ? atoi(b2) : 64;
            if (P.batch_min < 32) P.batch_min = 32;
            P.park_lanes = b3 ? atoi(b3) : 24;
            P.park_ns = b4 ? atoi(b4) : 2000;
        }
        const char* e3 = getenv("ER_WF_WATCHDOG_MS");
        P.watchdog = c->d_watchdog;
        P.watchdog_flag = c->d_wdflag;
        ER_CK(cudaMemsetAsync(c->d_wdflag, 0, sizeof(int), c->stream));
        ER_CK(cudaMemcpyToSymbolAsync(g_wf_watchdog, &c->d_watchdog, sizeof(int*), 0, cudaMemcpyHostToDevice, c->stream));
        P.watchdog_cycles = (long long)(e3 ? atof(e3) : 60000.0) * 2000000ll;  // ms at ~2 GHz
#ifdef WF_TIMELINE
        const char* tl_path = getenv("ER_WF_TIMELINE");
#else
        const char* tl_path = nullptr;  // the probes are compiled in with -DWF_TIMELINE only
#endif
        unsigned* d_tl = nullptr;
        const size_t tl_bytes = (size_t)grid * 512 * 8 * sizeof(unsigned);
        const size_t np_tl = (size_t)c->S.x_res * c->S.y_res;
        P.timeline = nullptr;
        P.pixel_done = nullptr;
        P.timeline_shift = 22;  // 2^22 cycles ~ 2.1 ms
        if (tl_path) {  // debugging aid: allocates and synchronises
            if (const char* ts = getenv("ER_WF_TIMELINE_SHIFT")) P.timeline_shift = atoi(ts);
            ER_CK(cudaMalloc((void**)&d_tl, tl_bytes + np_tl * 4));
            ER_CK(cudaMemsetAsync(d_tl, 0, tl_bytes + np_tl * 4, c->stream));
            P.timeline = d_tl;
            P.pixel_done = d_tl + tl_bytes / 4;
        }
        if (chunked) {
            if (count)
                k_render_wavefront<true, true><<<grid, WF_THREADS, WF_SMEM_BYTES, c->stream>>>(c->S, P);
            else
                k_render_wavefront<false, true><<<grid, WF_THREADS, WF_SMEM_BYTES, c->stream>>>(c->S, P);
        } else {
            if (count)
                k_render_wavefront<true, false><<<grid, WF_THREADS, WF_SMEM_BYTES, c->stream>>>(c->S, P);
            else
                k_render_wavefront<false, false><<<grid, WF_THREADS, WF_SMEM_BYTES, c->stream>>>(c->S, P);
        }
        if (d_tl) {
            std::vector<unsigned> tl(tl_bytes / 4 + np_tl);
            cudaError_t e = cudaMemcpyAsync(tl.data(), d_tl, tl_bytes + np_tl * 4, cudaMemcpyDeviceToHost, c->stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
            cudaFree(d_tl);
            if (e == cudaSuccess)
                if (FILE* f = fopen(tl_path, "wb")) {
                    const unsigned hdr[4] = {grid, 512u, (unsigned)P.timeline_shift | ((unsigned)c->S.x_res << 8), P.work_total};
                    fwrite(hdr, 4, 4, f);
                    fwrite(tl.data(), 4, tl.size(), f);
                    fclose(f);
                }
        }
    } else if (kernel == 3) {
        rc = launch_staged(c, P, spp, count);
        if (rc) {
            c->ev_pool.push_back(ev);  // the event pair goes back to the pool
            return rc;
        }
    } else {
        const unsigned grid = (P.work_total + 127) / 128;
        if (kernel == 1) {
            if (count)
                k_render_pixel<true, true><<<grid, 128, 0, c->stream>>>(c->S, P);
            else
                k_render_pixel<false, true><<<grid, 128, 0, c->stream>>>(c->S, P);
        } else {  // 2: reference visit order, no pruning, every shadow ray traced as a nearest query
            if (count)
                k_render_pixel<true, false><<<grid, 128, 0, c->stream>>>(c->S, P);
            else
                k_render_pixel<false, false><<<grid, 128, 0, c->stream>>>(c->S, P);
        }
    }
    ER_CK(cudaGetLastError());
    ER_CK(cudaEventRecord(ev.second, c->stream));
    c->ev_live.push_back(ev);
    c->launches++;
    return ER_OK;
}

int render_blocking(ErCtx* c, unsigned samples, const ErRenderOpts& o, bool sync_each) {
    unsigned left = samples;
    while (left > 0) {
        if (c->abort.load()) return fail(ER_ERR_ABORTED, "render aborted");
        const unsigned n = left < (unsigned)o.spp_per_launch ? left : (unsigned)o.spp_per_launch;
        {
            std::lock_guard<std::mutex> lk(c->batch_mutex);
            int rc = launch_render(c, (int)n, o);
            if (rc) return rc;
            if (sync_each) {
                ER_CK(cudaStreamSynchronize(c->stream));
                rc = check_watchdog(c);
                if (rc) return rc;
            }
        }
        c->completed += n;
        left -= n;
    }
    return ER_OK;
}

void join_worker(ErCtx* c) {
    if (c->worker.joinable()) c->worker.join();
}

int validate_desc(const ErSceneDesc* d) {
    if (!d) return fail(ER_ERR_INVALID, "null scene");
    if (d->x_res == 0 || d->y_res == 0 || (unsigned long long)d->x_res * d->y_res > 0x7fffffffull)
        return fail(ER_ERR_INVALID, "bad resolution");
    if (!d->camera) return fail(ER_ERR_INVALID, "camera missing");
    if (d->material_count < 1 || !d->materials) return fail(ER_ERR_INVALID, "at least one material is required (Scene.h:43)");
    if (d->texture_count < 0 || (d->texture_count && !d->textures)) return fail(ER_ERR_INVALID, "textures missing");
    if (!d->hdri.data || !d->hdri_cdf) return fail(ER_ERR_INVALID, "hdri texture / cdf missing");
    for (int i = 0; i < d->material_count; i++) {
        const ErMaterial& m = d->materials[i];
        const int ids[7] = {m.albedoTextureID, m.emissionTextureID, m.roughnessTextureID, m.metallicTextureID,
                            m.normalTextureID, m.opacityTextureID, m.transmissionTextureID};
        for (int k = 0; k < 7; k++)
            if (ids[k] >= d->texture_count) return fail(ER_ERR_INVALID, "material references a texture that was not supplied");
    }
    // (triangle materialIDs are checked where the triangles are read anyway: erh::flatten_geometry / k_flatten_tris)
    return ER_OK;
}

// The flattened geometry is staged in host vectors that live with the context (a re-upload touches no fresh
// pages). From the SECOND upload of a scene of the same shape on, the staging is page-locked in place
// (cudaHostRegister), so the H2D copies are plain DMA at PCIe rate instead of driver-staged copies from
// pageable memory (measured 9.6 GB/s before); pinning costs about as much as one pageable copy, so a scene
// that is uploaded once is not pinned.
void unpin_flat(ErCtx* c) {
    for (auto& r : c->pinned) cudaHostUnregister(r.first);
    c->pinned.clear();
}

int stage_flat(ErCtx* c, const ErSceneDesc* d, erh::FlatScene& fs) {
    const size_t T = d->tri_count;
    const bool same_shape = fs.tri_v.size() == 3 * T && fs.tri_n.size() == 3 * T && fs.tri_a.size() == 5 * T && !fs.pairs.empty();
    if (!same_shape) unpin_flat(c);  // the vectors are about to reallocate
    std::string err;
    int rc = erh::flatten_geometry(d, fs, err);
    if (rc) return fail(rc, "er_upload_scene: " + err);
    if (same_shape && c->pinned.empty() && !getenv("ER_NO_PIN")) {
        std::vector<erh::F4>* v[5] = {&fs.pairs, &fs.tri_v, &fs.tri_n, &fs.tri_a, &fs.tri_g};
        for (auto* x : v) {
            if (x->empty()) continue;
            if (cudaHostRegister(x->data(), x->size() * sizeof(erh::F4), cudaHostRegisterPortable) == cudaSuccess)
                c->pinned.push_back({(void*)x->data(), x->size() * sizeof(erh::F4)});
            else
                cudaGetLastError();  // pageable copies still work
        }
    }
    return ER_OK;
}

// Scenes whose triangles do not fit in L2 (10 M triangles: 1.8 GB against 126 MB) stream them through it and
// evict the child-pair records every ray walks; a persisting access-policy window over `pairs` (16.8 MB) keeps
// those resident. ER_L2_PERSIST=0/1 forces it off/on; default: on when the geometry exceeds half the L2.
void set_l2_window(ErCtx* c) {
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, c->device) != cudaSuccess) return;
    const char* e = getenv("ER_L2_PERSIST");
    const size_t geom = (size_t)c->S.tri_count * (48 + 48 + 80) + c->pairs_bytes;
    const bool on = e ? atoi(e) != 0 : false;
    (void)geom;
    cudaStreamAttrValue v{};
    if (on && p.persistingL2CacheMaxSize > 0 && c->pairs_bytes > 0) {
        size_t bytes = c->pairs_bytes;
        if (bytes > (size_t)p.accessPolicyMaxWindowSize) bytes = (size_t)p.accessPolicyMaxWindowSize;
        size_t carve = bytes < (size_t)p.persistingL2CacheMaxSize ? bytes : (size_t)p.persistingL2CacheMaxSize;
        cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, carve);
        v.accessPolicyWindow.base_ptr = (void*)c->S.pairs;
        v.accessPolicyWindow.num_bytes = bytes;
        v.accessPolicyWindow.hitRatio = carve >= bytes ? 1.0f : (float)carve / (float)bytes;
        v.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        v.accessPolicyWindow.missProp = cudaAccessPropertyNormal;
    } else {
        v.accessPolicyWindow.num_bytes = 0;
    }
    if (cudaStreamSetAttribute(c->stream, cudaStreamAttributeAccessPolicyWindow, &v) != cudaSuccess) cudaGetLastError();
    c->l2_window = on;
}

// ---------------------------------------------------------------------------------------------------
// Device-side flattening (round 2). When the caller's triangle array is page-locked (cudaHostAlloc /
// er_host_alloc / er_host_register) the raw records go over PCIe as they are (DMA straight from the caller's
// pages) and the per-slot SoA records of erh::flatten_geometry — gather by tri_indices, vertices / normals /
// attributes, the slack bound — are built by this kernel; only the per-slot slack (4 B) comes back, for the tree
// the host still builds (erh::flatten_geometry with device_slack). Same arithmetic in the same order as the
// host loop (double, no contraction): the records are the same bits (tests/test_gpu_upload.py).
// 10 M triangles: host flatten 50-70 ms of 16 threads + 1.81 GB H2D  ->  1.56 GB H2D + ~2 ms of kernel.
// ---------------------------------------------------------------------------------------------------
__global__ void k_flatten_tris(const char* __restrict__ raw, size_t stride, const int* __restrict__ idx, size_t T, int material_count,
                               int legacy_slack, float4* __restrict__ tv, float4* __restrict__ tn, float4* __restrict__ ta,
                               float* __restrict__ slack, int* flag) {
    const size_t s = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (s >= T) return;
    const int t = idx[s];
    if (t < 0 || (size_t)t >= T) {
        atomicOr(flag, 1);
        return;
    }
    const float* tr = (const float*)(raw + (size_t)t * stride);
    float w[39];
#pragma unroll
    for (int k = 0; k < 39; k++) w[k] = __ldg(tr + k);
    const float* V = w;        // vertices[3][3]
    const float* UV = w + 9;   // uv[3][3]
    const float* N = w + 18;   // normals[3][3]
    const float* TG = w + 27;  // tangents[3][3]
    const int material = __float_as_int(w[38]);
    if (material < 0 || material >= material_count) atomicOr(flag, 2);
#pragma unroll
    for (int k = 0; k < 3; k++) {
        tv[3 * s + k] = make_float4(V[3 * k], V[3 * k + 1], V[3 * k + 2], 0.f);
        tn[3 * s + k] = make_float4(N[3 * k], N[3 * k + 1], N[3 * k + 2], 0.f);
    }
    ta[5 * s + 0] = make_float4(UV[0], UV[1], UV[3], UV[4]);
    ta[5 * s + 1] = make_float4(UV[6], UV[7], w[36], w[37]);
    ta[5 * s + 2] = make_float4(TG[0], TG[1], TG[2], w[38]);
    ta[5 * s + 3] = make_float4(TG[3], TG[4], TG[5], 0.f);
    ta[5 * s + 4] = make_float4(TG[6], TG[7], TG[8], 0.f);
    // the slack bound of er_host.cpp (flatten_geometry), statement for statement
    double e2 = 0, n2 = 0, tight = 0;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const float* a = V + 3 * k;
        const float* b = V + 3 * ((k + 1) % 3);
        const double dx = (double)a[0] - b[0], dy = (double)a[1] - b[1], dz = (double)a[2] - b[2];
        e2 = fmax(e2, dx * dx + dy * dy + dz * dz);
        const float* nn = N + 3 * k;
        const double nk2 = (double)nn[0] * nn[0] + (double)nn[1] * nn[1] + (double)nn[2] * nn[2];
        n2 = fmax(n2, nk2);
#pragma unroll
        for (int j = 0; j < 3; j++) {
            const float* c = V + 3 * j;
            const double dt = ((double)c[0] - a[0]) * nn[0] + ((double)c[1] - a[1]) * nn[1] + ((double)c[2] - a[2]) * nn[2];
            tight = fmax(tight, fabs(dt) * sqrt(nk2));
        }
    }
    double sl = sqrt(e2) * n2 * 1.001 + 1e-30;
    if (!legacy_slack && n2 <= 1.21) sl = fmin(sl, tight * 1.001 + 1e-30);
    if (!(sl == sl) || !(e2 == e2) || !(n2 == n2) || !(tight == tight)) sl = (double)INFINITY;
    float f = (float)fmin(sl, (double)INFINITY);  // round to nearest, then up if that fell below
    if ((double)f < sl) f = nextafterf(f, INFINITY);
    slack[s] = f;
}

// is [p, p + bytes) page-locked host memory the device can read by DMA?
bool host_range_pinned(const void* p) {
    cudaPointerAttributes a{};
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeHost;
}

bool device_flatten_wanted(const ErSceneDesc* d) {
    if (getenv("ER_NO_DEVICE_FLATTEN")) return false;
    if (const char* gbx = getenv("ER_GROUP_BOXES"); gbx && atoi(gbx) != 0) return false;  // the group boxes are built from the host records
    if (d->tri_count < 65536 || d->tri_stride % 4 != 0 || ((uintptr_t)d->tris & 3u) != 0) return false;
    return host_range_pinned(d->tris) && host_range_pinned((const char*)d->tris + d->tri_count * d->tri_stride - 1);
}

// builds tri_v / tri_n / tri_a on the device from the caller's page-locked records, and the pair records in fs.pairs
// (host); leaves the device pointers in *tv, *tn, *ta
int flatten_on_device(ErCtx* c, const ErSceneDesc* d, erh::FlatScene& fs, const float4** tv, const float4** tn, const float4** ta) {
    const size_t T = d->tri_count;
    int rc;
    void *raw, *idx, *v, *n, *a, *sl;
    if ((rc = dev_alloc(c, T * d->tri_stride, &raw))) return rc;
    if ((rc = dev_alloc(c, T * sizeof(int), &idx))) return rc;
    if ((rc = dev_alloc(c, sizeof(float4) * 3 * T, &v))) return rc;
    if ((rc = dev_alloc(c, sizeof(float4) * 3 * T, &n))) return rc;
    if ((rc = dev_alloc(c, sizeof(float4) * 5 * T, &a))) return rc;
    if ((rc = dev_alloc(c, sizeof(float) * T, &sl))) return rc;
    if (!c->d_flag) ER_CK(cudaMalloc((void**)&c->d_flag, sizeof(int)));
    if (c->h_slack_n < T) {
        if (c->h_slack) cudaFreeHost(c->h_slack);
        c->h_slack = nullptr;
        c->h_slack_n = 0;
        ER_CK(cudaHostAlloc((void**)&c->h_slack, sizeof(float) * T + sizeof(int), cudaHostAllocDefault));
        c->h_slack_n = T;
    }
    ER_CK(cudaMemsetAsync(c->d_flag, 0, sizeof(int), c->stream));
    ER_CK(cudaMemcpyAsync(raw, d->tris, T * d->tri_stride, cudaMemcpyHostToDevice, c->stream));
    ER_CK(cudaMemcpyAsync(idx, d->tri_indices, T * sizeof(int), cudaMemcpyHostToDevice, c->stream));
    const char* slack_env = getenv("ER_SLACK");
    const int legacy = slack_env && !strcmp(slack_env, "legacy");
    k_flatten_tris<<<(unsigned)((T + 255) / 256), 256, 0, c->stream>>>((const char*)raw, d->tri_stride, (const int*)idx, T, d->material_count,
                                                                   legacy, (float4*)v, (float4*)n, (float4*)a, (float*)sl, c->d_flag);
    ER_CK(cudaGetLastError());
    ER_CK(cudaMemcpyAsync(c->h_slack, sl, sizeof(float) * T, cudaMemcpyDeviceToHost, c->stream));
    int* h_flag = (int*)(c->h_slack + T);
    ER_CK(cudaMemcpyAsync(h_flag, c->d_flag, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    ER_CK(cudaStreamSynchronize(c->stream));
    if (*h_flag & 1) return fail(ER_ERR_INVALID, "er_upload_scene: tri_indices entry out of range");
    if (*h_flag & 2) return fail(ER_ERR_INVALID, "er_upload_scene: triangle materialID out of range");
    std::string err;
    rc = erh::flatten_geometry(d, fs, err, c->h_slack);  // the tree: pair records with subtree slack (host, ~3 ms)
    if (rc) return fail(rc, "er_upload_scene: " + err);
    *tv = (const float4*)v, *tn = (const float4*)n, *ta = (const float4*)a;
    return ER_OK;
}

// copy_scene (src/SYCLCopy.cpp:3-104): the flattened geometry `fs` and the rest of `d` to the device of `c`,
// then setupKernel. Copies are enqueued on c->stream; the caller synchronises before the host arrays go away.
// device_flatten: `fs` is this context's own staging and holds nothing yet — the per-slot records are built on the
// device from the caller's page-locked triangles (flatten_on_device), the pair records in fs.pairs on the host.
int upload_flat(ErCtx* c, const ErSceneDesc* d, erh::FlatScene& fs, bool device_flatten = false) {
    int rc;
    ER_CK(cudaStreamSynchronize(c->stream));
    free_scene(c);
    DScene& S = c->S;
    const float4 *pairs, *tv, *tn, *ta;
    if (device_flatten) {
        if ((rc = flatten_on_device(c, d, fs, &tv, &tn, &ta))) return rc;
        if ((rc = dev_upload<float4>(c, (const float4*)fs.pairs.data(), fs.pairs.size(), &pairs))) return rc;
    } else {
        if ((rc = dev_upload<float4>(c, (const float4*)fs.pairs.data(), fs.pairs.size(), &pairs))) return rc;
        if ((rc = dev_upload<float4>(c, (const float4*)fs.tri_v.data(), fs.tri_v.size(), &tv))) return rc;
        if ((rc = dev_upload<float4>(c, (const float4*)fs.tri_n.data(), fs.tri_n.size(), &tn))) return rc;
        if ((rc = dev_upload<float4>(c, (const float4*)fs.tri_a.data(), fs.tri_a.size(), &ta))) return rc;
    }
    S.pairs = pairs, S.tri_v = tv, S.tri_n = tn, S.tri_a = ta;
    {
        const float4* tg;
        if ((rc = dev_upload<float4>(c, (const float4*)fs.tri_g.data(), fs.tri_g.size(), &tg))) return rc;
        S.tri_g = tg;
    }
    S.tri_count = (int)d->tri_count;
    c->pairs_bytes = fs.pairs.size() * sizeof(float4);

    std::vector<DMaterial> mats((size_t)d->material_count);
    for (int i = 0; i < d->material_count; i++) {
        static_assert(offsetof(DMaterial, roughness_pow) == sizeof(ErMaterial), "DMaterial extends ErMaterial");
        memcpy(&mats[i], &d->materials[i], sizeof(ErMaterial));
        // kernel.cpp:152-153 for untextured channels
        mats[i].roughness_pow = powf(d->materials[i].roughness, 2.2f);
        mats[i].metallic_pow = powf(d->materials[i].metallic, 2.2f);
        mats[i].pad = 0.f;
    }
    const DMaterial* dm;
    if ((rc = dev_upload<DMaterial>(c, mats.data(), mats.size(), &dm))) return rc;
    ER_CK(cudaStreamSynchronize(c->stream));  // `mats` is a local
    S.materials = dm;
    S.material_count = d->material_count;

    c->h_textures.resize((size_t)d->texture_count);
    for (int i = 0; i < d->texture_count; i++)
        if ((rc = upload_texture(c, d->textures[i], c->h_textures[i], "textures[]"))) return rc;
    const DTexture* dt;
    if ((rc = dev_upload<DTexture>(c, c->h_textures.data(), c->h_textures.size(), &dt))) return rc;
    S.textures = dt;
    S.texture_count = d->texture_count;

    if ((rc = upload_texture(c, d->hdri, S.hdri, "hdri"))) return rc;
    const float* cdf;
    // the full w*h+1 entries (the reference copies one byte more than w*h floats, SYCLCopy.cpp:95)
    if ((rc = dev_upload<float>(c, d->hdri_cdf, (size_t)d->hdri.width * d->hdri.height + 1, &cdf))) return rc;
    S.cdf = cdf;
    S.radianceSum = d->hdri_radiance_sum;

    S.x_res = (int)d->x_res;
    S.y_res = (int)d->y_res;
    S.cam = make_camera(*d->camera);
    // PointLight[] (src/SYCLCopy.cpp:64-66): inert in the reference's estimator, read by ER_EXT_POINT_LIGHTS
    S.point_lights = nullptr;
    S.point_light_count = 0;
    if (d->point_light_count > 0 && d->point_lights) {
        const float* pl;
        if ((rc = dev_upload<float>(c, (const float*)d->point_lights, (size_t)d->point_light_count * 6, &pl))) return rc;
        S.point_lights = pl;
        S.point_light_count = d->point_light_count;
    }
    const size_t P = (size_t)d->x_res * d->y_res;
    void* p;
    if ((rc = dev_alloc(c, sizeof(float4) * 5 * P, &p))) return rc;
    S.passes = (float4*)p;
    if ((rc = dev_alloc(c, sizeof(uint4) * P, &p))) return rc;
    S.pix = (uint4*)p;
    drain_pool(c);
    c->uploaded = true;
    set_l2_window(c);
    return launch_setup(c);
}

}  // namespace

// ---- scratch of er_trace_noalloc (the entry point is further down, with the other ray queries) -----------------
// File-scope device arrays: in host code these names are host-side stand-ins, not device addresses; the addresses
// come from cudaGetSymbolAddress, asked once per device with that device current (er_create, at set-up).
static __device__ __align__(256) float g_na_rays[6 * (size_t)ER_TRACE_NOALLOC_MAX];
static __device__ __align__(256) DHit g_na_hits[ER_TRACE_NOALLOC_MAX];

namespace {
struct NoAllocScratch {
    float* rays = nullptr;
    DHit* hits = nullptr;
};
std::mutex g_na_mu;
NoAllocScratch g_na_dev[64];

// `device` is the current device
int noalloc_scratch(int device, NoAllocScratch& out) {
    if (device < 0 || device >= 64) return fail(ER_ERR_INVALID, "er_trace_noalloc: device index out of range");
    std::lock_guard<std::mutex> lk(g_na_mu);
    NoAllocScratch& s = g_na_dev[device];
    if (!s.rays || !s.hits) {
        void *pr = nullptr, *ph = nullptr;
        ER_CK(cudaGetSymbolAddress(&pr, g_na_rays));
        ER_CK(cudaGetSymbolAddress(&ph, g_na_hits));
        if (!pr || !ph) return fail(ER_ERR_CUDA, "er_trace_noalloc: cudaGetSymbolAddress returned a null address");
        s.rays = (float*)pr;
        s.hits = (DHit*)ph;
    }
    out = s;
    return ER_OK;
}
}  // namespace


extern "C" {

const char* er_last_error(void) { return g_err.c_str(); }

int er_device_count(int* count) {
    if (!count) return fail(ER_ERR_INVALID, "null argument");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        *count = 0;
        cudaGetLastError();
        return fail(ER_ERR_NO_DEVICE, std::string("no CUDA device: ") + cudaGetErrorString(e));
    }
    *count = n;
    return ER_OK;
}

int er_device_info(int index, ErDeviceInfo* info) {
    if (!info) return fail(ER_ERR_INVALID, "null argument");
    int n = 0;
    int rc = er_device_count(&n);
    if (rc) return rc;
    if (index < 0 || index >= n) return fail(ER_ERR_INVALID, "device index out of range");
    cudaDeviceProp p;
    ER_CK(cudaGetDeviceProperties(&p, index));
    memset(info, 0, sizeof *info);
    snprintf(info->name, sizeof info->name, "%s", p.name);
    snprintf(info->platform, sizeof info->platform, "NVIDIA CUDA");
    info->memory_bytes = (unsigned long long)p.totalGlobalMem;
    info->max_compute_units = p.multiProcessorCount;
    info->cc_major = p.major;
    info->cc_minor = p.minor;
    info->is_compatible = p.major == 10 ? 1 : 0;
    return ER_OK;
}

int er_create(int device_id, ErCtx** out) {
    if (!out) return fail(ER_ERR_INVALID, "null argument");
    *out = nullptr;
    ErDeviceInfo info;
    int rc = er_device_info(device_id, &info);
    if (rc) return rc;
    if (!info.is_compatible)
        return fail(ER_ERR_NO_DEVICE, std::string("device '") + info.name + "' is not compute capability 10.x; kernels are built for sm_100a only");
    ER_CK(cudaSetDevice(device_id));
    ErCtx* c = new ErCtx();
    c->device = device_id;
    c->sm_count = info.max_compute_units;
    cudaError_t e = cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->read_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaMalloc((void**)&c->d_ticket, 2 * sizeof(unsigned));
    if (e == cudaSuccess) e = cudaMalloc((void**)&c->d_counters, sizeof(DCounters));
    if (e == cudaSuccess) e = cudaMemset(c->d_counters, 0, sizeof(DCounters));
    if (e == cudaSuccess) e = cudaMalloc((void**)&c->d_wdflag, sizeof(int));
    if (e == cudaSuccess) e = cudaHostAlloc((void**)&c->h_watchdog, 4096, cudaHostAllocMapped);
    if (e == cudaSuccess) {
        memset(c->h_watchdog, 0, 4096);
        e = cudaHostGetDevicePointer((void**)&c->d_watchdog, c->h_watchdog, 0);
    }
    // the render kernel asks for more dynamic shared memory than the default limit; the attribute is
    // per DEVICE, so it is set for every context's device (a second context on another GPU of the same
    // process needs it too)
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_render_wavefront<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, WF_SMEM_BYTES);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_render_wavefront<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, WF_SMEM_BYTES);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_render_wavefront<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, WF_SMEM_BYTES);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_render_wavefront<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, WF_SMEM_BYTES);
    if (e != cudaSuccess) {
        delete c;
        return fail(ER_ERR_CUDA, std::string("context setup failed: ") + cudaGetErrorString(e));
    }
    c->stream = c->own_stream;
    {
        // set-up of er_trace_noalloc: the device addresses of its file-scope scratch, once per device (a failure
        // here is not a failure of er_create: the entry point asks again and reports it)
        NoAllocScratch na;
        if (noalloc_scratch(device_id, na) != ER_OK) cudaGetLastError();
    }
    *out = c;
    return ER_OK;
}

int er_destroy(ErCtx* c) {
    if (!c) return ER_OK;
    c->abort = true;
    join_worker(c);
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    free_scene(c);
    drain_pool(c);
    unpin_flat(c);
    for (auto& pr : c->ev_live) cudaEventDestroy(pr.first), cudaEventDestroy(pr.second);
    for (auto& pr : c->ev_pool) cudaEventDestroy(pr.first), cudaEventDestroy(pr.second);
    cudaFree(c->d_ticket);
    cudaFree(c->d_counters);
    if (c->h_watchdog) cudaFreeHost(c->h_watchdog);
    cudaFree(c->d_wdflag);
    cudaFree(c->d_scratch);
    cudaFree(c->d_staged);
    if (c->h_live) cudaFreeHost(c->h_live);
    cudaFree(c->d_ring);
    cudaFree(c->d_row_cost);
    cudaFree(c->d_block_order);
    cudaFree(c->d_flag);
    if (c->h_slack) cudaFreeHost(c->h_slack);
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    if (c->read_stream) cudaStreamDestroy(c->read_stream);
    delete c;
    return ER_OK;
}

int er_set_shard(ErCtx* c, int shard_index, int shard_count, int rows_per_block) {
    int rc = use(c);
    if (rc) return rc;
    if (c->running) return fail(ER_ERR_INVALID, "render in progress");
    if (shard_count < 1 || shard_index < 0 || shard_index >= shard_count) return fail(ER_ERR_INVALID, "bad shard index/count");
    if (rows_per_block <= 0) rows_per_block = 8;
    if (rows_per_block % 4) return fail(ER_ERR_INVALID, "rows_per_block must be a multiple of 4");
    c->block_order_n = 0;
    c->first_block = shard_index;
    c->block_step = shard_count;
    c->rows_per_block = rows_per_block;
    c->y_limit = 0;
    if (c->uploaded) return launch_setup(c);
    return ER_OK;
}

int er_set_shard_stripe(ErCtx* c, int y_begin, int y_end, int rows_per_block) {
    int rc = use(c);
    if (rc) return rc;
    if (c->running) return fail(ER_ERR_INVALID, "render in progress");
    if (rows_per_block <= 0) rows_per_block = 8;
    if (rows_per_block % 4) return fail(ER_ERR_INVALID, "rows_per_block must be a multiple of 4");
    if (y_begin < 0 || y_end < y_begin || y_begin % rows_per_block) return fail(ER_ERR_INVALID, "stripe must start at a multiple of rows_per_block and not end before it starts");
    c->block_order_n = 0;
    c->first_block = y_begin / rows_per_block;
    c->block_step = 1;
    c->rows_per_block = rows_per_block;
    // an empty stripe owns nothing: first_block beyond every block
    if (y_end == y_begin) c->first_block = 0x3fffffff / rows_per_block;
    c->y_limit = y_end > 0 ? y_end : 1;
    if (c->uploaded) return launch_setup(c);
    return ER_OK;
}

int er_set_block_order(ErCtx* c, const int* order, int n) {
    int rc = use(c);
    if (rc) return rc;
    if (c->running) return fail(ER_ERR_INVALID, "render in progress");
    if (!order || n <= 0) {
        c->block_order_n = 0;
        return ER_OK;
    }
    std::vector<char> seen((size_t)n, 0);
    for (int i = 0; i < n; i++) {
        if (order[i] < 0 || order[i] >= n || seen[(size_t)order[i]]) return fail(ER_ERR_INVALID, "er_set_block_order: not a permutation of 0..n-1");
        seen[(size_t)order[i]] = 1;
    }
    if (c->block_order_cap < n) {
        if (c->d_block_order) ER_CK(cudaFree(c->d_block_order));
        c->d_block_order = nullptr;
        c->block_order_cap = 0;
        ER_CK(cudaMalloc((void**)&c->d_block_order, sizeof(int) * (size_t)n));
        c->block_order_cap = n;
    }
    ER_CK(cudaMemcpyAsync(c->d_block_order, order, sizeof(int) * (size_t)n, cudaMemcpyHostToDevice, c->stream));
    ER_CK(cudaStreamSynchronize(c->stream));  // `order` is the caller's
    c->block_order_n = n;
    return ER_OK;
}

// Host only: the row blocks of [y_begin, y_end) in descending order of their cost (block indices relative to the
// first block of the range); blocks the prepass did not sample take the nearest sampled block's cost.
int er_order_blocks(const uint32_t* cost_per_row, int y_res, int rows_per_block, int y_begin, int y_end, int* order_out, int* n_out) {
    if (!cost_per_row || !order_out || !n_out || y_res < 1) return fail(ER_ERR_INVALID, "bad argument");
    if (rows_per_block <= 0) rows_per_block = 8;
    if (y_begin < 0 || y_end > y_res || y_end < y_begin || y_begin % rows_per_block) return fail(ER_ERR_INVALID, "er_order_blocks: bad row range");
    const int nb = (y_res + rows_per_block - 1) / rows_per_block;
    std::vector<double> bc((size_t)nb, -1.0);
    for (int b = 0; b < nb; b++) {
        double s = 0;
        bool any = false;
        for (int y = b * rows_per_block; y < y_res && y < (b + 1) * rows_per_block; y++) {
            s += cost_per_row[y];
            any = any || cost_per_row[y] != 0;
        }
        if (any) bc[(size_t)b] = s;
    }
    std::vector<double> filled(bc);
    for (int b = 0; b < nb; b++) {
        if (bc[(size_t)b] >= 0) continue;
        double v = 0.0;
        for (int d = 1; d < nb; d++) {
            if (b - d >= 0 && bc[(size_t)(b - d)] >= 0) { v = bc[(size_t)(b - d)]; break; }
            if (b + d < nb && bc[(size_t)(b + d)] >= 0) { v = bc[(size_t)(b + d)]; break; }
        }
        filled[(size_t)b] = v;
    }
    const int b0 = y_begin / rows_per_block, b1 = (y_end + rows_per_block - 1) / rows_per_block;
    const int n = b1 - b0;
    std::vector<int> idx((size_t)n);
    for (int i = 0; i < n; i++) idx[(size_t)i] = i;
    std::stable_sort(idx.begin(), idx.end(), [&](int a, int b) { return filled[(size_t)(b0 + a)] > filled[(size_t)(b0 + b)]; });
    for (int i = 0; i < n; i++) order_out[i] = idx[(size_t)i];
    *n_out = n;
    return ER_OK;
}

int er_set_stream(ErCtx* c, void* cuda_stream) {
    int rc = use(c);
    if (rc) return rc;
    if (c->running) return fail(ER_ERR_INVALID, "render in progress");
    ER_CK(cudaStreamSynchronize(c->stream));
    c->stream = cuda_stream ? (cudaStream_t)cuda_stream : c->own_stream;
    return ER_OK;
}

int er_bvh_build(const void* tris, size_t tri_count, size_t tri_stride, ErNode* nodes_out, int32_t* tri_indices_out,
                 int nthreads) {
    std::string err;
    int rc = erh::bvh_build(tris, tri_count, tri_stride, nodes_out, tri_indices_out, nthreads, err);
    return rc ? fail(rc, err) : ER_OK;
}

int er_hdri_build_cdf(const ErTexture* tex, float* cdf_out, float* radiance_sum_out) {
    std::string err;
    int rc = erh::hdri_build_cdf(tex, cdf_out, radiance_sum_out, err);
    return rc ? fail(rc, err) : ER_OK;
}

int er_flatten_roundtrip(const ErSceneDesc* d, ErNode* nodes_out, float* slot_tris_out, float* max_slack_out) {
    if (!d || !nodes_out) return fail(ER_ERR_INVALID, "null argument");
    erh::FlatScene fs;
    std::string err;
    int rc = erh::flatten_geometry(d, fs, err);
    if (rc) return fail(rc, "er_flatten_roundtrip: " + err);
    erh::unflatten_nodes(fs, nodes_out);
    if (max_slack_out) *max_slack_out = fs.max_slack;
    if (slot_tris_out) {
        for (size_t s = 0; s < d->tri_count; s++) {
            float* w = slot_tris_out + 39 * s;
            const erh::F4* v = &fs.tri_v[3 * s];
            const erh::F4* n = &fs.tri_n[3 * s];
            const erh::F4* a = &fs.tri_a[5 * s];
            for (int k = 0; k < 3; k++) {
                w[3 * k] = v[k].x, w[3 * k + 1] = v[k].y, w[3 * k + 2] = v[k].z;
                w[18 + 3 * k] = n[k].x, w[18 + 3 * k + 1] = n[k].y, w[18 + 3 * k + 2] = n[k].z;
                w[27 + 3 * k] = a[2 + k].x, w[27 + 3 * k + 1] = a[2 + k].y, w[27 + 3 * k + 2] = a[2 + k].z;
            }
            w[9] = a[0].x, w[10] = a[0].y, w[11] = 0.f;
            w[12] = a[0].z, w[13] = a[0].w, w[14] = 0.f;
            w[15] = a[1].x, w[16] = a[1].y, w[17] = 0.f;
            w[36] = a[1].z;
            w[37] = a[1].w;  // objectID bits
            w[38] = a[2].w;  // materialID bits
        }
    }
    return ER_OK;
}

int er_upload_scene(ErCtx* c, const ErSceneDesc* d) {
    NvtxRange nvtx_("er_upload_scene");
    int rc = use(c);
    if (rc) return rc;
    if (c->running) return fail(ER_ERR_INVALID, "render in progress");
    if ((rc = validate_desc(d))) return rc;
    const bool timing = getenv("ER_TIMING") != nullptr;
    const auto t0 = std::chrono::steady_clock::now();
    const bool on_device = device_flatten_wanted(d);
    if (on_device) {
        // the host staging of the per-slot records is not used: give its pages (and their page-locking) back
        if (!c->flat.tri_v.empty()) {
            unpin_flat(c);
            std::vector<erh::F4>().swap(c->flat.tri_v);
            std::vector<erh::F4>().swap(c->flat.tri_n);
            std::vector<erh::F4>().swap(c->flat.tri_a);
        }
    } else if ((rc = stage_flat(c, d, c->flat))) {
        return rc;
    }
    if (timing) fprintf(stderr, "er_upload_scene %-12s %8.2f ms\n", on_device ? "(on device)" : "flatten", std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
    if ((rc = upload_flat(c, d, c->flat, on_device))) return rc;
    ER_CK(cudaStreamSynchronize(c->stream));  // host buffers are borrowed only for this call
    if (timing) fprintf(stderr, "er_upload_scene %-12s %8.2f ms\n", "total", std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
    return ER_OK;
}

int er_set_camera(ErCtx* c, const ErCamera* camera) {
    int rc = use(c);
    if (rc) return rc;
    if (!camera) return fail(ER_ERR_INVALID, "null camera");
    if (c->running) return fail(ER_ERR_INVALID, "render in progress");
    c->S.cam = make_camera(*camera);
    return ER_OK;
}

int er_render_reset(ErCtx* c) {
    int rc = use(c);
    if (rc) return rc;
    if (!c->uploaded) return fail(ER_ERR_INVALID, "no scene uploaded");
    if (c->running) return fail(ER_ERR_INVALID, "render in progress");
    return launch_setup(c);
}

int er_render_samples(ErCtx* c, unsigned samples, const ErRenderOpts* opts) {
    int rc = use(c);
    if (rc) return rc;
    if (c->running) return fail(ER_ERR_INVALID, "render in progress");
    c->abort = false;
    rc = render_blocking(c, samples, norm_opts(opts), false);
    if (rc) return rc;
    ER_CK(cudaStreamSynchronize(c->stream));
    return check_watchdog(c);
}

int er_render_enqueue(ErCtx* c, unsigned samples, const ErRenderOpts* opts) {
    int rc = use(c);
    if (rc) return rc;
    if (c->running) return fail(ER_ERR_INVALID, "render in progress");
    c->abort = false;
    return render_blocking(c, samples, norm_opts(opts), false);
}

int er_render_start(ErCtx* c, unsigned target_samples, const ErRenderOpts* opts) {
    int rc = use(c);
    if (rc) return rc;
    if (!c->uploaded) return fail(ER_ERR_INVALID, "no scene uploaded");
    if (c->running) return fail(ER_ERR_INVALID, "render in progress");
    join_worker(c);
    const ErRenderOpts o = norm_opts(opts);
    c->abort = false;
    c->running = true;
    c->worker_status = 0;
    c->worker = std::thread([c, target_samples, o]() {
        cudaSetDevice(c->device);
        int rc2 = render_blocking(c, target_samples, o, true);
        if (rc2 && rc2 != ER_ERR_ABORTED) c->worker_error = g_err;
        c->worker_status = rc2;
        c->running = false;
    });
    return ER_OK;
}

int er_render_poll(ErCtx* c, unsigned* samples, int* running) {
    if (!c) return fail(ER_ERR_INVALID, "null context");
    if (samples) *samples = c->completed.load() + 1;  // N+1 semantics of dev_samples
    if (running) *running = c->running.load() ? 1 : 0;
    return ER_OK;
}

int er_render_wait(ErCtx* c) {
    if (!c) return fail(ER_ERR_INVALID, "null context");
    join_worker(c);
    const int st = c->worker_status.load();
    if (st == ER_ERR_ABORTED) return fail(st, "render aborted");
    if (st) return fail(st, c->worker_error);
    return ER_OK;
}

int er_render_abort(ErCtx* c) {
    if (!c) return fail(ER_ERR_INVALID, "null context");
    c->abort = true;
    return ER_OK;
}

int er_sync(ErCtx* c) {
    int rc = use(c);
    if (rc) return rc;
    ER_CK(cudaStreamSynchronize(c->stream));
    return check_watchdog(c);
}

int er_get_render_info(ErCtx* c, unsigned* samples) {
    int rc = use(c);
    if (rc) return rc;
    if (!samples) return fail(ER_ERR_INVALID, "null argument");
    if (!c->uploaded) return fail(ER_ERR_INVALID, "no scene uploaded");
    if (c->first_block != 0) {  // pixel 0 is not this shard's
        *samples = c->completed.load() + 1;
        return ER_OK;
    }
    // dev_samples[0] read on a second stream while rendering goes on (Managers.cpp:217-218)
    unsigned v = 0;
    ER_CK(cudaMemcpyAsync(&v, &c->S.pix[0].y, sizeof v, cudaMemcpyDeviceToHost, c->read_stream));
    ER_CK(cudaStreamSynchronize(c->read_stream));
    *samples = v;
    return ER_OK;
}

int er_get_pass_device(ErCtx* c, int pass, void** device_ptr) {
    int rc = use(c);
    if (rc) return rc;
    if (!device_ptr || pass < 0 || pass >= ER_PASSES_COUNT) return fail(ER_ERR_INVALID, "bad pass");
    if (!c->uploaded) return fail(ER_ERR_INVALID, "no scene uploaded");
    *device_ptr = (void*)(c->S.passes + (size_t)pass * c->S.x_res * c->S.y_res);
    return ER_OK;
}

int er_get_pass(ErCtx* c, int pass, float* out_rgba) {
    NvtxRange nvtx_("er_get_pass");
    int rc = use(c);
    if (rc) return rc;
    if (!out_rgba || pass < 0 || pass >= ER_PASSES_COUNT) return fail(ER_ERR_INVALID, "bad pass");
    if (!c->uploaded) return fail(ER_ERR_INVALID, "no scene uploaded");
    const size_t P = (size_t)c->S.x_res * c->S.y_res;
    std::lock_guard<std::mutex> lk(c->batch_mutex);  // whole batches only
    ER_CK(cudaMemcpyAsync(out_rgba, c->S.passes + (size_t)pass * P, sizeof(float4) * P, cudaMemcpyDeviceToHost, c->stream));
    ER_CK(cudaStreamSynchronize(c->stream));
    return check_watchdog(c);
}

int er_get_pass_owned(ErCtx* c, int pass, float* frame_rgba) {
    NvtxRange nvtx_("er_get_pass_owned");
    int rc = use(c);
    if (rc) return rc;
    if (!frame_rgba || pass < 0 || pass >= ER_PASSES_COUNT) return fail(ER_ERR_INVALID, "bad pass");
    if (!c->uploaded) return fail(ER_ERR_INVALID, "no scene uploaded");
    const size_t P = (size_t)c->S.x_res * c->S.y_res, W = (size_t)c->S.x_res;
    std::lock_guard<std::mutex> lk(c->batch_mutex);  // whole batches only
    // one copy per run of consecutive owned rows (rows are contiguous in memory)
    int run0 = -1, run1 = -1;
    cudaError_t e = cudaSuccess;
    auto flush = [&]() {
        if (run0 < 0 || e != cudaSuccess) return;
        e = cudaMemcpyAsync(frame_rgba + 4 * W * (size_t)run0, c->S.passes + (size_t)pass * P + W * (size_t)run0,
                            sizeof(float4) * W * (size_t)(run1 - run0), cudaMemcpyDeviceToHost, c->stream);
    };
    for_owned_ranges(c, [&](int y0, int y1) {
        if (y0 == run1) {
            run1 = y1;
            return;
        }
        flush();
        run0 = y0;
        run1 = y1;
    });
    flush();
    ER_CK(e);
    ER_CK(cudaStreamSynchronize(c->stream));
    return check_watchdog(c);
}

int er_row_cost(ErCtx* c, int block_first, int block_stride, const ErRenderOpts* opts, uint32_t* cost_per_row) {
    NvtxRange nvtx_("er_row_cost");
    int rc = use(c);
    if (rc) return rc;
    if (!cost_per_row || block_stride < 1 || block_first < 0) return fail(ER_ERR_INVALID, "bad argument");
    if (!c->uploaded) return fail(ER_ERR_INVALID, "no scene uploaded");
    if (c->running) return fail(ER_ERR_INVALID, "render in progress");
    const size_t rows = (size_t)c->S.y_res;
    if (c->row_cost_rows < rows) {
        if (c->d_row_cost) ER_CK(cudaFree(c->d_row_cost));
        c->d_row_cost = nullptr;
        c->row_cost_rows = 0;
        ER_CK(cudaMalloc((void**)&c->d_row_cost, rows * sizeof(unsigned)));
        c->row_cost_rows = rows;
    }
    ER_CK(cudaMemsetAsync(c->d_row_cost, 0, rows * sizeof(unsigned), c->stream));
    // every block_stride-th row block of the WHOLE frame, 1 sample, counting build
    const int fb = c->first_block, bs = c->block_step, yl = c->y_limit;
    c->first_block = block_first, c->block_step = block_stride, c->y_limit = 0;
    ErRenderOpts o = norm_opts(opts);
    o.kernel = 0;
    o.count_stats = 2;
    o.spp_per_launch = 1;
    rc = launch_setup(c);
    if (!rc) rc = launch_render(c, 1, o);
    if (!rc) {
        cudaError_t e = cudaMemcpyAsync(cost_per_row, c->d_row_cost, rows * sizeof(unsigned), cudaMemcpyDeviceToHost, c->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
        if (e != cudaSuccess) rc = fail(ER_ERR_CUDA, std::string("er_row_cost: ") + cudaGetErrorString(e));
    }
    if (!rc) rc = check_watchdog(c);
    c->first_block = fb, c->block_step = bs, c->y_limit = yl;
    const int rc2 = launch_setup(c);  // pixel state as after er_render_reset, for the shard's own pixels
    return rc ? rc : rc2;
}

int er_balance_stripes(const uint32_t* cost_per_row, int y_res, int rows_per_block, int n, int* y_bounds) {
    if (!cost_per_row || !y_bounds || y_res < 1 || n < 1) return fail(ER_ERR_INVALID, "bad argument");
    if (rows_per_block <= 0) rows_per_block = 8;
    const int nb = (y_res + rows_per_block - 1) / rows_per_block;
    // cost per row block; blocks that were not sampled (all rows 0) take the nearest sampled block's cost
    std::vector<double> bc((size_t)nb, -1.0);
    for (int b = 0; b < nb; b++) {
        double s = 0;
        bool any = false;
        for (int y = b * rows_per_block; y < y_res && y < (b + 1) * rows_per_block; y++) {
            s += cost_per_row[y];
            any = any || cost_per_row[y] != 0;
        }
        if (any) bc[(size_t)b] = s;
    }
    std::vector<double> filled(bc);
    for (int b = 0; b < nb; b++) {
        if (bc[(size_t)b] >= 0) continue;
        double v = 1.0;
        for (int d = 1; d < nb; d++) {
            if (b - d >= 0 && bc[(size_t)(b - d)] >= 0) { v = bc[(size_t)(b - d)]; break; }
            if (b + d < nb && bc[(size_t)(b + d)] >= 0) { v = bc[(size_t)(b + d)]; break; }
        }
        filled[(size_t)b] = v;
    }
    double total = 0;
    for (double v : filled) total += v + 1e-9;
    // boundary k = the first block edge where the running cost reaches k/n of the total
    y_bounds[0] = 0;
    double run = 0;
    int k = 1;
    for (int b = 0; b < nb && k < n; b++) {
        run += filled[(size_t)b] + 1e-9;
        while (k < n && run >= total * k / n) {
            // the edge after block b, or before it when that is closer to the target
            const double over = run - total * k / n, under = total * k / n - (run - filled[(size_t)b] - 1e-9);
            int edge = (over <= under || b * rows_per_block <= y_bounds[k - 1]) ? (b + 1) : b;
            int y = edge * rows_per_block;
            if (y > y_res) y = y_res;
            if (y < y_bounds[k - 1]) y = y_bounds[k - 1];
            y_bounds[k++] = y;
        }
    }
    while (k < n) y_bounds[k++] = y_res;
    y_bounds[n] = y_res;
    return ER_OK;
}

// Feedback step of the stripe balance: the stripes y_bounds[0..n] took ms[0..n-1]; spread every stripe's time evenly
// over its rows and put boundary k where the running time reaches k/n of the total (block edges). A launch has a part
// that does not shrink with its rows (its longest pixels), so one step under-corrects and never overshoots; two or
// three steps settle.
#define ER_REBALANCE_GAIN 1.5
int er_rebalance_stripes(const int* y_bounds, const float* ms, int n, int rows_per_block, int* y_bounds_out) {
    if (!y_bounds || !ms || !y_bounds_out || n < 1) return fail(ER_ERR_INVALID, "bad argument");
    if (rows_per_block <= 0) rows_per_block = 8;
    const int y_res = y_bounds[n];
    double total = 0;
    for (int i = 0; i < n; i++) {
        if (y_bounds[i + 1] < y_bounds[i] || y_bounds[0] != 0 || !(ms[i] >= 0.f)) return fail(ER_ERR_INVALID, "er_rebalance_stripes: bad stripes or times");
        total += ms[i];
    }
    std::vector<int> out((size_t)n + 1);
    out[0] = 0, out[(size_t)n] = y_res;
    if (!(total > 0)) {
        memcpy(y_bounds_out, y_bounds, sizeof(int) * ((size_t)n + 1));
        return ER_OK;
    }
    int i = 0;
    double before = 0;  // time of the stripes before stripe i
    for (int k = 1; k < n; k++) {
        const double target = total * k / n;
        while (i < n - 1 && before + ms[i] < target) before += ms[i++];
        const int rows = y_bounds[i + 1] - y_bounds[i];
        double y = y_bounds[i] + (ms[i] > 0 ? (target - before) / ms[i] * rows : 0.0);
        // over-relaxation: a third to a half of a stripe's time does not move with its rows (measured on the 8-way
        // split of C3), so the proportional step covers only 0.5-0.7 of the way; 1.5 x stays below 2 x (no growth)
        // even for stripes whose time is all proportional
        y = y_bounds[k] + ER_REBALANCE_GAIN * (y - y_bounds[k]);
        int yy = (int)floor(y / rows_per_block + 0.5) * rows_per_block;
        // every stripe keeps at least one row block
        const int lo = out[(size_t)k - 1] + rows_per_block, hi = y_res - (n - k) * rows_per_block;
        if (yy > hi) yy = hi;
        if (yy < lo) yy = lo;
        if (yy > y_res) yy = y_res;
        out[(size_t)k] = yy;
    }
    memcpy(y_bounds_out, out.data(), sizeof(int) * ((size_t)n + 1));
    return ER_OK;
}

int er_get_pixel_state(ErCtx* c, uint32_t* rng_out, uint32_t* samples_out) {
    int rc = use(c);
    if (rc) return rc;
    if (!c->uploaded) return fail(ER_ERR_INVALID, "no scene uploaded");
    const size_t P = (size_t)c->S.x_res * c->S.y_res;
    std::lock_guard<std::mutex> lk(c->batch_mutex);
    // column x / y of the 16-B pixel records
    if (rng_out) ER_CK(cudaMemcpy2DAsync(rng_out, 4, &c->S.pix[0].x, 16, 4, P, cudaMemcpyDeviceToHost, c->stream));
    if (samples_out) ER_CK(cudaMemcpy2DAsync(samples_out, 4, &c->S.pix[0].y, 16, 4, P, cudaMemcpyDeviceToHost, c->stream));
    ER_CK(cudaStreamSynchronize(c->stream));
    return ER_OK;
}

int er_get_counters(ErCtx* c, ErCounters* out) {
    int rc = use(c);
    if (rc) return rc;
    if (!out) return fail(ER_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> lk(c->batch_mutex);
    ER_CK(cudaMemcpyAsync(out, c->d_counters, sizeof *out, cudaMemcpyDeviceToHost, c->stream));
    ER_CK(cudaStreamSynchronize(c->stream));
    return ER_OK;
}

int er_reset_counters(ErCtx* c) {
    int rc = use(c);
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(c->batch_mutex);
    ER_CK(cudaMemsetAsync(c->d_counters, 0, sizeof(DCounters), c->stream));
    rc = fold_events(c);
    c->kernel_ms = 0.f;
    c->launches = 0;
    return rc;
}

int er_get_kernel_time(ErCtx* c, float* ms, int* launches) {
    int rc = use(c);
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(c->batch_mutex);
    rc = fold_events(c);
    if (rc) return rc;
    if (ms) *ms = c->kernel_ms;
    if (launches) *launches = c->launches;
    return ER_OK;
}

}  // extern "C"

// ---- nearest-hit queries and device-function test hooks ------------------------------------------
namespace {

__global__ void k_test_rng(int idx, int n, uint32_t* states, float* values) {
    uint32_t s = er_rng_seed((uint32_t)idx);
    for (int i = 0; i < n; i++) {
        values[i] = er_rng_next(s);
        states[i] = s;
    }
}

__global__ void k_test_texture(const __grid_constant__ DScene S, int tex, int mode, const float* ab, size_t n, float* out3) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const DTexture T = tex < 0 ? S.hdri : S.textures[tex];
    const float a = ab[2 * i], b = ab[2 * i + 1];
    V3 v;
    if (mode == 0)
        v = er_tex_coords(T, er_f2i(a), er_f2i(b));
    else if (mode == 1)
        v = er_tex_uv(T, a, b);
    else if (mode == 2)
        v = er_tex_bilinear(T, a, b);
    else
        v = er_tex_filtered(T, a, b);
    out3[3 * i] = v.x, out3[3 * i + 1] = v.y, out3[3 * i + 2] = v.z;
}

// fn: 0 sin, 1 cos, 2 acos, 3 atan2(a, b), 4 pow(a, b), 5 log
__global__ void k_test_libm(int fn, const float* a, const float* b, size_t n, float* out) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float x = a[i], y = b ? b[i] : 0.f;
    float r = 0.f;
    switch (fn) {
        case 0: r = er_sinf(x); break;
        case 1: r = er_cosf(x); break;
        case 2: r = er_acosf(x); break;
        case 3: r = er_atan2f(x, y); break;
        case 4: r = er_powf(x, y); break;
        default: r = er_logf(x); break;
    }
    out[i] = r;
}

__global__ void k_test_spherical(const float* dirs3, size_t n, float* uv2) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    float u, v;
    er_spherical_mapping(v3(dirs3[3 * i], dirs3[3 * i + 1], dirs3[3 * i + 2]), u, v);
    uv2[2 * i] = u, uv2[2 * i + 1] = v;
}

// the env NEE chain of kernel.cpp:543-553,558,565
__global__ void k_test_hdri_sample(const __grid_constant__ DScene S, const float* r, size_t n, float* out10, float* pdf) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    int count = er_hdri_search(S.cdf, r[i], S.hdri.width * S.hdri.height);
    float tcx = (float)(count % S.hdri.width), tcy = (float)(count / S.hdri.width);
    float nu = tcx / (float)S.hdri.width, nv = tcy / (float)S.hdri.height;
    float iu, iv;
    er_inverse_transform_uv(S.hdri, nu, nv, iu, iv);
    V3 wi = -normalized(er_reverse_spherical(iu, iv));
    V3 val = er_tex_uv(S.hdri, iu, iv);
    float* o = out10 + 10 * i;
    o[0] = tcx, o[1] = tcy, o[2] = iu, o[3] = iv, o[4] = wi.x, o[5] = wi.y, o[6] = wi.z, o[7] = val.x, o[8] = val.y, o[9] = val.z;
    pdf[i] = er_hdri_pdf(S, er_f2i(iu * (float)S.hdri.width), er_f2i(iv * (float)S.hdri.height));
}

__global__ void k_test_disney(const float* hitdata37, const float* V3p, const float* L3p, const float* r3, size_t n,
                              float* sample3, float* eval3, float* pdf) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    HitData hd;
    float* w = (float*)&hd;
    for (int k = 0; k < 37; k++) w[k] = hitdata37[37 * i + k];
    V3 V = v3(V3p[3 * i], V3p[3 * i + 1], V3p[3 * i + 2]);
    V3 L = v3(L3p[3 * i], L3p[3 * i + 1], L3p[3 * i + 2]);
    V3 s = er_disney_sample(hd, V, hd.normal, r3[3 * i], r3[3 * i + 1], r3[3 * i + 2]);
    V3 e = er_disney_eval(hd, V, hd.normal, L);
    pdf[i] = er_disney_pdf(hd, V, hd.normal, L);
    sample3[3 * i] = s.x, sample3[3 * i + 1] = s.y, sample3[3 * i + 2] = s.z;
    eval3[3 * i] = e.x, eval3[3 * i + 1] = e.y, eval3[3 * i + 2] = e.z;
}

__global__ void k_test_hitdata(const __grid_constant__ DScene S, const DHit* hits, size_t n, float* hitdata37) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const DHit& h = hits[i];
    HitFull hf;
    hf.position = v3(h.position[0], h.position[1], h.position[2]);
    hf.normal = v3(h.normal[0], h.normal[1], h.normal[2]);
    hf.tangent = v3(h.tangent[0], h.tangent[1], h.tangent[2]);
    hf.bitangent = v3(h.bitangent[0], h.bitangent[1], h.bitangent[2]);
    hf.gnormal = v3(h.gnormal[0], h.gnormal[1], h.gnormal[2]);
    hf.tu = h.tu, hf.tv = h.tv;
    hf.objectID = h.objectID, hf.material_id = h.material_id, hf.triIdx = h.triIdx, hf.valid = true;
    HitData hd;
    memset(&hd, 0, sizeof hd);
    er_generate_hitdata(S, S.materials[hf.material_id], hf, hd);
    const float* w = (const float*)&hd;
    for (int k = 0; k < 37; k++) hitdata37[37 * i + k] = w[k];
}

__global__ void k_camera_rays(const __grid_constant__ DScene S, const int* xy, const float* r5, size_t n, float* rays) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    V3 o, d;
    er_camera_ray(S, xy[2 * i], xy[2 * i + 1], r5[5 * i], r5[5 * i + 1], r5[5 * i + 2], r5[5 * i + 3], r5[5 * i + 4], o, d);
    float* q = rays + 6 * i;
    q[0] = o.x, q[1] = o.y, q[2] = o.z, q[3] = d.x, q[4] = d.y, q[5] = d.z;
}

// small RAII scratch buffers for the hooks
struct Scratch {
    std::vector<void*> ptrs;
    ~Scratch() {
        for (void* p : ptrs) cudaFree(p);
    }
    template <typename T>
    int in(const T* host, size_t n, T** out, cudaStream_t st) {
        void* p = nullptr;
        ER_CK(cudaMalloc(&p, sizeof(T) * (n ? n : 1)));
        ptrs.push_back(p);
        if (n && host) ER_CK(cudaMemcpyAsync(p, host, sizeof(T) * n, cudaMemcpyHostToDevice, st));
        *out = (T*)p;
        return ER_OK;
    }
};

int trace_launch(ErCtx* c, const float* d_rays, size_t n, DHit* d_hits, int mode) {
    if (n == 0) return ER_OK;
    const int grid = grid_for(c, n, 128);
    if (mode == 0)
        k_trace<false, true><<<grid, 128, 0, c->stream>>>(c->S, d_rays, n, d_hits, c->d_counters);
    else if (mode == 1)
        k_trace<true, true><<<grid, 128, 0, c->stream>>>(c->S, d_rays, n, d_hits, c->d_counters);
    else if (mode == 2)
        k_trace<false, false><<<grid, 128, 0, c->stream>>>(c->S, d_rays, n, d_hits, c->d_counters);
    else
        k_trace<true, false><<<grid, 128, 0, c->stream>>>(c->S, d_rays, n, d_hits, c->d_counters);
    ER_CK(cudaGetLastError());
    return ER_OK;
}

int ready(ErCtx* c) {
    int rc = use(c);
    if (rc) return rc;
    if (!c->uploaded) return fail(ER_ERR_INVALID, "no scene uploaded");
    if (c->running) return fail(ER_ERR_INVALID, "render in progress");
    return ER_OK;
}

#define ER_OUT(dst, src, count) ER_CK(cudaMemcpyAsync(dst, src, sizeof(*(dst)) * (count), cudaMemcpyDeviceToHost, c->stream))

}  // namespace

extern "C" {

// mode: bit0 = count stats, bit1 = reference visit order without pruning
int er_trace_mode(ErCtx* c, const float* rays, size_t n, ErHit* hits_out, int mode) {
    NvtxRange nvtx_("er_trace");
    int rc = ready(c);
    if (rc) return rc;
    if (n && (!rays || !hits_out)) return fail(ER_ERR_INVALID, "null argument");
    Scratch s;
    float* d_rays;
    DHit* d_hits;
    if ((rc = s.in<float>(rays, 6 * n, &d_rays, c->stream))) return rc;
    if ((rc = s.in<DHit>(nullptr, n, &d_hits, c->stream))) return rc;
    if ((rc = trace_launch(c, d_rays, n, d_hits, mode))) return rc;
    if (n) ER_CK(cudaMemcpyAsync(hits_out, d_hits, sizeof(DHit) * n, cudaMemcpyDeviceToHost, c->stream));
    ER_CK(cudaStreamSynchronize(c->stream));
    return ER_OK;
}

int er_trace(ErCtx* c, const float* rays, size_t n, ErHit* hits_out) { return er_trace_mode(c, rays, n, hits_out, 0); }

int er_trace_device(ErCtx* c, const void* d_rays, size_t n, void* d_hits) {
    int rc = ready(c);
    if (rc) return rc;
    if (n && (!d_rays || !d_hits)) return fail(ER_ERR_INVALID, "null argument");
    std::pair<cudaEvent_t, cudaEvent_t> ev;
    if ((rc = get_event_pair(c, ev))) return rc;
    ER_CK(cudaEventRecord(ev.first, c->stream));
    if ((rc = trace_launch(c, (const float*)d_rays, n, (DHit*)d_hits, 0))) return rc;
    ER_CK(cudaEventRecord(ev.second, c->stream));
    c->ev_live.push_back(ev);
    c->launches++;
    return ER_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------
// er_trace_noalloc: er_trace for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
// loops, harnesses that forbid it). No counterpart in the reference.
// The device staging that er_trace allocates per call (the rays the kernel reads, the hit records it writes) lives
// here in two file-scope device arrays of ER_TRACE_NOALLOC_MAX entries (6.75 MiB per GPU that loads the library,
// held until exit). In host code `g_na_rays` / `g_na_hits` are host-side stand-ins, not device addresses: the
// addresses the copies and the kernel get come from cudaGetSymbolAddress, asked once per device with that device
// current. Every call shares the arrays and they keep the last call's contents: the hit records of a call are
// cleared before its kernel runs; calls that overlap on the GPU (two streams, two contexts of one device, two
// threads) race on them. (The arrays and the per-device address cache are defined above er_create, which resolves
// the addresses at set-up; the entry point resolves them itself if that did not happen.)
// ---------------------------------------------------------------------------------------------------
extern "C" {

int er_trace_noalloc(ErCtx* c, const float* rays, size_t n, ErHit* hits_out) {
    int rc = ready(c);
    if (rc) return rc;
    if (n > (size_t)ER_TRACE_NOALLOC_MAX)
        return fail(ER_ERR_INVALID, "er_trace_noalloc: " + std::to_string(n) + " rays exceed ER_TRACE_NOALLOC_MAX (" +
                                        std::to_string((int)ER_TRACE_NOALLOC_MAX) + "); split the batch or call er_trace");
    if (n == 0) return ER_OK;
    if (!rays || !hits_out) return fail(ER_ERR_INVALID, "null argument");
    NoAllocScratch sc;
    if ((rc = noalloc_scratch(c->device, sc))) return rc;  // cached by er_create; asked here if that failed
    ER_CK(cudaMemcpyAsync(sc.rays, rays, sizeof(float) * 6 * n, cudaMemcpyHostToDevice, c->stream));
    ER_CK(cudaMemsetAsync(sc.hits, 0, sizeof(DHit) * n, c->stream));
    if ((rc = trace_launch(c, sc.rays, n, sc.hits, 0))) return rc;
    ER_CK(cudaMemcpyAsync(hits_out, sc.hits, sizeof(DHit) * n, cudaMemcpyDeviceToHost, c->stream));
    ER_CK(cudaStreamSynchronize(c->stream));
    return ER_OK;
}

// test hook: the device-resident geometry arrays as uploaded (0 pair records, 1 tri_v, 2 tri_n, 3 tri_a), n_floats at most
int er_test_geometry(ErCtx* c, int which, float* out, size_t n_floats, size_t* total_floats) {
    int rc = ready(c);
    if (rc) return rc;
    const size_t T = (size_t)c->S.tri_count;
    const float4* src = which == 0 ? c->S.pairs : which == 1 ? c->S.tri_v : which == 2 ? c->S.tri_n : which == 3 ? c->S.tri_a : nullptr;
    const size_t total = which == 0 ? c->pairs_bytes / 4 : which == 1 || which == 2 ? 12 * T : which == 3 ? 20 * T : 0;
    if (!src) return fail(ER_ERR_INVALID, "er_test_geometry: which must be 0..3");
    if (total_floats) *total_floats = total;
    const size_t n = n_floats < total ? n_floats : total;
    if (n && out) {
        ER_CK(cudaMemcpyAsync(out, src, sizeof(float) * n, cudaMemcpyDeviceToHost, c->stream));
        ER_CK(cudaStreamSynchronize(c->stream));
    }
    return ER_OK;
}

int er_camera_rays(ErCtx* c, const int32_t* xy, const float* r5, size_t n, float* rays_out) {
    int rc = ready(c);
    if (rc) return rc;
    Scratch s;
    int* d_xy;
    float *d_r, *d_o;
    if ((rc = s.in<int>(xy, 2 * n, &d_xy, c->stream))) return rc;
    if ((rc = s.in<float>(r5, 5 * n, &d_r, c->stream))) return rc;
    if ((rc = s.in<float>(nullptr, 6 * n, &d_o, c->stream))) return rc;
    if (n) k_camera_rays<<<(unsigned)((n + 127) / 128), 128, 0, c->stream>>>(c->S, d_xy, d_r, n, d_o);
    ER_CK(cudaGetLastError());
    if (n) ER_OUT(rays_out, d_o, 6 * n);
    ER_CK(cudaStreamSynchronize(c->stream));
    return ER_OK;
}

int er_test_rng(ErCtx* c, int idx, int n, uint32_t* states, float* values) {
    int rc = use(c);
    if (rc) return rc;
    Scratch s;
    uint32_t* d_s;
    float* d_v;
    if ((rc = s.in<uint32_t>(nullptr, n, &d_s, c->stream))) return rc;
    if ((rc = s.in<float>(nullptr, n, &d_v, c->stream))) return rc;
    k_test_rng<<<1, 1, 0, c->stream>>>(idx, n, d_s, d_v);
    ER_CK(cudaGetLastError());
    if (n > 0) {
        ER_OUT(states, d_s, n);
        ER_OUT(values, d_v, n);
    }
    ER_CK(cudaStreamSynchronize(c->stream));
    return ER_OK;
}

int er_test_texture(ErCtx* c, int tex, int mode, const float* ab, size_t n, float* out3) {
    int rc = ready(c);
    if (rc) return rc;
    if (tex >= c->S.texture_count) return fail(ER_ERR_INVALID, "texture index out of range");
    Scratch s;
    float *d_ab, *d_o;
    if ((rc = s.in<float>(ab, 2 * n, &d_ab, c->stream))) return rc;
    if ((rc = s.in<float>(nullptr, 3 * n, &d_o, c->stream))) return rc;
    if (n) k_test_texture<<<(unsigned)((n + 127) / 128), 128, 0, c->stream>>>(c->S, tex, mode, d_ab, n, d_o);
    ER_CK(cudaGetLastError());
    if (n) ER_OUT(out3, d_o, 3 * n);
    ER_CK(cudaStreamSynchronize(c->stream));
    return ER_OK;
}

int er_test_libm(ErCtx* c, int fn, const float* a, const float* b, size_t n, float* out) {
    int rc = use(c);
    if (rc) return rc;
    if (fn < 0 || fn > 5 || ((fn == 3 || fn == 4) && !b)) return fail(ER_ERR_INVALID, "er_test_libm: bad function or missing operand");
    Scratch s;
    float *d_a, *d_b = nullptr, *d_o;
    if ((rc = s.in<float>(a, n, &d_a, c->stream))) return rc;
    if (b && (rc = s.in<float>(b, n, &d_b, c->stream))) return rc;
    if ((rc = s.in<float>(nullptr, n, &d_o, c->stream))) return rc;
    if (n) k_test_libm<<<(unsigned)((n + 127) / 128), 128, 0, c->stream>>>(fn, d_a, d_b, n, d_o);
    ER_CK(cudaGetLastError());
    if (n) ER_OUT(out, d_o, n);
    ER_CK(cudaStreamSynchronize(c->stream));
    return ER_OK;
}

int er_test_spherical(ErCtx* c, const float* dirs3, size_t n, float* uv2) {
    int rc = use(c);
    if (rc) return rc;
    Scratch s;
    float *d_d, *d_o;
    if ((rc = s.in<float>(dirs3, 3 * n, &d_d, c->stream))) return rc;
    if ((rc = s.in<float>(nullptr, 2 * n, &d_o, c->stream))) return rc;
    if (n) k_test_spherical<<<(unsigned)((n + 127) / 128), 128, 0, c->stream>>>(d_d, n, d_o);
    ER_CK(cudaGetLastError());
    if (n) ER_OUT(uv2, d_o, 2 * n);
    ER_CK(cudaStreamSynchronize(c->stream));
    return ER_OK;
}

int er_test_hdri_sample(ErCtx* c, const float* r, size_t n, float* out10, float* pdf) {
    int rc = ready(c);
    if (rc) return rc;
    Scratch s;
    float *d_r, *d_o, *d_p;
    if ((rc = s.in<float>(r, n, &d_r, c->stream))) return rc;
    if ((rc = s.in<float>(nullptr, 10 * n, &d_o, c->stream))) return rc;
    if ((rc = s.in<float>(nullptr, n, &d_p, c->stream))) return rc;
    if (n) k_test_hdri_sample<<<(unsigned)((n + 127) / 128), 128, 0, c->stream>>>(c->S, d_r, n, d_o, d_p);
    ER_CK(cudaGetLastError());
    if (n) {
        ER_OUT(out10, d_o, 10 * n);
        ER_OUT(pdf, d_p, n);
    }
    ER_CK(cudaStreamSynchronize(c->stream));
    return ER_OK;
}

int er_test_disney(ErCtx* c, const float* hitdata37, const float* V3p, const float* L3p, const float* r3, size_t n,
                   float* sample3, float* eval3, float* pdf) {
    int rc = use(c);
    if (rc) return rc;
    Scratch s;
    float *d_h, *d_v, *d_l, *d_r, *d_s, *d_e, *d_p;
    if ((rc = s.in<float>(hitdata37, 37 * n, &d_h, c->stream))) return rc;
    if ((rc = s.in<float>(V3p, 3 * n, &d_v, c->stream))) return rc;
    if ((rc = s.in<float>(L3p, 3 * n, &d_l, c->stream))) return rc;
    if ((rc = s.in<float>(r3, 3 * n, &d_r, c->stream))) return rc;
    if ((rc = s.in<float>(nullptr, 3 * n, &d_s, c->stream))) return rc;
    if ((rc = s.in<float>(nullptr, 3 * n, &d_e, c->stream))) return rc;
    if ((rc = s.in<float>(nullptr, n, &d_p, c->stream))) return rc;
    if (n) k_test_disney<<<(unsigned)((n + 127) / 128), 128, 0, c->stream>>>(d_h, d_v, d_l, d_r, n, d_s, d_e, d_p);
    ER_CK(cudaGetLastError());
    if (n) {
        ER_OUT(sample3, d_s, 3 * n);
        ER_OUT(eval3, d_e, 3 * n);
        ER_OUT(pdf, d_p, n);
    }
    ER_CK(cudaStreamSynchronize(c->stream));
    return ER_OK;
}

int er_test_hitdata(ErCtx* c, const ErHit* hits, size_t n, float* hitdata37) {
    int rc = ready(c);
    if (rc) return rc;
    for (size_t i = 0; i < n; i++)
        if ((int)hits[i].material_id >= c->S.material_count) return fail(ER_ERR_INVALID, "material_id out of range");
    Scratch s;
    DHit* d_h;
    float* d_o;
    if ((rc = s.in<DHit>((const DHit*)hits, n, &d_h, c->stream))) return rc;
    if ((rc = s.in<float>(nullptr, 37 * n, &d_o, c->stream))) return rc;
    if (n) k_test_hitdata<<<(unsigned)((n + 127) / 128), 128, 0, c->stream>>>(c->S, d_h, n, d_o);
    ER_CK(cudaGetLastError());
    if (n) ER_OUT(hitdata37, d_o, 37 * n);
    ER_CK(cudaStreamSynchronize(c->stream));
    return ER_OK;
}

}  // extern "C"

// ---- page-locked host memory ------------------------------------------------------------------------
extern "C" {

int er_host_alloc(void** ptr, size_t bytes) {
    if (!ptr) return fail(ER_ERR_INVALID, "null argument");
    *ptr = nullptr;
    ER_CK(cudaHostAlloc(ptr, bytes ? bytes : 1, cudaHostAllocPortable));
    return ER_OK;
}

int er_host_free(void* ptr) {
    if (ptr) ER_CK(cudaFreeHost(ptr));
    return ER_OK;
}

int er_host_register(void* ptr, size_t bytes) {
    if (!ptr || !bytes) return fail(ER_ERR_INVALID, "null argument");
    ER_CK(cudaHostRegister(ptr, bytes, cudaHostRegisterPortable));
    return ER_OK;
}

int er_host_unregister(void* ptr) {
    if (ptr) ER_CK(cudaHostUnregister(ptr));
    return ER_OK;
}

}  // extern "C"

// ---- device groups (multi-GPU from one host thread) ------------------------------------
struct ErGroup {
    std::vector<ErCtx*> m;
    int rows_per_block = 8;
    std::vector<int> bounds;  // stripes installed by er_group_balance / er_group_rebalance (empty: none)
    std::thread worker;
    std::atomic<bool> abort{false};
    std::atomic<bool> running{false};
    std::atomic<unsigned> completed{0};
    std::atomic<int> worker_status{0};
    std::string worker_error;
};

namespace {

void group_join(ErGroup* g) {
    if (g->worker.joinable()) g->worker.join();
}

// every member's launches are enqueued before any of them is waited for: the GPUs run side by side
int group_batches(ErGroup* g, unsigned samples, const ErRenderOpts& o, bool sync_each) {
    unsigned left = samples;
    while (left > 0) {
        if (g->abort.load()) return fail(ER_ERR_ABORTED, "render aborted");
        const unsigned n = left < (unsigned)o.spp_per_launch ? left : (unsigned)o.spp_per_launch;
        for (ErCtx* c : g->m) {
            int rc = use(c);
            if (rc) return rc;
            std::lock_guard<std::mutex> lk(c->batch_mutex);
            if ((rc = launch_render(c, (int)n, o))) return rc;
        }
        if (sync_each || left == n)
            for (ErCtx* c : g->m) {
                int rc = use(c);
                if (rc) return rc;
                ER_CK(cudaStreamSynchronize(c->stream));
                if ((rc = check_watchdog(c))) return rc;
            }
        for (ErCtx* c : g->m) c->completed += n;
        g->completed += n;
        left -= n;
    }
    return ER_OK;
}

int group_ok(ErGroup* g) {
    if (!g || g->m.empty()) return fail(ER_ERR_INVALID, "null group");
    if (g->running) return fail(ER_ERR_INVALID, "render in progress");
    return ER_OK;
}

}  // namespace

extern "C" {

int er_group_create(const int* device_ids, int n, ErGroup** out) {
    if (!device_ids || n < 1 || !out) return fail(ER_ERR_INVALID, "bad argument");
    *out = nullptr;
    ErGroup* g = new ErGroup();
    for (int i = 0; i < n; i++) {
        ErCtx* c = nullptr;
        int rc = er_create(device_ids[i], &c);
        if (!rc) rc = er_set_shard(c, i, n, g->rows_per_block);
        if (rc) {
            const std::string keep = g_err;
            if (c) er_destroy(c);
            er_group_destroy(g);
            return fail(rc, keep);
        }
        g->m.push_back(c);
    }
    *out = g;
    return ER_OK;
}

int er_group_destroy(ErGroup* g) {
    if (!g) return ER_OK;
    g->abort = true;
    group_join(g);
    for (ErCtx* c : g->m) er_destroy(c);
    delete g;
    return ER_OK;
}

int er_group_size(ErGroup* g, int* n) {
    if (!g || !n) return fail(ER_ERR_INVALID, "null argument");
    *n = (int)g->m.size();
    return ER_OK;
}

int er_group_ctx(ErGroup* g, int member, ErCtx** ctx) {
    if (!g || !ctx || member < 0 || member >= (int)g->m.size()) return fail(ER_ERR_INVALID, "bad member");
    *ctx = g->m[(size_t)member];
    return ER_OK;
}

int er_group_upload_scene(ErGroup* g, const ErSceneDesc* d) {
    NvtxRange nvtx_("er_group_upload_scene");
    int rc = group_ok(g);
    if (rc) return rc;
    if ((rc = validate_desc(d))) return rc;
    ErCtx* c0 = g->m[0];
    if ((rc = use(c0))) return rc;
    if ((rc = stage_flat(c0, d, c0->flat))) return rc;  // flattened ONCE, staged (and pinned) with member 0
    for (ErCtx* c : g->m) {  // H2D copies of all members in flight together, one PCIe link each
        if ((rc = use(c))) return rc;
        if ((rc = upload_flat(c, d, c0->flat))) return rc;
    }
    for (ErCtx* c : g->m) {
        if ((rc = use(c))) return rc;
        ER_CK(cudaStreamSynchronize(c->stream));
    }
    g->completed = 0;
    return ER_OK;
}

int er_group_balance(ErGroup* g, const ErRenderOpts* opts, int* y_bounds_out) {
    int rc = group_ok(g);
    if (rc) return rc;
    const int n = (int)g->m.size();
    if (!g->m[0]->uploaded) return fail(ER_ERR_INVALID, "no scene uploaded");
    const int y_res = g->m[0]->S.y_res;
    // member i samples the row blocks 4i, 4i + 4n, ...: together every 4th block of the frame, 1 spp
    std::vector<uint32_t> cost((size_t)y_res, 0u), part((size_t)y_res);
    for (int i = 0; i < n; i++) {
        // (sequential: each prepass is 1/(4n) of a 1-spp frame)
        ErCtx* c = g->m[(size_t)i];
        c->rows_per_block = g->rows_per_block;
        if ((rc = er_row_cost(c, 4 * i, 4 * n, opts, part.data()))) return rc;
        for (int y = 0; y < y_res; y++) cost[(size_t)y] += part[(size_t)y];
    }
    std::vector<int> yb((size_t)n + 1);
    if ((rc = er_balance_stripes(cost.data(), y_res, g->rows_per_block, n, yb.data()))) return rc;
    for (int i = 0; i < n; i++)
        if ((rc = er_set_shard_stripe(g->m[(size_t)i], yb[(size_t)i], yb[(size_t)i + 1], g->rows_per_block))) return rc;
    if (y_bounds_out) memcpy(y_bounds_out, yb.data(), sizeof(int) * ((size_t)n + 1));
    g->bounds = yb;
    g->completed = 0;
    return ER_OK;
}

// Feedback balance: new stripes from the kernel time every GPU spent since its counters were last reset (the frame
// or the progressive batches just rendered); restarts accumulation like er_group_balance.
int er_group_rebalance(ErGroup* g, int* y_bounds_out) {
    int rc = group_ok(g);
    if (rc) return rc;
    const int n = (int)g->m.size();
    if (!g->m[0]->uploaded) return fail(ER_ERR_INVALID, "no scene uploaded");
    if ((int)g->bounds.size() != n + 1) return fail(ER_ERR_INVALID, "er_group_rebalance: the group is not sharded by stripes (er_group_balance first)");
    std::vector<int> yb(g->bounds), nb((size_t)n + 1);
    std::vector<float> ms((size_t)n);
    for (int i = 0; i < n; i++) {
        int launches = 0;
        if ((rc = er_get_kernel_time(g->m[(size_t)i], &ms[(size_t)i], &launches))) return rc;
        if (launches == 0 && yb[(size_t)i + 1] > yb[(size_t)i])
            return fail(ER_ERR_INVALID, "er_group_rebalance: nothing rendered since the counters were reset");
    }
    if ((rc = er_rebalance_stripes(yb.data(), ms.data(), n, g->rows_per_block, nb.data()))) return rc;
    for (int i = 0; i < n; i++) {
        ErCtx* c = g->m[(size_t)i];
        if ((rc = er_set_shard_stripe(c, nb[(size_t)i], nb[(size_t)i + 1], g->rows_per_block))) return rc;
        if ((rc = er_reset_counters(c))) return rc;
    }
    if (y_bounds_out) memcpy(y_bounds_out, nb.data(), sizeof(int) * ((size_t)n + 1));
    g->bounds = nb;
    g->completed = 0;
    return ER_OK;
}

int er_group_render_reset(ErGroup* g) {
    int rc = group_ok(g);
    if (rc) return rc;
    for (ErCtx* c : g->m)
        if ((rc = er_render_reset(c))) return rc;
    g->completed = 0;
    return ER_OK;
}

int er_group_render_samples(ErGroup* g, unsigned samples, const ErRenderOpts* opts) {
    int rc = group_ok(g);
    if (rc) return rc;
    g->abort = false;
    return group_batches(g, samples, norm_opts(opts), false);
}

int er_group_render_start(ErGroup* g, unsigned target_samples, const ErRenderOpts* opts) {
    int rc = group_ok(g);
    if (rc) return rc;
    group_join(g);
    const ErRenderOpts o = norm_opts(opts);
    g->abort = false;
    g->running = true;
    g->worker_status = 0;
    g->worker = std::thread([g, target_samples, o]() {
        int rc2 = group_batches(g, target_samples, o, true);
        if (rc2 && rc2 != ER_ERR_ABORTED) g->worker_error = g_err;
        g->worker_status = rc2;
        g->running = false;
    });
    return ER_OK;
}

int er_group_render_poll(ErGroup* g, unsigned* samples, int* running) {
    if (!g) return fail(ER_ERR_INVALID, "null group");
    if (samples) *samples = g->completed.load() + 1;
    if (running) *running = g->running.load() ? 1 : 0;
    return ER_OK;
}

int er_group_render_wait(ErGroup* g) {
    if (!g) return fail(ER_ERR_INVALID, "null group");
    group_join(g);
    const int st = g->worker_status.load();
    if (st == ER_ERR_ABORTED) return fail(st, "render aborted");
    if (st) return fail(st, g->worker_error);
    return ER_OK;
}

int er_group_render_abort(ErGroup* g) {
    if (!g) return fail(ER_ERR_INVALID, "null group");
    g->abort = true;
    return ER_OK;
}

int er_group_get_render_info(ErGroup* g, unsigned* samples) {
    if (!g || g->m.empty() || !samples) return fail(ER_ERR_INVALID, "null argument");
    // a frame has n samples when every shard has n (get_render_info reads dev_samples[0], Managers.cpp:217)
    unsigned lo = 0xffffffffu;
    for (ErCtx* c : g->m) {
        const unsigned v = c->completed.load() + 1;
        if (v < lo) lo = v;
    }
    *samples = lo;
    return ER_OK;
}

int er_group_get_pass(ErGroup* g, int pass, float* out_rgba) {
    NvtxRange nvtx_("er_group_get_pass");
    if (!g || g->m.empty()) return fail(ER_ERR_INVALID, "null group");
    if (!out_rgba || pass < 0 || pass >= ER_PASSES_COUNT) return fail(ER_ERR_INVALID, "bad pass");
    // every member copies the rows it owns straight into the caller's frame: all PCIe links at once, 1/n of
    // the bytes each. One host thread per member because a copy into pageable memory blocks its caller.
    const size_t n = g->m.size();
    std::vector<int> rcs(n, 0);
    std::vector<std::string> errs(n);
    std::vector<std::thread> th;
    for (size_t i = 1; i < n; i++)
        th.emplace_back([&, i]() {
            rcs[i] = er_get_pass_owned(g->m[i], pass, out_rgba);
            if (rcs[i]) errs[i] = g_err;
        });
    rcs[0] = er_get_pass_owned(g->m[0], pass, out_rgba);
    if (rcs[0]) errs[0] = g_err;
    for (auto& t : th) t.join();
    for (size_t i = 0; i < n; i++)
        if (rcs[i]) return fail(rcs[i], errs[i]);
    return ER_OK;
}

}  // extern "C"
