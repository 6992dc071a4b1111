/*
 * er_cuda.h — C ABI of the B200 (sm_100a) path-tracing hot path of ElevenRender.
 *
 * This is the drop-in boundary: it takes the place of the reference's SYCL device layer
 *   - src/sycl.h, src/SYCLCopy.{h,cpp}          (scene deep-copy to the device)
 *   - renderSetup / kernel_render_enqueue        (src/kernel.h:124-125, src/kernel.cpp:651-706)
 *   - renderingKernel / setupKernel              (src/kernel.h:126, src/kernel.cpp:176-213,477-646)
 *   - RenderingManager::get_pass/get_render_info (src/Managers.cpp:211-232,287-302)
 *   - is_compatible / get_sycl_info device list  (src/kernel.cpp:708-720, src/CommandManager.cpp:303-362)
 * Plain pointers and sizes only; no exceptions cross it. Every function returns ER_OK (0) or a
 * negative ErStatus; er_last_error() gives the thread-local message. There is NO CPU fallback:
 * without a CUDA device every compute entry point fails with ER_ERR_NO_DEVICE.
 *
 * Record layouts are the leading POD words of the reference structs so a maintainer can pass the
 * reference's own arrays with a stride.
 */
#ifndef ER_CUDA_H
#define ER_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ER_API __attribute__((visibility("default")))

typedef enum ErStatus {
    ER_OK = 0,
    ER_ERR_INVALID = -1,   /* bad argument / call order */
    ER_ERR_NO_DEVICE = -2, /* no usable CUDA device (never falls back to the CPU) */
    ER_ERR_CUDA = -3,      /* CUDA runtime error, message in er_last_error() */
    ER_ERR_OOM = -4,
    ER_ERR_ABORTED = -5
} ErStatus;

/* Reference constants (src/Definitions.h:4-13, src/kernel.h:8) */
#define ER_BVH_DEPTH 18
#define ER_BVH_SAHBINS 14
#define ER_BVH_NODE_COUNT ((2 << ER_BVH_DEPTH) - 1) /* 524287 pre-order nodes, leaves at depth 18 */
#define ER_PASSES_COUNT 5
enum ErPass { ER_PASS_BEAUTY = 0, ER_PASS_DENOISE = 1, ER_PASS_NORMAL = 2, ER_PASS_TANGENT = 3, ER_PASS_BITANGENT = 4 };

/* RNG argument-evaluation order of the two call sites that draw inside an argument list
 * (src/kernel.cpp:492-493 and :545): left-to-right is what DPC++/clang (the shipped build) does,
 * right-to-left is what a g++ build of the reference does. */
enum ErRngOrder { ER_RNG_LTR = 0, ER_RNG_RTL = 1 };

/* ---- records -------------------------------------------------------------------------------- */

/* Leading 156 bytes of the reference `Tri` (src/Tri.h:11-20; sizeof(Tri)==192 with its std::string). */
typedef struct ErTri {
    float vertices[3][3];
    float uv[3][3];
    float normals[3][3];
    float tangents[3][3];
    float tangentsSign;
    int32_t objectID;
    int32_t materialID;
} ErTri; /* 39 words */

/* The reference `Node` (src/BVH.h:24-33), 44 bytes. */
typedef struct ErNode {
    float b1[3], b2[3];
    int32_t from, to, idx, depth;
    int32_t valid; /* bool + padding in the reference */
} ErNode;

/* POD view of the reference `Material` without its std::string members (src/Material.h:21-47). */
typedef struct ErMaterial {
    int32_t albedoTextureID, emissionTextureID, roughnessTextureID, metallicTextureID, normalTextureID,
        opacityTextureID, transmissionTextureID, albedoShaderID;
    float albedo[3];
    float emission[3];
    float opacity, roughness, metallic, clearcoatGloss, clearcoat, anisotropic, eta, transmission, specular,
        specularTint, sheenTint, subsurface, sheen, ax, ay;
} ErMaterial; /* 29 words */

/* The reference `Camera` (src/Camera.h:9-21), 48 bytes. rotation in degrees. */
typedef struct ErCamera {
    float focalLength, sensorWidth, sensorHeight, aperture, focusDistance;
    float rotation[3];
    int32_t bokeh;
    float position[3];
} ErCamera;

typedef struct ErPointLight { /* src/PointLight.h:8-9 — uploaded, inert in the reference (never sampled) */
    float position[3];
    float radiance[3];
} ErPointLight;

enum ErTexFilter { ER_FILTER_NONE = 0, ER_FILTER_BILINEAR = 1 };
typedef struct ErTexture { /* src/Texture.h:18-28; interleaved float texels, row-major */
    int32_t width, height, channels, filter;
    const float* data;
} ErTexture;

/* The reference `Hit` (src/Hit.h:7-12), 84 bytes. On a miss valid=0, triIdx=-2, the rest 0. */
typedef struct ErHit {
    float position[3], normal[3], tangent[3], bitangent[3], gnormal[3];
    int32_t valid;
    uint32_t objectID, material_id;
    float tu, tv;
    int32_t triIdx; /* slot in BVH order (index into triIndices), src/BVH.cpp:116 */
} ErHit;

/* Everything `copy_scene` deep-copies (src/SYCLCopy.cpp:3-104). Host pointers are borrowed only for
 * the duration of er_upload_scene. */
typedef struct ErSceneDesc {
    uint32_t x_res, y_res;
    const void* tris;        /* tri_count records, first 156 bytes laid out as ErTri */
    size_t tri_count;
    size_t tri_stride;       /* bytes; 192 for the reference's Tri[], 156 for packed ErTri[] */
    const void* nodes;       /* ER_BVH_NODE_COUNT records laid out as ErNode (reference pre-order) */
    size_t node_stride;      /* bytes; 44 */
    const int32_t* tri_indices; /* tri_count entries: BVH slot -> index into tris */
    const ErMaterial* materials;
    int32_t material_count;
    const ErTexture* textures;
    int32_t texture_count;
    ErTexture hdri;          /* environment texture */
    const float* hdri_cdf;   /* width*height+1 floats (HDRI::generateCDF, src/HDRI.cpp:62-83) */
    float hdri_radiance_sum;
    const ErCamera* camera;
    const ErPointLight* point_lights;
    int32_t point_light_count;
} ErSceneDesc;

typedef struct ErRenderOpts {
    int32_t max_depth;      /* bounce count; the reference hard-codes 5 (src/kernel.cpp:508). 0 -> 5 */
    int32_t rng_order;      /* ErRngOrder */
    int32_t spp_per_launch; /* samples traced per pixel per kernel launch (0 -> library default) */
    int32_t kernel;         /* 0 = default (persistent, path-regenerating); 1 = one-thread-per-pixel megakernel */
    int32_t count_stats;    /* 1 = accumulate ErCounters (slower build of the same kernels) */
    int32_t extensions;     /* ErExtension bits, 0 = the reference's estimator (default). Any bit selects the
                             * one-thread-per-pixel kernel: the extensions change the image, they are not on the hot path */
    int32_t reserved[2];
} ErRenderOpts;

/* Off-by-default corrections of the reference's estimator. Each changes the image. */
enum ErExtension {
    ER_EXT_MIS = 1,          /* balance-heuristic weights hw / bw (computed and dropped at src/kernel.cpp:571-572) applied
                              * to the environment NEE term and to the environment seen by the BRDF-sampled ray */
    ER_EXT_POINT_LIGHTS = 2, /* point-light NEE (pointLight(), src/kernel.cpp:269-301, never called there): one light per
                              * opaque hit, pdf count/(2 pi), quadratic attenuation, Disney BRDF, shadow ray */
    ER_EXT_ROULETTE = 4      /* Russian roulette from the third bounce on: survive with p = clamp(max(reduction), .05, 1) */
};

typedef struct ErCounters { /* SURVEY §8(d) definitions; zero unless count_stats */
    unsigned long long rays, box_tests, tri_tests, candidates, shaded_hits, nee_samples, samples;
} ErCounters;

typedef struct ErDeviceInfo { /* feeds the get_sycl_info JSON keys (src/CommandManager.cpp:303-362) */
    char name[256];
    char platform[64];
    unsigned long long memory_bytes;
    int32_t max_compute_units; /* SM count */
    int32_t cc_major, cc_minor;
    int32_t is_compatible;     /* 1 iff compute capability 10.x (kernels are sm_100a only) */
} ErDeviceInfo;

typedef struct ErCtx ErCtx;

/* ---- devices / context ---------------------------------------------------------------------- */
ER_API const char* er_last_error(void);
ER_API int er_device_count(int* count);
ER_API int er_device_info(int index, ErDeviceInfo* info);
/* One context drives one GPU. Multi-GPU = one context (one process) per GPU, each rendering its
 * shard of pixel rows; passes are combined at readout by the caller's collective (NCCL reduce of
 * er_get_pass_device buffers: non-owned pixels are exactly 0). */
ER_API int er_create(int device_id, ErCtx** ctx);
ER_API int er_destroy(ErCtx* ctx);
/* Row-block sharding: rows [k*rows_per_block, (k+1)*rows_per_block) belong to shard k % shard_count. */
ER_API int er_set_shard(ErCtx* ctx, int shard_index, int shard_count, int rows_per_block);
/* Contiguous-stripe sharding: this context owns image rows [y_begin, y_end); y_begin must be a multiple of
 * rows_per_block. Ownership stays a pure function of the pixel, so a sum over shards is still the 1-GPU frame
 * bit for bit (src/kernel.cpp:489-490 — the pixel -> idx map — is untouched). Stripes keep a GPU's rays inside
 * one band of the scene (L1/L2 locality at 1/N of the frame); er_row_cost + er_balance_stripes place the
 * boundaries so that the bands cost the same. */
ER_API int er_set_shard_stripe(ErCtx* ctx, int y_begin, int y_end, int rows_per_block);
/* Cost prepass for cost-balanced stripes: traces ONE sample of the row blocks block_first, block_first +
 * block_stride, ... of the whole frame with the counting build of the render kernel and writes, per image row,
 * the traversal + shading work of its pixels (0 for rows not sampled) into cost_per_row[y_res]. Restarts
 * accumulation afterwards (setupKernel): the pixel state is that of er_render_reset. */
ER_API int er_row_cost(ErCtx* ctx, int block_first, int block_stride, const struct ErRenderOpts* opts, uint32_t* cost_per_row);
/* Host only. y_bounds[0..n]: stripe k = rows [y_bounds[k], y_bounds[k+1]), boundaries at multiples of
 * rows_per_block, summed cost split evenly; row blocks without a sample take the nearest sampled block's cost. */
ER_API int er_balance_stripes(const uint32_t* cost_per_row, int y_res, int rows_per_block, int n, int* y_bounds);
/* Host only. Feedback step of the balance: stripe k = rows [y_bounds[k], y_bounds[k+1]) took ms[k] (er_get_kernel_time
 * of its GPU for the same job); y_bounds_out[0..n] = new boundaries (block edges, every stripe at least one block) that
 * would equalise the times if a stripe's time were spread evenly over its rows, moved 1.5 x as far (a launch has a
 * part that does not shrink with its rows, so the proportional step alone under-corrects); two or three steps settle.
 * The reference has one device and no counterpart (its frame is never split). */
ER_API int er_rebalance_stripes(const int* y_bounds, const float* ms, int n, int rows_per_block, int* y_bounds_out);
/* Order in which the context hands its own row blocks to the render kernel: the i-th block rendered is owned block
 * order[i] (a permutation of 0..n-1, n = the number of row blocks the context owns; anything else is ignored at
 * launch). A launch ends with its slowest pixel and a pixel's samples are sequential, so the costly blocks should
 * start first; the image does not depend on the order. NULL or n = 0: top to bottom. Reset by er_set_shard*. */
ER_API int er_set_block_order(ErCtx* ctx, const int* order, int n);
/* Host only. The row blocks of rows [y_begin, y_end) in descending order of their er_row_cost cost (indices relative
 * to the first block of the range; unsampled blocks take the nearest sampled block's cost): order_out[0..*n_out). */
ER_API int er_order_blocks(const uint32_t* cost_per_row, int y_res, int rows_per_block, int y_begin, int y_end, int* order_out, int* n_out);
/* Launch on a caller-owned cudaStream_t (e.g. torch's current stream); NULL -> the context's own stream. */
ER_API int er_set_stream(ErCtx* ctx, void* cuda_stream);

/* ---- host prerequisites (result-identical to the reference's host code) ---------------------- */
/* BVH::build (src/BVH.cpp:132-156,239-281,327-415,447-534): fixed depth-18 pre-order tree, 14-bin SAH.
 * Writes ER_BVH_NODE_COUNT nodes and tri_count slot->triangle indices. Multi-threaded over subtrees. */
ER_API int er_bvh_build(const void* tris, size_t tri_count, size_t tri_stride, ErNode* nodes_out,
                        int32_t* tri_indices_out, int nthreads);
/* HDRI::generateCDF (src/HDRI.cpp:62-83). cdf_out has width*height+1 floats. */
ER_API int er_hdri_build_cdf(const ErTexture* tex, float* cdf_out, float* radiance_sum_out);

/* Upload-layout round trip (test hook, host only): flattens desc's BVH into the device pair-record layout
 * and maps it back; nodes_out[ER_BVH_NODE_COUNT] receives b1,b2 of every depth>=1 node and from/to of the
 * leaves, slot_tris_out[tri_count*39] the ErTri words per BVH slot as the device buffers hold them. */
ER_API int er_flatten_roundtrip(const ErSceneDesc* desc, ErNode* nodes_out, float* slot_tris_out, float* max_slack_out);

/* Scene ingest, host only (ObjLoader::loadObjsRapid's stream path, src/ObjLoader.cpp:66-164: rapidobj parse +
 * triangulation, z flip, optional face-weighted normals, MikkTSpace tangents): Wavefront OBJ text + the MTL text
 * whose `newmtl` names `usemtl` refers to -> ErTri records in load order (objectID = index of the shape among the
 * loaded objects, materialID 0: Scene::pair_materials pairs by name). tris_out may be NULL to ask for the count.
 * mat_names_out receives one '\n'-terminated material name per triangle (Tri::matName). */
ER_API int er_load_obj(const char* obj_text, size_t obj_len, const char* mtl_text, size_t mtl_len, int recompute_normals,
                       ErTri* tris_out, size_t tri_capacity, size_t* tri_count, int32_t* object_of_tri,
                       char* mat_names_out, size_t mat_names_capacity, size_t* mat_names_len);

/* ---- scene upload + render ------------------------------------------------------------------ */
/* copy_scene + setupKernel (renderSetup, src/kernel.cpp:651-677): flattens to the device SoA layout,
 * seeds the per-pixel RNG, zeroes the passes (alpha 1) and sets samples to 1. */
/* When desc->tris points into page-locked host memory (cudaHostAlloc, er_host_alloc, er_host_register) and holds 65536
 * triangles or more, the records cross PCIe as they are (DMA from the caller's pages) and the device builds its own
 * layout; otherwise the host does and the result is staged. Same device arrays bit for bit either way. */
ER_API int er_upload_scene(ErCtx* ctx, const ErSceneDesc* desc);
ER_API int er_set_camera(ErCtx* ctx, const ErCamera* camera);
/* setupKernel only: restart accumulation on the uploaded scene. */
ER_API int er_render_reset(ErCtx* ctx);
/* kernel_render_enqueue (src/kernel.cpp:680-706): trace `samples` more samples for every owned pixel,
 * exactly one renderingKernel-equivalent per pixel per sample. Synchronous variant returns when the
 * GPU work is complete; the async pair runs it on the context's render thread. */
ER_API int er_render_samples(ErCtx* ctx, unsigned samples, const ErRenderOpts* opts);
ER_API int er_render_enqueue(ErCtx* ctx, unsigned samples, const ErRenderOpts* opts); /* no host sync */
ER_API int er_render_start(ErCtx* ctx, unsigned target_samples, const ErRenderOpts* opts);
ER_API int er_render_poll(ErCtx* ctx, unsigned* samples, int* running);
ER_API int er_render_wait(ErCtx* ctx);
ER_API int er_render_abort(ErCtx* ctx);
ER_API int er_sync(ErCtx* ctx);

/* ---- readout -------------------------------------------------------------------------------- */
/* RenderingManager::get_render_info: dev_samples[0], i.e. completed samples + 1. */
ER_API int er_get_render_info(ErCtx* ctx, unsigned* samples);
/* RenderingManager::get_pass: x_res*y_res RGBA floats into caller-owned memory. */
ER_API int er_get_pass(ErCtx* ctx, int pass, float* out_rgba);
/* Multi-GPU readout, shaped for disjoint ownership: copies ONLY the pixels this context's shard owns into the
 * caller's FULL-FRAME buffer (x_res*y_res float4; the other pixels are not touched). N contexts (threads of a
 * process, or processes sharing the buffer) each move 1/N of the frame over their own PCIe link; no reduce of
 * zeros. Pinned memory (er_host_alloc / er_host_register) makes the copies asynchronous DMA. */
ER_API int er_get_pass_owned(ErCtx* ctx, int pass, float* frame_rgba);
/* Page-locked host memory for frames and scene arrays; er_host_register pins caller memory in place (e.g. a
 * POSIX shared-memory frame that several single-GPU processes read their stripes into). */
ER_API int er_host_alloc(void** ptr, size_t bytes);
ER_API int er_host_free(void* ptr);
ER_API int er_host_register(void* ptr, size_t bytes);
ER_API int er_host_unregister(void* ptr);
/* Device pointer of the same data (x_res*y_res float4); with sharding, non-owned pixels hold 0 (alpha 0)
 * so that a sum over shards reproduces the single-GPU buffer bit for bit. */
ER_API int er_get_pass_device(ErCtx* ctx, int pass, void** device_ptr);
ER_API int er_get_pixel_state(ErCtx* ctx, uint32_t* rng_out, uint32_t* samples_out); /* either may be NULL */
ER_API int er_get_counters(ErCtx* ctx, ErCounters* out);
ER_API int er_reset_counters(ErCtx* ctx);
/* GPU time (ms, CUDA events on the launch stream) and launch count of render kernels since last reset */
ER_API int er_get_kernel_time(ErCtx* ctx, float* ms, int* launches);

/* ---- device groups: one host thread driving several GPUs (RenderingManager with "gpus" > 1) ------- */
/* A group owns one context per device and shards the frame over them (interleaved row blocks until
 * er_group_balance installs cost-balanced stripes). It is the multi-GPU form of the calls above:
 *   er_group_upload_scene   flattens ONCE, then uploads to every member over its own PCIe link (copy_scene)
 *   er_group_render_*       kernel_render_enqueue on every member; batches advance in lock step
 *   er_group_get_render_info  min over members (a frame has n samples when every shard has n)
 *   er_group_get_pass       every member copies its owned rows straight into the caller's frame, all links
 *                           in parallel (RenderingManager::get_pass, src/Managers.cpp:287-302); the result is
 *                           bit-identical to the 1-GPU frame
 * There is no collective in the render loop. */
typedef struct ErGroup ErGroup;
ER_API int er_group_create(const int* device_ids, int n, ErGroup** group);
ER_API int er_group_destroy(ErGroup* group);
ER_API int er_group_size(ErGroup* group, int* n);
ER_API int er_group_ctx(ErGroup* group, int member, ErCtx** ctx);
ER_API int er_group_upload_scene(ErGroup* group, const ErSceneDesc* desc);
ER_API int er_group_balance(ErGroup* group, const ErRenderOpts* opts, int* y_bounds_out /* n+1, may be NULL */);
/* New stripes from the kernel time every GPU of the group spent since its counters were last reset (the frame or the
 * batches just rendered with the stripes of er_group_balance); restarts accumulation. */
ER_API int er_group_rebalance(ErGroup* group, int* y_bounds_out /* n+1, may be NULL */);
ER_API int er_group_render_reset(ErGroup* group);
ER_API int er_group_render_samples(ErGroup* group, unsigned samples, const ErRenderOpts* opts);
ER_API int er_group_render_start(ErGroup* group, unsigned target_samples, const ErRenderOpts* opts);
ER_API int er_group_render_poll(ErGroup* group, unsigned* samples, int* running);
ER_API int er_group_render_wait(ErGroup* group);
ER_API int er_group_render_abort(ErGroup* group);
ER_API int er_group_get_render_info(ErGroup* group, unsigned* samples);
ER_API int er_group_get_pass(ErGroup* group, int pass, float* out_rgba);

/* ---- nearest-hit queries (BVH::transverse via throwRay, src/kernel.cpp:218-240) ------------------ */
/* rays: n * (origin xyz, direction xyz); direction is normalised as the reference Ray ctor does. */
ER_API int er_trace(ErCtx* ctx, const float* rays, size_t n, ErHit* hits_out);
/* mode bit0: accumulate ErCounters; bit1: reference visit order (left-first, no distance pruning, empty-subtree
 * skipping only) — the cross-check of the ordered/pruned traversal that er_trace uses */
ER_API int er_trace_mode(ErCtx* ctx, const float* rays, size_t n, ErHit* hits_out, int mode);
/* same with device-resident rays/hits (for throughput measurement) */
ER_API int er_trace_device(ErCtx* ctx, const void* d_rays, size_t n, void* d_hits);
/* er_trace without any device allocation. NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid
 * it). Same inputs, outputs and results as er_trace (host pointers; the call returns when hits_out is filled); the
 * device staging is a fixed file-scope array in the library (6.75 MiB per GPU that loads it, held until exit), so
 * n <= ER_TRACE_NOALLOC_MAX — more is ER_ERR_INVALID with the limit in er_last_error(). The entry point calls no
 * cudaMalloc / cudaFree / allocating library routine. All calls on one GPU share the array: calls that overlap on the
 * GPU (two streams, contexts or threads on one device) race and must be serialised by the caller. */
#define ER_TRACE_NOALLOC_MAX 65536
ER_API int er_trace_noalloc(ErCtx* ctx, const float* rays, size_t n, ErHit* hits_out);
/* camera rays of kernel.cpp:371-473 for given pixels and the five uniforms r1..r5 (unit test hook) */
ER_API int er_camera_rays(ErCtx* ctx, const int32_t* xy, const float* r5, size_t n, float* rays_out);

/* ---- device-function test hooks (run the device code on given inputs; used by tests only) ----- */
/* the geometry arrays as they sit on the device after er_upload_scene (which: 0 child-pair records, 1 vertices, 2 normals,
 * 3 attributes; 16-B vectors as floats): up to n_floats into out, the array's size into *total_floats (may be NULL) */
ER_API int er_test_geometry(ErCtx* ctx, int which, float* out, size_t n_floats, size_t* total_floats);
ER_API int er_test_rng(ErCtx* ctx, int idx, int n, uint32_t* states, float* values);
ER_API int er_test_texture(ErCtx* ctx, int tex /* -1 = hdri */, int mode, const float* ab, size_t n, float* out3);
/* the device's sin/cos/acos/atan2/pow/log (fn 0..5; b = second operand of atan2(a,b) / pow(a,b), else NULL):
 * bit-compared with the host libm the reference's sycl::sin etc. resolve to (src/sycl.h:8) */
ER_API int er_test_libm(ErCtx* ctx, int fn, const float* a, const float* b, size_t n, float* out);
ER_API int er_test_spherical(ErCtx* ctx, const float* dirs3, size_t n, float* uv2);
ER_API int er_test_hdri_sample(ErCtx* ctx, const float* r, size_t n, float* out10, float* pdf);
ER_API int er_test_disney(ErCtx* ctx, const float* hitdata37, const float* V3, const float* L3, const float* r3,
                          size_t n, float* sample3, float* eval3, float* pdf);
ER_API int er_test_hitdata(ErCtx* ctx, const ErHit* hits, size_t n, float* hitdata37);

#ifdef __cplusplus
}
#endif
#endif /* ER_CUDA_H */
