/*
 * api.cu -- host side of the drop-in boundary: the rt_* handle API and the legacy
 * launchKernel() shim over it (replaces the reference's host wrapper,
 * reference kernel.cu:406-462).
 *
 * Differences from the reference wrapper, by design:
 *   - scene data lives in HBM as an LBVH + SoA records (rt_device.h) instead of
 *     133 Objects in __constant__ memory; per-frame uniforms travel as one kernel
 *     parameter block instead of five cudaMemcpyToSymbol calls (kernel.cu:447-451);
 *   - sky maps are texture OBJECTS (point filter, clamp, normalised coordinates --
 *     the defaults the reference's texture references end up with, kernel.cu:19-22,
 *     429-436); legacy texture references no longer exist in CUDA 12;
 *   - rt_set_scene classifies the change against the previous upload (nothing /
 *     materials and planes only / moved primitives -> refit / topology -> rebuild).
 */
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "rt_device.h"


extern "C" {
cudaError_t rt_build_bvh(const RtBuildBuffers* b, int mode, int force_radix, int builder, cudaStream_t stream);
cudaError_t rt_collapse_wide(const float* nodes, int n_prims, float4* wide, cudaStream_t stream);
int rt_build_by_value(const RtBuildBuffers* b, int mode);
cudaError_t rt_radix_sort_pairs(unsigned int* keys, unsigned int* vals, unsigned int* tmp_keys, unsigned int* tmp_vals,
                                unsigned int* hist, int n, int force_radix, cudaStream_t stream);
int rt_radix_hist_entries(int n);
cudaError_t rt_radix_sort_pairs_static(unsigned int* keys, unsigned int* vals, int n, int device, cudaStream_t stream,
                                       const char** why);
cudaError_t rt_launch_trace(const RtSceneDev* sc, const RtFrameDev* f, unsigned int* image, int* hit_ids,
                            RtCountersDev* counters, unsigned int* err, int brute, int count, int packet,
                            cudaStream_t stream);
cudaError_t rt_launch_trace_big(const RtSceneDev* sc, const RtFrameDev* f, unsigned int* image, int* hit_ids,
                                RtCountersDev* counters, unsigned int* tile_counter, unsigned int* err, int count,
                                int packet, int smem_stack, int n_sm, cudaStream_t stream);
cudaError_t rt_launch_trace_small(const RtSceneDev* sc, const RtFrameDev* f, unsigned int* image, int* hit_ids,
                                  RtCountersDev* counters, unsigned int* tile_counter, unsigned short* tile_lists,
                                  int count, int force_linear, int n_sm, const unsigned int* order_r,
                                  unsigned int* order_w, int after_own_post, uint4* lgrid, int build_lgrid,
                                  int mirror_bins, int compact, cudaStream_t stream);
size_t rt_trace_small_order_bytes(int imgw, int imgh);
size_t rt_trace_small_patches(int imgw, int imgh);
size_t rt_trace_small_list_bytes(int imgw, int imgh);
cudaError_t rt_lgb_finish(const unsigned int* total, const unsigned int* fail, unsigned int cap, unsigned int* state,
                          cudaStream_t stream);
size_t rt_lgb_bins(void);
size_t rt_lgb_sums(void);
cudaError_t rt_lgb_count(const RtSceneDev* sc, const float* light_pos, unsigned int* counts, unsigned int* offsets,
                         unsigned int* sums, unsigned int* fail, cudaStream_t stream);
cudaError_t rt_lgb_fill(const RtSceneDev* sc, const float* light_pos, unsigned int* counts, const unsigned int* offsets,
                        unsigned int* entries, unsigned int cap, unsigned int* fail, cudaStream_t stream);
int rt_trace_small_supported(int n_bvh_prims, int n_objects, int n_planes);
cudaError_t rt_trace_small_init(void);
cudaError_t rt_fxaa_init(void);
cudaError_t rt_launch_fxaa_smem(const unsigned int* in, unsigned int* out, int w, int h, const rt_shard* sh,
                                const RtPeerSync* peer, cudaStream_t stream);
cudaError_t rt_launch_fxaa_tma(const CUtensorMap* map, const unsigned int* in, unsigned int* out, int w, int h,
                               const rt_shard* sh, int n_sm, const RtPeerSync* peer, cudaStream_t stream);
cudaError_t rt_launch_copy(const unsigned int* in, unsigned int* out, int w, int h, const rt_shard* sh,
                           const RtPeerSync* peer, cudaStream_t stream);
cudaError_t rt_launch_pack_tiles(const unsigned int* frame, unsigned int* packed, int w, int h, const rt_shard* sh,
                                 int unpack, cudaStream_t stream);
cudaError_t rt_launch_ffma_peak(int n_sm, cudaStream_t stream, float* ms, double* flops);
cudaError_t rt_launch_present(const unsigned int* in, unsigned int* out, int w, int h, int factor, int mode,
                              cudaStream_t stream);
cudaError_t rt_launch_peer_signal(unsigned int* flags, int rank, unsigned int seq, cudaStream_t stream);
cudaError_t rt_launch_peer_wait(unsigned int* flags, int n_ranks, unsigned int seq, unsigned int timeout_ms,
                                cudaStream_t stream);
cudaError_t rt_launch_peer_release(unsigned int* flags, unsigned int seq, cudaStream_t stream);
cudaError_t rt_launch_peer_acquire(unsigned int* flags, unsigned int seq, unsigned int timeout_ms, cudaStream_t stream);
cudaError_t rt_launch_peer_reset_error(unsigned int* flags, cudaStream_t stream);
}

/* The four sky maps of a device: 512 MiB for the reference's 8192x4096 maps, so contexts of one device share one
 * block (rt_share_sky).  The block lives until its last user lets go of it; a user lets go only after its own
 * stream has drained (rt_set_sky, rt_share_sky, rt_destroy), so no frame in flight can sample freed textures. */
struct RtSkyBlock {
    cudaArray_t arr[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaTextureObject_t tex[4] = {0, 0, 0, 0};
    int device = 0;
    int users = 1;
};

/* One slot of the asynchronous read-back pipeline: a device staging frame (rt_render_to_host_async) and the two events
 * that order tracer stream -> copy stream -> host. */
#define RT_MAX_SLOTS 8
struct RtSlot {
    unsigned int* dev = nullptr;
    size_t cap = 0;                 /* pixels */
    cudaEvent_t ready = nullptr;    /* the frame is complete on the context's stream */
    cudaEvent_t copied = nullptr;   /* the frame has reached host memory */
    bool used = false;
};

unsigned long long g_rt_launches = 0;      /* RT_LAUNCHED (rt_device.h) */

struct rt_context {
    int device = 0;
    int n_sm = 148;
    cudaStream_t stream = 0;
    std::string err;
    int opt[RT_OPT_N] = {0};

    /* sky */
    RtSkyBlock* sky = nullptr;

    /* scene */
    std::vector<rtb_Object> host_objs;      /* last upload, for change classification */
    std::vector<int> host_planes;
    rtb_Light lights[RT_MAX_LIGHTS];
    int n_lights = 0;
    bool have_scene = false;
    int cap_objects = 0, cap_bvh = 0, cap_planes = 0;
    float* d_objs = nullptr;
    int* d_plane_idx = nullptr;
    RtBuildBuffers bb;
    RtSceneDev scene;

    /* frame buffers */
    unsigned int* d_base = nullptr;
    size_t cap_base = 0;
    unsigned int* d_stage_out = nullptr;    /* for rt_render_to_host */
    int* d_stage_ids = nullptr;
    size_t cap_stage = 0;
    RtCountersDev* d_counters = nullptr;
    /* every write to the scene buffers completes before the last kernel enqueued by rt_render runs to its end:
     * the next frame's binning pre-pass may overlap that kernel (programmatic dependent launch, trace_small.cu) */
    bool scene_settled = false;
    unsigned int* last_base = nullptr;       /* where the last frame's tracer wrote: d_base, or the caller's buffer */
    unsigned int* d_order = nullptr;         /* two order-history buffers of the persistent tracer (trace_small.cu) */
    size_t cap_order = 0;                    /* uints */
    int order_key[7] = {0, 0, 0, 0, 0, 0, 0};   /* frame size + shard the history belongs to */
    int order_cur = 0;                       /* buffer holding it */
    unsigned short* d_tile_lists = nullptr;  /* per-tile primary candidate lists (trace_small.cu) */
    uint4* d_lgrid = nullptr;                /* light bins of the shared-memory tracer (trace_small.cu), RT_LGRID_BYTES */
    /* The light bins depend on the scene and the lights only, not on the camera: a frame that finds both unchanged
     * since the bins were built (static scene, moving camera) reuses them and its pre-pass runs the screen bins alone. */
    bool lgrid_valid = false;
    /* light grid of big scenes (light_grid.cu): scratch + per light offsets / entries; rebuilt when stale */
    unsigned int *d_lgb_counts = nullptr, *d_lgb_sums = nullptr, *d_lgb_fail = nullptr;
    unsigned int* d_lgb_offsets[RT_LGB_LIGHTS] = {nullptr, nullptr};
    unsigned int* d_lgb_entries[RT_LGB_LIGHTS] = {nullptr, nullptr};
    size_t cap_lgb_entries[RT_LGB_LIGHTS] = {0, 0};
    unsigned int* d_lgb_state = nullptr;     /* 4 uints per light: the build's own verdict (rt_lgb_finish_kernel) */
    unsigned int* h_lgb_state = nullptr;     /* page-locked copy of it, read back asynchronously */
    cudaEvent_t lgb_read[RT_LGB_LIGHTS] = {nullptr, nullptr};
    bool lgb_read_pending[RT_LGB_LIGHTS] = {false, false};
    bool lgb_stale = true;
    bool lgb_waited = false;                 /* a frame has been rendered since the grid went stale */
    size_t cap_tile_lists = 0;
    /* 4-wide view of the hierarchy for the plain tracer of big scenes (RT_OPT_WIDE_BVH), rebuilt after every
     * rt_set_scene that touched the hierarchy */
    float4* d_wide = nullptr;
    size_t cap_wide = 0;                     /* float4s */
    bool wide_valid = false;
    unsigned int* d_tile_counter = nullptr;  /* [0] next tile, [1] finished CTAs (persistent tracer; self re-arming) */

    /* TMA descriptor cache */
    CUtensorMap tmap;
    const void* tmap_ptr = nullptr;
    int tmap_w = 0, tmap_h = 0;

    cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    bool timing_valid = false, build_timed = false;

    /* asynchronous read-back: device -> pinned host copies on a stream of their own, RT_MAX_SLOTS frames in flight */
    cudaStream_t copy_stream = nullptr;
    RtSlot slots[RT_MAX_SLOTS];
};

namespace {

#define RT_CHECK(ctx, call)                                                                     \
    do {                                                                                        \
        cudaError_t e_ = (call);                                                                \
        if (e_ != cudaSuccess) {                                                                \
            (ctx)->err = std::string(#call) + ": " + cudaGetErrorString(e_);                    \
            return RT_ERR_CUDA;                                                                 \
        }                                                                                       \
    } while (0)

int fail(rt_context* ctx, int code, const char* msg) {
    ctx->err = msg;
    return code;
}

/* Replace *p by a fresh allocation of `count` elements (contents are not kept).  The new block is obtained first:
 * when that fails the old one -- and with it every capacity the caller recorded -- stays valid. */
template <class T>
cudaError_t regrow(T** p, size_t count) {
    T* fresh = nullptr;
    if (count != 0) {
        const cudaError_t e = cudaMalloc((void**)&fresh, sizeof(T) * count);
        if (e != cudaSuccess) return e;
    }
    if (*p) cudaFree(*p);
    *p = fresh;
    return cudaSuccess;
}

void sky_free(RtSkyBlock* b) {
    for (int i = 0; i < 4; i++) {
        if (b->tex[i]) cudaDestroyTextureObject(b->tex[i]);
        if (b->arr[i]) cudaFreeArray(b->arr[i]);
    }
    delete b;
}

/* the context stops using its sky block: only after everything it has enqueued has run */
void sky_release(rt_context* ctx) {
    if (!ctx->sky) return;
    cudaStreamSynchronize(ctx->stream);
    if (--ctx->sky->users == 0) sky_free(ctx->sky);
    ctx->sky = nullptr;
}

typedef CUresult (*encode_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

encode_tiled_fn get_encode_tiled() {
    static encode_tiled_fn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (encode_tiled_fn)p;
    }
    return fn;
}

/* 2-D uint32 image [h][w] -> 72x34 boxes (64x32 FXAA tile + halo; see FX_HALO_X in fxaa.cu) */
bool make_fxaa_tmap(rt_context* ctx, const unsigned int* img, int w, int h) {
    if (ctx->tmap_ptr == img && ctx->tmap_w == w && ctx->tmap_h == h) return true;
    encode_tiled_fn enc = get_encode_tiled();
    if (!enc) return false;
    cuuint64_t dims[2] = {(cuuint64_t)w, (cuuint64_t)h};
    cuuint64_t strides[1] = {(cuuint64_t)w * 4};
    cuuint32_t box[2] = {72, 34};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(&ctx->tmap, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, (void*)img, dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return false;
    ctx->tmap_ptr = img;
    ctx->tmap_w = w;
    ctx->tmap_h = h;
    return true;
}

int run_fxaa(rt_context* ctx, const unsigned int* in, unsigned int* out, int w, int h, const rt_shard* sh,
             const RtPeerSync* peer) {
    const bool shard_ok = !sh || sh->n_shards <= 1 || (sh->tile_w % 64 == 0 && sh->tile_h % 32 == 0);
    const bool tma_ok = ctx->opt[RT_OPT_FXAA_IMPL] == 0 && (w % 4 == 0) && (((uintptr_t)in) % 16 == 0) && shard_ok &&
                        make_fxaa_tmap(ctx, in, w, h);
    if (tma_ok) RT_CHECK(ctx, rt_launch_fxaa_tma(&ctx->tmap, in, out, w, h, sh, ctx->n_sm, peer, ctx->stream));
    else        RT_CHECK(ctx, rt_launch_fxaa_smem(in, out, w, h, sh, peer, ctx->stream));
    return RT_OK;
}

bool geometry_equal(const rtb_Object& a, const rtb_Object& b) {
    return a.type == b.type && memcmp(&a.pos, &b.pos, 36) == 0;   /* pos, size, third are contiguous */
}

}  // namespace

extern "C" {

int rt_create(int device, rt_context** out) {
    if (!out) return RT_ERR_ARG;
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || device < 0 || device >= count) return RT_ERR_CUDA;
    rt_context* ctx = new rt_context();
    ctx->device = device;
    memset(&ctx->bb, 0, sizeof(ctx->bb));
    memset(&ctx->scene, 0, sizeof(ctx->scene));
    ctx->scene.root = RT_NODE_SENTINEL;
    if (cudaSetDevice(device) != cudaSuccess) { delete ctx; return RT_ERR_CUDA; }
    cudaDeviceGetAttribute(&ctx->n_sm, cudaDevAttrMultiProcessorCount, device);
    bool ok = cudaMalloc((void**)&ctx->d_counters, sizeof(RtCountersDev)) == cudaSuccess &&
              cudaMemset(ctx->d_counters, 0, sizeof(RtCountersDev)) == cudaSuccess &&
              cudaMalloc((void**)&ctx->d_tile_counter, 64 * sizeof(unsigned int)) == cudaSuccess &&
              cudaMemset(ctx->d_tile_counter, 0, 64 * sizeof(unsigned int)) == cudaSuccess &&
              rt_trace_small_init() == cudaSuccess && rt_fxaa_init() == cudaSuccess;
    for (int i = 0; i < 6 && ok; i++) ok = cudaEventCreate(&ctx->ev[i]) == cudaSuccess;
    if (!ok) {
        rt_destroy(ctx);       /* frees whatever has been allocated so far */
        return RT_ERR_CUDA;
    }
    *out = ctx;
    return RT_OK;
}

int rt_destroy(rt_context* ctx) {
    if (!ctx) return RT_ERR_ARG;
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    sky_release(ctx);
    RtBuildBuffers& b = ctx->bb;
    void* ptrs[] = {ctx->d_objs, ctx->d_plane_idx, b.mats, b.recs, b.planes, b.bvh_src, b.prim_lo, b.prim_hi, b.bounds,
                    b.keys, b.vals, b.tmp_keys, b.tmp_vals, b.hist, b.children, b.node_parent, b.leaf_parent, b.node_other, b.flags, b.node_aux,
                    b.nodes, b.prims, b.depth, ctx->d_base, ctx->d_stage_out, ctx->d_stage_ids, ctx->d_counters, ctx->d_tile_counter, ctx->d_tile_lists, ctx->d_order, ctx->d_lgrid, ctx->d_lgb_counts, ctx->d_lgb_sums, ctx->d_lgb_fail, ctx->d_lgb_offsets[0],
                    ctx->d_lgb_offsets[1], ctx->d_lgb_entries[0], ctx->d_lgb_entries[1], ctx->d_lgb_state, ctx->d_wide};
    for (void* p : ptrs) if (p) cudaFree(p);
    if (ctx->h_lgb_state) cudaFreeHost(ctx->h_lgb_state);
    for (int i = 0; i < RT_LGB_LIGHTS; i++) if (ctx->lgb_read[i]) cudaEventDestroy(ctx->lgb_read[i]);
    for (int i = 0; i < 6; i++) if (ctx->ev[i]) cudaEventDestroy(ctx->ev[i]);
    for (RtSlot& sl : ctx->slots) {
        if (sl.dev) cudaFree(sl.dev);
        if (sl.ready) cudaEventDestroy(sl.ready);
        if (sl.copied) cudaEventDestroy(sl.copied);
    }
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    delete ctx;
    return RT_OK;
}

const char* rt_last_error(rt_context* ctx) { return ctx ? ctx->err.c_str() : "null context"; }

int rt_set_option(rt_context* ctx, int key, int value) {
    if (!ctx || key < 0 || key >= RT_OPT_N) return RT_ERR_ARG;
    ctx->scene_settled = false;
    ctx->lgrid_valid = false;
    if ((key == RT_OPT_BUILDER || key == RT_OPT_FORCE_RADIX) && ctx->opt[key] != value)
        ctx->have_scene = false;   /* the next rt_set_scene rebuilds the hierarchy */
    ctx->opt[key] = value;
    return RT_OK;
}

int rt_set_stream(rt_context* ctx, void* cuda_stream) {
    if (!ctx) return RT_ERR_ARG;
    ctx->stream = (cudaStream_t)cuda_stream;
    ctx->scene_settled = false;
    return RT_OK;
}

int rt_set_sky(rt_context* ctx, const unsigned char* const tex[4], int w, int h) {
    if (!ctx || !tex || w <= 0 || h <= 0) return RT_ERR_ARG;
    for (int i = 0; i < 4; i++)
        if (!tex[i]) return fail(ctx, RT_ERR_ARG, "null sky map");
    RT_CHECK(ctx, cudaSetDevice(ctx->device));
    /* build the new block completely, then swap: a failure leaves the context with the maps it had */
    RtSkyBlock* b = new RtSkyBlock;
    b->device = ctx->device;
    cudaError_t e = cudaSuccess;
    for (int i = 0; i < 4 && e == cudaSuccess; i++) {
        cudaChannelFormatDesc desc = cudaCreateChannelDesc<uchar4>();
        e = cudaMallocArray(&b->arr[i], &desc, w, h);
        if (e == cudaSuccess)
            e = cudaMemcpy2DToArray(b->arr[i], 0, 0, tex[i], (size_t)w * 4, (size_t)w * 4, h, cudaMemcpyHostToDevice);
        cudaResourceDesc rd;
        memset(&rd, 0, sizeof(rd));
        rd.resType = cudaResourceTypeArray;
        rd.res.array.array = b->arr[i];
        cudaTextureDesc td;
        memset(&td, 0, sizeof(td));
        td.addressMode[0] = td.addressMode[1] = cudaAddressModeClamp;
        td.filterMode = cudaFilterModePoint;
        td.readMode = cudaReadModeElementType;
        td.normalizedCoords = 1;
        if (e == cudaSuccess) e = cudaCreateTextureObject(&b->tex[i], &rd, &td, nullptr);
    }
    if (e != cudaSuccess) {
        sky_free(b);
        RT_CHECK(ctx, e);
    }
    sky_release(ctx);      /* drains this context's stream; sharers of the old block keep it alive */
    ctx->sky = b;
    return RT_OK;
}

int rt_share_sky(rt_context* ctx, rt_context* owner) {
    if (!ctx || !owner || ctx == owner) return RT_ERR_ARG;
    if (!owner->sky) return fail(ctx, RT_ERR_STATE, "owner has no sky maps");
    if (owner->device != ctx->device) return fail(ctx, RT_ERR_ARG, "sky maps live on another device");
    if (ctx->sky == owner->sky) return RT_OK;
    RT_CHECK(ctx, cudaSetDevice(ctx->device));
    RtSkyBlock* b = owner->sky;
    b->users++;            /* before the release: the two may already share a block */
    sky_release(ctx);
    ctx->sky = b;
    return RT_OK;
}

int rt_measure_fp32_peak(rt_context* ctx, float* tflops) {
    if (!ctx || !tflops) return RT_ERR_ARG;
    RT_CHECK(ctx, cudaSetDevice(ctx->device));
    float ms = 0.f;
    double flops = 0.0;
    RT_CHECK(ctx, rt_launch_ffma_peak(ctx->n_sm, ctx->stream, &ms, &flops));
    *tflops = ms > 0.f ? (float)(flops / (ms * 1e-3) / 1e12) : 0.f;
    return RT_OK;
}

int rt_set_scene(rt_context* ctx, const rtb_Object* objects, int n_objects, const rtb_Light* lights, int n_lights) {
    if (!ctx || n_objects < 0 || (n_objects > 0 && !objects) || n_lights < 0 || n_lights > RT_MAX_LIGHTS ||
        (n_lights > 0 && !lights) || (unsigned int)n_objects > RT_TAG_INDEX)
        return RT_ERR_ARG;
    RT_CHECK(ctx, cudaSetDevice(ctx->device));
    if (n_lights != ctx->n_lights || (n_lights > 0 && memcmp(ctx->lights, lights, sizeof(rtb_Light) * (size_t)n_lights) != 0))
        ctx->lgb_stale = true, ctx->lgb_waited = false, ctx->lgrid_valid = false;   /* light grid / light bins follow the lights */
    for (int i = 0; i < n_lights; i++) ctx->lights[i] = lights[i];
    ctx->n_lights = n_lights;

    /* classify the change */
    int mode;   /* -1 nothing to upload, 1 refit (also covers material-only), 0 rebuild */
    if (!ctx->have_scene || (int)ctx->host_objs.size() != n_objects) {
        mode = 0;
    } else if (n_objects == 0 || memcmp(ctx->host_objs.data(), objects, sizeof(rtb_Object) * (size_t)n_objects) == 0) {
        mode = -1;
    } else {
        mode = 1;
        for (int i = 0; i < n_objects; i++)
            if (objects[i].type != ctx->host_objs[i].type) { mode = 0; break; }
    }
    if (mode < 0) return RT_OK;
    /* From here to the end the device copy is in flux: an error return leaves the context without a scene, so that
     * rt_render refuses to run and the next rt_set_scene rebuilds everything instead of classifying the change
     * against host data that never reached the device. */
    ctx->have_scene = false;
    ctx->lgb_stale = true, ctx->lgb_waited = false, ctx->lgrid_valid = false;
    ctx->wide_valid = false;
    for (int i = 0; i < RT_LGB_LIGHTS; i++) ctx->scene.lg_offsets[i] = ctx->scene.lg_entries[i] = nullptr;
    const bool settled_before = ctx->scene_settled;   /* the stream ends in this context's own FXAA / copy kernel */
    ctx->scene_settled = false;

    bool bvh_geometry_changed = (mode == 0);
    if (mode == 1) {
        for (int i = 0; i < n_objects && !bvh_geometry_changed; i++)
            if (objects[i].type != RTB_PLANE && !geometry_equal(objects[i], ctx->host_objs[i])) bvh_geometry_changed = true;
    }

    if (mode == 0) {
        ctx->host_planes.clear();
        for (int i = 0; i < n_objects; i++)
            if (objects[i].type != RTB_SPHERE && objects[i].type != RTB_TRIANGLE) ctx->host_planes.push_back(i);
    }
    const int n_planes = (int)ctx->host_planes.size();
    const int n_bvh = n_objects - n_planes;
    RtBuildBuffers& b = ctx->bb;

    if (n_objects > ctx->cap_objects) {
        const size_t c = (size_t)n_objects;
        RT_CHECK(ctx, regrow(&ctx->d_objs, c * 17));
        RT_CHECK(ctx, regrow(&b.mats, c * 2));
        RT_CHECK(ctx, regrow(&b.recs, c * 3));
        ctx->cap_objects = n_objects;
    }
    if (n_planes > ctx->cap_planes || !ctx->d_plane_idx) {
        const size_t c = (size_t)(n_planes > 0 ? n_planes : 1);
        RT_CHECK(ctx, regrow(&ctx->d_plane_idx, c));
        RT_CHECK(ctx, regrow(&b.planes, c * 2));
        ctx->cap_planes = (int)c;
    }
    if (!b.bounds) RT_CHECK(ctx, regrow(&b.bounds, (size_t)8));
    if (!b.depth) {
        RT_CHECK(ctx, regrow(&b.depth, (size_t)1));
        RT_CHECK(ctx, cudaMemsetAsync(b.depth, 0, sizeof(int), ctx->stream));
    }
    if (n_bvh > ctx->cap_bvh) {
        const size_t c = (size_t)n_bvh;
        RT_CHECK(ctx, regrow(&b.bvh_src, c));
        RT_CHECK(ctx, regrow(&b.prim_lo, c));
        RT_CHECK(ctx, regrow(&b.prim_hi, c));
        RT_CHECK(ctx, regrow(&b.keys, c));
        RT_CHECK(ctx, regrow(&b.vals, c));
        RT_CHECK(ctx, regrow(&b.tmp_keys, c));
        RT_CHECK(ctx, regrow(&b.tmp_vals, c));
        RT_CHECK(ctx, regrow(&b.hist, (size_t)rt_radix_hist_entries(n_bvh)));
        RT_CHECK(ctx, regrow(&b.children, c));
        RT_CHECK(ctx, regrow(&b.node_parent, c));
        RT_CHECK(ctx, regrow(&b.leaf_parent, c));
        RT_CHECK(ctx, regrow(&b.node_other, c));
        RT_CHECK(ctx, regrow(&b.flags, c));
        RT_CHECK(ctx, regrow(&b.node_aux, c * 12));
        RT_CHECK(ctx, cudaMemsetAsync(b.node_aux, 0, sizeof(float) * c * 12, ctx->stream));   /* empty arrival slots */
        b.refit_epoch = 0;
        RT_CHECK(ctx, regrow(&b.nodes, c * 16));
        RT_CHECK(ctx, regrow(&b.prims, c * 3));
        ctx->cap_bvh = n_bvh;
    }

    ctx->host_objs.assign(objects, objects + n_objects);
    b.objs = ctx->d_objs;
    b.n_objects = n_objects;
    b.plane_idx = ctx->d_plane_idx;
    b.n_planes = n_planes;
    b.n_bvh = n_bvh;
    /* Per-frame refit of a small scene (the launchKernel path: 133 objects move / change colour every frame):
     * the objects travel in the kernel parameters, no copy; d_objs is refreshed by the next full build. */
    b.objs_host = rt_build_by_value(&b, mode == 0 ? 0 : 1) ? (const float*)ctx->host_objs.data() : nullptr;
    b.after_own_post = settled_before ? 1 : 0;
    if (n_objects > 0 && !b.objs_host)
        RT_CHECK(ctx, cudaMemcpyAsync(ctx->d_objs, ctx->host_objs.data(), sizeof(rtb_Object) * (size_t)n_objects, cudaMemcpyHostToDevice,
                                      ctx->stream));
    if (mode == 0 && n_planes > 0)
        RT_CHECK(ctx, cudaMemcpyAsync(ctx->d_plane_idx, ctx->host_planes.data(), sizeof(int) * (size_t)n_planes,
                                      cudaMemcpyHostToDevice, ctx->stream));
    /* host_objs is the staging source of the async copies above: it is only rewritten by the next
     * rt_set_scene, and pageable cudaMemcpyAsync returns after the source has been staged */

    const bool timing = ctx->opt[RT_OPT_TIMING] != 0;
    if (timing) cudaEventRecord(ctx->ev[4], ctx->stream);
    /* material / plane-only changes still run the (cheap) setup + refit: one code path */
    b.refit_epoch = (b.refit_epoch + 1u) & 0xFFFFFu;
    if (b.refit_epoch == 0u) {             /* the stamp wrapped: no slot may still carry this value from 2^20 refits ago */
        if (b.node_aux && ctx->cap_bvh > 0)
            RT_CHECK(ctx, cudaMemsetAsync(b.node_aux, 0, sizeof(float) * (size_t)ctx->cap_bvh * 12, ctx->stream));
        b.refit_epoch = 1u;
    }
    RT_CHECK(ctx, rt_build_bvh(&b, mode == 0 ? 0 : 1, ctx->opt[RT_OPT_FORCE_RADIX], ctx->opt[RT_OPT_BUILDER], ctx->stream));
    if (timing) { cudaEventRecord(ctx->ev[5], ctx->stream); ctx->build_timed = true; }
    (void)bvh_geometry_changed;
    if (mode == 0 && n_bvh > 1 && !rt_trace_small_supported(n_bvh, n_objects, n_planes)) {
        /* The HBM tracer keeps RT_STACK_SIZE entries per thread; a hierarchy deeper than that (possible with
         * Karras splits over heavily duplicated Morton codes) would lose pushes.  New topology is the one path
         * that already costs a build, so the 4-byte read-back is affordable here: fail loudly instead. */
        int depth = 0;
        RT_CHECK(ctx, cudaMemcpyAsync(&depth, b.depth, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
        RT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
        if (depth > RT_STACK_SIZE) {
            ctx->have_scene = false;
            char msg[160];
            snprintf(msg, sizeof(msg), "LBVH depth %d exceeds the traversal stack (%d entries)", depth, RT_STACK_SIZE);
            return fail(ctx, RT_ERR_UNSUPPORTED, msg);
        }
    }

    ctx->scene.nodes = (const float4*)b.nodes;
    ctx->scene.prims = b.prims;
    ctx->scene.planes = b.planes;
    ctx->scene.mats = b.mats;
    ctx->scene.recs = b.recs;
    ctx->scene.depth = b.depth;
    ctx->scene.n_bvh_prims = n_bvh;
    ctx->scene.n_planes = n_planes;
    ctx->scene.n_objects = n_objects;
    ctx->scene.root = n_bvh == 0 ? RT_NODE_SENTINEL : (n_bvh == 1 ? ~0 : 0);
    ctx->have_scene = true;
    return RT_OK;
}

static int fill_frame(rt_context* ctx, const rt_frame* fr, const rt_shard* sh, RtFrameDev* f) {
    memset(f, 0, sizeof(*f));
    memcpy(f->pos, &fr->cam.pos, 12);
    memcpy(f->LD, &fr->cam.LD, 12);
    memcpy(f->LU, &fr->cam.LU, 12);
    memcpy(f->RD, &fr->cam.RD, 12);
    memcpy(f->RU, &fr->cam.RU, 12);
    memcpy(f->ambient, fr->ambient, 12);
    memcpy(f->skyVars, fr->skyVars, 16);
    f->dayTime = fr->dayTime;
    f->imgw = fr->imgw;
    f->imgh = fr->imgh;
    f->n_lights = ctx->n_lights;
    for (int i = 0; i < ctx->n_lights; i++) {
        const rtb_Light& l = ctx->lights[i];
        f->lights[i].px = l.pos.x; f->lights[i].py = l.pos.y; f->lights[i].pz = l.pos.z;
        f->lights[i].intensity = l.intensity;
        f->lights[i].cr = l.color.x; f->lights[i].cg = l.color.y; f->lights[i].cb = l.color.z;
    }
    for (int i = 0; i < 4; i++) f->sky[i] = ctx->sky->tex[i];
    f->sky_all4 = ctx->opt[RT_OPT_SKY_ALL4];
    if (sh && sh->n_shards > 1) {
        /* tiles of at least 2x2 pixels: a pixel's 3x3 neighbourhood then spans at most 2x2 tiles, which is what
         * rt_shard_needs (trace_common.cuh) tests */
        if (sh->tile_w < 2 || sh->tile_h < 2 || sh->shard_id < 0 || sh->shard_id >= sh->n_shards) return RT_ERR_ARG;
        f->tile_w = sh->tile_w; f->tile_h = sh->tile_h;
        f->n_shards = sh->n_shards; f->shard_id = sh->shard_id; f->stride = sh->stride;
    } else {
        f->tile_w = f->tile_h = 1; f->n_shards = 1;
    }
    return RT_OK;
}

/* Light grid of a big scene (light_grid.cu): per light count pass + scan + fill pass.  No host synchronisation once a
 * light's entry buffer exists: the fill runs with the buffer's capacity, a one-thread kernel leaves the verdict
 * (complete / overflowed) on the device where the tracer reads it, and the host learns of it through an asynchronous
 * copy it looks at when the next frame is rendered (lgb_poll) -- an overflowed light traverses meanwhile and gets a
 * bigger buffer then.  Only the first build of a light reads the entry count back: it has to size an allocation, and
 * cudaMalloc synchronises anyway. */
static int build_light_grid(rt_context* ctx) {
    for (int i = 0; i < RT_LGB_LIGHTS; i++) ctx->scene.lg_offsets[i] = ctx->scene.lg_entries[i] = nullptr;
    ctx->lgb_stale = false;
    const int n = ctx->scene.n_bvh_prims;
    if (n < 2) return RT_OK;
    const size_t bins = rt_lgb_bins();
    if (!ctx->d_lgb_counts) {
        RT_CHECK(ctx, regrow(&ctx->d_lgb_counts, bins));
        RT_CHECK(ctx, regrow(&ctx->d_lgb_sums, rt_lgb_sums()));
        RT_CHECK(ctx, regrow(&ctx->d_lgb_fail, (size_t)1));
        RT_CHECK(ctx, regrow(&ctx->d_lgb_state, (size_t)4 * RT_LGB_LIGHTS));
        RT_CHECK(ctx, cudaMemsetAsync(ctx->d_lgb_state, 0, sizeof(unsigned int) * 4 * RT_LGB_LIGHTS, ctx->stream));
        RT_CHECK(ctx, cudaHostAlloc((void**)&ctx->h_lgb_state, sizeof(unsigned int) * 4 * RT_LGB_LIGHTS, cudaHostAllocDefault));
        for (int i = 0; i < RT_LGB_LIGHTS; i++) RT_CHECK(ctx, cudaEventCreateWithFlags(&ctx->lgb_read[i], cudaEventDisableTiming));
    }
    ctx->scene.lg_state = ctx->d_lgb_state;
    const size_t budget = (size_t)64 << 20;          /* entries per light: 256 MB */
    const int lit = ctx->n_lights < RT_LGB_LIGHTS ? ctx->n_lights : RT_LGB_LIGHTS;
    for (int li = 0; li < lit; li++) {
        if (!ctx->d_lgb_offsets[li]) RT_CHECK(ctx, regrow(&ctx->d_lgb_offsets[li], bins + 1));
        const float lp[3] = {ctx->lights[li].pos.x, ctx->lights[li].pos.y, ctx->lights[li].pos.z};
        RT_CHECK(ctx, rt_lgb_count(&ctx->scene, lp, ctx->d_lgb_counts, ctx->d_lgb_offsets[li], ctx->d_lgb_sums,
                                   ctx->d_lgb_fail, ctx->stream));
        if (!ctx->d_lgb_entries[li]) {
            unsigned int total = 0, failed = 0;
            RT_CHECK(ctx, cudaMemcpyAsync(&total, ctx->d_lgb_offsets[li] + bins, sizeof(unsigned int), cudaMemcpyDeviceToHost, ctx->stream));
            RT_CHECK(ctx, cudaMemcpyAsync(&failed, ctx->d_lgb_fail, sizeof(unsigned int), cudaMemcpyDeviceToHost, ctx->stream));
            RT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
            if (failed || (size_t)total > budget) continue;          /* this light keeps traversing */
            size_t c = (size_t)total + (size_t)total / 4 + 1024;     /* head-room for lights that move */
            if (c > budget) c = budget;
            RT_CHECK(ctx, regrow(&ctx->d_lgb_entries[li], c));
            ctx->cap_lgb_entries[li] = c;
        }
        RT_CHECK(ctx, rt_lgb_fill(&ctx->scene, lp, ctx->d_lgb_counts, ctx->d_lgb_offsets[li], ctx->d_lgb_entries[li],
                                  (unsigned int)ctx->cap_lgb_entries[li], ctx->d_lgb_fail, ctx->stream));
        RT_CHECK(ctx, rt_lgb_finish(ctx->d_lgb_offsets[li] + bins, ctx->d_lgb_fail, (unsigned int)ctx->cap_lgb_entries[li],
                                    ctx->d_lgb_state + 4 * li, ctx->stream));
        RT_CHECK(ctx, cudaMemcpyAsync(ctx->h_lgb_state + 4 * li, ctx->d_lgb_state + 4 * li, sizeof(unsigned int) * 4,
                                      cudaMemcpyDeviceToHost, ctx->stream));
        RT_CHECK(ctx, cudaEventRecord(ctx->lgb_read[li], ctx->stream));
        ctx->lgb_read_pending[li] = true;
        ctx->scene.lg_offsets[li] = ctx->d_lgb_offsets[li];
        ctx->scene.lg_entries[li] = ctx->d_lgb_entries[li];
    }
    return RT_OK;
}

/* Has a build's verdict arrived, and did a light's lists outgrow its buffer?  Then drop the buffer (after the frames
 * that may still read it) and have the next frame build the grid again, sized afresh.  Never waits for a verdict. */
static int lgb_poll(rt_context* ctx) {
    for (int li = 0; li < RT_LGB_LIGHTS; li++) {
        if (!ctx->lgb_read_pending[li] || cudaEventQuery(ctx->lgb_read[li]) != cudaSuccess) continue;
        ctx->lgb_read_pending[li] = false;
        const unsigned int* st = ctx->h_lgb_state + 4 * li;
        if (st[0] == 0u && st[2] == 0u && (size_t)st[1] > ctx->cap_lgb_entries[li]) {
            RT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
            RT_CHECK(ctx, regrow(&ctx->d_lgb_entries[li], (size_t)0));
            ctx->cap_lgb_entries[li] = 0;
            ctx->lgb_stale = true;
            ctx->lgb_waited = true;
        }
    }
    return RT_OK;
}

static int render_impl(rt_context* ctx, const rt_frame* frame, const rt_shard* shard, unsigned int* out_dev,
                       int* hit_ids_dev, const RtPeerSync* peer) {
    if (!ctx || !frame || !out_dev || frame->imgw < 2 || frame->imgh < 2) return RT_ERR_ARG;
    if (!ctx->have_scene) return fail(ctx, RT_ERR_STATE, "rt_render before rt_set_scene");
    if (!ctx->sky) return fail(ctx, RT_ERR_STATE, "rt_render before rt_set_sky");
    RT_CHECK(ctx, cudaSetDevice(ctx->device));
    const size_t npx = (size_t)frame->imgw * frame->imgh;
    if (npx > ctx->cap_base) {
        RT_CHECK(ctx, regrow(&ctx->d_base, npx));
        ctx->cap_base = npx;
        ctx->tmap_ptr = nullptr;
    }
    RtFrameDev f;
    if (fill_frame(ctx, frame, shard, &f) != RT_OK) return fail(ctx, RT_ERR_ARG, "bad shard");
    const bool timing = ctx->opt[RT_OPT_TIMING] != 0;
    const int count = ctx->opt[RT_OPT_COUNTERS] != 0;
    /* AA off, whole frame: the reference's second kernel only copies baseImage to the result (kernel.cu:275-278),
     * so the tracer writes the caller's buffer itself and no post kernel runs.  (Shards keep the copy: it is what
     * confines a rank's stores into a shared / peer frame to its own tiles.) */
    const bool direct = !frame->alias && f.n_shards <= 1 && !peer;
    unsigned int* const image = direct ? out_dev : ctx->d_base;
    ctx->last_base = image;
    if (count) RT_CHECK(ctx, cudaMemsetAsync(ctx->d_counters, 0, sizeof(RtCountersDev), ctx->stream));
    if (timing) cudaEventRecord(ctx->ev[0], ctx->stream);
    /* RT_OPT_TRACE_IMPL: 0 = auto (shared-memory tracer when the scene fits, else the plain one-thread-per-pixel tracer
     * over the HBM-resident LBVH), 1 = the plain tracer for any scene (tests: it has no lists, no packets), 2 =
     * shared-memory tracer with its deep-hierarchy fallback forced (tests), 3 = the persistent tracer of trace_big.cu
     * (packets per RT_OPT_PACKET; measured slower than the plain kernel on the terrain, hence not the default), 4 = the
     * plain tracer with packet traversal of the primary rays */
    const int impl = ctx->opt[RT_OPT_TRACE_IMPL];
    const bool brute = ctx->opt[RT_OPT_TRAVERSAL] == 1;
    const bool small = !brute && impl != 1 && impl != 3 && impl != 4 &&
                       rt_trace_small_supported(ctx->scene.n_bvh_prims, ctx->scene.n_objects, ctx->scene.n_planes);
    if (small) {
        const size_t need = rt_trace_small_list_bytes(frame->imgw, frame->imgh) / sizeof(unsigned short);
        if (need > ctx->cap_tile_lists) {
            RT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));   /* a frame in flight may still read the old lists */
            RT_CHECK(ctx, regrow(&ctx->d_tile_lists, need));
            ctx->cap_tile_lists = need;
        }
        if (!ctx->d_lgrid && ctx->opt[RT_OPT_SHADOW_BINS] != 1) {
            RT_CHECK(ctx, regrow(&ctx->d_lgrid, RT_LGRID_BYTES / sizeof(uint4)));
            ctx->lgrid_valid = false;
        }
        const bool use_lgrid = ctx->opt[RT_OPT_SHADOW_BINS] != 1;
        const bool build_lgrid = use_lgrid && !(ctx->lgrid_valid && ctx->opt[RT_OPT_LIGHT_BINS_REUSE] == 0);
        /* Order history: a frame files every patch under its measured cost; the next frame of the same size
         * and shard starts with the most expensive ones (frame-to-frame coherence of an interactive renderer;
         * the first frame of a kind walks bottom-up).  Pure scheduling: pixels do not depend on it. */
        const unsigned int* ord_r = nullptr;
        unsigned int* ord_w = nullptr;
        /* Only where the order matters: below ~32 patches per resident warp (1080p: 14).  A 4K frame (55 per
         * warp) balances by itself bottom-up, and there filing + indirection cost 3 % (measured). */
        const size_t ob = rt_trace_small_order_bytes(frame->imgw, frame->imgh) / sizeof(unsigned int);
        const size_t patches = rt_trace_small_patches(frame->imgw, frame->imgh);
        const int key[7] = {frame->imgw, frame->imgh, f.n_shards, f.shard_id, f.tile_w, f.tile_h, f.stride};
        if (ctx->opt[RT_OPT_ORDER] == 0 && patches / (size_t)f.n_shards < (size_t)32 * ctx->n_sm * 32) {
            if (2 * ob > ctx->cap_order) {
                RT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
                RT_CHECK(ctx, regrow(&ctx->d_order, 2 * ob));
                ctx->cap_order = 2 * ob;
                ctx->order_key[0] = 0;
            }
            const bool same = memcmp(ctx->order_key, key, sizeof(key)) == 0;
            if (!same) {
                RT_CHECK(ctx, cudaMemsetAsync(ctx->d_order, 0, RT_SMALL_ORDER_HEAD * sizeof(unsigned int), ctx->stream));
                RT_CHECK(ctx, cudaMemsetAsync(ctx->d_order + ob, 0, RT_SMALL_ORDER_HEAD * sizeof(unsigned int), ctx->stream));
            }
            if (same) ord_r = ctx->d_order + (size_t)ctx->order_cur * ob;
            ord_w = ctx->d_order + (size_t)(1 - ctx->order_cur) * ob;
            ctx->order_cur = 1 - ctx->order_cur;
            memcpy(ctx->order_key, key, sizeof(key));
        }
        RT_CHECK(ctx, rt_launch_trace_small(&ctx->scene, &f, image, hit_ids_dev, ctx->d_counters,
                                            ctx->d_tile_counter, ctx->d_tile_lists, count,
                                            ctx->opt[RT_OPT_TRACE_IMPL] == 2, ctx->n_sm, ord_r, ord_w,
                                            ctx->scene_settled ? 1 : 0, use_lgrid ? ctx->d_lgrid : (uint4*)nullptr,
                                            build_lgrid ? 1 : 0, ctx->opt[RT_OPT_MIRROR_BINS] == 0,
                                            ctx->opt[RT_OPT_COMPACTION] != 0, ctx->stream));
        if (use_lgrid) ctx->lgrid_valid = true;
    }
    else {
        /* The grid costs ~2 ms to build for a million triangles and saves ~0.4 ms per 4K frame: it is built when a
         * second frame finds scene and lights unchanged (RT_OPT_SHADOW_BINS 2: at once), so lights that move every
         * frame never pay for it and keep traversing. */
        const int sb = ctx->opt[RT_OPT_SHADOW_BINS];
        bool grid = sb != 1 && ctx->opt[RT_OPT_TRAVERSAL] != 1;
        if (grid) {
            const int prc = lgb_poll(ctx);
            if (prc != RT_OK) return prc;
        }
        if (grid && ctx->lgb_stale) {
            if (sb == 2 || ctx->lgb_waited) {
                const int rc = build_light_grid(ctx);
                if (rc != RT_OK) return rc;
            } else {
                ctx->lgb_waited = true;
                grid = false;
            }
        }
        RtSceneDev sc = ctx->scene;
        if (!grid)
            for (int i = 0; i < RT_LGB_LIGHTS; i++) sc.lg_offsets[i] = sc.lg_entries[i] = nullptr;
        unsigned int* const errw = ctx->d_tile_counter + RT_DEVICE_ERROR_WORD;
        sc.err = errw;
        /* RT_OPT_WIDE_BVH 1: the plain tracer walks the 4-wide view of the hierarchy (measured 1.42 ms against the binary
         * walk's 1.16 ms on the terrain at 3840x2160, profiles/r02_big_probe.jsonl: opt-in) */
        if (!brute && impl != 3 && ctx->opt[RT_OPT_WIDE_BVH] == 1 && sc.n_bvh_prims > 1) {
            const size_t need = (size_t)RT_WIDE_F4 * (size_t)(sc.n_bvh_prims - 1);
            if (need > ctx->cap_wide) {
                RT_CHECK(ctx, regrow(&ctx->d_wide, need));
                ctx->cap_wide = need;
                ctx->wide_valid = false;
            }
            if (!ctx->wide_valid) {
                RT_CHECK(ctx, rt_collapse_wide((const float*)ctx->scene.nodes, sc.n_bvh_prims, ctx->d_wide, ctx->stream));
                ctx->wide_valid = true;
            }
            sc.wide = ctx->d_wide;
        }
        if (brute || impl != 3)
            RT_CHECK(ctx, rt_launch_trace(&sc, &f, image, hit_ids_dev, ctx->d_counters, errw, brute, count,
                                          impl == 4 && !brute && ctx->opt[RT_OPT_PACKET] == 0, ctx->stream));
        else
            RT_CHECK(ctx, rt_launch_trace_big(&sc, &f, image, hit_ids_dev, ctx->d_counters, ctx->d_tile_counter, errw, count,
                                              ctx->opt[RT_OPT_PACKET] == 0, ctx->opt[RT_OPT_BIG_STACK] != 0, ctx->n_sm,
                                              ctx->stream));
    }
    if (timing) { cudaEventRecord(ctx->ev[1], ctx->stream); cudaEventRecord(ctx->ev[2], ctx->stream); }
    int rc = RT_OK;
    if (frame->alias) rc = run_fxaa(ctx, ctx->d_base, out_dev, frame->imgw, frame->imgh, shard, peer);
    else if (!direct) RT_CHECK(ctx, rt_launch_copy(ctx->d_base, out_dev, frame->imgw, frame->imgh, shard, peer, ctx->stream));
    if (rc != RT_OK) return rc;
    if (timing) { cudaEventRecord(ctx->ev[3], ctx->stream); ctx->timing_valid = true; }
    /* `settled` promises the next frame that the stream ends in this context's own post kernel (its pre-pass and
     * refit then overlap it); after a direct frame it ends in the tracer, which still reads what they write */
    ctx->scene_settled = !direct;
    return RT_OK;
}

int rt_render(rt_context* ctx, const rt_frame* frame, const rt_shard* shard, unsigned int* out_dev, int* hit_ids_dev) {
    return render_impl(ctx, frame, shard, out_dev, hit_ids_dev, nullptr);
}

int rt_render_to_host(rt_context* ctx, const rt_frame* frame, const rt_shard* shard, unsigned int* out_host,
                      int* hit_ids_host) {
    if (!ctx || !frame || !out_host) return RT_ERR_ARG;
    RT_CHECK(ctx, cudaSetDevice(ctx->device));
    const size_t npx = (size_t)frame->imgw * frame->imgh;
    if (npx > ctx->cap_stage) {
        RT_CHECK(ctx, regrow(&ctx->d_stage_out, npx));
        RT_CHECK(ctx, regrow(&ctx->d_stage_ids, npx));
        ctx->cap_stage = npx;
    }
    if (shard && shard->n_shards > 1) RT_CHECK(ctx, cudaMemsetAsync(ctx->d_stage_out, 0, npx * 4, ctx->stream));
    int rc = rt_render(ctx, frame, shard, ctx->d_stage_out, hit_ids_host ? ctx->d_stage_ids : nullptr);
    if (rc != RT_OK) return rc;
    RT_CHECK(ctx, cudaMemcpyAsync(out_host, ctx->d_stage_out, npx * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (hit_ids_host)
        RT_CHECK(ctx, cudaMemcpyAsync(hit_ids_host, ctx->d_stage_ids, npx * 4, cudaMemcpyDeviceToHost, ctx->stream));
    RT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return RT_OK;
}

/* ---------------------------------------------------------------- asynchronous read-back
 * The reference hands the finished frame to the display without it ever leaving the device (mapped PBO,
 * main.cpp:141-165).  A headless consumer wants it in host memory; so that the copy of frame i (8.3 MB at 1080p:
 * about as long over PCIe as the frame takes to render) hides behind the rendering of frames i+1.., copies run on
 * a stream of their own: slot k = one frame in flight.
 *   rt_render_to_host_async(ctx, frame, shard, pinned, k)    render into the slot's staging frame, then copy
 *   rt_readback_async(ctx, dev, n, pinned, k)                copy a frame the caller rendered itself (launchKernel)
 *   rt_wait_slot(ctx, k)                                     the slot's frame is in host memory */

static int slot_prepare(rt_context* ctx, int slot) {
    if (slot < 0 || slot >= RT_MAX_SLOTS) return fail(ctx, RT_ERR_ARG, "slot out of range");
    RT_CHECK(ctx, cudaSetDevice(ctx->device));
    if (!ctx->copy_stream) RT_CHECK(ctx, cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    RtSlot& sl = ctx->slots[slot];
    if (!sl.ready) RT_CHECK(ctx, cudaEventCreateWithFlags(&sl.ready, cudaEventDisableTiming));
    if (!sl.copied) RT_CHECK(ctx, cudaEventCreateWithFlags(&sl.copied, cudaEventDisableTiming));
    return RT_OK;
}

int rt_readback_async(rt_context* ctx, const unsigned int* src_dev, size_t n_pixels, unsigned int* dst_host, int slot) {
    if (!ctx || !src_dev || !dst_host || n_pixels == 0) return RT_ERR_ARG;
    int rc = slot_prepare(ctx, slot);
    if (rc != RT_OK) return rc;
    RtSlot& sl = ctx->slots[slot];
    RT_CHECK(ctx, cudaEventRecord(sl.ready, ctx->stream));
    RT_CHECK(ctx, cudaStreamWaitEvent(ctx->copy_stream, sl.ready, 0));
    RT_CHECK(ctx, cudaMemcpyAsync(dst_host, src_dev, n_pixels * sizeof(unsigned int), cudaMemcpyDeviceToHost, ctx->copy_stream));
    RT_CHECK(ctx, cudaEventRecord(sl.copied, ctx->copy_stream));
    sl.used = true;
    return RT_OK;
}

int rt_render_to_host_async(rt_context* ctx, const rt_frame* frame, const rt_shard* shard, unsigned int* out_host,
                            int slot) {
    if (!ctx || !frame || !out_host) return RT_ERR_ARG;
    int rc = slot_prepare(ctx, slot);
    if (rc != RT_OK) return rc;
    RtSlot& sl = ctx->slots[slot];
    const size_t npx = (size_t)frame->imgw * frame->imgh;
    /* the staging frame is rewritten: its previous copy must have drained (device-side wait, the host runs on) */
    if (sl.used) RT_CHECK(ctx, cudaStreamWaitEvent(ctx->stream, sl.copied, 0));
    if (npx > sl.cap) {
        RT_CHECK(ctx, regrow(&sl.dev, npx));
        sl.cap = npx;
    }
    if (shard && shard->n_shards > 1) RT_CHECK(ctx, cudaMemsetAsync(sl.dev, 0, npx * 4, ctx->stream));
    rc = rt_render(ctx, frame, shard, sl.dev, nullptr);
    if (rc != RT_OK) return rc;
    return rt_readback_async(ctx, sl.dev, npx, out_host, slot);
}

int rt_wait_slot(rt_context* ctx, int slot) {
    if (!ctx || slot < 0 || slot >= RT_MAX_SLOTS) return RT_ERR_ARG;
    RtSlot& sl = ctx->slots[slot];
    if (!sl.used) return RT_OK;
    RT_CHECK(ctx, cudaSetDevice(ctx->device));
    RT_CHECK(ctx, cudaEventSynchronize(sl.copied));
    return RT_OK;
}

/* page-locked host memory for the read-back targets (a pageable target makes cudaMemcpyAsync synchronous) */
int rt_host_alloc(void** out, size_t bytes) {
    if (!out || bytes == 0) return RT_ERR_ARG;
    return cudaHostAlloc(out, bytes, cudaHostAllocDefault) == cudaSuccess ? RT_OK : RT_ERR_CUDA;
}

int rt_host_free(void* p) {
    if (!p) return RT_OK;
    return cudaFreeHost(p) == cudaSuccess ? RT_OK : RT_ERR_CUDA;
}

int rt_get_base_image(rt_context* ctx, unsigned int** base_dev) {
    if (!ctx || !base_dev) return RT_ERR_ARG;
    *base_dev = ctx->last_base ? ctx->last_base : ctx->d_base;   /* AA off, whole frame: the caller's own buffer */
    return RT_OK;
}

int rt_fxaa(rt_context* ctx, const unsigned int* in_dev, unsigned int* out_dev, int imgw, int imgh,
            const rt_shard* shard) {
    if (!ctx || !in_dev || !out_dev || imgw < 1 || imgh < 1) return RT_ERR_ARG;
    RT_CHECK(ctx, cudaSetDevice(ctx->device));
    return run_fxaa(ctx, in_dev, out_dev, imgw, imgh, shard, nullptr);
}

int rt_shard_tile_count(const rt_shard* sh, int imgw, int imgh) {
    if (!sh || sh->n_shards < 1 || sh->tile_w < 1 || sh->tile_h < 1) return -1;
    const int tiles_x = (imgw + sh->tile_w - 1) / sh->tile_w, tiles_y = (imgh + sh->tile_h - 1) / sh->tile_h;
    return ((tiles_x + sh->n_shards - 1) / sh->n_shards) * tiles_y;   /* slots, some may be empty */
}

int rt_present(rt_context* ctx, const unsigned int* in_dev, int out_w, int out_h, int factor, int mode,
               unsigned int* out_dev) {
    if (!ctx || !in_dev || !out_dev || out_w < 1 || out_h < 1 || factor < 1 || factor > 16 || mode < 0 || mode > 1)
        return RT_ERR_ARG;
    RT_CHECK(ctx, cudaSetDevice(ctx->device));
    RT_CHECK(ctx, rt_launch_present(in_dev, out_dev, out_w, out_h, factor, mode, ctx->stream));
    return RT_OK;
}

/* ---------------------------------------------------------------- peer frame (peer.cu) */

struct rt_peer_frame {
    unsigned int* base = nullptr;    /* n_buffers frames, then the flag block */
    size_t n_pixels = 0;
    int n_buffers = 0;
    bool owner = false;
    unsigned int timeout_ms = RT_PEER_TIMEOUT_MS_DEFAULT;
    unsigned int* flags() const { return base + (size_t)n_buffers * n_pixels; }
};

int rt_peer_frame_create(rt_context* ctx, size_t n_pixels, int n_buffers, rt_peer_frame** out, unsigned char* handle) {
    if (!ctx || !out || !handle || n_pixels == 0 || n_buffers < 1 || n_buffers > 4) return RT_ERR_ARG;
    static_assert(sizeof(cudaIpcMemHandle_t) == RT_PEER_HANDLE_BYTES, "handle size");
    RT_CHECK(ctx, cudaSetDevice(ctx->device));
    rt_peer_frame* pf = new rt_peer_frame;
    pf->n_pixels = n_pixels;
    pf->n_buffers = n_buffers;
    pf->owner = true;
    const size_t bytes = ((size_t)n_buffers * n_pixels + RT_PEER_FLAG_UINTS) * sizeof(unsigned int);
    cudaError_t e = cudaMalloc((void**)&pf->base, bytes);
    if (e == cudaSuccess) e = cudaMemset(pf->base, 0, bytes);
    cudaIpcMemHandle_t h;
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, pf->base);
    if (e != cudaSuccess) {
        if (pf->base) cudaFree(pf->base);
        delete pf;
        RT_CHECK(ctx, e);
    }
    memcpy(handle, &h, sizeof(h));
    *out = pf;
    return RT_OK;
}

int rt_peer_frame_open(rt_context* ctx, const unsigned char* handle, size_t n_pixels, int n_buffers,
                       rt_peer_frame** out) {
    if (!ctx || !out || !handle || n_pixels == 0 || n_buffers < 1 || n_buffers > 4) return RT_ERR_ARG;
    RT_CHECK(ctx, cudaSetDevice(ctx->device));
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    void* p = nullptr;
    RT_CHECK(ctx, cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    rt_peer_frame* pf = new rt_peer_frame;
    pf->base = (unsigned int*)p;
    pf->n_pixels = n_pixels;
    pf->n_buffers = n_buffers;
    pf->owner = false;
    *out = pf;
    return RT_OK;
}

unsigned int* rt_peer_frame_buffer(rt_peer_frame* pf, unsigned int seq) {
    if (!pf) return nullptr;
    return pf->base + (size_t)(seq % (unsigned int)pf->n_buffers) * pf->n_pixels;
}

int rt_peer_frame_acquire(rt_context* ctx, rt_peer_frame* pf, unsigned int seq) {
    if (!ctx || !pf) return RT_ERR_ARG;
    RT_CHECK(ctx, cudaSetDevice(ctx->device));
    RT_CHECK(ctx, rt_launch_peer_acquire(pf->flags(), seq - (unsigned int)pf->n_buffers, pf->timeout_ms, ctx->stream));
    return RT_OK;
}

/* acquire + sharded render into the frame's buffer + signal, the two halves of the hand-shake carried by the frame's
 * last kernel (RtPeerSync): three launches per producer frame */
int rt_render_peer(rt_context* ctx, const rt_frame* frame, const rt_shard* shard, rt_peer_frame* pf, int rank,
                   unsigned int seq, int* hit_ids_dev) {
    if (!ctx || !pf || !frame || rank < 0 || rank >= 16) return RT_ERR_ARG;
    if ((size_t)frame->imgw * (size_t)frame->imgh > pf->n_pixels) return fail(ctx, RT_ERR_ARG, "frame larger than the peer buffer");
    RtPeerSync ps;
    ps.flags = pf->flags();
    ps.ticket = ctx->d_tile_counter + 48;      /* its own 32-byte sector of the context's counter block */
    ps.acquire_seq = seq - (unsigned int)pf->n_buffers;
    ps.seq = seq;
    ps.rank = rank;
    ps.timeout_ms = pf->timeout_ms;
    return render_impl(ctx, frame, shard, rt_peer_frame_buffer(pf, seq), hit_ids_dev, &ps);
}

int rt_peer_frame_set_timeout(rt_peer_frame* pf, unsigned int timeout_ms) {
    if (!pf || timeout_ms == 0) return RT_ERR_ARG;
    pf->timeout_ms = timeout_ms;
    return RT_OK;
}

int rt_peer_frame_reset_error(rt_context* ctx, rt_peer_frame* pf) {
    if (!ctx || !pf) return RT_ERR_ARG;
    RT_CHECK(ctx, cudaSetDevice(ctx->device));
    RT_CHECK(ctx, rt_launch_peer_reset_error(pf->flags(), ctx->stream));
    return RT_OK;
}

int rt_peer_frame_signal(rt_context* ctx, rt_peer_frame* pf, int rank, unsigned int seq) {
    if (!ctx || !pf || rank < 0 || rank >= 16) return RT_ERR_ARG;
    RT_CHECK(ctx, cudaSetDevice(ctx->device));
    RT_CHECK(ctx, rt_launch_peer_signal(pf->flags(), rank, seq, ctx->stream));
    return RT_OK;
}

int rt_peer_frame_wait(rt_context* ctx, rt_peer_frame* pf, int n_ranks, unsigned int seq) {
    if (!ctx || !pf || n_ranks < 1 || n_ranks > 16) return RT_ERR_ARG;
    RT_CHECK(ctx, cudaSetDevice(ctx->device));
    RT_CHECK(ctx, rt_launch_peer_wait(pf->flags(), n_ranks, seq, pf->timeout_ms, ctx->stream));
    return RT_OK;
}

int rt_peer_frame_release(rt_context* ctx, rt_peer_frame* pf, unsigned int seq) {
    if (!ctx || !pf) return RT_ERR_ARG;
    RT_CHECK(ctx, cudaSetDevice(ctx->device));
    RT_CHECK(ctx, rt_launch_peer_release(pf->flags(), seq, ctx->stream));
    return RT_OK;
}

int rt_peer_frame_error(rt_context* ctx, rt_peer_frame* pf, unsigned int* err) {
    if (!ctx || !pf || !err) return RT_ERR_ARG;
    RT_CHECK(ctx, cudaSetDevice(ctx->device));
    RT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    RT_CHECK(ctx, cudaMemcpy(err, pf->flags() + RT_PEER_FLAG_ERROR, sizeof(unsigned int), cudaMemcpyDeviceToHost));
    return RT_OK;
}

int rt_peer_frame_close(rt_context* ctx, rt_peer_frame* pf) {
    if (!ctx || !pf) return RT_ERR_ARG;
    RT_CHECK(ctx, cudaSetDevice(ctx->device));
    RT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    if (pf->owner) cudaFree(pf->base);
    else cudaIpcCloseMemHandle(pf->base);
    delete pf;
    return RT_OK;
}

int rt_pack_tiles(rt_context* ctx, const unsigned int* frame_dev, int imgw, int imgh, const rt_shard* sh,
                  unsigned int* packed_dev) {
    if (!ctx || !frame_dev || !packed_dev || !sh) return RT_ERR_ARG;
    RT_CHECK(ctx, cudaSetDevice(ctx->device));
    RT_CHECK(ctx, rt_launch_pack_tiles(frame_dev, packed_dev, imgw, imgh, sh, 0, ctx->stream));
    return RT_OK;
}

int rt_unpack_tiles(rt_context* ctx, const unsigned int* packed_dev, int imgw, int imgh, const rt_shard* sh,
                    unsigned int* frame_dev) {
    if (!ctx || !frame_dev || !packed_dev || !sh) return RT_ERR_ARG;
    RT_CHECK(ctx, cudaSetDevice(ctx->device));
    RT_CHECK(ctx, rt_launch_pack_tiles(frame_dev, (unsigned int*)packed_dev, imgw, imgh, sh, 1, ctx->stream));
    return RT_OK;
}

int rt_synchronize(rt_context* ctx) {
    if (!ctx) return RT_ERR_ARG;
    RT_CHECK(ctx, cudaSetDevice(ctx->device));
    RT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return RT_OK;
}

int rt_get_counters(rt_context* ctx, unsigned long long out[RT_CNT_COUNT]) {
    if (!ctx || !out) return RT_ERR_ARG;
    RT_CHECK(ctx, cudaSetDevice(ctx->device));
    RT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    RT_CHECK(ctx, cudaMemcpy(out, ctx->d_counters, sizeof(RtCountersDev), cudaMemcpyDeviceToHost));
    return RT_OK;
}

/* diagnostics: the light bins of the last frame of the shared-memory tracer (RT_LGRID_LIGHTS x 6 x G x G bins of
 * 8 ushorts: [0] count or 0xFFFF, [1..7] leaf offsets).  Returns the bytes copied (<= bytes), 0 if there are none,
 * minus the status code on an error. */
long long rt_debug_light_bins(rt_context* ctx, void* out, size_t bytes) {
    if (!ctx || !out) return -(long long)RT_ERR_ARG;
    if (!ctx->d_lgrid) return 0;
    if (cudaSetDevice(ctx->device) != cudaSuccess || cudaStreamSynchronize(ctx->stream) != cudaSuccess) return -(long long)RT_ERR_CUDA;
    const size_t n = bytes < RT_LGRID_BYTES ? bytes : RT_LGRID_BYTES;
    if (cudaMemcpy(out, ctx->d_lgrid, n, cudaMemcpyDeviceToHost) != cudaSuccess) return -(long long)RT_ERR_CUDA;
    return (long long)n;
}

unsigned long long rt_launch_count(void) { return __atomic_load_n(&g_rt_launches, __ATOMIC_RELAXED); }

/* Error word of the device code (bit 0: a traversal stack ran out -- the hierarchy is deeper than rt_set_scene's
 * check allows, pixels of that frame are not to be trusted).  Synchronises; the word is cleared by the read. */
int rt_get_device_errors(rt_context* ctx, unsigned int* out) {
    if (!ctx || !out) return RT_ERR_ARG;
    RT_CHECK(ctx, cudaSetDevice(ctx->device));
    RT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    RT_CHECK(ctx, cudaMemcpy(out, ctx->d_tile_counter + RT_DEVICE_ERROR_WORD, sizeof(unsigned int), cudaMemcpyDeviceToHost));
    if (*out) RT_CHECK(ctx, cudaMemset(ctx->d_tile_counter + RT_DEVICE_ERROR_WORD, 0, sizeof(unsigned int)));
    return RT_OK;
}

int rt_get_timings(rt_context* ctx, float out_ms[3]) {
    if (!ctx || !out_ms) return RT_ERR_ARG;
    out_ms[0] = out_ms[1] = out_ms[2] = 0.0f;
    RT_CHECK(ctx, cudaSetDevice(ctx->device));
    RT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    if (ctx->timing_valid) {
        cudaEventElapsedTime(&out_ms[0], ctx->ev[0], ctx->ev[1]);
        cudaEventElapsedTime(&out_ms[1], ctx->ev[2], ctx->ev[3]);
    }
    if (ctx->build_timed) cudaEventElapsedTime(&out_ms[2], ctx->ev[4], ctx->ev[5]);
    return RT_OK;
}

int rt_bvh_info(rt_context* ctx, int* n_prims_in_bvh, int* n_nodes, int* n_planes) {
    if (!ctx) return RT_ERR_ARG;
    if (n_prims_in_bvh) *n_prims_in_bvh = ctx->scene.n_bvh_prims;
    if (n_nodes) *n_nodes = ctx->scene.n_bvh_prims > 1 ? ctx->scene.n_bvh_prims - 1 : 0;
    if (n_planes) *n_planes = ctx->scene.n_planes;
    return RT_OK;
}

int rt_bvh_download(rt_context* ctx, float* nodes16, int* sorted_prim_ids, unsigned int* morton_sorted) {
    if (!ctx || !ctx->have_scene) return RT_ERR_ARG;
    RT_CHECK(ctx, cudaSetDevice(ctx->device));
    RT_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    const int n = ctx->scene.n_bvh_prims;
    if (nodes16 && n > 1) RT_CHECK(ctx, cudaMemcpy(nodes16, ctx->bb.nodes, sizeof(float) * 16 * (size_t)(n - 1), cudaMemcpyDeviceToHost));
    if (sorted_prim_ids && n > 0) {
        std::vector<float4> prims((size_t)n * 3);
        RT_CHECK(ctx, cudaMemcpy(prims.data(), ctx->bb.prims, sizeof(float4) * 3 * (size_t)n, cudaMemcpyDeviceToHost));
        for (int i = 0; i < n; i++) {
            unsigned int tag;
            memcpy(&tag, &prims[3 * (size_t)i].w, 4);
            sorted_prim_ids[i] = (int)(tag & RT_TAG_INDEX);
        }
    }
    if (morton_sorted && n > 0) RT_CHECK(ctx, cudaMemcpy(morton_sorted, ctx->bb.keys, sizeof(unsigned int) * (size_t)n, cudaMemcpyDeviceToHost));
    return RT_OK;
}

int rt_sort_pairs(rt_context* ctx, unsigned int* keys_dev, unsigned int* vals_dev, int n) {
    if (!ctx || n < 0) return RT_ERR_ARG;
    if (n <= 1) return RT_OK;
    RT_CHECK(ctx, cudaSetDevice(ctx->device));
    unsigned int *tk = nullptr, *tv = nullptr, *hist = nullptr;
    cudaError_t e = cudaMalloc((void**)&tk, sizeof(unsigned int) * (size_t)n);
    if (e == cudaSuccess) e = cudaMalloc((void**)&tv, sizeof(unsigned int) * (size_t)n);
    if (e == cudaSuccess) e = cudaMalloc((void**)&hist, sizeof(unsigned int) * (size_t)rt_radix_hist_entries(n));
    if (e == cudaSuccess) e = rt_radix_sort_pairs(keys_dev, vals_dev, tk, tv, hist, n, 1, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(tk); cudaFree(tv); cudaFree(hist);
    RT_CHECK(ctx, e);
    return RT_OK;
}

/* rt_sort_pairs without any allocation: scratch in a file-scope device array (bvh_build.cu).  No cudaMalloc, no
 * cudaFree, no synchronisation: asynchronous on the context's stream. */
int rt_sort_pairs_static(rt_context* ctx, unsigned int* keys_dev, unsigned int* vals_dev, int n) {
    if (!ctx || n < 0 || (n > 0 && (!keys_dev || !vals_dev))) return RT_ERR_ARG;
    if (n > RT_SORT_STATIC_MAX) {
        char msg[160];
        snprintf(msg, sizeof(msg), "rt_sort_pairs_static: n = %d exceeds the static scratch capacity of %d keys "
                 "(use rt_sort_pairs)", n, RT_SORT_STATIC_MAX);
        return fail(ctx, RT_ERR_ARG, msg);
    }
    if (n <= 1) return RT_OK;
    RT_CHECK(ctx, cudaSetDevice(ctx->device));
    const char* why = "";
    const cudaError_t e = rt_radix_sort_pairs_static(keys_dev, vals_dev, n, ctx->device, ctx->stream, &why);
    if (e != cudaSuccess) {
        ctx->err = std::string("rt_sort_pairs_static: ") + why + ": " + cudaGetErrorString(e);
        return RT_ERR_CUDA;
    }
    return RT_OK;
}

/* ------------------------------------------------------------------ legacy drop-in */

static rt_context* g_legacy = nullptr;
static bool g_legacy_sky = false;

static void legacy_die(const char* what, rt_context* ctx, int rc) {
    /* the reference's getLastCudaError: print and exit (common/inc/helper_cuda.h:584-596) */
    fprintf(stderr, "launchKernel: %s failed (%d): %s\n", what, rc, ctx ? rt_last_error(ctx) : "no context");
    exit(EXIT_FAILURE);
}

void launchKernel(unsigned int* result, int imgw, int imgh, rtb_Camera cam, rtb_Object* objects, rtb_Light* lights,
                  rtb_float3 ambient, float* h_skyVars, unsigned char* h_tex1, unsigned char* h_tex2,
                  unsigned char* h_tex3, unsigned char* h_tex4, int w, int h, float h_dayTime, bool alias) {
    int rc;
    if (!g_legacy) {
        int dev = 0;
        cudaGetDevice(&dev);
        rc = rt_create(dev, &g_legacy);
        if (rc != RT_OK) legacy_die("rt_create", g_legacy, rc);
    }
    if (!g_legacy_sky) {   /* textures are read on the first call only (reference kernel.cu:414-442) */
        const unsigned char* tex[4] = {h_tex1, h_tex2, h_tex3, h_tex4};
        rc = rt_set_sky(g_legacy, tex, w, h);
        if (rc != RT_OK) legacy_die("rt_set_sky", g_legacy, rc);
        g_legacy_sky = true;
    }
    rc = rt_set_scene(g_legacy, objects, RTB_OBJECTS_NUMBER, lights, RTB_LIGHTS_NUMBER);
    if (rc != RT_OK) legacy_die("rt_set_scene", g_legacy, rc);
    rt_frame fr;
    memset(&fr, 0, sizeof(fr));
    fr.cam = cam;
    fr.ambient[0] = ambient.x; fr.ambient[1] = ambient.y; fr.ambient[2] = ambient.z;
    memcpy(fr.skyVars, h_skyVars, 16);
    fr.dayTime = h_dayTime;
    fr.alias = alias ? 1 : 0;
    fr.imgw = imgw;
    fr.imgh = imgh;
    rc = rt_render(g_legacy, &fr, nullptr, result, nullptr);
    if (rc != RT_OK) legacy_die("rt_render", g_legacy, rc);
    /* asynchronous like the reference; the caller synchronises (reference scene.cpp:825) */
}

/* the context behind launchKernel (tests: options, counters) */
rt_context* rt_legacy_context(void) { return g_legacy; }

/* The reference never frees what launchKernel allocated (kernel.cu:419-441).  A host that wants the memory back, or
 * other sky maps, calls this: the next launchKernel starts from scratch and reads its texture arguments again. */
int rtb_legacy_shutdown(void) {
    int rc = RT_OK;
    if (g_legacy) rc = rt_destroy(g_legacy);
    g_legacy = nullptr;
    g_legacy_sky = false;
    return rc;
}

void rtb_scene_launch(rtb_scene* s, unsigned int* out_dev, int imgw, int imgh, unsigned char* tex[4], int texw,
                      int texh) {
    rtb_scene_update_camera(s, imgw, imgh);
    rt_frame fr;
    rtb_scene_frame(s, imgw, imgh, &fr);
    int n = 0, nl = 0;
    rtb_Object* objs = rtb_scene_objects(s, &n);
    rtb_Light* lights = rtb_scene_lights(s, &nl);
    rtb_float3 amb;
    amb.x = fr.ambient[0]; amb.y = fr.ambient[1]; amb.z = fr.ambient[2];
    launchKernel(out_dev, imgw, imgh, fr.cam, objs, lights, amb, fr.skyVars, tex[0], tex[1], tex[2], tex[3], texw, texh,
                 fr.dayTime, fr.alias != 0);
    cudaDeviceSynchronize();
}

}  // extern "C"
