/*
 * rtb_api.h -- C-ABI of the B200-native ray-tracing hot path.
 *
 * Two layers are exported by libraytracing_b200.so:
 *
 *  (1) launchKernel(...)  -- the legacy drop-in.  Exactly the symbol the
 *      reference's scene.cpp declares (reference scene.cpp:79-85) and calls
 *      once per frame (reference scene.cpp:823-824); defined by the reference
 *      in kernel.cu:406-462.  `result` is a DEVICE pointer, everything else is
 *      host memory; sky textures are read on the first call only; errors print
 *      to stderr and exit(1) like the reference's getLastCudaError
 *      (reference common/inc/helper_cuda.h:584-596).
 *
 *  (2) rt_*  -- a handle API for what the legacy form cannot express: any number
 *      of primitives (LBVH in HBM instead of 133 objects in __constant__),
 *      explicit lifetime, several devices (one context per device), screen-tile
 *      shards, hit-id / counter read-back, status codes instead of exit().
 *      launchKernel is a thin shim over one process-global rt_context.
 *
 *  (3) rtb_scene_* -- the host-side scene layer (reference scene.cpp) restated
 *      with explicit state instead of globals; produces the Object/Light/Camera
 *      inputs of the boundary for any view, time of day and sea level.
 *
 * No torch types, no C++ types: plain pointers, ints and the PODs of
 * rtb_structs.h.
 */
#ifndef RTB_API_H
#define RTB_API_H

#include "rtb_structs.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ legacy */

#if defined(__cplusplus) && !defined(RTB_NO_LEGACY_DECL)
/* replaces reference kernel.cu:406-411 */
void launchKernel(unsigned int* result, int imgw, int imgh, rtb_Camera cam,
                  rtb_Object* objects, rtb_Light* lights, rtb_float3 ambient,
                  float* h_skyVars,
                  unsigned char* h_tex1, unsigned char* h_tex2,
                  unsigned char* h_tex3, unsigned char* h_tex4, int w, int h,
                  float h_dayTime, bool alias);
#endif

/* The reference never frees what launchKernel allocates on its first call (kernel.cu:419-441).  Extension: destroy
 * the process-global context behind launchKernel; the next call starts from scratch and reads its sky arguments
 * again. */
int  rtb_legacy_shutdown(void);

/* ------------------------------------------------------------ handle API */

typedef struct rt_context rt_context;

enum {
    RT_OK = 0,
    RT_ERR_CUDA = 1,       /* a CUDA runtime/driver call failed; see rt_last_error */
    RT_ERR_ARG = 2,        /* bad argument */
    RT_ERR_STATE = 3,      /* call order (e.g. render before set_scene/set_sky) */
    RT_ERR_UNSUPPORTED = 4
};

/* rt_set_option keys */
enum {
    RT_OPT_TRAVERSAL = 0,   /* 0 = LBVH (default), 1 = brute force over all primitives (GPU cross-check) */
    RT_OPT_COUNTERS  = 1,   /* 1 = instrumented kernels: count rays / node / primitive tests */
    RT_OPT_FXAA_IMPL = 2,   /* 0 = TMA-staged (default when width%4==0), 1 = plain shared-memory */
    RT_OPT_TIMING    = 3,   /* 1 = record CUDA events around trace and fxaa */
    RT_OPT_SKY_ALL4  = 4,   /* 1 = fetch all four sky maps like the reference (default 0: only non-zero weights; bit-identical) */
    RT_OPT_TRACE_IMPL = 5,  /* 0 = auto: shared-memory tracer when the scene fits (<= ~165 primitives), else the plain one-thread-per-pixel tracer over the HBM-resident LBVH; 1 = the plain tracer for any scene (no lists, no packets: the cross-check of the others); 2 = shared-memory tracer with its deep-hierarchy fallback forced (tests); 3 = the persistent tracer of big scenes (trace_big.cu: warps pull 8x4 patches, primary rays as one packet per RT_OPT_PACKET, stack per RT_OPT_BIG_STACK) -- measured 8..21 % slower than the plain kernel on the million-triangle terrain, hence opt-in; 4 = the plain tracer with packet traversal of the primary rays (13 % slower).  Identical pixels */
    RT_OPT_BUILDER = 6,     /* hierarchy over the Morton-sorted primitives: 0 = auto (PLOC up to 1024 primitives, Karras above), 1 = Karras, 2 = PLOC */
    RT_OPT_FORCE_RADIX = 7, /* 1 = use the multi-pass radix sort even for tiny scenes (tests) */
    RT_OPT_ORDER = 8,       /* work order of the shared-memory tracer on whole frames: 0 = most expensive patches of the previous frame of this size first (default; first frame bottom-up), 1 = always bottom-up.  Scheduling only: identical pixels */
    RT_OPT_SHADOW_BINS = 9, /* shadow rays: 0 = per-light direction bins (default; small scenes: built by the pre-pass of the first frame after a change of scene or lights, see RT_OPT_LIGHT_BINS_REUSE; big scenes: a CSR grid built once a second frame finds scene and lights unchanged), 1 = traverse the hierarchy, 2 = like 0, big scenes build the grid at once.  Which primitives get tested only: identical pixels */
    RT_OPT_MIRROR_BINS = 10, /* rays that primary rays send off plane 0 (the sea) in the shared-memory tracer: 0 = candidate lists of the bin's mirrored frustum (default), 1 = traverse the hierarchy.  Which primitives get tested only: identical pixels */
    RT_OPT_LIGHT_BINS_REUSE = 11, /* light bins of the shared-memory tracer: 0 = rebuilt only when the scene or the lights changed since they were built (default; they do not depend on the camera), 1 = rebuilt by every frame's pre-pass.  Identical pixels */
    RT_OPT_COMPACTION = 12, /* ray compaction in the shared-memory tracer: 0 = every lane follows its own pixel's chain to the end (default), 1 = rays of the second and later generations that have to traverse the hierarchy are parked in a per-CTA shared-memory queue (ballot + popc slot assignment) and traced 32 at a time by full warps.  Measured SLOWER on every BASELINE view (full warps, but rays of unrelated patches: the traversal loses the coherence an 8x4 patch has), hence opt-in.  Scheduling only: identical pixels */
    RT_OPT_PACKET = 13, /* persistent tracer of big scenes: 0 = the 32 primary rays of an 8x4 patch walk the hierarchy as ONE packet (warp-uniform node, one stack per warp, broadcast node loads; default), 1 = every lane traverses for itself.  Identical pixels */
    RT_OPT_BIG_STACK = 14, /* persistent tracer of big scenes, per-lane traversal stack: 0 = local memory (default), 1 = first 16 entries in shared memory.  Identical pixels */
    RT_OPT_WIDE_BVH = 15,  /* plain tracer of big scenes: 0 = walks the binary LBVH (default), 1 = walks a 4-wide view of it (one 128-byte node per visit holding the boxes of the grandchildren; rebuilt after every rt_set_scene, 128 bytes per primitive) -- measured 22 % slower on the million-triangle terrain, hence opt-in.  Identical pixels */
    RT_OPT_N = 16
};

/* One frame's uniforms: what launchKernel receives by value / as small host arrays. */
typedef struct rt_frame {
    rtb_Camera cam;
    float      ambient[3];
    float      skyVars[4];
    float      dayTime;      /* dayNightTime / 24 (reference scene.cpp:822) */
    int        alias;        /* FXAA on/off */
    int        imgw, imgh;
} rt_frame;

/* Interleaved screen-tile shard: tile (i,j) belongs to shard (i + j*stride) % n_shards.
 * A shard traces its tiles plus a 1-pixel apron so FXAA needs no halo exchange. */
typedef struct rt_shard {
    int tile_w, tile_h;
    int n_shards, shard_id;
    int stride;
} rt_shard;

/* counters returned by rt_get_counters (RT_OPT_COUNTERS=1) */
enum {
    RT_CNT_RAYS_PRIMARY = 0, RT_CNT_RAYS_SHADOW = 1, RT_CNT_RAYS_MIRROR = 2,
    RT_CNT_NODE_TESTS = 3, RT_CNT_TRI_TESTS = 4, RT_CNT_SPHERE_TESTS = 5,
    RT_CNT_PLANE_TESTS = 6, RT_CNT_SKY_LOOKUPS = 7, RT_CNT_SHADOW_SKIPPED = 8,
    /* triangle tests by exit (reference kernel.cu:100-123): u outside / v outside / through to t (hit or t < 0) /
     * hits; the remaining RT_CNT_TRI_TESTS were culled by det < 0.001.  Sphere tests that hit (kernel.cu:49-66). */
    RT_CNT_TRI_EXIT_U = 9, RT_CNT_TRI_EXIT_V = 10, RT_CNT_TRI_FULL = 11, RT_CNT_TRI_ACCEPT = 12,
    RT_CNT_SPHERE_HITS = 13,
    RT_CNT_COUNT = 16
};

int  rt_create(int device, rt_context** out);
int  rt_destroy(rt_context* ctx);
const char* rt_last_error(rt_context* ctx);
int  rt_set_option(rt_context* ctx, int key, int value);
/* cudaStream_t to launch on (default: the legacy default stream) */
int  rt_set_stream(rt_context* ctx, void* cuda_stream);

/* Four equirectangular RGBA8 sky maps, HOST memory, w*h*4 bytes each
 * (reference kernel.cu:418-436).  Copied to device once. */
int  rt_set_sky(rt_context* ctx, const unsigned char* const tex[4], int w, int h);
/* Let `ctx` sample the sky maps `owner` uploaded (same device; 512 MiB is uploaded once, `owner` must
 * outlive `ctx`). */
int  rt_share_sky(rt_context* ctx, rt_context* owner);

/* Upload primitives + lights (HOST memory) and build / refit the LBVH.
 * Classifies the change against the previous call: identical geometry ->
 * materials only; same topology with moved primitives -> refit; else rebuild. */
int  rt_set_scene(rt_context* ctx, const rtb_Object* objects, int n_objects,
                  const rtb_Light* lights, int n_lights);

/* Render one frame into a DEVICE buffer of imgw*imgh uint32 (0x00BBGGRR).
 * shard may be NULL (whole frame).  hit_ids_dev may be NULL; else imgw*imgh int32
 * primary hit object index (-1 = sky).  Asynchronous on the context's stream. */
int  rt_render(rt_context* ctx, const rt_frame* frame, const rt_shard* shard,
               unsigned int* out_dev, int* hit_ids_dev);
/* Same, result copied to HOST memory (pinned or pageable); synchronous. */
int  rt_render_to_host(rt_context* ctx, const rt_frame* frame, const rt_shard* shard,
                       unsigned int* out_host, int* hit_ids_host);
/* Asynchronous read-back (replaces the display hand-off of reference main.cpp:141-165 for a headless consumer):
 * device -> host copies run on a stream of the context's own, so the copy of one frame hides behind the rendering
 * of the next ones.  slot in [0, RT_MAX_SLOTS) names one frame in flight; out_host should be page-locked
 * (rt_host_alloc) -- a pageable target turns the copy synchronous.
 *   rt_render_to_host_async  render into the slot's own device frame, then copy it to out_host
 *   rt_readback_async        copy a device frame the caller rendered itself (rt_render, launchKernel on this
 *                            context's stream) once everything enqueued so far has run; the caller keeps src_dev
 *                            untouched until rt_wait_slot
 *   rt_wait_slot             returns when the slot's last copy has landed in host memory */
#define RT_MAX_SLOTS 8
int  rt_render_to_host_async(rt_context* ctx, const rt_frame* frame, const rt_shard* shard, unsigned int* out_host,
                             int slot);
int  rt_readback_async(rt_context* ctx, const unsigned int* src_dev, size_t n_pixels, unsigned int* dst_host, int slot);
int  rt_wait_slot(rt_context* ctx, int slot);
int  rt_host_alloc(void** out, size_t bytes);
int  rt_host_free(void* p);
/* the process-global context behind launchKernel (NULL before its first call): options, counters, read-back slots */
rt_context* rt_legacy_context(void);

/* The pre-FXAA image of the last rt_render (device pointer owned by ctx; for a whole frame with alias = 0 the
 * tracer writes the caller's out_dev directly and that pointer is returned). */
int  rt_get_base_image(rt_context* ctx, unsigned int** base_dev);
/* FXAA alone: in_dev -> out_dev (reference kernel.cu:262-403 with alias=true). */
int  rt_fxaa(rt_context* ctx, const unsigned int* in_dev, unsigned int* out_dev,
             int imgw, int imgh, const rt_shard* shard);

/* The reference's `factor` render-scale knob (main.cpp:40-47): in_dev holds a frame of (out_w*factor) x
 * (out_h*factor) pixels; mode 0 = the GL_NEAREST blit of the reference (window pixel (x,y) shows texel
 * (x f + f/2, y f + f/2), main.cpp:317-318), mode 1 = f x f box average (supersampling). factor 1..16. */
int  rt_present(rt_context* ctx, const unsigned int* in_dev, int out_w, int out_h, int factor, int mode,
                unsigned int* out_dev);

/* Tile gather helpers for multi-GPU compositing: pack the tiles a shard owns
 * into a dense [n_tiles][tile_h][tile_w] buffer / scatter them back. */
int  rt_shard_tile_count(const rt_shard* shard, int imgw, int imgh);
int  rt_pack_tiles(rt_context* ctx, const unsigned int* frame_dev, int imgw, int imgh,
                   const rt_shard* shard, unsigned int* packed_dev);
int  rt_unpack_tiles(rt_context* ctx, const unsigned int* packed_dev, int imgw, int imgh,
                     const rt_shard* shard, unsigned int* frame_dev);

/* Peer frame: ONE frame over N GPUs (one process each) without a gather.  The owner allocates n_buffers frames
 * of n_pixels and exports a 64-byte handle (cudaIpc); the other processes open it and pass
 * rt_peer_frame_buffer(pf, seq) as `out_dev` of their sharded rt_render, so the FXAA / copy kernel stores the
 * rank's tiles straight into the owner's memory over NVLink ("peer-mapped stores straight into
 * GPU 0's framebuffer").  Hand-shake, all asynchronous on the contexts' streams, frame numbers seq = 1, 2, ...:
 *   every rank:  rt_peer_frame_acquire(seq)  -- buffer seq % n_buffers is free (frame seq - n_buffers released)
 *                rt_render(..., shard, rt_peer_frame_buffer(pf, seq), ...)
 *                rt_peer_frame_signal(rank, seq)
 *   the owner:   rt_peer_frame_wait(n_ranks, seq) ... use the frame ... rt_peer_frame_release(seq)
 * rt_render_peer(...) = acquire + rt_render into rt_peer_frame_buffer(pf, seq) + signal in ONE call: the frame's last
 * kernel (FXAA / copy) carries both halves of the hand-shake, so a producer frame is three kernel launches.
 * A wait that is not satisfied within the frame's timeout (default 10 s, rt_peer_frame_set_timeout; per process)
 * sets the error word instead of hanging: 1 + k = the owner gave up on rank k, 0x100 + k = rank k gave up on the
 * owner's release.  The frame is then INVALID (a producer that gave up stores nothing).  rt_peer_frame_error
 * synchronises and reads the word -- check it before using a frame; it stays set until rt_peer_frame_reset_error. */
#define RT_PEER_HANDLE_BYTES 64
typedef struct rt_peer_frame rt_peer_frame;
int  rt_peer_frame_create(rt_context* ctx, size_t n_pixels, int n_buffers, rt_peer_frame** out,
                          unsigned char* handle /* RT_PEER_HANDLE_BYTES */);
int  rt_peer_frame_open(rt_context* ctx, const unsigned char* handle, size_t n_pixels, int n_buffers,
                        rt_peer_frame** out);
unsigned int* rt_peer_frame_buffer(rt_peer_frame* pf, unsigned int seq);
int  rt_peer_frame_acquire(rt_context* ctx, rt_peer_frame* pf, unsigned int seq);
int  rt_peer_frame_signal(rt_context* ctx, rt_peer_frame* pf, int rank, unsigned int seq);
int  rt_peer_frame_wait(rt_context* ctx, rt_peer_frame* pf, int n_ranks, unsigned int seq);
int  rt_peer_frame_release(rt_context* ctx, rt_peer_frame* pf, unsigned int seq);
int  rt_render_peer(rt_context* ctx, const rt_frame* frame, const rt_shard* shard, rt_peer_frame* pf, int rank,
                    unsigned int seq, int* hit_ids_dev);
int  rt_peer_frame_set_timeout(rt_peer_frame* pf, unsigned int timeout_ms);
int  rt_peer_frame_error(rt_context* ctx, rt_peer_frame* pf, unsigned int* err);
int  rt_peer_frame_reset_error(rt_context* ctx, rt_peer_frame* pf);
int  rt_peer_frame_close(rt_context* ctx, rt_peer_frame* pf);

int  rt_synchronize(rt_context* ctx);
int  rt_get_counters(rt_context* ctx, unsigned long long out[RT_CNT_COUNT]);
/* diagnostics: copy the light bins the last frame of the shared-memory tracer used (4 lights x 6 faces x 64 x 64
 * bins of 8 ushorts: [0] count or 0xFFFF = too many, [1..7] leaf offsets); returns the bytes copied, 0 if none,
 * -RT_ERR_* on an error */
long long rt_debug_light_bins(rt_context* ctx, void* out, size_t bytes);
/* Error word of the device code: bit 0 = a traversal stack ran out (hierarchy deeper than rt_set_scene's check lets
 * through; the frame is not to be trusted).  Synchronises; reading clears the word. */
int  rt_get_device_errors(rt_context* ctx, unsigned int* out);
/* kernels this library has launched in this process so far (all contexts): callers difference it around a region */
unsigned long long rt_launch_count(void);
/* ms of the last frame's kernels (RT_OPT_TIMING=1): [0]=trace [1]=fxaa [2]=bvh build/refit */
int  rt_get_timings(rt_context* ctx, float out_ms[3]);
/* FP32 FFMA throughput of the device in TFLOP/s (2 flops per FFMA), measured with a register-resident
 * FMA chain kernel: the denominator of the tracer's FP32 roofline fraction. */
int  rt_measure_fp32_peak(rt_context* ctx, float* tflops);

/* LBVH introspection (tests): copies to HOST.  nodes: 16 floats per node. */
int  rt_bvh_info(rt_context* ctx, int* n_prims_in_bvh, int* n_nodes, int* n_planes);
int  rt_bvh_download(rt_context* ctx, float* nodes16, int* sorted_prim_ids,
                     unsigned int* morton_sorted);
/* Stand-alone key sort used by the builder (tests compare it with a library sort):
 * sorts (key,val) pairs in DEVICE memory, stable, ascending. */
int  rt_sort_pairs(rt_context* ctx, unsigned int* keys_dev, unsigned int* vals_dev, int n);
/* No counterpart in the reference.  The same stable ascending sort of
 * (key, val) pairs in DEVICE memory for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it): it calls no cudaMalloc, cudaFree or allocating library routine -- the
 * scratch rt_sort_pairs allocates per call (two temporary arrays + digit histograms) lives in a file-scope device
 * array of RT_SORT_STATIC_MAX keys' capacity (8.5 MiB per GPU that loads the library, held until the process exits).
 * n > RT_SORT_STATIC_MAX is rejected with RT_ERR_ARG and a message in rt_last_error.  Asynchronous on the context's
 * stream (rt_sort_pairs synchronises before it frees; this one does not).  Every call on a GPU shares the one array,
 * which keeps the last call's contents (nothing in it has to start at zero: every pass overwrites what it reads):
 * calls that OVERLAP on the GPU -- other streams, contexts or threads -- race on it; serialise them. */
#define RT_SORT_STATIC_MAX (1 << 20)
int  rt_sort_pairs_static(rt_context* ctx, unsigned int* keys_dev, unsigned int* vals_dev, int n);

/* ----------------------------------------------------- host scene layer */

typedef struct rtb_scene rtb_scene;

/* initCamera + initLights + initObjects (reference scene.cpp:165-173, 634-652, 444-488) */
rtb_scene*  rtb_scene_create(void);
void        rtb_scene_destroy(rtb_scene* s);
rtb_Object* rtb_scene_objects(rtb_scene* s, int* n);
rtb_Light*  rtb_scene_lights(rtb_scene* s, int* n);
rtb_Camera* rtb_scene_camera(rtb_scene* s);
float*      rtb_scene_sky_vars(rtb_scene* s);      /* 4 floats */
float*      rtb_scene_ambient(rtb_scene* s);       /* 3 floats */
float       rtb_scene_time(rtb_scene* s);          /* dayNightTime, hours */
/* Steady state of animate() at a fixed hour: calcSkyVars -> recolorObjects ->
 * moveLights (reference scene.cpp:778-804, 674-687, 758-776; order per SURVEY 3.2). */
void        rtb_scene_set_time(rtb_scene* s, float hours);
/* key '5' (island) / '6' (mountains) presets (reference scene.cpp:736-747) */
int         rtb_scene_set_view(rtb_scene* s, int key);
void        rtb_scene_set_pose(rtb_scene* s, float px, float py, float pz, float horAngle, float verAngle);
/* aspect = (1.0f*w)/h then cameraHelperAngles (reference scene.cpp:821, 100-126) */
void        rtb_scene_update_camera(rtb_scene* s, int imgw, int imgh);
void        rtb_scene_set_sea_level(rtb_scene* s, float y);   /* objects[0].pos.y, reference scene.cpp:709 */
void        rtb_scene_set_antialiasing(rtb_scene* s, int on);
/* Headless stand-in for one animate() tick (reference scene.cpp:806-816) with the
 * key state given explicitly.  keys: bitmask of RTB_KEY_* */
void        rtb_scene_animate(rtb_scene* s, float deltaTime, unsigned int keys);
/* reference mouseMotion (scene.cpp:128-140) with explicit deltas */
void        rtb_scene_mouse(rtb_scene* s, int deltaX, int deltaY);
const char* rtb_scene_clock(rtb_scene* s);                    /* getTime(), reference scene.cpp:93-96 */
/* Fill an rt_frame from the scene state (what launch() passes, reference scene.cpp:820-824) */
void        rtb_scene_frame(rtb_scene* s, int imgw, int imgh, rt_frame* out);
/* launch(): update aspect, call the legacy boundary, synchronise (reference scene.cpp:820-826).
 * tex pointers as loaded by the caller. */
void        rtb_scene_launch(rtb_scene* s, unsigned int* out_dev, int imgw, int imgh,
                             unsigned char* tex[4], int texw, int texh);

enum {
    RTB_KEY_W = 1u << 0, RTB_KEY_S = 1u << 1, RTB_KEY_A = 1u << 2, RTB_KEY_D = 1u << 3,
    RTB_KEY_Q = 1u << 4, RTB_KEY_E = 1u << 5, RTB_KEY_SHIFT = 1u << 6,
    RTB_KEY_LEFT = 1u << 7, RTB_KEY_RIGHT = 1u << 8, RTB_KEY_UP = 1u << 9, RTB_KEY_DOWN = 1u << 10,
    RTB_KEY_1 = 1u << 11, RTB_KEY_2 = 1u << 12, RTB_KEY_3 = 1u << 13, RTB_KEY_4 = 1u << 14,
    RTB_KEY_5 = 1u << 15, RTB_KEY_6 = 1u << 16, RTB_KEY_P = 1u << 17, RTB_KEY_O = 1u << 18,
    RTB_KEY_B = 1u << 19, RTB_KEY_V = 1u << 20
};

/* Synthetic inputs (no reference counterpart; BASELINE.json config 4 and the
 * asset-free sky used on the GPU box). */
/* Tessellated terrain: cells x cells grid, 2 triangles per cell, fBm heights, plus the
 * water plane at index 0 and sun/moon spheres as the last two.  Returns object count
 * (2*cells*cells + 3); out may be NULL to query. */
int  rtb_terrain_build(int cells, unsigned int seed, rtb_Object* out, int capacity);
/* Scene ingest (csrc/ingest.cpp).  Heightmap: nx x nz samples (row-major, z outer) over [x0,x1] x [z0,z1] ->
 * 2 (nx-1)(nz-1) up-facing triangles, cells split like the reference's island top (scene.cpp:369-373).
 * Wavefront OBJ text ("v", "f" with v, v/vt, v//vn, v/vt/vn and negative indices; polygons fan-triangulated).
 * Material fields (shine, specular, mirror, color, light) are copied from *material (NULL: the mountain
 * material, scene.cpp:338-342).  out = NULL queries the count; -1 = capacity too small, -2 = malformed face. */
int  rtb_heightmap_objects(const float* heights, int nx, int nz, float x0, float x1, float z0, float z1,
                           const rtb_Object* material, rtb_Object* out, int capacity);
int  rtb_obj_parse(const char* text, size_t len, const rtb_Object* material, rtb_Object* out, int capacity);
/* Procedural equirectangular sky map `which` (0 morning,1 day,2 evening,3 night), RGBA8. */
void rtb_sky_synthetic(int which, int w, int h, unsigned char* out);

/* Image output (replaces the GL blit, reference main.cpp:141-165): packed 0x00BBGGRR
 * host pixels -> binary PPM / PNG.  Return 0 on success. */
int  rtb_write_ppm(const char* path, const unsigned int* pixels, int w, int h);
int  rtb_write_png(const char* path, const unsigned int* pixels, int w, int h);
/* frame `index` of a sequence dump: <prefix><index, six digits>.ppm (format 0) / .png (format 1) */
int  rtb_write_frame(const char* prefix, int index, int format, const unsigned int* pixels, int w, int h);
/* Image input (replaces stbi_load(path, &w, &h, &n, 4), reference scene.cpp:626-632): PNG file -> RGBA8 rows, top
 * row first.  out = NULL queries *w, *h; capacity = bytes at out (>= 4 w h).  8-bit grey / grey+alpha / RGB / RGBA /
 * palette, non-interlaced; anything else RT_ERR_UNSUPPORTED. */
int  rtb_read_png(const char* path, int* w, int* h, unsigned char* out, size_t capacity);

#ifdef __cplusplus
}
#endif
#endif /* RTB_API_H */
