/*
 * rt_device.h -- device-side data layout shared by the builder, the tracer and
 * the host wrapper.
 *
 * HBM layout (all 16-byte aligned, float4 granularity):
 *
 *   nodes   [n_nodes][4] float4   LBVH internal nodes, BOTH children's boxes embedded as centre c and
 *                                  half-extent e (the slab test is then 9 FMAs per box and no per-axis
 *                                  min/max: t_near/far = (c-o)*inv -/+ e*|inv|)
 *        words 0..5   child0: c.x c.y c.z e.x e.y e.z
 *        words 6..11  child1: c.x c.y c.z e.x e.y e.z
 *        words 12,13  child0, child1 as int bits; child >= 0: internal node,
 *                     child < 0: leaf, slot = ~child into `prims`;  words 14,15 unused
 *   prims   [n_bvh_prims][3] float4   leaf records in Morton order
 *        triangle: q0 = (v0.xyz, tag)  q1 = (e1.xyz, 0)  q2 = (e2.xyz, 0)     e1 = v1-v0, e2 = v2-v0
 *        sphere  : q0 = (c.xyz,  tag)  q1 = (r, r*r, 0, 0)
 *        tag = original object index | RT_TAG_SPHERE | RT_TAG_EMISSIVE
 *   planes  [n_planes][2] float4  (point.xyz, tag) (normal.xyz, 0) -- infinite planes are
 *                                  tested analytically, outside the BVH
 *   mats    [n_objects][2] float4 (color.rgb, mirror) (shine, specular, emissive, type)
 *   recs    [n_objects][3] float4 leaf records in ORIGINAL order (brute-force mode, build input)
 *
 * The reference keeps all of this as 133 x 68-byte AoS Objects in __constant__
 * memory (reference kernel.cu:14) and loops over them linearly per ray
 * (reference kernel.cu:144-151, 188-193).
 */
#ifndef RTB_RT_DEVICE_H
#define RTB_RT_DEVICE_H

#include <cuda_runtime.h>
#include "../../include/rtb_api.h"

#define RT_TAG_SPHERE   0x80000000u
#define RT_TAG_EMISSIVE 0x40000000u
#define RT_TAG_PLANE    0x20000000u
#define RT_TAG_INDEX    0x1FFFFFFFu

#define RT_MAX_LIGHTS 8
#define RT_NODE_SENTINEL 0x7FFFFFFF   /* empty-stack marker / empty tree */
#define RT_STACK_SIZE 64
/* shared-memory tracer (trace_small.cu): link offsets 64*(n-1) + 64*n must fit a signed 16-bit stack entry */
#define RT_SMALL_MAX_PRIMS 256
/* peer frame (peer.cu): flag block behind the frames, uints with 32-byte spacing */
#define RT_PEER_FLAG_UINTS 256
#define RT_PEER_FLAG_DONE(k) (8 * (k))     /* frame number rank k (< 16) has finished */
#define RT_PEER_FLAG_RELEASED 128          /* frame number the owner has released */
#define RT_PEER_FLAG_ERROR 136             /* set by a wait that gave up: 1 + k = the owner waiting for rank k,
                                            * 0x100 + k = rank k waiting for the owner's release */
#define RT_PEER_TIMEOUT_MS_DEFAULT 10000u

/* Hand-shake of a rank's sharded frame with the owner of a peer frame, carried by the frame's LAST kernel (FXAA or
 * copy, fxaa.cu) instead of one-thread kernels around it: before its first store into the peer buffer every CTA
 * waits until the owner has released the buffer's previous use (flags[RELEASED] >= acquire_seq), and the last CTA to
 * finish (ticket counter in this GPU's memory) publishes flags[DONE(rank)] = seq behind a system-scope fence.
 * flags == NULL: no hand-shake. */
struct RtPeerSync {
    unsigned int* flags;       /* flag block behind the owner's frames (peer mapping, or local on the owner) */
    unsigned int* ticket;      /* this context's CTA ticket counter (self re-arming) */
    unsigned int acquire_seq;  /* wait for released >= this */
    unsigned int seq;          /* publish this */
    int rank;
    unsigned int timeout_ms;
};
/* the context's 64-uint counter block (api.cu d_tile_counter): [0] next work unit, [1] finished warps, [32] listed
 * tiles of a shard (trace_small.cu), [40] device error word (rt_get_device_errors), [48] CTA ticket of the peer
 * hand-shake (fxaa.cu) */
#define RT_DEVICE_ERROR_WORD 40
#define RT_BYVALUE_MAX 304             /* objects passed in kernel parameters (68 B each) by the per-frame refit */
/* light bins of the small-scene tracer (trace_small.cu): per light a cube map of RT_LGRID_G x RT_LGRID_G direction
 * bins per face, each bin one uint4 = 8 ushorts: [0] count (0xFFFF: too many, traverse), [1..7] leaf offsets */
#define RT_LGRID_G 64
#define RT_LGRID_LIGHTS 4             /* lights with bins; further lights traverse the hierarchy */
#define RT_LGRID_BYTES ((size_t)RT_LGRID_LIGHTS * 6 * RT_LGRID_G * RT_LGRID_G * 16)
/* light grid of big scenes (light_grid.cu): cube map of RT_LGB_G x RT_LGB_G direction bins per face and light */
#ifndef RT_LGB_G
#define RT_LGB_G 2048
#endif
#define RT_LGB_LIGHTS 2               /* lights with a grid; further lights traverse the hierarchy */
#define RT_SMALL_ORDER_CLASSES 8      /* patch cost classes of the tracer's order history */
#define RT_SMALL_ORDER_HEAD 32        /* uints before the class lists of an order buffer: [0..7] class counts */
#ifndef RT_SMALL_DYN_SMEM_MAX
#define RT_SMALL_DYN_SMEM_MAX (24 * 1024)
#endif

/* kernels this library has launched in this process (rt_launch_count, api.cu): every launcher adds what it enqueued */
extern "C" unsigned long long g_rt_launches;
#define RT_LAUNCHED(n) ((void)__atomic_fetch_add(&g_rt_launches, (unsigned long long)(n), __ATOMIC_RELAXED))

struct RtSceneDev {
    const float4* nodes;
    const float4* prims;
    const float4* planes;
    const float4* mats;
    const float4* recs;
    const int* depth;  /* device: height of the hierarchy (edges root -> deepest leaf), written by the refit */
    int root;          /* encoded child: >=0 node, <0 leaf, RT_NODE_SENTINEL = empty */
    int n_bvh_prims;
    int n_planes;
    int n_objects;
    /* light grid of big scenes (light_grid.cu): per light with a grid, offsets[6 G G + 1] into entries (Morton slots
     * of the possible occluders per direction bin); NULL: that light's shadow rays traverse the hierarchy */
    const unsigned int* lg_offsets[RT_LGB_LIGHTS];
    const unsigned int* lg_entries[RT_LGB_LIGHTS];
    /* 4 uints per light, written by the build itself (rt_lgb_finish_kernel): [0] = 1 when the light's lists are
     * complete and may be used, else its shadow rays traverse; [1] entries, [2] overflow flag */
    const unsigned int* lg_state;
    /* 4-wide view of the hierarchy (rt_collapse_wide, bvh_build.cu), NULL: none.  Wide node i (RT_WIDE_F4 float4,
     * 128 bytes) belongs to binary node i and holds the boxes and links of its grandchildren (a child that is a
     * leaf stands for itself), so a ray makes one dependent fetch where the binary walk makes two. */
    const float4* wide;
    unsigned int* err;     /* device error word (RT_DEVICE_ERROR_WORD of the context's counter block) */
};
/* wide node: [0..5] four boxes, 6 floats each (centre xyz, half-extent xyz) packed back to back; [6] the four links as
 * ints (>= 0 wide node, < 0 ~leaf slot, RT_NODE_SENTINEL: no child); [7] unused (one node = one 128-byte line) */
#define RT_WIDE_F4 8

struct RtLightDev {
    float px, py, pz, intensity;
    float cr, cg, cb, _pad;
};

/* Per-frame uniforms, passed by value as a kernel parameter (the reference uploads the
 * same data with five cudaMemcpyToSymbol calls per frame, reference kernel.cu:447-451). */
struct RtFrameDev {
    float pos[3];
    float LD[3], LU[3], RD[3], RU[3];
    float ambient[3];
    float skyVars[4];
    float dayTime;
    int imgw, imgh;
    int n_lights;
    RtLightDev lights[RT_MAX_LIGHTS];
    cudaTextureObject_t sky[4];
    int sky_all4;
    /* shard: n_shards <= 1 means whole frame */
    int tile_w, tile_h, n_shards, shard_id, stride;
};

struct RtCountersDev {
    unsigned long long v[RT_CNT_COUNT];
};

/* scratch + outputs of the LBVH build (bvh_build.cu) */
struct RtBuildBuffers {
    const float* objs; int n_objects;
    const int* plane_idx; int n_planes;
    float4 *mats, *recs, *planes;
    int* bvh_src;
    float4 *prim_lo, *prim_hi;
    unsigned int* bounds;
    unsigned int *keys, *vals, *tmp_keys, *tmp_vals, *hist;
    int2* children; int *node_parent, *leaf_parent;
    int* node_other;   /* Karras: node i covers the sorted leaves between i and node_other[i]; -1: unknown (PLOC) */
    unsigned int* flags;
    float* nodes; float* node_aux; float4* prims;
    int* depth;
    int n_bvh;
    /* per-frame refit of a small scene: the objects on the host (NULL: already uploaded to `objs`), and whether the
     * previous operation on the stream is this context's own FXAA / copy kernel (bvh_build.cu, by-value refit) */
    const float* objs_host;
    int after_own_post;
    /* stamp of the current refit in the arrival slots of `node_aux` (rt_refit_kernel): a slot whose stamp is older
     * is empty, so the slots are cleared when they are allocated and when the 20-bit stamp wraps, not per refit */
    unsigned int refit_epoch;
};

/* owner of the tile containing pixel (x,y) */
__host__ __device__ inline int rt_tile_owner(int x, int y, int tile_w, int tile_h, int n_shards, int stride) {
    int i = x / tile_w, j = y / tile_h;
    return (i + j * stride) % n_shards;
}

#endif
