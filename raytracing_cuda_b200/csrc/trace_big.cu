/*
 * trace_big.cu -- persistent, warp-coherent tracer for scenes that do not fit shared memory (BASELINE config 4: a
 * million triangles).  The LBVH and the leaf records stay in HBM / the 126 MB L2 and are read with float4 loads
 * through the read-only path.  Same per-pixel arithmetic as trace.cu and trace_small.cu (trace_pixel.cuh), so
 * all three kernels agree bit for bit; only WHICH primitives get tested, and in what order, differs.
 *
 * What changes against the plain one-thread-per-pixel kernel (trace.cu):
 *   - persistent CTAs (4 per SM), every WARP autonomous: it pulls single 8x4-pixel patches from an atomic counter,
 *     the frame walked from its last patch to its first, no CTA barrier in the loop (as in trace_small.cu); for a
 *     screen-tile shard a patch none of whose pixels the shard needs costs four owner look-ups, not a launch slot;
 *   - PACKET traversal for the primary rays.  The 32 primary rays of a patch form a thin bundle and visit almost
 *     the same nodes, so the warp walks the hierarchy ONCE for all of them: one warp-uniform node index, one stack
 *     per warp in shared memory, every lane tests ITS ray against the node's two child boxes, a child is entered
 *     when ANY lane's ray meets it (ballot), the child most lanes reach first goes first.  All 32 lanes load the
 *     same node / leaf record at the same time -- one broadcast LDG.128 wavefront instead of up to 32 -- and there
 *     is no divergence and no per-lane stack traffic.  A lane may so see primitives its own traversal would have
 *     skipped; the exact leaf tests decide, and the (t, index) minimum over a superset of the primitives a ray can
 *     hit is the reference's minimum over all objects (kernel.cu:144-151);
 *   - the per-lane traversal (bounces; shadow rays of a light without a grid) keeps its stack in SHARED memory:
 *     RB_STACK_SMEM 32-bit entries per thread, deeper pushes continue in a local-memory array, and a push beyond
 *     both (a hierarchy deeper than the host-side check lets through, api.cu) raises the context's device error word
 *     (rt_get_device_errors) instead of dropping the subtree silently.
 */
#include "trace_global.cuh"

#define RB_THREADS 256
#define RB_WARPS (RB_THREADS / 32)
#ifndef RB_CTAS_PER_SM
#define RB_CTAS_PER_SM 4
#endif
#define RB_STACK_SMEM 16                      /* per-thread stack entries in shared memory */
#define RB_PACKET_STACK (RT_STACK_SIZE + 2)   /* per-warp stack of the packet traversal: one push per level at most */
#define RB_CNT_PATCH 0                        /* tile_counter[]: next patch, finished warps (api.cu: d_tile_counter) */
#define RB_CNT_DONE 1

__device__ __forceinline__ void rb_griddep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

/* LBVH traversal of ONE ray with the short stack in shared memory.  `stk` = this thread's column (entry e at
 * stk[32 e]), entries beyond RB_STACK_SMEM go to `deep`.  ANY: first non-emissive hit with t < tmax
 * (kernel.cu:188-193); else closest hit into `best`. */
template <bool ANY, bool COUNT, bool ROBUST>
__device__ __forceinline__ bool rb_traverse_loop(const RtSceneDev& sc, const RtRayBox& rb, const vec3& o, const vec3& d,
                                                 RtHit& best, float tmax, RtLocalCounters& cnt, int* __restrict__ stk,
                                                 unsigned int* __restrict__ err) {
    int node = sc.root;
    int deep[RT_STACK_SIZE];
    int sp = 0;
    const float4* __restrict__ nodes = sc.nodes;
    while (true) {
        if (node >= 0) {
            const float4 n0 = __ldg(nodes + 4 * node);
            const float4 n1 = __ldg(nodes + 4 * node + 1);
            const float4 n2 = __ldg(nodes + 4 * node + 2);
            const float4 n3 = __ldg(nodes + 4 * node + 3);
            const float lim = ANY ? tmax : fmaxf(best.t, 0.0f);
            float t0, t1;
            const bool h0 = ROBUST ? rt_slab_robust(rb, n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, lim, t0)
                                   : rt_slab(rb, n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, lim, t0);
            const bool h1 = ROBUST ? rt_slab_robust(rb, n1.z, n1.w, n2.x, n2.y, n2.z, n2.w, lim, t1)
                                   : rt_slab(rb, n1.z, n1.w, n2.x, n2.y, n2.z, n2.w, lim, t1);
            if (COUNT) cnt.node_tests += 2;
            int c0 = __float_as_int(n3.x), c1 = __float_as_int(n3.y);
            if (h0 && h1) {
                if (t1 < t0) { int tmp = c0; c0 = c1; c1 = tmp; }
                if (sp < RB_STACK_SMEM) {
                    stk[32 * sp] = c1;
                    sp++;
                } else if (sp < RB_STACK_SMEM + RT_STACK_SIZE) {
                    deep[sp - RB_STACK_SMEM] = c1;
                    sp++;
                } else {
                    atomicOr(err, 1u);           /* deeper than any hierarchy the host lets through: say so */
                }
                node = c0;
                continue;
            }
            if (h0) { node = c0; continue; }
            if (h1) { node = c1; continue; }
        } else {
            const int slot = ~node;
            float t;
            unsigned int tag;
            if (rt_test_record<COUNT>(sc.prims + 3 * (size_t)slot, o, d, t, tag, cnt)) {
                if (ANY) {
                    if (!(tag & RT_TAG_EMISSIVE) && t < tmax) return true;
                } else {
                    rt_consider(best, t, tag, slot);
                }
            }
        }
        if (sp == 0) break;
        sp--;
        node = sp < RB_STACK_SMEM ? stk[32 * sp] : deep[sp - RB_STACK_SMEM];
    }
    return false;
}

template <bool ANY, bool COUNT>
__device__ __forceinline__ bool rb_traverse(const RtSceneDev& sc, const vec3& o, const vec3& d, RtHit& best, float tmax,
                                            RtLocalCounters& cnt, int* __restrict__ stk, unsigned int* __restrict__ err) {
    if (sc.root == RT_NODE_SENTINEL) return false;
    const RtRayBox rb = rt_ray_box(o, d);
    if (rb.slack > RT_ROBUST_SLACK) return rb_traverse_loop<ANY, COUNT, true>(sc, rb, o, d, best, tmax, cnt, stk, err);
    return rb_traverse_loop<ANY, COUNT, false>(sc, rb, o, d, best, tmax, cnt, stk, err);
}

/* Packet traversal: closest hits of the warp's 32 primary rays in one walk.  All 32 lanes call; `live` = this lane
 * has a ray.  `wstk` = the warp's stack in shared memory.  Every slab test is the per-axis-slack one
 * (rt_slab_robust): it is at least as tight as the common-slack test for every ray and needs no per-ray choice. */
template <bool COUNT>
__device__ __forceinline__ void rb_packet_closest(const RtSceneDev& sc, bool live, const vec3& o, const vec3& d, RtHit& best,
                                                  RtLocalCounters& cnt, int* __restrict__ wstk, int lane,
                                                  unsigned int* __restrict__ err) {
    if (sc.root == RT_NODE_SENTINEL) return;
    const RtRayBox rb = rt_ray_box(o, d);
    const float4* __restrict__ nodes = sc.nodes;
    int node = sc.root;                           /* warp-uniform from here on */
    int sp = 0;
    for (;;) {
        if (node >= 0) {
            const float4 n0 = __ldg(nodes + 4 * node);
            const float4 n1 = __ldg(nodes + 4 * node + 1);
            const float4 n2 = __ldg(nodes + 4 * node + 2);
            const float4 n3 = __ldg(nodes + 4 * node + 3);
            const float lim = fmaxf(best.t, 0.0f);
            float t0, t1;
            const bool h0 = live && rt_slab_robust(rb, n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, lim, t0);
            const bool h1 = live && rt_slab_robust(rb, n1.z, n1.w, n2.x, n2.y, n2.z, n2.w, lim, t1);
            if (COUNT && live) cnt.node_tests += 2;
            const unsigned int b0 = __ballot_sync(0xffffffffu, h0), b1 = __ballot_sync(0xffffffffu, h1);
            const int c0 = __float_as_int(n3.x), c1 = __float_as_int(n3.y);
            if (b0 && b1) {
                /* the child most rays reach first goes first (rays that meet only one child vote for it) */
                const unsigned int first1 = __ballot_sync(0xffffffffu, h1 && (!h0 || t1 < t0));
                const bool one_first = 2 * __popc(first1) > __popc(b0 | b1);
                if (sp < RB_PACKET_STACK) {
                    if (lane == 0) wstk[sp] = one_first ? c0 : c1;
                    sp++;
                } else if (lane == 0) {
                    atomicOr(err, 1u);
                }
                node = one_first ? c1 : c0;
                continue;
            }
            if (b0) { node = c0; continue; }
            if (b1) { node = c1; continue; }
        } else {
            const int slot = ~node;
            float t;
            unsigned int tag;
            if (live && rt_test_record<COUNT>(sc.prims + 3 * (size_t)slot, o, d, t, tag, cnt)) rt_consider(best, t, tag, slot);
        }
        if (sp == 0) break;
        __syncwarp();                             /* lane 0's pushes are visible to everybody */
        sp--;
        node = wstk[sp];
        __syncwarp();                             /* ... and read before the slot is pushed over */
    }
}

/* scene accessor of the persistent tracer: RtGlobalScene's materials / normals / light-grid look-up; the closest hit
 * of the PRIMARY ray is handed in (packet traversal), everything else traverses with the shared-memory stack */
template <bool COUNT>
struct RtBigScene {
    const RtSceneDev& sc;
    int* stk;                         /* this thread's stack column */
    unsigned int* err;
    RtHit primary;                    /* closest hit of this lane's primary ray, when have_primary */
    bool have_primary;
    __device__ __forceinline__ RtBigScene(const RtSceneDev& s, int* st, unsigned int* e)
        : sc(s), stk(st), err(e), have_primary(false) {}

    __device__ __forceinline__ void planes_closest(const vec3& o, const vec3& d, RtHit& best, RtLocalCounters& cnt) const {
        for (int p = 0; p < sc.n_planes; p++) {
            const float4 q0 = __ldg(sc.planes + 2 * p), q1 = __ldg(sc.planes + 2 * p + 1);
            float t;
            if (COUNT) cnt.plane_tests++;
            if (rt_hit_plane(o, d, f4xyz(q0), f4xyz(q1), t)) rt_consider(best, t, __float_as_uint(q0.w), ~p);
        }
    }
    __device__ __forceinline__ RtHit closest(const vec3& o, const vec3& d, int depth, bool, RtLocalCounters& cnt) {
        if (depth == 0 && have_primary) return primary;
        RtHit best;
        best.t = __int_as_float(0x7f800000);
        best.tag = RT_MISS;
        best.slot = 0;
        planes_closest(o, d, best, cnt);
        rb_traverse<false, COUNT>(sc, o, d, best, 0.0f, cnt, stk, err);
        return best;
    }
    /* to_light = L - P (the vector the caller normalised into d), li = index of the light */
    __device__ __forceinline__ bool occluded(const vec3& o, const vec3& d, float tmax, const vec3& to_light, int li,
                                             RtLocalCounters& cnt) {
        const int g = rt_light_grid_occluded<COUNT>(sc, o, d, tmax, to_light, li, cnt);
        if (g >= 0) return g != 0;
        for (int p = 0; p < sc.n_planes; p++) {
            const float4 q0 = __ldg(sc.planes + 2 * p), q1 = __ldg(sc.planes + 2 * p + 1);
            float t;
            if (COUNT) cnt.plane_tests++;
            if (__float_as_uint(q0.w) & RT_TAG_EMISSIVE) continue;
            if (rt_hit_plane(o, d, f4xyz(q0), f4xyz(q1), t) && t < tmax) return true;
        }
        RtHit dummy;
        return rb_traverse<true, COUNT>(sc, o, d, dummy, tmax, cnt, stk, err);
    }
    __device__ __forceinline__ vec3 normal(const RtHit& h, const vec3& P) { return rt_hit_normal<false>(sc, h, P); }
    __device__ __forceinline__ void material(unsigned int idx, float4& m0, float4& m1) {
        m0 = __ldg(sc.mats + 2 * idx);
        m1 = __ldg(sc.mats + 2 * idx + 1);
    }
};

/* Which pixels of the patch [x0, x0+8) x [y0, y0+4) does this shard trace (own tiles + the 1-pixel FXAA apron)?
 * Whole frames: all of them.  Shards whose tiles are at least 10x6 pixels: the patch grown by one pixel spans at most
 * 2x2 tiles, so four owner look-ups (lanes 0..3, one corner each) settle the common cases -- no corner owned: nothing to
 * trace; all four owned: everything; only a patch on a tile border asks per pixel. */
__device__ __forceinline__ bool rb_patch_pixel_needed(const RtFrameDev& f, int x0, int y0, int x, int y, int lane) {
    const bool inside = x < f.imgw && y < f.imgh;
    if (f.n_shards <= 1) return inside;
    if (f.tile_w >= 10 && f.tile_h >= 6) {
        const int cx = (lane & 1) ? min(x0 + 8, f.imgw - 1) : max(x0 - 1, 0);
        const int cy = (lane & 2) ? min(y0 + 4, f.imgh - 1) : max(y0 - 1, 0);
        const unsigned int own = __ballot_sync(0xffffffffu, lane < 4 && rt_shard_owns(f, cx, cy)) & 15u;
        if (own == 0u) return false;
        if (own == 15u) return inside;
    }
    return inside && rt_shard_needs(f, x, y);
}

template <bool COUNT>
__global__ void __launch_bounds__(RB_THREADS, RB_CTAS_PER_SM)
rt_trace_big_kernel(const RtSceneDev sc, const RtFrameDev f, unsigned int* __restrict__ image, int* __restrict__ hit_ids,
                    RtCountersDev* __restrict__ counters, unsigned int* __restrict__ tile_counter,
                    unsigned int* __restrict__ err, int packet) {
    __shared__ float4 s_pending[RTB_MAX_DEPTH * RB_THREADS];
    __shared__ int s_stack[RB_WARPS][RB_STACK_SMEM][32];     /* per-lane stacks: entry e of lane l at [warp][e][l] */
    __shared__ int s_wstack[RB_WARPS][RB_PACKET_STACK];       /* per-warp stack of the packet traversal */
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int px_w = (f.imgw + 7) / 8, px_h = (f.imgh + 3) / 4;
    const unsigned int n_patches = (unsigned int)(px_w * px_h);

    unsigned int u = 0;
    if (lane == 0) u = atomicAdd(tile_counter + RB_CNT_PATCH, 1u);
    u = __shfl_sync(0xffffffffu, u, 0);

    RtBigScene<COUNT> acc(sc, &s_stack[warp][0][lane], err);
    RtPixelCounters pc = {{0, 0, 0, 0, 0, 0, 0, 0, 0}, 0, 0, 0, 0};
    unsigned int n_pixels = 0;

    while (u < n_patches) {
        const unsigned int p = n_patches - 1u - u;            /* bottom-up, like trace_small.cu */
        const int py = (int)(p / (unsigned int)px_w), pxi = (int)p - py * px_w;
        const int x = pxi * 8 + (lane & 7), y = py * 4 + (lane >> 3);
        const bool mine = rb_patch_pixel_needed(f, pxi * 8, py * 4, x, y, lane);
        if (__any_sync(0xffffffffu, mine)) {
            acc.have_primary = false;
            if (packet) {
                const vec3 o = v3(f.pos[0], f.pos[1], f.pos[2]);
                const vec3 d = mine ? rt_primary_dir(f, x, y) : v3(0.0f, 0.0f, 1.0f);
                RtHit best;
                best.t = __int_as_float(0x7f800000);
                best.tag = RT_MISS;
                best.slot = 0;
                if (mine) acc.planes_closest(o, d, best, pc.t);
                rb_packet_closest<COUNT>(sc, mine, o, d, best, pc.t, s_wstack[warp], lane, err);
                acc.primary = best;
                acc.have_primary = true;
            }
            if (mine) {
                image[y * f.imgw + x] = rt_trace_pixel<RtBigScene<COUNT>, COUNT>(acc, f, x, y, s_pending, RB_THREADS, tid,
                                                                               hit_ids, pc);
                n_pixels++;
            }
            __syncwarp();
        }
        unsigned int un = 0;
        if (lane == 0) un = atomicAdd(tile_counter + RB_CNT_PATCH, 1u);
        u = __shfl_sync(0xffffffffu, un, 0);
    }
    rb_griddep_launch();
    if (COUNT) rt_flush_counters(pc, n_pixels, counters);
    /* the last warp to finish re-arms the patch counter for the next launch on this stream */
    if (lane == 0) {
        __threadfence();
        if (atomicAdd(tile_counter + RB_CNT_DONE, 1u) == gridDim.x * RB_WARPS - 1) {
            tile_counter[RB_CNT_PATCH] = 0;
            tile_counter[RB_CNT_DONE] = 0;
        }
    }
}

extern "C" cudaError_t rt_launch_trace_big(const RtSceneDev* sc, const RtFrameDev* f, unsigned int* image, int* hit_ids,
                                           RtCountersDev* counters, unsigned int* tile_counter, unsigned int* err,
                                           int count, int packet, int n_sm, cudaStream_t stream) {
    const int patches = ((f->imgw + 7) / 8) * ((f->imgh + 3) / 4);
    int grid = n_sm * RB_CTAS_PER_SM;
    if (grid > (patches + RB_WARPS - 1) / RB_WARPS) grid = (patches + RB_WARPS - 1) / RB_WARPS;
    if (grid < 1) grid = 1;
    RT_LAUNCHED(1);
    if (count) rt_trace_big_kernel<true><<<grid, RB_THREADS, 0, stream>>>(*sc, *f, image, hit_ids, counters, tile_counter, err, packet);
    else       rt_trace_big_kernel<false><<<grid, RB_THREADS, 0, stream>>>(*sc, *f, image, hit_ids, counters, tile_counter, err, packet);
    return cudaGetLastError();
}
