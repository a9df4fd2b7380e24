/*
 * trace_big.cu -- persistent, warp-coherent tracer for scenes that do not fit shared memory (BASELINE config 4: a
 * million triangles).  The LBVH and the leaf records stay in HBM / the 126 MB L2 and are read with float4 loads
 * through the read-only path.  Same per-pixel arithmetic as trace.cu and trace_small.cu (trace_pixel.cuh), so
 * all three kernels agree bit for bit; only WHICH primitives get tested, and in what order, differs.
 *
 * What changes against the plain one-thread-per-pixel kernel (trace.cu):
 *   - persistent CTAs (4 per SM), every WARP autonomous: a CTA reserves chunks of 8x4 patches (64x16 pixels) from a
 *     global counter, its warps pull single 8x4-pixel patches of the chunk from a counter in shared memory; the
 *     frame is walked from its last chunk to its first, no CTA barrier in the loop; for a screen-tile shard a
 *     patch none of whose pixels the shard needs costs four owner look-ups, not a launch slot;
 *   - PACKET traversal for the primary rays.  The 32 primary rays of a patch form a thin bundle and visit almost
 *     the same nodes, so the warp walks the hierarchy ONCE for all of them: one warp-uniform node index, one stack
 *     per warp in shared memory, every lane tests ITS ray against the node's two child boxes, a child is entered
 *     when ANY lane's ray meets it (ballot), the child most lanes reach first goes first.  All 32 lanes load the
 *     same node / leaf record at the same time -- one broadcast LDG.128 wavefront instead of up to 32 -- and there
 *     is no divergence and no per-lane stack traffic.  A lane may so see primitives its own traversal would have
 *     skipped; the exact leaf tests decide, and the (t, index) minimum over a superset of the primitives a ray can
 *     hit is the reference's minimum over all objects (kernel.cu:144-151);
 *   - the per-lane traversal (bounces; shadow rays of a light without a grid) can keep the first RB_STACK_SMEM
 *     entries of its stack in SHARED memory (template SS; deeper pushes continue in a local-memory array).  Measured:
 *     the 16 KB per CTA this takes out of the unified L1 cost more than the conflict-free stack accesses gain -- the
 *     kernel lives on its 97 % L1 hit rate -- so the default keeps the whole stack in local memory (RT_OPT_BIG_STACK).
 *     Either way a push beyond the stack (a hierarchy deeper than the host-side check lets through, api.cu) raises
 *     the context's device error word (rt_get_device_errors) instead of dropping the subtree silently.
 */
#include "trace_global.cuh"

#define RB_THREADS 256
#define RB_WARPS (RB_THREADS / 32)
#ifndef RB_CTAS_PER_SM
#define RB_CTAS_PER_SM 4
#endif
#define RB_STACK_SMEM 16                      /* per-thread stack entries in shared memory (SS = 0: none) */
#define RB_CNT_PATCH 0                        /* tile_counter[]: next patch, finished warps (api.cu: d_tile_counter) */
#define RB_CNT_DONE 1

__device__ __forceinline__ void rb_griddep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

/* LBVH traversal of ONE ray with the short stack in shared memory.  `stk` = this thread's column (entry e at
 * stk[32 e]), entries beyond SS go to `deep`.  ANY: first non-emissive hit with t < tmax
 * (kernel.cu:188-193); else closest hit into `best`. */
template <bool ANY, bool COUNT, bool ROBUST, int SS>
__device__ __forceinline__ bool rb_traverse_loop(const RtSceneDev& sc, const RtRayBox& rb, const vec3& o, const vec3& d,
                                                 RtHit& best, float tmax, RtLocalCounters& cnt, int* __restrict__ stk,
                                                 unsigned int* __restrict__ err) {
    int node = sc.root;
    int deep[RT_STACK_SIZE];
    int sp = 0;
    bool overflow = false;                       /* reported by one atomic BEHIND the loop (trace_common.cuh) */
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
                if (sp < SS) {
                    stk[32 * sp] = c1;
                    sp++;
                } else if (sp < SS + RT_STACK_SIZE) {
                    deep[sp - SS] = c1;
                    sp++;
                } else {
                    overflow = true;             /* deeper than any hierarchy the host lets through: say so */
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
        node = sp < SS ? stk[32 * sp] : deep[sp - SS];
    }
    if (overflow) atomicOr(err, 1u);
    return false;
}

template <bool ANY, bool COUNT, int SS>
__device__ __forceinline__ bool rb_traverse(const RtSceneDev& sc, const vec3& o, const vec3& d, RtHit& best, float tmax,
                                            RtLocalCounters& cnt, int* __restrict__ stk, unsigned int* __restrict__ err) {
    if (sc.root == RT_NODE_SENTINEL) return false;
    const RtRayBox rb = rt_ray_box(o, d);
    if (rb.slack > RT_ROBUST_SLACK) return rb_traverse_loop<ANY, COUNT, true, SS>(sc, rb, o, d, best, tmax, cnt, stk, err);
    return rb_traverse_loop<ANY, COUNT, false, SS>(sc, rb, o, d, best, tmax, cnt, stk, err);
}

/* scene accessor of the persistent tracer: RtGlobalScene's materials / normals / light-grid look-up; the closest hit
 * of the PRIMARY ray is handed in (packet traversal), everything else traverses with the shared-memory stack */
template <bool COUNT, int SS>
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
        rb_traverse<false, COUNT, SS>(sc, o, d, best, 0.0f, cnt, stk, err);
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
        return rb_traverse<true, COUNT, SS>(sc, o, d, dummy, tmax, cnt, stk, err);
    }
    __device__ __forceinline__ vec3 normal(const RtHit& h, const vec3& P) { return rt_hit_normal<false>(sc, h, P); }
    __device__ __forceinline__ void material(unsigned int idx, float4& m0, float4& m1) {
        m0 = __ldg(sc.mats + 2 * idx);
        m1 = __ldg(sc.mats + 2 * idx + 1);
    }
};

template <bool COUNT, int SS>
__global__ void __launch_bounds__(RB_THREADS, RB_CTAS_PER_SM)
rt_trace_big_kernel(const RtSceneDev sc, const RtFrameDev f, unsigned int* __restrict__ image, int* __restrict__ hit_ids,
                    RtCountersDev* __restrict__ counters, unsigned int* __restrict__ tile_counter,
                    unsigned int* __restrict__ err, int packet) {
    __shared__ float4 s_pending[RTB_MAX_DEPTH * RB_THREADS];
    __shared__ int s_stack[RB_WARPS][SS > 0 ? SS : 1][SS > 0 ? 32 : 1];   /* per-lane stacks: entry e of lane l at [warp][e][l] */
    __shared__ int s_wstack[RB_WARPS][RT_PACKET_STACK];       /* per-warp stack of the packet traversal */
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    /* Work units.  The frame is cut into CHUNKS of 8x4 patches (64x16 pixels); a CTA reserves one chunk at a time from
     * the global counter and its eight warps take the chunk's patches one by one from a counter in shared memory,
     * each warp for itself.  The warps of an SM so work on neighbouring patches at about the same time and share
     * the nodes and leaf records they pull through L1 -- handing single patches to whichever warp of the GPU asks
     * next scatters neighbours over all SMs and was measured 38 % slower than the plain kernel.
     * s_pool = (chunk << 6) | k: k < 32 = next patch of the chunk, k = 32 = used up (whoever draws it refills),
     * k > 32 = a refill is under way (at most one draw per waiting warp: k <= 40). */
    __shared__ unsigned int s_pool;
    const int cw = (f.imgw + 63) / 64, ch = (f.imgh + 15) / 16;
    const unsigned int n_chunks = (unsigned int)(cw * ch);
    if (tid == 0) s_pool = 32u;
    __syncthreads();

    RtBigScene<COUNT, SS> acc(sc, &s_stack[warp][0][SS > 0 ? lane : 0], err);
    RtPixelCounters pc = {{0, 0, 0, 0, 0, 0, 0, 0, 0}, 0, 0, 0, 0};
    unsigned int n_pixels = 0;

    for (;;) {
        unsigned int unit = 0;
        if (lane == 0) {
            for (;;) {
                const unsigned int old = atomicAdd(&s_pool, 1u);
                const unsigned int k = old & 63u;
                if (k < 32u) { unit = old; break; }
                if (k == 32u) {                                   /* this warp refills the CTA's pool */
                    const unsigned int g = atomicAdd(tile_counter + RB_CNT_PATCH, 1u);
                    unit = g << 6;                                /* ... and takes the new chunk's first patch */
                    atomicExch(&s_pool, (g << 6) | 1u);
                    break;
                }
                while ((*(volatile unsigned int*)&s_pool & 63u) >= 32u) __nanosleep(64);
            }
        }
        unit = __shfl_sync(0xffffffffu, unit, 0);
        const unsigned int g = unit >> 6, k = unit & 63u;
        if (g >= n_chunks) break;                                 /* every later chunk is past the end too */
        const unsigned int c = n_chunks - 1u - g;                /* bottom-up, like trace_small.cu */
        const int cy = (int)(c / (unsigned int)cw), cx = (int)c - cy * cw;
        const int x0 = cx * 64 + (int)(k & 7u) * 8, y0 = cy * 16 + (int)(k >> 3) * 4;
        if (x0 >= f.imgw || y0 >= f.imgh) continue;
        const int x = x0 + (lane & 7), y = y0 + (lane >> 3);
        const bool mine = rt_patch_pixel_needed(f, x0, y0, x, y, lane);
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
                rt_packet_closest<COUNT>(sc, mine, o, d, best, pc.t, s_wstack[warp], lane, err);
                acc.primary = best;
                acc.have_primary = true;
            }
            if (mine) {
                image[y * f.imgw + x] = rt_trace_pixel<RtBigScene<COUNT, SS>, COUNT>(acc, f, x, y, s_pending, RB_THREADS, tid,
                                                                               hit_ids, pc);
                n_pixels++;
            }
            __syncwarp();
        }
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
                                           int count, int packet, int smem_stack, int n_sm, cudaStream_t stream) {
    const int chunks = ((f->imgw + 63) / 64) * ((f->imgh + 15) / 16);     /* a CTA works through whole chunks */
    int grid = n_sm * RB_CTAS_PER_SM;
    if (grid > chunks) grid = chunks;
    if (grid < 1) grid = 1;
    RT_LAUNCHED(1);
#define RB_LAUNCH(C, S) rt_trace_big_kernel<C, S><<<grid, RB_THREADS, 0, stream>>>(*sc, *f, image, hit_ids, counters, tile_counter, err, packet)
    if (smem_stack) { if (count) RB_LAUNCH(true, RB_STACK_SMEM); else RB_LAUNCH(false, RB_STACK_SMEM); }
    else            { if (count) RB_LAUNCH(true, 0); else RB_LAUNCH(false, 0); }
#undef RB_LAUNCH
    return cudaGetLastError();
}
